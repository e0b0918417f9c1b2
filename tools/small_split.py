"""Test-kernel / gather-kernel split of the single-view cull at small sizes (0.1, 1 and 4 M entities, config-2 distribution),
from the library's CUDA events: where the single-launch path and the two-kernel path cross."""
import sys, numpy as np, torch
sys.path.insert(0, "/root/repo")
from orion_b200 import lib, scenes
for n in (100_000, 1_000_000, 4_000_000):
    s = scenes.config2(n)
    stream = torch.cuda.Stream()
    w = lib.World(0); w.set_stream(stream.cuda_stream)
    f = w.add_entities(s["aabb"], s["flags"]); w.set_world_aabbs(f, s["aabb"])
    planes = lib.view_planes(s["view"])
    out = torch.zeros(n + 1, dtype=torch.int32, device="cuda")
    scratch = torch.zeros(64 * 1024 * 1024, dtype=torch.int32, device="cuda")
    for _ in range(5): w.cull_planes_device(planes, out.data_ptr() + 4, out.data_ptr())
    w.set_timing(True)
    with torch.cuda.stream(stream):
        for _ in range(30):
            scratch.add_(1)
            w.cull_planes_device(planes, out.data_ptr() + 4, out.data_ptr())
    t, g = w.read_timing()
    print(n, "cold  test us %.2f gather us %.2f" % (np.median(t) * 1e3, np.median(g) * 1e3))
    with torch.cuda.stream(stream):
        for _ in range(30):
            w.cull_planes_device(planes, out.data_ptr() + 4, out.data_ptr())
    t, g = w.read_timing()
    print(n, "warm  test us %.2f gather us %.2f" % (np.median(t) * 1e3, np.median(g) * 1e3))
    w.close()
