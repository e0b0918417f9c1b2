#!/usr/bin/env python
"""Test-kernel / gather-kernel split of the single-view cull per OVIS_CULL_VARIANT (library CUDA events), 12.5 M entities.
Tuning aid, run on the GPU box: python tools/split_cull.py [variant ...]"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r'''
import json, sys, numpy as np, torch
sys.path.insert(0, %r)
from orion_b200 import lib, scenes
res = {}
stream = torch.cuda.Stream()
for n, extent in ((12_500_000, 4000.0), (12_500_000, 1000.0), (2_000_000, 1000.0)):
    aabb, flags = scenes.static_boxes(n, 4000, extent)
    planes = lib.view_planes(scenes.default_camera())
    w = lib.World(0); w.set_stream(stream.cuda_stream)
    f = w.add_entities(aabb, flags); w.set_world_aabbs(f, aabb)
    out = torch.zeros(n + 1, dtype=torch.int32, device="cuda")
    for _ in range(5): w.cull_planes_device(planes, out.data_ptr() + 4, out.data_ptr())
    w.set_timing(True)
    for _ in range(50): w.cull_planes_device(planes, out.data_ptr() + 4, out.data_ptr())
    t, g = w.read_timing(); w.set_timing(False)
    # back to back, no events in between: what a frame loop sees
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        a.record(stream)
        for _ in range(100): w.cull_planes_device(planes, out.data_ptr() + 4, out.data_ptr())
        b.record(stream)
    stream.synchronize()
    vis = int(out[0].item())
    res[f"{n}@{extent:g}"] = {"test_us": round(float(np.median(t)) * 1e3, 2), "gather_us": round(float(np.median(g)) * 1e3, 2),
                              "frame_us": round(a.elapsed_time(b) * 10, 2), "vis": vis,
                              "test_GBs": round((24 * n + 4 * vis) / float(np.median(t)) / 1e6)}
    w.close()
print(json.dumps(res))
''' % ROOT
variants = [int(v) for v in sys.argv[1:]] or list(range(6))
for v in variants:
    env = dict(os.environ, OVIS_CULL_VARIANT=str(v))
    p = subprocess.run([sys.executable, "-c", CHILD], env=env, capture_output=True, text=True)
    line = p.stdout.strip().splitlines()[-1] if p.stdout.strip() else p.stderr[-600:]
    print(f"variant {v}: {line}", flush=True)
