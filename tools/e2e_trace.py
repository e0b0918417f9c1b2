"""Host-side trace of one end-to-end cull (ovis_cull with host output, 12.5 M entities): how long the two launches take to
queue and when the result is there (OVIS_MV_TRACE=1 prints the library's own marks). Numbers in profiles/r02_cull_notes.md."""
import os, sys, time
sys.path.insert(0, '/root/repo')
import numpy as np, torch
from orion_b200 import lib, scenes
n = 12_500_000
aabb, flags = scenes.static_boxes(n, 4242, 4000.0)
w = lib.World(0); w.reserve(n); f = w.add_entities(aabb, flags); w.set_world_aabbs(f, aabb)
view = scenes.default_camera()
pinned = lib.PinnedBuffer(1 << 16)
step = w.bind_cull(view, pinned.array)
for _ in range(20): step()
os.environ['X']='1'
t0=time.perf_counter()
for _ in range(200): step()
print("e2e us", (time.perf_counter()-t0)/200*1e6, file=sys.stderr)
