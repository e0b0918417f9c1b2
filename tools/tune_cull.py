#!/usr/bin/env python
"""Time the single-view cull kernel for every launch shape (OVIS_CULL_VARIANT) at a few sizes and check each
against the oracle. Tuning aid, run on the GPU box: python tools/tune_cull.py [variant ...]"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

CHILD = r'''
import json, os, sys, numpy as np, torch
sys.path.insert(0, %r)
from orion_b200 import lib, scenes
from oracle import pyoracle
oracle = pyoracle.load("best")
res = {}
stream = torch.cuda.Stream()
for n, extent in ((1_000_000, 1000.0), (12_500_000, 4000.0), (12_500_000, 1000.0)):
    aabb, flags = scenes.static_boxes(n, 4000, extent)
    view = scenes.default_camera()
    planes = lib.view_planes(view)
    w = lib.World(0); w.set_stream(stream.cuda_stream)
    f = w.add_entities(aabb, flags); w.set_world_aabbs(f, aabb)
    out = torch.zeros(n + 1, dtype=torch.int32, device="cuda")
    def run(req=0):
        w.cull_planes_device(planes, out.data_ptr() + 4, out.data_ptr(), req)
    for _ in range(5): run()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(50)]
    scratch = torch.zeros(64 * 1024 * 1024, dtype=torch.int32, device="cuda")
    with torch.cuda.stream(stream):
        for a, b in evs:
            if n < 5_000_000: scratch.add_(1)
            a.record(stream); run(); b.record(stream)
    stream.synchronize()
    ms = np.array([a.elapsed_time(b) for a, b in evs])
    vis = int(out[0].item())
    got = out[1:1 + vis].cpu().numpy().view(np.uint32)
    ok = bool(np.array_equal(got, oracle.cull_aabb(aabb, planes, threads=16)))
    res[f"{n}@{extent:g}"] = {"ms_med": float(np.median(ms)), "ms_min": float(ms.min()), "vis": vis, "ok": ok,
                              "GBs": (24 * n + 4 * vis) / np.median(ms) / 1e6}
    w.close(); del scratch
print(json.dumps(res))
''' % ROOT

variants = [int(v) for v in sys.argv[1:]] or list(range(6))
for v in variants:
    env = dict(os.environ, OVIS_CULL_VARIANT=str(v))
    p = subprocess.run([sys.executable, "-c", CHILD], env=env, capture_output=True, text=True)
    line = p.stdout.strip().splitlines()[-1] if p.stdout.strip() else p.stderr[-400:]
    print(f"variant {v}: {line}", flush=True)
