#!/usr/bin/env python
"""BASELINE.json's metric names "cull ms/frame, 1M-100M entities": the single-view cull of one GPU over that range.
Per size: device-resident ms per frame (CUDA events over back-to-back frames), end to end through ovis_cull into pinned host
memory (wall clock), bytes per second against the measured HBM peak, parity of the list against the oracle, and the
reference's arithmetic on all host cores beside it. One JSON document.

    python tools/size_sweep.py [sizes in millions ...]"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from orion_b200 import lib, scenes  # noqa: E402
from oracle import pyoracle  # noqa: E402  (the checker and the CPU baseline, as in bench.py)


def peak_gbs():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        return 6546.2


def one(n, oracle, cores, frames=60):
    extent = 4000.0 * (n / 12.5e6) ** (1.0 / 3.0)                 # the bench's density: the visible fraction stays ~0.09 %
    aabb, flags = scenes.static_boxes(n, 4242, extent)
    view = scenes.default_camera()
    planes = lib.view_planes(view)
    w = lib.World(0)
    stream = torch.cuda.Stream()
    w.set_stream(stream.cuda_stream)
    w.reserve(n)
    first = w.add_entities(aabb, flags)
    w.set_world_aabbs(first, aabb)
    out = torch.zeros(n + 1, dtype=torch.int32, device="cuda")
    for _ in range(5):
        w.cull_planes_device(planes, out.data_ptr() + 4, out.data_ptr())
    stream.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(stream)
    for _ in range(frames):
        w.cull_planes_device(planes, out.data_ptr() + 4, out.data_ptr())
    b.record(stream)
    stream.synchronize()
    ms = a.elapsed_time(b) / frames
    got = out.cpu().numpy().view(np.uint32)
    ids = got[1:1 + got[0]]
    pinned = lib.PinnedBuffer(max(int(got[0]) * 2, 1024))
    step = w.bind_cull(view, pinned.array)
    for _ in range(5):
        step()
    t0 = time.perf_counter()
    for _ in range(frames):
        k, _ = step()
    e2e_ms = (time.perf_counter() - t0) / frames * 1e3
    t0 = time.perf_counter()
    exp = oracle.cull_aabb(aabb, planes, threads=cores)
    cpu_ms = (time.perf_counter() - t0) * 1e3
    row = {"entities": n, "visible": int(len(exp)), "device_ms_per_frame": ms, "tests_per_s": n / (ms * 1e-3),
           "achieved_gbs": (24.0 * n + 4.0 * len(exp)) / (ms * 1e-3) / 1e9, "e2e_ms_per_frame": e2e_ms,
           "list_exact": bool(np.array_equal(ids, exp) and np.array_equal(pinned.array[:k], exp)),
           "reference_cpu_ms": cpu_ms, "cores": cores}
    row["frac_of_hbm_peak"] = row["achieved_gbs"] / peak_gbs()
    w.close()
    return row


def main():
    sizes = [float(x) for x in sys.argv[1:]] or [1, 2, 5, 12.5, 25, 50, 100]
    oracle = pyoracle.load("best")
    cores = os.cpu_count() or 1
    rows = []
    for m in sizes:
        rows.append(one(int(m * 1e6), oracle, cores))
        print(json.dumps(rows[-1]), file=sys.stderr)
        torch.cuda.empty_cache()
    print(json.dumps({"what": "single-view cull on one B200 over BASELINE's size range; bounds larger than L2 from 5.25 M entities up",
                      "hbm_peak_gbs": peak_gbs(), "oracle": oracle.name, "rows": rows}))


if __name__ == "__main__":
    main()
