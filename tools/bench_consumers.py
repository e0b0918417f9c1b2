#!/usr/bin/env python
"""Timing + parity of the consumers of the visible list on one GPU: draw-list
expansion and EntityUniforms packing, right after the cull, on the device. Evidence for DESIGN.md, not the bench.py line.

  static:   configs[4]'s 12.5M-entity shard (static boxes), one camera: cull -> draw lists -> uniforms
  dynamic:  a 2M-node tree (every entity on its own node): cull -> uniforms gathered from the world TRS

Each stage is timed with CUDA events on the world's stream (median of the iterations, inputs larger than L2 for the
cull; the consumers work on the visible list only). The CPU oracle runs the same stages beside it on one core.
    python tools/bench_consumers.py [--small]
"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from orion_b200 import lib, scenes  # noqa: E402
from oracle import pyoracle  # noqa: E402
import vis_testutil as util  # noqa: E402


def timed(stream, fn, iters=20, warm=3):
    for _ in range(warm):
        fn()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for a, b in evs:
        a.record(stream)
        fn()
        b.record(stream)
    stream.synchronize()
    return float(np.median([a.elapsed_time(b) for a, b in evs]))


def cpu_time(fn, reps=3):
    best = None
    out = None
    for _ in range(reps):
        t0 = time.perf_counter()
        out = fn()
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return best * 1e3, out


def static_case(n, oracle):
    s = scenes.config5(n=n, seed=4)
    rng = np.random.default_rng(8)
    pc = util.random_pass_counts(rng, n)
    stream = torch.cuda.Stream()
    w = lib.World(0)
    w.set_stream(stream.cuda_stream)
    w.reserve(n, 0)
    first = w.add_entities(s["aabb"], s["flags"])
    w.set_world_aabbs(first, s["aabb"])
    w.set_pass_counts(np.arange(n), pc & 0xff, (pc >> 8) & 0xff, pc >> 16)
    planes = lib.view_planes(s["view"])
    with torch.cuda.stream(stream):
        d_list = torch.zeros(n + 1, dtype=torch.int32, device="cuda")
        cap = n
        d0 = torch.zeros((cap, 2), dtype=torch.int32, device="cuda")
        d1 = torch.zeros((cap, 2), dtype=torch.int32, device="cuda")
        cnt = torch.zeros(2, dtype=torch.int32, device="cuda")
        uni = torch.zeros((n // 4, 20), dtype=torch.float32, device="cuda")
    stream.synchronize()

    def cull():
        w.cull_planes_device(planes, d_list.data_ptr() + 4, d_list.data_ptr())

    cull()
    stream.synchronize()
    m = int(d_list[0].item())

    def draws(mode):
        w.build_draw_lists_device(mode, d0.data_ptr(), cap, d1.data_ptr(), cap, d_list.data_ptr() + 4, d_list.data_ptr(), m,
                                  cnt.data_ptr())

    def uniforms():
        w.pack_entity_uniforms_device(uni.data_ptr(), len(uni), d_list.data_ptr() + 4, d_list.data_ptr(), m)

    def frame():
        cull()
        draws(lib.DRAWS_MAIN)
        uniforms()

    t = {"cull_ms": timed(stream, cull), "draws_main_ms": timed(stream, lambda: draws(lib.DRAWS_MAIN)),
         "draws_shadow_ms": timed(stream, lambda: draws(lib.DRAWS_SHADOW)), "uniforms_ms": timed(stream, uniforms),
         "frame_ms": timed(stream, frame)}
    draws(lib.DRAWS_MAIN)
    stream.synchronize()
    c = cnt.cpu().numpy()
    visible = d_list[1:1 + m].cpu().numpy().view(np.uint32)
    cpu_ms, (e0, e1) = cpu_time(lambda: oracle.draw_lists(visible, pc, s["flags"], 0))
    ok = (int(c[0]), int(c[1])) == (len(e0), len(e1)) and np.array_equal(d0[:c[0]].cpu().numpy().view(np.uint32), e0) and \
        np.array_equal(d1[:c[1]].cpu().numpy().view(np.uint32), e1)
    ident = np.tile(np.array([0, 0, 0, 1, 0, 0, 0, 1, 1, 1], dtype=np.float32), (m, 1))
    cpu_u_ms, eu = cpu_time(lambda: oracle.entity_uniforms(ident))
    ok_u = util.same(uni[:m].cpu().numpy(), eu)
    n_draws = int(c[0]) + int(c[1])
    t.update({"entities": n, "visible": m, "draws": [int(c[0]), int(c[1])], "parity_draws": bool(ok), "parity_uniforms": bool(ok_u),
              "cpu_draws_ms_1core": cpu_ms, "cpu_uniforms_ms_1core": cpu_u_ms,
              # list read + pass-count gather + draw write; list read + node gather + 80 B write (no TRS: static geometry)
              "draws_alg_bytes": 8 * m * 2 + 8 * n_draws, "uniforms_alg_bytes": m * (4 + 4 + 80)})
    t["draws_main_GBps"] = t["draws_alg_bytes"] / t["draws_main_ms"] / 1e6
    t["uniforms_GBps"] = t["uniforms_alg_bytes"] / t["uniforms_ms"] / 1e6
    w.close()
    return t


def dynamic_case(n, oracle):
    rng = np.random.default_rng(9)
    fan = 100
    groups = n // fan
    parent = np.empty(n, dtype=np.int32)
    parent[:groups] = -1
    parent[groups:] = rng.integers(0, groups, n - groups)
    local = util.random_trs(rng, n, pos=30.0)
    local[:, 7:10] = rng.uniform(0.5, 1.5, (n, 3))
    local[:groups, 0:3] = rng.uniform(-600, 600, (groups, 3))
    boxes = util.random_boxes(rng, n, extent=1.0, hmax=1.0)
    stream = torch.cuda.Stream()
    w = lib.World(0)
    w.set_stream(stream.cuda_stream)
    w.reserve(n, n)
    e0 = w.add_entities(boxes)
    n0 = w.add_nodes(parent, local)
    w.bind_entities(np.arange(n0, n0 + n, dtype=np.uint32), np.arange(e0, e0 + n, dtype=np.uint32))
    w.update_transforms()
    planes = lib.view_planes(scenes.default_camera())
    with torch.cuda.stream(stream):
        d_list = torch.zeros(n + 1, dtype=torch.int32, device="cuda")
        uni = torch.zeros((n, 20), dtype=torch.float32, device="cuda")
    stream.synchronize()
    w.cull_planes_device(planes, d_list.data_ptr() + 4, d_list.data_ptr())
    stream.synchronize()
    m = int(d_list[0].item())

    def uniforms():
        w.pack_entity_uniforms_device(uni.data_ptr(), len(uni), d_list.data_ptr() + 4, d_list.data_ptr(), m)

    t = {"entities": n, "visible": m, "uniforms_ms": timed(stream, uniforms)}
    visible = d_list[1:1 + m].cpu().numpy().view(np.uint32)
    wt = w.node_world_trs()
    cpu_ms, eu = cpu_time(lambda: oracle.entity_uniforms(wt[visible]))
    t["parity_uniforms"] = bool(util.same(uni[:m].cpu().numpy(), eu))
    t["cpu_uniforms_ms_1core"] = cpu_ms
    t["uniforms_alg_bytes"] = m * (4 + 4 + 40 + 80)
    t["uniforms_GBps"] = t["uniforms_alg_bytes"] / t["uniforms_ms"] / 1e6
    # every entity (the reference rewrites the uniforms of everything that moved)
    everyone = torch.arange(n, dtype=torch.int32, device="cuda")
    ecnt = torch.tensor([n], dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()

    def uniforms_all():
        w.pack_entity_uniforms_device(uni.data_ptr(), len(uni), everyone.data_ptr(), ecnt.data_ptr(), n)

    t["uniforms_all_ms"] = timed(stream, uniforms_all)
    t["uniforms_all_GBps"] = n * 128 / t["uniforms_all_ms"] / 1e6
    w.close()
    return t


def main():
    small = "--small" in sys.argv
    oracle = pyoracle.load("best")
    out = {"oracle": oracle.name, "gpu": torch.cuda.get_device_name(0),
           "static": static_case(1_000_000 if small else 12_500_000, oracle),
           "dynamic": dynamic_case(200_000 if small else 2_000_000, oracle)}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
