#!/usr/bin/env python
"""Timing + parity of BASELINE.json configs[2] and configs[3] on one GPU (not the bench.py line; evidence for DESIGN.md).

  config 3: 10M dynamic entities with a depth-3 transform hierarchy + 4096 point lights: per frame all depth-1 nodes
            move -> full tree update -> main cull (+ lights) -> shadow-caster lists of the visible lights.
  config 4: 10M static entities x (main camera + 4 cascade views + 64 point lights x 6 cube faces) = 389 views.

Every stage is timed with CUDA events on the world's stream; results are checked against the CPU oracle (full size for
transforms / bounds / main view, a sample of the shadow views). Prints one JSON object.
    python tools/bench_configs.py [3] [4] [--small]
"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from orion_b200 import lib, scenes  # noqa: E402
from oracle import pyoracle  # noqa: E402

THREADS = len(os.sched_getaffinity(0))


def timed(stream, fn, iters=5, warm=2):
    for _ in range(warm):
        fn()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    out = None
    for a, b in evs:
        a.record(stream)
        out = fn()
        b.record(stream)
    stream.synchronize()
    return float(np.median([a.elapsed_time(b) for a, b in evs])), out


def config3(small):
    oracle = pyoracle.load("best")
    kw = dict(n_groups=1000, n_parts=900_000, n_lights=512) if small else {}
    t0 = time.perf_counter()
    s = scenes.config3(**kw)
    gen_s = time.perf_counter() - t0
    n_ent, n_nodes = s["n_entities"], s["n_nodes"]
    stream = torch.cuda.Stream()
    w = lib.World(0)
    w.set_stream(stream.cuda_stream)
    w.reserve(n_ent, n_nodes)
    t0 = time.perf_counter()
    e0 = w.add_entities(s["local_aabb"], s["flags"])
    n0 = w.add_nodes(s["parent"], s["local_trs"])
    w.bind_entities(s["entity_node"], np.arange(e0, e0 + n_ent, dtype=np.uint32))
    w.add_lights(s["lights"])
    w.update_transforms()
    stream.synchronize()
    build_s = time.perf_counter() - t0
    g0, g1 = s["group_range"]
    group_ids = np.arange(g0, g1, dtype=np.uint32)
    frame = [0]

    def update():
        frame[0] += 1
        w.set_node_local_trs(group_ids, scenes.move_groups(s, frame[0]))
        w.update_transforms()

    # pre-generate the moves so host RNG time is not inside the events
    moves = [scenes.move_groups(s, f) for f in range(1, 9)]

    def update_pre():
        frame[0] = frame[0] % 8 + 1
        w.set_node_local_trs(group_ids, moves[frame[0] - 1])
        w.update_transforms()

    upd_ms, _ = timed(stream, update_pre, iters=7)
    # a frame in which only 1 % of the groups move (dirty-subtree update)
    few = group_ids[:: 100]
    few_moves = [m[:: 100] for m in moves]

    def update_few():
        frame[0] = frame[0] % 8 + 1
        w.set_node_local_trs(few, few_moves[frame[0] - 1])
        w.update_transforms()

    upd_few_ms, _ = timed(stream, update_few, iters=7)
    update_pre()                       # leave the tree in a fully known state for the parity check below
    planes = lib.view_planes(s["view"])
    out = torch.zeros(n_ent + 1, dtype=torch.int32, device="cuda")
    cull_ms, _ = timed(stream, lambda: w.cull_planes_device(planes, out.data_ptr() + 4, out.data_ptr()), iters=10)
    t0 = time.perf_counter()
    ents, ls = w.cull(s["view"])
    full_cull_host_ms = (time.perf_counter() - t0) * 1e3
    shadow_ids = ls[(s["lights"]["flags"][ls] & 1) == 1]
    lists_ms, res = timed(stream, lambda: w.cull_shadow_views_device(shadow_ids), iters=5)
    ll, lv, off, d_off, d_ent = res
    t0 = time.perf_counter()
    full = w.shadow_lists(s["view"])
    lists_host_ms = (time.perf_counter() - t0) * 1e3
    n_views = 1 + len(ll)

    # ---- parity (state after the last update)
    local = s["local_trs"].copy()
    local[g0:g1] = moves[frame[0] - 1]
    t0 = time.perf_counter()
    wt = oracle.world_trs(s["parent"], local)
    wb = oracle.world_aabb(wt[s["entity_node"]], s["local_aabb"])
    cpu_update_s = time.perf_counter() - t0
    ok_trs = bool(np.array_equal(w.node_world_trs(), wt, equal_nan=True))
    ok_aabb = bool(np.array_equal(w.world_aabbs(), wb, equal_nan=True))
    t0 = time.perf_counter()
    exp = oracle.cull_aabb(wb, planes, threads=THREADS)
    cpu_cull_s = time.perf_counter() - t0
    ok_main = bool(np.array_equal(ents, exp)) and bool(np.array_equal(ls, oracle.light_cull(s["lights"], planes)))
    ent_all = w.read_device(d_ent, int(off[-1]))
    sample = list(range(0, len(ll), max(1, len(ll) // 24)))[:24]
    ok_lists = True
    cpu_view_s = 0.0
    for k in sample:
        v = oracle.light_shadow_views(s["lights"][ll[k]])[lv[k]]
        t0 = time.perf_counter()
        e = oracle.cull_aabb(wb, oracle.view_planes(v), s["flags"], 1, threads=THREADS)
        cpu_view_s += time.perf_counter() - t0
        ok_lists &= bool(np.array_equal(ent_all[off[k]:off[k + 1]], e))
    ok_lists &= all(np.array_equal(a, b) for a, b in zip(full, (ll, lv, off, ent_all)))
    w.close()
    return {
        "workload": f"configs[2]: {n_ent / 1e6:g}M dynamic entities, depth-3 hierarchy ({n_nodes} nodes), {len(s['lights'])} point lights",
        "scene_generation_s": gen_s, "world_build_s": build_s,
        "update_ms": upd_ms, "update_1pct_of_groups_ms": upd_few_ms, "update_bytes_per_node": 132,
        "update_achieved_gbs": 132 * n_nodes / (upd_ms * 1e-3) / 1e9,
        "main_cull_ms": cull_ms, "visible_entities": int(len(ents)), "visible_lights": int(len(ls)),
        "cull_with_lights_host_api_ms": full_cull_host_ms,
        "shadow_views": int(len(ll)), "shadow_pairs": int(off[-1]) if len(off) else 0,
        "shadow_lists_device_ms": lists_ms, "shadow_lists_host_api_ms": lists_host_ms,
        "frame_ms_update_cull_lists": upd_ms + cull_ms + lists_ms,
        "tests_per_s_brute_force_equivalent": n_ent * n_views / ((upd_ms + cull_ms + lists_ms) * 1e-3),
        "parity": {"world_trs_exact": ok_trs, "world_aabb_exact": ok_aabb, "main_view_and_lights": ok_main,
                   "shadow_lists_sampled_and_api_consistent": bool(ok_lists), "sampled_lists": len(sample), "oracle": oracle.name},
        "cpu": {"threads": THREADS, "update_1thread_s": cpu_update_s, "main_cull_s": cpu_cull_s,
                "per_shadow_view_cull_s": cpu_view_s / max(len(sample), 1),
                "frame_estimate_s": cpu_update_s + cpu_cull_s + cpu_view_s / max(len(sample), 1) * len(ll)},
    }


def config4(small):
    oracle = pyoracle.load("best")
    s = scenes.config4(n=1_000_000 if small else 10_000_000)
    n = len(s["aabb"])
    stream = torch.cuda.Stream()
    w = lib.World(0)
    w.set_stream(stream.cuda_stream)
    w.reserve(n)
    f = w.add_entities(s["aabb"], s["flags"])
    w.set_world_aabbs(f, s["aabb"])
    w.add_lights(s["lights"])
    cam_views = np.concatenate([np.asarray(s["view"]).reshape(1), s["cascades"]])
    cam_planes = np.stack([lib.view_planes(v).reshape(24) for v in cam_views])
    cam_ms, (coff, _, cd_ent) = timed(stream, lambda: w.cull_views_device(cam_planes), iters=5)
    cam_ent = w.read_device(cd_ent, int(coff[-1]))
    ok_cam = all(np.array_equal(cam_ent[coff[i]:coff[i + 1]], oracle.cull_aabb(s["aabb"], cam_planes[i], threads=THREADS))
                 for i in range(len(cam_views)))
    light_ids = np.arange(len(s["lights"]), dtype=np.uint32)
    lists_ms, res = timed(stream, lambda: w.cull_shadow_views_device(light_ids), iters=5)
    ll, lv, off, d_off, d_ent = res
    ent_all = w.read_device(d_ent, int(off[-1]))
    sample = list(range(0, len(ll), 16))
    ok_lists = True
    cpu_view_s = 0.0
    for k in sample:
        v = oracle.light_shadow_views(s["lights"][ll[k]])[lv[k]]
        t0 = time.perf_counter()
        e = oracle.cull_aabb(s["aabb"], oracle.view_planes(v), s["flags"], 1, threads=THREADS)
        cpu_view_s += time.perf_counter() - t0
        ok_lists &= bool(np.array_equal(ent_all[off[k]:off[k + 1]], e))
    # brute-force comparison point on the GPU: the same 384 views without the pre-filter
    face_planes = []
    for l in s["lights"]:
        for v in lib.light_shadow_views(np.asarray(l)):
            face_planes.append(lib.view_planes(v).reshape(24))
    face_planes = np.stack(face_planes)
    brute_ms, (boff, _, bd_ent) = timed(stream, lambda: w.cull_views_device(face_planes, np.ones(len(face_planes), dtype=np.uint32)),
                                        iters=3, warm=1)
    ok_brute = bool(np.array_equal(boff, off)) and bool(np.array_equal(w.read_device(bd_ent, int(boff[-1])), ent_all))
    n_views = len(cam_views) + len(ll)
    w.close()
    return {
        "workload": f"configs[3]: {n / 1e6:g}M static entities x (1 camera + 4 cascade views + {len(s['lights'])} point lights x 6 faces) = {n_views} views",
        "camera_and_cascades_ms": cam_ms, "camera_visible": [int(coff[i + 1] - coff[i]) for i in range(len(cam_views))],
        "light_faces_ms": lists_ms, "light_face_pairs": int(off[-1]),
        "frame_ms": cam_ms + lists_ms,
        "tests_per_s_brute_force_equivalent": n * n_views / ((cam_ms + lists_ms) * 1e-3),
        "light_faces_without_prefilter_ms": brute_ms,
        "parity": {"camera_views": bool(ok_cam), "light_faces_sampled": bool(ok_lists), "sampled_lists": len(sample),
                   "prefilter_equals_brute_force_all_384": ok_brute, "oracle": oracle.name},
        "cpu": {"threads": THREADS, "per_view_cull_s": cpu_view_s / len(sample), "frame_estimate_s": cpu_view_s / len(sample) * n_views},
    }


if __name__ == "__main__":
    small = "--small" in sys.argv
    which = [a for a in sys.argv[1:] if a in ("3", "4")] or ["3", "4"]
    out = {}
    if "3" in which:
        out["config3"] = config3(small)
    if "4" in which:
        out["config4"] = config4(small)
    print(json.dumps(out, indent=1))
