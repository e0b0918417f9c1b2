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


def hbm_peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        return 6650.0


def issue_peak():
    """The no-FMA fp32 issue-rate peak measured by tools/issue_probe.cu on this pool's B200 (None if not recorded)."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "r02_issue_peak.json")))
    except Exception:
        return None


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


def config3(small, full_parity=True):
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
    # the tree kernels alone (K1 + K2): the moved groups are uploaded and scattered first, untimed
    kern = []
    for _ in range(7):
        frame[0] = frame[0] % 8 + 1
        w.set_node_local_trs(group_ids, moves[frame[0] - 1])
        stream.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        w.update_transforms()
        b.record(stream)
        stream.synchronize()
        kern.append(a.elapsed_time(b))
    upd_kernels_ms = float(np.median(kern))
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
    ents, ls = w.cull(s["view"])                      # first call: uploads the 4096 lights' records and uniform blocks
    first_cull_host_ms = (time.perf_counter() - t0) * 1e3
    pinned_e, pinned_l = lib.PinnedBuffer(n_ent), np.empty(len(s["lights"]), dtype=np.uint32)
    frame_cull = w.bind_cull(s["view"], pinned_e.array, flags=lib.CULL_LIGHTS, out_lights=pinned_l)
    frame_cull()
    t0 = time.perf_counter()
    for _ in range(20):
        frame_cull()
    full_cull_host_ms = (time.perf_counter() - t0) / 20 * 1e3
    shadow_ids = ls[(s["lights"]["flags"][ls] & 1) == 1]
    # timed through calls whose arguments were marshalled once (the C-ABI cost; the convenience wrappers allocate per call)
    lists_ms, _ = timed(stream, w.bind_cull_shadow_views_device(shadow_ids), iters=9)
    ll, lv, off, d_off, d_ent = w.cull_shadow_views_device(shadow_ids)
    full = w.shadow_lists(s["view"])
    frame_lists = w.bind_shadow_lists(s["view"], cap_entities=max(4 * len(full[3]), 1 << 16), pinned=True)
    frame_lists()
    t0 = time.perf_counter()
    for _ in range(9):
        frame_lists()
    lists_host_ms = (time.perf_counter() - t0) / 9 * 1e3
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
    sample = list(range(len(ll))) if full_parity else list(range(0, len(ll), max(1, len(ll) // 24)))[:24]
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
    peak = hbm_peak()
    frame_ms = upd_ms + cull_ms + lists_ms
    cpu_frame_s = cpu_update_s + cpu_cull_s + cpu_view_s / max(len(sample), 1) * len(ll)
    return {
        "workload": f"configs[2]: {n_ent / 1e6:g}M dynamic entities, depth-3 hierarchy ({n_nodes} nodes), {len(s['lights'])} point lights",
        "scene_generation_s": gen_s, "world_build_s": build_s,
        "update_ms": upd_ms, "update_1pct_of_groups_ms": upd_few_ms, "update_bytes_per_node": 132,
        "update_achieved_gbs": 132 * n_nodes / (upd_ms * 1e-3) / 1e9,
        "main_cull_ms": cull_ms, "visible_entities": int(len(ents)), "visible_lights": int(len(ls)),
        "cull_with_lights_host_api_ms": full_cull_host_ms, "first_cull_with_lights_host_api_ms": first_cull_host_ms,
        "shadow_views": int(len(ll)), "shadow_pairs": int(off[-1]) if len(off) else 0,
        "shadow_lists_device_ms": lists_ms, "shadow_lists_host_api_ms": lists_host_ms,
        "frame_ms_update_cull_lists": frame_ms,
        "tests_per_s_brute_force_equivalent": n_ent * n_views / (frame_ms * 1e-3),
        "update_kernels_only_ms": upd_kernels_ms,
        "roofline_update": {"bound": "hbm", "achieved": 132 * n_nodes / (upd_kernels_ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                            "frac": 132 * n_nodes / (upd_kernels_ms * 1e-3) / 1e9 / peak,
                            "formula": "132 B per dirty node (SURVEY.md 8(d)) over the level kernels of one update (one launch per tree level)",
                            "whole_update_frac": 132 * n_nodes / (upd_ms * 1e-3) / 1e9 / peak,
                            "whole_update_note": "update_ms also holds the 400 KB upload + scatter of the moved groups and the host gap "
                                                 "between the two API calls"},
        "parity": {"bit_exact_vs_oracle": bool(ok_trs and ok_aabb and ok_main and ok_lists),
                   "world_trs_exact": ok_trs, "world_aabb_exact": ok_aabb, "main_view_and_lights": ok_main,
                   "shadow_lists_exact_and_api_consistent": bool(ok_lists), "lists_checked": len(sample), "lists": int(len(ll)),
                   "oracle": oracle.name},
        "cpu_baseline": {"value": n_ent * n_views / cpu_frame_s, "unit": "tests/s", "cores": THREADS, "kind": oracle.name,
                         "sample": f"one whole frame: tree update (1 thread, {cpu_update_s:.2f} s) + main cull + {len(sample)} of {len(ll)} "
                                   f"shadow-view culls timed ({THREADS} threads), scaled to all {len(ll)}",
                         "frame_s": cpu_frame_s, "update_1thread_s": cpu_update_s, "main_cull_s": cpu_cull_s,
                         "per_shadow_view_cull_s": cpu_view_s / max(len(sample), 1)},
    }


def config1():
    """configs[0]: the apps/cubes demo scene - 4 render entities, 20 views per frame (camera + 1 spot + 3 point lights x 6).
    Times one frame's culls three ways: the reference's own SimpleRenderWorld (tier A, as DeferredRenderPipeline drives it:
    one cull() per view), this library called the same way (ovis_cull per view: 20 host round trips), and this library's
    frame call (ovis_cull for the camera + one ovis_build_shadow_lists). The GPU loses here and the numbers say so."""
    oracle = pyoracle.load("best")
    s = scenes.cubes_scene()
    n = len(s["local_aabb"])
    w = lib.World(0)
    e0 = w.add_entities(s["local_aabb"], s["flags"])
    w.set_transforms(np.arange(e0, e0 + n, dtype=np.uint32), s["trs"])
    w.add_lights(s["lights"])
    ents, ls = w.cull(s["view"])
    lists = w.shadow_lists(s["view"])
    views = [s["view"]]
    for k in range(len(lists[0])):
        views.append(lib.light_shadow_views(s["lights"][lists[0][k]])[lists[1][k]])
    n_views = len(views)

    def per_view():
        w.cull(s["view"])
        for v in views[1:]:
            w.cull(v, flags=0)

    # the frame as a C++ caller makes it: both calls with their arguments marshalled once (the convenience wrappers above
    # allocate their result arrays per call, which at this size costs more than the library does)
    pin_e, out_l = lib.PinnedBuffer(max(n, 1)), np.zeros(max(len(s["lights"]), 1), dtype=np.uint32)
    frame_cull = w.bind_cull(s["view"], pin_e.array, flags=lib.CULL_LIGHTS, out_lights=out_l)
    frame_lists = w.bind_shadow_lists(s["view"], cap_entities=n * 6 * max(len(s["lights"]), 1), pinned=True)

    def per_frame():
        frame_cull()
        return frame_lists()

    def wall(fn, reps=200):
        for _ in range(10):
            fn()
        t0 = time.perf_counter()
        for _ in range(reps):
            fn()
        return (time.perf_counter() - t0) / reps * 1e6

    per_view_us, per_frame_us = wall(per_view), wall(per_frame)
    out = {"workload": f"configs[0]: apps/cubes demo scene, {n} render entities x {n_views} views (1 camera + 1 spot + 3 point lights x 6 faces)",
           "entities": n, "views": n_views,
           "gpu_per_view_calls_us_per_frame": per_view_us, "gpu_frame_call_us_per_frame": per_frame_us,
           "tests_per_s_gpu_frame_call": n * n_views / (per_frame_us * 1e-6),
           "note": "launch- and round-trip-bound: 80 entity-view tests per frame; the CPU reference does them in a few microseconds"}
    ok = None
    if oracle.has_world:
        rw = oracle.world()
        rw.add_entities(s["trs"], s["local_aabb"], s["flags"])
        rw.add_lights(s["lights"])
        r_ents, r_ls = rw.cull(s["view"])
        r_lists = rw.shadow_lists(s["view"])
        ok = bool(np.array_equal(ents, r_ents) and np.array_equal(ls, r_ls) and all(np.array_equal(a, b) for a, b in zip(lists, r_lists)))

        def ref_frame():
            rw.cull(s["view"])
            rw.shadow_lists(s["view"])

        ref_us = wall(ref_frame, reps=2000)
        rw.close()
        out["cpu_baseline"] = {"value": n * n_views / (ref_us * 1e-6), "unit": "tests/s", "cores": 1, "kind": oracle.name,
                               "sample": "the whole frame: SimpleRenderWorld::cull for the camera + the prepareLights loop over every "
                                         "shadow view (tier A: the reference's own classes), through ctypes, 2000 frames",
                               "us_per_frame": ref_us}
        out["gpu_over_cpu_time_ratio_frame_call"] = per_frame_us / ref_us
    out["parity"] = {"bit_exact_vs_oracle": ok, "oracle": oracle.name if ok is not None else None}
    w.close()
    return out


def dense_visible(n=12_500_000):
    """The bench shard with the camera inside a dense cloud (config-2 distribution at 1/4 of the extent): ~5 % of 12.5 M
    entities visible, so the gather kernel and the host copy carry real traffic (the headline frustum sees 0.09 %)."""
    oracle = pyoracle.load("best")
    aabb, flags = scenes.static_boxes(n, 4000, 1000.0)
    view = scenes.default_camera()
    planes = lib.view_planes(view)
    stream = torch.cuda.Stream()
    w = lib.World(0)
    w.set_stream(stream.cuda_stream)
    w.reserve(n)
    f = w.add_entities(aabb, flags)
    w.set_world_aabbs(f, aabb)
    out = torch.zeros(n + 1, dtype=torch.int32, device="cuda")

    def fn():
        w.cull_planes_device(planes, out.data_ptr() + 4, out.data_ptr())

    for _ in range(5):
        fn()
    w.set_timing(True)
    for _ in range(30):
        fn()
    test_ms, gather_ms = w.read_timing()
    w.set_timing(False)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(stream)
    for _ in range(100):
        fn()
    b.record(stream)
    stream.synchronize()
    frame_ms = a.elapsed_time(b) / 100
    vis = int(out[0].item())
    got = out[1:1 + vis].cpu().numpy().view(np.uint32)
    t0 = time.perf_counter()
    exp = oracle.cull_aabb(aabb, planes, threads=THREADS)
    cpu_s = time.perf_counter() - t0
    pinned = lib.PinnedBuffer(max(2 * vis, 1024))
    frame = w.bind_cull(view, pinned.array, flags=0)
    for _ in range(3):
        frame()
    t0 = time.perf_counter()
    for _ in range(50):
        got_n = frame()[0]
    e2e_ms = (time.perf_counter() - t0) / 50 * 1e3
    ok = bool(np.array_equal(got, exp)) and got_n == vis and bool(np.array_equal(pinned.array[:vis], exp))
    w.close()
    peak = hbm_peak()
    alg = 24 * n + 4 * vis
    return {"workload": f"{n / 1e6:g}M entities (config-2 distribution in [-1000,1000]^3), 1 camera frustum: {100.0 * vis / n:.1f} % visible",
            "entities": n, "visible": vis, "test_kernel_ms": float(np.median(test_ms)), "gather_kernel_ms": float(np.median(gather_ms)),
            "cull_ms_per_frame": frame_ms, "tests_per_s": n / (frame_ms * 1e-3),
            "roofline": {"bound": "hbm", "achieved": alg / (frame_ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                         "frac": alg / (frame_ms * 1e-3) / 1e9 / peak, "formula": "24 B/entity + 4 B/visible id over the whole frame (both kernels)"},
            "e2e_host_api": {"ms_per_frame": e2e_ms, "tests_per_s": n / (e2e_ms * 1e-3), "d2h_bytes_per_step": 8 + 4 * vis},
            "parity": {"bit_exact_vs_oracle": ok, "oracle": oracle.name},
            "cpu_baseline": {"value": n / cpu_s, "unit": "tests/s", "cores": THREADS, "kind": oracle.name,
                             "sample": "one pass over the same 12.5 M entities, flat arrays", "ms_per_frame": cpu_s * 1e3}}


def stress_all_lights(small=False):
    """configs[2]'s world with ALL 4096 point lights casting shadows at once: 24 576 shadow views over 10 M entities in one
    call (the light grid holds 4096 cubes). Timed on the device; a sample of lists is checked against the oracle."""
    oracle = pyoracle.load("best")
    kw = dict(n_groups=1000, n_parts=900_000, n_lights=512) if small else {}
    s = scenes.config3(**kw)
    n_ent, n_nodes = s["n_entities"], s["n_nodes"]
    stream = torch.cuda.Stream()
    w = lib.World(0)
    w.set_stream(stream.cuda_stream)
    w.reserve(n_ent, n_nodes)
    e0 = w.add_entities(s["local_aabb"], s["flags"])
    w.add_nodes(s["parent"], s["local_trs"])
    w.bind_entities(s["entity_node"], np.arange(e0, e0 + n_ent, dtype=np.uint32))
    w.add_lights(s["lights"])
    w.update_transforms()
    ids = np.arange(len(s["lights"]), dtype=np.uint32)
    ms, _ = timed(stream, w.bind_cull_shadow_views_device(ids), iters=5, warm=1)
    res = w.cull_shadow_views_device(ids)
    ll, lv, off, _, d_ent = res
    ent_all = w.read_device(d_ent, int(off[-1]))
    wb = w.world_aabbs()
    rng = np.random.default_rng(5)
    sample = rng.choice(len(ll), 48, replace=False)
    ok = True
    for k in sample:
        v = oracle.light_shadow_views(s["lights"][ll[k]])[lv[k]]
        ok &= bool(np.array_equal(ent_all[off[k]:off[k + 1]], oracle.cull_aabb(wb, oracle.view_planes(v), s["flags"], 1, threads=THREADS)))
    w.close()
    return {"workload": f"{n_ent / 1e6:g}M entities x {len(ll)} shadow views ({len(s['lights'])} point lights, all of them)",
            "shadow_lists_ms": ms, "pairs": int(off[-1]), "tests_per_s_brute_force_equivalent": n_ent * len(ll) / (ms * 1e-3),
            "parity": {"sampled_lists_exact": bool(ok), "sampled": int(len(sample)), "oracle": oracle.name}}


def config4(small, full_parity=True, brute=True):
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
    cam_ms, _ = timed(stream, w.bind_cull_views_device(cam_planes), iters=9)
    coff, _, cd_ent = w.cull_views_device(cam_planes)
    cam_ent = w.read_device(cd_ent, int(coff[-1]))
    ok_cam = all(np.array_equal(cam_ent[coff[i]:coff[i + 1]], oracle.cull_aabb(s["aabb"], cam_planes[i], threads=THREADS))
                 for i in range(len(cam_views)))
    light_ids = np.arange(len(s["lights"]), dtype=np.uint32)
    lists_ms, _ = timed(stream, w.bind_cull_shadow_views_device(light_ids), iters=9)
    ll, lv, off, d_off, d_ent = w.cull_shadow_views_device(light_ids)
    ent_all = w.read_device(d_ent, int(off[-1]))
    sample = list(range(len(ll))) if full_parity else list(range(0, len(ll), 16))
    ok_lists = True
    cpu_view_s = 0.0
    for k in sample:
        v = oracle.light_shadow_views(s["lights"][ll[k]])[lv[k]]
        t0 = time.perf_counter()
        e = oracle.cull_aabb(s["aabb"], oracle.view_planes(v), s["flags"], 1, threads=THREADS)
        cpu_view_s += time.perf_counter() - t0
        ok_lists &= bool(np.array_equal(ent_all[off[k]:off[k + 1]], e))
    # brute-force comparison point on the GPU: the same 384 views without the pre-filter (the dense pass)
    brute_ms, ok_brute = None, None
    if brute:
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
    issue = issue_peak()
    cam_tests = n * len(cam_views)
    return {
        "workload": f"configs[3]: {n / 1e6:g}M static entities x (1 camera + 4 cascade views + {len(s['lights'])} point lights x 6 faces) = {n_views} views",
        "camera_and_cascades_ms": cam_ms, "camera_visible": [int(coff[i + 1] - coff[i]) for i in range(len(cam_views))],
        "light_faces_ms": lists_ms, "light_face_pairs": int(off[-1]),
        "frame_ms": cam_ms + lists_ms,
        "tests_per_s_brute_force_equivalent": n * n_views / ((cam_ms + lists_ms) * 1e-3),
        "light_faces_without_prefilter_ms": brute_ms,
        "roofline_camera_and_cascades": None if not issue else {
            "bound": "fp32 issue rate without FMA", "unit": "plane tests/s",
            "achieved": 6 * cam_tests / (cam_ms * 1e-3), "peak": issue["cull_mix_plane_tests_per_s"],
            "frac": 6 * cam_tests / (cam_ms * 1e-3) / issue["cull_mix_plane_tests_per_s"],
            "peak_source": "profiles/r02_issue_peak.json (tools/issue_probe.cu: the inner loop's instruction mix, nothing else)",
            "note": "6 plane tests per entity-view counted (no early-out credit); includes the scan and expand kernels. The five "
                    "views share their four side planes, which the kernel evaluates once: it executes 14 plane tests per entity "
                    "where 30 are counted here, so this fraction says how far the pass is from a kernel that shares nothing",
            "executed_plane_tests_frac": (14 / 30) * 6 * cam_tests / (cam_ms * 1e-3) / issue["cull_mix_plane_tests_per_s"]},
        "parity": {"bit_exact_vs_oracle": bool(ok_cam and ok_lists and (ok_brute is not False)),
                   "camera_views": bool(ok_cam), "light_faces_exact": bool(ok_lists), "lists_checked": len(sample), "lists": int(len(ll)),
                   "prefilter_equals_brute_force_all_384": ok_brute, "oracle": oracle.name},
        "cpu_baseline": {"value": n * n_views / (cpu_view_s / len(sample) * n_views), "unit": "tests/s", "cores": THREADS,
                         "kind": oracle.name,
                         "sample": f"{len(sample)} of the {len(ll)} light-face culls timed ({THREADS} threads, flat arrays), scaled to all {n_views} views",
                         "per_view_cull_s": cpu_view_s / len(sample), "frame_s": cpu_view_s / len(sample) * n_views},
    }


if __name__ == "__main__":
    small = "--small" in sys.argv
    which = [a for a in sys.argv[1:] if a in ("1", "3", "4", "dense", "stress")] or ["1", "3", "4", "dense", "stress"]
    out = {}
    if "1" in which:
        out["config1"] = config1()
    if "dense" in which:
        out["dense_visible"] = dense_visible(1_000_000 if small else 12_500_000)
    if "stress" in which:
        out["all_4096_lights"] = stress_all_lights(small)
    if "3" in which:
        out["config3"] = config3(small)
    if "4" in which:
        out["config4"] = config4(small)
    print(json.dumps(out, indent=1))
