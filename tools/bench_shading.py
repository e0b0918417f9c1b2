#!/usr/bin/env python
"""Timing + parity of the deferred light pass and the colour post effects on one GPU.
Evidence for DESIGN.md, not the bench.py line.

  1920 x 1080 headless G-buffer (fp32), L lights of all four types, with and without 512^2 shadow maps; then tonemap +
  gamma fused. Timed with CUDA events on the world's stream (median). The CPU oracle (oracle/ovis_shading.c, one core)
  shades a 320 x 180 crop of the same buffers beside it, and the GPU result on that crop is checked against it.
    python tools/bench_shading.py [--small]
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


def timed(stream, fn, iters=10, warm=2):
    for _ in range(warm):
        fn()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for a, b in evs:
        a.record(stream)
        fn()
        b.record(stream)
    stream.synchronize()
    return float(np.median([a.elapsed_time(b) for a, b in evs]))


def rel_err(got, want):
    return float(np.nanmax(np.abs(got.astype(np.float64) - want) / np.maximum(np.abs(want), 1e-3)))


def main():
    small = "--small" in sys.argv
    w, h = (480, 270) if small else (1920, 1080)
    cw, ch = (96, 54) if small else (320, 180)
    size = 128 if small else 512
    oracle = pyoracle.load("best")
    rng = np.random.default_rng(99)
    view = scenes.make_view((0, 2, 10), scenes.quat_yaw_pitch(10.0, -5.0), 75.0, 0.1, 500.0, w, h)
    gbuf = util.random_gbuffer(rng, w, h)
    stream = torch.cuda.Stream()
    world = lib.World(0)
    world.set_stream(stream.cuda_stream)
    out = {"gpu": torch.cuda.get_device_name(0), "oracle": oracle.name, "pixels": w * h, "shadow_map_size": size, "cases": []}
    with torch.cuda.stream(stream):
        d = [torch.from_numpy(x).cuda() for x in gbuf]
        d_out = torch.zeros((h * w, 4), dtype=torch.float32, device="cuda")
        d_post = torch.zeros_like(d_out)
        vb = oracle.view_uniforms(view)
        d_vb = torch.from_numpy(vb.copy()).cuda()
    n_max = 256
    lights, att, bias = util.lights_for_shading(rng, n_max, view)
    world.add_lights(lights)
    for i in range(n_max):
        world.set_light_shading(i, att[i], float(bias[i]))
    blocks = np.stack([oracle.light_uniforms(l, a, float(b)) for l, a, b in zip(lights, att, bias)])
    maps_host = {}
    ptrs = np.zeros(n_max, dtype=np.uint64)
    keep = []
    for i, l in enumerate(lights[:64]):                         # shadow maps for the shadow casters among the first 64
        if l["flags"] & 1 and l["type"] in (scenes.LIGHT_SPOT, scenes.LIGHT_POINT):
            shape = (size, size) if l["type"] == scenes.LIGHT_SPOT else (6, size, size)
            maps_host[i] = rng.uniform(0.9, 1.0, shape).astype(np.float32)
            t = torch.from_numpy(maps_host[i]).cuda()
            keep.append(t)
            ptrs[i] = t.data_ptr()
    d_ptrs = torch.from_numpy(ptrs.view(np.int64)).cuda()
    # a crop for the CPU side: the shader uses the pixel coordinate, so the crop is the top-left corner rows/columns
    crop = [np.ascontiguousarray(x[:ch, :cw]) for x in gbuf]

    for n_lights, shadows in ((5, False), (64, False), (64, True), (256, False)):
        ids = torch.arange(n_lights, dtype=torch.int32, device="cuda")
        cnt = torch.tensor([n_lights], dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()

        def shade():
            world.shade_deferred_device(w, h, d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(), d[3].data_ptr(), d_vb.data_ptr(),
                                        d_out.data_ptr(), ids.data_ptr(), cnt.data_ptr(), n_lights,
                                        d_ptrs.data_ptr() if shadows else 0, size if shadows else 0)

        ms = timed(stream, shade)
        got = d_out.view(h, w, 4)[:ch, :cw].cpu().numpy()
        maps = [maps_host.get(i) for i in range(n_lights)] if shadows else None
        # the crop keeps the full-frame viewport in the view block, so pixel (x, y) reconstructs the same position
        t0 = time.perf_counter()
        want = oracle.shade_deferred(*crop, vb, blocks[:n_lights], lights["type"][:n_lights], maps, size)
        cpu_ms = (time.perf_counter() - t0) * 1e3
        case = {"lights": n_lights, "shadow_maps": int(sum(m is not None for m in maps)) if maps else 0, "gpu_ms": ms,
                "ns_per_pixel_light": ms * 1e6 / (w * h * n_lights), "pixel_lights_per_s": w * h * n_lights / (ms * 1e-3),
                "gbuffer_gbs": 68.0 * w * h / (ms * 1e-3) / 1e9,
                "cpu_1core_ms_scaled_to_frame": cpu_ms * (w * h) / (cw * ch), "cpu_sample": f"{cw}x{ch} crop",
                "max_rel_err_vs_oracle": rel_err(got, want), "nan_pixels": int(np.isnan(want[..., 0]).sum())}
        case["speedup_vs_1core"] = case["cpu_1core_ms_scaled_to_frame"] / ms
        out["cases"].append(case)

    # the same frame size with lights that reach only part of the scene: view distances 1..300, 256 lights of range 5..40
    gbuf2 = util.random_gbuffer(rng, w, h, linear_z=(0.1, 500.0, 1.0, 300.0))
    lights2, att2, bias2 = util.lights_for_shading(rng, n_max, view, extent=150.0)
    lights2["range"] = rng.uniform(5.0, 40.0, n_max)
    world2 = lib.World(0)
    world2.set_stream(stream.cuda_stream)
    world2.add_lights(lights2)
    for i in range(n_max):
        world2.set_light_shading(i, att2[i], float(bias2[i]))
    blocks2 = np.stack([oracle.light_uniforms(l, a, float(b)) for l, a, b in zip(lights2, att2, bias2)])
    with torch.cuda.stream(stream):
        d2 = [torch.from_numpy(x).cuda() for x in gbuf2]
    ids = torch.arange(n_max, dtype=torch.int32, device="cuda")
    cnt = torch.tensor([n_max], dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()

    def shade_sparse():
        world2.shade_deferred_device(w, h, d2[0].data_ptr(), d2[1].data_ptr(), d2[2].data_ptr(), d2[3].data_ptr(), d_vb.data_ptr(),
                                     d_out.data_ptr(), ids.data_ptr(), cnt.data_ptr(), n_max)

    ms = timed(stream, shade_sparse)
    crop2 = [np.ascontiguousarray(x[:ch, :cw]) for x in gbuf2]
    t0 = time.perf_counter()
    want = oracle.shade_deferred(*crop2, vb, blocks2, lights2["type"])
    cpu_ms = (time.perf_counter() - t0) * 1e3
    got = d_out.view(h, w, 4)[:ch, :cw].cpu().numpy()
    out["sparse_lights"] = {"lights": n_max, "light_range": "5..40", "view_distance": "1..300", "gpu_ms": ms,
                            "ns_per_pixel_light": ms * 1e6 / (w * h * n_max), "cpu_1core_ms_scaled_to_frame": cpu_ms * (w * h) / (cw * ch),
                            "max_rel_err_vs_oracle": rel_err(got, want),
                            "note": "pixel-light pairs beyond the light's range are skipped where that is provably exact"}
    world2.close()

    # a spatially coherent frame: the ground plane in perspective, 256 point / spot lights of range 5..25 standing on it
    view3 = scenes.make_view((0, 6, 40), scenes.quat_yaw_pitch(0.0, -12.0), 75.0, 0.1, 500.0, w, h)
    gbuf3 = util.ground_gbuffer(rng, oracle, view3, w, h)
    lights3, att3, bias3 = util.lights_for_shading(rng, n_max, view3)
    lights3["type"] = np.where(rng.random(n_max) < 0.5, scenes.LIGHT_POINT, scenes.LIGHT_SPOT)
    lights3["position"][:, 0] = rng.uniform(-120, 120, n_max)
    lights3["position"][:, 1] = rng.uniform(0.5, 5.0, n_max)
    lights3["position"][:, 2] = rng.uniform(-250, 35, n_max)
    lights3["direction"][:, 1] = -np.abs(lights3["direction"][:, 1]) - 0.5
    lights3["range"] = rng.uniform(5.0, 25.0, n_max)
    world3 = lib.World(0)
    world3.set_stream(stream.cuda_stream)
    world3.add_lights(lights3)
    for i in range(n_max):
        world3.set_light_shading(i, att3[i], float(bias3[i]))
    blocks3 = np.stack([oracle.light_uniforms(l, a, float(b)) for l, a, b in zip(lights3, att3, bias3)])
    vb3 = oracle.view_uniforms(view3)
    with torch.cuda.stream(stream):
        d3 = [torch.from_numpy(x).cuda() for x in gbuf3]
        d_vb3 = torch.from_numpy(vb3.copy()).cuda()
    torch.cuda.synchronize()

    def shade_ground():
        world3.shade_deferred_device(w, h, d3[0].data_ptr(), d3[1].data_ptr(), d3[2].data_ptr(), d3[3].data_ptr(), d_vb3.data_ptr(),
                                     d_out.data_ptr(), ids.data_ptr(), cnt.data_ptr(), n_max)

    ms = timed(stream, shade_ground)
    y0 = h // 2                                                          # a crop from the middle rows, where the lights are
    crop3 = [np.ascontiguousarray(x[y0:y0 + ch, :cw]) for x in gbuf3]
    got = d_out.view(h, w, 4)[y0:y0 + ch, :cw].cpu().numpy()
    # the oracle derives positions from pixel coordinates: shade the crop as a frame whose viewport starts at (0, -y0)
    vb3_crop = oracle.view_uniforms(view3, (0, -y0))
    t0 = time.perf_counter()
    want = oracle.shade_deferred(*crop3, vb3_crop, blocks3, lights3["type"])
    cpu_ms = (time.perf_counter() - t0) * 1e3
    out["ground_plane_256_lights"] = {"lights": n_max, "light_range": "5..25", "gpu_ms": ms,
                                      "ns_per_pixel_light": ms * 1e6 / (w * h * n_max),
                                      "cpu_1core_ms_scaled_to_frame": cpu_ms * (w * h) / (cw * ch),
                                      "lit_fraction_of_crop": float((want[..., :3] > 0).any(axis=2).mean()),
                                      "max_rel_err_vs_oracle": rel_err(got, want),
                                      "note": "16x16-pixel tiles reject lights whose range ends before the tile's position bounds"}
    world3.close()

    def post():
        world.post_effects_device(w * h, d_out.data_ptr(), d_post.data_ptr(), lib.POST_TONEMAP | lib.POST_GAMMA, 1.5, 11.2, 2.2)

    ms = timed(stream, post)
    hdr = d_out[:200_000].cpu().numpy()
    t0 = time.perf_counter()
    want = oracle.gamma(oracle.tonemap(hdr, 1.5, 11.2), 2.2)
    cpu_ms = (time.perf_counter() - t0) * 1e3
    out["post_effects"] = {"gpu_ms": ms, "gbs": 32.0 * w * h / (ms * 1e-3) / 1e9, "cpu_1core_ms_scaled_to_frame": cpu_ms * w * h / 200_000,
                           "max_rel_err_vs_oracle": rel_err(d_post[:200_000].cpu().numpy(), want)}
    world.close()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
