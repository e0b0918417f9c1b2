"""Randomised check that the C restatement of the light pass (oracle/ovis_shading.c - the checker of the CUDA light
pass) is bit-equal to the reference's own shader source compiled as C++ against GLM (oracle/glsl_harness.cc:
lighting.h, deferred_light_frag.glsl, the post effects), on random cameras, G-buffers, lights of every type, shadow maps
and unclean values (NaN / Inf / huge / negative). NaNs must sit in the same places. Runs on the CPU.

    python tools/fuzz_glsl.py [--cases 300] [--seed 1]"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from orion_b200 import scenes  # noqa: E402
import vis_testutil as util  # noqa: E402


def same_bits(a, b):
    a, b = np.asarray(a, dtype=np.float32), np.asarray(b, dtype=np.float32)
    both_nan = np.isnan(a) & np.isnan(b)
    return a.shape == b.shape and bool(((a.view(np.uint32) == b.view(np.uint32)) | both_nan).all())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", type=int, default=300)
    ap.add_argument("--seed", type=int, default=1)
    a = ap.parse_args()
    import warnings
    warnings.simplefilter("ignore")
    from oracle import pyoracle
    ref = pyoracle.load("reference")
    rng = np.random.default_rng(a.seed)
    pixels = 0
    for case in range(a.cases):
        w, h = [(1, 1), (5, 3), (16, 16), (33, 21)][int(rng.integers(0, 4))]
        n = int(rng.choice([1, 4, 25]))
        if rng.random() < 0.5:
            view = scenes.make_view((float(rng.uniform(-20, 20)), float(rng.uniform(1, 15)), float(rng.uniform(0, 60))),
                                    scenes.quat_yaw_pitch(float(rng.uniform(-40, 40)), float(rng.uniform(-35, -3))), float(rng.uniform(40, 110)),
                                    0.1, float(rng.choice([100.0, 2000.0])), w, h)
            gbuf = util.ground_gbuffer(rng, ref, view, w, h)
        else:
            view = util.random_views(rng, 1, extent=30.0)[0]
            view["vp_width"], view["vp_height"] = w, h
            zn, zf = float(view["znear"]), float(view["zfar"])
            gbuf = util.random_gbuffer(rng, w, h, sky_fraction=float(rng.choice([0.0, 0.2])), linear_z=(zn, zf, zn * 1.01, zn + (zf - zn) * 0.5))
        lights, att, bias = util.lights_for_shading(rng, n, view, extent=float(rng.choice([10.0, 100.0])))
        lights["range"] = np.exp(rng.uniform(np.log(0.05), np.log(2000.0), n)).astype(np.float32)
        if rng.random() < 0.5:
            ab, bb, cb, depth = gbuf
            for _ in range(int(rng.integers(1, 6))):
                y, x = int(rng.integers(0, h)), int(rng.integers(0, w))
                bad = float(rng.choice([np.nan, np.inf, -np.inf, 1e30, -3.0, 1e9, 0.0]))
                [ab, bb, cb][int(rng.integers(0, 3))][y, x, int(rng.integers(0, 3))] = bad
                if rng.random() < 0.3:
                    depth[y, x] = float(rng.choice([np.nan, -0.5, 1.5, 0.0, 1.0]))
                if rng.random() < 0.3:
                    cb[y, x, 3] = float(rng.choice([0.0, -2.0, 1e9]))
            for _ in range(int(rng.integers(1, 4))):
                i = int(rng.integers(0, n))
                k = int(rng.integers(0, 5))
                if k == 0: att[i] = (0.0, 0.0, 0.0)
                elif k == 1: att[i] = (float(rng.choice([np.nan, -1.0, 1e30])), 0.1, 0.01)
                elif k == 2: lights["intensity"][i] = float(rng.choice([np.nan, np.inf, 1e30, -1.0]))
                elif k == 3: lights["position"][i] = (float(rng.choice([1e30, np.nan, np.inf])), 0.0, 0.0)
                else: lights["range"][i] = float(rng.choice([-5.0, 0.0, np.nan, np.inf]))
        blocks = np.stack([ref.light_uniforms(l, x, float(y)) for l, x, y in zip(lights, att, bias)])
        vb = ref.view_uniforms(view, (int(rng.integers(0, 4)), int(rng.integers(0, 4))))
        size = int(rng.choice([4, 16]))
        maps = None
        if rng.random() < 0.6:
            maps = [rng.uniform(0.0, 1.0, (size, size) if l["type"] == scenes.LIGHT_SPOT else (6, size, size)).astype(np.float32)
                    if (l["flags"] & 1 and l["type"] >= scenes.LIGHT_POINT) else None for l in lights]
        c_out = ref.shade_deferred(*gbuf, vb, blocks, lights["type"], maps, size if maps else 0)
        g_out = ref.shade_deferred(*gbuf, vb, blocks, lights["type"], maps, size if maps else 0, glsl=True)
        if not same_bits(c_out, g_out):
            bad = np.argwhere(~((c_out.view(np.uint32) == g_out.view(np.uint32)) | (np.isnan(c_out) & np.isnan(g_out))))[0]
            print(json.dumps({"case": case, "what": "light pass", "at": bad.tolist(), "c": float(c_out[tuple(bad)]), "glsl": float(g_out[tuple(bad)])}))
            sys.exit(1)
        hdr = np.nan_to_num(c_out.reshape(-1, 4), nan=0.5, posinf=100.0, neginf=0.0)
        exposure, white = float(rng.uniform(0.2, 4.0)), float(rng.uniform(1.0, 20.0))
        if not same_bits(ref.tonemap(hdr, exposure, white), ref.tonemap(hdr, exposure, white, glsl=True)):
            print(json.dumps({"case": case, "what": "tonemap"}))
            sys.exit(1)
        ldr = np.abs(ref.tonemap(hdr, exposure, white))
        if not same_bits(ref.gamma(ldr, 2.2), ref.gamma(ldr, 2.2, glsl=True)):
            print(json.dumps({"case": case, "what": "gamma"}))
            sys.exit(1)
        pixels += w * h
    # the rasteriser's shader stages: lit_vtx / shadow_vtx / unlit, lit_deferred_frag in all eight keyword combinations
    port = pyoracle.load("port")
    stages = 0
    for case in range(a.cases * 4):
        trs = util.random_trs(rng, 1, pos=float(rng.choice([1.0, 50.0, 1e5])))
        if rng.random() < 0.2:
            trs[0, rng.integers(0, 10)] = float(rng.choice([np.nan, np.inf, 0.0, 1e30, -1e30]))
        transform = ref.entity_uniforms(trs)[0, :16]
        view = scenes.make_view(rng.uniform(-30, 30, 3), scenes.random_unit_quats(rng, 1)[0], float(rng.uniform(10, 150)),
                                float(rng.uniform(0.01, 2.0)), float(rng.uniform(3.0, 3000.0)), int(rng.choice([1, 17, 640])), int(rng.choice([1, 9, 360])))
        vp = ref.view_matrices(view)[0][2]
        pos = (rng.uniform(-3, 3, 3) * float(rng.choice([1.0, 1e-6, 1e6]))).astype(np.float32)
        nrm = (rng.standard_normal(3) * float(rng.choice([1.0, 0.0, 1e-30, 1e20]))).astype(np.float32)
        tan4 = np.concatenate([rng.standard_normal(3) * float(rng.choice([1.0, 0.0, 1e-25])), [rng.choice([-1.0, 1.0])]]).astype(np.float32)
        g_clip, g_n = ref.raster_vertex(transform, vp, pos, nrm, glsl=True)
        s_clip, _ = ref.raster_vertex(transform, vp, pos, nrm, glsl=True, shadow=True)
        u_clip, _ = ref.glsl_unlit(transform, vp, pos, np.ones(3, np.float32))
        gn, gt, gb = ref.glsl_vertex_normalmap(transform, vp, pos, nrm, tan4)
        textures = [(rng.integers(0, 256, (int(rng.choice([1, 3, 8])), int(rng.choice([1, 5, 16])), 4), dtype=np.uint8), bool(rng.random() < 0.5),
                     bool(rng.random() < 0.5)) for _ in range(2)]
        uv = (rng.uniform(-3, 4, 2) * float(rng.choice([1.0, 100.0, 1e-6]))).astype(np.float32)
        for o in (port, ref):
            clip, n = o.raster_vertex(transform, vp, pos, nrm)
            t = o.raster_vertex(transform, vp, pos, tan4[:3])[1]
            ok = same_bits(clip, g_clip) and same_bits(n, g_n) and same_bits(clip, s_clip) and same_bits(clip, u_clip) and \
                same_bits(n, gn) and same_bits(t, gt) and same_bits(o.raster_bitangent(n, t, float(tan4[3])), gb)
            for flags in range(8):
                mat8 = np.concatenate([rng.uniform(0, 1, 6), [rng.uniform(1, 128), flags & 1]]).astype(np.float32)
                mat10 = np.concatenate([mat8[:7], [flags, 0, 1]]).astype(np.float32)
                ok = ok and all(same_bits(x, y) for x, y in zip(o.raster_fragment(gn, mat8), ref.raster_fragment(gn, mat8, glsl=True)))
                ok = ok and all(same_bits(x, y) for x, y in zip(o.raster_fragment_textured(gn, uv, gt, gb, mat10, textures),
                                                               ref.raster_fragment_textured(gn, uv, gt, gb, mat10, textures, glsl=True)))
            if not ok:
                print(json.dumps({"case": case, "what": "rasteriser shader stages", "oracle": o.name}))
                sys.exit(1)
        stages += 1
    print(json.dumps({"cases": a.cases, "pixels": pixels, "shader_stage_cases": stages, "differences": 0}))


if __name__ == "__main__":
    main()
