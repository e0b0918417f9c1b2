"""Randomised differential check of the deferred light pass (ovis_shade_deferred_device) against its CPU restatement
(oracle/ovis_shading.c, itself held bit-equal to the reference's shader source): random cameras, incoherent and coherent
G-buffers, lights of every type with ranges from centimetres to the whole scene, shadow maps, and unclean values sprinkled
over pixels and lights (NaN, Inf, huge, negative) - the kernel's per-tile and per-pixel range rejection must never change
a value. Tolerance: that of tests/test_shading_gpu.py (CUDA's powf), NaNs in the same places.

    python tools/fuzz_shading.py [--cases 60] [--seed 1] [--only SEED]"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from orion_b200 import lib, scenes  # noqa: E402
import vis_testutil as util  # noqa: E402
import test_shading_gpu as tsg  # noqa: E402


def run_case(oracle, seed):
    import torch
    rng = np.random.default_rng(seed)
    w, h = [(1, 1), (16, 16), (17, 5), (33, 47), (160, 90), (250, 131)][int(rng.integers(0, 6))]
    n = int(rng.choice([1, 3, 24, 97, 260]))
    coherent = bool(rng.random() < 0.5)
    if coherent:
        view = scenes.make_view((float(rng.uniform(-20, 20)), float(rng.uniform(1, 15)), float(rng.uniform(0, 60))),
                                scenes.quat_yaw_pitch(float(rng.uniform(-40, 40)), float(rng.uniform(-35, -3))), float(rng.uniform(40, 110)),
                                0.1, float(rng.choice([100.0, 500.0, 2000.0])), w, h)
        gbuf = util.ground_gbuffer(rng, oracle, view, w, h)
    else:
        view = util.random_views(rng, 1, extent=30.0)[0]
        view["vp_width"], view["vp_height"] = w, h
        zn, zf = float(view["znear"]), float(view["zfar"])
        gbuf = util.random_gbuffer(rng, w, h, sky_fraction=float(rng.choice([0.0, 0.1, 0.9])),
                                   linear_z=(zn, zf, zn * 1.01, zn + (zf - zn) * float(rng.choice([0.05, 0.5, 0.99]))))
    extent = float(rng.choice([10.0, 60.0, 300.0]))
    lights, att, bias = util.lights_for_shading(rng, n, view, extent=extent)
    range_kind = str(rng.choice(["short", "mixed", "long"]))
    lo, hi = {"short": (0.05, 8.0), "mixed": (0.5, 300.0), "long": (100.0, 3000.0)}[range_kind]
    lights["range"] = np.exp(rng.uniform(np.log(lo), np.log(hi), n)).astype(np.float32)
    if coherent:
        lights["position"][:, 1] = rng.uniform(0.2, 6.0, n)
        lights["direction"][:, 1] = -np.abs(lights["direction"][:, 1]) - 0.3
    unclean = bool(rng.random() < 0.5)
    if unclean:
        a, b, c, depth = gbuf
        for _ in range(int(rng.integers(1, 8))):
            y, x = int(rng.integers(0, h)), int(rng.integers(0, w))
            k = rng.integers(0, 7)
            bad = float(rng.choice([np.nan, np.inf, -np.inf, 1e30, -3.0, 1e9]))
            if k == 0: a[y, x, int(rng.integers(0, 3))] = bad
            elif k == 1: b[y, x, int(rng.integers(0, 3))] = bad
            elif k == 2: c[y, x, int(rng.integers(0, 4))] = bad
            elif k == 3: depth[y, x] = float(rng.choice([np.nan, -0.5, 1.5, 0.0, 1.0]))
            elif k == 4: a[y, x, :3] = (3.0, 3.0, 3.0)
            elif k == 5: c[y, x, 3] = float(rng.choice([0.0, -2.0, 1e9]))
            else: b[y, x, :3] = 0.0
        for _ in range(int(rng.integers(1, 6))):
            i = int(rng.integers(0, n))
            k = rng.integers(0, 8)
            if k == 0: att[i] = (0.0, 0.0, 0.0)
            elif k == 1: att[i] = (1.0, float(rng.uniform(-1, 0)), 0.0)
            elif k == 2: att[i] = (float(rng.choice([np.nan, -1.0, 1e30])), 0.1, 0.01)
            elif k == 3: lights["intensity"][i] = float(rng.choice([np.nan, np.inf, 1e30, -1.0]))
            elif k == 4: lights["colour"][i] = (float(rng.choice([np.inf, np.nan, -1.0])), 1.0, 1.0)
            elif k == 5: lights["position"][i] = (float(rng.choice([1e30, np.nan, np.inf])), 0.0, 0.0)
            elif k == 6: lights["range"][i] = float(rng.choice([-5.0, 0.0, np.nan, np.inf]))
            else:
                lights["cutoff"][i] = float(rng.choice([0.0, 45.0, 1e-3]))
                lights["type"][i] = scenes.LIGHT_SPOT
    desc = dict(size=[w, h], lights=n, coherent=coherent, ranges=range_kind, unclean=unclean)
    world = lib.World(0)
    world.add_lights(lights)
    for i in range(n):
        world.set_light_shading(i, att[i], float(bias[i]))
    blocks = np.stack([oracle.light_uniforms(l, x, float(y)) for l, x, y in zip(lights, att, bias)])
    vb = oracle.view_uniforms(view, (int(rng.integers(0, 5)), int(rng.integers(0, 5))))
    ids = rng.permutation(n).astype(np.uint32)[:max(1, int(rng.integers(1, n + 1)))]
    size = int(rng.choice([8, 32]))
    with_maps = bool(rng.random() < 0.5)
    maps = None
    if with_maps:
        maps = [rng.uniform(float(rng.choice([0.0, 0.9])), 1.0, (size, size) if l["type"] == scenes.LIGHT_SPOT else (6, size, size)).astype(np.float32)
                if (l["flags"] & 1 and l["type"] >= scenes.LIGHT_POINT) else None for l in lights]
    desc["maps"] = with_maps
    sel_maps = [maps[i] for i in ids] if maps is not None else None
    all_maps_by_light = maps
    want = oracle.shade_deferred(*gbuf, vb, blocks[ids], lights["type"][ids], sel_maps, size if with_maps else 0)
    got = tsg._shade(torch, world, w, h, gbuf, vb, ids, all_maps_by_light, size if with_maps else 0)
    world.close()
    if not np.array_equal(np.isnan(got), np.isnan(want)):
        return desc, f"NaNs in different places ({int(np.isnan(got).sum())} vs {int(np.isnan(want).sum())})"
    try:
        tsg._close(got, want)
    except AssertionError as e:
        return desc, str(e)
    desc["lit"] = round(float((np.nan_to_num(want[..., :3]) > 0).any(axis=2).mean()), 3)
    return desc, None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", type=int, default=60)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--only", type=int, default=None)
    ap.add_argument("--seconds", type=float, default=600.0)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    from oracle import pyoracle
    oracle = pyoracle.load("best")
    seeds = [a.only] if a.only is not None else [a.seed * 10_000 + i for i in range(a.cases)]
    t0 = time.time()
    done, pixels, failures = 0, 0, []
    import warnings
    warnings.simplefilter("ignore")
    for s in seeds:
        desc, err = run_case(oracle, s)
        done += 1
        print(s, json.dumps(desc), "FAIL: " + err if err else "ok", flush=True)
        if err:
            failures.append({"seed": s, "case": desc, "error": err})
            if len(failures) >= 3:
                break
        else:
            pixels += desc["size"][0] * desc["size"][1]
        if time.time() - t0 > a.seconds:
            break
    summary = {"cases": done, "pixels_compared": pixels, "failures": failures, "oracle": oracle.name, "seconds": round(time.time() - t0, 1)}
    print(json.dumps(summary))
    if a.out:
        with open(a.out, "w") as f:
            json.dump(summary, f, indent=1)
    sys.exit(1 if failures else 0)


if __name__ == "__main__":
    main()
