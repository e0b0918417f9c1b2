"""Randomised differential check of the library's host-side math (what stays on the CPU: view matrices, frustum planes and
corners, a spot light's cone box, shadow views, the ViewUniforms / LightUniforms blocks) against the reference's own
compiled classes (oracle/_ref: RenderView, Frustum, RenderLight driven for real). NaNs must sit in the same places; every
other value must be the same bits. Runs on the CPU (no device work).

    python tools/fuzz_hostmath.py [--cases 20000] [--seed 1]"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from orion_b200 import lib, scenes  # noqa: E402


def same(a, b):
    a, b = np.asarray(a), np.asarray(b)
    return a.shape == b.shape and bool(np.array_equal(a, b, equal_nan=True))


def draw_float(rng, scale):
    k = rng.integers(0, 12)
    if k == 0:
        return 0.0
    if k == 1:
        return float(rng.choice([-0.0, 1.0, -1.0, 1e-30, -1e-30, 1e30, -1e30, 3.4e38, np.inf, -np.inf, np.nan]))
    if k == 2:
        return float(np.float32(rng.normal()) * np.float32(10.0) ** rng.integers(-20, 20))
    return float(rng.uniform(-scale, scale))


def draw_view(rng):
    v = np.zeros((), dtype=scenes.VIEW_DTYPE)
    v["position"] = [draw_float(rng, 1000.0) for _ in range(3)]
    q = rng.normal(size=4)
    kind = rng.integers(0, 6)
    if kind == 0:
        q = np.array([1.0, 0, 0, 0])
    elif kind == 1:
        q = q                                                      # not normalised
    elif kind == 2:
        q = np.array([draw_float(rng, 2.0) for _ in range(4)])
    else:
        q = q / np.linalg.norm(q)
    v["orientation"] = q.astype(np.float32)
    v["hfov_deg"] = float(rng.choice([rng.uniform(1, 179), rng.uniform(1, 179), 90.0, 0.0, 180.0, 360.0, -45.0, draw_float(rng, 400.0)]))
    v["znear"] = float(rng.choice([rng.uniform(0.001, 10), 0.1, 0.0, -1.0, draw_float(rng, 100.0)]))
    v["zfar"] = float(rng.choice([rng.uniform(0.01, 5000), 1000.0, float(v["znear"]), 0.0, draw_float(rng, 1e5)]))
    v["vp_width"] = int(rng.choice([1, 2, 512, 1440, 1920, 7, 100000, 0]))
    v["vp_height"] = int(rng.choice([1, 2, 512, 900, 1080, 3, 100000, 0]))
    return v


def draw_light(rng):
    l = np.zeros((), dtype=scenes.LIGHT_DTYPE)
    l["type"] = int(rng.integers(0, 4))
    l["flags"] = int(rng.integers(0, 2))
    l["position"] = [draw_float(rng, 500.0) for _ in range(3)]
    d = rng.normal(size=3)
    k = rng.integers(0, 6)
    if k == 0:
        d = np.array([[0, 0, -1], [0, 0, 1], [0, 1, 0], [0, -1, 0], [1, 0, 0], [-1, 0, 0]][int(rng.integers(0, 6))], dtype=np.float64)
    elif k == 1:
        d = np.array([draw_float(rng, 3.0) for _ in range(3)])
    elif k == 2:
        d = np.array([0.0, 0.0, -1.0]) + rng.normal(size=3) * 1e-4   # almost the rotation's special case
    elif k == 3:
        d = np.array([0.0, 0.0, 1.0]) + rng.normal(size=3) * 1e-4
    l["direction"] = d.astype(np.float32)
    l["colour"] = [float(rng.choice([0.0, 1.0, rng.uniform(0, 1)])) for _ in range(3)]
    l["intensity"] = float(rng.choice([0.0, 1.0, rng.uniform(0, 3), draw_float(rng, 5.0)]))
    l["range"] = float(rng.choice([rng.uniform(0.01, 500), 0.1, 0.05, 0.0, draw_float(rng, 1000.0)]))
    l["cutoff"] = float(rng.choice([rng.uniform(0.5, 45), 45.0, 0.0, 1e-3, -draw_float(rng, 200.0) ** 2]))
    if l["cutoff"] > 45.0 or np.isnan(l["cutoff"]):               # the reference asserts cutoff <= 45 (render_light.cc:104); so does the library
        l["cutoff"] = 45.0
    return l


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", type=int, default=20000)
    ap.add_argument("--seed", type=int, default=1)
    a = ap.parse_args()
    from oracle import pyoracle
    o = pyoracle.load("reference")
    rng = np.random.default_rng(a.seed)
    import warnings
    warnings.simplefilter("ignore")
    checks = 0
    for case in range(a.cases):
        v = draw_view(rng)
        l = draw_light(rng)
        att = tuple(float(x) for x in rng.uniform(0, 1, 3))
        bias = float(rng.uniform(-0.01, 0.01))
        pos = (int(rng.integers(-10, 2000)), int(rng.integers(-10, 2000)))
        what = None
        if v["vp_width"] > 0 and v["vp_height"] > 0:               # (a zero viewport is refused by the library's checks)
            if not same(lib.view_planes(v), o.view_planes(v)):
                what = "view_planes"
            m, c = lib.view_matrices(v)
            rm, rc = o.view_matrices(v)
            if not same(m, rm):
                what = "view_matrices"
            elif not same(c, rc):
                what = "frustum corners"
            if not same(lib.view_uniforms(v, pos).view(np.float32), o.view_uniforms(v, pos).view(np.float32)):
                what = what or "view_uniforms"
            checks += 4
        if l["type"] == scenes.LIGHT_SPOT and not same(lib.light_spot_aabb(l), o.light_spot_aabb(l)):
            what = what or "spot cone box"
        sv, rsv = lib.light_shadow_views(l), o.light_shadow_views(l)
        if len(sv) != len(rsv) or not all(same(np.frombuffer(x.tobytes(), np.float32), np.frombuffer(y.tobytes(), np.float32))
                                          for x, y in zip(sv, rsv)):
            what = what or "shadow views"
        if not same(lib.light_uniforms(l, att, bias).view(np.float32), o.light_uniforms(l, att, bias).view(np.float32)):
            what = what or "light_uniforms"
        checks += 3
        if what:
            print(json.dumps({"case": case, "what": what, "view": [x.tolist() if hasattr(x, "tolist") else x for x in v.tolist()],
                              "light": [x.tolist() if hasattr(x, "tolist") else x for x in l.tolist()]}, default=str))
            sys.exit(1)
    print(json.dumps({"cases": a.cases, "comparisons": checks, "differences": 0}))


if __name__ == "__main__":
    main()
