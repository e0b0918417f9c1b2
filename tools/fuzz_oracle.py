"""Randomised differential check of the two CPU oracles against each other: the plain-C restatement (oracle/ovis_oracle.c,
"port") and the reference's own compiled classes (oracle/_ref, "reference") must agree value for value on transform
chains, matrices, world boxes, culls, light culls, shadow views and shadow-caster lists - on ordinary and on extreme
inputs (NaN, Inf, huge, tiny, zero and negative scales, unnormalised quaternions, inverted boxes). This is what lets the
port stand in for the reference. Runs on the CPU.

    python tools/fuzz_oracle.py [--cases 300] [--seed 1]"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "tools"))

import vis_testutil as util  # noqa: E402
import fuzz_hostmath  # noqa: E402


def same(a, b):
    a, b = np.asarray(a), np.asarray(b)
    return a.shape == b.shape and bool(np.array_equal(a, b, equal_nan=True))


def wild(rng, a, p=0.02):
    """Sprinkle extreme values over a float array."""
    a = a.copy()
    m = rng.random(a.shape) < p
    a[m] = rng.choice(np.array([np.nan, np.inf, -np.inf, 0.0, -0.0, 1e30, -1e30, 1e-30, 3.4e38, -3.4e38], dtype=np.float32), int(m.sum()))
    return a


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", type=int, default=300)
    ap.add_argument("--seed", type=int, default=1)
    a = ap.parse_args()
    import warnings
    warnings.simplefilter("ignore")
    from oracle import pyoracle
    port, ref = pyoracle.load("port"), pyoracle.load("reference")
    rng = np.random.default_rng(a.seed)
    checks = 0
    for case in range(a.cases):
        n = int(rng.choice([1, 7, 300, 5000]))
        parent = np.minimum((rng.random(n) * np.arange(n)).astype(np.int32), np.arange(n) - 1).astype(np.int32)
        parent[0] = -1
        local = util.random_trs(rng, n, pos=float(rng.choice([1.0, 100.0, 1e6])))
        kind = int(rng.integers(0, 4))
        if kind == 1:
            local[:, 3:7] = rng.normal(0, 3, (n, 4))
        elif kind == 2:
            local[:, 7:10] = np.exp(rng.uniform(-30, 30, (n, 3))) * rng.choice([-1.0, 1.0], (n, 3))
        elif kind == 3:
            local = wild(rng, local, 0.05)
        what = None
        wt_p, wt_r = port.world_trs(parent, local), ref.world_trs(parent, local)
        if not same(wt_p, wt_r):
            what = "world_trs"
        if not same(port.trs_matrix(wt_r), ref.trs_matrix(wt_r)):
            what = what or "trs_matrix"
        boxes = util.random_boxes(rng, n, extent=float(rng.choice([1.0, 50.0, 1e5])), hmax=float(rng.choice([0.01, 3.0, 1e4])))
        if rng.random() < 0.5:
            boxes = wild(rng, boxes, 0.05)
        e = util.edge_boxes()
        boxes[:min(len(e), n)] = e[:min(len(e), n)]
        wb_p, wb_r = port.world_aabb(wt_r, boxes), ref.world_aabb(wt_r, boxes)
        if not same(wb_p, wb_r):
            what = what or "world_aabb"
        if not same(port.entity_uniforms(wt_r), ref.entity_uniforms(wt_r)):
            what = what or "entity_uniforms"
        flags = rng.integers(0, 4, n).astype(np.uint32)
        view = fuzz_hostmath.draw_view(rng)
        if view["vp_width"] <= 0 or view["vp_height"] <= 0:
            view["vp_width"], view["vp_height"] = 16, 9
        planes_p, planes_r = port.view_planes(view), ref.view_planes(view)
        if not same(planes_p, planes_r):
            what = what or "view_planes"
        for req in (0, 1):
            if not np.array_equal(port.cull_aabb(wb_r, planes_r, flags, req), ref.cull_aabb(wb_r, planes_r, flags, req)):
                what = what or "cull_aabb"
        nl = int(rng.choice([1, 9, 60]))
        lights = np.array([fuzz_hostmath.draw_light(rng) for _ in range(nl)], dtype=pyoracle.LIGHT_DTYPE)
        if not np.array_equal(port.light_cull(lights, planes_r), ref.light_cull(lights, planes_r)):
            what = what or "light_cull"
        if n <= 300:
            for x, y in zip(port.shadow_lists(wb_r, flags, lights, planes_r), ref.shadow_lists(wb_r, flags, lights, planes_r)):
                if not np.array_equal(x, y):
                    what = what or "shadow_lists"
        if ref.has_world and n <= 300 and np.isfinite(wb_r).all() and case % 5 == 0:
            w = ref.world()                                      # tier A: the reference's own SimpleRenderWorld on real objects
            w.add_entities(wt_r, boxes, flags & 1)
            if np.isfinite(boxes).all() and np.isfinite(wt_r).all():
                ents, _ = w.cull(view, flags=0)
                if not np.array_equal(ents, ref.cull_aabb(wb_r, planes_r)):
                    what = what or "tier A world cull vs flat cull"
        # the consumers' restatements: draw lists (the reference's DrawList::add on real RenderEntity objects) and uniform blocks
        visible = np.flatnonzero(rng.random(n) < rng.choice([0.0, 0.3, 1.0])).astype(np.uint32)
        pc = util.random_pass_counts(rng, n)
        for mode in (0, 1):
            for x, y in zip(port.draw_lists(visible, pc, flags, mode), ref.draw_lists(visible, pc, flags, mode)):
                if not np.array_equal(x, y):
                    what = what or "draw_lists"
        att = tuple(float(x) for x in rng.uniform(0, 1, 3))
        for l in lights[:6]:
            if not same(port.light_uniforms(l, att, -0.002).view(np.float32), ref.light_uniforms(l, att, -0.002).view(np.float32)):
                what = what or "light_uniforms"
        if not same(port.view_uniforms(view, (3, 4)).view(np.float32), ref.view_uniforms(view, (3, 4)).view(np.float32)):
            what = what or "view_uniforms"
        checks += 12
        if what:
            print(json.dumps({"case": case, "what": what, "n": n, "kind": kind}))
            sys.exit(1)
    print(json.dumps({"cases": a.cases, "comparisons": checks, "differences": 0}))


if __name__ == "__main__":
    main()
