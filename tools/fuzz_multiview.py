"""Randomised differential check of the many-view paths (K4 dense pass, K6 light grid) against the CPU oracle.

Every case draws a world (size, extent, box sizes, clustering, edge boxes, caster flags, removals), a set of lights (count,
placement inside / outside / on top of each other, ranges from tiny to larger than the world, spot and point) and a mode
(size rule, grid forced, dense forced, calls split into rounds), runs the shadow-view culls and a batch of plain views
through the library and compares EVERY list with the oracle's brute-force cull of the same view (ids and order).

    python tools/fuzz_multiview.py [--cases 60] [--seed 1] [--budget 2e8]

Prints one line per case and a JSON summary; exit status 1 on the first mismatch (the case's parameters are printed so it
can be replayed with --only SEED). The oracle here is the checker, never the product."""
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


def draw_world(rng, budget_views):
    n = int(rng.choice([1, 2, 31, 33, 257, 1025, 5000, 8193, 60_000, 200_000, 500_000]))
    extent = float(rng.choice([1.0, 50.0, 400.0, 2000.0, 1.0e5]))
    kind = rng.choice(["uniform", "clustered", "one_point", "huge_boxes", "flat", "tiny_boxes"])
    if kind == "uniform":
        boxes = util.random_boxes(rng, n, extent=extent, hmax=extent * 0.01 + 0.1)
    elif kind == "clustered":
        centres = rng.uniform(-extent, extent, (max(n // 500, 1), 3))
        c = centres[rng.integers(0, len(centres), n)] + rng.normal(0, extent * 0.002, (n, 3))
        h = rng.uniform(0.0, extent * 0.003, (n, 3))
        boxes = np.concatenate([c - h, c + h], axis=1).astype(np.float32)
    elif kind == "one_point":
        c = np.tile(rng.uniform(-extent, extent, (1, 3)), (n, 1))
        h = rng.uniform(0.0, 1e-3, (n, 3))
        boxes = np.concatenate([c - h, c + h], axis=1).astype(np.float32)
    elif kind == "huge_boxes":
        c = rng.uniform(-extent, extent, (n, 3))
        h = rng.uniform(0.0, extent * 2.0, (n, 3)) * (rng.random((n, 1)) < 0.02) + rng.uniform(0, 1, (n, 3))
        boxes = np.concatenate([c - h, c + h], axis=1).astype(np.float32)
    elif kind == "flat":
        c = rng.uniform(-extent, extent, (n, 3))
        c[:, 1] = 0.0
        h = rng.uniform(0.0, extent * 0.01, (n, 3))
        h[:, 1] = 0.0
        boxes = np.concatenate([c - h, c + h], axis=1).astype(np.float32)
    else:
        c = rng.uniform(-extent, extent, (n, 3))
        boxes = np.concatenate([c, c], axis=1).astype(np.float32)
    if rng.random() < 0.5:
        e = util.edge_boxes()
        k = min(len(e), n)
        where = rng.choice(n, k, replace=False)
        boxes[where] = e[:k]
    flag_mode = rng.choice(["all", "none", "random", "rare"])
    flags = {"all": np.ones(n), "none": np.zeros(n), "random": rng.random(n) < 0.6, "rare": rng.random(n) < 0.01}[flag_mode]
    flags = flags.astype(np.uint32) | (rng.integers(0, 4, n).astype(np.uint32) << 1)      # other bits must not matter
    max_lights = max(1, int(budget_views // max(n, 1) // 6))
    nl = int(min(rng.choice([1, 2, 7, 33, 64, 300, 1000, 4096]), max_lights))
    place = rng.choice(["inside", "outside", "stacked", "on_entities"])
    if place == "inside":
        pos = rng.uniform(-extent, extent, (nl, 3))
    elif place == "outside":
        pos = rng.uniform(-extent, extent, (nl, 3)) + np.sign(rng.normal(size=(nl, 3))) * extent * 3.0
    elif place == "stacked":
        pos = np.tile(rng.uniform(-extent, extent, (1, 3)), (nl, 1))
    else:
        pick = rng.integers(0, n, nl)
        pos = np.nan_to_num(boxes[pick, :3], nan=0.0, posinf=extent, neginf=-extent).astype(np.float64)
        pos = np.clip(pos, -extent * 10, extent * 10)
    range_kind = rng.choice(["tiny", "small", "mixed", "large", "huge"])
    lo, hi = {"tiny": (1e-3, 0.2), "small": (extent * 0.002 + 0.11, extent * 0.02 + 0.2), "mixed": (0.11, extent * 2.0),
              "large": (extent * 0.3 + 0.2, extent + 0.3), "huge": (extent * 5 + 1, extent * 50 + 2)}[range_kind]
    lights = np.zeros(nl, dtype=lib.LIGHT_DTYPE)
    lights["type"] = scenes.LIGHT_POINT
    n_spot = int(rng.integers(0, nl + 1)) if rng.random() < 0.4 else 0
    if n_spot:
        lights["type"][rng.choice(nl, n_spot, replace=False)] = scenes.LIGHT_SPOT
    lights["position"] = pos.astype(np.float32)
    d = rng.normal(size=(nl, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    lights["direction"] = d.astype(np.float32)
    lights["colour"] = 1.0
    lights["intensity"] = 1.0
    lights["range"] = np.exp(rng.uniform(np.log(lo), np.log(hi), nl)).astype(np.float32)
    lights["cutoff"] = rng.uniform(5.0, 44.0, nl).astype(np.float32)
    lights["flags"] = 1
    n_gone = int(rng.choice([0, 0, 1, n // 10, n // 2])) if n > 1 else 0
    gone = rng.choice(n, n_gone, replace=False).astype(np.uint32) if n_gone else np.zeros(0, dtype=np.uint32)
    desc = dict(n=n, extent=extent, kind=str(kind), flags=str(flag_mode), lights=nl, spots=n_spot, place=str(place),
                ranges=str(range_kind), removed=n_gone)
    return boxes, flags, lights, gone, desc


def run_case(oracle, seed, budget_views):
    rng = np.random.default_rng(seed)
    boxes, flags, lights, gone, desc = draw_world(rng, budget_views)
    mode = rng.choice(["rule", "grid", "dense", "rounds"])
    for k in ("OVIS_MV_DENSE", "OVIS_MV_ROUND_VIEWS"):
        os.environ.pop(k, None)
    if mode == "grid":
        os.environ["OVIS_MV_DENSE"] = "0"
    elif mode == "dense":
        os.environ["OVIS_MV_DENSE"] = "1"
    elif mode == "rounds":
        os.environ["OVIS_MV_DENSE"] = str(int(rng.integers(0, 2)))
        os.environ["OVIS_MV_ROUND_VIEWS"] = str(int(rng.choice([6, 12, 60])))
    desc["mode"] = str(mode)
    w = lib.World(0)
    w.set_world_aabbs(w.add_entities(boxes, flags), boxes)
    w.add_lights(lights)
    if len(gone):
        w.remove_entities(gone)
    keep = np.ones(len(boxes), dtype=bool)
    keep[gone] = False
    ids = np.arange(len(lights), dtype=np.uint32)
    lists = tests = 0
    for call in range(2):                                   # the second call reuses the scratch the first one sized
        ll, lv, off, _, d_ent = w.cull_shadow_views_device(ids)
        ent = w.read_device(d_ent, int(off[-1]))
        k = 0
        for i in ids:
            for face, v in enumerate(oracle.light_shadow_views(lights[i])):
                assert ll[k] == i and lv[k] == face, ("list order", k, ll[k], lv[k])
                exp = oracle.cull_aabb(boxes, oracle.view_planes(v), flags, 1, threads=8)
                exp = exp[keep[exp]]
                got = ent[off[k]:off[k + 1]]
                if not np.array_equal(got, exp):
                    missing, extra = np.setdiff1d(exp, got)[:4], np.setdiff1d(got, exp)[:4]
                    desc["env"] = {k_: os.environ.get(k_) for k_ in ("OVIS_MV_DENSE", "OVIS_MV_ROUND_VIEWS")}
                    desc["light"] = [float(x) for x in lights[i]["position"]] + [float(lights[i]["range"])]
                    desc["missing"] = {int(e): boxes[e].tolist() for e in missing}
                    desc["extra"] = {int(e): boxes[e].tolist() for e in extra}
                    return desc, f"shadow list {k} (light {i} face {face}) call {call}: {len(got)} ids, expected {len(exp)}"
                k += 1
                if call == 0:
                    lists += 1
                    tests += len(boxes)
        if k != len(ll):
            return desc, f"{len(ll)} lists, expected {k}"
        if call == 1:                                        # the host-output form of the same call: the same CSR
            h_ll, h_lv, h_off, h_ent = w.cull_shadow_views(ids)
            if not (np.array_equal(h_ll, ll) and np.array_equal(h_lv, lv) and np.array_equal(h_off, off) and np.array_equal(h_ent, ent)):
                return desc, f"host-output shadow lists differ from the device ones ({len(h_ent)} vs {len(ent)} ids)"
    # plain views: random cameras, duplicates and near-duplicates (runs of views that share P-vertex codes), per-view filters
    nv = int(min(rng.choice([1, 3, 5, 17, 64]), max(1, budget_views // max(len(boxes), 1) // 2)))
    views = util.random_views(rng, nv, extent=desc["extent"])
    if nv >= 3:
        views[1] = views[0]
        views[2] = views[0]
        views[2]["zfar"] = views[0]["zfar"] * 0.5
    planes = np.stack([oracle.view_planes(v) for v in views])
    require = rng.integers(0, 2, nv).astype(np.uint32)
    off, _, d_ent = w.cull_views_device(planes, require)
    ent = w.read_device(d_ent, int(off[-1]))
    for k in range(nv):
        exp = oracle.cull_aabb(boxes, planes[k], flags, int(require[k]), threads=8)
        exp = exp[keep[exp]]
        if not np.array_equal(ent[off[k]:off[k + 1]], exp):
            return desc, f"plain view {k}: {off[k + 1] - off[k]} ids, expected {len(exp)}"
        lists += 1
        tests += len(boxes)
    w.close()
    desc["lists"], desc["tests"] = lists, tests
    return desc, None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", type=int, default=60)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--only", type=int, default=None)
    ap.add_argument("--budget", type=float, default=2e8, help="entity-view tests per case the oracle is asked for")
    ap.add_argument("--seconds", type=float, default=600.0)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    from oracle import pyoracle
    oracle = pyoracle.load("best")
    seeds = [a.only] if a.only is not None else [a.seed * 10_000 + i for i in range(a.cases)]
    t0 = time.time()
    done, lists, tests, failures = 0, 0, 0, []
    for s in seeds:
        desc, err = run_case(oracle, s, a.budget)
        done += 1
        print(s, json.dumps(desc), "FAIL: " + err if err else "ok", flush=True)
        if err:
            failures.append({"seed": s, "case": desc, "error": err})
            break
        lists += desc["lists"]
        tests += desc["tests"]
        if time.time() - t0 > a.seconds:
            break
    summary = {"cases": done, "lists_compared": lists, "entity_view_tests": tests, "failures": failures,
               "oracle": oracle.name, "seconds": round(time.time() - t0, 1)}
    print(json.dumps(summary))
    if a.out:
        with open(a.out, "w") as f:
            json.dump(summary, f, indent=1)
    sys.exit(1 if failures else 0)


if __name__ == "__main__":
    main()
