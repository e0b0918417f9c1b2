"""Randomised differential check of the single-process sharded world (ovis_sharded_*: one world over several devices,
a worker thread per shard) against the CPU oracle: host-output cull, many views, shadow-caster lists; worlds smaller than
the shard count (empty shards), removals, caller buffers that start too small. On a one-GPU box every shard sits on
device 0.

    python tools/fuzz_sharded.py [--cases 40] [--seed 1] [--only SEED]

The oracle here is the checker, never the product."""
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


def run_case(oracle, seed):
    import torch
    rng = np.random.default_rng(seed)
    g = torch.cuda.device_count()
    n_shards = int(rng.integers(1, 7))
    n = int(rng.choice([1, 2, 5, 100, 4097, 70_000, 400_000]))
    extent = float(rng.choice([5.0, 300.0, 3000.0]))
    boxes = util.random_boxes(rng, n, extent=extent, hmax=float(rng.choice([0.1, 3.0, extent])))
    if rng.random() < 0.5:
        e = util.edge_boxes()
        k = min(len(e), n)
        boxes[rng.choice(n, k, replace=False)] = e[:k]
    flags = rng.integers(0, 4, n).astype(np.uint32)
    n_lights = int(rng.choice([0, 1, 12, 60]))
    while n_lights * n > 6_000_000:
        n_lights //= 2
    lights = util.random_lights(rng, n_lights, extent=extent)
    if n_lights:
        lights["range"] = np.exp(rng.uniform(np.log(0.05), np.log(extent * 2.0), n_lights)).astype(np.float32)
    os.environ.pop("OVIS_MV_DENSE", None)
    mode = str(rng.choice(["rule", "grid"]))
    if mode == "grid":
        os.environ["OVIS_MV_DENSE"] = "0"
    desc = dict(shards=n_shards, n=n, extent=extent, lights=n_lights, mode=mode)
    sw = lib.ShardedWorld([r % g for r in range(n_shards)])
    sw.add_entities(boxes, flags)
    if n_lights:
        sw.add_lights(lights)
    keep = np.ones(n, dtype=bool)
    for frame in range(3):
        if frame and n > 1:
            alive = np.flatnonzero(keep)
            gone = rng.choice(alive, min(len(alive) - 1, int(rng.choice([0, 1, len(alive) // 3]))), replace=False).astype(np.uint32)
            if len(gone):
                sw.remove_entities(gone)
                keep[gone] = False
        views = util.random_views(rng, 5, extent=extent)
        views[0]["zfar"] = extent * 10.0                       # one wide view: long lists
        views[0]["hfov_deg"] = 120.0
        planes = np.stack([oracle.view_planes(v) for v in views])
        for v in range(2):
            require = int(rng.integers(0, 2))
            exp = oracle.cull_aabb(boxes, planes[v], flags, require, threads=8)
            exp = exp[keep[exp]]
            got, counts = sw.cull(views[v], require)
            if not np.array_equal(got, exp) or counts.sum() != len(exp) or len(counts) != n_shards:
                return desc, f"frame {frame}: cull of view {v}: {len(got)} ids, expected {len(exp)}"
        require = rng.integers(0, 2, 5).astype(np.uint32)
        off, ent = sw.cull_views(planes, require, capacity=int(rng.choice([1, 16, 1 << 16])))
        for v in range(5):
            exp = oracle.cull_aabb(boxes, planes[v], flags, int(require[v]), threads=8)
            exp = exp[keep[exp]]
            if not np.array_equal(ent[off[v]:off[v + 1]], exp):
                return desc, f"frame {frame}: cull_views list {v}: {off[v + 1] - off[v]} ids, expected {len(exp)}"
        if n_lights:
            o_flags = np.where(keep, flags, flags & ~np.uint32(1)).astype(np.uint32)
            got = sw.shadow_lists(views[0], n_lights, capacity=int(rng.choice([1, 16, 1 << 16])))
            want = oracle.shadow_lists(boxes, o_flags, lights, planes[0])
            for name, x, y in zip(("list_light", "list_view", "offsets", "entities"), got, want):
                if not np.array_equal(x, y):
                    return desc, f"frame {frame}: shadow lists differ in {name} ({len(x)} vs {len(y)})"
            desc["lists"] = desc.get("lists", 0) + len(got[0])
    sw.close()
    return desc, None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", type=int, default=40)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--only", type=int, default=None)
    ap.add_argument("--seconds", type=float, default=600.0)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    from oracle import pyoracle
    oracle = pyoracle.load("best")
    seeds = [a.only] if a.only is not None else [a.seed * 10_000 + i for i in range(a.cases)]
    t0 = time.time()
    done, lists, failures = 0, 0, []
    for s in seeds:
        desc, err = run_case(oracle, s)
        done += 1
        print(s, json.dumps(desc), "FAIL: " + err if err else "ok", flush=True)
        if err:
            failures.append({"seed": s, "case": desc, "error": err})
            break
        lists += desc.get("lists", 0)
        if time.time() - t0 > a.seconds:
            break
    summary = {"cases": done, "shadow_lists_compared": lists, "failures": failures, "oracle": oracle.name,
               "seconds": round(time.time() - t0, 1)}
    print(json.dumps(summary))
    if a.out:
        with open(a.out, "w") as f:
            json.dump(summary, f, indent=1)
    sys.exit(1 if failures else 0)


if __name__ == "__main__":
    main()
