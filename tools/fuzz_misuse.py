"""Random sequences of VALID and INVALID calls on one world (ids that were never issued, removed, repeated, far out of
range; parents that are descendants; empty and huge argument arrays; stale ids after a compaction). Every invalid call
must be refused with an error code and leave the world as it was; every valid one is mirrored in a host model; after each
batch the world's cull must still equal the oracle's on the model. The process must never die: the batches run in a child
process and a signal is a failure.

    python tools/fuzz_misuse.py [--cases 40] [--seed 1]

The oracle here is the checker, never the product."""
import argparse
import json
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def wild_ids(rng, n, k):
    """k ids of which about half are wrong in some way; the right ones are distinct (a batch is a parallel scatter: a
    repeated id stores one of its values, unspecified which)."""
    good = rng.permutation(max(n, 1))[np.arange(k) % max(n, 1)].astype(np.int64) if k <= max(n, 1) else rng.integers(0, max(n, 1), k).astype(np.int64)
    if k > max(n, 1):
        good[max(n, 1):] = n                                   # more ids than entities: the surplus is wrong
    bad = rng.choice(np.array([n, n + 1, n + 1000, 2**31 - 1, 2**31, 2**32 - 1], dtype=np.int64), k)
    return np.where(rng.random(k) < 0.5, good, bad).astype(np.uint32)


def run_case(oracle, seed):
    from orion_b200 import lib, scenes
    import vis_testutil as util
    rng = np.random.default_rng(seed)
    n = int(rng.choice([1, 5, 300, 20_000]))
    nodes = int(rng.choice([1, 4, 200, 5000]))
    boxes = util.random_boxes(rng, n, extent=50.0, hmax=2.0)
    flags = rng.integers(0, 2, n).astype(np.uint32)
    parent = np.minimum((rng.random(nodes) * np.arange(nodes)).astype(np.int32), np.arange(nodes) - 1).astype(np.int32)
    parent[0] = -1
    local = util.random_trs(rng, nodes, pos=20.0)
    local[:, 7:10] = rng.uniform(0.5, 1.5, (nodes, 3))
    w = lib.World(0)
    w.add_entities(boxes, flags)
    w.add_nodes(parent, local)
    n_bound = min(n, nodes)
    ent_node = np.full(n, -1, dtype=np.int64)
    ent_node[:n_bound] = rng.integers(0, nodes, n_bound)
    w.bind_entities(ent_node[:n_bound].astype(np.uint32), np.arange(n_bound, dtype=np.uint32))
    static = np.arange(n_bound, n)
    if len(static):
        w.set_world_aabbs_by_id(static.astype(np.uint32), boxes[static])
    lights = util.random_lights(rng, 12, extent=40.0)
    w.add_lights(lights)
    light_alive = np.ones(len(lights), dtype=bool)
    keep = np.ones(n, dtype=bool)
    refused = accepted = 0

    def expect_refusal(fn, *args):
        nonlocal refused
        try:
            fn(*args)
        except lib.OvisError as e:
            assert e.code in (lib.OVIS_ERR_INVALID, lib.OVIS_ERR_CAPACITY), e.code
            refused += 1
            return True
        return False

    def check(stage):
        w.update_transforms()
        import fuzz_world
        wt = fuzz_world.world_trs_any_order(oracle, parent, local)
        wb = boxes.copy()
        bound = ent_node >= 0
        wb[bound] = oracle.world_aabb(wt[ent_node[bound]], boxes[bound])
        view = np.asarray(util.random_views(rng, 1, extent=40.0)[0])
        planes = oracle.view_planes(view)
        exp = oracle.cull_aabb(wb, planes, flags, 0)
        exp = exp[keep[exp]]
        ents, ls = w.cull(view)
        assert np.array_equal(ents, exp), (stage, len(ents), len(exp))
        exp_l = oracle.light_cull(lights, planes)
        assert np.array_equal(ls, exp_l[light_alive[exp_l]]), stage

    check("initial")
    for step in range(60):
        k = int(rng.integers(0, 16))
        m = int(rng.choice([0, 1, 7, 200]))
        ids = wild_ids(rng, n, m)
        all_good = bool(len(ids) == 0 or (ids < n).all() and keep[np.minimum(ids, n - 1)].all())
        if k == 0:                                             # entity setters with a mixed id list: all or nothing
            f = rng.integers(0, 2, m).astype(np.uint32)
            if not expect_refusal(w.set_flags, ids, f):
                assert all_good, ("set_flags accepted", ids.tolist()[:8])
                flags[ids] = f
                accepted += 1
        elif k == 1:
            b = util.random_boxes(rng, m, extent=2.0)
            if not expect_refusal(w.set_local_aabbs, ids, b):
                assert all_good
                # a static entity's world box is its local box (no node): mirror what the reference's setBoundingBox does
                for i, bb in zip(ids, b):
                    boxes[i] = bb
                accepted += 1
        elif k == 2:
            if not expect_refusal(w.remove_entities, ids):
                assert all_good or len(ids) == 0, ("remove accepted", ids.tolist()[:8])
                keep[ids] = False
                accepted += 1
        elif k == 3:
            node_ids = wild_ids(rng, nodes, m)
            t = util.random_trs(rng, m, pos=20.0)
            t[:, 7:10] = rng.uniform(0.5, 1.5, (m, 3))
            if not expect_refusal(w.set_node_local_trs, node_ids, t):
                assert (node_ids < nodes).all()
                for i, tt in zip(node_ids, t):
                    local[i] = tt
                accepted += 1
        elif k == 4:                                           # a parent that is the node itself or one of its descendants
            node = int(rng.integers(0, nodes))
            desc = [i for i in range(nodes) if i != node and _is_ancestor(parent, node, i)]
            target = int(rng.choice(desc)) if desc and rng.random() < 0.7 else int(rng.choice([node, nodes, nodes + 7, 2**31 - 1]))
            assert expect_refusal(w.set_node_parent, node, target), ("set_node_parent accepted", node, target)
        elif k == 5:
            assert expect_refusal(w.remove_node, int(rng.choice([nodes, nodes + 1, 2**31 - 1, 2**32 - 1])))
        elif k == 6:
            lid = int(rng.choice([len(lights), len(lights) + 3, 2**31 - 1, 2**32 - 1]))
            assert expect_refusal(w.update_light, lid, lights[0])
            assert expect_refusal(w.remove_light, lid)
            assert expect_refusal(w.set_light_shading, lid, (1.0, 0.0, 0.0), 0.0)
            assert expect_refusal(w.bind_light_to_node, lid, 0)
            assert expect_refusal(w.bind_light_to_node, 0, int(rng.choice([nodes, 2**31 - 1])))
        elif k == 7:                                           # a spot light wider than the reference allows
            bad = lights[0].copy()
            bad["type"], bad["cutoff"] = scenes.LIGHT_SPOT, 46.0
            assert expect_refusal(w.update_light, 0, bad)
            assert expect_refusal(w.add_lights, np.array([bad]))
        elif k == 8:
            lid = wild_ids(rng, len(lights), int(rng.choice([1, 3, 40])))
            ok_ids = bool((lid < len(lights)).all() and light_alive[np.minimum(lid, len(lights) - 1)].all())
            try:
                w.cull_shadow_views_device(lid)
                assert ok_ids, ("shadow views of unknown lights accepted", lid.tolist())
                accepted += 1
            except lib.OvisError:
                refused += 1
        elif k == 9:
            assert expect_refusal(w.bind_entities, wild_ids(rng, nodes, 3) | np.uint32(1 << 30), np.arange(3, dtype=np.uint32))
            assert expect_refusal(w.bind_entities, np.zeros(2, dtype=np.uint32), np.array([n, 2**32 - 1], dtype=np.uint32))
        elif k == 10:
            assert expect_refusal(w.set_world_aabbs, n, util.random_boxes(rng, 2))            # a range that runs off the end
            if n > 1:
                assert expect_refusal(w.set_world_aabbs, n - 1, util.random_boxes(rng, 2))
        elif k == 11:
            assert expect_refusal(w.add_nodes, np.array([nodes + 5], dtype=np.int32), util.random_trs(rng, 1))    # a parent that does not exist
            assert expect_refusal(w.add_nodes, np.array([-7], dtype=np.int32), util.random_trs(rng, 1))
        elif k == 12:
            pid = wild_ids(rng, n, m)
            ok_ids = bool(len(pid) == 0 or (pid < n).all() and keep[np.minimum(pid, n - 1)].all())
            if not expect_refusal(w.set_pass_counts, pid, np.ones(m, np.uint32), np.ones(m, np.uint32), np.ones(m, np.uint32)):
                assert ok_ids
                accepted += 1
        elif k == 13 and light_alive.sum() > 2:
            i = int(rng.choice(np.flatnonzero(light_alive)))
            w.remove_light(i)
            light_alive[i] = False
            assert expect_refusal(w.remove_light, i)                                           # twice
            assert expect_refusal(w.update_light, i, lights[i])
            accepted += 1
        elif k == 14 and not all_good:
            assert expect_refusal(w.set_transforms, ids, util.random_trs(rng, m)), ("set_transforms accepted", ids.tolist()[:8])
        if step % 10 == 9:
            check(f"step {step}")
    w.close()
    return dict(n=n, nodes=nodes, refused=refused, accepted=accepted)


def _is_ancestor(parent, a, b):
    while b >= 0:
        if b == a:
            return True
        b = parent[b]
    return False


def child(seeds):
    from oracle import pyoracle
    oracle = pyoracle.load("best")
    for s in seeds:
        print(s, json.dumps(run_case(oracle, s)), "ok", flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", type=int, default=40)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--child", default=None)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    if a.child:
        child([int(x) for x in a.child.split(",")])
        return
    seeds = [a.seed * 10_000 + i for i in range(a.cases)]
    t0 = time.time()
    done, refused, failures = 0, 0, []
    for b in range(0, len(seeds), 10):
        batch = seeds[b:b + 10]
        r = subprocess.run([sys.executable, os.path.abspath(__file__), "--child", ",".join(map(str, batch))], capture_output=True, text=True)
        lines = [l for l in r.stdout.splitlines() if l.endswith(" ok")]
        done += len(lines)
        for l in lines:
            refused += json.loads(l.split(" ", 1)[1].rsplit(" ", 1)[0])["refused"]
        if r.returncode != 0:
            failures.append({"batch": batch, "after": len(lines), "returncode": r.returncode, "stderr": r.stderr[-1500:]})
            break
    summary = {"cases": done, "calls_refused": refused, "failures": failures, "seconds": round(time.time() - t0, 1)}
    print(json.dumps(summary))
    if a.out:
        with open(a.out, "w") as f:
            json.dump(summary, f, indent=1)
    sys.exit(1 if failures else 0)


if __name__ == "__main__":
    main()
