"""Randomised differential check of a whole frame through one world against the CPU oracle: transform tree -> world boxes
(K1 + K2), cull with lights (K3 + K5), shadow-caster lists from a main view (K5 + K4/K6), with mutations between frames
(moved nodes, changed boxes and flags, removed entities, updated and removed lights).

Each case draws a forest (flat, bushy or chains hundreds of levels deep), local transforms (ordinary, negative / zero /
huge / tiny scales, unnormalised quaternions), entities bound to nodes (several per node), entities moved directly, and
entities with a given world box; lights of all four types with zero intensities, black colours and axis directions.
Everything is compared value for value (boxes, NaNs in the same places) or id for id (lists, in order).

    python tools/fuzz_world.py [--cases 40] [--seed 1] [--only SEED]

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


def draw_trs(rng, n, kind, pos):
    t = util.random_trs(rng, n, pos=pos)
    if kind == "ordinary":
        t[:, 7:10] = rng.uniform(0.5, 1.5, (n, 3))
    elif kind == "signed":
        pass                                                   # util.random_trs: scales in [-3, 3]
    elif kind == "zeros":
        t[:, 7:10] *= rng.random((n, 3)) < 0.8
    elif kind == "huge":
        t[:, 7:10] = np.exp(rng.uniform(-20, 20, (n, 3)))
    elif kind == "raw_quats":
        t[:, 3:7] = rng.normal(0, 2.0, (n, 4))
        t[:, 7:10] = rng.uniform(0.5, 1.5, (n, 3))
    return t.astype(np.float32)


def draw_case(rng):
    n_nodes = int(rng.choice([1, 2, 40, 700, 3000, 40_000, 120_000]))
    shape = rng.choice(["flat", "bushy", "chains", "mixed"])
    parent = np.empty(n_nodes, dtype=np.int32)
    roots = max(1, int(rng.choice([1, 5, 50])))
    idx = np.arange(n_nodes)
    if shape == "flat":
        parent[:] = -1
        parent[roots:] = rng.integers(0, roots, max(n_nodes - roots, 0))
    elif shape == "bushy":
        parent[:] = (rng.random(n_nodes) * idx).astype(np.int32)
    elif shape == "chains":
        parent[:] = idx - roots                                # `roots` chains, each n / roots deep
        if n_nodes > 20_000:                                   # keep the level count (= launches) in the hundreds
            parent[:] = np.where(idx % 200 < roots, (rng.random(n_nodes) * np.minimum(idx, roots)).astype(np.int32), idx - roots)
    else:
        parent[:] = np.where(rng.random(n_nodes) < 0.5, idx - 1, (rng.random(n_nodes) * idx).astype(np.int32))
    parent[:min(roots, n_nodes)] = -1
    parent = np.minimum(parent, idx - 1).astype(np.int32)
    parent[parent < 0] = -1
    trs_kind = str(rng.choice(["ordinary", "ordinary", "signed", "zeros", "huge", "raw_quats"]))
    pos = float(rng.choice([1.0, 30.0, 1000.0]))
    local = draw_trs(rng, n_nodes, trs_kind, pos)
    n_bound = n_nodes + int(rng.choice([0, 1, n_nodes // 3]))
    n_direct = int(rng.choice([0, 1, 500]))
    n_static = int(rng.choice([0, 1, 2000]))
    n_ent = n_bound + n_direct + n_static
    boxes = util.random_boxes(rng, n_ent, extent=float(rng.choice([0.5, 5.0, 100.0])), hmax=float(rng.choice([0.1, 2.0, 30.0])))
    if rng.random() < 0.5:
        e = util.edge_boxes()
        k = min(len(e), n_ent)
        boxes[rng.choice(n_ent, k, replace=False)] = e[:k]
    flags = rng.integers(0, 8, n_ent).astype(np.uint32)
    ent_node = np.concatenate([np.arange(n_nodes), rng.integers(0, n_nodes, n_bound - n_nodes)]).astype(np.uint32)
    n_lights = int(rng.choice([0, 1, 9, 40, 150]))
    while n_lights * n_ent > 4_000_000:
        n_lights //= 2
    lights = util.random_lights(rng, n_lights, extent=pos * 2.0)
    if n_lights and rng.random() < 0.5:
        lights["range"] = np.exp(rng.uniform(np.log(0.05), np.log(pos * 6.0 + 1.0), n_lights)).astype(np.float32)
    desc = dict(nodes=n_nodes, shape=str(shape), trs=trs_kind, pos=pos, entities=n_ent, direct=n_direct, static=n_static,
                lights=n_lights)
    return parent, local, boxes, flags, ent_node, n_bound, n_direct, n_static, lights, desc


def levels_of(parent):
    level = np.full(len(parent), -1, dtype=np.int64)
    for i in range(len(parent)):
        path, a = [], i
        while a >= 0 and level[a] < 0:
            path.append(a)
            a = parent[a]
        l = level[a] if a >= 0 else -1
        for n in reversed(path):
            l += 1
            level[n] = l
    return level


def world_trs_any_order(oracle, parent, local):
    """oracle.world_trs wants parents before children; after re-parenting that need not hold: sort by level first."""
    order = np.argsort(levels_of(parent), kind="stable")
    rank = np.empty(len(order), dtype=np.int64)
    rank[order] = np.arange(len(order))
    p = parent[order]
    p = np.where(p >= 0, rank[np.maximum(p, 0)], -1).astype(np.int32)
    out = np.empty_like(local)
    out[order] = oracle.world_trs(p, local[order])
    return out


def expected_boxes(oracle, parent, local, boxes, ent_node, n_bound, direct_trs, static_boxes):
    wt = world_trs_any_order(oracle, parent, local)
    parts = [oracle.world_aabb(wt[ent_node], boxes[:n_bound])]
    if len(direct_trs):
        parts.append(oracle.world_aabb(direct_trs, boxes[n_bound:n_bound + len(direct_trs)]))
    if len(static_boxes):
        parts.append(static_boxes)
    return wt, np.concatenate(parts)


def run_case(oracle, seed):
    rng = np.random.default_rng(seed)
    parent, local, boxes, flags, ent_node, n_bound, n_direct, n_static, lights, desc = draw_case(rng)
    n_nodes, n_ent = len(parent), len(boxes)
    w = lib.World(0)
    w.add_entities(boxes, flags)
    w.add_nodes(parent, local)
    w.bind_entities(ent_node, np.arange(n_bound))
    direct_ids = np.arange(n_bound, n_bound + n_direct, dtype=np.uint32)
    direct_trs = draw_trs(rng, n_direct, "signed", desc["pos"])
    if n_direct:
        w.set_transforms(direct_ids, direct_trs)
    static_boxes = boxes[n_bound + n_direct:].copy()
    if n_static:
        w.set_world_aabbs(n_bound + n_direct, static_boxes)
    if len(lights):
        w.add_lights(lights)
    pass_counts = util.random_pass_counts(rng, n_ent)
    w.set_pass_counts(np.arange(n_ent), pass_counts & 0xff, (pass_counts >> 8) & 0xff, pass_counts >> 16)
    identity = np.array([0, 0, 0, 1, 0, 0, 0, 1, 1, 1], dtype=np.float32)
    keep = np.ones(n_ent, dtype=bool)
    node_alive = np.ones(n_nodes, dtype=bool)
    light_alive = np.ones(len(lights), dtype=bool)
    frames = 0
    for frame in range(4):
        if frame:
            # ---- mutations
            if n_nodes <= 3000 and rng.random() < 0.6:         # tree surgery (python-side bookkeeping: small trees only)
                level = levels_of(parent)
                for _ in range(int(rng.choice([1, 5, 40]))):
                    live_nodes = np.flatnonzero(node_alive)
                    if len(live_nodes) < 2:
                        break
                    node = int(rng.choice(live_nodes))
                    if rng.random() < 0.75:                    # new parent: a live node that is not below `node`, or none
                        cand = live_nodes[(level[live_nodes] <= level[node]) & (live_nodes != node)]
                        new_parent = int(rng.choice(cand)) if len(cand) and rng.random() < 0.9 else -1
                        w.set_node_parent(node, new_parent)
                        parent[node] = new_parent
                        level = levels_of(parent)
                    elif node_alive.sum() > 2:                 # the subtree goes, with every entity bound into it
                        sub = np.zeros(n_nodes, dtype=bool)
                        sub[node] = True
                        for i in np.argsort(level, kind="stable"):
                            if parent[i] >= 0 and sub[parent[i]]:
                                sub[i] = True
                        sub &= node_alive
                        gone_entities = keep[:n_bound] & sub[ent_node]
                        r_nodes, r_entities = w.remove_node(node)
                        if (r_nodes, r_entities) != (int(sub.sum()), int(gone_entities.sum())):
                            return desc, f"frame {frame}: remove_node reported {(r_nodes, r_entities)}, " \
                                         f"expected {(int(sub.sum()), int(gone_entities.sum()))}"
                        node_alive &= ~sub
                        keep[:n_bound] &= ~gone_entities
                        parent[sub & (parent >= 0) & ~sub[np.maximum(parent, 0)]] = -1   # (the bookkeeping's: detached)
                desc["surgery"] = True
            live_nodes = np.flatnonzero(node_alive)
            k = int(rng.choice([1, max(1, n_nodes // 50), n_nodes]))
            moved = rng.choice(live_nodes, min(k, len(live_nodes)), replace=False).astype(np.uint32)
            local[moved] = draw_trs(rng, len(moved), desc["trs"], desc["pos"])
            w.set_node_local_trs(moved, local[moved])
            live_bound = np.flatnonzero(keep[:n_bound])            # (the library refuses ids of removed entities)
            changed = rng.choice(live_bound, min(len(live_bound), int(rng.choice([0, 1, 40]))), replace=False).astype(np.uint32)
            if len(changed):
                boxes[changed] = util.random_boxes(rng, len(changed), extent=2.0)
                w.set_local_aabbs(changed, boxes[changed])
            live_direct = np.flatnonzero(keep[n_bound:n_bound + n_direct])
            if len(live_direct) and rng.random() < 0.7:
                pick = rng.choice(live_direct, max(1, len(live_direct) // 3), replace=False)
                direct_trs[pick] = draw_trs(rng, len(pick), "signed", desc["pos"])
                w.set_transforms(direct_ids[pick], direct_trs[pick])
            live = np.flatnonzero(keep)
            reflag = rng.choice(live, min(len(live), int(rng.choice([0, 1, 300]))), replace=False).astype(np.uint32)
            if len(reflag):
                flags[reflag] = rng.integers(0, 8, len(reflag)).astype(np.uint32)
                w.set_flags(reflag, flags[reflag])
            alive = np.flatnonzero(keep)
            gone = rng.choice(alive, min(len(alive) - 1, int(rng.choice([0, 1, len(alive) // 4]))), replace=False).astype(np.uint32) \
                if len(alive) > 1 else np.zeros(0, dtype=np.uint32)
            if len(gone):
                w.remove_entities(gone)
                keep[gone] = False
            if not keep.all() and rng.random() < 0.4:
                # slot reclamation: the survivors move up in order, the remap says where; every id after this is a new one
                remap, new_n = w.compact()
                survivors = np.flatnonzero(keep)
                want = np.full(n_ent, 0xFFFFFFFF, dtype=np.uint32)
                want[survivors] = np.arange(len(survivors), dtype=np.uint32)
                if new_n != len(survivors) or len(remap) != n_ent or not np.array_equal(remap, want):
                    return desc, f"frame {frame}: compaction left {new_n} entities, expected {len(survivors)}, or a wrong remap"
                kb, kd, ks = keep[:n_bound], keep[n_bound:n_bound + n_direct], keep[n_bound + n_direct:]
                boxes, flags, pass_counts = boxes[keep], flags[keep], pass_counts[keep]
                ent_node, direct_trs, static_boxes = ent_node[kb], direct_trs[kd], static_boxes[ks]
                n_bound, n_direct, n_ent = int(kb.sum()), int(kd.sum()), len(boxes)
                direct_ids = np.arange(n_bound, n_bound + n_direct, dtype=np.uint32)
                keep = np.ones(n_ent, dtype=bool)
                desc["compacted"] = desc.get("compacted", 0) + 1
            for _ in range(int(rng.integers(0, 4))):
                if not light_alive.any():
                    break
                i = int(rng.choice(np.flatnonzero(light_alive)))
                if rng.random() < 0.5:
                    lights[i] = util.random_lights(rng, 1, extent=desc["pos"] * 2.0)[0]
                    w.update_light(i, lights[i])
                else:
                    w.remove_light(i)
                    light_alive[i] = False
        w.update_transforms()
        wt, wb = expected_boxes(oracle, parent, local, boxes, ent_node, n_bound, direct_trs, static_boxes)
        if not util.same(w.node_world_trs(0, n_nodes)[node_alive], wt[node_alive]):   # (entities moved directly own private nodes after these)
            return desc, f"frame {frame}: node world transforms differ"
        got_b = w.world_aabbs()
        if not util.same(got_b[keep], wb[keep]):
            bad = np.flatnonzero(~np.all((got_b == wb) | (np.isnan(got_b) & np.isnan(wb)), axis=1) & keep)
            return desc, f"frame {frame}: {len(bad)} world boxes differ, first entity {int(bad[0])}"
        view = np.asarray(util.random_views(rng, 1, extent=desc["pos"])[0])
        planes = oracle.view_planes(view)
        ents, ls = w.cull(view)
        exp = oracle.cull_aabb(wb, planes)
        exp = exp[keep[exp]]
        if not np.array_equal(ents, exp):
            return desc, f"frame {frame}: cull returned {len(ents)} entities, expected {len(exp)}"
        # the consumers of the visible list: draw lists (f1) and EntityUniforms blocks (f2), from the cull's own device
        # result and from the list handed in
        for mode in (lib.DRAWS_MAIN, lib.DRAWS_SHADOW):
            e0, e1 = oracle.draw_lists(exp, pass_counts, flags, mode)
            for vis_arg in (None, exp):
                d0, d1, counts = util.gpu_draw_lists(w, mode, visible=vis_arg)
                if counts != (len(e0), len(e1)) or not np.array_equal(d0, e0) or not np.array_equal(d1, e1):
                    return desc, f"frame {frame}: draw lists of mode {mode} differ ({counts} vs {(len(e0), len(e1))})"
        trs_of = np.tile(identity, (n_ent, 1))
        trs_of[:n_bound] = wt[ent_node]
        trs_of[n_bound:n_bound + n_direct] = direct_trs
        want_u = oracle.entity_uniforms(trs_of[exp])
        for vis_arg in (None, exp):
            got_u = util.gpu_entity_uniforms(w, vis_arg)[:len(exp)]
            if not util.same(got_u, want_u):
                return desc, f"frame {frame}: entity uniform blocks differ"
        exp_l = oracle.light_cull(lights, planes) if len(lights) else np.zeros(0, dtype=np.uint32)
        exp_l = exp_l[light_alive[exp_l]] if len(exp_l) else exp_l
        if not np.array_equal(ls, exp_l):
            return desc, f"frame {frame}: cull returned {len(ls)} lights, expected {len(exp_l)}"
        if len(lights):
            o_lights = lights.copy()
            o_lights["intensity"][~light_alive] = 0.0           # a removed light: no effect, no lists; slots keep their ids
            o_flags = np.where(keep, flags, flags & ~np.uint32(1)).astype(np.uint32)
            got = w.shadow_lists(view)
            want = oracle.shadow_lists(wb, o_flags, o_lights, planes)
            for name, x, y in zip(("list_light", "list_view", "offsets", "entities"), got, want):
                if not np.array_equal(x, y):
                    return desc, f"frame {frame}: shadow lists differ in {name} ({len(x)} vs {len(y)})"
            desc["lists"] = desc.get("lists", 0) + len(got[0])
        frames += 1
    w.close()
    desc["frames"] = frames
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
    done, frames, lists, failures = 0, 0, 0, []
    for s in seeds:
        desc, err = run_case(oracle, s)
        done += 1
        print(s, json.dumps(desc), "FAIL: " + err if err else "ok", flush=True)
        if err:
            failures.append({"seed": s, "case": desc, "error": err})
            break
        frames += desc["frames"]
        lists += desc.get("lists", 0)
        if time.time() - t0 > a.seconds:
            break
    summary = {"cases": done, "frames_compared": frames, "shadow_lists_compared": lists, "failures": failures,
               "oracle": oracle.name, "seconds": round(time.time() - t0, 1)}
    print(json.dumps(summary))
    if a.out:
        with open(a.out, "w") as f:
            json.dump(summary, f, indent=1)
    sys.exit(1 if failures else 0)


if __name__ == "__main__":
    main()
