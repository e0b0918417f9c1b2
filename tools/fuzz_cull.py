"""Randomised differential check of the single-view cull (K3 + K5) through its host-facing entry point, ovis_cull:
sizes around every boundary of the implementation (warp tiles of 128, gather blocks of 8192, the single-launch limit of
512 K, capacity multiples of 1024), visible fractions from nothing to everything, page-locked and pageable output
buffers, buffers that are too small, lights on and off, removals between frames, views changed in place between
back-to-back calls (consecutive culls overlap on the device), several worlds alive at once.

    python tools/fuzz_cull.py [--cases 60] [--seed 1] [--only SEED]

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

SIZES = [0, 1, 31, 32, 33, 127, 128, 129, 1023, 1024, 1025, 8191, 8192, 8193, 65_537, 262_144, 524_287, 524_288, 524_289,
         700_001, 1_048_577, 2_500_000]


def run_case(oracle, seed, others):
    rng = np.random.default_rng(seed)
    n = int(rng.choice(SIZES)) + int(rng.choice([0, 0, 0, int(rng.integers(0, 5000))]))
    extent = float(rng.choice([20.0, 300.0, 3000.0]))
    boxes = util.random_boxes(rng, n, extent=extent, hmax=float(rng.choice([0.01, 4.0, extent * 0.3])))
    if n and rng.random() < 0.5:
        e = util.edge_boxes()
        k = min(len(e), n)
        boxes[rng.choice(n, k, replace=False)] = e[:k]
    flags = rng.integers(0, 4, n).astype(np.uint32)
    n_lights = int(rng.choice([0, 0, 3, 200, 4096]))
    lights = util.random_lights(rng, n_lights, extent=extent)
    pinned = bool(rng.random() < 0.6)
    desc = dict(n=n, extent=extent, lights=n_lights, pinned=pinned)
    w = lib.World(0)
    if n:
        w.set_world_aabbs(w.add_entities(boxes, flags), boxes)
    if n_lights:
        w.add_lights(lights)
    others.append(w)                                           # a few earlier worlds stay alive (their scratch, their streams)
    if len(others) > 3:
        others.pop(0).close()
    hold = lib.PinnedBuffer(n + 8) if pinned else None
    whole = hold.array if pinned else np.empty(n + 8, dtype=np.uint32)
    out = whole[:n + 4]                                        # the capacity the library is told; four guard words behind it
    out_l = np.empty(max(n_lights, 1), dtype=np.uint32)
    view = np.zeros(1, dtype=scenes.VIEW_DTYPE)
    step = w.bind_cull(view, out, flags=lib.CULL_LIGHTS if n_lights else 0, out_lights=out_l if n_lights else None)
    keep = np.ones(n, dtype=bool)
    frames = 0
    for frame in range(int(rng.choice([2, 6, 30]))):
        kind = rng.integers(0, 4)
        if kind == 0:                                          # sees (nearly) everything
            view[0] = scenes.make_view((0, 0, extent * 4.0), (1, 0, 0, 0), 120.0, 0.1, extent * 20.0, 1600, 900)
        elif kind == 1:                                        # sees nothing
            view[0] = scenes.make_view((0, 0, extent * 50.0), scenes.quat_yaw_pitch(180.0, 0.0), 30.0, 0.1, 1.0, 800, 600)
        else:
            view[0] = util.random_views(rng, 1, extent=extent)[0]
        if frame and n > 1 and rng.random() < 0.3:
            alive = np.flatnonzero(keep)
            gone = rng.choice(alive, min(len(alive) - 1, int(rng.choice([1, len(alive) // 5]))), replace=False).astype(np.uint32)
            w.remove_entities(gone)
            keep[gone] = False
        planes = oracle.view_planes(view[0])
        exp = oracle.cull_aabb(boxes, planes, threads=8) if n else np.zeros(0, dtype=np.uint32)
        exp = exp[keep[exp]]
        whole[:] = 0xDEADBEEF
        ne, nl = step()
        if ne != len(exp) or not np.array_equal(out[:ne], exp):
            return desc, f"frame {frame}: {ne} entities, expected {len(exp)}"
        if np.any(whole[n + 4:] != 0xDEADBEEF):                # (between the count and the capacity: unspecified)
            return desc, f"frame {frame}: wrote past the capacity"
        if n_lights:
            exp_l = oracle.light_cull(lights, planes)
            if nl != len(exp_l) or not np.array_equal(out_l[:nl], exp_l):
                return desc, f"frame {frame}: {nl} lights, expected {len(exp_l)}"
        if frame == 0 and len(exp) > 2:                        # a caller buffer that is too small: the size comes back, nothing past it
            cap = int(rng.integers(1, len(exp)))
            small_hold = lib.PinnedBuffer(cap + 4) if pinned else None
            small = small_hold.array if pinned else np.empty(cap + 4, dtype=np.uint32)
            small[:] = 0xDEADBEEF
            try:
                got, _ = w.cull(view[0], flags=0, out_entities=small[:cap])
                return desc, f"a buffer of {cap} for {len(exp)} ids was accepted"
            except lib.OvisError as e:
                if e.code != lib.OVIS_ERR_CAPACITY:
                    return desc, f"undersized buffer: error {e.code}"
            if np.any(small[cap:] != 0xDEADBEEF):
                return desc, "undersized buffer: wrote past its end"
        frames += 1
    if 0 < n <= 1_048_576:
        # the entry point that may not allocate (caller-owned SoA bounds, scratch inside the library), on the last view:
        # the same list (it knows no removals), also when the output is clipped by its capacity
        import torch
        stride = (n + 255) // 256 * 256
        soa = np.zeros((6, stride), dtype=np.float32)
        soa[:, :n] = boxes.T
        d_bounds = torch.from_numpy(soa).cuda()
        exp_all = oracle.cull_aabb(boxes, planes, threads=8)
        for cap in (n, max(1, len(exp_all) // 2)):
            d_out = torch.full((cap + 4,), -1, dtype=torch.int32, device="cuda")
            d_count = torch.zeros(1, dtype=torch.int32, device="cuda")
            base = int(rng.choice([0, 7, 1_000_000]))
            lib.cull_planes_noalloc(0, 0, d_bounds.data_ptr(), stride, n, planes.reshape(-1), d_out.data_ptr(), cap, d_count.data_ptr(),
                                    id_base=base)
            torch.cuda.synchronize()
            count = int(d_count.item())
            got = d_out.cpu().numpy()
            if count != len(exp_all) or not np.array_equal(got[:min(count, cap)].view(np.uint32), exp_all[:cap] + np.uint32(base)):
                return desc, f"no-allocation cull: {count} ids (capacity {cap}), expected {len(exp_all)}"
            if np.any(got[cap:] != -1):
                return desc, "no-allocation cull wrote past its capacity"
        desc["noalloc"] = True
    desc["frames"] = frames
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
    done, frames, failures, others = 0, 0, [], []
    for s in seeds:
        desc, err = run_case(oracle, s, others)
        done += 1
        print(s, json.dumps(desc), "FAIL: " + err if err else "ok", flush=True)
        if err:
            failures.append({"seed": s, "case": desc, "error": err})
            break
        frames += desc["frames"]
        if time.time() - t0 > a.seconds:
            break
    summary = {"cases": done, "frames_compared": frames, "failures": failures, "oracle": oracle.name, "seconds": round(time.time() - t0, 1)}
    print(json.dumps(summary))
    if a.out:
        with open(a.out, "w") as f:
            json.dump(summary, f, indent=1)
    sys.exit(1 if failures else 0)


if __name__ == "__main__":
    main()
