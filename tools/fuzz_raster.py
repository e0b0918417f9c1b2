"""Randomised differential check of the headless rasteriser (G-buffer pass, depth-only shadow pass) against its CPU
restatement (oracle/ovis_raster.c): every pixel of every target must hold the same bits.

Each case draws meshes (a cube, a quad, triangle soups with slivers, repeated indices and zero-area triangles), instances
with signed / tiny / large scales, a camera anywhere (outside, among or inside the geometry: triangles cross w = 0, the
rasteriser never clips), and a target size from 1 x 1 to a few hundred pixels with odd widths.

    python tools/fuzz_raster.py [--cases 60] [--seed 1] [--only SEED]

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


def cube_mesh():
    faces = [((1, 0, 0), (0, 1, 0), (0, 0, 1)), ((-1, 0, 0), (0, 0, 1), (0, 1, 0)), ((0, 1, 0), (0, 0, 1), (1, 0, 0)),
             ((0, -1, 0), (1, 0, 0), (0, 0, 1)), ((0, 0, 1), (1, 0, 0), (0, 1, 0)), ((0, 0, -1), (0, 1, 0), (1, 0, 0))]
    pos, nrm, idx = [], [], []
    for n, u, v in faces:
        n, u, v = np.array(n, np.float32), np.array(u, np.float32), np.array(v, np.float32)
        base = len(pos)
        for su, sv in ((-1, -1), (1, -1), (1, 1), (-1, 1)):
            pos.append(0.5 * (n + su * u + sv * v))
            nrm.append(n)
        idx += [base, base + 1, base + 2, base, base + 2, base + 3]
    return np.array(pos, np.float32), np.array(nrm, np.float32), np.array(idx, np.uint32)


def run_textured_case(oracle, seed):
    """The TEXTURED / NORMALMAP variants: random texture sizes (1 x 1 and non-square included), sRGB or linear, repeat or
    clamp, texture coordinates far outside [0, 1], all eight keyword combinations, generated or random tangents."""
    import torch
    rng = np.random.default_rng(seed)
    meshes = []
    cp, cn, ci = cube_mesh()
    meshes.append((cp, cn, ci, (rng.uniform(-3, 3, (len(cp), 2)) * float(rng.choice([1.0, 40.0]))).astype(np.float32)))
    for _ in range(int(rng.integers(0, 3))):
        p_, n_, i_ = soup_mesh(rng)
        uv = rng.uniform(-2, 2, (len(p_), 2)).astype(np.float32)
        if rng.random() < 0.3:
            uv[:] = uv[0]                                           # degenerate mapping: the tangent generator divides by ~0
        meshes.append((p_, n_, i_, uv))
    tangents = []
    for p_, n_, i_, uv in meshes:
        if rng.random() < 0.6:
            t = lib.mesh_tangents(p_, n_, uv, i_)
            t = np.nan_to_num(t, nan=0.0, posinf=1.0, neginf=-1.0).astype(np.float32)      # (NaN payloads differ between CPU and GPU)
        else:
            t = np.concatenate([rng.normal(size=(len(p_), 3)), rng.choice([-1.0, 1.0], (len(p_), 1))], axis=1).astype(np.float32)
        tangents.append(t)
    textures = []
    for _ in range(int(rng.integers(2, 5))):
        tw, th = int(rng.choice([1, 2, 3, 8, 17, 64])), int(rng.choice([1, 2, 5, 8, 33]))
        textures.append((rng.integers(0, 256, (th, tw, 4), dtype=np.uint8), bool(rng.random() < 0.5), bool(rng.random() < 0.5)))
    n = int(rng.choice([1, 6, 40]))
    spread = float(rng.choice([2.0, 8.0]))
    trs = np.zeros((n, 10), dtype=np.float32)
    trs[:, 0:3] = rng.uniform(-spread, spread, (n, 3))
    trs[:, 3:7] = scenes.random_unit_quats(rng, n)
    trs[:, 7:10] = rng.uniform(0.5, 4.0, (n, 3)) * rng.choice([-1.0, 1.0], (n, 3))
    mesh_of = rng.integers(0, len(meshes), n).astype(np.uint32)
    width, height = [(1, 1), (16, 16), (33, 17), (160, 90), (97, 131)][int(rng.integers(0, 5))]
    view = scenes.make_view((0, 0.2 * spread, 2.5 * spread + 2.0), scenes.quat_yaw_pitch(float(rng.uniform(-15, 15)), float(rng.uniform(-15, 5))),
                            float(rng.uniform(40, 100)), 0.1, 40.0 * spread + 10, width, height)
    w = lib.World(0)
    e0 = w.add_entities(np.tile(np.array([[-1, -1, -1, 1, 1, 1]], np.float32), (n, 1)))
    ids = np.arange(e0, e0 + n, dtype=np.uint32)
    w.set_transforms(ids, trs)
    r = lib.Raster(w)
    mids = np.array([r.add_mesh(p_, n_, i_, uv) for p_, n_, i_, uv in meshes], dtype=np.uint32)
    for m, t in zip(mids, tangents):
        r.set_mesh_tangents(int(m), t)
    tids = [r.add_texture(*t) for t in textures]
    n_mat = 8
    mat_rows = []
    for flags in range(n_mat):
        d, sp, sh = rng.uniform(0, 1, 3), rng.uniform(0, 1, 3), float(rng.uniform(1, 64))
        t0, t1 = int(rng.choice(tids)), int(rng.choice(tids))
        r.add_material(d, sp, sh, flags, t0, t1)
        mat_rows.append(list(d) + list(sp) + [sh, flags, t0, t1])
    mat_of = rng.integers(0, n_mat, n).astype(np.uint32)
    r.bind_entities(ids, mids[mesh_of], mat_of)
    order = rng.permutation(n).astype(np.int32)
    d_list = torch.tensor(np.concatenate([[n], order]).astype(np.int32), device="cuda")
    a, b, c = (torch.full((height, width, 4), -1.0, dtype=torch.float32, device="cuda") for _ in range(3))
    depth = torch.full((height, width), -1.0, dtype=torch.float32, device="cuda")
    r.gbuffer(view, d_list.data_ptr() + 4, d_list.data_ptr(), width, height, a.data_ptr(), b.data_ptr(), c.data_ptr(), depth.data_ptr())
    torch.cuda.synchronize()
    vertices = np.concatenate([np.concatenate([p_, n_, uv], axis=1) for p_, n_, i_, uv in meshes])
    all_tangents = np.concatenate(tangents)
    indices = np.concatenate([i_ for _, _, i_, _ in meshes])
    first_v = np.cumsum([0] + [len(m[0]) for m in meshes])
    first_i = np.cumsum([0] + [len(m[2]) for m in meshes])
    mesh_ref = np.array([[first_v[m], first_i[m], len(meshes[m][2]) // 3] for m in range(len(meshes))], dtype=np.int64)
    sel = order.astype(np.int64)
    material = np.array(mat_rows, dtype=np.float32)[mat_of[sel]]
    matrix = oracle.entity_uniforms(trs[sel])[:, :16]
    ea, eb, ec, ed = oracle.raster_textured(matrix, mesh_ref[mesh_of[sel]], material, vertices, all_tangents, indices, textures,
                                            oracle.view_matrices(view)[0][2], width, height)
    desc = dict(textured=True, instances=n, meshes=len(meshes), textures=[list(t[0].shape[:2]) for t in textures], size=[width, height],
                covered=round(float((ed < 1.0).mean()), 3))
    r.close()
    w.close()
    for name, got, exp in (("depth", depth, ed), ("a", a, ea), ("b", b, eb), ("c", c, ec)):
        g = got.cpu().numpy()
        both_nan = np.isnan(g) & np.isnan(exp)                      # a zero tangent normalises to NaN on both sides; payloads differ
        bad = (g.view(np.uint32) != exp.view(np.uint32)) & ~both_nan
        if bad.any():
            where = np.argwhere(bad)[0]
            return desc, f"textured G-buffer target {name}: {int(bad.sum())} words differ, first at {where.tolist()}: {g[tuple(where)]!r} vs {exp[tuple(where)]!r}"
    return desc, None


def soup_mesh(rng):
    nv = int(rng.integers(3, 60))
    nt = int(rng.integers(1, 80))
    pos = rng.uniform(-1, 1, (nv, 3)).astype(np.float32)
    if rng.random() < 0.3:
        pos[:, int(rng.integers(0, 3))] = 0.0                       # all in one plane: edge-on views, slivers
    nrm = rng.normal(size=(nv, 3))
    nrm = (nrm / np.linalg.norm(nrm, axis=1, keepdims=True)).astype(np.float32)
    idx = rng.integers(0, nv, (nt, 3)).astype(np.uint32)            # repeated indices happen: zero-area triangles
    return pos, nrm, idx.reshape(-1)


def run_case(oracle, seed):
    import torch
    rng = np.random.default_rng(seed)
    meshes = [cube_mesh()] + [soup_mesh(rng) for _ in range(int(rng.integers(0, 4)))]
    n = int(rng.choice([1, 2, 9, 60, 300]))
    spread = float(rng.choice([0.5, 4.0, 30.0]))
    trs = np.zeros((n, 10), dtype=np.float32)
    trs[:, 0:3] = rng.uniform(-spread, spread, (n, 3))
    trs[:, 3:7] = scenes.random_unit_quats(rng, n)
    scale_kind = str(rng.choice(["ordinary", "signed", "tiny", "large"]))
    trs[:, 7:10] = {"ordinary": rng.uniform(0.5, 3.0, (n, 3)), "signed": rng.uniform(-3.0, 3.0, (n, 3)),
                    "tiny": rng.uniform(1e-4, 1e-2, (n, 3)), "large": rng.uniform(5.0, 80.0, (n, 3))}[scale_kind]
    mesh_of = rng.integers(0, len(meshes), n).astype(np.uint32)
    materials = np.zeros(3, dtype=lib.MATERIAL_DTYPE)
    materials[0] = ((0.8, 0.2, 0.1), (0.5, 0.5, 0.5), 16.0, lib.MATERIAL_SPECULAR, 0, 0)
    materials[1] = ((0.1, 0.6, 0.3), (0.0, 0.0, 0.0), 0.0, 0, 0, 0)
    materials[2] = ((0.4, 0.4, 0.9), (0.25, 0.5, 1.0), 64.0, lib.MATERIAL_SPECULAR, 0, 0)
    material_of = rng.integers(0, 3, n).astype(np.uint32)
    width, height = [(1, 1), (7, 5), (16, 16), (17, 33), (64, 64), (200, 120), (333, 111), (130, 257)][int(rng.integers(0, 8))]
    cam = str(rng.choice(["outside", "among", "random"]))
    if cam == "outside":
        view = scenes.make_view((0, 0.3 * spread, 3.0 * spread + 2.0), (1, 0, 0, 0), float(rng.uniform(40, 100)), 0.1, 50.0 * spread + 10,
                                width, height)
    elif cam == "among":
        view = scenes.make_view(rng.uniform(-spread, spread, 3) * 0.5, scenes.random_unit_quats(rng, 1)[0], float(rng.uniform(60, 120)),
                                float(rng.uniform(0.01, 1.0)), 20.0 * spread + 5, width, height)
    else:
        view = scenes.make_view(rng.uniform(-spread, spread, 3) * 2.0, scenes.random_unit_quats(rng, 1)[0], float(rng.uniform(10, 150)),
                                float(rng.uniform(0.01, 2.0)), float(rng.uniform(3.0, 400.0)), width, height)
    desc = dict(instances=n, meshes=len(meshes), triangles=int(sum(len(meshes[m][2]) // 3 for m in mesh_of)), spread=spread,
                scale=scale_kind, size=[width, height], camera=cam)

    w = lib.World(0)
    local_aabb = np.tile(np.array([[-1, -1, -1, 1, 1, 1]], np.float32), (n, 1))
    e0 = w.add_entities(local_aabb, np.ones(n, dtype=np.uint32))
    ids = np.arange(e0, e0 + n, dtype=np.uint32)
    w.set_transforms(ids, trs)
    r = lib.Raster(w)
    mesh_ids = np.array([r.add_mesh(p, nr, i) for p, nr, i in meshes], dtype=np.uint32)
    for m in materials:
        r.add_material(m["diffuse"], m["specular"], float(m["shininess"]), int(m["flags"]))
    r.bind_entities(ids, mesh_ids[mesh_of], material_of)
    # every instance, in a shuffled draw order (later draws win depth ties: the order is part of the picture)
    order = rng.permutation(n).astype(np.int32)
    d_list = torch.tensor(np.concatenate([[n], order]).astype(np.int32), device="cuda")
    a, b, c = (torch.full((height, width, 4), -1.0, dtype=torch.float32, device="cuda") for _ in range(3))
    depth = torch.full((height, width), -1.0, dtype=torch.float32, device="cuda")
    shadow = torch.full((height, width), -1.0, dtype=torch.float32, device="cuda")
    for _ in range(2):
        r.gbuffer(view, d_list.data_ptr() + 4, d_list.data_ptr(), width, height, a.data_ptr(), b.data_ptr(), c.data_ptr(), depth.data_ptr())
    torch.cuda.synchronize()

    vertices = np.concatenate([np.concatenate([p, nr, np.zeros((len(p), 2), np.float32)], axis=1) for p, nr, _ in meshes])
    indices = np.concatenate([i for _, _, i in meshes])
    first_v = np.cumsum([0] + [len(p) for p, _, _ in meshes])
    first_i = np.cumsum([0] + [len(i) for _, _, i in meshes])
    mesh_ref = np.array([[first_v[m], first_i[m], len(meshes[m][2]) // 3] for m in range(len(meshes))], dtype=np.int64)
    sel = order.astype(np.int64)
    uniforms = oracle.entity_uniforms(trs[sel])[:, :16]
    mats = np.zeros((n, 8), dtype=np.float32)
    mm = materials[material_of[sel]]
    mats[:, 0:3], mats[:, 3:6], mats[:, 6], mats[:, 7] = mm["diffuse"], mm["specular"], mm["shininess"], mm["flags"]
    vp = oracle.view_matrices(view)[0][2]
    ea, eb, ec, ed = oracle.raster(uniforms, mesh_ref[mesh_of[sel]], mats, vertices, indices, vp, width, height)
    desc["covered"] = round(float((ed < 1.0).mean()), 3)

    def differ(got, exp):
        g, e = got.cpu().numpy().view(np.uint32), exp.view(np.uint32)
        return int((g != e).sum())
    for name, got, exp in (("depth", depth, ed), ("a", a, ea), ("b", b, eb), ("c", c, ec)):
        bad = differ(got, exp)
        if bad:
            g, e = got.cpu().numpy(), exp
            where = np.argwhere(g.view(np.uint32) != e.view(np.uint32))[0]
            return desc, f"G-buffer target {name}: {bad} words differ, first at {where.tolist()}: {g[tuple(where)]!r} vs {e[tuple(where)]!r}"
    if width == height:                                             # the depth-only pass wants a square map
        off = np.array([0, n], dtype=np.int64)
        d_ent = torch.tensor(order.astype(np.int32), device="cuda")
        sv = view.copy()
        r.shadow_maps(np.stack([sv]), off, d_ent.data_ptr(), width, [shadow.data_ptr()])
        torch.cuda.synchronize()
        _, _, _, em = oracle.raster(uniforms, mesh_ref[mesh_of[sel]], mats, vertices, indices, vp, width, height, depth_only=True)
        bad = differ(shadow, em)
        if bad:
            return desc, f"shadow map: {bad} texels differ"
        desc["shadow"] = True
    r.close()
    w.close()
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
    done, failures, pixels = 0, [], 0
    for s in seeds:
        desc, err = run_textured_case(oracle, s) if s % 3 == 2 else run_case(oracle, s)      # every third case: textured materials
        done += 1
        print(s, json.dumps(desc), "FAIL: " + err if err else "ok", flush=True)
        if err:
            failures.append({"seed": s, "case": desc, "error": err})
            if len(failures) >= 5:
                break
        else:
            pixels += desc["size"][0] * desc["size"][1]
        if time.time() - t0 > a.seconds:
            break
    summary = {"cases": done, "pixels_compared": pixels, "failures": failures, "oracle": oracle.name,
               "seconds": round(time.time() - t0, 1)}
    print(json.dumps(summary))
    if a.out:
        with open(a.out, "w") as f:
            json.dump(summary, f, indent=1)
    sys.exit(1 if failures else 0)


if __name__ == "__main__":
    main()
