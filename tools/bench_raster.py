#!/usr/bin/env python
"""Times the headless rasteriser: a G-buffer pass at 1920 x 1080 over N instances of a tessellated cube and the six
shadow-map passes of a point light over the same instances. One JSON line.

    python tools/bench_raster.py [instances] [subdivisions per cube edge]"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from orion_b200 import lib, scenes  # noqa: E402


def cube_mesh(k):
    """Unit cube, every face k x k quads (2 k^2 triangles), a normal per face."""
    faces = [((1, 0, 0), (0, 1, 0), (0, 0, 1)), ((-1, 0, 0), (0, 0, 1), (0, 1, 0)), ((0, 1, 0), (0, 0, 1), (1, 0, 0)),
             ((0, -1, 0), (1, 0, 0), (0, 0, 1)), ((0, 0, 1), (1, 0, 0), (0, 1, 0)), ((0, 0, -1), (0, 1, 0), (1, 0, 0))]
    pos, nrm, idx = [], [], []
    for n, u, v in faces:
        n, u, v = np.array(n, np.float32), np.array(u, np.float32), np.array(v, np.float32)
        base = len(pos)
        for j in range(k + 1):
            for i in range(k + 1):
                pos.append(0.5 * n + (i / k - 0.5) * u + (j / k - 0.5) * v)
                nrm.append(n)
        for j in range(k):
            for i in range(k):
                a = base + j * (k + 1) + i
                idx += [a, a + 1, a + k + 2, a, a + k + 2, a + k + 1]
    return np.array(pos, np.float32), np.array(nrm, np.float32), np.array(idx, np.uint32)


def run(n=2000, k=8):
    width, height, shadow = 1920, 1080, 1024
    rng = np.random.default_rng(11)
    trs = np.zeros((n, 10), dtype=np.float32)
    trs[:, 0] = rng.uniform(-40, 40, n)
    trs[:, 1] = rng.uniform(-20, 20, n)
    trs[:, 2] = rng.uniform(-90, -10, n)
    trs[:, 3:7] = scenes.random_unit_quats(rng, n)
    trs[:, 7:10] = rng.uniform(0.8, 2.5, (n, 1))
    w = lib.World(0)
    e0 = w.add_entities(np.tile(np.array([[-0.5, -0.5, -0.5, 0.5, 0.5, 0.5]], np.float32), (n, 1)), np.ones(n, dtype=np.uint32))
    ids = np.arange(e0, e0 + n, dtype=np.uint32)
    w.set_transforms(ids, trs)
    r = lib.Raster(w)
    pos, nrm, idx = cube_mesh(k)
    mesh = r.add_mesh(pos, nrm, idx)
    mat = r.add_material((0.8, 0.6, 0.4), (0.5, 0.5, 0.5), 32.0, lib.MATERIAL_SPECULAR)
    r.bind_entities(ids, mesh, mat)
    light = scenes.make_light(scenes.LIGHT_POINT, position=(0, 0, -50), range=120.0, flags=1)
    w.add_lights(np.array([light]))
    view = scenes.make_view((0, 0, 10), (1, 0, 0, 0), 75.0, 0.1, 500.0, width, height)
    d_vis = torch.zeros(n + 1, dtype=torch.int32, device="cuda")
    w.cull_planes_device(lib.view_planes(view), d_vis.data_ptr() + 4, d_vis.data_ptr())
    a, b, c = (torch.zeros((height, width, 4), dtype=torch.float32, device="cuda") for _ in range(3))
    depth = torch.zeros((height, width), dtype=torch.float32, device="cuda")
    maps = torch.zeros((6, shadow, shadow), dtype=torch.float32, device="cuda")
    ll, lv, off, _, d_ent = w.cull_shadow_views_device(np.array([0], dtype=np.uint32), shadow)
    views = np.stack([lib.light_shadow_views(light, shadow)[lv[q]] for q in range(len(ll))])

    def gbuffer():
        r.gbuffer(view, d_vis.data_ptr() + 4, d_vis.data_ptr(), width, height, a.data_ptr(), b.data_ptr(), c.data_ptr(), depth.data_ptr())
        torch.cuda.synchronize()

    def shadows():
        r.shadow_maps(views, off, d_ent, shadow, [maps[q].data_ptr() for q in range(6)])
        torch.cuda.synchronize()

    def wall(fn, reps=10):
        fn()
        fn()
        t0 = time.perf_counter()
        for _ in range(reps):
            fn()
        return (time.perf_counter() - t0) / reps * 1e3

    g_ms, s_ms = wall(gbuffer), wall(shadows)
    visible = int(d_vis[0].item())
    tris = len(idx) // 3
    covered = float((depth < 1.0).float().mean().item())
    out = {"workload": f"{n} instances of a cube with {tris} triangles, 1920 x 1080 G-buffer, one point light with six 1024^2 shadow maps",
           "visible_instances": visible, "gbuffer_triangles": visible * tris, "gbuffer_ms": g_ms,
           "gbuffer_mtriangles_per_s": visible * tris / (g_ms * 1e-3) / 1e6, "pixels_covered": covered,
           "shadow_instances": int(off[-1]), "shadow_triangles": int(off[-1]) * tris, "shadow_maps_ms": s_ms,
           "shadow_mtriangles_per_s": int(off[-1]) * tris / (s_ms * 1e-3) / 1e6,
           "note": "wall clock per call, synchronous (each call sizes its scratch from device counts: three round trips)"}
    r.close()
    w.close()
    return out


if __name__ == "__main__":
    print(json.dumps(run(int(sys.argv[1]) if len(sys.argv) > 1 else 2000, int(sys.argv[2]) if len(sys.argv) > 2 else 8)))
