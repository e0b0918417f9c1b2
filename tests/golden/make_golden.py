#!/usr/bin/env python
"""Generates tests/golden/visibility_golden.npz from the REFERENCE-COMPILED oracle (oracle/_ref, i.e. the reference's
own sources built in place from /root/reference). Run in the container that has the reference:

    python tests/golden/make_golden.py

The fixture holds seeded inputs and the reference's outputs for every function on the path, so that the C
restatement (and, on a GPU box, the CUDA path) can be pinned where /root/reference and oracle/_ref do not exist.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import pyoracle  # noqa: E402
from orion_b200 import scenes  # noqa: E402
import vis_testutil as util  # noqa: E402


def main():
    ref = pyoracle.load("reference")
    assert ref.name == "reference"
    rng = np.random.default_rng(2026)
    g = {}
    # views
    views = np.concatenate([util.random_views(rng, 24), np.asarray(scenes.default_camera()).reshape(1),
                            np.asarray(scenes.cubes_scene()["view"]).reshape(1)])
    g["views"] = views
    g["view_planes"] = np.stack([ref.view_planes(v) for v in views])
    mats, corners = zip(*[ref.view_matrices(v) for v in views])
    g["view_mats"], g["view_corners"] = np.stack(mats), np.stack(corners)
    # transform chain + bounds
    n = 400
    parent = np.empty(n, dtype=np.int32)
    parent[0] = -1
    parent[1:] = (rng.random(n - 1) * np.arange(1, n)).astype(np.int32)
    parent[1:20] = np.arange(0, 19)
    local = util.random_trs(rng, n)
    local[:, 7:10] = np.where(rng.random((n, 3)) < 0.1, -1, 1) * rng.uniform(0.5, 1.5, (n, 3))
    boxes = util.random_boxes(rng, n, extent=2.0, hmax=2.0)
    e = util.edge_boxes()
    boxes[:len(e)] = e
    g["parent"], g["local_trs"], g["local_aabb"] = parent, local, boxes
    g["world_trs"] = ref.world_trs(parent, local)
    g["matrices"] = ref.trs_matrix(g["world_trs"])
    g["world_aabb"] = ref.world_aabb(g["world_trs"], boxes)
    # flat cull
    cb = np.concatenate([util.edge_boxes(), util.random_boxes(rng, 3000, extent=120.0)])
    cf = rng.integers(0, 2, len(cb)).astype(np.uint32)
    g["cull_boxes"], g["cull_flags"] = cb, cf
    cull_views = util.random_views(rng, 6, extent=40.0)
    g["cull_views"] = cull_views
    for i, v in enumerate(cull_views):
        pl = ref.view_planes(v)
        g[f"cull_all_{i}"] = ref.cull_aabb(cb, pl)
        g[f"cull_casters_{i}"] = ref.cull_aabb(cb, pl, cf, 1)
    # lights
    lights = util.random_lights(rng, 60, extent=60.0)
    g["lights"] = lights
    g["spot_aabb"] = np.stack([ref.light_spot_aabb(l) if l["type"] == scenes.LIGHT_SPOT else np.zeros(6, np.float32) for l in lights])
    sv = [ref.light_shadow_views(l) for l in lights]
    g["shadow_view_counts"] = np.array([len(x) for x in sv], dtype=np.int32)
    g["shadow_views"] = np.concatenate([x for x in sv if len(x)])
    for i, v in enumerate(cull_views):
        g[f"light_cull_{i}"] = ref.light_cull(lights, ref.view_planes(v))
    # shadow lists (flat) and the cubes scene through the as-shipped world
    cam = scenes.make_view((0, 0, 70), (1, 0, 0, 0), 90.0, 0.1, 300.0, 1600, 900)
    g["lists_view"] = np.asarray(cam).reshape(1)
    for k, x in zip(("ll", "lv", "off", "ent"), ref.shadow_lists(cb, cf, lights, ref.view_planes(cam))):
        g[f"lists_{k}"] = x
    s = scenes.cubes_scene()
    w = ref.world()
    w.add_entities(s["trs"], s["local_aabb"], s["flags"])
    w.add_lights(s["lights"])
    g["cubes_bounds"] = w.bounds()
    g["cubes_entities"], g["cubes_lights"] = w.cull(s["view"])
    for k, x in zip(("ll", "lv", "off", "ent"), w.shadow_lists(s["view"])):
        g[f"cubes_lists_{k}"] = x
    w.close()
    # consumers of the visible lists (section 8(f) rows 1-2): draw-list expansion and EntityUniforms packing
    rng2 = np.random.default_rng(2027)                     # own stream: the arrays above keep their values
    nb = len(cb)
    deferred = np.where(rng2.random(nb) < 0.6, rng2.integers(1, 4, nb), 0)
    basic = np.where(rng2.random(nb) < 0.8, rng2.integers(1, 3, nb), 0)
    shadow = np.where(rng2.random(nb) < 0.7, rng2.integers(1, 3, nb), 0)
    pc = (deferred | (basic << 8) | (shadow << 16)).astype(np.uint32)
    g["draw_pass_counts"] = pc
    for i in range(2):
        g[f"draws_deferred_{i}"], g[f"draws_basic_{i}"] = ref.draw_lists(g[f"cull_all_{i}"], pc, cf, 0)
        g[f"draws_shadow_{i}"], _ = ref.draw_lists(g[f"cull_all_{i}"], pc, cf, 1)
    g["entity_uniforms"] = ref.entity_uniform_blocks(g["world_trs"])      # read back from the RenderEntity's uniform buffer
    assert np.array_equal(g["entity_uniforms"], ref.entity_uniforms(g["world_trs"]), equal_nan=True)
    g["view_viewport_pos"] = rng2.integers(0, 400, (len(views), 2)).astype(np.int32)
    g["view_uniforms"] = np.stack([ref.view_uniforms(v, p) for v, p in zip(views, g["view_viewport_pos"])])
    g["light_attenuation"] = rng2.uniform(0, 1, (len(lights), 3)).astype(np.float32)
    g["light_shadow_bias"] = rng2.uniform(-0.01, 0, len(lights)).astype(np.float32)
    g["light_uniforms"] = np.stack([ref.light_uniforms(l, a, float(b)) for l, a, b in
                                    zip(lights, g["light_attenuation"], g["light_shadow_bias"])])
    out = os.path.join(HERE, "visibility_golden.npz")
    np.savez_compressed(out, **g)
    print(out, os.path.getsize(out), "bytes,", len(g), "arrays")


if __name__ == "__main__":
    main()
