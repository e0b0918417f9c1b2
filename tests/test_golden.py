"""Committed golden vectors (tests/golden/visibility_golden.npz, produced by tests/golden/make_golden.py from the
reference's own compiled sources) pin the path where /root/reference is absent: the C restatement on any box, the
CUDA library's host math on any box, and the CUDA path itself on a GPU box."""
import os

import numpy as np
import pytest

from orion_b200 import lib, scenes
import vis_testutil as util
from vis_testutil import same

G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "visibility_golden.npz"))


def _lists(prefix):
    return tuple(G[f"{prefix}_{k}"] for k in ("ll", "lv", "off", "ent"))


def test_port_views_and_transforms(port):
    for i, v in enumerate(G["views"]):
        assert same(port.view_planes(v), G["view_planes"][i])
        m, c = port.view_matrices(v)
        assert same(m, G["view_mats"][i]) and same(c, G["view_corners"][i])
    wt = port.world_trs(G["parent"], G["local_trs"])
    assert same(wt, G["world_trs"])
    assert same(port.trs_matrix(wt), G["matrices"])
    assert same(port.world_aabb(wt, G["local_aabb"]), G["world_aabb"])


def test_port_culls_lights_and_lists(port):
    for i, v in enumerate(G["cull_views"]):
        pl = port.view_planes(v)
        assert np.array_equal(port.cull_aabb(G["cull_boxes"], pl), G[f"cull_all_{i}"])
        assert np.array_equal(port.cull_aabb(G["cull_boxes"], pl, G["cull_flags"], 1), G[f"cull_casters_{i}"])
        assert np.array_equal(port.light_cull(G["lights"], pl), G[f"light_cull_{i}"])
    k = 0
    for i, l in enumerate(G["lights"]):
        if l["type"] == scenes.LIGHT_SPOT:
            assert same(port.light_spot_aabb(l), G["spot_aabb"][i])
        sv = port.light_shadow_views(l)
        assert len(sv) == G["shadow_view_counts"][i]
        assert sv.tobytes() == G["shadow_views"][k:k + len(sv)].tobytes()
        k += len(sv)
    got = port.shadow_lists(G["cull_boxes"], G["cull_flags"], G["lights"], port.view_planes(G["lists_view"][0]))
    for x, y in zip(got, _lists("lists")):
        assert np.array_equal(x, y)
    s = scenes.cubes_scene()
    wb = port.world_aabb(s["trs"], s["local_aabb"])
    assert same(wb, G["cubes_bounds"])
    pl = port.view_planes(s["view"])
    assert np.array_equal(port.cull_aabb(wb, pl), G["cubes_entities"])
    assert np.array_equal(port.light_cull(s["lights"], pl), G["cubes_lights"])
    assert len(G["cubes_lists_ll"]) + 1 == 20          # 1 + 1 + 3 x 6 views per frame
    for x, y in zip(port.shadow_lists(wb, s["flags"], s["lights"], pl), _lists("cubes_lists")):
        assert np.array_equal(x, y)


def test_port_draw_lists_and_uniforms(port):
    for i in range(2):
        d0, d1 = port.draw_lists(G[f"cull_all_{i}"], G["draw_pass_counts"], G["cull_flags"], 0)
        assert np.array_equal(d0, G[f"draws_deferred_{i}"]) and np.array_equal(d1, G[f"draws_basic_{i}"])
        s0, s1 = port.draw_lists(G[f"cull_all_{i}"], G["draw_pass_counts"], G["cull_flags"], 1)
        assert np.array_equal(s0, G[f"draws_shadow_{i}"]) and len(s1) == 0
    assert same(port.entity_uniforms(G["world_trs"]), G["entity_uniforms"])
    for v, pos, want in zip(G["views"], G["view_viewport_pos"], G["view_uniforms"]):
        assert port.view_uniforms(v, pos).tobytes() == want.tobytes()
    for l, att, bias, want in zip(G["lights"], G["light_attenuation"], G["light_shadow_bias"], G["light_uniforms"]):
        assert port.light_uniforms(l, att, float(bias)).tobytes() == want.tobytes()


def test_library_host_math_against_golden():
    for v, pos, want in zip(G["views"], G["view_viewport_pos"], G["view_uniforms"]):
        assert lib.view_uniforms(v, pos).tobytes() == want.tobytes()
    for l, att, bias, want in zip(G["lights"], G["light_attenuation"], G["light_shadow_bias"], G["light_uniforms"]):
        assert lib.light_uniforms(l, att, float(bias)).tobytes() == want.tobytes()
    for i, v in enumerate(G["views"]):
        assert same(lib.view_planes(v), G["view_planes"][i])
        m, c = lib.view_matrices(v)
        assert same(m, G["view_mats"][i]) and same(c, G["view_corners"][i])
    k = 0
    for i, l in enumerate(G["lights"]):
        if l["type"] == scenes.LIGHT_SPOT:
            assert same(lib.light_spot_aabb(l), G["spot_aabb"][i])
        sv = lib.light_shadow_views(l)
        assert sv.tobytes() == G["shadow_views"][k:k + len(sv)].tobytes()
        k += len(sv)


@pytest.mark.gpu
def test_cuda_path_against_golden():
    n = len(G["parent"])
    w = lib.World(0)
    e0 = w.add_entities(G["local_aabb"])
    n0 = w.add_nodes(G["parent"], G["local_trs"])
    w.bind_entities(np.arange(n0, n0 + n), np.arange(e0, e0 + n))
    assert same(w.node_world_trs(), G["world_trs"])
    assert same(w.world_aabbs(), G["world_aabb"])
    assert same(util.gpu_entity_uniforms(w, np.arange(n, dtype=np.uint32)), G["entity_uniforms"])
    w.close()
    w = lib.World(0)
    f = w.add_entities(G["cull_boxes"], G["cull_flags"])
    w.set_world_aabbs(f, G["cull_boxes"])
    w.add_lights(G["lights"])
    for i, v in enumerate(G["cull_views"]):
        ents, ls = w.cull(v)
        assert np.array_equal(ents, G[f"cull_all_{i}"]) and np.array_equal(ls, G[f"light_cull_{i}"])
        assert np.array_equal(w.cull_planes(lib.view_planes(v), require=1), G[f"cull_casters_{i}"])
    w.set_pass_counts(np.arange(len(G["cull_boxes"])), G["draw_pass_counts"] & 0xff, (G["draw_pass_counts"] >> 8) & 0xff,
                      G["draw_pass_counts"] >> 16)
    for i in range(2):
        w.cull(G["cull_views"][i])                               # draw lists straight from the cull's device result
        d0, d1, _ = util.gpu_draw_lists(w, lib.DRAWS_MAIN)
        assert np.array_equal(d0, G[f"draws_deferred_{i}"]) and np.array_equal(d1, G[f"draws_basic_{i}"])
        s0, s1, _ = util.gpu_draw_lists(w, lib.DRAWS_SHADOW)
        assert np.array_equal(s0, G[f"draws_shadow_{i}"]) and len(s1) == 0
    for x, y in zip(w.shadow_lists(G["lists_view"][0]), _lists("lists")):
        assert np.array_equal(x, y)
    for i, (att, bias) in enumerate(zip(G["light_attenuation"], G["light_shadow_bias"])):
        w.set_light_shading(i, att, float(bias))
    _, ls = w.cull(G["cull_views"][0])
    assert len(ls) > 5
    assert util.gpu_light_uniforms(w)[:len(ls)].tobytes() == G["light_uniforms"][ls].tobytes()
    got = util.gpu_view_uniforms(w, G["views"], G["view_viewport_pos"])
    assert got.tobytes() == G["view_uniforms"].tobytes()
    w.close()
    s = scenes.cubes_scene()
    w = lib.World(0)
    w.add_entities(s["local_aabb"], s["flags"])
    w.set_transforms(np.arange(4), s["trs"])
    w.add_lights(s["lights"])
    assert same(w.world_aabbs(), G["cubes_bounds"])
    ents, ls = w.cull(s["view"])
    assert np.array_equal(ents, G["cubes_entities"]) and np.array_equal(ls, G["cubes_lights"])
    for x, y in zip(w.shadow_lists(s["view"]), _lists("cubes_lists")):
        assert np.array_equal(x, y)
    w.close()
