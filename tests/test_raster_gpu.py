"""The headless rasteriser (ovis_raster_*: G-buffer pass and shadow-map passes as CUDA kernels) against its CPU
restatement (oracle/ovis_raster.c): every pixel of every buffer must be the same value - the two evaluate the same IEEE
operations in the same order. The instance matrices on the oracle side come from the oracle's own EntityUniforms
(Transform::matrix), the view-projection from its own view matrices."""
import numpy as np
import pytest

from orion_b200 import lib, scenes
import vis_testutil as util

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def oracle():
    from oracle import pyoracle
    return pyoracle.load("best")


def cube_mesh():
    """Unit cube, 24 vertices (a normal per face), 12 triangles, counter-clockwise seen from outside."""
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


def quad_mesh():
    """A quad in the xz plane facing +y (the cubes demo's floor)."""
    pos = np.array([[-0.5, 0, -0.5], [-0.5, 0, 0.5], [0.5, 0, 0.5], [0.5, 0, -0.5]], np.float32)
    nrm = np.tile(np.array([[0, 1, 0]], np.float32), (4, 1))
    return pos, nrm, np.array([0, 1, 2, 0, 2, 3], np.uint32)


class Scene:
    def __init__(self, rng, n_cubes, spread, with_floor=True):
        self.meshes = [cube_mesh(), quad_mesh()]
        self.materials = np.zeros(3, dtype=lib.MATERIAL_DTYPE)
        self.materials[0] = ((0.8, 0.2, 0.1), (0.5, 0.5, 0.5), 16.0, lib.MATERIAL_SPECULAR, 0, 0)
        self.materials[1] = ((0.1, 0.6, 0.3), (0.0, 0.0, 0.0), 0.0, 0, 0, 0)
        self.materials[2] = ((0.4, 0.4, 0.9), (0.25, 0.5, 1.0), 64.0, lib.MATERIAL_SPECULAR, 0, 0)
        n = n_cubes + (1 if with_floor else 0)
        trs = np.zeros((n, 10), dtype=np.float32)
        trs[:, 0:3] = rng.uniform(-spread, spread, (n, 3))
        trs[:, 3:7] = scenes.random_unit_quats(rng, n)
        trs[:, 7:10] = rng.uniform(0.5, 3.0, (n, 3))
        self.mesh_of = np.zeros(n, dtype=np.uint32)
        self.material_of = rng.integers(0, 3, n).astype(np.uint32)
        if with_floor:
            trs[-1] = (0, -spread, 0, 1, 0, 0, 0, 8 * spread, 1, 8 * spread)
            self.mesh_of[-1] = 1
            self.material_of[-1] = 1
        self.trs = trs
        self.local_aabb = np.tile(np.array([[-0.5, -0.5, -0.5, 0.5, 0.5, 0.5]], np.float32), (n, 1))
        if with_floor:
            self.local_aabb[-1] = (-0.5, 0, -0.5, 0.5, 0, 0.5)
        self.flags = np.ones(n, dtype=np.uint32)
        # the oracle's flat arrays
        self.vertices = np.concatenate([np.concatenate([p, nr, np.zeros((len(p), 2), np.float32)], axis=1) for p, nr, _ in self.meshes])
        self.indices = np.concatenate([i for _, _, i in self.meshes])
        first_v = np.cumsum([0] + [len(p) for p, _, _ in self.meshes])
        first_i = np.cumsum([0] + [len(i) for _, _, i in self.meshes])
        self.mesh_ref = np.array([[first_v[m], first_i[m], len(self.meshes[m][2]) // 3] for m in range(len(self.meshes))], dtype=np.int64)

    def build(self):
        w = lib.World(0)
        e0 = w.add_entities(self.local_aabb, self.flags)
        ids = np.arange(e0, e0 + len(self.trs), dtype=np.uint32)
        w.set_transforms(ids, self.trs)
        r = lib.Raster(w)
        mesh_ids = [r.add_mesh(p, n, i) for p, n, i in self.meshes]
        mat_ids = [r.add_material(m["diffuse"], m["specular"], float(m["shininess"]), int(m["flags"])) for m in self.materials]
        assert mesh_ids == [0, 1] and mat_ids == [0, 1, 2]
        r.bind_entities(ids, self.mesh_of, self.material_of)
        return w, r

    def oracle_inputs(self, oracle, entity_ids):
        uniforms = oracle.entity_uniforms(self.trs[entity_ids])
        mats = np.zeros((len(entity_ids), 8), dtype=np.float32)
        m = self.materials[self.material_of[entity_ids]]
        mats[:, 0:3], mats[:, 3:6], mats[:, 6], mats[:, 7] = m["diffuse"], m["specular"], m["shininess"], m["flags"]
        return uniforms[:, :16], self.mesh_ref[self.mesh_of[entity_ids]], mats


def _same(a, b):
    return np.array_equal(a.view(np.uint32), b.view(np.uint32))


@pytest.mark.parametrize("case", ["outside_looking_in", "camera_among_the_cubes"])
def test_gbuffer_pass_equals_the_restatement(oracle, case):
    import torch
    rng = np.random.default_rng(900 if case == "outside_looking_in" else 901)
    sc = Scene(rng, 40, 6.0)
    w, r = sc.build()
    width, height = 200, 120
    if case == "outside_looking_in":
        view = scenes.make_view((0, 2, 22), (1, 0, 0, 0), 70.0, 0.1, 200.0, width, height)
    else:                                   # cubes cross the near plane and surround the camera: no clipping anywhere
        view = scenes.make_view((0.5, 0.2, 0.3), scenes.quat_yaw_pitch(30.0, -10.0), 95.0, 0.3, 60.0, width, height)
    planes = lib.view_planes(view)
    d_vis = torch.zeros(len(sc.trs) + 1, dtype=torch.int32, device="cuda")
    w.cull_planes_device(planes, d_vis.data_ptr() + 4, d_vis.data_ptr())
    a, b, c = (torch.full((height, width, 4), -1.0, dtype=torch.float32, device="cuda") for _ in range(3))
    depth = torch.full((height, width), -1.0, dtype=torch.float32, device="cuda")
    for _ in range(2):                      # the second call reuses every scratch buffer
        r.gbuffer(view, d_vis.data_ptr() + 4, d_vis.data_ptr(), width, height, a.data_ptr(), b.data_ptr(), c.data_ptr(), depth.data_ptr())
    torch.cuda.synchronize()
    vis = d_vis.cpu().numpy().view(np.uint32)
    ids = vis[1:1 + vis[0]]
    assert len(ids) > 10
    matrix, mesh, material = sc.oracle_inputs(oracle, ids)
    vp = oracle.view_matrices(view)[0][2]
    ea, eb, ec, ed = oracle.raster(matrix, mesh, material, sc.vertices, sc.indices, vp, width, height)
    covered = ed < 1.0
    assert 0.2 < covered.mean() < 1.0       # the picture is neither empty nor trivial
    assert _same(depth.cpu().numpy(), ed)
    assert _same(a.cpu().numpy(), ea) and _same(b.cpu().numpy(), eb) and _same(c.cpu().numpy(), ec)
    # normals decode to unit vectors wherever something was drawn; the cleared pixels hold the clear values
    n = ea[covered][:, :3] * 2.0 - 1.0
    assert np.allclose(np.linalg.norm(n, axis=1), 1.0, atol=1e-5)
    assert np.all(ea[~covered] == 0.0) and np.all(ed[~covered] == 1.0)
    r.close()
    w.close()


def test_shadow_map_passes_equal_the_restatement(oracle):
    import torch
    rng = np.random.default_rng(902)
    sc = Scene(rng, 60, 8.0)
    sc.flags[::3] = 0                        # a third of the entities cast no shadow: not in the lists, not in the maps
    w, r = sc.build()
    lights = np.zeros(2, dtype=scenes.LIGHT_DTYPE)
    lights[0] = scenes.make_light(scenes.LIGHT_POINT, position=(1, 1, -2), range=40.0, flags=1)
    lights[1] = scenes.make_light(scenes.LIGHT_SPOT, position=(0, 14, 3), direction=(0.1, -1, -0.2), range=50.0, cutoff=35.0, flags=1)
    w.add_lights(lights)
    size = 96
    ll, lv, off, d_off, d_ent = w.cull_shadow_views_device(np.arange(2, dtype=np.uint32), shadow_res=size)
    assert len(ll) == 7
    views = np.stack([lib.light_shadow_views(lights[ll[k]], size)[lv[k]] for k in range(len(ll))])
    maps = torch.full((len(ll), size, size), -1.0, dtype=torch.float32, device="cuda")
    r.shadow_maps(views, off, d_ent, size, [maps[k].data_ptr() for k in range(len(ll))])
    torch.cuda.synchronize()
    ent = w.read_device(d_ent, int(off[-1]))
    got = maps.cpu().numpy()
    drawn = 0
    for k in range(len(ll)):
        ids = ent[off[k]:off[k + 1]]
        assert np.all(sc.flags[ids] == 1)
        matrix, mesh, material = sc.oracle_inputs(oracle, ids)
        vp = oracle.view_matrices(views[k])[0][2]
        _, _, _, ed = oracle.raster(matrix, mesh, material, sc.vertices, sc.indices, vp, size, size, depth_only=True)
        assert _same(got[k], ed), k
        drawn += int((ed < 1.0).sum())
    assert drawn > 2000
    r.close()
    w.close()


def test_whole_frame_against_the_restatements(oracle):
    """tools/render_frame.py's built-in scene, every stage of the frame held against its CPU restatement: visible lists ->
    shadow maps and G-buffer (bit for bit) -> light pass with shadows (the tolerance of the shading test: CUDA powf) ->
    tonemap + gamma."""
    import os
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
    import render_frame
    st = {}
    width, height = 160, 90
    render_frame.render(width=width, height=height, keep_state=st, shadow_size=64)
    d = render_frame.builtin_scene.description
    pos, nrm, idx = d["mesh"]
    vertices = np.concatenate([pos, nrm, np.zeros((len(pos), 2), np.float32)], axis=1)
    mat_table = np.array([list(m[0]) + list(m[1]) + [m[2], float(m[3])] for m in d["materials"]], dtype=np.float32)

    def inputs(ids):
        return (oracle.entity_uniforms(d["trs"][ids])[:, :16], np.tile(np.array([[0, 0, len(idx) // 3]], np.int64), (len(ids), 1)),
                mat_table[d["material_of"][ids]])

    view = st["view"]
    mats = oracle.view_matrices(view)[0]
    # the visible list and the G-buffer
    boxes = st["world"].world_aabbs()
    assert np.array_equal(st["visible_entities"], oracle.cull_aabb(boxes, oracle.view_planes(view)))
    ea, eb, ec, ed = oracle.raster(*inputs(st["visible_entities"]), vertices, idx, mats[2], width, height)
    assert _same(st["depth"], ed) and _same(st["a"], ea) and _same(st["b"], eb) and _same(st["c"], ec)
    assert (ed < 1.0).mean() > 0.5
    # the shadow maps of the visible shadow-casting lights
    ll, lv, off, ent = st["shadow_lists"]
    size = st["shadow_size"]
    assert len(ll) == 13                                          # one spot + two point lights
    for k in range(len(ll)):
        sv = lib.light_shadow_views(st["lights"][ll[k]], size)[lv[k]]
        _, _, _, em = oracle.raster(*inputs(ent[off[k]:off[k + 1]]), vertices, idx, oracle.view_matrices(sv)[0][2], size, size, depth_only=True)
        assert _same(st["shadow_maps"][int(ll[k])][lv[k]], em), k
    # the light pass over the G-buffer and the maps, then the post effects
    w = st["world"]
    vis = st["visible_lights"]
    lu = np.stack([oracle.light_uniforms(w.light(int(i))[0], w.light(int(i))[1], float(w.light(int(i))[2])) for i in vis])
    types = np.array([st["lights"][int(i)]["type"] for i in vis], dtype=np.int32)
    maps = [st["shadow_maps"].get(int(i)) for i in vis]
    exp = oracle.shade_deferred(ea, eb, ec, ed, oracle.view_uniforms(view), lu, types, maps, size)
    err = np.abs(st["hdr"] - exp) / np.maximum(np.abs(exp), 1e-3)
    assert err.max() <= 1e-5, err.max()
    lit = exp[..., :3].sum(axis=2)
    assert (lit > 0.05).mean() > 0.3
    no_shadows = oracle.shade_deferred(ea, eb, ec, ed, oracle.view_uniforms(view), lu, types, None, 0)
    assert (no_shadows[..., :3].sum(axis=2) > lit + 1e-3).mean() > 0.02       # the shadow maps do darken part of the picture
    ldr = oracle.gamma(oracle.tonemap(exp.reshape(-1, 4), 1.0, 11.2), 2.2).reshape(height, width, 4)
    assert np.allclose(st["ldr"], ldr, rtol=1e-4, atol=1e-5)
    w.close()


def test_draw_list_input_and_basic_pass(oracle):
    """The deferred draw list ((entity, pass) pairs from ovis_build_draw_lists_device) feeds the G-buffer pass, the basic
    draw list the unlit pass, which draws into the colour target and depth buffer the earlier passes left
    (deferred_render_pipeline.cc:318-330, 506-522)."""
    import torch
    rng = np.random.default_rng(903)
    sc = Scene(rng, 50, 6.0)
    w, r = sc.build()
    n = len(sc.trs)
    basic = np.zeros(n, dtype=bool)
    basic[1::4] = True                                      # a quarter of the entities use a Basic (unlit) material
    ids = np.arange(n, dtype=np.uint32)
    w.set_pass_counts(ids, (~basic).astype(np.uint32), basic.astype(np.uint32), np.ones(n, np.uint32))
    width, height = 160, 100
    view = scenes.make_view((0, 2, 22), (1, 0, 0, 0), 70.0, 0.1, 200.0, width, height)
    d_vis = torch.zeros(n + 1, dtype=torch.int32, device="cuda")
    w.cull_planes_device(lib.view_planes(view), d_vis.data_ptr() + 4, d_vis.data_ptr())
    d0 = torch.zeros((n, 2), dtype=torch.int32, device="cuda")
    d1 = torch.zeros((n, 2), dtype=torch.int32, device="cuda")
    counts = torch.zeros(2, dtype=torch.int32, device="cuda")
    w.build_draw_lists_device(0, d0.data_ptr(), n, d1.data_ptr(), n, d_vis.data_ptr() + 4, d_vis.data_ptr(), d_draw_counts=counts.data_ptr())
    a, b, c = (torch.zeros((height, width, 4), dtype=torch.float32, device="cuda") for _ in range(3))
    depth = torch.zeros((height, width), dtype=torch.float32, device="cuda")
    r.gbuffer_draws(view, d0.data_ptr(), counts.data_ptr(), width, height, a.data_ptr(), b.data_ptr(), c.data_ptr(), depth.data_ptr())
    colour = torch.full((height, width, 4), 0.25, dtype=torch.float32, device="cuda")       # "what the light pass wrote"
    r.basic(view, d1.data_ptr(), counts.data_ptr() + 4, width, height, colour.data_ptr(), depth.data_ptr(), draws=True)
    torch.cuda.synchronize()
    vis = d_vis.cpu().numpy().view(np.uint32)
    visible = vis[1:1 + vis[0]]
    deferred, unlit = visible[~basic[visible]], visible[basic[visible]]
    assert np.array_equal(d0.cpu().numpy().view(np.uint32)[:len(deferred), 0], deferred) and len(unlit) > 3
    vp = oracle.view_matrices(view)[0][2]
    ea, eb, ec, ed = oracle.raster(*sc.oracle_inputs(oracle, deferred), sc.vertices, sc.indices, vp, width, height)
    assert _same(a.cpu().numpy(), ea) and _same(b.cpu().numpy(), eb) and _same(c.cpu().numpy(), ec)
    e_rgba, e_depth = oracle.raster_basic(*sc.oracle_inputs(oracle, unlit), sc.vertices, sc.indices, vp,
                                          np.full((height, width, 4), 0.25, np.float32), ed)
    assert _same(colour.cpu().numpy(), e_rgba) and _same(depth.cpu().numpy(), e_depth)
    touched = (e_rgba[..., 3] == 1.0)
    assert 0.01 < touched.mean() < 0.9 and np.any(e_depth < ed)
    r.close()
    w.close()


def test_textured_and_normal_mapped_materials(oracle):
    """The TEXTURED / NORMALMAP variants of lit_deferred_frag.glsl: sRGB diffuse texture with repeat addressing, a linear
    normal map, tangents as the reference's mesh loader generates them; all four keyword combinations in one picture."""
    import torch
    rng = np.random.default_rng(904)
    width, height = 192, 120
    qp, qn, qi = quad_mesh()
    quv = np.array([[0, 0], [0, 4], [4, 4], [4, 0]], np.float32)               # repeats four times across the quad
    cp, cn, ci = cube_mesh()
    cuv = np.tile(np.array([[0, 0], [1, 0], [1, 1], [0, 1]], np.float32), (6, 1))
    meshes = [(qp, qn, qi, quv), (cp, cn, ci, cuv)]
    tangents = [lib.mesh_tangents(p, n, uv, i) for p, n, i, uv in meshes]
    assert np.allclose(np.linalg.norm(tangents[1][:, :3], axis=1), 1.0, atol=1e-6) and set(np.unique(tangents[1][:, 3])) <= {-1.0, 1.0}
    tex_diffuse = (rng.integers(0, 256, (16, 16, 4), dtype=np.uint8), True, True)
    nm = rng.integers(96, 160, (8, 8, 4), dtype=np.uint8)
    nm[..., 2] = 255
    tex_normal = (nm, False, True)
    w = lib.World(0)
    n_ent = 13
    trs = np.zeros((n_ent, 10), dtype=np.float32)
    trs[:, 0:3] = rng.uniform(-5, 5, (n_ent, 3))
    trs[:, 2] -= 14
    trs[:, 3:7] = scenes.random_unit_quats(rng, n_ent)
    trs[:, 7:10] = rng.uniform(1.5, 3.0, (n_ent, 3))
    trs[0] = (0, -4, -14, 1, 0, 0, 0, 30, 1, 30)
    mesh_of = np.array([0] + [1] * (n_ent - 1), dtype=np.uint32)
    e0 = w.add_entities(np.tile(np.array([[-0.5, -0.5, -0.5, 0.5, 0.5, 0.5]], np.float32), (n_ent, 1)))
    ids = np.arange(e0, e0 + n_ent, dtype=np.uint32)
    w.set_transforms(ids, trs)
    r = lib.Raster(w)
    mids = [r.add_mesh(p, n, i, uv) for p, n, i, uv in meshes]
    for m, t in zip(mids, tangents):
        r.set_mesh_tangents(m, t)
    t0, t1 = r.add_texture(*tex_diffuse), r.add_texture(*tex_normal)
    with pytest.raises(lib.OvisError):
        r.add_material((1, 1, 1), flags=lib.MATERIAL_TEXTURED, diffuse_texture=7)
    mat_rows = []
    for flags in (0, 1, 2, 3, 4, 5, 6, 7):
        d, sp, sh = rng.uniform(0, 1, 3), rng.uniform(0, 1, 3), float(rng.uniform(4, 64))
        r.add_material(d, sp, sh, flags, t0, t1)
        mat_rows.append(list(d) + list(sp) + [sh, flags, t0, t1])
    mat_of = np.array([7] + [k % 8 for k in range(n_ent - 1)], dtype=np.uint32)
    r.bind_entities(ids, np.array(mids, np.uint32)[mesh_of], mat_of)
    view = scenes.make_view((0, 1, 4), scenes.quat_yaw_pitch(0.0, -12.0), 75.0, 0.1, 100.0, width, height)
    d_list = torch.tensor([n_ent] + list(range(n_ent)), dtype=torch.int32, device="cuda")
    a, b, c = (torch.zeros((height, width, 4), dtype=torch.float32, device="cuda") for _ in range(3))
    depth = torch.zeros((height, width), dtype=torch.float32, device="cuda")
    r.gbuffer(view, d_list.data_ptr() + 4, d_list.data_ptr(), width, height, a.data_ptr(), b.data_ptr(), c.data_ptr(), depth.data_ptr())
    torch.cuda.synchronize()
    vertices = np.concatenate([np.concatenate([p, n, uv], axis=1) for p, n, i, uv in meshes])
    all_tangents = np.concatenate(tangents)
    indices = np.concatenate([qi, ci])
    mesh_ref = np.array([[0, 0, 2] if m == 0 else [4, 6, 12] for m in mesh_of], dtype=np.int64)
    material = np.array(mat_rows, dtype=np.float32)[mat_of]
    matrix = oracle.entity_uniforms(trs)[:, :16]
    ea, eb, ec, ed = oracle.raster_textured(matrix, mesh_ref, material, vertices, all_tangents, indices, [tex_diffuse, tex_normal],
                                            oracle.view_matrices(view)[0][2], width, height)
    assert (ed < 1.0).mean() > 0.4
    assert _same(depth.cpu().numpy(), ed) and _same(b.cpu().numpy(), eb) and _same(c.cpu().numpy(), ec) and _same(a.cpu().numpy(), ea)
    # the texture shows: many distinct diffuse values on the floor (material 7), and its normals leave the geometric one
    floor = (eb[..., 3] == 1.0) & (np.abs(ed - ed) == 0) & (ec[..., 3] == np.float32(mat_rows[7][6]))
    assert len(np.unique(eb[floor][:, 0])) > 50
    geometric = oracle.raster(matrix, mesh_ref, material[:, :8] * np.array([1, 1, 1, 1, 1, 1, 1, 0], np.float32) + np.array([0] * 7 + [1], np.float32),
                              vertices, indices, oracle.view_matrices(view)[0][2], width, height)[0]
    assert np.abs(ea[floor] - geometric[floor]).max() > 0.02
    r.close()
    w.close()


def test_raster_edge_cases(oracle):
    """Coplanar duplicates (LessOrEqual: the later draw wins), a degenerate triangle, a NaN vertex, an entity without a
    mesh, an empty visible list: the same picture as the restatement, and nothing faults."""
    import torch
    width, height = 96, 64
    qp, qn, qi = quad_mesh()
    bad_p = qp.copy()
    bad_p[2, 1] = np.nan
    deg_p = np.zeros_like(qp)                                   # all four vertices in one point
    meshes = [(qp, qn, qi), (bad_p, qn, qi), (deg_p, qn, qi)]
    trs = np.zeros((6, 10), dtype=np.float32)
    trs[:, 3] = 1.0
    trs[:, 7:10] = (6, 1, 6)
    trs[:, 0:3] = [(0, -1, -8), (0, -1, -8), (3, -1, -8), (-3, -1, -8), (0, -1, -5), (0, 0.5, -7)]
    mesh_of = np.array([0, 0, 1, 2, 0, 0], dtype=np.uint32)         # entity 5 stays unbound
    mat_of = np.array([0, 1, 0, 0, 2, 0], dtype=np.uint32)
    w = lib.World(0)
    e0 = w.add_entities(np.tile(np.array([[-0.5, 0, -0.5, 0.5, 0, 0.5]], np.float32), (6, 1)))
    ids = np.arange(e0, e0 + 6, dtype=np.uint32)
    w.set_transforms(ids, trs)
    r = lib.Raster(w)
    mids = [r.add_mesh(p, n, i) for p, n, i in meshes]
    mats = np.zeros(3, dtype=lib.MATERIAL_DTYPE)
    mats[0] = ((1, 0, 0), (0, 0, 0), 0.0, 0, 0, 0)
    mats[1] = ((0, 1, 0), (0.5, 0.5, 0.5), 8.0, lib.MATERIAL_SPECULAR, 0, 0)
    mats[2] = ((0, 0, 1), (0, 0, 0), 0.0, 0, 0, 0)
    for m in mats:
        r.add_material(m["diffuse"], m["specular"], float(m["shininess"]), int(m["flags"]))
    r.bind_entities(ids[:5], np.array(mids, np.uint32)[mesh_of[:5]], mat_of[:5])
    view = scenes.make_view((0, 1.5, 0), scenes.quat_yaw_pitch(0.0, -15.0), 80.0, 0.1, 100.0, width, height)
    d_list = torch.tensor([6, 0, 1, 2, 3, 4, 5], dtype=torch.int32, device="cuda")     # count, then every entity in order
    a, b, c = (torch.full((height, width, 4), -1.0, dtype=torch.float32, device="cuda") for _ in range(3))
    depth = torch.full((height, width), -1.0, dtype=torch.float32, device="cuda")
    r.gbuffer(view, d_list.data_ptr() + 4, d_list.data_ptr(), width, height, a.data_ptr(), b.data_ptr(), c.data_ptr(), depth.data_ptr())
    torch.cuda.synchronize()
    vertices = np.concatenate([np.concatenate([p, n, np.zeros((4, 2), np.float32)], axis=1) for p, n, _ in meshes])
    indices = np.concatenate([qi, qi, qi])
    drawn = np.arange(5)
    matrix = oracle.entity_uniforms(trs[drawn])[:, :16]
    mesh_ref = np.array([[4 * m, 6 * m, 2] for m in mesh_of[drawn]], dtype=np.int64)
    material = np.zeros((5, 8), dtype=np.float32)
    mm = mats[mat_of[drawn]]
    material[:, 0:3], material[:, 3:6], material[:, 6], material[:, 7] = mm["diffuse"], mm["specular"], mm["shininess"], mm["flags"]
    ea, eb, ec, ed = oracle.raster(matrix, mesh_ref, material, vertices, indices, oracle.view_matrices(view)[0][2], width, height)
    assert _same(depth.cpu().numpy(), ed) and _same(a.cpu().numpy(), ea) and _same(b.cpu().numpy(), eb) and _same(c.cpu().numpy(), ec)
    got_b = b.cpu().numpy()
    assert (got_b[..., 1] == 1.0).sum() > 200 and (got_b[..., 2] == 1.0).sum() > 50    # the duplicate drawn later (green) shows, and the nearer quad
    # an empty list clears the target
    d_list[0] = 0
    r.gbuffer(view, d_list.data_ptr() + 4, d_list.data_ptr(), width, height, a.data_ptr(), b.data_ptr(), c.data_ptr(), depth.data_ptr())
    torch.cuda.synchronize()
    assert float(depth.min()) == 1.0 and float(a.abs().max()) == 0.0
    r.close()
    w.close()


def test_raster_argument_checks():
    w = lib.World(0)
    w.add_entities(np.array([[-1, -1, -1, 1, 1, 1]], np.float32))
    r = lib.Raster(w)
    p, n, i = cube_mesh()
    with pytest.raises(lib.OvisError):
        r.add_mesh(p, n, np.array([0, 1, 99], np.uint32))          # index out of range
    with pytest.raises(lib.OvisError):
        r.add_mesh(p, n, i[:5])                                      # not a multiple of three
    m = r.add_mesh(p, n, i)
    with pytest.raises(lib.OvisError):
        r.bind_entities([0], [m], [0])                               # no material yet
    mat = r.add_material((1, 1, 1))
    with pytest.raises(lib.OvisError):
        r.bind_entities([5], [m], [mat])                             # unknown entity
    r.bind_entities([0], [m], [mat])
    r.close()
    w.close()
