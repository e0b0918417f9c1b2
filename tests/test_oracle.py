"""The C restatement (oracle/ovis_oracle.c) must equal the reference-compiled oracle (oracle/_ref)
value-for-value on every function of the path. Runs only where oracle/_ref is built (the container
that has /root/reference, or a box the prebuilt .so travelled to); tests/test_golden.py pins the same
functions against committed fixtures everywhere else."""
import numpy as np
import pytest

from orion_b200 import scenes
import vis_testutil as util
from vis_testutil import same


def test_view_planes_and_matrices(port, ref):
    rng = np.random.default_rng(11)
    views = util.random_views(rng, 300)
    for v in list(views) + [scenes.default_camera(), scenes.cubes_scene()["view"]]:
        v = np.asarray(v)
        assert same(port.view_planes(v), ref.view_planes(v))
        pm, pc = port.view_matrices(v)
        rm, rc = ref.view_matrices(v)
        assert same(pm, rm) and same(pc, rc)


def test_world_trs_chain(port, ref):
    rng = np.random.default_rng(12)
    n = 5000
    parent = np.empty(n, dtype=np.int32)
    parent[0] = -1
    parent[1:] = (rng.random(n - 1) * np.arange(1, n)).astype(np.int32)   # any earlier node
    parent[1:50] = np.arange(0, 49)                                        # a 50-deep chain
    local = util.random_trs(rng, n)
    assert same(port.world_trs(parent, local), ref.world_trs(parent, local))


def test_matrix_and_world_aabb(port, ref):
    rng = np.random.default_rng(13)
    n = 20000
    trs = util.random_trs(rng, n)
    boxes = util.random_boxes(rng, n)
    trs[:16] = scenes.IDENTITY_TRS
    e = util.edge_boxes()
    boxes[:len(e)] = e
    trs[16, 0] = np.nan
    trs[17, 3] = np.inf
    trs[18, 7:10] = 0
    assert same(port.trs_matrix(trs), ref.trs_matrix(trs))
    assert same(port.world_aabb(trs, boxes), ref.world_aabb(trs, boxes))


def test_cull_flat(port, ref):
    rng = np.random.default_rng(14)
    boxes = np.concatenate([util.edge_boxes(), util.random_boxes(rng, 50000, extent=300.0)])
    flags = rng.integers(0, 4, len(boxes)).astype(np.uint32)
    for v in util.random_views(rng, 20, extent=50.0):
        planes = ref.view_planes(np.asarray(v))
        a = ref.cull_aabb(boxes, planes)
        assert np.array_equal(port.cull_aabb(boxes, planes), a)
        assert np.array_equal(port.cull_aabb(boxes, planes, threads=3), a)
        assert np.array_equal(ref.cull_aabb(boxes, planes, threads=4), a)
        b = ref.cull_aabb(boxes, planes, flags, 1)
        assert np.array_equal(port.cull_aabb(boxes, planes, flags, 1), b)
        assert np.array_equal(b, a[(flags[a] & 1) == 1])
    # axis-aligned planes with zero normal components and boxes touching them exactly
    planes = np.array([[1, 0, 0, -10], [-1, 0, 0, -10], [0, -1, 0, -10], [0, 1, 0, -10], [0, 0, -1, -10],
                       [0, 0, 1, -10]], dtype=np.float32)
    touch = np.array([[10, 0, 0, 11, 1, 1], [-11, 0, 0, -10, 1, 1], [10.000001, 0, 0, 11, 1, 1],
                      [-12, 0, 0, -10.000001, 1, 1], [0, 0, 0, 1, 1, 1]], dtype=np.float32)
    assert np.array_equal(port.cull_aabb(touch, planes), ref.cull_aabb(touch, planes))
    planes[:, :3] *= -0.0   # all-zero normals with negative zero sign: p-vertex picks max, distance = -w
    assert np.array_equal(port.cull_aabb(touch, planes), ref.cull_aabb(touch, planes))


def test_lights(port, ref):
    rng = np.random.default_rng(15)
    lights = util.random_lights(rng, 400)
    for l in lights[lights["type"] == scenes.LIGHT_SPOT]:
        assert same(port.light_spot_aabb(np.asarray(l)), ref.light_spot_aabb(np.asarray(l)))
    for l in lights:
        pv, rv = port.light_shadow_views(np.asarray(l)), ref.light_shadow_views(np.asarray(l))
        assert pv.tobytes() == rv.tobytes()
    for v in util.random_views(rng, 10, extent=50.0):
        planes = ref.view_planes(np.asarray(v))
        assert np.array_equal(port.light_cull(lights, planes), ref.light_cull(lights, planes))


def test_shadow_lists_flat(port, ref):
    rng = np.random.default_rng(16)
    boxes = util.random_boxes(rng, 4000, extent=60.0, hmax=3.0)
    flags = rng.integers(0, 2, len(boxes)).astype(np.uint32)
    lights = util.random_lights(rng, 24, extent=40.0)
    cam = scenes.make_view((0, 0, 70), (1, 0, 0, 0), 90.0, 0.1, 300.0, 1600, 900)
    planes = ref.view_planes(cam)
    a, b = port.shadow_lists(boxes, flags, lights, planes), ref.shadow_lists(boxes, flags, lights, planes)
    assert len(a[0]) > 0
    for x, y in zip(a, b):
        assert np.array_equal(x, y)


def test_tier_a_world_equals_flat(port, ref):
    """SimpleRenderWorld as shipped (std::list of RenderEntity) == the flat restated loop, including
    removal keeping order, light culling and the prepareLights shadow lists."""
    rng = np.random.default_rng(17)
    n = 3000
    trs = util.random_trs(rng, n, pos=60.0)
    local = util.random_boxes(rng, n, extent=2.0, hmax=2.0)
    flags = rng.integers(0, 2, n).astype(np.uint32)
    lights = util.random_lights(rng, 16, extent=40.0)
    w = ref.world()
    w.add_entities(trs, local, flags)
    w.add_lights(lights)
    wb = w.bounds()
    assert same(wb, port.world_aabb(trs, local))
    assert same(w.matrices(), port.trs_matrix(trs))
    cam = scenes.make_view((0, 0, 80), (1, 0, 0, 0), 90.0, 0.1, 300.0, 1600, 900)
    planes = port.view_planes(cam)
    ents, ls = w.cull(cam)
    assert np.array_equal(ents, port.cull_aabb(wb, planes))
    assert np.array_equal(ls, port.light_cull(lights, planes))
    got = w.shadow_lists(cam)
    exp = port.shadow_lists(wb, flags, lights, planes)
    for x, y in zip(got, exp):
        assert np.array_equal(x, y)
    # removal preserves the order of the survivors (std::list::remove, simple_render_world.cc:65-67)
    removed = rng.choice(n, 500, replace=False)
    for i in removed:
        w.remove_entity(int(i))
    keep = np.ones(n, dtype=bool)
    keep[removed] = False
    ents2, _ = w.cull(cam, flags=0)
    assert np.array_equal(ents2, ents[keep[ents]])
    w.close()


def test_cubes_scene_twenty_views(port, ref):
    s = scenes.cubes_scene()
    w = ref.world()
    w.add_entities(s["trs"], s["local_aabb"], s["flags"])
    w.add_lights(s["lights"])
    ents, ls = w.cull(s["view"])
    ll, lv, off, ent = w.shadow_lists(s["view"])
    assert len(ll) + 1 == 20 or len(ls) < 5      # 1 + 1 + 3*6 views when every light is visible
    wb = port.world_aabb(s["trs"], s["local_aabb"])
    planes = port.view_planes(s["view"])
    assert np.array_equal(ents, port.cull_aabb(wb, planes))
    exp = port.shadow_lists(wb, s["flags"], s["lights"], planes)
    for x, y in zip((ll, lv, off, ent), exp):
        assert np.array_equal(x, y)
    w.close()


def test_draw_lists_and_entity_uniforms(port, ref):
    """Section 8(f) rows 1-2: draw-list expansion (prepareEntities / shadow-caster loop through the reference's own
    DrawList::add control flow) and EntityUniforms packing (Transform::matrix())."""
    rng = np.random.default_rng(29)
    n = 5000
    pc = util.random_pass_counts(rng, n)
    flags = rng.integers(0, 2, n).astype(np.uint32)
    for m in (0, 1, 7, 2500, n):
        visible = np.sort(rng.choice(n, m, replace=False)).astype(np.uint32)
        for mode in (0, 1):
            p0, p1 = port.draw_lists(visible, pc, flags, mode)
            r0, r1 = ref.draw_lists(visible, pc, flags, mode)
            assert np.array_equal(p0, r0) and np.array_equal(p1, r1)
            if mode == 1:
                assert len(p1) == 0
    # The reference tier builds real RenderEntity objects and calls the reference's DrawList::add up to 200 000 visible
    # entities, and falls back to its restated loop beyond: hold both equal to the port.
    big = 400_000
    pc = util.random_pass_counts(rng, big)
    flags = rng.integers(0, 2, big).astype(np.uint32)
    for m in (150_000, 250_000):
        visible = np.sort(rng.choice(big, m, replace=False)).astype(np.uint32)
        for mode in (0, 1):
            p0, p1 = port.draw_lists(visible, pc, flags, mode)
            r0, r1 = ref.draw_lists(visible, pc, flags, mode)
            assert np.array_equal(p0, r0) and np.array_equal(p1, r1)
    trs = util.random_trs(rng, 4000)
    trs[:5, 7:10] = 0.0                                                     # degenerate scale
    trs[5, 3:7] = (1, 0, 0, 0)
    assert same(port.entity_uniforms(trs), ref.entity_uniforms(trs))


def test_uniform_blocks_as_shipped(port, ref):
    """Row 2 of section 8(f), pinned through the reference's own objects: the reference tier reads the 80 / 288 / 208-byte
    blocks back from the (headless) GPU buffers that RenderEntity / RenderView / RenderLight flush into."""
    assert ref.uniform_sizes() == port.uniform_sizes() == (80, 288, 208)
    rng = np.random.default_rng(31)
    trs = util.random_trs(rng, 2000)
    blocks = ref.entity_uniform_blocks(trs)                                  # RenderEntity::setTransform + getResources
    assert same(blocks, ref.entity_uniforms(trs)) and same(blocks, port.entity_uniform_blocks(trs))
    for v in list(util.random_views(rng, 200)) + [scenes.default_camera(), scenes.cubes_scene()["view"]]:
        pos = rng.integers(0, 1000, 2)
        assert port.view_uniforms(v, pos).tobytes() == ref.view_uniforms(v, pos).tobytes()
    lights = np.concatenate([util.random_lights(rng, 1500), scenes.cubes_scene()["lights"]])
    for l in lights:
        att = rng.uniform(0, 1, 3).astype(np.float32)
        bias = float(np.float32(rng.uniform(-0.01, 0)))
        assert port.light_uniforms(l, att, bias).tobytes() == ref.light_uniforms(l, att, bias).tobytes()


def test_shader_restatement_equals_reference_glsl(port, ref):
    """Section 8(f) row 4: oracle/ovis_shading.c against the reference's own shader SOURCE (engine/shaders/lighting.h,
    deferred_light_frag.glsl, tonemap_effect_frag.glsl, gamma_correction_effect_frag.glsl) compiled as C++ against GLM by
    oracle/glsl_harness.cc. Same float operations in the same order: compared for equality."""
    rng = np.random.default_rng(53)
    w, h, size = 80, 45, 24
    view = scenes.make_view((0, 2, 10), (1, 0, 0, 0), 75.0, 0.1, 500.0, w, h)
    gbuf = util.random_gbuffer(rng, w, h)
    vb = ref.view_uniforms(view, (2, 1))
    lights, att, bias = util.lights_for_shading(rng, 40, view)
    blocks = np.stack([ref.light_uniforms(l, a, float(b)) for l, a, b in zip(lights, att, bias)])
    maps = []
    for l in lights:
        if l["flags"] & 1 and l["type"] == scenes.LIGHT_SPOT:
            maps.append(rng.uniform(0.9, 1.0, (size, size)).astype(np.float32))
        elif l["flags"] & 1 and l["type"] == scenes.LIGHT_POINT:
            maps.append(rng.uniform(0.9, 1.0, (6, size, size)).astype(np.float32))
        else:
            maps.append(None)
    for t in range(4):                                                       # one shader variant at a time, then all together
        m = lights["type"] == t
        assert m.sum() >= 3
        sub = [x for x, k in zip(maps, m) if k]
        for shadow in (None, sub):
            glsl = ref.shade_deferred(*gbuf, vb, blocks[m], lights["type"][m], shadow, size, glsl=True)
            assert same(ref.shade_deferred(*gbuf, vb, blocks[m], lights["type"][m], shadow, size), glsl)
            assert same(port.shade_deferred(*gbuf, vb, blocks[m], lights["type"][m], shadow, size), glsl)
    glsl = ref.shade_deferred(*gbuf, vb, blocks, lights["type"], maps, size, glsl=True)
    assert same(port.shade_deferred(*gbuf, vb, blocks, lights["type"], maps, size), glsl) and np.nanmax(glsl[..., :3]) > 0.1
    hdr = glsl.reshape(-1, 4)
    tm = ref.tonemap(hdr, 1.5, 11.2, glsl=True)
    assert same(port.tonemap(hdr, 1.5, 11.2), tm)
    assert same(port.gamma(tm, 2.2), ref.gamma(tm, 2.2, glsl=True))


def test_raster_shader_stages_equal_reference_glsl(port, ref):
    """The shader arithmetic inside the rasteriser's restatement (oracle/ovis_raster.c: vertex transform, world normal,
    G-buffer encode) against the reference's own shader SOURCE - engine/shaders/lit_vtx.glsl, shadow_vtx.glsl and
    lit_deferred_frag.glsl (the variants without TEXTURED / NORMALMAP, with and without SPECULAR) compiled as C++ against
    GLM by oracle/glsl_harness.cc. Same float operations in the same order: compared for equality."""
    rng = np.random.default_rng(54)
    trs = util.random_trs(rng, 200, pos=50.0)
    transforms = ref.entity_uniforms(trs)[:, :16]
    for k in range(200):
        view = scenes.make_view(rng.uniform(-30, 30, 3), scenes.random_unit_quats(rng, 1)[0], float(rng.uniform(30, 120)), 0.1,
                                float(rng.uniform(50, 500)), 640, 360)
        vp = ref.view_matrices(view)[0][2]
        pos = rng.uniform(-3, 3, 3).astype(np.float32)
        nrm = rng.standard_normal(3).astype(np.float32)
        g_clip, g_n = ref.raster_vertex(transforms[k], vp, pos, nrm, glsl=True)
        s_clip, _ = ref.raster_vertex(transforms[k], vp, pos, nrm, glsl=True, shadow=True)
        for o in (port, ref):
            clip, n = o.raster_vertex(transforms[k], vp, pos, nrm)
            assert same(clip, g_clip) and same(n, g_n) and same(clip, s_clip)
        for flags in (0.0, 1.0):
            mat = np.concatenate([rng.uniform(0, 1, 6), [rng.uniform(1, 128), flags]]).astype(np.float32)
            normal = g_n if k % 7 else np.zeros(3, np.float32)          # a degenerate normal: NaN in both
            ga, gb, gc = ref.raster_fragment(normal, mat, glsl=True)
            for o in (port, ref):
                a, b, c = o.raster_fragment(normal, mat)
                assert same(a, ga) and same(b, gb) and same(c, gc)


def test_basic_pass_shaders_equal_reference_glsl(port, ref):
    """unlit_vtx.glsl + unlit_frag.glsl (the Basic pass) from the reference's source: the clip position is the one the
    restatement computes for every pass, the colour is vec4(diffuseColour, 1)."""
    rng = np.random.default_rng(55)
    transforms = ref.entity_uniforms(util.random_trs(rng, 50, pos=50.0))[:, :16]
    for k in range(50):
        view = scenes.make_view(rng.uniform(-30, 30, 3), scenes.random_unit_quats(rng, 1)[0], 70.0, 0.1, 300.0, 640, 360)
        vp = ref.view_matrices(view)[0][2]
        pos = rng.uniform(-3, 3, 3).astype(np.float32)
        diffuse = rng.uniform(0, 1, 3).astype(np.float32)
        clip, colour = ref.glsl_unlit(transforms[k], vp, pos, diffuse)
        for o in (port, ref):
            assert same(o.raster_vertex(transforms[k], vp, pos, np.zeros(3, np.float32))[0], clip)
        assert same(colour, np.concatenate([diffuse, [np.float32(1.0)]]).astype(np.float32))


def test_textured_shader_variants_equal_reference_glsl(port, ref):
    """lit_vtx.glsl with NORMALMAP and all eight keyword combinations of lit_deferred_frag.glsl (SPECULAR / TEXTURED /
    NORMALMAP) from the reference's source against the rasteriser's restatement; the texture sampler is the library's
    definition on both sides (orc_texture_sample)."""
    rng = np.random.default_rng(56)
    textures = [(rng.integers(0, 256, (8, 16, 4), dtype=np.uint8), True, True), (rng.integers(0, 256, (4, 4, 4), dtype=np.uint8), False, False)]
    transforms = ref.entity_uniforms(util.random_trs(rng, 40, pos=20.0))[:, :16]
    vp = ref.view_matrices(scenes.make_view((0, 0, 30), (1, 0, 0, 0), 70.0, 0.1, 300.0, 640, 360))[0][2]
    for k in range(40):
        pos, nrm = rng.uniform(-2, 2, 3).astype(np.float32), rng.standard_normal(3).astype(np.float32)
        tan4 = np.concatenate([rng.standard_normal(3), [rng.choice([-1.0, 1.0])]]).astype(np.float32)
        gn, gt, gb = ref.glsl_vertex_normalmap(transforms[k], vp, pos, nrm, tan4)
        _, n = port.raster_vertex(transforms[k], vp, pos, nrm)
        t = port.raster_vertex(transforms[k], vp, pos, tan4[:3])[1]              # vec3(transform * vec4(tangent, 0))
        assert same(n, gn) and same(t, gt) and same(port.raster_bitangent(n, t, float(tan4[3])), gb)
        uv = rng.uniform(-2, 3, 2).astype(np.float32)
        for flags in range(8):
            mat = np.concatenate([rng.uniform(0, 1, 6), [rng.uniform(1, 128), flags, 0, 1]]).astype(np.float32)
            g = ref.raster_fragment_textured(gn, uv, gt, gb, mat, textures, glsl=True)
            for o in (port, ref):
                got = o.raster_fragment_textured(gn, uv, gt, gb, mat, textures)
                assert all(same(x, y) for x, y in zip(got, g)), flags
    # the sampler itself: texel centres reproduce the texel, wrap and clamp addressing
    rgba, _, _ = textures[1]
    lin = rgba.astype(np.float32) / np.float32(255.0)
    assert same(port.texture_sample((rgba, False, False), 0.125, 0.125), lin[0, 0])
    assert same(port.texture_sample((rgba, False, False), -3.0, 7.0), lin[3, 0])             # clamp
    assert same(port.texture_sample((rgba, False, True), 1.125, -0.875), lin[0, 0])          # repeat
