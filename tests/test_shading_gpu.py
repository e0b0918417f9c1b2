"""GPU parity for the deferred light pass and the colour post effects, through the C ABI,
against oracle/ovis_shading.c - which tests/test_oracle.py shows to be bit-identical to the reference's own shader source
compiled against GLM.

Tolerance: everything except pow() is the same IEEE operation sequence on both sides; CUDA's powf is accurate to a few ulp
rather than correctly rounded, so the specular and gamma terms may differ in the last bits. The bound asserted is
|gpu - oracle| <= 1e-5 * max(|oracle|, 1e-3) per channel (north_star's 1e-5 relative, with a floor for values near zero);
passes without pow (ambient-only, tonemap) must be exact."""
import numpy as np
import pytest

from orion_b200 import lib, scenes
import vis_testutil as util

pytestmark = pytest.mark.gpu

RTOL, FLOOR = 1e-5, 1e-3


@pytest.fixture(scope="module")
def oracle():
    from oracle import pyoracle
    return pyoracle.load("best")


def _close(got, want):
    err = np.abs(got.astype(np.float64) - want.astype(np.float64))
    bound = RTOL * np.maximum(np.abs(want.astype(np.float64)), FLOOR)
    ok = (err <= bound) | (np.isnan(got) & np.isnan(want)) | (got == want)
    assert ok.all(), f"{(~ok).sum()} of {ok.size} values off; worst {np.nanmax(err / np.maximum(np.abs(want), FLOOR)):.3g} relative"


def _shade(torch, world, w, h, gbuf, view_block, out_visible=None, maps=None, size=0):
    a, b, c, depth = (torch.from_numpy(x).cuda() for x in gbuf)
    vu = torch.from_numpy(view_block.copy()).cuda()
    out = torch.full((h * w + 1, 4), -3.0, dtype=torch.float32, device="cuda")
    kw = {}
    keep = []
    if out_visible is not None:
        vis = torch.from_numpy(np.ascontiguousarray(out_visible, dtype=np.uint32).view(np.int32)).cuda()
        cnt = torch.tensor([len(out_visible)], dtype=torch.int32, device="cuda")
        keep += [vis, cnt]
        kw.update(d_visible=vis.data_ptr() if len(out_visible) else out.data_ptr(), d_visible_count=cnt.data_ptr(),
                  max_visible=len(out_visible))
    if maps is not None:
        ptrs = np.zeros(len(maps), dtype=np.uint64)
        for i, m in enumerate(maps):
            if m is not None:
                t = torch.from_numpy(m).cuda()
                keep.append(t)
                ptrs[i] = t.data_ptr()
        table = torch.from_numpy(ptrs.view(np.int64)).cuda()
        keep.append(table)
        kw.update(d_shadow_maps=table.data_ptr(), shadow_size=size)
    world.shade_deferred_device(w, h, a.data_ptr(), b.data_ptr(), c.data_ptr(), depth.data_ptr(), vu.data_ptr(), out.data_ptr(), **kw)
    torch.cuda.synchronize()
    assert bool((out[h * w] == -3.0).all())
    return out[:h * w].cpu().numpy().reshape(h, w, 4)


def _scene(rng, oracle, w, h, n_lights):
    view = scenes.make_view((0, 2, 10), scenes.quat_yaw_pitch(10.0, -5.0), 75.0, 0.1, 500.0, w, h)
    gbuf = util.random_gbuffer(rng, w, h)
    lights, att, bias = util.lights_for_shading(rng, n_lights, view)
    world = lib.World(0)
    world.add_lights(lights)
    for i in range(n_lights):
        world.set_light_shading(i, att[i], float(bias[i]))
    blocks = np.stack([oracle.light_uniforms(l, a, float(b)) for l, a, b in zip(lights, att, bias)])
    return view, gbuf, lights, blocks, world


@pytest.mark.parametrize("w,h", [(1, 1), (17, 5), (160, 90), (333, 111)])
def test_deferred_lights_without_shadows(oracle, w, h):
    import torch
    rng = np.random.default_rng(w * 1000 + h)
    view, gbuf, lights, blocks, world = _scene(rng, oracle, w, h, 40)
    pos = (3, 2) if w > 16 else (0, 0)
    vb = oracle.view_uniforms(view, pos)
    ids = np.arange(40, dtype=np.uint32)
    got = _shade(torch, world, w, h, gbuf, vb, ids)
    want = oracle.shade_deferred(*gbuf, vb, blocks, lights["type"])
    _close(got, want)
    assert (got[..., 3] == 1.0).all() and np.nanmax(want[..., :3]) > 0.1
    # any list order is a different sum: the result follows the list, not the light ids
    order = rng.permutation(40).astype(np.uint32)[:23]
    got = _shade(torch, world, w, h, gbuf, vb, order)
    _close(got, oracle.shade_deferred(*gbuf, vb, blocks[order], lights["type"][order]))
    # the light list of the world's own cull
    _, ls = world.cull(view)
    got = _shade(torch, world, w, h, gbuf, vb)
    _close(got, oracle.shade_deferred(*gbuf, vb, blocks[ls], lights["type"][ls]))
    # no lights: the cleared target
    got = _shade(torch, world, w, h, gbuf, vb, np.empty(0, np.uint32))
    assert (got[..., :3] == 0).all() and (got[..., 3] == 1).all()
    world.close()


def test_ambient_and_directional_only_is_exact(oracle):
    """No pow-free path exists for directional lights (specular), but ambient light is pure multiplication: exact."""
    import torch
    rng = np.random.default_rng(12)
    w, h = 64, 48
    view, gbuf, lights, blocks, world = _scene(rng, oracle, w, h, 30)
    vb = oracle.view_uniforms(view)
    amb = np.flatnonzero(lights["type"] == scenes.LIGHT_AMBIENT).astype(np.uint32)
    assert len(amb) >= 3
    got = _shade(torch, world, w, h, gbuf, vb, amb)
    assert np.array_equal(got, oracle.shade_deferred(*gbuf, vb, blocks[amb], lights["type"][amb]))
    world.close()


def test_deferred_lights_with_shadow_maps(oracle):
    import torch
    rng = np.random.default_rng(77)
    w, h, size = 120, 67, 32
    view, gbuf, lights, blocks, world = _scene(rng, oracle, w, h, 48)
    vb = oracle.view_uniforms(view)
    maps = []
    for l in lights:
        casts = bool(l["flags"] & 1)
        if casts and l["type"] == scenes.LIGHT_SPOT:
            maps.append(rng.uniform(0.9, 1.0, (size, size)).astype(np.float32))
        elif casts and l["type"] == scenes.LIGHT_POINT:
            maps.append(rng.uniform(0.9, 1.0, (6, size, size)).astype(np.float32))
        else:
            maps.append(None)
    assert sum(m is not None for m in maps) >= 8
    ids = np.arange(48, dtype=np.uint32)
    got = _shade(torch, world, w, h, gbuf, vb, ids, maps, size)
    want = oracle.shade_deferred(*gbuf, vb, blocks, lights["type"], maps, size)
    _close(got, want)
    unshadowed = oracle.shade_deferred(*gbuf, vb, blocks, lights["type"])
    assert (want[..., :3] <= unshadowed[..., :3] + 1e-6).all() and (want < unshadowed).any()   # the maps do darken
    # a subset in another order: maps are looked up by light id
    order = rng.permutation(48).astype(np.uint32)[:20]
    got = _shade(torch, world, w, h, gbuf, vb, order, maps, size)
    _close(got, oracle.shade_deferred(*gbuf, vb, blocks[order], lights["type"][order], [maps[i] for i in order], size))
    world.close()


def test_post_effects(oracle):
    import torch
    rng = np.random.default_rng(5)
    n = 100_003
    hdr = np.concatenate([rng.uniform(0, 12, (n, 3)), rng.uniform(0, 1, (n, 1))], axis=1).astype(np.float32)
    hdr[:7, :3] = [[0, 0, 0], [1e-8, 1, 1e4], [11.2, 11.2, 11.2], [1, 1, 1], [0.5, 2, 8], [100, 1000, 1e6], [3, 3, 3]]
    d_in = torch.from_numpy(hdr).cuda()
    d_out = torch.zeros_like(d_in)
    w = lib.World(0)
    w.post_effects_device(n, d_in.data_ptr(), d_out.data_ptr(), lib.POST_TONEMAP, exposure=1.5, white_point=11.2)
    torch.cuda.synchronize()
    tm = oracle.tonemap(hdr, 1.5, 11.2)
    assert np.array_equal(d_out.cpu().numpy(), tm)                           # no pow: exact
    w.post_effects_device(n, d_in.data_ptr(), d_out.data_ptr(), lib.POST_GAMMA, gamma=2.2)
    torch.cuda.synchronize()
    _close(d_out.cpu().numpy(), oracle.gamma(hdr, 2.2))
    w.post_effects_device(n, d_in.data_ptr(), d_out.data_ptr(), lib.POST_TONEMAP | lib.POST_GAMMA, 1.5, 11.2, 2.2)
    torch.cuda.synchronize()
    _close(d_out.cpu().numpy(), oracle.gamma(tm, 2.2))
    w.post_effects_device(n, d_in.data_ptr(), d_in.data_ptr(), lib.POST_TONEMAP, 1.5, 11.2)     # in place
    torch.cuda.synchronize()
    assert np.array_equal(d_in.cpu().numpy(), tm)
    with pytest.raises(lib.OvisError):
        w.post_effects_device(n, d_in.data_ptr(), d_out.data_ptr(), 8)
    w.close()


def test_sparse_lights_and_unclean_inputs(oracle):
    """Most pixel-light pairs out of range (the kernel skips them where that is provably exact), plus lights and pixels
    that must NOT take the shortcut: non-finite or out-of-bounds values, zero / negative attenuation terms."""
    import torch
    rng = np.random.default_rng(2024)
    w, h, n = 150, 80, 96
    view = scenes.make_view((0, 2, 10), scenes.quat_yaw_pitch(10.0, -5.0), 75.0, 0.1, 500.0, w, h)
    gbuf = util.random_gbuffer(rng, w, h, linear_z=(0.1, 500.0, 1.0, 300.0))
    a, b, c, depth = gbuf
    a[3, 5, :3] = np.nan                       # pixels that are not "clean"
    b[4, 6, 0] = np.inf
    b[5, 7, 1] = 1e30
    c[6, 8, 3] = -2.0                          # negative shininess: pow(0, -2) = inf
    c[7, 9, 3] = 1e9
    a[8, 10, :3] = (3.0, 3.0, 3.0)             # |normal| > 1
    depth[9, 11] = np.nan
    lights, att, bias = util.lights_for_shading(rng, n, view, extent=150.0)
    lights["range"] = rng.uniform(5.0, 40.0, n)
    att[0] = (0.0, 0.0, 0.0)                   # lights that are not "skippable"
    att[1] = (1.0, -0.5, 0.0)
    att[2] = (1.0, 0.0, -0.001)
    lights["intensity"][3] = np.nan
    lights["colour"][4] = (np.inf, 1.0, 1.0)
    lights["intensity"][5] = 1e30
    lights["cutoff"][6] = 0.0                  # cosCutoff = 1 for a spot light: division by zero in the cone term
    lights["type"][6] = scenes.LIGHT_SPOT
    lights["position"][7] = (1e30, 0.0, 0.0)
    lights["range"][8] = -5.0
    world = lib.World(0)
    world.add_lights(lights)
    for i in range(n):
        world.set_light_shading(i, att[i], float(bias[i]))
    blocks = np.stack([oracle.light_uniforms(l, x, float(y)) for l, x, y in zip(lights, att, bias)])
    vb = oracle.view_uniforms(view)
    size = 16
    maps = [rng.uniform(0.9, 1.0, (size, size) if l["type"] == scenes.LIGHT_SPOT else (6, size, size)).astype(np.float32)
            if (l["flags"] & 1 and l["type"] >= scenes.LIGHT_POINT) else None for l in lights]
    clean = np.arange(9, n, dtype=np.uint32)
    want = oracle.shade_deferred(*gbuf, vb, blocks[clean], lights["type"][clean])
    in_range = 0
    for i in clean[:20]:                       # the scene really is sparse: a light reaches few pixels
        one = oracle.shade_deferred(*gbuf, vb, blocks[i:i + 1], lights["type"][i:i + 1])
        if lights["type"][i] >= scenes.LIGHT_POINT:
            in_range += int((one[..., :3] != 0).any(axis=2).sum())
    assert in_range < 0.2 * 20 * w * h
    _close(_shade(torch, world, w, h, gbuf, vb, clean), want)
    everyone = np.arange(n, dtype=np.uint32)
    for m, sz in ((None, 0), (maps, size)):
        want = oracle.shade_deferred(*gbuf, vb, blocks, lights["type"], m, sz)
        got = _shade(torch, world, w, h, gbuf, vb, everyone, m, sz)
        assert np.array_equal(np.isnan(got), np.isnan(want))
        _close(got, want)
    one_by_one = [9, 0, 1, 2, 3, 4, 5, 6, 7, 8]
    for i in one_by_one:                       # each unclean light alone, so that a NaN from another cannot mask it
        ids = np.array([i], dtype=np.uint32)
        want = oracle.shade_deferred(*gbuf, vb, blocks[ids], lights["type"][ids])
        got = _shade(torch, world, w, h, gbuf, vb, ids)
        assert np.array_equal(np.isnan(got), np.isnan(want)), i
        _close(got, want)
    world.close()


def test_coherent_scene_tile_rejection(oracle):
    """A ground plane seen in perspective with many short-range lights standing on it: most (tile, light) pairs are out of
    range and are rejected per 16 x 16 tile. One tile carries an unclean pixel (so it must not take the shortcut), the
    frame size is not a multiple of the tile size."""
    import torch
    rng = np.random.default_rng(808)
    w, h, n = 333, 187, 160
    view = scenes.make_view((0, 6, 40), scenes.quat_yaw_pitch(0.0, -12.0), 75.0, 0.1, 500.0, w, h)
    gbuf = util.ground_gbuffer(rng, oracle, view, w, h)
    gbuf[1][100, 200, 0] = np.nan                                       # unclean pixel inside tile (12, 6)
    gbuf[2][40, 17, 3] = 1e9
    lights, att, bias = util.lights_for_shading(rng, n, view)
    lights["type"] = np.where(rng.random(n) < 0.5, scenes.LIGHT_POINT, scenes.LIGHT_SPOT)
    lights["type"][:4] = (scenes.LIGHT_AMBIENT, scenes.LIGHT_DIRECTIONAL, scenes.LIGHT_POINT, scenes.LIGHT_SPOT)
    lights["position"][:, 0] = rng.uniform(-120, 120, n)
    lights["position"][:, 1] = rng.uniform(0.5, 5.0, n)
    lights["position"][:, 2] = rng.uniform(-250, 35, n)
    lights["direction"][:, 1] = -np.abs(lights["direction"][:, 1]) - 0.5   # pointing down at the ground
    lights["range"] = rng.uniform(4.0, 25.0, n)
    lights["range"][2] = 400.0                                          # one light that reaches everything
    world = lib.World(0)
    world.add_lights(lights)
    for i in range(n):
        world.set_light_shading(i, att[i], float(bias[i]))
    blocks = np.stack([oracle.light_uniforms(l, x, float(y)) for l, x, y in zip(lights, att, bias)])
    vb = oracle.view_uniforms(view)
    ids = np.arange(n, dtype=np.uint32)
    want = oracle.shade_deferred(*gbuf, vb, blocks, lights["type"])
    got = _shade(torch, world, w, h, gbuf, vb, ids)
    assert np.array_equal(np.isnan(got), np.isnan(want)) and np.isnan(want[100, 200, 0])
    _close(got, want)
    lit = (want[..., :3] > 0).any(axis=2)
    assert 0.3 < lit.mean() < 1.0
    size = 32
    maps = [rng.uniform(0.9, 1.0, (size, size) if l["type"] == scenes.LIGHT_SPOT else (6, size, size)).astype(np.float32)
            if l["flags"] & 1 and l["type"] >= scenes.LIGHT_POINT else None for l in lights]
    want = oracle.shade_deferred(*gbuf, vb, blocks, lights["type"], maps, size)
    got = _shade(torch, world, w, h, gbuf, vb, ids, maps, size)
    assert np.array_equal(np.isnan(got), np.isnan(want))
    _close(got, want)
    world.close()


def test_light_ids_that_name_no_light_are_skipped(oracle):
    import torch
    rng = np.random.default_rng(99)
    w, h = 64, 40
    view, gbuf, lights, blocks, world = _scene(rng, oracle, w, h, 20)
    vb = oracle.view_uniforms(view)
    good = np.array([4, 0, 19, 7], dtype=np.uint32)
    mixed = np.array([4, 20, 0, 2**32 - 1, 19, 7, 12345], dtype=np.uint32)
    want = oracle.shade_deferred(*gbuf, vb, blocks[good], lights["type"][good])
    _close(_shade(torch, world, w, h, gbuf, vb, mixed), want)
    size = 8
    maps = [rng.uniform(0.0, 1.0, (size, size) if l["type"] == scenes.LIGHT_SPOT else (6, size, size)).astype(np.float32)
            if (l["flags"] & 1 and l["type"] >= scenes.LIGHT_POINT) else None for l in lights]
    want = oracle.shade_deferred(*gbuf, vb, blocks[good], lights["type"][good], [maps[i] for i in good], size)
    _close(_shade(torch, world, w, h, gbuf, vb, mixed, maps, size), want)
    world.close()
