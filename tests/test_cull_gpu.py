"""GPU parity: the CUDA path, called through the C ABI, against the CPU oracle on identical inputs.
Visible sets and lists must be identical (same ids, same order); transforms and bounds are compared
for exact value equality (the tolerance north_star allows is 1e-5 relative; we hold 0)."""
import numpy as np
import pytest

from orion_b200 import lib, scenes
import vis_testutil as util
from vis_testutil import same

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def oracle():
    from oracle import pyoracle
    return pyoracle.load("best")


def _world_static(aabb, flags=None):
    w = lib.World(0)
    first = w.add_entities(aabb, flags)
    w.set_world_aabbs(first, aabb)
    return w


@pytest.mark.parametrize("n", [0, 1, 3, 4, 5, 1023, 1024, 1025, 4097, 100_000])
def test_single_view_sizes(oracle, n):
    rng = np.random.default_rng(100 + n)
    aabb = util.random_boxes(rng, n, extent=300.0)
    flags = rng.integers(0, 2, n).astype(np.uint32)
    w = _world_static(aabb, flags)
    for v in util.random_views(rng, 3, extent=50.0):
        planes = oracle.view_planes(np.asarray(v))
        assert np.array_equal(w.cull_planes(planes), oracle.cull_aabb(aabb, planes))
        assert np.array_equal(w.cull_planes(planes, require=1), oracle.cull_aabb(aabb, planes, flags, 1))
        ents, _ = w.cull(np.asarray(v), flags=0)
        assert np.array_equal(ents, oracle.cull_aabb(aabb, planes))
    w.close()


def test_config2_one_million_static(oracle):
    s = scenes.config2()
    w = _world_static(s["aabb"], s["flags"])
    planes = oracle.view_planes(s["view"])
    exp = oracle.cull_aabb(s["aabb"], planes, threads=8)
    got, _ = w.cull(s["view"], flags=0)
    assert 0.02 * len(s["aabb"]) < len(exp) < 0.3 * len(s["aabb"])
    assert np.array_equal(got, exp)
    # repeated launches reuse the tile-status scratch through the epoch tag: results must not change
    for _ in range(5):
        assert np.array_equal(w.cull_planes(planes), exp)
    w.close()


def test_edge_boxes_and_planes(oracle):
    rng = np.random.default_rng(31)
    aabb = np.concatenate([util.edge_boxes(), util.random_boxes(rng, 2000, extent=30.0)])
    w = _world_static(aabb)
    for v in util.random_views(rng, 10, extent=20.0):
        planes = oracle.view_planes(np.asarray(v))
        assert np.array_equal(w.cull_planes(planes), oracle.cull_aabb(aabb, planes))
    axis = np.array([[1, 0, 0, -10], [-1, 0, 0, -10], [0, -1, 0, -10], [0, 1, 0, -10], [0, 0, -1, -10],
                     [0, 0, 1, -10]], dtype=np.float32)
    touch = np.array([[10, 0, 0, 11, 1, 1], [-11, 0, 0, -10, 1, 1], [10.000001, 0, 0, 11, 1, 1],
                      [-12, 0, 0, -10.000001, 1, 1], [0, 0, 0, 1, 1, 1]], dtype=np.float32)
    w2 = _world_static(touch)
    assert np.array_equal(w2.cull_planes(axis), oracle.cull_aabb(touch, axis))
    neg0 = axis.copy()
    neg0[:, :3] *= -0.0
    assert np.array_equal(w2.cull_planes(neg0), oracle.cull_aabb(touch, neg0))
    w.close()
    w2.close()


def test_remove_keeps_order_and_flags_update(oracle):
    rng = np.random.default_rng(32)
    n = 20_000
    aabb = util.random_boxes(rng, n, extent=200.0)
    flags = rng.integers(0, 2, n).astype(np.uint32)
    w = _world_static(aabb, flags)
    v = scenes.make_view((0, 0, 250), (1, 0, 0, 0), 90.0, 0.1, 1000.0, 1600, 900)
    planes = oracle.view_planes(v)
    full = oracle.cull_aabb(aabb, planes)
    removed = rng.choice(n, 3000, replace=False).astype(np.uint32)
    w.remove_entities(removed)
    keep = np.ones(n, dtype=bool)
    keep[removed] = False
    assert np.array_equal(w.cull_planes(planes), full[keep[full]])
    flags2 = flags.copy()
    flip = rng.choice(np.flatnonzero(keep), 2000, replace=False).astype(np.uint32)
    flags2[flip] ^= 1
    w.set_flags(flip, flags2[flip])
    exp = oracle.cull_aabb(aabb, planes, flags2, 1)
    assert np.array_equal(w.cull_planes(planes, require=1), exp[keep[exp]])
    # entities appended later come after every earlier id
    more = util.random_boxes(rng, 500, extent=200.0)
    first = w.add_entities(more)
    w.set_world_aabbs(first, more)
    allb = np.concatenate([aabb, more])
    keep2 = np.concatenate([keep, np.ones(500, dtype=bool)])
    e2 = oracle.cull_aabb(allb, planes)
    assert np.array_equal(w.cull_planes(planes), e2[keep2[e2]])
    with pytest.raises(lib.OvisError):
        w.remove_entities(removed[:1])      # already gone
    w.close()


def test_transform_tree_and_bounds(oracle):
    rng = np.random.default_rng(33)
    n = 30_000
    parent = np.empty(n, dtype=np.int32)
    parent[0] = -1
    parent[1:] = (rng.random(n - 1) * np.arange(1, n)).astype(np.int32)      # arbitrary earlier parent: ragged levels
    parent[1:40] = np.arange(0, 39)                                           # deep chain
    local = util.random_trs(rng, n)
    local[:, 7:10] = rng.uniform(0.5, 1.5, (n, 3))                            # keep magnitudes finite down the chain
    boxes = util.random_boxes(rng, n, extent=2.0, hmax=2.0)
    flags = rng.integers(0, 2, n).astype(np.uint32)
    w = lib.World(0)
    e0 = w.add_entities(boxes, flags)
    n0 = w.add_nodes(parent, local)
    w.bind_entities(np.arange(n0, n0 + n), np.arange(e0, e0 + n))
    wt = oracle.world_trs(parent, local)
    assert same(w.node_world_trs(), wt)
    wb = oracle.world_aabb(wt, boxes)
    assert same(w.world_aabbs(), wb)
    v = scenes.make_view((0, 0, 150), (1, 0, 0, 0), 90.0, 0.1, 1000.0, 1600, 900)
    planes = oracle.view_planes(v)
    assert np.array_equal(w.cull_planes(planes), oracle.cull_aabb(wb, planes))
    # move some nodes: everything below must follow
    moved = rng.choice(n, 500, replace=False).astype(np.uint32)
    local[moved] = util.random_trs(rng, 500)
    local[moved, 7:10] = rng.uniform(0.5, 1.5, (500, 3))
    w.set_node_local_trs(moved, local[moved])
    wt = oracle.world_trs(parent, local)
    wb = oracle.world_aabb(wt, boxes)
    assert same(w.node_world_trs(), wt)
    assert same(w.world_aabbs(), wb)
    assert np.array_equal(w.cull_planes(planes), oracle.cull_aabb(wb, planes))
    w.close()


def test_two_entities_on_one_node(oracle):
    """A MeshRenderer owns one RenderEntity per sub-mesh, all following the same Entity."""
    rng = np.random.default_rng(38)
    parent = np.array([-1, 0, 1], dtype=np.int32)
    local = util.random_trs(rng, 3)
    boxes = util.random_boxes(rng, 5, extent=2.0)
    w = lib.World(0)
    w.add_entities(boxes)
    w.add_nodes(parent, local)
    w.bind_entities([2, 2, 2, 1, 0], [0, 1, 2, 3, 4])
    wt = oracle.world_trs(parent, local)
    exp = oracle.world_aabb(wt[[2, 2, 2, 1, 0]], boxes)
    assert same(w.world_aabbs(), exp)
    w.close()


def test_direct_transforms_and_edge_cases(oracle):
    """RenderEntity::setTransform path incl. skybox box, negative scale, NaN / Inf transforms."""
    rng = np.random.default_rng(34)
    n = 5000
    trs = util.random_trs(rng, n)
    boxes = util.random_boxes(rng, n, extent=2.0, hmax=2.0)
    e = util.edge_boxes()
    boxes[:len(e)] = e
    trs[:len(e)] = scenes.IDENTITY_TRS
    trs[20] = scenes.IDENTITY_TRS
    boxes[20] = e[0]
    trs[20, 0:3] = (5, 6, 7)                 # translated skybox box -> Inf/NaN bounds
    trs[21, 0] = np.nan
    trs[22, 3] = np.inf
    trs[23, 7:10] = 0
    w = lib.World(0)
    w.add_entities(boxes)
    # before any transform: world box = local box under the identity transform
    assert same(w.world_aabbs(), oracle.world_aabb(np.tile(scenes.IDENTITY_TRS, (n, 1)), boxes))
    w.set_transforms(np.arange(n), trs)
    wb = oracle.world_aabb(trs, boxes)
    assert same(w.world_aabbs(), wb)
    for v in util.random_views(rng, 5, extent=50.0):
        planes = oracle.view_planes(np.asarray(v))
        assert np.array_equal(w.cull_planes(planes), oracle.cull_aabb(wb, planes))
    # second update of a subset, and a changed local box
    sub = rng.choice(n, 700, replace=False).astype(np.uint32)
    trs[sub] = util.random_trs(rng, 700)
    w.set_transforms(sub, trs[sub])
    boxes[sub[:50]] = util.random_boxes(rng, 50, extent=1.0)
    w.set_local_aabbs(sub[:50], boxes[sub[:50]])
    assert same(w.world_aabbs(), oracle.world_aabb(trs, boxes))
    w.close()


def test_lights_and_full_cull(oracle):
    rng = np.random.default_rng(35)
    aabb = util.random_boxes(rng, 10_000, extent=100.0)
    lights = util.random_lights(rng, 700, extent=120.0)
    w = _world_static(aabb)
    w.add_lights(lights)
    for v in util.random_views(rng, 8, extent=60.0):
        v = np.asarray(v)
        planes = oracle.view_planes(v)
        ents, ls = w.cull(v)
        assert np.array_equal(ents, oracle.cull_aabb(aabb, planes))
        assert np.array_equal(ls, oracle.light_cull(lights, planes))
    lights2 = lights.copy()
    lights2[5] = util.random_lights(rng, 1)[0]
    w.update_light(5, lights2[5])
    w.remove_light(7)
    v = np.asarray(util.random_views(rng, 1, extent=60.0)[0])
    planes = oracle.view_planes(v)
    exp = oracle.light_cull(lights2, planes)
    _, ls = w.cull(v)
    assert np.array_equal(ls, exp[exp != 7])
    w.close()


def test_shadow_lists_and_cubes_scene(oracle):
    s = scenes.cubes_scene()
    w = lib.World(0)
    w.add_entities(s["local_aabb"], s["flags"])
    w.set_transforms(np.arange(4), s["trs"])
    w.add_lights(s["lights"])
    wb = oracle.world_aabb(s["trs"], s["local_aabb"])
    assert same(w.world_aabbs(), wb)
    planes = oracle.view_planes(s["view"])
    ents, ls = w.cull(s["view"])
    assert np.array_equal(ents, oracle.cull_aabb(wb, planes))
    assert np.array_equal(ls, oracle.light_cull(s["lights"], planes))
    got = w.shadow_lists(s["view"])
    exp = oracle.shadow_lists(wb, s["flags"], s["lights"], planes)
    for x, y in zip(got, exp):
        assert np.array_equal(x, y)
    w.close()
    rng = np.random.default_rng(36)
    boxes = util.random_boxes(rng, 20_000, extent=60.0, hmax=3.0)
    flags = rng.integers(0, 2, len(boxes)).astype(np.uint32)
    lights = util.random_lights(rng, 40, extent=40.0)
    w = _world_static(boxes, flags)
    w.add_lights(lights)
    cam = scenes.make_view((0, 0, 70), (1, 0, 0, 0), 90.0, 0.1, 300.0, 1600, 900)
    got = w.shadow_lists(cam)
    exp = oracle.shadow_lists(boxes, flags, lights, oracle.view_planes(cam))
    assert len(exp[0]) > 6
    for x, y in zip(got, exp):
        assert np.array_equal(x, y)
    w.close()


def test_multi_view_csr(oracle):
    rng = np.random.default_rng(37)
    boxes = util.random_boxes(rng, 50_000, extent=200.0)
    flags = rng.integers(0, 2, len(boxes)).astype(np.uint32)
    w = _world_static(boxes, flags)
    views = util.random_views(rng, 9, extent=80.0)
    planes = np.stack([oracle.view_planes(np.asarray(v)).reshape(24) for v in views])
    require = np.array([0, 1, 0, 1, 1, 0, 0, 1, 0], dtype=np.uint32)
    off, d_off, d_ent = w.cull_views_device(planes, require)
    total = int(off[-1])
    ent = w.read_device(d_ent, total)
    assert np.array_equal(w.read_device(d_off, len(views) + 1, np.int64), off)
    for i in range(len(views)):
        exp = oracle.cull_aabb(boxes, planes[i], flags, int(require[i]))
        assert np.array_equal(ent[off[i]:off[i + 1]], exp)
    w.close()


def test_shadow_lists_prefilter_is_conservative(oracle, monkeypatch):
    """(OVIS_MV_DENSE=0: a world this small would otherwise skip the pre-filter and test every view densely.)
    The cube pre-filter in front of the exact shadow-view tests must never drop a pair the reference keeps:
    giant boxes (the P-vertex test accepts them far outside the cube), skybox / infinite / NaN boxes, inverted
    boxes, boxes hugging cube faces and corners, lights at large coordinates. More than 1024 views forces
    several passes; removed entities and non-casters must stay out."""
    monkeypatch.setenv("OVIS_MV_DENSE", "0")
    rng = np.random.default_rng(39)
    n = 30_000
    boxes = util.random_boxes(rng, n, extent=300.0, hmax=4.0)
    e = util.edge_boxes()
    boxes[:len(e)] = e
    big = rng.choice(n, 300, replace=False)
    boxes[big] = util.random_boxes(rng, 300, extent=300.0, hmax=500.0)          # giant boxes
    boxes[100] = [-1e6, -1e6, -1e6, 1e6, 1e6, 1e6]
    boxes[101] = [5, 5, 5, -5, -5, -5]                                            # inverted
    flags = (rng.random(n) < 0.7).astype(np.uint32)
    flags[:len(e)] = 1
    lights = scenes.point_lights(260, 77, 280.0, 5.0, 120.0)
    lights["flags"][::7] = 0                                                       # some cast no shadows
    spot = util.random_lights(rng, 30, extent=250.0)
    spot["type"] = scenes.LIGHT_SPOT
    spot["flags"] = 1
    spot["intensity"] = 1.0
    spot["colour"] = 1.0
    lights = np.concatenate([lights, spot])
    # boxes hugging the first light's cube faces / corners from outside and inside
    p, r = lights["position"][0], lights["range"][0]
    k = 200
    for sx in (-1, 1):
        for d in (0.999, 1.0, 1.001, 1.02):
            c = p + np.array([sx * r * d, r * d * 0.99, -r * d])
            boxes[k] = np.concatenate([c - 0.01, c + 0.01])
            k += 1
    w = _world_static(boxes, flags)
    w.add_lights(lights)
    removed = rng.choice(np.arange(300, n), 2000, replace=False).astype(np.uint32)
    w.remove_entities(removed)
    cam = scenes.make_view((0, 0, 900), (1, 0, 0, 0), 90.0, 0.1, 3000.0, 1600, 900)   # sees every light
    got = w.shadow_lists(cam)
    keep = np.ones(n, dtype=bool)
    keep[removed] = False
    oflags = flags * keep                                                             # a removed entity casts nothing
    exp = oracle.shadow_lists(boxes, oflags, lights, oracle.view_planes(cam))
    assert len(exp[0]) > 1024                                                         # several passes
    for x, y in zip(got, exp):
        assert np.array_equal(x, y)
    w.close()


def test_multi_view_many_views_and_empty(oracle):
    rng = np.random.default_rng(40)
    boxes = util.random_boxes(rng, 7001, extent=150.0)
    flags = rng.integers(0, 2, len(boxes)).astype(np.uint32)
    w = _world_static(boxes, flags)
    views = util.random_views(rng, 40, extent=100.0)
    planes = np.stack([oracle.view_planes(np.asarray(v)).reshape(24) for v in views])
    require = rng.integers(0, 2, 40).astype(np.uint32)
    off, d_off, d_ent = w.cull_views_device(planes, require)
    ent = w.read_device(d_ent, int(off[-1]))
    for i in range(len(views)):
        assert np.array_equal(ent[off[i]:off[i + 1]], oracle.cull_aabb(boxes, planes[i], flags, int(require[i])))
    # a second call reuses every scratch buffer
    off2, _, d_ent2 = w.cull_views_device(planes[::-1].copy(), require[::-1].copy())
    ent2 = w.read_device(d_ent2, int(off2[-1]))
    for i in range(len(views)):
        j = len(views) - 1 - i
        assert np.array_equal(ent2[off2[i]:off2[i + 1]], ent[off[j]:off[j + 1]])
    w.close()
    w0 = lib.World(0)
    off0, _, _ = w0.cull_views_device(planes[:3])
    assert np.array_equal(off0, np.zeros(4, dtype=np.int64))
    w0.close()


def test_partial_updates_touch_only_moved_subtrees(oracle):
    """Only nodes below a moved node are recomputed (Entity::transformed runs from the moved entity down); the rest must
    keep their values, including second entities on a node and a changed local box of a node-bound entity."""
    rng = np.random.default_rng(41)
    n = 20_000
    parent = np.empty(n, dtype=np.int32)
    parent[:50] = -1
    parent[50:] = (rng.random(n - 50) * np.arange(50, n)).astype(np.int32)
    local = util.random_trs(rng, n)
    local[:, 7:10] = rng.uniform(0.6, 1.4, (n, 3))
    boxes = util.random_boxes(rng, n + 300, extent=2.0, hmax=2.0)
    w = lib.World(0)
    w.add_entities(boxes)
    w.add_nodes(parent, local)
    ent_node = np.concatenate([np.arange(n), rng.integers(0, n, 300)]).astype(np.uint32)   # 300 second entities
    w.bind_entities(ent_node, np.arange(n + 300))
    for step in range(4):
        moved = rng.choice(n, [1, 30, 400, 15000][step], replace=False).astype(np.uint32)   # the last one flips to "all"
        local[moved] = util.random_trs(rng, len(moved))
        local[moved, 7:10] = rng.uniform(0.6, 1.4, (len(moved), 3))
        w.set_node_local_trs(moved, local[moved])
        if step == 1:
            changed = rng.choice(n + 300, 40, replace=False).astype(np.uint32)
            boxes[changed] = util.random_boxes(rng, 40, extent=1.0)
            w.set_local_aabbs(changed, boxes[changed])
        wt = oracle.world_trs(parent, local)
        assert same(w.node_world_trs(), wt)
        assert same(w.world_aabbs(), oracle.world_aabb(wt[ent_node], boxes))
    w.close()


@pytest.mark.parametrize("n", [0, 3, 4097, 300_000, 1_200_000])
def test_pinned_output_buffer_gets_the_list_without_a_copy(oracle, n):
    """A page-locked output buffer (ovis_host_alloc) is written by the gather kernel itself; a pageable one goes through
    the copy path. Both must hold the same list; a buffer that is too small reports the full count and is not overrun."""
    rng = np.random.default_rng(900 + n)
    boxes = util.random_boxes(rng, n, extent=150.0)
    flags = rng.integers(0, 2, n).astype(np.uint32)
    w = _world_static(boxes, flags)
    w.add_lights(util.random_lights(rng, 40, extent=60.0))
    pinned = lib.PinnedBuffer(n + 8)
    for v in util.random_views(rng, 4, extent=60.0):
        planes = lib.view_planes(v)
        for require in (0, 1):
            expected = oracle.cull_aabb(boxes, planes, flags, require)
            pinned.array[:] = 0xdeadbeef
            got = w.cull_planes(planes, require=require, out=pinned.array)
            assert np.array_equal(got, expected)
            assert (pinned.array[len(expected):] == 0xdeadbeef).all()
            assert np.array_equal(w.cull_planes(planes, require=require), expected)          # pageable
        ents, ls = w.cull(v, out_entities=pinned.array)                                      # with the light cull beside it
        e2, l2 = w.cull(v)
        assert np.array_equal(ents, e2) and np.array_equal(ls, l2)
        assert np.array_equal(ents, oracle.cull_aabb(boxes, planes))
    # too small: error + the count that is needed, nothing written past the capacity handed in
    full = oracle.cull_aabb(boxes, lib.view_planes(scenes.make_view((0, 0, 400), (1, 0, 0, 0), 100.0, 0.1, 2000.0, 1600, 900)))
    if len(full) > 8:
        planes = lib.view_planes(scenes.make_view((0, 0, 400), (1, 0, 0, 0), 100.0, 0.1, 2000.0, 1600, 900))
        pinned.array[:] = 0xdeadbeef
        cap = len(full) // 2
        ne = lib.C.c_int64()
        rc = w.lib.ovis_cull_planes(w.h, lib._p(np.ascontiguousarray(planes, dtype=np.float32)), lib.C.c_uint32(0),
                                    lib.C.c_void_p(pinned.ptr.value), lib.C.c_int64(cap), lib.C.byref(ne))
        assert rc == lib.OVIS_ERR_CAPACITY and ne.value == len(full)
        assert np.array_equal(pinned.array[:cap], full[:cap]) and (pinned.array[cap:] == 0xdeadbeef).all()
        assert np.array_equal(w.cull_planes(planes, out=pinned.array), full)                 # and the next call is fine
    w.close()
    pinned.free()


def test_read_segments_host_pinned_and_pageable():
    """The reader side of the sharded exchange on one GPU: segments and flags prepared locally."""
    import torch
    rng = np.random.default_rng(77)
    n_seg, words = 5, 4001
    seg = rng.integers(0, 2**31, (n_seg, words)).astype(np.uint32)
    counts = np.array([0, 1, 1234, 4000, 9000], dtype=np.uint32)     # the last one overflowed its segment: clipped copy
    seg[:, 0] = counts
    d_seg = torch.from_numpy(seg.view(np.int32)).cuda()
    d_flags = torch.full((n_seg,), 7, dtype=torch.int32, device="cuda")
    w = lib.World(0)
    pinned = lib.PinnedBuffer(n_seg * words)
    pinned.array[:] = 0xdeadbeef
    w.read_segments_host(d_seg.data_ptr(), words, n_seg, pinned.array, d_flags.data_ptr(), 7)
    host = pinned.array.reshape(n_seg, words)
    for r in range(n_seg):
        k = 1 + min(int(counts[r]), words - 1)
        assert np.array_equal(host[r, :k], seg[r, :k])
        assert (host[r, k:] == 0xdeadbeef).all()                      # exactly count ids are moved
    pageable = np.zeros(n_seg * words, dtype=np.uint32)
    w.read_segments_host(d_seg.data_ptr(), words, n_seg, pageable, d_flags.data_ptr(), 6)    # flags already past 6
    assert np.array_equal(pageable.reshape(n_seg, words), seg)
    w.read_segments_host(d_seg.data_ptr(), words, n_seg, pageable)                          # no flags at all
    assert np.array_equal(pageable.reshape(n_seg, words), seg)
    assert w.exchange_error() == 0
    w.close()
    pinned.free()


@pytest.mark.parametrize("variant", list(range(6)))
def test_every_test_kernel_variant_is_exact(oracle, variant, monkeypatch):
    """OVIS_CULL_VARIANT selects the form of the entity test kernel (register-fed with tiles of 128 or 256 entities, or
    the TMA ring with 2-4 stages of 1024 entities). Whichever runs, the list is the reference's: edge-case boxes, ragged
    sizes, removed entities and the castsShadow filter included."""
    monkeypatch.setenv("OVIS_CULL_VARIANT", str(variant))
    monkeypatch.setenv("OVIS_CULL_FUSED", "0")                 # small worlds would otherwise take the fused kernel
    rng = np.random.default_rng(4000 + variant)
    n = 70_001
    boxes = np.concatenate([util.edge_boxes(), util.random_boxes(rng, n - len(util.edge_boxes()), extent=150.0)])
    flags = rng.integers(0, 2, n).astype(np.uint32)
    w = _world_static(boxes, flags)
    views = list(util.random_views(rng, 3, extent=60.0)) + [scenes.make_view((0, 0, 300), (1, 0, 0, 0), 100.0, 0.1, 2000.0, 1600, 900)]
    for v in views:
        planes = lib.view_planes(v)
        assert np.array_equal(w.cull_planes(planes), oracle.cull_aabb(boxes, planes))
        assert np.array_equal(w.cull_planes(planes, require=1), oracle.cull_aabb(boxes, planes, flags, 1))
    gone = rng.choice(n, 5000, replace=False).astype(np.uint32)
    w.remove_entities(gone)
    alive = np.ones(n, dtype=bool)
    alive[gone] = False
    planes = lib.view_planes(views[-1])
    expected = oracle.cull_aabb(boxes, planes)
    assert np.array_equal(w.cull_planes(planes), expected[alive[expected]])
    w.close()


def test_light_updates_upload_only_what_changed(oracle):
    """Lights are re-sent to the device by dirty slot range: single updates, growth past the allocated capacity, shading
    changes and removals must all leave the device copies (cull records and LightUniforms blocks) current."""
    rng = np.random.default_rng(61)
    lights = util.random_lights(rng, 300, extent=80.0)
    att = np.zeros((300, 3), dtype=np.float32)
    att[lights["type"] == scenes.LIGHT_POINT] = (1.0, 0.045, 0.0075)
    att[lights["type"] == scenes.LIGHT_SPOT] = (1.0, 0.09, 0.032)
    bias = np.full(300, -0.001, dtype=np.float32)
    w = _world_static(util.random_boxes(rng, 1000))
    w.add_lights(lights)
    view = scenes.make_view((0, 0, 120), (1, 0, 0, 0), 80.0, 0.1, 500.0, 1280, 720)
    planes = lib.view_planes(view)
    alive = np.ones(300, dtype=bool)

    def check():
        _, ls = w.cull(view)
        want = oracle.light_cull(lights, planes)
        assert np.array_equal(ls, want[alive[want]])
        ids = np.flatnonzero(alive).astype(np.uint32)
        blocks = util.gpu_light_uniforms(w, ids)
        exp = np.stack([oracle.light_uniforms(lights[i], att[i], float(bias[i])) for i in ids])
        assert blocks.tobytes() == exp.tobytes()

    check()
    lights[150] = util.random_lights(rng, 1, extent=80.0)[0]              # one light in the middle
    lights[150]["type"] = scenes.LIGHT_POINT
    att[150] = (1.0, 0.045, 0.0075)
    w.update_light(150, lights[150])
    w.set_light_shading(150, att[150], -0.001)
    check()
    more = util.random_lights(rng, 500, extent=80.0)                       # growth: new device buffers
    more_att = np.zeros((500, 3), dtype=np.float32)
    more_att[more["type"] == scenes.LIGHT_POINT] = (1.0, 0.045, 0.0075)
    more_att[more["type"] == scenes.LIGHT_SPOT] = (1.0, 0.09, 0.032)
    assert w.add_lights(more) == 300
    lights = np.concatenate([lights, more])
    att = np.concatenate([att, more_att])
    bias = np.concatenate([bias, np.full(500, -0.001, dtype=np.float32)])
    alive = np.ones(800, dtype=bool)
    check()
    for i in (10, 700):                                                    # two far-apart updates in one frame
        lights[i] = util.random_lights(rng, 1, extent=80.0)[0]
        w.update_light(i, lights[i])
        att[i] = rng.uniform(0.1, 1.0, 3)
        bias[i] = -0.004
        w.set_light_shading(i, att[i], float(bias[i]))
    w.remove_light(5)
    alive[5] = False
    check()
    w.close()
