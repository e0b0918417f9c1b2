"""GPU parity for the consumers of the visible lists, through the C ABI:
draw-list expansion (DeferredRenderPipeline::prepareEntities and the shadow-caster loop of prepareLights, with
DrawList::add) and EntityUniforms packing (RenderEntity::setTransform). Draw lists are integer work: identical
(entity, pass) pairs in identical order. Uniforms are floats computed in the reference's operation order: compared
for exact value equality (tolerance 0)."""
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


def _static_world(rng, n, extent=120.0):
    boxes = util.random_boxes(rng, n, extent=extent)
    flags = rng.integers(0, 2, n).astype(np.uint32)
    pc = util.random_pass_counts(rng, n)
    w = lib.World(0)
    first = w.add_entities(boxes, flags)
    w.set_world_aabbs(first, boxes)
    if n:
        w.set_pass_counts(np.arange(n), pc & 0xff, (pc >> 8) & 0xff, pc >> 16)
    return w, boxes, flags, pc


@pytest.mark.parametrize("n", [0, 1, 5, 1023, 1025, 4099, 200_000])
def test_draw_lists_from_cull_result(oracle, n):
    rng = np.random.default_rng(100 + n)
    w, boxes, flags, pc = _static_world(rng, n)
    views = [scenes.make_view((0, 0, 150), (1, 0, 0, 0), 90.0, 0.1, 1000.0, 1600, 900),
             scenes.make_view((0, 0, 500), scenes.quat_yaw_pitch(180.0, 0.0), 60.0, 0.1, 100.0, 800, 600)]   # sees nothing
    for v in views:
        planes = lib.view_planes(v)
        visible = w.cull_planes(planes)
        expected = oracle.cull_aabb(boxes, planes)
        assert np.array_equal(visible, expected)
        for mode in (lib.DRAWS_MAIN, lib.DRAWS_SHADOW):
            e0, e1 = oracle.draw_lists(expected, pc, flags, mode)
            d0, d1, counts = util.gpu_draw_lists(w, mode)                 # straight from the cull's device result
            assert counts == (len(e0), len(e1))
            assert np.array_equal(d0, e0) and np.array_equal(d1, e1)
            h0, h1, _ = util.gpu_draw_lists(w, mode, visible=expected)    # same list handed in by the caller
            assert np.array_equal(h0, e0) and np.array_equal(h1, e1)
    w.close()


def test_draw_lists_internal_counts_overflow_and_flag_changes(oracle):
    rng = np.random.default_rng(7)
    n = 20_000
    w, boxes, flags, pc = _static_world(rng, n)
    planes = lib.view_planes(scenes.make_view((0, 0, 150), (1, 0, 0, 0), 90.0, 0.1, 1000.0, 1600, 900))
    visible = w.cull_planes(planes)
    e0, e1 = oracle.draw_lists(visible, pc, flags, 0)
    assert len(e0) > 3000 and len(e1) > 500
    # capacity smaller than the list: the prefix is written, the counts still say how many there are
    d0, d1, counts = util.gpu_draw_lists(w, lib.DRAWS_MAIN, cap=1000)
    assert counts == (len(e0), len(e1))
    assert np.array_equal(d0, e0[:1000]) and np.array_equal(d1, e1[:1000])
    # counts into the library's own pair
    import torch
    buf = torch.zeros((len(e0), 2), dtype=torch.int32, device="cuda")
    w.build_draw_lists_device(lib.DRAWS_MAIN, buf.data_ptr(), len(e0))   # list 1 dropped entirely (capacity 0)
    assert w.draw_counts() == (len(e0), len(e1))
    assert np.array_equal(buf.cpu().numpy().view(np.uint32), e0)
    # castsShadow changes and removals are seen by the shadow-caster expansion
    flip = rng.choice(n, 3000, replace=False).astype(np.uint32)
    flags[flip] ^= 1
    w.set_flags(flip, flags[flip])
    gone = rng.choice(n, 2000, replace=False).astype(np.uint32)
    w.remove_entities(gone)
    alive = np.ones(n, dtype=bool)
    alive[gone] = False
    visible = w.cull_planes(planes)
    expected = oracle.cull_aabb(boxes, planes)
    expected = expected[alive[expected]]
    assert np.array_equal(visible, expected)
    s0, s1 = oracle.draw_lists(expected, pc, flags, 1)
    g0, g1, _ = util.gpu_draw_lists(w, lib.DRAWS_SHADOW)
    assert np.array_equal(g0, s0) and len(g1) == 0 and len(s1) == 0
    # new pass counts for some entities (material swap)
    swap = rng.choice(np.flatnonzero(alive), 5000, replace=False).astype(np.uint32)
    pc[swap] = util.random_pass_counts(rng, len(swap))
    w.set_pass_counts(swap, pc[swap] & 0xff, (pc[swap] >> 8) & 0xff, pc[swap] >> 16)
    e0, e1 = oracle.draw_lists(expected, pc, flags, 0)
    d0, d1, _ = util.gpu_draw_lists(w, lib.DRAWS_MAIN)
    assert np.array_equal(d0, e0) and np.array_equal(d1, e1)
    w.close()


def test_shadow_draw_lists_per_view_slice_and_id_base(oracle):
    """The shadow-caster loop runs once per shadow view: expand each view's slice of the multi-view CSR. With an id
    base (sharded world) the ids in the lists - and in the draws - are global."""
    import torch
    rng = np.random.default_rng(19)
    n = 30_000
    base = 5_000_000
    w, boxes, flags, pc = _static_world(rng, n)
    w.set_id_base(base)
    views = util.random_views(rng, 9, extent=60.0)
    planes = np.stack([lib.view_planes(v) for v in views])
    off, d_off, d_ent = w.cull_views_device(planes, require=np.ones(len(views), dtype=np.uint32))
    cap = 4 * n
    draws = torch.zeros((cap, 2), dtype=torch.int32, device="cuda")
    cnt = torch.zeros(2, dtype=torch.int32, device="cuda")
    total = 0
    for i in range(len(views)):
        m = int(off[i + 1] - off[i])
        expected_list = oracle.cull_aabb(boxes, planes[i], flags, 1)
        assert m == len(expected_list)
        vcnt = torch.tensor([m], dtype=torch.int32, device="cuda")
        w.build_draw_lists_device(lib.DRAWS_SHADOW, draws.data_ptr(), cap, d_visible=d_ent + 4 * int(off[i]),
                                  d_visible_count=vcnt.data_ptr(), max_visible=m, d_draw_counts=cnt.data_ptr())
        torch.cuda.synchronize()
        e0, _ = oracle.draw_lists(expected_list, pc, flags, 1)
        k = int(cnt[0].item())
        assert k == len(e0) and int(cnt[1].item()) == 0
        got = draws[:k].cpu().numpy().view(np.uint32)
        assert np.array_equal(got[:, 0], e0[:, 0] + np.uint32(base)) and np.array_equal(got[:, 1], e0[:, 1])
        total += k
    assert total > 1000
    w.close()


def test_entity_uniforms_follow_the_transform_tree(oracle):
    rng = np.random.default_rng(23)
    n = 20_000
    parent = np.empty(n, dtype=np.int32)
    parent[0] = -1
    parent[1:] = (rng.random(n - 1) * np.arange(1, n)).astype(np.int32)
    local = util.random_trs(rng, n)
    local[:, 7:10] = rng.uniform(0.5, 1.5, (n, 3))
    n_ent = n + 500                                           # 300 second entities on shared nodes + 200 never moved
    boxes = util.random_boxes(rng, n_ent, extent=2.0, hmax=2.0)
    w = lib.World(0)
    e0 = w.add_entities(boxes)
    n0 = w.add_nodes(parent, local)
    ent_node = np.full(n_ent, -1, dtype=np.int64)
    ent_node[:n] = np.arange(n)
    ent_node[n:n + 300] = rng.choice(n, 300, replace=False)
    bound = np.flatnonzero(ent_node >= 0)
    w.bind_entities(n0 + ent_node[bound], e0 + bound)
    static = np.arange(n + 300, n_ent)
    w.set_world_aabbs_by_id(static, util.random_boxes(rng, len(static), extent=50.0))

    def expected_uniforms(visible, local_trs):
        wt = oracle.world_trs(parent, local_trs)
        identity = np.array([0, 0, 0, 1, 0, 0, 0, 1, 1, 1], dtype=np.float32)
        trs = np.where((ent_node[visible] >= 0)[:, None], wt[np.maximum(ent_node[visible], 0)], identity)
        return oracle.entity_uniforms(trs)

    planes = lib.view_planes(scenes.make_view((0, 0, 150), (1, 0, 0, 0), 90.0, 0.1, 1000.0, 1600, 900))
    visible = w.cull_planes(planes)
    assert len(visible) > 1000 and (ent_node[visible] < 0).any() and (visible >= n).any()
    want = expected_uniforms(visible, local)
    got = util.gpu_entity_uniforms(w)[:len(visible)]          # from the cull's device result
    assert same(got, want)
    assert same(util.gpu_entity_uniforms(w, visible), want)   # same list handed in
    # capacity below the list length: a prefix
    assert same(util.gpu_entity_uniforms(w, visible, cap=777), want[:777])
    # empty list, and every entity
    assert util.gpu_entity_uniforms(w, np.empty(0, np.uint32), cap=4).shape == (4, 20)
    everyone = np.arange(n_ent, dtype=np.uint32)
    assert same(util.gpu_entity_uniforms(w, everyone), expected_uniforms(everyone, local))
    # moved nodes: the pending update is flushed before packing
    moved = rng.choice(n, 400, replace=False).astype(np.uint32)
    local[moved] = util.random_trs(rng, 400)
    local[moved, 7:10] = rng.uniform(0.5, 1.5, (400, 3))
    w.set_node_local_trs(moved, local[moved])
    assert same(util.gpu_entity_uniforms(w, everyone), expected_uniforms(everyone, local))
    w.close()


def test_entity_uniforms_direct_transforms_edge_values(oracle):
    """RenderEntity::setTransform with world transforms, including zero/negative scale and non-finite positions."""
    rng = np.random.default_rng(5)
    n = 3000
    trs = util.random_trs(rng, n)
    trs[0, 7:10] = 0.0
    trs[1, 0:3] = (np.inf, -np.inf, np.nan)
    trs[2, 3:7] = (0.0, 0.0, 0.0, 0.0)
    trs[3, 7:10] = (-1.0, 1e30, 1e-30)
    w = lib.World(0)
    w.add_entities(util.random_boxes(rng, n, extent=1.0))
    w.set_transforms(np.arange(n), trs)
    ids = rng.permutation(n).astype(np.uint32)               # any list order
    assert same(util.gpu_entity_uniforms(w, ids), oracle.entity_uniforms(trs[ids]))
    w.close()


def test_config2_one_million_draws_and_uniforms(oracle):
    """BASELINE configs[1] size: the visible list of 1M static boxes expanded and packed; besides the oracle comparison,
    a size-independent property: draws per list = sum of the pass counts over the visible list."""
    s = scenes.config2(seed=3)
    n = len(s["aabb"])
    rng = np.random.default_rng(3)
    pc = util.random_pass_counts(rng, n)
    w = lib.World(0)
    first = w.add_entities(s["aabb"], s["flags"])
    w.set_world_aabbs(first, s["aabb"])
    w.set_pass_counts(np.arange(n), pc & 0xff, (pc >> 8) & 0xff, pc >> 16)
    planes = lib.view_planes(s["view"])
    visible = w.cull_planes(planes)
    assert len(visible) > 10_000
    d0, d1, counts = util.gpu_draw_lists(w, lib.DRAWS_MAIN, cap=4 * n)
    deferred, basic = pc[visible] & 0xff, (pc[visible] >> 8) & 0xff
    assert counts == (int(deferred.sum()), int(basic[deferred == 0].sum()))
    e0, e1 = oracle.draw_lists(visible, pc, s["flags"], 0)
    assert np.array_equal(d0, e0) and np.array_equal(d1, e1)
    u = util.gpu_entity_uniforms(w)[:len(visible)]
    ident = oracle.entity_uniforms(np.tile(np.array([0, 0, 0, 1, 0, 0, 0, 1, 1, 1], dtype=np.float32), (1, 1)))
    assert same(u, np.repeat(ident, len(visible), axis=0))    # static geometry: Transform() of an entity never moved
    w.close()


def test_light_and_view_uniform_blocks(oracle):
    """LightUniforms of the visible lights (list order) gathered on the device, ViewUniforms uploaded per view."""
    rng = np.random.default_rng(41)
    n_lights = 700
    lights = util.random_lights(rng, n_lights, extent=80.0)
    w, boxes, flags, pc = _static_world(rng, 5000)
    first = w.add_lights(lights)
    assert first == 0
    # default shading parameters of the Light components, then explicit ones for half of the lights
    att = np.zeros((n_lights, 3), dtype=np.float32)
    att[lights["type"] == scenes.LIGHT_POINT] = (1.0, 0.045, 0.0075)
    att[lights["type"] == scenes.LIGHT_SPOT] = (1.0, 0.09, 0.032)
    bias = np.full(n_lights, -0.001, dtype=np.float32)
    for i in range(0, n_lights, 2):
        att[i] = rng.uniform(0, 1, 3)
        bias[i] = rng.uniform(-0.01, 0)
        w.set_light_shading(i, att[i], float(bias[i]))

    def expected(ids):
        if len(ids) == 0:
            return np.zeros((0, 208), dtype=np.uint8)
        return np.stack([oracle.light_uniforms(lights[i], att[i], float(bias[i])) for i in ids])

    for v in util.random_views(rng, 3, extent=40.0):
        _, ls = w.cull(v)
        assert np.array_equal(ls, oracle.light_cull(lights, lib.view_planes(v)))
        got = util.gpu_light_uniforms(w)                                   # from the cull's device light list
        assert got[:len(ls)].tobytes() == expected(ls).tobytes()
        assert (got[len(ls):] == 0x55).all()
    ids = rng.permutation(n_lights).astype(np.uint32)[:300]
    assert util.gpu_light_uniforms(w, ids).tobytes() == expected(ids).tobytes()     # any list the caller hands in
    assert util.gpu_light_uniforms(w, ids, cap=100).tobytes() == expected(ids[:100]).tobytes()
    assert util.gpu_light_uniforms(w, np.empty(0, np.uint32), cap=3).shape == (3, 208)
    # a light that changes is re-packed
    lights[ids[0]] = util.random_lights(rng, 1, extent=80.0)[0]
    w.update_light(int(ids[0]), lights[ids[0]])
    assert util.gpu_light_uniforms(w, ids).tobytes() == expected(ids).tobytes()
    # views
    views = np.concatenate([util.random_views(rng, 50), np.asarray(scenes.default_camera()).reshape(1)])
    pos = rng.integers(0, 2000, (len(views), 2)).astype(np.int32)
    got = util.gpu_view_uniforms(w, views, pos)
    want = np.stack([oracle.view_uniforms(v, p) for v, p in zip(views, pos)])
    assert got.tobytes() == want.tobytes()
    assert util.gpu_view_uniforms(w, views[:2]).tobytes() == np.stack([oracle.view_uniforms(v) for v in views[:2]]).tobytes()
    w.close()


def test_draw_lists_with_an_undersized_bound_say_so(oracle):
    """max_visible sizes the grids; when the device count is larger than the caller promised the lists cannot be
    complete, and the counts must say that (0xffffffff) instead of passing the first part off as the whole."""
    import torch
    rng = np.random.default_rng(77)
    n = 20_000
    boxes = util.random_boxes(rng, n, extent=50.0)
    w = lib.World(0)
    first = w.add_entities(boxes, np.ones(n, dtype=np.uint32))
    w.set_world_aabbs(first, boxes)
    pc = util.random_pass_counts(rng, n)
    w.set_pass_counts(np.arange(n, dtype=np.uint32), pc & 0xff, (pc >> 8) & 0xff, (pc >> 16) & 0xff)
    visible = np.arange(0, n, 2, dtype=np.uint32)                          # 10 000 entries
    vis = torch.from_numpy(visible.view(np.int32)).cuda()
    vcnt = torch.tensor([len(visible)], dtype=torch.int32, device="cuda")
    cap = 8 * n
    d0 = torch.zeros((cap, 2), dtype=torch.int32, device="cuda")
    d1 = torch.zeros((cap, 2), dtype=torch.int32, device="cuda")
    cnt = torch.zeros(2, dtype=torch.int32, device="cuda")
    for bound, complete in ((len(visible), True), (3000, False), (500, False)):
        w.build_draw_lists_device(lib.DRAWS_MAIN, d0.data_ptr(), cap, d1.data_ptr(), cap, vis.data_ptr(), vcnt.data_ptr(), bound,
                                  cnt.data_ptr())
        torch.cuda.synchronize()
        c = cnt.cpu().numpy().view(np.uint32)
        if complete:
            e0, e1 = oracle.draw_lists(visible, pc, np.ones(n, dtype=np.uint32), 0)
            assert (int(c[0]), int(c[1])) == (len(e0), len(e1))
            assert np.array_equal(d0[:len(e0)].cpu().numpy().view(np.uint32), e0)
        else:
            assert int(c[0]) == 0xffffffff and int(c[1]) == 0xffffffff
    w.close()


def test_ids_that_name_no_slot_index_nothing(oracle):
    """A caller's device list may hold ids that are no longer (or never were) slots of this world - stale after a
    compaction, another shard's. The consumers must not index their tables with them: no draws, zero blocks, no light."""
    import torch
    rng = np.random.default_rng(61)
    n = 5000
    w, boxes, flags, pc = _static_world(rng, n)
    lights = util.random_lights(rng, 30, extent=60.0)
    w.add_lights(lights)
    good = rng.choice(n, 700, replace=False).astype(np.uint32)
    wild = np.array([n, n + 1, 2**31, 2**32 - 1, 123456789], dtype=np.uint32)
    mixed = np.concatenate([good[:300], wild[:2], good[300:], wild[2:]])
    for mode in (lib.DRAWS_MAIN, lib.DRAWS_SHADOW):
        e0, e1 = oracle.draw_lists(good, pc, flags, mode)
        d0, d1, counts = util.gpu_draw_lists(w, mode, visible=mixed)
        assert counts == (len(e0), len(e1)) and np.array_equal(d0, e0) and np.array_equal(d1, e1)
    blocks = util.gpu_entity_uniforms(w, mixed)
    is_wild = np.isin(mixed, wild)
    assert np.all(blocks[is_wild] == 0.0)
    assert same(blocks[~is_wild], oracle.entity_uniforms(np.tile(np.array([0, 0, 0, 1, 0, 0, 0, 1, 1, 1], np.float32), (len(good), 1))))
    l_mixed = np.array([3, 30, 7, 2**32 - 1, 29], dtype=np.uint32)
    lb = util.gpu_light_uniforms(w, l_mixed)
    assert np.all(lb[[1, 3]] == 0)
    assert np.array_equal(lb[[0, 2, 4]], util.gpu_light_uniforms(w, np.array([3, 7, 29], dtype=np.uint32)))
    w.close()
