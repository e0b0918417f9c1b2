"""World mutation on the GPU path against the oracle: slot reclamation (ovis_world_compact) with spawn / despawn churn
held against the reference's own SimpleRenderWorld (tier A), re-parenting and removal of tree nodes, lights that follow a
node. Ids and order must stay the reference's; transforms and bounds are compared for exact equality."""
import numpy as np
import pytest

from orion_b200 import lib, scenes
import vis_testutil as util
from vis_testutil import same

pytestmark = pytest.mark.gpu
NONE = 0xFFFFFFFF


@pytest.fixture(scope="module")
def oracle():
    from oracle import pyoracle
    return pyoracle.load("best")


def _levels(parent):
    level = np.full(len(parent), -1, dtype=np.int64)
    for i in range(len(parent)):
        path, a = [], i
        while a >= 0 and level[a] < 0:
            path.append(a)
            a = parent[a]
        l = level[a] if a >= 0 else -1
        for n in reversed(path):
            l += 1
            level[n] = l
    return level


def _world_trs_any_order(oracle, parent, local):
    """oracle.world_trs wants parents before children; after re-parenting that need not hold: sort by level first."""
    order = np.argsort(_levels(parent), kind="stable")
    rank = np.empty(len(order), dtype=np.int64)
    rank[order] = np.arange(len(order))
    p = parent[order]
    p = np.where(p >= 0, rank[np.maximum(p, 0)], -1).astype(np.int32)
    out = np.empty_like(local)
    out[order] = oracle.world_trs(p, local[order])
    return out


def test_churn_and_compaction_against_the_reference_world(oracle):
    """Interleaved adds, removals and moves on CudaRenderWorld's C ABI and on the reference's SimpleRenderWorld (a
    std::list that shrinks on remove); after every step both must list the same entities in the same order - before a
    compaction through the stable ids, after it through the returned remap."""
    if not oracle.has_world:
        pytest.skip("needs the reference-compiled oracle")
    rng = np.random.default_rng(801)
    view = scenes.make_view((0, 0, 140), (1, 0, 0, 0), 90.0, 0.1, 1000.0, 1600, 900)
    ref = oracle.world()
    w = lib.World(0)
    ours_of_ref = []                                        # our current id of the reference's k-th added entity (NONE = removed)
    alive_ref = []

    def add(n):
        trs = util.random_trs(rng, n, pos=60.0)
        box = util.random_boxes(rng, n, extent=1.0, hmax=2.0)
        flags = rng.integers(0, 2, n).astype(np.uint32)
        ref.add_entities(trs, box, flags)
        first = w.add_entities(box, flags)
        ids = np.arange(first, first + n, dtype=np.uint32)
        w.set_transforms(ids, trs)
        ours_of_ref.extend(ids.tolist())
        alive_ref.extend([True] * n)

    def remove(k):
        live = [i for i, a in enumerate(alive_ref) if a]
        for i in rng.choice(live, k, replace=False):
            ref.remove_entity(int(i))
            w.remove_entities(np.array([ours_of_ref[i]], dtype=np.uint32))
            alive_ref[i] = False
            ours_of_ref[i] = NONE

    def check():
        r_ents, _ = ref.cull(view, flags=0)
        g_ents, _ = w.cull(view, flags=0)
        want = np.array([ours_of_ref[i] for i in r_ents], dtype=np.uint32)
        assert np.array_equal(g_ents, want)
        assert len(g_ents) > 0

    add(4000)
    check()
    remove(1500)
    check()
    add(1000)
    remove(500)
    check()
    live, slots = w.live_counts()
    assert (live, slots) == (3000, 5000)
    remap, new_slots = w.compact()
    assert new_slots == 3000 and len(remap) == 5000
    for i in range(len(ours_of_ref)):
        if ours_of_ref[i] != NONE:
            ours_of_ref[i] = int(remap[ours_of_ref[i]])
    assert sorted(x for x in ours_of_ref if x != NONE) == list(range(3000))
    assert w.live_counts() == (3000, 3000)
    check()
    # life goes on in the compacted world: moves, adds, removals, another compaction
    live = [i for i, a in enumerate(alive_ref) if a]
    moved = rng.choice(live, 800, replace=False)
    trs = util.random_trs(rng, len(moved), pos=60.0)
    ref.set_transforms(moved.astype(np.uint32), trs)
    w.set_transforms(np.array([ours_of_ref[i] for i in moved], dtype=np.uint32), trs)
    check()
    add(700)
    remove(900)
    check()
    remap, new_slots = w.compact()
    for i in range(len(ours_of_ref)):
        if ours_of_ref[i] != NONE:
            ours_of_ref[i] = int(remap[ours_of_ref[i]])
    assert new_slots == 2800
    check()
    remap, new_slots = w.compact()                          # nothing dead: identity
    assert new_slots == 2800 and np.array_equal(remap, np.arange(2800, dtype=np.uint32))
    ref.abandon()
    w.close()


def test_compaction_keeps_tree_bindings(oracle):
    """Entities bound to tree nodes (two on some nodes) keep following them after a compaction moved their slots."""
    rng = np.random.default_rng(802)
    n_nodes, n_ent = 3000, 3600
    parent = np.full(n_nodes, -1, dtype=np.int32)
    parent[50:] = rng.integers(0, np.arange(50, n_nodes))
    local = util.random_trs(rng, n_nodes, pos=30.0)
    boxes = util.random_boxes(rng, n_ent, extent=1.0, hmax=1.5)
    flags = rng.integers(0, 2, n_ent).astype(np.uint32)
    node_of = np.concatenate([np.arange(n_nodes), rng.integers(0, n_nodes, n_ent - n_nodes)]).astype(np.uint32)
    w = lib.World(0)
    e0 = w.add_entities(boxes, flags)
    n0 = w.add_nodes(parent, local)
    w.bind_entities(node_of + n0, np.arange(e0, e0 + n_ent, dtype=np.uint32))
    gone = rng.choice(n_ent, 1300, replace=False).astype(np.uint32)
    w.remove_entities(gone)
    remap, new_slots = w.compact()
    keep = np.flatnonzero(remap != NONE)
    assert new_slots == len(keep) == n_ent - 1300 and np.array_equal(remap[keep], np.arange(new_slots))
    # move every root: all bounds change
    roots = np.arange(50, dtype=np.uint32)
    local[:50] = util.random_trs(rng, 50, pos=30.0)
    w.set_node_local_trs(roots, local[:50])
    wt = oracle.world_trs(parent, local)
    want = oracle.world_aabb(wt[node_of[keep]], boxes[keep])
    assert same(w.world_aabbs(), want)
    view = scenes.make_view((0, 0, 120), (1, 0, 0, 0), 90.0, 0.1, 1000.0, 1600, 900)
    planes = oracle.view_planes(view)
    ents, _ = w.cull(view, flags=0)
    assert np.array_equal(ents, oracle.cull_aabb(want, planes))
    assert np.array_equal(w.cull_planes(planes, require=1), oracle.cull_aabb(want, planes, flags[keep], 1))
    w.close()


def test_reparent_and_remove_nodes(oracle):
    rng = np.random.default_rng(803)
    n = 2000
    parent = np.full(n, -1, dtype=np.int32)
    parent[20:] = rng.integers(0, np.arange(20, n))
    local = util.random_trs(rng, n, pos=20.0)
    boxes = util.random_boxes(rng, n, extent=1.0, hmax=1.5)
    w = lib.World(0)
    e0 = w.add_entities(boxes)
    n0 = w.add_nodes(parent, local)
    w.bind_entities(np.arange(n0, n0 + n, dtype=np.uint32), np.arange(e0, e0 + n, dtype=np.uint32))
    assert same(w.node_world_trs(), oracle.world_trs(parent, local))

    def is_ancestor(a, b):                                  # a is b or above b
        while b >= 0:
            if b == a:
                return True
            b = parent[b]
        return False

    moved = 0
    for node in rng.choice(np.arange(20, n), 60, replace=False):
        new_parent = int(rng.integers(-1, n))              # may be a LATER node, or -1 = becomes a root
        if new_parent >= 0 and is_ancestor(int(node), new_parent):
            with pytest.raises(lib.OvisError):
                w.set_node_parent(int(node), new_parent)
            continue
        w.set_node_parent(int(node), new_parent)
        parent[node] = new_parent
        moved += 1
    assert moved > 30
    wt = _world_trs_any_order(oracle, parent, local)
    assert same(w.node_world_trs(), wt)
    assert same(w.world_aabbs(), oracle.world_aabb(wt, boxes))
    # destroy a few subtrees: their entities leave every list, everything else is untouched
    dead = np.zeros(n, dtype=bool)
    for node in rng.choice(n, 15, replace=False):
        if dead[node]:
            continue
        sub = np.array([is_ancestor(int(node), int(i)) for i in range(n)])
        removed_nodes, removed_entities = w.remove_node(int(node))
        assert removed_nodes == int((sub & ~dead).sum()) == removed_entities
        dead |= sub
    with pytest.raises(lib.OvisError):
        w.remove_node(int(np.flatnonzero(dead)[0]))
    view = scenes.make_view((0, 0, 100), (1, 0, 0, 0), 100.0, 0.1, 1000.0, 1600, 900)
    exp = oracle.cull_aabb(oracle.world_aabb(wt, boxes), oracle.view_planes(view))
    ents, _ = w.cull(view, flags=0)
    assert np.array_equal(ents, exp[~dead[exp]]) and len(ents) > 0
    # the survivors still follow their (moved) parents
    roots = np.flatnonzero((parent < 0) & ~dead).astype(np.uint32)
    local[roots] = util.random_trs(rng, len(roots), pos=20.0)
    w.set_node_local_trs(roots, local[roots])
    wt = _world_trs_any_order(oracle, parent, local)
    got = w.node_world_trs()
    assert same(got[~dead], wt[~dead])
    w.close()


def test_lights_follow_their_nodes(oracle):
    rng = np.random.default_rng(804)
    parent = np.array([-1, 0, 1, -1], dtype=np.int32)
    local = util.random_trs(rng, 4, pos=10.0)
    w = lib.World(0)
    w.add_nodes(parent, local)
    lights = np.zeros(3, dtype=scenes.LIGHT_DTYPE)
    lights[0] = scenes.make_light(scenes.LIGHT_POINT, position=(0, 0, 0), range=30.0, flags=1)
    lights[1] = scenes.make_light(scenes.LIGHT_SPOT, position=(0, 0, 0), direction=(0, 0, -1), range=25.0, cutoff=30.0, flags=1)
    lights[2] = scenes.make_light(scenes.LIGHT_POINT, position=(5, 5, 5), range=10.0)
    w.add_lights(lights)
    w.bind_light_to_node(0, 2)
    w.bind_light_to_node(1, 3)
    view = scenes.make_view((0, 0, 80), (1, 0, 0, 0), 90.0, 0.1, 500.0, 1600, 900)
    bound = {0: 2, 1: 3}
    for step in range(4):
        wt = oracle.world_trs(parent, local)
        want = lights.copy()
        for li, node in bound.items():
            want["position"][li] = wt[node, 0:3]
            want["direction"][li] = scenes.quat_rotate(wt[node, 3:7].astype(np.float32), np.array([0, 0, -1], dtype=np.float32))
        _, ls = w.cull(view)                                 # the update (and the follow) happens on demand
        for li in range(3):
            got, _, _ = w.light(li)
            assert np.array_equal(got["position"], want["position"][li])
            assert np.allclose(got["direction"], want["direction"][li], atol=1e-6)
            want["direction"][li] = got["direction"]         # (quat_rotate here is float64; the library uses GLM's float32 form)
        assert np.array_equal(ls, oracle.light_cull(want, oracle.view_planes(view)))
        local[[0, 3]] = util.random_trs(rng, 2, pos=10.0)
        w.set_node_local_trs(np.array([0, 3], dtype=np.uint32), local[[0, 3]])
        if step == 2:
            w.bind_light_to_node(0, -1)                      # unbound: stays where it was
            del bound[0]
            lights["position"][0] = want["position"][0]
            lights["direction"][0] = want["direction"][0]
    w.close()
