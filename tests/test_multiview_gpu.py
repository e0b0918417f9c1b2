"""K4 / K6 on the GPU against the CPU oracle: the dense pass and the sparse light-grid pass must produce the reference's
lists (ids and order), agree with each other, and the sparse pass must hand over to the dense one when a block of
entities is too crowded for it. BASELINE.json configs[2] and configs[3] run at FULL size with every list checked.
Also the back-to-back stress of the single-view cull (programmatic dependent launch between consecutive culls)."""
import os
import sys

import numpy as np
import pytest

from orion_b200 import lib, scenes
import vis_testutil as util

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def oracle():
    from oracle import pyoracle
    return pyoracle.load("best")


def _world_static(aabb, flags=None):
    w = lib.World(0)
    first = w.add_entities(aabb, flags)
    w.set_world_aabbs(first, aabb)
    return w


def _shadow_lists(w, light_ids):
    ll, lv, off, _, d_ent = w.cull_shadow_views_device(np.asarray(light_ids, dtype=np.uint32))
    return ll, lv, off, w.read_device(d_ent, int(off[-1]))


def _scene(rng, n, n_point, n_spot, extent, rmin, rmax):
    boxes = util.random_boxes(rng, n, extent=extent, hmax=4.0)
    e = util.edge_boxes()
    boxes[:len(e)] = e
    flags = (rng.random(n) < 0.8).astype(np.uint32)
    lights = scenes.point_lights(n_point, 5, extent * 0.9, rmin, rmax)
    if n_spot:
        spot = util.random_lights(rng, n_spot, extent=extent * 0.8)
        spot["type"] = scenes.LIGHT_SPOT
        spot["flags"] = 1
        spot["intensity"] = 1.0
        spot["colour"] = 1.0
        lights = np.concatenate([lights, spot])
    return boxes, flags, lights


def _expected_lists(oracle, boxes, flags, lights, ids):
    out = []
    for i in ids:
        for v in oracle.light_shadow_views(lights[i]):
            out.append(oracle.cull_aabb(boxes, oracle.view_planes(v), flags, 1, threads=8))
    return out


def test_sparse_and_dense_passes_give_the_reference_lists(oracle, monkeypatch):
    rng = np.random.default_rng(501)
    boxes, flags, lights = _scene(rng, 150_003, 48, 6, 400.0, 10.0, 90.0)
    ids = np.arange(len(lights), dtype=np.uint32)
    expected = _expected_lists(oracle, boxes, flags, lights, ids)
    results = []
    for dense in ("0", "1"):
        monkeypatch.setenv("OVIS_MV_DENSE", dense)            # read when the world is created
        w = _world_static(boxes, flags)
        w.add_lights(lights)
        ll, lv, off, ent = _shadow_lists(w, ids)
        assert len(ll) == len(expected) == 48 * 6 + 6
        for k, exp in enumerate(expected):
            assert np.array_equal(ent[off[k]:off[k + 1]], exp), (dense, k)
        # removal and a second call on the same scratch
        gone = rng.choice(np.arange(100, len(boxes)), 5000, replace=False).astype(np.uint32)
        w.remove_entities(gone)
        keep = np.ones(len(boxes), dtype=bool)
        keep[gone] = False
        ll2, lv2, off2, ent2 = _shadow_lists(w, ids)
        for k, exp in enumerate(expected):
            assert np.array_equal(ent2[off2[k]:off2[k + 1]], exp[keep[exp]]), (dense, k)
        results.append((off, ent))
        w.close()
    assert np.array_equal(results[0][0], results[1][0]) and np.array_equal(results[0][1], results[1][1])


def test_sparse_pass_hands_over_when_a_block_is_too_crowded(oracle, monkeypatch):
    """64 lights whose cubes all contain all 20 000 entities: 64 candidate pairs per entity, far beyond what a block's
    shared-memory pair list holds. The call must notice, rerun those views through the dense pass and still be exact.
    (OVIS_MV_DENSE=0 keeps the size rule from sending so small a world to the dense pass at once.)"""
    monkeypatch.setenv("OVIS_MV_DENSE", "0")
    rng = np.random.default_rng(502)
    n = 20_000
    boxes = util.random_boxes(rng, n, extent=20.0, hmax=1.0)
    flags = np.ones(n, dtype=np.uint32)
    lights = scenes.point_lights(64, 9, 15.0, 200.0, 300.0)
    w = _world_static(boxes, flags)
    w.add_lights(lights)
    ids = np.arange(64, dtype=np.uint32)
    ll, lv, off, ent = _shadow_lists(w, ids)
    expected = _expected_lists(oracle, boxes, flags, lights, ids)
    assert int(off[-1]) > 64 * n                              # most entities are in several faces of every light
    for k, exp in enumerate(expected):
        assert np.array_equal(ent[off[k]:off[k + 1]], exp), k
    # and a sparse frame afterwards on the same world is still served correctly
    far = scenes.point_lights(8, 10, 15.0, 2.0, 4.0)
    first = w.add_lights(far)
    ids2 = np.arange(first, first + 8, dtype=np.uint32)
    ll, lv, off, ent = _shadow_lists(w, ids2)
    for k, exp in enumerate(_expected_lists(oracle, boxes, flags, far, range(8))):
        assert np.array_equal(ent[off[k]:off[k + 1]], exp), k
    w.close()


@pytest.mark.parametrize("dense", ["0", "1"])
def test_point_lights_whose_range_is_below_the_shadow_near_plane(oracle, monkeypatch, dense):
    """perspective(90, 0.1, range) with range <= 0.1 turns the far plane round: the reference's face frusta then open
    outwards and list entities any distance from the light (render_light.cc:319 has no guard). Such lights must not be
    pre-filtered by a cube of their range; ordinary lights in the same call still are."""
    monkeypatch.setenv("OVIS_MV_DENSE", dense)
    rng = np.random.default_rng(505)
    n = 40_000
    boxes = util.random_boxes(rng, n, extent=120.0, hmax=2.0)
    flags = np.ones(n, dtype=np.uint32)
    lights = scenes.point_lights(12, 11, 100.0, 20.0, 60.0)
    lights["range"][:8] = [0.02, 0.05, 0.07, 0.1, 0.0, -1.0, np.float32(0.1) + np.float32(1e-6), 0.0999]
    w = _world_static(boxes, flags)
    w.add_lights(lights)
    ids = np.arange(len(lights), dtype=np.uint32)
    ll, lv, off, ent = _shadow_lists(w, ids)
    expected = _expected_lists(oracle, boxes, flags, lights, ids)
    assert sum(len(e) for e in expected[:6]) > n // 2          # range 0.02: nearly every entity lands in some face
    for k, exp in enumerate(expected):
        assert np.array_equal(ent[off[k]:off[k + 1]], exp), (dense, k)
    w.close()


def test_inverted_boxes_reach_the_lists_through_the_light_grid(oracle, monkeypatch):
    """min > max on some axes: the reference tests such a box like any other (its P-vertex is then the least favourable
    corner) and lists it when the whole spanned box is inside a face frustum. The grid must file it under the cells of the
    spanned box, not under an empty range."""
    monkeypatch.setenv("OVIS_MV_DENSE", "0")
    rng = np.random.default_rng(506)
    n = 30_000
    boxes = util.random_boxes(rng, n, extent=300.0, hmax=3.0)
    flip = rng.random((n, 3)) < 0.3
    lo, hi = boxes[:, :3].copy(), boxes[:, 3:].copy()
    boxes[:, :3] = np.where(flip, hi, lo)
    boxes[:, 3:] = np.where(flip, lo, hi)
    flags = np.ones(n, dtype=np.uint32)
    lights = scenes.point_lights(24, 12, 280.0, 40.0, 320.0)
    w = _world_static(boxes, flags)
    w.add_lights(lights)
    ids = np.arange(len(lights), dtype=np.uint32)
    ll, lv, off, ent = _shadow_lists(w, ids)
    expected = _expected_lists(oracle, boxes, flags, lights, ids)
    inverted = flip.any(axis=1)
    assert sum(int(inverted[e].sum()) for e in expected) > 1000
    for k, exp in enumerate(expected):
        assert np.array_equal(ent[off[k]:off[k + 1]], exp), k
    w.close()


def test_small_worlds_take_the_dense_pass_and_agree(oracle, monkeypatch):
    """Below ~1e8 entity-view tests the cube groups are tested densely (cheaper than the grid); same lists either way."""
    rng = np.random.default_rng(504)
    boxes, flags, lights = _scene(rng, 20_000, 12, 2, 60.0, 10.0, 40.0)
    ids = np.arange(len(lights), dtype=np.uint32)
    got = []
    for mode in (None, "0"):
        if mode is not None:
            monkeypatch.setenv("OVIS_MV_DENSE", mode)
        w = _world_static(boxes, flags)
        w.add_lights(lights)
        got.append(_shadow_lists(w, ids))
        w.close()
    for a, b in zip(got[0], got[1]):
        assert np.array_equal(a, b)
    ll, lv, off, ent = got[0]
    for k, exp in enumerate(_expected_lists(oracle, boxes, flags, lights, ids)):
        assert np.array_equal(ent[off[k]:off[k + 1]], exp), k


def test_many_lights_small_world_and_mixed_plain_views(oracle):
    """4096 point lights over a small world (every light its own cube group: the per-call group limit), spot lights in
    between (plain views among cube groups), and a camera + cascades call on the same world."""
    rng = np.random.default_rng(503)
    boxes, flags, lights = _scene(rng, 60_000, 4096, 0, 600.0, 5.0, 30.0)
    spot = util.random_lights(rng, 40, extent=500.0)
    spot["type"] = scenes.LIGHT_SPOT
    spot["flags"] = 1
    spot["intensity"] = 1.0
    spot["colour"] = 1.0
    lights = np.concatenate([lights[:2000], spot, lights[2000:]])
    w = _world_static(boxes, flags)
    w.add_lights(lights)
    ids = np.arange(len(lights), dtype=np.uint32)
    ll, lv, off, ent = _shadow_lists(w, ids)
    assert len(ll) == 4096 * 6 + 40
    check = rng.choice(len(ll), 300, replace=False)
    for k in check:
        v = oracle.light_shadow_views(lights[ll[k]])[lv[k]]
        assert np.array_equal(ent[off[k]:off[k + 1]], oracle.cull_aabb(boxes, oracle.view_planes(v), flags, 1)), k
    cam = scenes.default_camera()
    views = np.concatenate([np.asarray(cam).reshape(1), scenes.cascade_views(cam)])
    planes = np.stack([lib.view_planes(v).reshape(24) for v in views])
    off, _, d_ent = w.cull_views_device(planes)
    ent = w.read_device(d_ent, int(off[-1]))
    for i in range(len(views)):
        assert np.array_equal(ent[off[i]:off[i + 1]], oracle.cull_aabb(boxes, planes[i]))
    w.close()


def test_shadow_views_host_output_with_and_without_the_caster_filter(oracle):
    """ovis_cull_shadow_views: the batched pass the adapter serves RenderWorld::cull(shadowView, results, 0) from - with
    require = 0 every intersecting entity is listed (SimpleRenderWorld's answer; the pipeline filters castsShadow() itself),
    with OVIS_CASTS_SHADOW the casters only. Several calls in a row: the speculative copy of the ids must never cut a list."""
    rng = np.random.default_rng(505)
    boxes, flags, lights = _scene(rng, 400_000, 20, 4, 300.0, 20.0, 120.0)
    w = _world_static(boxes, flags)
    w.add_lights(lights)
    for ids in (np.arange(24, dtype=np.uint32), np.array([3, 21], dtype=np.uint32), np.arange(24, dtype=np.uint32)):
        for require in (0, 1):
            ll, lv, off, ent = w.cull_shadow_views(ids, require)
            k = 0
            for i in ids:
                for v in oracle.light_shadow_views(lights[i]):
                    exp = oracle.cull_aabb(boxes, oracle.view_planes(v), flags if require else None, require, threads=8)
                    assert np.array_equal(ent[off[k]:off[k + 1]], exp), (int(i), require, k)
                    k += 1
            assert k == len(ll) and int(off[-1]) == len(ent)
    w.close()


@pytest.mark.parametrize("dense", ["0", "1"])
def test_calls_split_into_rounds_give_the_same_lists(oracle, monkeypatch, dense):
    """A call whose scratch would not fit is processed in rounds of views (here forced: 8 views per round). Offsets continue
    from round to round, the overflow word outlives a round, and every round's offsets reach the host."""
    monkeypatch.setenv("OVIS_MV_ROUND_VIEWS", "8")
    monkeypatch.setenv("OVIS_MV_DENSE", dense)
    rng = np.random.default_rng(506)
    boxes, flags, lights = _scene(rng, 120_000, 9, 5, 250.0, 20.0, 100.0)
    w = _world_static(boxes, flags)
    w.add_lights(lights)
    ids = np.arange(len(lights), dtype=np.uint32)
    for _ in range(2):
        ll, lv, off, ent = _shadow_lists(w, ids)
        exp = _expected_lists(oracle, boxes, flags, lights, ids)
        assert len(ll) == len(exp) == 9 * 6 + 5
        for k, e in enumerate(exp):
            assert np.array_equal(ent[off[k]:off[k + 1]], e), k
    w.close()


def _bench_configs():
    if ROOT not in sys.path:
        sys.path.insert(0, ROOT)
    from tools import bench_configs
    return bench_configs


def test_config3_full_size_every_list(oracle):
    """BASELINE.json configs[2] at full size: 10 M dynamic entities under a depth-3 tree, 4096 point lights; transforms,
    bounds, main cull, light cull and EVERY shadow list against the oracle."""
    r = _bench_configs().config3(False, full_parity=True)
    p = r["parity"]
    assert p["world_trs_exact"] and p["world_aabb_exact"] and p["main_view_and_lights"]
    assert p["shadow_lists_exact_and_api_consistent"] and p["lists_checked"] == p["lists"] > 100
    assert r["visible_entities"] > 10_000


def test_config4_full_size_every_list(oracle):
    """BASELINE.json configs[3] at full size: 10 M entities x (camera + 4 cascades + 64 point lights x 6 faces); all 389
    lists against the oracle, and the sparse pass against the dense pass."""
    r = _bench_configs().config4(False, full_parity=True, brute=True)
    p = r["parity"]
    assert p["camera_views"] and p["light_faces_exact"] and p["lists_checked"] == p["lists"] == 384
    assert p["prefilter_equals_brute_force_all_384"]


@pytest.mark.parametrize("n", [600_000, 1_000_000])
def test_thousand_culls_back_to_back_with_alternating_views(oracle, n):
    """Consecutive culls overlap on the GPU (the next test kernel is a programmatic dependent of the previous gather
    kernel) and rotate their scratch sets. 1000 frames are queued without any host synchronisation, alternating between a
    wide and a narrow frustum with heavy visibility - at these sizes the two kernels take about equally long, the case
    in which a missing ordering would show. Every frame's count and a sample of whole lists are checked."""
    import torch
    rng = np.random.default_rng(600 + n % 7)
    boxes = util.random_boxes(rng, n, extent=200.0)
    w = _world_static(boxes)
    stream = torch.cuda.Stream()
    w.set_stream(stream.cuda_stream)
    views = [scenes.make_view((0, 0, 260), (1, 0, 0, 0), 100.0, 0.1, 2000.0, 1600, 900),
             scenes.make_view((50, 10, 100), scenes.quat_yaw_pitch(25.0, -5.0), 40.0, 0.5, 600.0, 1280, 720)]
    planes = [lib.view_planes(v) for v in views]
    expected = [oracle.cull_aabb(boxes, oracle.view_planes(v), threads=8) for v in views]
    assert len(expected[0]) > n // 4 and 0 < len(expected[1]) < len(expected[0])
    frames = 1000
    counts = torch.zeros(frames, dtype=torch.int32, device="cuda")
    ring = [torch.zeros(n, dtype=torch.int32, device="cuda") for _ in range(2)]
    kept = {f: torch.zeros(n, dtype=torch.int32, device="cuda") for f in range(0, frames, 97)}
    torch.cuda.synchronize()
    for f in range(frames):
        out = kept.get(f, ring[f & 1])
        w.cull_planes_device(planes[f & 1], out.data_ptr(), counts.data_ptr() + 4 * f)
    stream.synchronize()
    got_counts = counts.cpu().numpy()
    want = np.array([len(expected[f & 1]) for f in range(frames)], dtype=np.int32)
    assert np.array_equal(got_counts, want)
    for f, buf in kept.items():
        assert np.array_equal(buf[:want[f]].cpu().numpy().view(np.uint32), expected[f & 1]), f
    for f in (frames - 2, frames - 1):
        assert np.array_equal(ring[f & 1][:want[f]].cpu().numpy().view(np.uint32), expected[f & 1])
    w.close()


@pytest.mark.parametrize("n", [5_000, 400_000])
def test_depth_slices_of_one_frustum_share_their_side_planes(oracle, monkeypatch, n):
    """A camera and nine depth slices of its frustum (bit-identical left / right / top / bottom planes: runs of 7 + 3),
    two more cameras, per-view caster filters and removed entities: the shared-side runs must give the lists of the
    reference - and of the same call with the sharing turned off."""
    rng = np.random.default_rng(507)
    boxes = util.random_boxes(rng, n, extent=300.0, hmax=4.0)
    e = util.edge_boxes()
    boxes[:len(e)] = e
    flags = (rng.random(n) < 0.5).astype(np.uint32)
    cam = scenes.make_view((10, 5, 280), scenes.quat_yaw_pitch(8.0, -3.0), 80.0, 0.1, 900.0, 1600, 900)
    cuts = [0.1, 3.0, 20.0, 60.0, 120.0, 200.0, 300.0, 450.0, 600.0, 900.0]
    views = [cam] + [scenes.make_view(cam["position"], cam["orientation"], 80.0, cuts[i], cuts[i + 1], 1600, 900) for i in range(9)]
    views += [scenes.make_view((0, 0, 0), scenes.quat_yaw_pitch(120.0, 10.0), 60.0, 0.5, 500.0, 800, 600),
              scenes.make_view((10, 5, 280), scenes.quat_yaw_pitch(8.0, -3.0), 70.0, 0.1, 900.0, 1600, 900)]   # same pose, other fov
    planes = np.stack([oracle.view_planes(v) for v in views])
    assert all(np.array_equal(planes[0, :4], planes[k, :4]) for k in range(1, 10))       # the premise: identical side planes
    require = rng.integers(0, 2, len(views)).astype(np.uint32)
    gone = rng.choice(np.arange(20, n), n // 7, replace=False).astype(np.uint32)
    keep = np.ones(n, dtype=bool)
    keep[gone] = False
    got = []
    for shared in ("1", "0"):
        monkeypatch.setenv("OVIS_MV_SHARED_SIDES", shared)
        w = _world_static(boxes, flags)
        w.remove_entities(gone)
        off, _, d_ent = w.cull_views_device(planes, require)
        got.append((off.copy(), w.read_device(d_ent, int(off[-1]))))
        w.close()
    assert np.array_equal(got[0][0], got[1][0]) and np.array_equal(got[0][1], got[1][1])
    off, ent = got[0]
    listed = 0
    for k in range(len(views)):
        exp = oracle.cull_aabb(boxes, planes[k], flags, int(require[k]), threads=8)
        exp = exp[keep[exp]]
        assert np.array_equal(ent[off[k]:off[k + 1]], exp), k
        listed += len(exp)
    assert listed > n // 10
