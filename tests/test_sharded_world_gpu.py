"""The single-process sharded world of the C ABI (ovis_sharded_*): several shards, one host thread calling, one worker
thread per shard inside the library, every shard's gather kernel storing into one page-locked buffer. On a one-GPU box
all shards sit on device 0 (a device may be listed more than once), so the driver's -m gpu run exercises the whole
host-consumer path; with two or more GPUs the shards spread out and the NCCL all-gather for device consumers runs too."""
import numpy as np
import pytest

from orion_b200 import lib, scenes
import vis_testutil as util

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def oracle():
    from oracle import pyoracle
    return pyoracle.load("best")


def _devices(n_shards):
    import torch
    g = torch.cuda.device_count()
    return [r % g for r in range(n_shards)]


@pytest.mark.parametrize("n_shards", [1, 2, 3])
def test_sharded_world_merged_list_is_the_single_world_list(oracle, n_shards):
    rng = np.random.default_rng(700 + n_shards)
    n = 500_003
    boxes = np.concatenate([util.edge_boxes(), util.random_boxes(rng, n - len(util.edge_boxes()), extent=300.0)])
    flags = rng.integers(0, 2, n).astype(np.uint32)
    sw = lib.ShardedWorld(_devices(n_shards))
    sw.add_entities(boxes, flags)
    views = [scenes.make_view((0, 0, 350), (1, 0, 0, 0), 100.0, 0.1, 2000.0, 1600, 900),       # a third of the world
             scenes.make_view((-100, 20, 40), scenes.quat_yaw_pitch(-70.0, 10.0), 50.0, 0.5, 400.0, 1024, 768)]
    for view in views:
        for require in (0, 1):
            exp = oracle.cull_aabb(boxes, oracle.view_planes(view), flags, require, threads=8)
            got, counts = sw.cull(view, require)
            assert np.array_equal(got, exp)                     # ids and order; the wide view outgrows the first segments
            assert counts.sum() == len(exp) and len(counts) == n_shards
    # a frame loop on one bound buffer, then removals routed to the owning shards
    out = np.empty(n, dtype=np.uint32)
    step = sw.bind_cull(views[1], out)
    exp = oracle.cull_aabb(boxes, oracle.view_planes(views[1]), threads=8)
    for _ in range(20):
        k, _ = step()
        assert np.array_equal(out[:k], exp)
    gone = rng.choice(n, 20_000, replace=False).astype(np.uint32)
    sw.remove_entities(gone)
    alive = np.ones(n, dtype=bool)
    alive[gone] = False
    k, _ = step()
    assert np.array_equal(out[:k], exp[alive[exp]])
    # too small a caller buffer: the required size comes back
    with pytest.raises(lib.OvisError) as e:
        sw.cull(views[0], capacity=10)
    assert e.value.code == lib.OVIS_ERR_CAPACITY
    sw.close()


def test_sharded_world_nccl_allgather_for_device_consumers(oracle):
    """ncclAllGather of the per-shard [count | ids] blocks from C++ (NCCL loaded at run time): one rank per distinct GPU
    of the box, so a single rank on a one-GPU box and the real exchange with `gpurun --gpus 2`."""
    import torch
    g = min(torch.cuda.device_count(), 8)
    rng = np.random.default_rng(710)
    n = 200_001
    boxes = util.random_boxes(rng, n, extent=300.0)
    sw = lib.ShardedWorld(list(range(g)))
    sw.add_entities(boxes)
    view = scenes.make_view((0, 0, 350), (1, 0, 0, 0), 60.0, 0.1, 2000.0, 1600, 900)
    planes = lib.view_planes(view)
    exp = oracle.cull_aabb(boxes, oracle.view_planes(view), threads=8)
    cap = len(exp) + 16
    blocks, counts = sw.allgather_device(planes, cap)
    assert counts.sum() == len(exp)
    for d in range(g):                                          # every device holds every shard's block
        import ctypes as C
        host = np.empty(g * (1 + cap), dtype=np.uint32)
        lib._check(lib.load().ovis_read_device(sw.shard(d)[0], C.c_void_p(blocks[d]), C.c_int64(host.nbytes), host.ctypes.data_as(C.c_void_p)))
        seg = host.reshape(g, 1 + cap)
        merged = np.concatenate([seg[r, 1:1 + int(seg[r, 0])] for r in range(g)])
        assert np.array_equal(merged, exp), d
    sw.close()


@pytest.mark.parametrize("n_shards", [1, 3])
def test_sharded_many_views_and_shadow_lists(oracle, n_shards, monkeypatch):
    """BASELINE configs[4] "then 5 views", and the per-light shadow-caster lists, over a sharded world: every merged list
    must be the single-world (= reference) list. The sparse pass is forced so that both many-view passes run per shard."""
    monkeypatch.setenv("OVIS_MV_DENSE", "0")
    rng = np.random.default_rng(720 + n_shards)
    n = 300_007
    boxes = np.concatenate([util.edge_boxes(), util.random_boxes(rng, n - len(util.edge_boxes()), extent=300.0, hmax=3.0)])
    flags = (rng.random(n) < 0.7).astype(np.uint32)
    lights = scenes.point_lights(10, 5, 250.0, 20.0, 90.0)
    spot = util.random_lights(rng, 3, extent=200.0)
    spot["type"] = scenes.LIGHT_SPOT
    spot["flags"] = 1
    lights = np.concatenate([lights, spot])
    sw = lib.ShardedWorld(_devices(n_shards))
    sw.add_entities(boxes, flags)
    assert sw.add_lights(lights) == 0
    cam = scenes.make_view((0, 0, 350), (1, 0, 0, 0), 100.0, 0.1, 2000.0, 1600, 900)
    views = [cam,
             scenes.make_view((-100, 20, 40), scenes.quat_yaw_pitch(-70.0, 10.0), 50.0, 0.5, 400.0, 1024, 768),
             scenes.make_view((50, -30, 10), scenes.quat_yaw_pitch(120.0, -20.0), 70.0, 0.5, 300.0, 1024, 1024),
             scenes.make_view((0, 200, 0), scenes.quat_yaw_pitch(0.0, -89.0), 60.0, 1.0, 500.0, 2048, 2048),
             scenes.make_view((10, 10, 10), scenes.quat_yaw_pitch(45.0, 45.0), 30.0, 0.1, 50.0, 512, 512)]
    planes = np.stack([oracle.view_planes(v) for v in views])
    require = np.array([0, 1, 0, 1, 0], dtype=np.uint32)
    for _ in range(2):                                          # the second call finds the workers' buffers already grown
        off, ent = sw.cull_views(planes, require, capacity=16)
        assert len(off) == 6 and off[0] == 0 and off[-1] == len(ent)
        for v in range(5):
            exp = oracle.cull_aabb(boxes, planes[v], flags, int(require[v]), threads=8)
            assert np.array_equal(ent[off[v]:off[v + 1]], exp), v
    ll, lv, off, ent = sw.shadow_lists(cam, len(lights), capacity=16)
    e_ll, e_lv, e_off, e_ent = oracle.shadow_lists(boxes, flags, lights, oracle.view_planes(cam))
    assert np.array_equal(ll, e_ll) and np.array_equal(lv, e_lv)
    assert np.array_equal(off, e_off) and np.array_equal(ent, e_ent)
    assert len(ll) > 6 and len(ent) > 0
    sw.close()


def test_pinning_a_thread_next_to_its_gpu_stays_within_the_allowed_cpus():
    """ovis_host_pin_thread_near_device: the thread ends up on a non-empty subset of the CPUs it was allowed before (the
    GPU's local_cpulist within the cpuset), or where it was when the box has no topology files. Runs in a thread of its own
    so that the test process keeps its affinity."""
    import os
    import threading
    result = {}

    def body():
        before = os.sched_getaffinity(0)
        n = lib.pin_thread_near_device(0)
        after = os.sched_getaffinity(0)
        result.update(before=before, after=after, n=n)

    t = threading.Thread(target=body)
    t.start()
    t.join()
    assert result["after"] and result["after"] <= result["before"]
    assert result["n"] in (0, len(result["after"]))
    with pytest.raises(lib.OvisError):
        lib.pin_thread_near_device(99)
