"""ovis_cull_planes_noalloc - the cull for callers that may not allocate GPU memory.

A renderer calling from inside its frame loop cannot use cudaMalloc/cudaFree (they stall the device and
may fail under a fixed VRAM budget), so this entry point keeps its scratch in a fixed array inside the
library. The tests drive it through the C ABI with caller-owned device buffers and require exactly the
list the regular world path (ovis_cull_planes_device) produces - same ids, same order, same count - on two
calls in a row with different frusta, so that anything the first call left in the shared scratch would
show up in the second; and exactly the list the CPU oracle computes, including for the edge-case boxes,
for two calls queued back to back on one stream with no host synchronisation in between, at the size
limit, and when `capacity` clips the output. Sizes beyond the limit and misaligned bounds are refused
with a clear error.

This file is named to be collected last: a run that stops at the first failure has then already run every
other test. It needs nothing from the other tests and passes on its own in a fresh process.
"""
import numpy as np
import pytest

from orion_b200 import lib, scenes
import vis_testutil as util

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def oracle():
    from oracle import pyoracle
    return pyoracle.load("best")


def _soa_on_device(torch, aabb, stride):
    """Caller-side copy of the bounds in the library's SoA layout: component k of entity i at [k*stride + i]."""
    soa = np.zeros((6, stride), dtype=np.float32)
    soa[:, :len(aabb)] = aabb.T
    return torch.from_numpy(soa).cuda()


def _world_result(world, torch, planes, n):
    out = torch.zeros(n + 1, dtype=torch.int32, device="cuda")
    world.cull_planes_device(planes, out.data_ptr() + 4, out.data_ptr())
    torch.cuda.synchronize()
    count = int(out[0].item())
    return out[1:1 + count].cpu().numpy().view(np.uint32)


def test_noalloc_cull_matches_world_cull_twice_in_a_row():
    import torch
    rng = np.random.default_rng(71)
    n = 300_001                                   # ragged: not a multiple of the 256-entity warp tile
    aabb = util.random_boxes(rng, n, extent=400.0)
    stride = (n + 255) // 256 * 256
    d_bounds = _soa_on_device(torch, aabb, stride)
    d_out = torch.zeros(n, dtype=torch.int32, device="cuda")
    d_count = torch.zeros(1, dtype=torch.int32, device="cuda")

    world = lib.World(0)
    first = world.add_entities(aabb)
    world.set_world_aabbs(first, aabb)

    views = [scenes.make_view((0, 0, 450), (1, 0, 0, 0), 90.0, 0.1, 2000.0, 1600, 900),     # many visible
             scenes.make_view((300, 50, 0), scenes.quat_yaw_pitch(40.0, -5.0), 30.0, 1.0, 150.0, 512, 512)]  # few
    sizes = []
    for view in views:                            # two calls in a row on the same shared scratch
        planes = lib.view_planes(view)
        expected = _world_result(world, torch, planes, n)
        lib.cull_planes_noalloc(0, 0, d_bounds.data_ptr(), stride, n, planes, d_out.data_ptr(), n, d_count.data_ptr())
        torch.cuda.synchronize()
        count = int(d_count.item())
        got = d_out[:count].cpu().numpy().view(np.uint32)
        assert count == len(expected)
        assert np.array_equal(got, expected)      # integer ids: the tolerance is zero, as everywhere in these tests
        sizes.append(count)
    assert sizes[0] > 1000 and 0 < sizes[1] < sizes[0]

    # id_base is added to every id, as in the sharded world path
    lib.cull_planes_noalloc(0, 0, d_bounds.data_ptr(), stride, n, planes, d_out.data_ptr(), n, d_count.data_ptr(),
                            id_base=1_000_000)
    torch.cuda.synchronize()
    assert np.array_equal(d_out[:int(d_count.item())].cpu().numpy().view(np.uint32), expected + np.uint32(1_000_000))
    world.close()


def test_noalloc_cull_matches_the_oracle_back_to_back_on_one_stream(oracle):
    """Against the CPU oracle (not another CUDA path): edge-case boxes in front, two frusta queued on one stream with
    no host synchronisation between them (the second call reuses the scratch the first one's kernels are still
    using, in stream order), each into its own output buffer."""
    import torch
    rng = np.random.default_rng(72)
    edge = util.edge_boxes()
    n = 200_003
    aabb = np.concatenate([edge, util.random_boxes(rng, n - len(edge), extent=300.0)])
    stride = (n + 255) // 256 * 256
    d_bounds = _soa_on_device(torch, aabb, stride)
    views = [scenes.make_view((0, 0, 350), (1, 0, 0, 0), 100.0, 0.1, 2000.0, 1600, 900),
             scenes.make_view((-100, 20, 40), scenes.quat_yaw_pitch(-70.0, 10.0), 50.0, 0.5, 400.0, 1024, 768),
             scenes.make_view((0, 0, 350), (1, 0, 0, 0), 100.0, 0.1, 2000.0, 1600, 900)]
    outs = [torch.zeros(n, dtype=torch.int32, device="cuda") for _ in views]
    counts = [torch.zeros(1, dtype=torch.int32, device="cuda") for _ in views]
    stream = torch.cuda.Stream()
    torch.cuda.synchronize()
    for view, d_out, d_count in zip(views, outs, counts):                 # queued, never synchronised in between
        lib.cull_planes_noalloc(0, stream.cuda_stream, d_bounds.data_ptr(), stride, n, lib.view_planes(view),
                                d_out.data_ptr(), n, d_count.data_ptr())
    stream.synchronize()
    for view, d_out, d_count in zip(views, outs, counts):
        expected = oracle.cull_aabb(aabb, oracle.view_planes(view))
        count = int(d_count.item())
        assert count == len(expected) and count > 0
        assert np.array_equal(d_out[:count].cpu().numpy().view(np.uint32), expected)
    assert int(counts[0].item()) != int(counts[1].item())


def test_noalloc_cull_at_the_size_limit_and_with_a_clipped_capacity(oracle):
    import torch
    rng = np.random.default_rng(73)
    n = lib.NOALLOC_MAX_ENTITIES                  # exactly the limit: every tile and block slot of the scratch is used
    aabb = util.random_boxes(rng, n, extent=500.0)
    d_bounds = _soa_on_device(torch, aabb, n)
    view = scenes.make_view((0, 0, 600), (1, 0, 0, 0), 80.0, 0.1, 3000.0, 1600, 900)
    planes = lib.view_planes(view)
    expected = oracle.cull_aabb(aabb, oracle.view_planes(view), threads=8)
    assert len(expected) > 10_000
    d_out = torch.zeros(n, dtype=torch.int32, device="cuda")
    d_count = torch.zeros(1, dtype=torch.int32, device="cuda")
    lib.cull_planes_noalloc(0, 0, d_bounds.data_ptr(), n, n, planes, d_out.data_ptr(), n, d_count.data_ptr())
    torch.cuda.synchronize()
    assert int(d_count.item()) == len(expected)
    assert np.array_equal(d_out[:len(expected)].cpu().numpy().view(np.uint32), expected)

    # capacity below the visible count: the first `capacity` ids are written, nothing beyond, and the count is the true one
    cap = 1000
    d_small = torch.full((cap + 64,), -1, dtype=torch.int32, device="cuda")
    lib.cull_planes_noalloc(0, 0, d_bounds.data_ptr(), n, n, planes, d_small.data_ptr(), cap, d_count.data_ptr())
    torch.cuda.synchronize()
    assert int(d_count.item()) == len(expected)
    got = d_small.cpu().numpy()
    assert np.array_equal(got[:cap].view(np.uint32), expected[:cap])
    assert np.all(got[cap:] == -1)


def test_noalloc_cull_rejects_what_it_cannot_hold():
    import torch
    d = torch.zeros(6 * 256 + 4, dtype=torch.float32, device="cuda")
    out = torch.zeros(256, dtype=torch.int32, device="cuda")
    cnt = torch.zeros(1, dtype=torch.int32, device="cuda")
    planes = lib.view_planes(scenes.default_camera())
    too_many = lib.NOALLOC_MAX_ENTITIES + 1
    with pytest.raises(lib.OvisError) as e:
        lib.cull_planes_noalloc(0, 0, d.data_ptr(), (too_many + 255) // 256 * 256, too_many, planes, out.data_ptr(), 256,
                                cnt.data_ptr())
    assert e.value.code == lib.OVIS_ERR_CAPACITY and "exceed" in str(e.value)
    with pytest.raises(lib.OvisError) as e:      # stride must cover whole warp tiles
        lib.cull_planes_noalloc(0, 0, d.data_ptr(), 100, 100, planes, out.data_ptr(), 256, cnt.data_ptr())
    assert e.value.code == lib.OVIS_ERR_INVALID
    with pytest.raises(lib.OvisError) as e:      # the bounds are read as 128-bit vectors: 4 bytes off is refused, not a fault
        lib.cull_planes_noalloc(0, 0, d.data_ptr() + 4, 256, 100, planes, out.data_ptr(), 256, cnt.data_ptr())
    assert e.value.code == lib.OVIS_ERR_INVALID and "aligned" in str(e.value)
    # an empty set is fine and yields an empty list
    lib.cull_planes_noalloc(0, 0, d.data_ptr(), 256, 0, planes, out.data_ptr(), 256, cnt.data_ptr())
    torch.cuda.synchronize()
    assert int(cnt.item()) == 0
    # and the context is still healthy after the refusals
    lib.cull_planes_noalloc(0, 0, d.data_ptr(), 256, 100, planes, out.data_ptr(), 256, cnt.data_ptr())
    torch.cuda.synchronize()
    assert int(cnt.item()) >= 0
