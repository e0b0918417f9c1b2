"""ovis_cull_planes_noalloc - the cull for callers that may not allocate GPU memory.

A renderer calling from inside its frame loop cannot use cudaMalloc/cudaFree (they stall the device and
may fail under a fixed VRAM budget), so this entry point keeps its scratch in a fixed array inside the
library. The test drives it through the C ABI with caller-owned device buffers and requires exactly the
list the regular world path (ovis_cull_planes_device) produces - same ids, same order, same count - on two
calls in a row with different frusta, so that anything the first call left in the shared scratch would
show up in the second. It also checks that the size limit is enforced with a clear error.

This file is named to be collected last: a run that stops at the first failure has then already run every
other test. It needs nothing from the other tests and passes on its own in a fresh process.
"""
import numpy as np
import pytest

from orion_b200 import lib, scenes
import vis_testutil as util

pytestmark = pytest.mark.gpu


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


def test_noalloc_cull_rejects_what_it_cannot_hold():
    import torch
    d = torch.zeros(6 * 256, dtype=torch.float32, device="cuda")
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
    # an empty set is fine and yields an empty list
    lib.cull_planes_noalloc(0, 0, d.data_ptr(), 256, 0, planes, out.data_ptr(), 256, cnt.data_ptr())
    torch.cuda.synchronize()
    assert int(cnt.item()) == 0
