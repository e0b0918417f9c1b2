"""Two real GPUs: both exchange modes of the sharded world (NCCL all-gather, and the gather kernel storing straight into
the peers' symmetric-memory segments over NVLink) must reproduce the single-world visible list. Skipped on boxes with
fewer than two GPUs (the driver's -m gpu run has one); exercised with `gpurun --gpus 2`."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys, numpy as np, torch, torch.distributed as dist
sys.path.insert(0, %r)
from orion_b200 import lib, scenes
from orion_b200.sharded import ShardedWorld, merged_list, shard_range
from oracle import pyoracle
rank, size = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
oracle = pyoracle.load("best")
n_total = 600_001
aabb, flags = scenes.static_boxes(n_total, 9, 800.0)
lo, hi = shard_range(n_total, rank, size)
for mode in ("nccl", "p2p"):
    w = lib.World(rank)
    f = w.add_entities(aabb[lo:hi], flags[lo:hi]); w.set_world_aabbs(f, aabb[lo:hi])
    sw = ShardedWorld(w, rank, size, lo, hi - lo, stream=torch.cuda.Stream(), exchange=mode)
    for view, req in ((scenes.default_camera(), 0), (scenes.make_view((100, 0, 300), (1, 0, 0, 0), 60.0, 1.0, 900.0, 800, 600), 1),
                      (scenes.default_camera(), 1)):
        planes = lib.view_planes(view)
        blocks, counts = sw.cull(planes, req)
        got = merged_list(blocks, counts)
        exp = oracle.cull_aabb(aabb, planes, flags, req, threads=4)
        assert np.array_equal(got, exp), (mode, rank, len(got), len(exp))
        for _ in range(2):                                   # host path: speculative copy, then steady state
            parts = sw.cull_to_host(planes, req)
            assert np.array_equal(np.concatenate(parts), exp), (mode, rank, "host")
    assert mode == "nccl" or sw.exchange_mode == "p2p", "symmetric memory path did not come up"
    if mode == "p2p":
        # 300 frames back to back with the visible set changing every frame: each rank copies the segments the moment the
        # ready flags allow it; a flag that arrives before a peer's stores would show up as a stale or torn segment
        views = [scenes.default_camera(), scenes.make_view((100, 0, 300), (1, 0, 0, 0), 60.0, 1.0, 900.0, 800, 600),
                 scenes.make_view((-300, 50, 0), scenes.quat_yaw_pitch(70.0, 0.0), 90.0, 0.5, 700.0, 1600, 900)]
        planes = [lib.view_planes(v) for v in views]
        expected = [oracle.cull_aabb(aabb, p, threads=4) for p in planes]
        frames = 300
        with torch.cuda.stream(sw.stream):
            history = torch.zeros((frames, size, sw.capacity + 1), dtype=torch.int32, device="cuda")
            for i in range(frames):
                blocks = sw.cull_device(planes[i %% 3])
                sw.enqueue_ready_wait()
                history[i].copy_(blocks)
        sw.stream.synchronize()
        assert w.exchange_error() == 0
        host = history.cpu()
        for i in range(frames):
            got = merged_list(host[i])
            assert np.array_equal(got, expected[i %% 3]), (rank, "frame", i, len(got), len(expected[i %% 3]))
    if mode == "p2p":
        # host consumers: every rank's kernel stores its segment into one shared page-locked block, no exchange at all;
        # a small first block forces the regrow path on the wide frustum
        from orion_b200.sharded import SharedHostSegments
        seg = SharedHostSegments(rank, size, 2000, "ovis_test2gpu_%%d" %% os.getppid())
        cams = [np.asarray(v).reshape(1).copy() for v in views]
        cam = cams[0].copy()
        step = seg.bind(w, cam)
        for i in range(30):
            cam[:] = cams[i %% 3]                             # the view is read in place every frame
            parts = step()
            assert np.array_equal(np.concatenate(parts), expected[i %% 3]), (rank, "shared host frame", i)
        assert seg.generation >= 1
        dist.barrier()
        seg.close()
    if mode == "p2p":
        # segments far too small for the lists: every rank sees the same counts, all reopen larger symmetric buffers at the
        # same frame and redo it - through the device-consumer call and through the host one, several times over
        for first_call in ("cull", "cull_to_host"):
            w2 = lib.World(rank)
            f2 = w2.add_entities(aabb[lo:hi], flags[lo:hi]); w2.set_world_aabbs(f2, aabb[lo:hi])
            small = ShardedWorld(w2, rank, size, lo, hi - lo, stream=torch.cuda.Stream(), exchange="p2p", p2p_capacity=64)
            assert small.exchange_mode == "p2p" and small.capacity == 64
            for i in range(6):
                planes_i = lib.view_planes(views[i %% 3])
                if (first_call == "cull") == (i %% 2 == 0):
                    blocks, counts = small.cull(planes_i)
                    got = merged_list(blocks, counts)
                else:
                    got = np.concatenate(small.cull_to_host(planes_i))
                assert np.array_equal(got, expected[i %% 3]), (rank, "regrow", first_call, i, len(got), len(expected[i %% 3]))
            assert small.generation >= 2 and small.capacity >= max(len(e) for e in expected) // size
            dist.barrier()
            w2.close()
    dist.barrier()
    w.close()
if rank == 0:
    print("MULTI_GPU_OK")
dist.destroy_process_group()
''' % ROOT


@pytest.mark.gpu
def test_two_gpu_exchange_modes(tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29533", str(script)],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "MULTI_GPU_OK" in r.stdout, (r.stdout + r.stderr)[-4000:]


ONE_GPU_WORKER = r'''
import os, sys, numpy as np
sys.path.insert(0, %r)
rank, size, name = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3]
from orion_b200 import lib, scenes
from orion_b200.sharded import SharedHostSegments, shard_range
from oracle import pyoracle
oracle = pyoracle.load("best")
n_total = 300_001
aabb, flags = scenes.static_boxes(n_total, 9, 800.0)
lo, hi = shard_range(n_total, rank, size)
w = lib.World(0)                                             # every shard on device 0: the box has one GPU
f = w.add_entities(aabb[lo:hi], flags[lo:hi]); w.set_world_aabbs(f, aabb[lo:hi]); w.set_id_base(lo)
views = [scenes.default_camera(), scenes.make_view((100, 0, 300), (1, 0, 0, 0), 60.0, 1.0, 900.0, 800, 600),
         scenes.make_view((-300, 50, 0), scenes.quat_yaw_pitch(70.0, 0.0), 90.0, 0.5, 700.0, 1600, 900)]
expected = [oracle.cull_aabb(aabb, oracle.view_planes(v), threads=4) for v in views]
seg = SharedHostSegments(rank, size, 1500, name)
cam = np.asarray(views[0]).reshape(1).copy()
step = seg.bind(w, cam)
for i in range(40):
    cam[:] = np.asarray(views[i %% 3]).reshape(1)
    parts = step()
    assert np.array_equal(np.concatenate(parts), expected[i %% 3]), (rank, i)
assert seg.generation >= 1
seg.close()
w.close()
print("SHARED_HOST_OK", rank)
''' % ROOT


@pytest.mark.gpu
def test_shared_host_segments_three_shards_on_one_gpu(tmp_path):
    """The host-consumer merge of a sharded world (sharded.SharedHostSegments) with three processes that each hold one
    shard on device 0: no kernel waits for another process, so it runs on a one-GPU box. Ids, order, regrow."""
    script = tmp_path / "worker.py"
    script.write_text(ONE_GPU_WORKER)
    name = "ovis_test1gpu_%d" % os.getpid()
    procs = [subprocess.Popen([sys.executable, str(script), str(r), "3", name], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
             for r in range(3)]
    outs = [p.communicate(timeout=600)[0] for p in procs]
    for r, (p, out) in enumerate(zip(procs, outs)):
        assert p.returncode == 0 and "SHARED_HOST_OK" in out, (r, out[-3000:])
