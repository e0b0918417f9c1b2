"""N > 1 host logic on CPU: two gloo ranks each hold a contiguous shard, exchange [count | ids] blocks
with the same helpers the NCCL path uses, and the rank-order concatenation must equal the single-world
cull of the whole scene (the per-shard visible lists come from the oracle here: no GPU in this test)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from orion_b200 import scenes, sharded


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, size, port, n_total, capacity, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=size)
    from oracle import pyoracle
    oracle = pyoracle.load("port")
    aabb, _ = scenes.static_boxes(n_total, 5, 600.0)
    planes = oracle.view_planes(scenes.default_camera())
    lo, hi = sharded.shard_range(n_total, rank, size)
    local = oracle.cull_aabb(aabb[lo:hi], planes) + np.uint32(lo)          # what the shard's kernel emits (id_base = lo)
    block = torch.zeros(capacity + 1, dtype=torch.int32)
    block[0] = len(local)
    k = min(len(local), capacity)
    block[1:1 + k] = torch.from_numpy(local[:k].astype(np.int32))
    blocks = sharded.allgather_blocks(block, size)
    counts = sharded.segment_counts(blocks)
    merged = sharded.merged_list(blocks, np.minimum(counts, capacity))
    np.save(os.path.join(out_dir, f"merged{rank}.npy"), merged)
    np.save(os.path.join(out_dir, f"counts{rank}.npy"), counts)
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_merge_equals_single_world(tmp_path, port):
    n_total, size = 60_001, 2
    tcp = _free_port()
    mp.spawn(_worker, args=(size, tcp, n_total, 40_000, str(tmp_path)), nprocs=size, join=True)
    aabb, _ = scenes.static_boxes(n_total, 5, 600.0)
    exp = port.cull_aabb(aabb, port.view_planes(scenes.default_camera()))
    assert len(exp) > 100
    for r in range(size):
        assert np.array_equal(np.load(tmp_path / f"merged{r}.npy"), exp)
        assert np.load(tmp_path / f"counts{r}.npy").sum() == len(exp)


def test_overflow_is_detected_not_truncated(tmp_path):
    """A block smaller than a shard's count still carries the true count, so the caller can redo the frame."""
    n_total, size = 60_001, 2
    mp.spawn(_worker, args=(size, _free_port(), n_total, 16, str(tmp_path)), nprocs=size, join=True)
    counts = np.load(tmp_path / "counts0.npy")
    assert counts.max() > 16


def test_shard_ranges_tile_the_scene():
    for n in (0, 1, 7, 100_000_000):
        for g in (1, 2, 4, 8):
            edges = [sharded.shard_range(n, r, g) for r in range(g)]
            assert edges[0][0] == 0 and edges[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(edges, edges[1:]))


def _shared_host_worker(rank, size, name, n_total, capacity, frames, out_dir):
    """One rank of the shared-host exchange with the oracle standing in for the shard's kernel (plain shared memory: no
    CUDA in this test). Frame f uses camera f % 3, so the visible sets change and one of them overflows the block."""
    from oracle import pyoracle
    oracle = pyoracle.load("port")
    aabb, _ = scenes.static_boxes(n_total, 5, 600.0)
    lo, hi = sharded.shard_range(n_total, rank, size)
    cams = [scenes.default_camera(), scenes.make_view((0, 0, 900), (1, 0, 0, 0), 110.0, 0.1, 3000.0, 1600, 900),
            scenes.make_view((100, 0, 0), scenes.quat_yaw_pitch(90.0, 0.0), 40.0, 0.1, 500.0, 800, 600)]
    lists = [oracle.cull_aabb(aabb[lo:hi], oracle.view_planes(c)) + np.uint32(lo) for c in cams]
    seg = sharded.SharedHostSegments(rank, size, capacity, name, page_lock=False)
    cursor = {"f": 0}

    def make_culls():
        def make(k):
            def cull():
                ids = lists[cursor["f"] % 3]
                m = min(len(ids), seg.capacity)
                seg.sets[k, rank, 1:1 + m] = ids[:m]
                return len(ids)
            return cull
        return [make(0), make(1)]

    step = seg.bind_producer(make_culls)
    for f in range(frames):
        cursor["f"] = f
        parts = step()
        np.save(os.path.join(out_dir, f"r{rank}_f{f}.npy"), np.concatenate([np.asarray(p) for p in parts]))
    np.save(os.path.join(out_dir, f"gen{rank}.npy"), np.array([seg.generation, seg.capacity]))
    seg.close()


def test_shared_host_segments_two_ranks_with_regrow(tmp_path, port):
    n_total, size, frames = 40_001, 2, 7
    name = f"ovis_test_{os.getpid()}"
    mp.spawn(_shared_host_worker, args=(size, name, n_total, 2_000, frames, str(tmp_path)), nprocs=size, join=True)
    aabb, _ = scenes.static_boxes(n_total, 5, 600.0)
    cams = [scenes.default_camera(), scenes.make_view((0, 0, 900), (1, 0, 0, 0), 110.0, 0.1, 3000.0, 1600, 900),
            scenes.make_view((100, 0, 0), scenes.quat_yaw_pitch(90.0, 0.0), 40.0, 0.1, 500.0, 800, 600)]
    exp = [port.cull_aabb(aabb, port.view_planes(c)) for c in cams]
    assert max(len(e) for e in exp) > 2 * 2_000                 # the wide camera does not fit the first block
    for r in range(size):
        for f in range(frames):
            assert np.array_equal(np.load(tmp_path / f"r{r}_f{f}.npy"), exp[f % 3]), (r, f)
        gen, cap = np.load(tmp_path / f"gen{r}.npy")
        assert gen >= 1 and cap > 2_000                          # both ranks moved to the larger block
