"""Entity set sharded by contiguous id range across the GPUs of one box.

One process per GPU. Rank r owns global ids [base_r, base_r + n_r); its world reports ids already
globalised (``World.set_id_base``). A cull is: local single-pass cull into ``block[1:]`` with the count in
``block[0]`` (one kernel), then ONE NCCL all-gather of the fixed-size blocks. The merged result is kept in
segmented form - segment r = ``blocks[r, 1:1+blocks[r, 0]]`` - and because every shard's compaction is
stable and shards are contiguous id ranges, reading the segments in rank order is exactly the reference's
single-world list order (render/src/simple_render_world.cc:38-41).

The block capacity follows the previous frames' counts (visible sets are temporally coherent); a frame whose
count exceeds it is redone with a larger block - results are never truncated silently.

The merge helpers work on any torch device/backend so the host logic is covered on CPU with gloo.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist


def allgather_blocks(block: torch.Tensor, world_size: int, group=None) -> torch.Tensor:
    """All ranks contribute one int32 block [count, id0, id1, ...]; returns (world_size, len(block))."""
    flat = torch.empty(world_size * block.numel(), dtype=block.dtype, device=block.device)
    if world_size == 1:
        flat.copy_(block)
    else:
        dist.all_gather_into_tensor(flat, block.contiguous(), group=group)
    return flat.view(world_size, block.numel())


def segment_counts(blocks: torch.Tensor) -> np.ndarray:
    return blocks[:, 0].cpu().numpy().astype(np.int64)


def merged_list(blocks: torch.Tensor, counts: np.ndarray | None = None) -> np.ndarray:
    """Host copy of the merged visible list (rank-order concatenation of the segments), uint32."""
    if counts is None:
        counts = segment_counts(blocks)
    parts = [blocks[r, 1:1 + int(c)] for r, c in enumerate(counts)]
    if not parts:
        return np.empty(0, dtype=np.uint32)
    return torch.cat(parts).cpu().numpy().view(np.uint32)


def shard_range(n_total: int, rank: int, world_size: int) -> tuple[int, int]:
    """Contiguous index range [r*N/G, (r+1)*N/G) of SURVEY.md section 8(e)."""
    return n_total * rank // world_size, n_total * (rank + 1) // world_size


P2P_BUFFERS = 3


def open_symmetric_blocks(world_size: int, capacity: int, group=None):
    """Three (world_size, capacity + 1) int32 buffers in NVLink-mapped symmetric memory, used in rotation.

    Returns (tensors, handles) or None when torch's symmetric memory is unavailable (then NCCL all-gather is used).
    """
    try:
        import torch.distributed._symmetric_memory as symm_mem
        group = group or dist.group.WORLD
        tensors, handles = [], []
        for _ in range(P2P_BUFFERS):
            t = symm_mem.empty((world_size, capacity + 1), dtype=torch.int32, device=torch.device("cuda", torch.cuda.current_device()))
            h = symm_mem.rendezvous(t, group=group)
            t.zero_()
            tensors.append(t)
            handles.append(h)
        return tensors, handles
    except Exception as e:                  # pragma: no cover - depends on the torch build / topology
        import warnings
        warnings.warn(f"symmetric memory unavailable ({e!r}); sharded cull falls back to NCCL all-gather")
        return None


class ShardedWorld:
    """A lib.World holding one shard plus the merge of the per-shard visible lists.

    exchange="nccl": local cull, then one NCCL all-gather of [count | ids] blocks.
    exchange="p2p":  the gather kernel itself stores this shard's list into every peer's segment buffer over NVLink
                     (ovis_cull_planes_device_multi). The symmetric-memory barrier that makes a frame readable is
                     enqueued on a side stream, so the cull stream goes straight on to the next frame; three buffers
                     rotate and a frame waits for the barrier of two frames ago (long finished) before it stores into a
                     buffer again, so stores never land in a buffer a peer may still be reading. `ready` is the event
                     after which the returned blocks are complete.
    """

    def __init__(self, world, rank: int, world_size: int, id_base: int, n_local: int, group=None,
                 stream: "torch.cuda.Stream | None" = None, exchange: str = "nccl", p2p_capacity: int | None = None):
        self.world = world
        self.rank, self.size, self.group = rank, world_size, group
        self.n_local = n_local
        self.stream = stream or torch.cuda.current_stream()
        world.set_stream(self.stream.cuda_stream)
        world.set_id_base(id_base)
        # The cull kernel bounds its writes by the slot count, so the send block has room for a full
        # shard; only the first `capacity + 1` words of it are exchanged.
        self.block = torch.zeros(n_local + 1, dtype=torch.int32, device="cuda")
        self.capacity = 0
        self.resize(max(1024, n_local // 8))
        self.exchange_mode = "nccl"
        self.frame = 0
        self._pinned = None
        self._last_max = 0
        if exchange == "p2p" and world_size > 1:
            cap = int(min(n_local, p2p_capacity or max(1024, n_local // 8)))
            opened = open_symmetric_blocks(world_size, cap, group)
            if opened is not None:
                self.sym_tensors, self.sym_handles = opened
                self.exchange_mode = "p2p"
                self.capacity = cap
                seg = (cap + 1) * 4
                # destination 0 is this rank's own copy; the rest are the peers' copies of segment `rank`
                order = [rank] + [p for p in range(world_size) if p != rank]
                self.side = torch.cuda.Stream()
                self.ready_events = [None] * P2P_BUFFERS
                self.p2p_out, self.p2p_cnt = [], []
                for h in self.sym_handles:
                    ptrs = [int(h.buffer_ptrs[p]) + rank * seg for p in order]
                    self.p2p_cnt.append(ptrs)
                    self.p2p_out.append([q + 4 for q in ptrs])

    def resize(self, capacity: int):
        self.capacity = int(min(max(capacity, 1), self.n_local))

    def cull_local(self, planes, require: int = 0):
        """Enqueue the local cull only (one kernel): count -> block[0], global ids -> block[1:]."""
        self.world.cull_planes_device(planes, self.block.data_ptr() + 4, self.block.data_ptr(), require)

    def exchange(self) -> torch.Tensor:
        """Enqueue the all-gather of the blocks on the stream. Returns (size, capacity + 1) int32."""
        with torch.cuda.stream(self.stream):
            return allgather_blocks(self.block[:self.capacity + 1], self.size, self.group)

    def cull_device(self, planes, require: int = 0) -> torch.Tensor:
        """Enqueue one frame's cull + exchange; returns the (size, capacity + 1) blocks tensor (valid in stream order)."""
        if self.exchange_mode == "p2p":
            k = self.frame % P2P_BUFFERS
            self.frame += 1
            # peers have finished reading buffer k (frame - 3) once the barrier of frame - 2 has passed
            old = self.ready_events[(k + 1) % P2P_BUFFERS]
            if old is not None:
                self.stream.wait_event(old)
            self.world.cull_planes_device_multi(planes, self.p2p_out[k], self.p2p_cnt[k], self.capacity, require)
            stored = torch.cuda.Event()
            stored.record(self.stream)
            self.side.wait_event(stored)
            with torch.cuda.stream(self.side):
                self.sym_handles[k].barrier(channel=0)
                ready = torch.cuda.Event()
                ready.record(self.side)
            self.ready_events[k] = ready
            self.ready = ready
            return self.sym_tensors[k]
        self.cull_local(planes, require)
        blocks = self.exchange()
        self.ready = torch.cuda.Event()
        self.ready.record(self.stream)
        return blocks

    def cull_to_host(self, planes, require: int = 0):
        """One frame end to end: cull + exchange, then the merged list in pinned host memory with ONE strided copy
        sized from the previous frame's counts (a frame that grew past it pays a second copy). Returns a list of
        per-rank uint32 arrays (views of the pinned buffer, valid until the next call); rank order = list order."""
        blocks = self.cull_device(planes, require)
        spec = int(min(self.capacity, self._last_max + self._last_max // 4 + 256))
        if self._pinned is None or self._pinned.shape[1] < self.capacity + 1:
            self._pinned = torch.empty((self.size, self.capacity + 1), dtype=torch.int32).pin_memory()
        with torch.cuda.stream(self.stream):
            self.stream.wait_event(self.ready)
            self._pinned[:, :spec + 1].copy_(blocks[:, :spec + 1], non_blocking=True)
        self.stream.synchronize()
        host = self._pinned.numpy()
        counts = host[:, 0].astype(np.int64)
        top = int(counts.max(initial=0))
        if top > self.capacity:
            if self.exchange_mode == "p2p":
                raise RuntimeError(f"visible list of {top} ids exceeds the symmetric segment capacity {self.capacity}")
            self.resize(int(top * 1.25) + 1024)
            return self.cull_to_host(planes, require)
        if top > spec:
            with torch.cuda.stream(self.stream):
                self._pinned[:, spec + 1:top + 1].copy_(blocks[:, spec + 1:top + 1], non_blocking=True)
            self.stream.synchronize()
        self._last_max = top
        return [host[r, 1:1 + int(counts[r])].view(np.uint32) for r in range(self.size)]

    def wait_all(self):
        """Block the host until every enqueued frame (cull stream and barrier stream) has finished."""
        self.stream.synchronize()
        if self.exchange_mode == "p2p":
            self.side.synchronize()

    def cull(self, planes, require: int = 0):
        """Synchronous cull; returns (blocks on device, per-rank counts on host). Redone if a shard overflowed."""
        while True:
            blocks = self.cull_device(planes, require)
            self.ready.synchronize()
            counts = segment_counts(blocks)
            if counts.max(initial=0) <= self.capacity:
                return blocks, counts
            if self.exchange_mode == "p2p":
                raise RuntimeError(f"visible list of {int(counts.max())} ids exceeds the symmetric segment capacity "
                                   f"{self.capacity}; create the ShardedWorld with a larger p2p_capacity")
            self.resize(int(counts.max() * 1.25) + 1024)
