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


class ShardedWorld:
    """A lib.World holding one shard plus the NCCL merge of the per-shard visible lists."""

    def __init__(self, world, rank: int, world_size: int, id_base: int, n_local: int, group=None,
                 stream: "torch.cuda.Stream | None" = None):
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
        self.cull_local(planes, require)
        return self.exchange()

    def cull(self, planes, require: int = 0):
        """Synchronous cull; returns (blocks on device, per-rank counts on host). Redone if a shard overflowed."""
        while True:
            blocks = self.cull_device(planes, require)
            self.stream.synchronize()
            counts = segment_counts(blocks)
            if counts.max(initial=0) <= self.capacity:
                return blocks, counts
            self.resize(int(counts.max() * 1.25) + 1024)
