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

    Also one (world_size,) int32 flag array per rank: flags[src] = last frame whose segment from rank `src` is complete
    here. Returns (tensors, handles, flags, flag_handle), or None when torch's symmetric memory is unavailable (then NCCL
    all-gather is used).
    """
    try:
        import torch.distributed._symmetric_memory as symm_mem
        group = group or dist.group.WORLD
        flags = symm_mem.empty((world_size,), dtype=torch.int32, device=torch.device("cuda", torch.cuda.current_device()))
        flag_handle = symm_mem.rendezvous(flags, group=group)
        flags.zero_()
        tensors, handles = [], []
        for _ in range(P2P_BUFFERS):
            t = symm_mem.empty((world_size, capacity + 1), dtype=torch.int32, device=torch.device("cuda", torch.cuda.current_device()))
            h = symm_mem.rendezvous(t, group=group)
            t.zero_()
            tensors.append(t)
            handles.append(h)
        torch.cuda.synchronize()
        dist.barrier(group=group)            # every rank's flags are zero before anyone signals
        return tensors, handles, flags, flag_handle
    except Exception as e:                  # pragma: no cover - depends on the torch build / topology
        import warnings
        warnings.warn(f"symmetric memory unavailable ({e!r}); sharded cull falls back to NCCL all-gather")
        return None


class SharedHostSegments:
    """The merged visible list for HOST consumers of a sharded world, one process per GPU: a POSIX shared-memory block
    every rank maps and page-locks (ovis_host_register). Rank r's gather kernel stores its ids straight into segment r of
    that block - the same single-GPU `ovis_cull` call, no GPU-to-GPU exchange, no reader kernel, one synchronise - and
    then publishes the frame number in its flag word; a frame is complete for a rank once every flag shows it.

    Layout (uint32 words): flags[size] | pad to 64 | 2 x size x (1 + capacity) segments [count | ids...]. Two segment
    sets alternate by frame: a rank refills set (f & 1) in frame f + 2 only after it has seen every rank finish frame
    f + 1, i.e. after every rank has consumed frame f. A frame in which some shard's list does not fit is redone by all
    ranks in a larger block (every rank sees the same counts, so every rank takes the same decision).
    """

    def __init__(self, rank: int, size: int, capacity: int, name: str, page_lock: bool = True):
        self.rank, self.size, self.name = rank, size, name
        self.page_lock = page_lock                       # False: plain shared memory (the CPU tests of the protocol)
        self.generation = 0
        self.frame = 0
        self.shm = None
        self._open(int(capacity))

    def _open(self, capacity: int):
        import time
        from multiprocessing import shared_memory
        from . import lib as _lib
        self._lib = _lib
        self.capacity = capacity
        self.seg_words = 1 + capacity
        words = 64 + 2 * self.size * self.seg_words
        name = f"{self.name}_g{self.generation}"
        if self.rank == 0:
            self.shm = shared_memory.SharedMemory(name=name, create=True, size=4 * words)
            self.words = np.frombuffer(self.shm.buf, dtype=np.uint32, count=words)
            self.words[:64] = 0
            self.words[:self.size] = self.frame          # the frames so far count as delivered in the new block
            self.words[63] = 0x600DF00D                  # the block is initialised: the other ranks may use it
        else:
            t0 = time.perf_counter()
            while True:
                try:
                    self.shm = shared_memory.SharedMemory(name=name, create=False)
                    try:                                       # rank 0 owns the name; this process only maps it
                        from multiprocessing import resource_tracker
                        resource_tracker.unregister(self.shm._name, "shared_memory")
                    except Exception:
                        pass
                    if self.shm.size >= 4 * words and int(np.frombuffer(self.shm.buf, dtype=np.uint32, count=64)[63]) == 0x600DF00D:
                        self.words = np.frombuffer(self.shm.buf, dtype=np.uint32, count=words)
                        break
                    self.shm.close()
                except FileNotFoundError:
                    pass
                if time.perf_counter() - t0 > 30.0:
                    raise RuntimeError(f"shared result block {name} never appeared")
                time.sleep(0.0005)
        self.flags = self.words[:self.size]
        self.sets = self.words[64:].reshape(2, self.size, self.seg_words)
        if self.page_lock:
            _lib.host_register(self.words)

    def _release(self, unlink_after: bool = False):
        """Unmap the block. Rank 0 owns the name: it unlinks it (the other ranks' mappings stay valid until they close;
        on a regrow they have all seen the overflowing frame's flags, and none of them touches the old block again)."""
        if self.shm is None:
            return
        try:
            if self.page_lock:
                self._lib.host_unregister(self.words)
        except Exception:
            pass
        self.flags = self.sets = self.words = None
        shm, self.shm = self.shm, None
        try:
            if self.rank == 0:
                shm.unlink()
            shm.close()
        except BufferError:
            pass                                          # a caller still holds a list of an earlier frame: left to the GC
        except Exception:
            pass

    def bind(self, world, view):
        """Per-rank frame function: step() -> list of per-rank id arrays (views into the shared block, valid until the
        call after next)."""
        import ctypes as C
        import time
        lib = self._lib
        view = np.ascontiguousarray(view, dtype=lib.VIEW_DTYPE).reshape(1)
        fn, h = world.lib.ovis_cull, world.h
        ne, nl = C.c_int64(), C.c_int32()

        def make_culls():
            def make(k):
                args = (h, view.ctypes.data_as(C.c_void_p), C.c_uint32(0), self.sets[k, self.rank, 1:].ctypes.data_as(C.c_void_p),
                        C.c_int64(self.capacity), C.byref(ne), None, C.c_int32(0), C.byref(nl))

                def cull():
                    rc = fn(*args)
                    if rc not in (lib.OVIS_OK, lib.OVIS_ERR_CAPACITY):
                        lib._check(rc)
                    return ne.value
                return cull
            return [make(0), make(1)]

        step = self.bind_producer(make_culls)
        step.keep = (view, ne, nl)
        return step

    def bind_producer(self, make_culls):
        """The exchange protocol around any producer: make_culls() -> [cull0, cull1], where cull_k() fills
        self.sets[k, rank, 1:1 + capacity] and returns the TRUE number of ids (also when they did not all fit). Called
        again after the block was regrown."""
        import time
        lib = self._lib
        state = {"culls": make_culls()}

        import ctypes as C
        wait = lib.load().ovis_host_flags_wait
        n_ranks = C.c_int32(self.size)

        def step(timeout_s: float = 20.0):
            while True:
                flags_ptr = C.c_void_p(self.flags.ctypes.data)      # (the block may have been regrown)
                self.frame += 1
                f, k = self.frame, self.frame & 1
                self.sets[k, self.rank, 0] = state["culls"][k]()   # the true count, also when it did not fit
                self.flags[self.rank] = f                     # after the ids and the count (x86 keeps the store order)
                # every rank has delivered frame f (spin in C: a Python poll loop costs microseconds per look)
                if wait(flags_ptr, n_ranks, C.c_uint32(f), C.c_int64(int(timeout_s * 1e6))) != 0:
                    raise RuntimeError("a rank did not deliver its segment")
                counts = self.sets[k, :, 0].astype(np.int64)
                if int(counts.max()) <= self.capacity:
                    return [self.sets[k, r, 1:1 + int(counts[r])] for r in range(self.size)]
                # some shard overflowed: every rank sees that and moves to a larger block, then the frame is redone
                need = int(counts.max() * 1.25) + 1024
                del counts
                state["culls"] = None                        # (they hold pointers into the old block)
                self._release(unlink_after=True)
                self.generation += 1
                self._open(need)
                state["culls"] = make_culls()

        return step

    def close(self):
        self._release()


class ShardedWorld:
    """A lib.World holding one shard plus the merge of the per-shard visible lists.

    exchange="nccl": local cull, then one NCCL all-gather of [count | ids] blocks.
    exchange="p2p":  the gather kernel itself stores this shard's list into every peer's segment buffer over NVLink and,
                     when its last block has finished, writes the frame number into its slot of every rank's flag
                     array (ovis_cull_planes_device_exchange). No barrier and no collective is launched: a frame is
                     two kernels, exactly as on one GPU, so consecutive frames keep overlapping. Three segment buffers
                     rotate; before refilling one, the gather kernel waits until every rank has signalled the frame
                     two before the current one - which a rank only does after everything it had enqueued on the
                     frame being overwritten. Readers enqueue `enqueue_ready_wait()` (one tiny kernel) first.
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
        self.generation = 0                                      # symmetric blocks opened so far (p2p): > 1 after a regrow
        if exchange == "p2p" and world_size > 1:
            self._open_p2p(int(min(max(n_local, 1), p2p_capacity or max(1024, n_local // 8))))

    def _open_p2p(self, cap: int) -> bool:
        """(Re)open the symmetric segment buffers with room for `cap` ids per shard. Collective: every rank calls it at
        the same frame - they all see the same counts, so they all take the same decision."""
        opened = open_symmetric_blocks(self.size, cap, self.group)
        if opened is None:
            return False
        rank, world_size, world = self.rank, self.size, self.world
        self.sym_tensors, self.sym_handles, self.sym_flags, self.flag_handle = opened
        self.exchange_mode = "p2p"
        self.capacity = cap
        self.frame = 0                                           # the new flag words start at 0: frames are numbered afresh
        self.generation += 1
        self._pinned = None
        seg = (cap + 1) * 4
        # destination 0 is this rank's own copy; the rest are the peers' copies of segment `rank`
        order = [rank] + [p for p in range(world_size) if p != rank]
        self.p2p_out, self.p2p_cnt = [], []
        for h in self.sym_handles:
            ptrs = [int(h.buffer_ptrs[p]) + rank * seg for p in order]
            self.p2p_cnt.append(ptrs)
            self.p2p_out.append([q + 4 for q in ptrs])
        self.flag_local = int(self.flag_handle.buffer_ptrs[rank])
        self.flag_signal = [int(self.flag_handle.buffer_ptrs[p]) + rank * 4 for p in order]
        self.plans = [world.exchange_plan(self.p2p_out[k], self.p2p_cnt[k], self.flag_signal) for k in range(P2P_BUFFERS)]
        return True

    def _regrow_p2p(self, top: int):
        """Some shard's list did not fit its segment (the kernels cut the ids at the capacity but deliver the full count, to
        every rank alike). Every rank has waited for every flag of that frame, so nobody is still storing into anybody's
        old buffers; all move to larger ones and the frame is redone."""
        self.stream.synchronize()
        old = (self.sym_tensors, self.sym_handles, self.sym_flags, self.flag_handle)
        cap = int(min(max(self.n_local_max(), 1), top + top // 4 + 1024))
        if not self._open_p2p(max(cap, top)):
            raise RuntimeError("symmetric memory could not be reopened for a larger segment")
        del old

    def n_local_max(self) -> int:
        """The largest shard's size (segments are equally sized on every rank): shards differ by at most one entity."""
        return self.n_local + 1

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
            self.frame += 1
            f = self.frame                                       # frames are numbered from 1
            k = f % P2P_BUFFERS
            planes = np.ascontiguousarray(planes, dtype=np.float32).reshape(24)
            self.world.cull_exchange_planned(planes, self.plans[k], self.capacity, self.flag_local, self.size,
                                             max(f - 2, 0), f, require)
            return self.sym_tensors[k]
        self.cull_local(planes, require)
        return self.exchange()

    def enqueue_ready_wait(self):
        """Make the stream wait until the latest frame's segments from every rank are complete in local memory."""
        if self.exchange_mode == "p2p":
            self.world.wait_flags(self.flag_local, self.size, self.frame)

    def cull_to_host(self, planes, require: int = 0):
        """One frame end to end: cull + exchange, then the merged list in pinned host memory. Returns a list of per-rank
        uint32 arrays (views of the pinned buffer, valid until the next call); rank order = list order.

        p2p: two library calls - the cull with the fused exchange, and ovis_read_segments_host, whose single kernel waits
        for each rank's ready flag and stores that segment's count and ids straight into the pinned buffer.
        nccl: all-gather, then ONE strided copy sized from the previous frame's counts (a frame that grew past it pays
        a second copy)."""
        from . import lib as _lib
        blocks = self.cull_device(planes, require)
        words = self.capacity + 1
        if self.exchange_mode == "p2p":
            if self._pinned is None or self._pinned.array.size < self.size * words:
                self._pinned = _lib.PinnedBuffer(self.size * words)
            host = self._pinned.array[:self.size * words].reshape(self.size, words)
            self.world.read_segments_host(blocks.data_ptr(), words, self.size, self._pinned.array, self.flag_local, self.frame)
            counts = host[:, 0].astype(np.int64)
            top = int(counts.max(initial=0))
            if top > self.capacity:                              # every rank reads the same counts: all regrow, all redo
                self._regrow_p2p(top)
                return self.cull_to_host(planes, require)
            return [host[r, 1:1 + int(counts[r])] for r in range(self.size)]
        spec = int(min(self.capacity, self._last_max + self._last_max // 4 + 256))
        if self._pinned is None or self._pinned.shape[1] < words:
            self._pinned = torch.empty((self.size, words), dtype=torch.int32).pin_memory()
        with torch.cuda.stream(self.stream):
            self._pinned[:, :spec + 1].copy_(blocks[:, :spec + 1], non_blocking=True)
        self.stream.synchronize()
        host = self._pinned.numpy()
        counts = host[:, 0].astype(np.int64)
        top = int(counts.max(initial=0))
        if top > self.capacity:
            self.resize(int(top * 1.25) + 1024)
            self._pinned = None
            return self.cull_to_host(planes, require)
        if top > spec:
            with torch.cuda.stream(self.stream):
                self._pinned[:, spec + 1:top + 1].copy_(blocks[:, spec + 1:top + 1], non_blocking=True)
            self.stream.synchronize()
        self._last_max = top
        return [host[r, 1:1 + int(counts[r])].view(np.uint32) for r in range(self.size)]

    def wait_all(self):
        """Block the host until every enqueued frame (cull stream and barrier stream) has finished."""
        self.enqueue_ready_wait()
        self.stream.synchronize()

    def cull(self, planes, require: int = 0):
        """Synchronous cull; returns (blocks on device, per-rank counts on host). Redone if a shard overflowed."""
        while True:
            blocks = self.cull_device(planes, require)
            self.enqueue_ready_wait()
            self.stream.synchronize()
            if self.exchange_mode == "p2p" and self.world.exchange_error():
                raise RuntimeError("sharded exchange timed out waiting for a peer's ready flag")
            counts = segment_counts(blocks)
            if counts.max(initial=0) <= self.capacity:
                return blocks, counts
            if self.exchange_mode == "p2p":
                self._regrow_p2p(int(counts.max()))
                continue
            self.resize(int(counts.max() * 1.25) + 1024)
