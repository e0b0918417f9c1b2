#!/usr/bin/env python
"""bench.py - entity-view cull throughput of the B200 scene-visibility path (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W            # ours (CUDA, through the C ABI)
    python bench.py --impl reference --gpus N --steps K --warmup W   # the reference's CPU cull, host cores

Workload (config.workload): BASELINE.json configs[4] sharded by contiguous id range with a FIXED shard of
12.5 M entities per GPU (100 M at 8 GPUs; weak scaling), entities generated like configs[1]/[4]
(SURVEY.md section 8(d)), one camera frustum. A step = one frame's cull of the whole (sharded) entity set:
one streaming test kernel + one small gather kernel per GPU and, for N > 1, one NCCL all-gather of the compacted per-shard lists.
configs[1] itself (1 M entities, 24 MB: L2-resident and launch-latency-bound) is reported beside it under
"also" at N = 1, cold (L2 flushed between iterations) and warm.

One JSON line on stdout (rank 0). Timing: CUDA events on the launching stream, barrier + synchronize on
both sides, max over ranks. The 300 MB shard is larger than the 126 MB L2, so every step streams from HBM.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from orion_b200 import scenes  # noqa: E402

SHARD_ENTITIES = 12_500_000
SCENE_SEED = 4
EXTENT = 4000.0
METRIC = "entity-view cull tests/sec"
UNIT = "tests/s"


def shard_scene(shard: int, n: int):
    """Shard `shard` of the sharded configs[4] scene: every shard has its own seed so any rank can build its part."""
    return scenes.static_boxes(n, SCENE_SEED * 1000 + shard, EXTENT)


def workload_name(n_gpus: int, n: int) -> str:
    return (f"configs[4] sharded: {n / 1e6:g}M entities/GPU x {n_gpus} GPU = {n * n_gpus / 1e6:g}M entities, "
            f"1 camera frustum (hfov 75, 16:9, far 1000), config-2 distribution in [-{EXTENT:g},{EXTENT:g}]^3")


CACHE_NOTE = "inputs larger than L2 (300 MB of bounds per GPU vs 126 MB L2); no flush needed"


def bench_config(n_gpus: int, n: int, visible_total: int) -> dict:
    """`config` of the JSON line - the same keys and values in both arms (ours and --impl reference)."""
    return {"workload": workload_name(n_gpus, n), "entities_total": n * n_gpus, "views": 1, "visible_total": int(visible_total),
            "cache": CACHE_NOTE if n * 24 > 126e6 else "inputs fit L2 (sizes below the benched one are for debugging only)"}


def host_cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


# --------------------------------------------------------------------------------------------------
# clocks
# --------------------------------------------------------------------------------------------------

class ClockSampler(threading.Thread):
    """Samples SM clock and clock-event reasons through NVML while the timed region runs."""
    REASONS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown", 0x2: "applications_clocks_setting"}

    def __init__(self, device_index: int, period: float = 0.01):
        super().__init__(daemon=True)
        self.period = period
        self.samples = []          # (t, sm_mhz, reasons)
        self.stop_flag = threading.Event()
        self.ok = False
        try:
            import pynvml
            import torch
            pynvml.nvmlInit()
            uuid = str(torch.cuda.get_device_properties(device_index).uuid)
            if not uuid.startswith("GPU-"):
                uuid = "GPU-" + uuid
            try:
                self.h = pynvml.nvmlDeviceGetHandleByUUID(uuid)
            except Exception:
                self.h = pynvml.nvmlDeviceGetHandleByIndex(device_index)
            self.nv = pynvml
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception as e:   # NVML missing: report that instead of inventing numbers
            self.error = repr(e)

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        while not self.stop_flag.is_set():
            try:
                mhz = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                try:
                    reasons = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    reasons = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                self.samples.append((time.perf_counter(), mhz, reasons))
            except Exception:
                pass
            time.sleep(self.period)

    def count_between(self, t0: float, t1: float) -> int:
        return len([s for s in self.samples if t0 <= s[0] <= t1])

    def summary(self, t0: float, t1: float, note: str):
        if not self.ok:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "error": getattr(self, "error", "nvml unavailable")}
        sel = [s for s in self.samples if t0 <= s[0] <= t1]
        reasons = set()
        for _, _, r in sel:
            for bit, name in self.REASONS.items():
                if r & bit:
                    reasons.add(name)
        mhz = float(np.median([s[1] for s in sel])) if sel else None
        return {"sm_mhz": mhz, "sm_max_mhz": float(self.max_mhz), "reasons": sorted(reasons), "samples": len(sel),
                "sampled_over": note}


# --------------------------------------------------------------------------------------------------
# reference arm: the reference's CPU cull on the host cores
# --------------------------------------------------------------------------------------------------

def cpu_cull_timing(oracle, aabb, planes, threads: int, steps: int, warmup: int):
    out = None
    for _ in range(warmup):
        out = oracle.cull_aabb(aabb, planes, threads=threads)
    times = []
    for _ in range(steps):
        t0 = time.perf_counter()
        out = oracle.cull_aabb(aabb, planes, threads=threads)
        times.append(time.perf_counter() - t0)
    return np.array(times), out


def as_shipped_timing(oracle, aabb, view, iters=5):
    """SimpleRenderWorld::cull over std::list<RenderEntity*> exactly as shipped, 1 thread (tier A)."""
    if not oracle.has_world:
        return None
    w = oracle.world()
    n = len(aabb)
    w.add_entities(np.tile(scenes.IDENTITY_TRS, (n, 1)), aabb, None)
    ms, vis = w.time_cull(view, flags=0, iters=iters)
    w.abandon()     # entity-by-entity teardown is O(n^2) in the reference
    return {"entities": n, "ms_per_frame": float(np.median(ms)), "tests_per_s": n / (float(np.median(ms)) * 1e-3),
            "visible": int(vis), "threads": 1}


def run_reference(args, rank: int):
    """The reference's CPU cull of the SAME work as our arm: every one of the N shards, every step, on this host's cores."""
    if rank != 0:
        return
    from oracle import pyoracle
    oracle = pyoracle.load("best")
    cores = host_cores()
    n = args.entities_per_gpu
    shards = [shard_scene(r, n)[0] for r in range(args.gpus)]
    view = scenes.default_camera()
    planes = oracle.view_planes(view)
    visible = 0
    for _ in range(max(args.warmup, 1)):
        visible = sum(len(oracle.cull_aabb(a, planes, threads=cores)) for a in shards)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        for a in shards:
            oracle.cull_aabb(a, planes, threads=cores)
    total = time.perf_counter() - t0
    value = n * args.gpus * args.steps / total
    shipped = as_shipped_timing(oracle, shards[0][:1_000_000], view) if not args.skip_extras else None
    how = ("the arithmetic is the reference's own Math::intersect compiled from its sources, driven over flat AABB arrays"
           if oracle.name == "reference" else "C restatement of the reference")
    sample = (f"all {args.gpus} shard(s) of {n / 1e6:g}M entities, 1 frustum, per step ({n * args.gpus / 1e6:g}M entity-view tests); "
              f"{how} by {cores} host threads")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": total / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": bench_config(args.gpus, n, visible),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": oracle.name, "sample": sample,
                         "visible": int(visible), "as_shipped_std_list_1thread": shipped},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# --------------------------------------------------------------------------------------------------
# ours
# --------------------------------------------------------------------------------------------------

def ncu_traffic_bytes():
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch of the dominant kernel, from the committed
    `ncu --set full` summary of this same workload (profiles/r02_cull_test_kernel_ncu.csv); None if absent."""
    path = os.path.join(ROOT, "profiles", "r02_cull_test_kernel_ncu.csv")
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    try:
        total, seen = 0.0, set()
        for line in open(path):
            parts = [x.strip('"') for x in line.strip().split(",")]
            if len(parts) == 3 and parts[0] in ("dram__bytes_read.sum", "dram__bytes_write.sum") and parts[0] not in seen:
                total += float(parts[2]) * scale[parts[1]]
                seen.add(parts[0])
        return total if len(seen) == 2 else None
    except Exception:
        return None


def time_kernel_loop(torch, stream, fn, iters: int, flush=None):
    """Per-iteration device times (ms) of fn() with CUDA events on `stream`."""
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    with torch.cuda.stream(stream):
        for a, b in evs:
            if flush is not None:
                flush()
            a.record(stream)
            fn()
            b.record(stream)
    stream.synchronize()
    return np.array([a.elapsed_time(b) for a, b in evs])


def config2_extra(torch, lib, stream, scratch, peak_gbs):
    """configs[1]: 1M static entities, 1 frustum, 1 GPU - cold (L2 flushed) and warm, kernel only + e2e."""
    s = scenes.config2()
    n = len(s["aabb"])
    w = lib.World(torch.cuda.current_device())
    w.set_stream(stream.cuda_stream)
    first = w.add_entities(s["aabb"], s["flags"])
    w.set_world_aabbs(first, s["aabb"])
    planes = lib.view_planes(s["view"])
    out = torch.zeros(n + 1, dtype=torch.int32, device="cuda")

    def fn():
        w.cull_planes_device(planes, out.data_ptr() + 4, out.data_ptr())

    time_kernel_loop(torch, stream, fn, 5)
    cold = time_kernel_loop(torch, stream, fn, 30, flush=lambda: scratch.add_(1))   # 512 MB r+w evicts the 126 MB L2
    warm = time_kernel_loop(torch, stream, fn, 200)
    vis = int(out[0].item())
    pinned = lib.PinnedBuffer(n)
    frame = w.bind_cull(s["view"], pinned.array, flags=0)
    for _ in range(3):
        frame()
    t0 = time.perf_counter()
    reps = 50
    for _ in range(reps):
        frame()
    e2e_ms = (time.perf_counter() - t0) / reps * 1e3
    w.close()
    bytes_alg = 24 * n + 4 * vis

    def pack(ms):
        return {"ms_per_frame": ms, "tests_per_s": n / (ms * 1e-3), "achieved_gbs": bytes_alg / (ms * 1e-3) / 1e9,
                "frac_of_hbm_peak": bytes_alg / (ms * 1e-3) / 1e9 / peak_gbs}

    return {"workload": "configs[1]: synthetic 1M static entities, 1 camera frustum, 1 GPU", "entities": n, "visible": vis,
            "cold_l2_flushed": pack(float(np.median(cold))), "warm_l2_resident": pack(float(np.median(warm))),
            "e2e_host_api": {"ms_per_frame": e2e_ms, "tests_per_s": n / (e2e_ms * 1e-3)},
            "note": "24 MB of bounds = 3.7 us of HBM traffic at the measured peak: this size is launch-latency bound"}


def run_ours(args, rank: int, local_rank: int, world_size: int):
    import torch
    import torch.distributed as dist
    from orion_b200 import lib
    from orion_b200.sharded import ShardedWorld, merged_list

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device - the product path has no CPU fallback "
                         "(use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    pinned_cpus = 0
    if world_size > 1 and os.environ.get("OVIS_BENCH_PIN", "1") != "0":
        # one process per GPU: keep this rank on its GPU's socket (torchrun places ranks anywhere); before the world and its
        # page-locked buffers exist, so that they are allocated there too
        try:
            pinned_cpus = lib.pin_thread_near_device(local_rank)
        except Exception:                                        # placement is an optimisation: never let it stop the run
            pinned_cpus = 0
    if world_size > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak_gbs = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "MEASURED_PEAKS.json hbm_gbs (measured)" if "hbm_gbs" in peaks else "6650 GB/s fallback (B200_PROFILING.md)"

    n = args.entities_per_gpu
    aabb, flags = shard_scene(rank, n)
    view = scenes.default_camera()
    planes = lib.view_planes(view)

    stream = torch.cuda.Stream()
    world = lib.World(local_rank)
    world.reserve(n_entities=n)
    first = world.add_entities(aabb, flags)
    world.set_world_aabbs(first, aabb)
    sw = ShardedWorld(world, rank, world_size, rank * n, n, stream=stream, exchange=args.exchange)

    def barrier():
        torch.cuda.synchronize()
        if world_size > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- warm-up (also sizes the exchange block from the real counts)
    blocks, counts = sw.cull(planes)
    if sw.exchange_mode == "nccl":
        sw.resize(int(counts.max() * 1.25) + 1024)
    for _ in range(max(args.warmup, 3)):
        sw.cull_device(planes)
    barrier()

    sampler = ClockSampler(local_rank)
    sampler.start()
    time.sleep(0.05)

    # ---- timed region: exactly K steps, device-resident inputs
    launches0 = world.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t_wall0 = time.perf_counter()
    with torch.cuda.stream(stream):
        ev0.record(stream)
        for i in range(args.steps):
            if world_size > 1:
                blocks = sw.cull_device(planes)        # cull + exchange (fused P2P stores, or NCCL all-gather)
            else:
                sw.cull_local(planes)
        if world_size > 1:
            sw.enqueue_ready_wait()                    # the last frame is only done when every peer's segment has landed
        ev1.record(stream)
    barrier()
    t_wall1 = time.perf_counter()
    launches = world.launch_count() - launches0
    elapsed_ms = ev0.elapsed_time(ev1)
    t = torch.tensor([elapsed_ms], dtype=torch.float64, device="cuda")
    if world_size > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    elapsed_ms = float(t.item())

    # per-kernel durations for the roofline: the same K steps once more, with the library recording CUDA events around
    # each kernel on the launching stream (kept out of the timed region above because an event between two kernels
    # stops the next test kernel from overlapping the previous gather kernel)
    world.set_timing(True)
    with torch.cuda.stream(stream):
        for i in range(args.steps):
            sw.cull_local(planes)
    test_ms, gather_ms = world.read_timing()
    world.set_timing(False)

    # clocks: the timed region can be shorter than a few NVML samples; soak with the identical loop if so
    note = "timed region"
    if sampler.count_between(t_wall0, t_wall1) < 3:
        t_soak = time.perf_counter()
        while time.perf_counter() - t_soak < 1.0:
            for _ in range(50):
                sw.cull_local(planes)
            stream.synchronize()
        t_wall1 = time.perf_counter()
        note = "timed region + 1 s of the identical untimed loop (region shorter than 3 NVML samples)"
    clocks = sampler.summary(t_wall0, t_wall1, note)

    # ---- e2e: the call a user makes, host buffers, copies inside the timed region
    blocks, counts = sw.cull(planes)
    total_visible = int(counts.sum())
    pinned = lib.PinnedBuffer(max(total_visible * 2, 1024))
    if world_size == 1:
        cull_frame = world.bind_cull(view, pinned.array, flags=0)           # ovis_cull: view in, ids out (pinned host)

        def e2e_step():
            return cull_frame()[0]
        d2h = 8 + 4 * total_visible
    else:
        # Host consumers of a sharded world need no GPU-to-GPU exchange: every rank's gather kernel stores its segment
        # straight into one page-locked shared-memory block all ranks map (sharded.SharedHostSegments); a frame is one
        # ovis_cull per rank plus a flag word per rank.
        from orion_b200.sharded import SharedHostSegments
        seg = SharedHostSegments(rank, world_size, max(4096, int(counts.max()) * 2), f"ovis_bench_{os.environ.get('MASTER_PORT', '0')}_{os.getppid()}")
        e2e_frame = seg.bind(world, view)
        e2e_parts = [None]

        def e2e_step():
            e2e_parts[0] = e2e_frame()
            return int(sum(len(p) for p in e2e_parts[0]))
        d2h = 4 * total_visible + 4 * world_size
    for _ in range(3):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    got = 0
    for _ in range(args.steps):
        got = e2e_step()
    barrier()
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
    if world_size > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_s = float(t.item())
    assert got == total_visible
    e2e_ids = np.concatenate([np.asarray(p) for p in e2e_parts[0]]).copy() if world_size > 1 else pinned.array[:got].copy()
    if world_size > 1:
        barrier()
        e2e_parts[0] = None                          # the last frame's lists are views into the shared block: let go of them,
        e2e_frame = None                             # and of the frame function that holds the block's arrays, before unmapping
        seg.close()
    sampler.stop_flag.set()

    total_entities = n * world_size
    value = total_entities * args.steps / (elapsed_ms * 1e-3)
    local_visible = int(counts[rank])
    # algorithmic bytes: 24 B per entity read once + 4 B per emitted id
    alg_bytes = 24 * n + 4 * local_visible
    k_ms = float(test_ms.mean())
    g_ms = float(gather_ms.mean())
    achieved = alg_bytes / (k_ms * 1e-3) / 1e9
    step_achieved = alg_bytes / ((k_ms + g_ms) * 1e-3) / 1e9
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world_size, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": elapsed_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": bench_config(world_size, n, total_visible),
        "parallelism": ((f"entity shards x{world_size}, exchange of visible lists: " +
                         ("gather kernel stores to peers over NVLink + ready flags, no collective" if sw.exchange_mode == "p2p"
                          else "NCCL all-gather")) if world_size > 1 else "1 GPU"),
        "cull_ms_per_frame": elapsed_ms / args.steps,
        "host_placement": ({"rank0_pinned_to_cpus_next_to_its_gpu": pinned_cpus,
                            "how": "ovis_host_pin_thread_near_device (PCI local_cpulist); OVIS_BENCH_PIN=0 turns it off"}
                           if world_size > 1 else None),
        "roofline": {"bound": "hbm", "kernel": "ovis::cull_test_kernel<1,false,5> (dominant: reads every bound once)",
                     "achieved": achieved, "peak": peak_gbs, "unit": "GB/s", "frac": achieved / peak_gbs,
                     "traffic": ncu_traffic_bytes() if n == SHARD_ENTITIES else None,
                     "traffic_source": "profiles/r02_cull_test_kernel_ncu.csv (ncu --set full, one launch, same workload)",
                     "algorithmic_bytes_per_launch": alg_bytes, "kernel_ms_mean": k_ms, "kernel_ms_min": float(test_ms.min()),
                     "gather_kernel_ms_mean": g_ms,
                     "both_kernels": {"achieved": step_achieved, "frac": step_achieved / peak_gbs,
                                      "note": "same bytes over test + gather kernel time: the whole cull"},
                     "peak_source": peak_src, "formula": "24 B/entity + 4 B/visible id (SURVEY.md 8(d))",
                     "timed_with": "cudaEvents recorded by the library on the launching stream around each kernel, over the same K steps repeated right after the timed region"},
        "e2e": {"value": total_entities * args.steps / e2e_s, "unit": UNIT, "ms_per_step": e2e_s / args.steps * 1e3,
                "h2d_bytes_per_step": 96, "d2h_bytes_per_step": d2h,
                "what": "ovis_cull(view) -> visible ids in pinned host memory; the world is device-resident between frames, "
                        "like the reference's SimpleRenderWorld is host-resident; per-frame inputs are the view (planes ride "
                        "in the kernel parameters)" + ("; at N > 1 every rank's kernel stores its segment into one shared "
                        "page-locked block, a frame ends when every rank's flag shows it" if world_size > 1 else "")},
        "gpu_launches": int(launches),
        "clocks": clocks,
    }

    if rank == 0 and not args.skip_extras:
        # parity of the benched result: the merged list of ALL shards against the CPU oracle's cull of all shards
        from oracle import pyoracle
        oracle = pyoracle.load("best")
        cores = host_cores()
        got_ids = merged_list(blocks, counts)
        exp_parts, times = [], None
        for r in range(world_size):
            shard = aabb if r == rank else shard_scene(r, n)[0]
            if r == 0 and world_size == 1:
                times, e = cpu_cull_timing(oracle, shard, planes, cores, 5, 1)
            else:
                e = oracle.cull_aabb(shard, planes, threads=cores)
            exp_parts.append(e.astype(np.uint64) + np.uint64(r * n))
        exp = np.concatenate(exp_parts).astype(np.uint32)
        line["parity"] = {"bit_exact_vs_oracle": bool(np.array_equal(got_ids, exp)) and bool(np.array_equal(e2e_ids, exp)),
                          "device_list_exact": bool(np.array_equal(got_ids, exp)), "e2e_host_list_exact": bool(np.array_equal(e2e_ids, exp)),
                          "oracle": oracle.name, "visible": int(len(exp)), "shards_checked": world_size}
    if rank == 0 and world_size == 1 and not args.skip_extras:
        # CPU baseline on the same shard, same host
        t1, _ = cpu_cull_timing(oracle, aabb, planes, 1, 2, 1)
        line["cpu_baseline"] = {
            "value": n / float(np.median(times)), "unit": UNIT, "cores": cores, "kind": oracle.name,
            "sample": f"the whole benched shard ({n / 1e6:g}M entities, 1 frustum), 5 timed passes, flat AABB arrays, "
                      f"{cores} host threads over contiguous index ranges",
            "ms_per_frame": float(np.median(times)) * 1e3,
            "one_thread_flat": {"tests_per_s": n / float(np.median(t1)), "ms_per_frame": float(np.median(t1)) * 1e3},
            "as_shipped_std_list_1thread": as_shipped_timing(oracle, aabb[:1_000_000], view),
        }
        scratch = torch.zeros(128 * 1024 * 1024, dtype=torch.int32, device="cuda")

        def masked():
            sw.cull_local(planes, 1)

        time_kernel_loop(torch, stream, masked, 3)
        mk = time_kernel_loop(torch, stream, masked, 20)
        vis_m = int(sw.block[0].item())
        with torch.cuda.stream(stream):
            line["also"] = {
                "masked_cull_same_shard": {"kernel": "masked test kernel + gather", "ms_mean": float(mk.mean()),
                                           "achieved_gbs": (24 * n + n / 8 + 4 * vis_m) / (float(mk.mean()) * 1e-3) / 1e9},
                "config2_1M": config2_extra(torch, lib, stream, scratch, peak_gbs),
            }
        del scratch
        # the other BASELINE.json configs, each with parity against the oracle and its own CPU baseline (tools/bench_configs.py)
        world.close()
        world = None
        del sw
        torch.cuda.empty_cache()
        from tools import bench_configs, bench_raster, render_frame

        def whole_frame():
            _, covered, ms = render_frame.render(width=1920, height=1080, timing=True, textured=True)
            return {"what": "tools/render_frame.py's built-in scene at 1920 x 1080 (61 entities, spot + two point lights with 512^2 shadow maps, "
                            "textured floor): cull -> shadow lists -> 13 shadow-map passes -> G-buffer -> light pass -> tonemap + gamma",
                    "ms_per_frame": ms, "pixels_covered": covered,
                    "note": "synchronous host loop through the Python binding; every rasteriser call sizes its scratch from device counts"}

        def size_sweep():
            from tools import size_sweep as sweep
            return {"what": "the single-view cull of one GPU at other sizes of BASELINE's range (tools/size_sweep.py; 100 M: profiles/r02_size_sweep.json)",
                    "rows": [sweep.one(int(m * 1e6), oracle, cores, frames=40) for m in (1, 5, 25)]}

        for key, fn in (("size_sweep", size_sweep),
                        ("headless_rasteriser", lambda: bench_raster.run(2000, 8)),      # G-buffer + shadow-map passes (ovis_raster.cu)
                        ("headless_frame", whole_frame),
                        ("config1", bench_configs.config1),
                        ("config3", lambda: bench_configs.config3(False, full_parity=True)),
                        ("config4", lambda: bench_configs.config4(False, full_parity=True, brute=False)),
                        ("dense_visible", bench_configs.dense_visible)):
            try:
                line["also"][key] = fn()
            except Exception as e:            # an extra must not take the headline line with it
                line["also"][key] = {"error": repr(e)}

    if world_size > 1 and not args.skip_extras:
        # The same sharded cull driven by ONE process through the C ABI (ovis_sharded_*: a worker thread per device inside
        # the library, all gather kernels storing into one page-locked buffer) - what a single-process engine would call.
        # Rank 0 runs it over all N GPUs while the other ranks wait on a host-side (gloo) barrier.
        host_group = dist.new_group(backend="gloo")
        torch.cuda.synchronize()
        dist.barrier(group=host_group)
        if rank == 0:
            try:
                spw = lib.ShardedWorld(list(range(world_size)))
                spw.add_entities(np.concatenate([aabb if r == 0 else shard_scene(r, n)[0] for r in range(world_size)]))
                out1 = lib.PinnedBuffer(max(total_visible * 2, 1024))
                step1 = spw.bind_cull(view, out1.array)
                for _ in range(5):
                    step1()
                t0 = time.perf_counter()
                for _ in range(args.steps):
                    got1, _ = step1()
                dt = time.perf_counter() - t0
                line.setdefault("also", {})["single_process_sharded_world"] = {
                    "what": "ovis_sharded_cull(view) from one host thread over all N GPUs -> merged ids in host memory",
                    "ms_per_step": dt / args.steps * 1e3, "tests_per_s": total_entities * args.steps / dt,
                    "list_exact": bool(np.array_equal(out1.array[:got1], exp)) if "parity" in line else None}
                # configs[4]'s "then 5 views": the camera and four more frusta in one many-view pass per shard, lists spliced
                # by the calling thread (ovis_sharded_cull_views)
                try:
                    five = [view] + [scenes.make_view((x, 40.0, z), scenes.quat_yaw_pitch(yaw, -5.0), 60.0, 0.5, 900.0, 1920, 1080)
                                     for x, z, yaw in ((1500.0, -800.0, 40.0), (-2200.0, 600.0, 200.0), (300.0, 2500.0, -95.0), (-900.0, -3000.0, 130.0))]
                    planes5 = np.stack([lib.view_planes(v) for v in five])
                    off5, ent5 = spw.cull_views(planes5, capacity=max(total_visible * 8, 1 << 16))
                    t0 = time.perf_counter()
                    reps = max(args.steps // 10, 5)
                    for _ in range(reps):
                        off5, ent5 = spw.cull_views(planes5, capacity=len(ent5) + 1024)
                    dt5 = time.perf_counter() - t0
                    ok5 = None
                    if "parity" in line:
                        all_boxes = np.concatenate([aabb if r == 0 else shard_scene(r, n)[0] for r in range(world_size)])
                        ok5 = all(np.array_equal(ent5[off5[v]:off5[v + 1]], oracle.cull_aabb(all_boxes, planes5[v], threads=cores))
                                  for v in range(5))
                    line["also"]["single_process_sharded_world"]["five_views"] = {
                        "what": "ovis_sharded_cull_views: 5 frusta over all shards, merged CSR in host memory",
                        "ms_per_step": dt5 / reps * 1e3, "entity_view_tests_per_s": 5 * total_entities * reps / dt5,
                        "visible": [int(off5[v + 1] - off5[v]) for v in range(5)], "lists_exact": ok5}
                except Exception as e:
                    line["also"]["single_process_sharded_world"]["five_views"] = {"error": repr(e)}
                spw.close()
            except Exception as e:
                line.setdefault("also", {})["single_process_sharded_world"] = {"error": repr(e)}
        dist.barrier(group=host_group)
    if rank == 0:
        emit(line)
    if world is not None:
        world.close()
    if world_size > 1:
        dist.barrier()
        dist.destroy_process_group()


_RESULT_FD = None


def emit(line: dict):
    """The ONE JSON line, written to the process's original stdout."""
    text = (json.dumps(line) + "\n").encode()
    if _RESULT_FD is None:
        sys.stdout.write(text.decode())
        sys.stdout.flush()
    else:
        os.write(_RESULT_FD, text)


def main():
    # Libraries write banners to fd 1 from native code (NCCL prints its version there at NCCL_DEBUG=VERSION or WARN): keep
    # the original stdout for the result line and point fd 1 at stderr for everything else.
    global _RESULT_FD
    sys.stdout.flush()
    _RESULT_FD = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--entities-per-gpu", type=int, default=SHARD_ENTITIES)
    ap.add_argument("--exchange", choices=["p2p", "nccl"], default="p2p",
                    help="multi-GPU merge of the visible lists: fused peer stores (default) or NCCL all-gather")
    ap.add_argument("--skip-extras", action="store_true", help="only the headline numbers (no CPU baseline / also)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world_size = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    if world_size == 1 and args.gpus > 1:
        raise SystemExit(f"bench.py --gpus {args.gpus}: launch with torchrun --nproc-per-node {args.gpus} (one rank per GPU)")
    run_ours(args, rank, local_rank, world_size)


if __name__ == "__main__":
    main()
