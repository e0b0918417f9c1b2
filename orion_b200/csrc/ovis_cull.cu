// ovis_cull.cu - K3: one view over all entity slots with ORDER-PRESERVING compaction, in two launches.
//
// Replaces the entity loop of SimpleRenderWorld::cull (render/src/simple_render_world.cc:38-41) with
// Math::intersect(Frustum, BoundingBox) (core/src/math/intersection.cc:42-57) inside.
//
//   K3a test kernel    pure streaming: every entity's six plane tests, result = ONE BIT per entity slot in a survivor
//                      bitmap (4 bytes per 32 entities) plus an integer total per block of 8192 entities (atomicAdd,
//                      order-independent). Two forms:
//                        cull_test_ring_kernel  large worlds: a producer warp per CTA streams the bounds through a
//                                               shared-memory ring with TMA bulk copies (cp.async.bulk + mbarrier), eight
//                                               consumer warps test 128 entities each per stage, reading P-vertex rows
//                                               straight from shared memory;
//                        cull_test_kernel       small worlds: register-fed, one warp per 128 x GROUPS entities.
//   K3b gather kernel  one CTA per block of 8192 entities: base = sum of the earlier block totals, exclusive scan of the
//                      popcounts of its 256 bitmap words, ids laid out in shared memory, then written densely (also to
//                      peers' / the host's buffers).
//
// Output order = slot order = insertion order and is identical from run to run: positions come from sums of counts only.
// History (profiles/r01_cull_notes.md, profiles/r02_cull_notes.md): a single-launch decoupled look-back version was capped
// at ~2.1 TB/s by CTA-wide barrier stalls; per-warp staging lists + tile counts (round 1) reached 0.88 of the HBM peak but
// were latency-bound on 16 warps / SM; per-warp TMA rings removed the memory stalls but executed 50 % more instructions.
#include "ovis_kernels.h"

#include <cstdlib>

#include <cooperative_groups.h>

#include "ovis_device.cuh"

namespace ovis {

__device__ __forceinline__ uint32_t load_flag_sys(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// after a __threadfence_system(): the fence orders every earlier store before this one at system scope
__device__ __forceinline__ void store_flag_relaxed_sys(uint32_t *p, uint32_t v) {
    asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// Spin until flags[i] >= value for every i, at most ~2 s (a peer that never arrives must not hang the GPU).
__device__ __forceinline__ void wait_flags_at_least(const uint32_t *flags, int n, uint32_t value, uint32_t *error_flag) {
    const long long start = clock64();
    for (int i = 0; i < n; i++) {
        while ((int32_t)(load_flag_sys(flags + i) - value) < 0) {
            if (clock64() - start > 4000000000ll) {
                if (error_flag)
                    *error_flag = 1u;
                return;
            }
            __nanosleep(100);
        }
    }
}

// The same by one whole warp, lane i polling flag i (n <= 32): the system-scope loads overlap instead of queueing up.
__device__ __forceinline__ void wait_flags_warp(const uint32_t *flags, int n, uint32_t value, uint32_t *error_flag) {
    const int lane = threadIdx.x & 31;
    if (lane < n)
        wait_flags_at_least(flags + lane, 1, value, error_flag);
    __syncwarp();
}

// ---- register-fed test kernel (small worlds, the fused single-launch form, the no-allocation entry point) -------------
// A warp owns tiles of 128 x GROUPS consecutive entities (warp-strided walk); a lane owns 4 x GROUPS of them and reads six
// 128-bit loads per group. The next tile's loads are issued before the current tile's bits are assembled. The P-vertex
// corner depends only on the signs of the plane normal, which are the same for every entity: one warp-uniform switch per
// plane (sign code computed on the host) picks the code path; products are packed FMUL2; "any distance < 0" is
// "minimum distance < 0" (FMNMX3 ignores a NaN operand exactly as a NaN distance rejects nothing): ovis_device.cuh.
template <int GROUPS, bool MASKED>
__device__ __forceinline__ void cull_test_body(const CullArgs &a) {
    // Warp-strided walk: at any moment the grid's warps read one contiguous stretch of every component array. (Measured
    // alternatives, profiles/r02_cull_notes.md: one contiguous range of tiles per warp, or runs of 8 tiles per warp so
    // that a run's survivors need a single atomic, are 2.5-3 us slower per 12.5 M entities - 4736 scattered streams.)
    const int lane = threadIdx.x & 31;
    const int warps_total = gridDim.x * (kCullThreads / 32);
    const int gwarp = blockIdx.x * (kCullThreads / 32) + (threadIdx.x >> 5);
    const float4 *__restrict__ bounds = reinterpret_cast<const float4 *>(a.bounds);
    const int64_t vstride = a.stride >> 2;                 // component stride in float4 units
    const int num_tiles = (a.num_groups + GROUPS - 1) / GROUPS;

    // re-arm the block totals the NEXT launch will accumulate into (this launch uses another set)
    for (int i = blockIdx.x * kCullThreads + threadIdx.x; i < a.num_blocks; i += gridDim.x * kCullThreads)
        a.block_totals_next[i] = 0u;

    float4 mnx[GROUPS], mny[GROUPS], mnz[GROUPS], mxx[GROUPS], mxy[GROUPS], mxz[GROUPS];
    uint32_t mask_bits[GROUPS];
    auto issue_loads = [&](int wt) {
        if (wt >= num_tiles)
            return;
#pragma unroll
        for (int g = 0; g < GROUPS; g++) {
            const int64_t v = (int64_t)wt * (32 * GROUPS) + g * 32 + lane;     // float4 index: entities 4v .. 4v+3
            mnx[g] = __ldg(bounds + 0 * vstride + v); mny[g] = __ldg(bounds + 1 * vstride + v); mnz[g] = __ldg(bounds + 2 * vstride + v);
            mxx[g] = __ldg(bounds + 3 * vstride + v); mxy[g] = __ldg(bounds + 4 * vstride + v); mxz[g] = __ldg(bounds + 5 * vstride + v);
            if (MASKED)
                mask_bits[g] = __ldg(a.mask + (v >> 3));                       // the 4 bits of entity 4v sit at bit (4v & 31)
        }
    };

    int wt = gwarp;
    issue_loads(wt);
    for (; wt < num_tiles; wt += warps_total) {
        const float inf = __int_as_float(0x7f800000);
        float4 lowest[GROUPS], held[GROUPS];
#pragma unroll
        for (int g = 0; g < GROUPS; g++)
            lowest[g] = make_float4(inf, inf, inf, inf);
#pragma unroll
        for (int p = 0; p < 6; p++) {
            const float4 q = a.planes[p];
#pragma unroll
            for (int g = 0; g < GROUPS; g++) {
                const float4 dd = plane_dist4_coded(a.sign_codes >> (3 * p), q, mnx[g], mny[g], mnz[g], mxx[g], mxy[g], mxz[g]);
                if (p & 1)
                    lowest[g] = min3_4(lowest[g], held[g], dd);
                else
                    held[g] = dd;
            }
        }
        uint32_t keep[GROUPS];
        uint32_t count = 0;
#pragma unroll
        for (int g = 0; g < GROUPS; g++) {
            const int64_t e0 = ((int64_t)wt * (32 * GROUPS) + g * 32 + lane) << 2;
            keep[g] = candidate_bits<MASKED>(e0, a.n, MASKED ? mask_bits[g] : 0u) & ~negative_bits(lowest[g]);
            count += __popc(keep[g]);
        }
        issue_loads(wt + warps_total);                                        // this tile's registers are dead now
#pragma unroll
        for (int g = 0; g < GROUPS; g++)
            if ((int64_t)wt * GROUPS + g < a.num_groups)
                store_group_bits(a.survivors, (int64_t)wt * GROUPS + g, keep[g], lane);
        const uint32_t total = __reduce_add_sync(0xffffffffu, count);
        if (lane == 0 && total)
            atomicAdd(a.block_totals + ((int64_t)wt * GROUPS) / kMaskBlockGroups, total);
    }
}

template <int GROUPS, bool MASKED, int CTAS = 4 / GROUPS>
__global__ void __launch_bounds__(kCullThreads, CTAS)
cull_test_kernel(const __grid_constant__ CullArgs a) {
    // Sharded exchange: the segment buffer this frame's gather kernel will refill must have been released by every peer.
    // One warp of the whole grid checks that here, hidden behind the streaming pass (the gather kernel starts only after
    // this kernel has finished), instead of every gather block paying a round of system-scope loads before its stores.
    if (a.n_wait && blockIdx.x == 0 && threadIdx.x < 32)
        wait_flags_warp(a.wait_flags, a.n_wait, a.wait_value, a.error_flag);
    cull_test_body<GROUPS, MASKED>(a);
    // Launched with programmatic stream serialization this kernel may START while the previous cull's gather kernel is
    // still running (they share no scratch). It must not FINISH before that kernel has: the next gather kernel and the
    // next test kernel are ordered after this one only, and they reuse the previous gather's scratch set and outputs.
    // Waiting here, after all the work, costs nothing and makes "this kernel completed" imply "its predecessor completed
    // and its writes are visible". Without the launch attribute the instruction does nothing.
    asm volatile("griddepcontrol.wait;" ::: "memory");
}

// ---- TMA ring test kernel (large worlds) --------------------------------------------------------------------------------
// The register-fed kernel holds 48-96 registers of bounds per lane while they are in flight, which caps the SM at 16 warps,
// and has loads outstanding only while it is not computing: ncu shows it latency-bound (long_scoreboard 3.2 per issue),
// not bandwidth-bound. Here the bytes never pass through registers on their way in: warp 8 of each CTA is a producer whose
// elected lane asks the TMA unit for whole rows (cp.async.bulk global -> shared, completion counted in bytes on an
// mbarrier): six 4 KB rows = 1024 entities per stage, STAGES stages ahead of the consumers. The other warps are consumers:
// each waits for the stage's "full" barrier, moves its 128 entities into registers, arrives on the stage's "empty"
// barrier at once and then runs the plane tests. There is no CTA-wide barrier in the loop. (Reading the P-vertex rows
// straight from shared memory - an address choice instead of a switch - was built first and is slower: it holds the stage
// for the whole computation.) Sources are this GPU's own memory only.
// A consumer warp must touch every stage it is responsible for, so with one set of eight consumer warps the CTA's rate is
// one warp's serial latency per stage (ncu: 4 warps per scheduler, 0.9 eligible, producer and consumers waiting on each
// other). SETS sets of eight consumers take the stages in turn: SETS stages are tested at a time, the rest are in flight.
constexpr int kRingConsumers = 8;                                     // warps that share one stage
constexpr int kRingStageEntities = 128 * kRingConsumers;             // 1024
constexpr int kRingRowBytes = kRingStageEntities * 4;                 // 4096
__host__ __device__ constexpr int ring_threads(int sets) { return (kRingConsumers * sets + 1) * 32; }

template <int STAGES, bool MASKED>
struct Ring {
    static constexpr int kMaskBytes = MASKED ? kRingStageEntities / 8 : 0;
    static constexpr int kStageBytes = 6 * kRingRowBytes + kMaskBytes;
    static constexpr int kBarrierOffset = STAGES * kStageBytes;
    static constexpr int kCtaBytes = kBarrierOffset + 2 * STAGES * 8;
    static constexpr int kFit = (227 * 1024) / (kCtaBytes + 1024);
    static constexpr int kCtasPerSm = kFit > 4 ? 4 : (kFit < 1 ? 1 : kFit);
};

__device__ __forceinline__ uint32_t smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbarrier_init(uint32_t bar, uint32_t arrivals) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(arrivals) : "memory");
}
__device__ __forceinline__ void mbarrier_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbarrier_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// arrive once `after` has been computed (the operand only orders the instruction behind its producers)
__device__ __forceinline__ void mbarrier_arrive_after(uint32_t bar, float after) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];  // %1" ::"r"(bar), "f"(after) : "memory");
}
__device__ __forceinline__ void mbarrier_wait(uint32_t bar, uint32_t parity) {
    uint32_t done;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    } while (!done);
}
__device__ __forceinline__ void bulk_load(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

template <int STAGES, int SETS, bool MASKED>
__global__ void __launch_bounds__(ring_threads(SETS), Ring<STAGES, MASKED>::kCtasPerSm)
cull_test_ring_kernel(const __grid_constant__ CullArgs a) {
    using R = Ring<STAGES, MASKED>;
    constexpr int kRingThreads = ring_threads(SETS);
    extern __shared__ __align__(128) unsigned char s_ring[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t ring = smem_addr(s_ring);
    const uint32_t full0 = ring + R::kBarrierOffset, empty0 = full0 + STAGES * 8;
    const int num_stage_tiles = (a.num_groups + kRingConsumers - 1) / kRingConsumers;

    if (a.n_wait && blockIdx.x == 0 && warp == 0)
        wait_flags_warp(a.wait_flags, a.n_wait, a.wait_value, a.error_flag);
    for (int i = blockIdx.x * kRingThreads + threadIdx.x; i < a.num_blocks; i += gridDim.x * kRingThreads)
        a.block_totals_next[i] = 0u;
    if (threadIdx.x == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; s++) {
            mbarrier_init(full0 + 8 * s, 1);
            mbarrier_init(empty0 + 8 * s, kRingConsumers);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");   // the async proxy sees initialised barriers
    }
    __syncthreads();

    if (warp == kRingConsumers * SETS) {
        // ---- producer: one lane keeps STAGES stages requested
        if (lane == 0) {
            int stage = 0;
            uint32_t parity = 0;
            for (int t = blockIdx.x; t < num_stage_tiles; t += gridDim.x) {
                mbarrier_wait(empty0 + 8 * stage, parity ^ 1u);                // passes at once the first time round
                const uint32_t bar = full0 + 8 * stage, dst = ring + stage * R::kStageBytes;
                mbarrier_expect_tx(bar, R::kStageBytes);
                const float *src = a.bounds + (int64_t)t * kRingStageEntities;
#pragma unroll
                for (int c = 0; c < 6; c++)
                    bulk_load(dst + c * kRingRowBytes, src + c * a.stride, kRingRowBytes, bar);
                if (MASKED)
                    bulk_load(dst + 6 * kRingRowBytes, a.mask + (int64_t)t * (kRingStageEntities / 32), R::kMaskBytes, bar);
                if (++stage == STAGES) {
                    stage = 0;
                    parity ^= 1u;
                }
            }
        }
    } else {
        // ---- consumers: warp w of set s tests entities [128 w, 128 w + 128) of every SETS-th stage, starting with stage s.
        // A stage is held only as long as it takes to move the warp's 3 KB into registers: tools/bulk_probe.cu shows the
        // stream falling from 45 us to 50-56 us when consumers sit on a stage for 1000-2000 cycles (the time the plane tests
        // take at this occupancy), because a CTA cannot take stages faster than its warps hand them back.
        const int set = warp / kRingConsumers, sub = warp % kRingConsumers;
        int stage = set % STAGES;
        uint32_t parity = (set / STAGES) & 1;
        for (int t = blockIdx.x + set * gridDim.x; t < num_stage_tiles; t += SETS * gridDim.x) {
            mbarrier_wait(full0 + 8 * stage, parity);                          // the stage's bytes have landed
            const float4 *rows = reinterpret_cast<const float4 *>(s_ring + stage * R::kStageBytes) + sub * 32 + lane;
            constexpr int kRow4 = kRingRowBytes / 16;
            const float4 mnx = rows[0 * kRow4], mny = rows[1 * kRow4], mnz = rows[2 * kRow4];
            const float4 mxx = rows[3 * kRow4], mxy = rows[4 * kRow4], mxz = rows[5 * kRow4];
            uint32_t mask_word = 0;
            if (MASKED)
                mask_word = reinterpret_cast<const uint32_t *>(s_ring + stage * R::kStageBytes + 6 * kRingRowBytes)[sub * 4 + (lane >> 3)];
            // The values must be in registers before the slot goes back. `landed` depends on all of the warp's loads (a
            // register written by a shared load is ready for the warp as a whole), and the arrive takes it as an operand.
            const float landed = ((mnx.x + mny.x) + (mnz.x + mxx.x)) + ((mxy.x + mxz.x) + __uint_as_float(mask_word));
            if (lane == 0)
                mbarrier_arrive_after(empty0 + 8 * stage, landed);
            const float inf = __int_as_float(0x7f800000);
            float4 lowest = make_float4(inf, inf, inf, inf), held;
#pragma unroll
            for (int p = 0; p < 6; p++) {
                const float4 dd = plane_dist4_coded(a.sign_codes >> (3 * p), a.planes[p], mnx, mny, mnz, mxx, mxy, mxz);
                if (p & 1)
                    lowest = min3_4(lowest, held, dd);
                else
                    held = dd;
            }
            const int64_t group = (int64_t)t * kRingConsumers + sub;
            const int64_t e0 = (group * 32 + lane) << 2;
            const uint32_t keep = candidate_bits<MASKED>(e0, a.n, mask_word) & ~negative_bits(lowest);
            if (group < a.num_groups) {
                store_group_bits(a.survivors, group, keep, lane);
                const uint32_t total = __reduce_add_sync(0xffffffffu, __popc(keep));
                if (lane == 0 && total)
                    atomicAdd(a.block_totals + group / kMaskBlockGroups, total);
            }
            stage += SETS;                                                     // the set's next stage
            if (stage >= STAGES) {
                stage -= STAGES;
                parity ^= 1u;
            }
        }
    }
    // every requested stage was consumed above, so no copy is outstanding when the CTA leaves
    asm volatile("griddepcontrol.wait;" ::: "memory");                         // see cull_test_kernel
}

// ---- K3b ---------------------------------------------------------------------------------------------------------------
// Gather of one block of 8192 entity slots (256 bitmap words) by one 256-thread CTA; `num_gather_blocks` = blocks in the
// launch. Thread t owns word t: its ids go to shared memory at the rank the block scan gives them, then the list leaves
// with coalesced stores (to the result, and to the peers' / the host's copies), kGatherWindow ids at a time.
// Everything the block needs from global memory - its own total, the totals of the earlier blocks, its bitmap word, the
// chained base - is requested up front, so the kernel's critical path is one trip to L2, not four dependent ones.
constexpr int kGatherWindow = 4096;                       // ids staged in shared memory per pass (16 KB)

__device__ __forceinline__ void cull_gather_body(const GatherArgs &a, int block, int num_gather_blocks, uint32_t *s_ids) {
    __shared__ uint32_t s_warp_base[8], s_warp_count[8];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    const uint32_t own = a.block_totals[block];
    const int64_t w = (int64_t)block * kMaskBlockWords + tid;
    uint32_t word = w < a.num_words ? a.survivors[w] : 0u;
    const long long out_base = a.d_base ? *a.d_base : 0ll;
    uint32_t part = 0;                                     // base = survivors in all earlier blocks
    for (int i = tid; i < block; i += 256)
        part += a.block_totals[i];
    // A block without survivors has nothing to place (most blocks of a large world under a narrow frustum); only the last
    // block must go on, it publishes the grand total.
    if (block != num_gather_blocks - 1 && own == 0u)
        return;
    part = __reduce_add_sync(0xffffffffu, part);
    const uint32_t c = __popc(word);
    uint32_t incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t up = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o)
            incl += up;
    }
    if (lane == 0)
        s_warp_base[warp] = part;
    if (lane == 31)
        s_warp_count[warp] = incl;
    __syncthreads();
    uint32_t before = 0, total = 0, base = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        if (k < warp)
            before += s_warp_count[k];
        total += s_warp_count[k];
        base += s_warp_base[k];
    }
    const uint32_t first_rank = before + incl - c;
    const uint32_t id0 = (uint32_t)(w << 5) + a.id_base;
    const long long dst0 = out_base + (long long)base;
    if (block == num_gather_blocks - 1 && tid == 0) {       // the last block knows the grand total
        uint32_t grand = base + total;
        if (a.d_count) *a.d_count = grand;
        if (a.d_next) *a.d_next = out_base + (long long)grand;
        for (int x = 0; x < a.n_extra; x++)
            *a.extra_count[x] = grand;
    }
    for (uint32_t lo = 0; lo < total; lo += kGatherWindow) {
        uint32_t rank = first_rank;
        for (uint32_t bits = word; bits; bits &= bits - 1, rank++)
            if (rank - lo < (uint32_t)kGatherWindow)        // (unsigned: ranks below the window wrap around and fail too)
                s_ids[rank - lo] = id0 + (__ffs(bits) - 1);
        __syncthreads();
        const uint32_t n = min(total - lo, (uint32_t)kGatherWindow);
        for (uint32_t k = tid; k < n; k += 256) {
            const uint32_t v = s_ids[k];
            const long long pos = dst0 + lo + k;
            if (pos < a.capacity)
                a.out[pos] = v;
            if (pos < a.extra_capacity)
                for (int x = 0; x < a.n_extra; x++)            // peers' copies of this shard's segment / the host's copy
                    a.extra_out[x][pos] = v;
        }
        __syncthreads();                                    // the window is refilled by the next pass / the next block
    }
}

// The last block of a gather (or fused) launch to finish publishes the result: ready flags for peers, the frame word for a
// watching host.
__device__ __forceinline__ void signal_when_all_blocks_done(const GatherArgs &a) {
    if (a.n_signal) {                                         // the last block to finish publishes the segment
        __syncthreads();                                      // this block's stores happen-before thread 0's fence
        if (threadIdx.x == 0) {
            // Release this block's stores (the barrier above makes them cumulative with thread 0) to the rest of the GPU;
            // the last block's system fence below then orders every block's stores before the flags at system scope
            // (PTX causality order is transitive across the gpu-scope and the sys-scope synchronisation). A system fence
            // in every block was measured to cost ~4.4 us per frame: it stalls the next cull's test kernel on every SM.
            if (a.block_fence_sys)
                __threadfence_system();
            else
                __threadfence();
            unsigned int prev = atomicAdd(a.done_counter, 1u);
            if (prev == gridDim.x - 1) {
                *a.done_counter = 0u;
                __threadfence_system();
                for (int x = 0; x < a.n_signal; x++)
                    store_flag_relaxed_sys(a.signal_flags[x], a.signal_value);
            }
        }
    }
}

__global__ void __launch_bounds__(256)
cull_gather_kernel(const __grid_constant__ GatherArgs a) {
    __shared__ uint32_t s_ids[kGatherWindow];
    // Let the next cull's test kernel start now (if it was launched to allow that): it shares nothing with this kernel.
    asm volatile("griddepcontrol.launch_dependents;");
    if (a.n_wait) {                                           // peers must be done with the buffer we are about to refill
        if (threadIdx.x == 0)
            wait_flags_at_least(a.wait_flags, a.n_wait, a.wait_value, a.error_flag);
        __syncthreads();
    }
    // a CTA takes every gridDim.x-th block (blocks without survivors return at once inside the body)
    for (int b = blockIdx.x; b < a.num_blocks; b += gridDim.x) {
        cull_gather_body(a, b, a.num_blocks, s_ids);
        __syncthreads();                                      // the body's shared arrays are reused by the next block
    }
    signal_when_all_blocks_done(a);
}


__global__ void wait_flags_kernel(const uint32_t *flags, int n, uint32_t value, uint32_t *error_flag) {
    wait_flags_at_least(flags, n, value, error_flag);
}

// Reader side of the sharded exchange with the host as the consumer: CTA row r waits for rank r's ready flag alone, then
// moves that segment - its count word and exactly `count` ids - into mapped pinned host memory. One launch replaces the
// wait kernel plus a strided device-to-host copy sized by guesswork, and a segment leaves as soon as it has landed.
__global__ void __launch_bounds__(256)
segments_to_host_kernel(const uint32_t *flags, uint32_t value, const uint32_t *segments, long long segment_words,
                        uint32_t *host, uint32_t *error_flag) {
    const int r = blockIdx.y;
    if (flags) {
        if (threadIdx.x == 0)
            wait_flags_at_least(flags + r, 1, value, error_flag);
        __syncthreads();
    }
    const uint32_t *seg = segments + (long long)r * segment_words;
    uint32_t *dst = host + (long long)r * segment_words;
    const uint32_t count = __ldcg(seg);                       // L2 is where the peers' stores land: skip L1
    const long long words = 1 + (long long)(count < (uint32_t)(segment_words - 1) ? count : (uint32_t)(segment_words - 1));
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < words; i += (long long)gridDim.x * blockDim.x)
        dst[i] = __ldcg(seg + i);
}

void launch_segments_to_host(const uint32_t *flags, uint32_t value, const uint32_t *segments, long long segment_words,
                             int n_segments, uint32_t *host, uint32_t *error_flag, cudaStream_t s) {
    segments_to_host_kernel<<<dim3(8, (unsigned)n_segments), 256, 0, s>>>(flags, value, segments, segment_words, host, error_flag);
}

void launch_wait_flags(const uint32_t *flags, int n, uint32_t value, uint32_t *error_flag, cudaStream_t s) {
    wait_flags_kernel<<<1, 1, 0, s>>>(flags, n, value, error_flag);
}

// Small worlds are bound by launch and memory latency, not bandwidth: both phases in ONE cooperative launch with a grid
// barrier in between save a kernel boundary and a launch (configs[1], 1 M entities: see profiles/r01_cull_notes.md).
template <int GROUPS, bool MASKED>
__global__ void __launch_bounds__(kCullThreads, 4 / GROUPS)
cull_fused_kernel(const __grid_constant__ CullArgs a, const __grid_constant__ GatherArgs g) {
    __shared__ uint32_t s_ids[kGatherWindow];
    cull_test_body<GROUPS, MASKED>(a);
    cooperative_groups::this_grid().sync();
    for (int b = blockIdx.x; b < g.num_blocks; b += gridDim.x) {
        cull_gather_body(g, b, g.num_blocks, s_ids);
        __syncthreads();                                    // shared arrays are reused by the next block
    }
    signal_when_all_blocks_done(g);
}

// Launch shapes selectable at run time (OVIS_CULL_VARIANT, tuning only); the world picks by size (cull_default_variant).
static const CullVariant kVariants[] = {
    { 2, 2, 0, 0 },   // 0  register-fed, 256-entity warp tiles, 2 CTAs / SM
    { 1, 4, 0, 0 },   // 1  register-fed, 128-entity warp tiles, 4 CTAs / SM (also the fused single-launch form)
    { 1, 5, 0, 0 },   // 2  the same compiled for 5 CTAs / SM
    { 1, 6, 0, 0 },   // 3  the same compiled for 6 CTAs / SM
    { 1, 0, 4, 1 },   // 4  TMA ring, 4 stages of 1024 entities, 8 consumer warps (2 CTAs / SM)
    { 1, 0, 6, 3 },   // 5  TMA ring, 6 stages, 3 x 8 consumer warps (1 CTA / SM)
};

int cull_num_variants() { return (int)(sizeof(kVariants) / sizeof(kVariants[0])); }
CullVariant cull_variant(int v) { return kVariants[v]; }

static int ring_smem(int stages, bool masked) {
    switch (stages) {
    case 6: return masked ? Ring<6, true>::kCtaBytes : Ring<6, false>::kCtaBytes;
    default: return masked ? Ring<4, true>::kCtaBytes : Ring<4, false>::kCtaBytes;
    }
}

template <typename F>
static void with_variant(int variant, bool masked, F &&f) {
    const CullVariant &cv = kVariants[variant];
#define OVIS_RING(S, N) do { if (masked) f(cull_test_ring_kernel<S, N, true>, ring_threads(N)); else f(cull_test_ring_kernel<S, N, false>, ring_threads(N)); } while (0)
#define OVIS_REG(G, C) do { if (masked) f(cull_test_kernel<G, true, C>, kCullThreads); else f(cull_test_kernel<G, false, C>, kCullThreads); } while (0)
    if (cv.ring_stages == 4) { OVIS_RING(4, 1); return; }
    if (cv.ring_stages == 6) { OVIS_RING(6, 3); return; }
    if (cv.groups == 2) { OVIS_REG(2, 2); return; }
    if (cv.ctas == 5) { OVIS_REG(1, 5); return; }
    if (cv.ctas == 6) { OVIS_REG(1, 6); return; }
    OVIS_REG(1, 4);
#undef OVIS_RING
#undef OVIS_REG
}

// CTAs of the variant's test kernel that fit one SM, and the largest useful grid for `num_groups` groups of 128 entities.
int cull_test_max_blocks_per_sm(bool masked, int variant) {
    int blocks = 0;
    const int smem = kVariants[variant].ring_stages ? ring_smem(kVariants[variant].ring_stages, masked) : 0;
    with_variant(variant, masked, [&](auto kernel, int threads) {
        if (smem)
            cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks, kernel, threads, (size_t)smem);
    });
    return blocks;
}

int cull_test_grid(int variant, int num_groups, int resident_ctas) {
    const CullVariant &cv = kVariants[variant];
    const int groups_per_cta = cv.ring_stages ? kRingConsumers : (kCullThreads / 32) * cv.groups;
    const int needed = (num_groups + groups_per_cta - 1) / groups_per_cta;
    return needed < resident_ctas ? (needed < 1 ? 1 : needed) : resident_ctas;
}

// overlap_previous_gather: launch with programmatic stream serialization, so this test kernel may start while the
// PREVIOUS cull's gather kernel (which signals its dependents at once) is still running. Only legal when the two use
// different scratch sets and nothing else sits between them in the stream; the kernel ends with griddepcontrol.wait.
void launch_cull_test(const CullArgs &a, bool masked, int variant, int grid, cudaStream_t s, bool overlap_previous_gather) {
    const int smem = kVariants[variant].ring_stages ? ring_smem(kVariants[variant].ring_stages, masked) : 0;
    with_variant(variant, masked, [&](auto kernel, int threads) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)grid);
        cfg.blockDim = dim3((unsigned)threads);
        cfg.dynamicSmemBytes = (size_t)smem;
        if (smem)
            cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        cfg.stream = s;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr;
        cfg.numAttrs = overlap_previous_gather ? 1 : 0;
        cudaLaunchKernelEx(&cfg, kernel, a);
    });
}

// Cooperative single launch of both phases; grid must not exceed the co-resident capacity (the caller sizes it so).
cudaError_t launch_cull_fused(const CullArgs &a, const GatherArgs &g, bool masked, int variant, int grid, cudaStream_t s) {
    cudaError_t err = cudaSuccess;
    auto go = [&](auto kernel) {
        void *params[2] = { (void *)&a, (void *)&g };
        err = cudaLaunchCooperativeKernel((const void *)kernel, dim3((unsigned)grid), dim3(kCullThreads), params, 0, s);
    };
    if (kVariants[variant].ring_stages)
        return cudaErrorInvalidValue;
    switch (kVariants[variant].groups) {
    case 1: if (masked) go(cull_fused_kernel<1, true>); else go(cull_fused_kernel<1, false>); break;
    case 2: if (masked) go(cull_fused_kernel<2, true>); else go(cull_fused_kernel<2, false>); break;
    default: err = cudaErrorInvalidValue; break;
    }
    return err;
}

int cull_fused_max_blocks_per_sm(bool masked, int variant) {
    int blocks = 0;
    auto q = [&](auto kernel) { cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks, kernel, kCullThreads, 0); };
    if (kVariants[variant].ring_stages)
        return 0;                                           // the single-launch form exists for the register-fed kernels
    switch (kVariants[variant].groups) {
    case 1: if (masked) q(cull_fused_kernel<1, true>); else q(cull_fused_kernel<1, false>); break;
    case 2: if (masked) q(cull_fused_kernel<2, true>); else q(cull_fused_kernel<2, false>); break;
    }
    return blocks;
}

void launch_cull_gather(const GatherArgs &a, cudaStream_t s) {
    int blocks = a.num_blocks;
    static const int cap = getenv("OVIS_GATHER_CTAS") ? atoi(getenv("OVIS_GATHER_CTAS")) : 0;   // tuning
    if (cap > 0 && blocks > cap)
        blocks = cap;
    cull_gather_kernel<<<blocks, 256, 0, s>>>(a);
}

// n == 0 never launches the tile loop; this writes the empty result instead.
__global__ void empty_result_kernel(const long long *d_base, long long *d_next, uint32_t *d_count) {
    if (d_count) *d_count = 0u;
    if (d_next) *d_next = d_base ? *d_base : 0ll;
}

void launch_empty_result(const long long *d_base, long long *d_next, uint32_t *d_count, cudaStream_t s) {
    empty_result_kernel<<<1, 1, 0, s>>>(d_base, d_next, d_count);
}

// P-vertex selection bits of six planes: bit 3p + c set = component c of plane p's normal is >= 0 (calcPVertex picks the
// maximum there, core/include/math/bounding_box.h:60-71; -0.0 >= 0 holds, NaN >= 0 does not - as in the reference's select)
uint32_t plane_sign_codes(const float planes[24]) {
    uint32_t codes = 0;
    for (int p = 0; p < 6; p++)
        for (int c = 0; c < 3; c++)
            if (planes[4 * p + c] >= 0.0f)
                codes |= 1u << (3 * p + c);
    return codes;
}

}  // namespace ovis
