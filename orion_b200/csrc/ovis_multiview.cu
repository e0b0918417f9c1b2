// ovis_multiview.cu - K4: many views over one world, K6: light -> entity candidate binning.
//
// Replaces the reference's "one full world cull per shadow view" (render/src/deferred_render_pipeline.cc:259-284:
// 1 + sum of shadow views full scans per frame) and serves the main camera + cascade views the same way. Every exact
// test is evaluated ONCE (round 1 counted, scanned and then repeated the tests to write).
//
//   dense pass (views without a pre-filter: camera, cascades, spot lights)
//     mv_plain_test_kernel    a warp takes 128 entities (4 per lane, six 128-bit loads) and a chunk of the plain views; per
//                             view the six plane tests with K3's instructions, result = 128 survivor bits in the view's
//                             bitmap + the count added to the view's total for that block of 8192 entities. Small worlds
//                             split the views over more warps so that the machine fills.
//   sparse pass (cube groups: the six faces of a point light share the cube of half-side range + extent about the light)
//     mv_bin_lights_kernel    K6, light side: every group is linked into the cells of a uniform grid it covers (per-cell
//                             lists built with atomics) and marks them in an occupancy bitmap.
//     mv_classify_kernel      streaming: (1) every entity looks its cells up in the occupancy bitmap held in shared memory
//                             and leaves one bit - nearly all leave with 0
//     mv_sparse_kernel        one CTA per block of 8192 entities with bits set: (2) those entities walk their cells' lists
//                             and append (entity, light) candidate pairs that pass the cube test to a shared-memory list
//                             with shared-memory atomics; (3) one thread per pair runs the exact tests of the light's
//                             faces; (4) the surviving records are sorted by (light, entity) in shared memory, which makes
//                             everything that follows independent of the order the atomics happened in; (5) a segmented
//                             scan ranks each survivor inside its (view, block), the block's per-view totals are stored
//                             and the survivors leave as (view, rank, id) items.
//   mv_scan_kernel            per view: exclusive scan of the block totals; the last CTA turns the view totals into CSR
//                             offsets.
//   mv_expand_masks_kernel    dense pass: (block, view) bitmap -> ids at csr_offset + block prefix + rank.
//   mv_expand_items_kernel    sparse pass: items -> ids at csr_offset + block prefix + rank.
//
// Positions are sums of counts, so every list is in slot (= insertion) order and identical from run to run. The exact
// tests are the same instructions as K3 (ovis_device.cuh), so lists are bit-identical to brute force; the grid and the
// cube test only ever skip (entity, group) pairs that every one of the group's views would reject.
#include "ovis_kernels.h"

#include "ovis_device.cuh"

namespace ovis {

constexpr int kMvThreads = 256;
constexpr int kPlainThreads = 256, kPlainCtas = 2;   // mv_plain_test_kernel: 128 registers, 16 warps per SM (128 x 5 at 96 registers with a little spilling: 5 % slower)
constexpr int kSpThreads = 256;           // mv_sparse_kernel (128-thread CTAs, eleven per SM - one wave for a 10 M-entity world - were
                                          // measured slower on every workload: 0.130 -> 0.161 ms for configs[2]'s shadow lists)
constexpr int kMvBigSlots = 512;          // records per block that are tested against every group a warp each
constexpr int kMvRankSortBelow = 320;    // candidate pairs of a block up to which they are sorted by counting (<= kMvPairCapacity / 2)

// The kernels of one round are a chain of short launches. Each is launched with programmatic stream serialization and
// starts with these two instructions: "my dependents may be scheduled" and "wait until everything before me in the stream
// has completed and is visible". The dependency stays complete (every kernel waits before it touches memory or returns,
// so "kernel N finished" still implies "kernels before N finished"); what overlaps is the launch latency and the scheduling
// of the next grid's CTAs with the tail of this one. Without the launch attribute both instructions do nothing.
__device__ __forceinline__ void chain_enter() {
    asm volatile("griddepcontrol.launch_dependents;");
    asm volatile("griddepcontrol.wait;" ::: "memory");
}

template <typename Kernel>
static void launch_chained(Kernel kernel, dim3 grid, size_t smem, cudaStream_t s, const MvArgs &a, int threads = kMvThreads) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = dim3((unsigned)threads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cudaLaunchKernelEx(&cfg, kernel, a);
}

// ------------------------------------------------------------------------------------------------
// dense pass
// ------------------------------------------------------------------------------------------------

// The exact tests of R plain views that share their P-vertex sign codes (a camera and its cascades; the same face of
// every point light), for the lane's four entities. Which corner is the P-vertex is decided per plane by three selects on
// the (warp-uniform) code bits, shared by the R views: no branches, so the six planes of all R views are one basic block
// the scheduler can interleave (round 2's first version switched over the 8 codes per plane: 140 KB of code, a third of the
// issue slots spent on branches and compares, ncu in profiles/r02_k4_config4_v2_plain_runs_sparse_ncu.csv).
// Products are packed FMUL2, "sum - d" is a packed subtract, "any distance < 0" is "minimum distance < 0" (fminf ignores a
// NaN operand exactly as a NaN distance rejects nothing). Planes come with doubled components (plain_planes), so a packed
// operand is a register pair as loaded. Warp-uniform early-out after the two side planes.
template <int R>
__device__ __forceinline__ void test_view_run(const MvArgs &a, int pv, uint32_t codes, const float4 &ax, const float4 &ay, const float4 &az,
                                              const float4 &bx, const float4 &by, const float4 &bz, uint32_t alive, uint32_t caster,
                                              int64_t group, int lane) {
    const int64_t num_words = (int64_t)a.num_groups128 * 4;
    const int row = a.num_blocks + 1;
    const float4 *__restrict__ pl = a.plain_planes + (int64_t)pv * 12;
    uint32_t view[R], keep[R];
    uint32_t any = 0;
#pragma unroll
    for (int r = 0; r < R; r++) {
        const uint2 m = __ldg(a.plain_meta + pv + r);
        view[r] = m.y;
        keep[r] = (m.x >> 31) ? caster : alive;
        any |= keep[r];
    }
    const float inf = __int_as_float(0x7f800000);
    float4 lowest[R], held[R];
#pragma unroll
    for (int r = 0; r < R; r++)
        lowest[r] = make_float4(inf, inf, inf, inf);
    bool alive_somewhere = __any_sync(0xffffffffu, any != 0u);
    if (alive_somewhere) {
#pragma unroll
        for (int p = 0; p < 6; p++) {
            const bool sx = (codes >> (3 * p)) & 1u, sy = (codes >> (3 * p + 1)) & 1u, sz = (codes >> (3 * p + 2)) & 1u;
            const float4 px = sx ? bx : ax, py = sy ? by : ay, pz = sz ? bz : az;
            const unsigned long long x01 = pack2(px.x, px.y), x23 = pack2(px.z, px.w), y01 = pack2(py.x, py.y), y23 = pack2(py.z, py.w),
                                     z01 = pack2(pz.x, pz.y), z23 = pack2(pz.z, pz.w);
#pragma unroll
            for (int r = 0; r < R; r++) {
                const float4 dd = plane_dist4_packed(__ldg(pl + 12 * r + 2 * p), __ldg(pl + 12 * r + 2 * p + 1), x01, x23, y01, y23, z01, z23);
                if (p & 1)
                    lowest[r] = min3_4(lowest[r], held[r], dd);
                else
                    held[r] = dd;
            }
            if (p == 1) {                                              // the side planes reject most tiles of a coherent scene
                uint32_t left = 0;
#pragma unroll
                for (int r = 0; r < R; r++)
                    left |= keep[r] & ~negative_bits(lowest[r]);
                alive_somewhere = __any_sync(0xffffffffu, left != 0u);
                if (!alive_somewhere)
                    break;
            }

        }
    }
    // Results: per view four ballots (ballot j = entity 4 * lane + j of every lane), stored as they are - the dense pass's
    // bitmaps are "j-major" inside a group of 128 (word 4 * group + j, bit = lane), which mv_expand_masks_kernel undoes.
    // Nothing here waits on a shuffle, and the R views' instructions are independent of one another.
#pragma unroll
    for (int r = 0; r < R; r++) {
        const uint32_t k = alive_somewhere ? keep[r] : 0u;
        uint4 b;
        b.x = __ballot_sync(0xffffffffu, (k & 1u) && !(lowest[r].x < 0.0f));
        b.y = __ballot_sync(0xffffffffu, (k & 2u) && !(lowest[r].y < 0.0f));
        b.z = __ballot_sync(0xffffffffu, (k & 4u) && !(lowest[r].z < 0.0f));
        b.w = __ballot_sync(0xffffffffu, (k & 8u) && !(lowest[r].w < 0.0f));
        const uint32_t total = __popc(b.x) + __popc(b.y) + __popc(b.z) + __popc(b.w);
        if (lane == 0) {
            *reinterpret_cast<uint4 *>(a.masks + (int64_t)(pv + r) * num_words + group * 4) = b;
            if (total)
                atomicAdd(a.totals + (int64_t)view[r] * row + group / kMaskBlockGroups, total);
        }
    }
}

// A run of `count` views (2 .. 7) whose left / right / top / bottom planes are bit-identical - a camera and the depth slices
// of its frustum ("cascades"): only near and far differ; the host says so per run. The four side distances are computed
// once and serve every view of the run: 4 + 2 * count plane tests instead of 6 * count, and nothing is held per view, so
// the run can be longer than four. Same planes, same P-vertex, same operations: same bits.
__device__ __forceinline__ void test_shared_side_run(const MvArgs &a, int pv, int count, uint32_t codes, const float4 &ax, const float4 &ay,
                                                     const float4 &az, const float4 &bx, const float4 &by, const float4 &bz, uint32_t alive,
                                                     uint32_t caster, int64_t group, int lane) {
    const int64_t num_words = (int64_t)a.num_groups128 * 4;
    const int row = a.num_blocks + 1;
    const float4 *__restrict__ pl = a.plain_planes + (int64_t)pv * 12;
    const float inf = __int_as_float(0x7f800000);
    float4 side = make_float4(inf, inf, inf, inf), held = side;
    unsigned long long x01[2], x23[2], y01[2], y23[2], z01[2], z23[2];           // P-vertices of the near and the far plane
    bool alive_somewhere = __any_sync(0xffffffffu, (alive | caster) != 0u);
    if (alive_somewhere) {
#pragma unroll
        for (int p = 0; p < 6; p++) {
            const bool sx = (codes >> (3 * p)) & 1u, sy = (codes >> (3 * p + 1)) & 1u, sz = (codes >> (3 * p + 2)) & 1u;
            const float4 px = sx ? bx : ax, py = sy ? by : ay, pz = sz ? bz : az;
            if (p >= 4) {
                x01[p - 4] = pack2(px.x, px.y); x23[p - 4] = pack2(px.z, px.w); y01[p - 4] = pack2(py.x, py.y);
                y23[p - 4] = pack2(py.z, py.w); z01[p - 4] = pack2(pz.x, pz.y); z23[p - 4] = pack2(pz.z, pz.w);
                continue;
            }
            const float4 dd = plane_dist4_packed(__ldg(pl + 2 * p), __ldg(pl + 2 * p + 1), pack2(px.x, px.y), pack2(px.z, px.w),
                                                 pack2(py.x, py.y), pack2(py.z, py.w), pack2(pz.x, pz.y), pack2(pz.z, pz.w));
            if (p & 1)
                side = min3_4(side, held, dd);
            else
                held = dd;
        }
        alive_somewhere = __any_sync(0xffffffffu, ((alive | caster) & ~negative_bits(side)) != 0u);
    }
    for (int r = 0; r < count; r++) {
        const uint2 m = __ldg(a.plain_meta + pv + r);
        const uint32_t keep = alive_somewhere ? ((m.x >> 31) ? caster : alive) : 0u;
        uint4 b = make_uint4(0u, 0u, 0u, 0u);
        if (alive_somewhere) {
            const float4 dn = plane_dist4_packed(__ldg(pl + 12 * r + 8), __ldg(pl + 12 * r + 9), x01[0], x23[0], y01[0], y23[0], z01[0], z23[0]);
            const float4 df = plane_dist4_packed(__ldg(pl + 12 * r + 10), __ldg(pl + 12 * r + 11), x01[1], x23[1], y01[1], y23[1], z01[1], z23[1]);
            const float4 lowest = min3_4(side, dn, df);
            b.x = __ballot_sync(0xffffffffu, (keep & 1u) && !(lowest.x < 0.0f));
            b.y = __ballot_sync(0xffffffffu, (keep & 2u) && !(lowest.y < 0.0f));
            b.z = __ballot_sync(0xffffffffu, (keep & 4u) && !(lowest.z < 0.0f));
            b.w = __ballot_sync(0xffffffffu, (keep & 8u) && !(lowest.w < 0.0f));
        }
        const uint32_t total = __popc(b.x) + __popc(b.y) + __popc(b.z) + __popc(b.w);
        if (lane == 0) {
            *reinterpret_cast<uint4 *>(a.masks + (int64_t)(pv + r) * num_words + group * 4) = b;
            if (total)
                atomicAdd(a.totals + (int64_t)m.y * row + group / kMaskBlockGroups, total);
        }
    }
}

__global__ void __launch_bounds__(kPlainThreads, kPlainCtas)
mv_plain_test_kernel(const __grid_constant__ MvArgs a) {
    chain_enter();
    const int lane = threadIdx.x & 31;
    const long long warps_total = (long long)gridDim.x * (kPlainThreads / 32);
    const long long gwarp = (long long)blockIdx.x * (kPlainThreads / 32) + (threadIdx.x >> 5);
    const float4 *__restrict__ bounds = reinterpret_cast<const float4 *>(a.bounds);
    const int64_t vstride = a.stride >> 2;
    const int n_chunks = (a.n_plain + a.views_per_warp - 1) / a.views_per_warp;
    const long long items = (long long)a.num_groups128 * n_chunks;

    float4 ax, ay, az, bx, by, bz;
    auto issue_loads = [&](long long item) {
        if (item >= items)
            return;
        const int64_t v = (n_chunks == 1 ? item : item % a.num_groups128) * 32 + lane;   // float4 index: entities 4v .. 4v+3
        ax = __ldg(bounds + 0 * vstride + v); ay = __ldg(bounds + 1 * vstride + v); az = __ldg(bounds + 2 * vstride + v);
        bx = __ldg(bounds + 3 * vstride + v); by = __ldg(bounds + 4 * vstride + v); bz = __ldg(bounds + 5 * vstride + v);
    };
    long long item = gwarp;
    issue_loads(item);
    for (; item < items; item += warps_total) {
        const int chunk = n_chunks == 1 ? 0 : (int)(item / a.num_groups128);
        const int64_t group = n_chunks == 1 ? item : item % a.num_groups128;
        const int64_t e0 = (group * 32 + lane) << 2;
        uint32_t alive = candidate_bits<false>(e0, a.n, 0u);
        if (a.alive_mask)
            alive &= (__ldg(a.alive_mask + (e0 >> 5)) >> (e0 & 31)) & 0xfu;
        uint32_t caster = alive;
        if (a.caster_mask)
            caster &= (__ldg(a.caster_mask + (e0 >> 5)) >> (e0 & 31)) & 0xfu;
        const int pv1 = min(a.n_plain, (chunk + 1) * a.views_per_warp);
        // the host lists the plain views sorted by sign code and says how many equal codes follow: runs of 4, 2 or 1
        for (int pv = chunk * a.views_per_warp; pv < pv1;) {
            const uint32_t m = __ldg(&a.plain_meta[pv].x);
            const uint32_t codes = m & 0x3ffffu;
            const int shared = min((int)((m >> kPlainSharedShift) & 7u), pv1 - pv);
            if (shared >= 2) {
                test_shared_side_run(a, pv, shared, codes, ax, ay, az, bx, by, bz, alive, caster, group, lane);
                pv += shared;
                continue;
            }
            int run = min((int)((m >> 18) & 7u), pv1 - pv);
            if (run >= 4) {
                run = 4;
                test_view_run<4>(a, pv, codes, ax, ay, az, bx, by, bz, alive, caster, group, lane);
            } else if (run >= 2) {
                run = 2;
                test_view_run<2>(a, pv, codes, ax, ay, az, bx, by, bz, alive, caster, group, lane);
            } else {
                run = 1;
                test_view_run<1>(a, pv, codes, ax, ay, az, bx, by, bz, alive, caster, group, lane);
            }
            pv += run;
        }
        issue_loads(item + warps_total);
    }
}

// block-wide exclusive scan of one value per thread (256 threads); returns the rank, *total = the block sum
template <int THREADS = kMvThreads>
__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t c, uint32_t *s_warp, uint32_t *total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t up = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o)
            incl += up;
    }
    __syncthreads();                                       // s_warp may still be read from the previous use
    if (lane == 31)
        s_warp[warp] = incl;
    __syncthreads();
    uint32_t before = 0, sum = 0;
#pragma unroll
    for (int k = 0; k < THREADS / 32; k++) {
        if (k < warp)
            before += s_warp[k];
        sum += s_warp[k];
    }
    *total = sum;
    return before + incl - c;
}

__global__ void __launch_bounds__(kMvThreads)
mv_expand_masks_kernel(const __grid_constant__ MvArgs a) {
    chain_enter();
    __shared__ uint32_t s_ids[kMaskBlockEntities];
    __shared__ uint32_t s_warp[kMvThreads / 32];
    const int b = blockIdx.x, pv = blockIdx.y, tid = threadIdx.x;
    const int view = (int)a.plain_views[pv];
    const uint32_t *row = a.totals + (int64_t)view * (a.num_blocks + 1);
    const uint32_t before = row[b], count = row[b + 1] - before;
    if (count == 0)
        return;
    const int64_t num_words = (int64_t)a.num_groups128 * 4;
    // thread tid owns entities 32 * tid .. + 31 of the block: byte (tid & 3) of each of its group's four ballot words,
    // bit l of ballot j -> bit 4 * l + j (mv_plain_test_kernel stores ballots, not entity-ordered words)
    const int64_t w = (int64_t)b * kMaskBlockWords + tid;
    uint32_t word = 0u;
    if (w < num_words) {
        const uint4 q = __ldg(reinterpret_cast<const uint4 *>(a.masks + (int64_t)pv * num_words + (w & ~(int64_t)3)));
        const int sh = (tid & 3) * 8;
        auto spread = [](uint32_t x) {                     // bit l of a byte -> bit 4 * l
            x = (x | (x << 12)) & 0x000f000fu;
            x = (x | (x << 6)) & 0x03030303u;
            return (x | (x << 3)) & 0x11111111u;
        };
        word = spread((q.x >> sh) & 0xffu) | (spread((q.y >> sh) & 0xffu) << 1) | (spread((q.z >> sh) & 0xffu) << 2) |
               (spread((q.w >> sh) & 0xffu) << 3);
    }
    uint32_t total;
    uint32_t rank = block_exclusive_scan(__popc(word), s_warp, &total);
    const uint32_t id0 = (uint32_t)(w << 5) + a.id_base;
    while (word) {
        const int bit = __ffs(word) - 1;
        word &= word - 1;
        s_ids[rank++] = id0 + bit;
    }
    __syncthreads();
    const long long dst0 = a.csr_offsets[view] + (long long)before;
    for (uint32_t k = tid; k < total; k += kMvThreads)
        if (dst0 + k < a.capacity)
            a.out[dst0 + k] = s_ids[k];
}

// ------------------------------------------------------------------------------------------------
// reset: everything a round's kernels accumulate into, in one launch (it was a memset per array: each one a submission on
// the critical path of a call that is otherwise five short kernels)
// ------------------------------------------------------------------------------------------------

__global__ void __launch_bounds__(kMvThreads)
mv_reset_kernel(const MvResetArgs r) {
    const int64_t t0 = (int64_t)blockIdx.x * kMvThreads + threadIdx.x, step = (int64_t)gridDim.x * kMvThreads;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        uint4 *p = reinterpret_cast<uint4 *>(r.ptr[k]);
        const uint32_t v = r.value[k];
        const uint4 fill = make_uint4(v, v, v, v);
        for (int64_t i = t0; i < r.quads[k]; i += step)
            p[i] = fill;
    }
    for (int64_t i = t0; i < r.copy_quads; i += step)
        r.copy_dst[i] = r.copy_src[i];
    if (t0 == 0) {
        r.state[1] = r.state[2] = r.state[3] = 0u;           // (the first 16 bytes are not among the arrays: word 0 outlives a round)
        if (r.first_round) {
            r.state[0] = 0u;
            r.csr_offsets[0] = 0ll;
        }
    }
}

void launch_mv_reset(const MvResetArgs &r, int num_sms, cudaStream_t s) {
    int64_t most = r.copy_quads > 1 ? r.copy_quads : 1;
    for (int k = 0; k < 4; k++)
        most = most > r.quads[k] ? most : r.quads[k];
    int64_t grid = (most + kMvThreads - 1) / kMvThreads;
    if (grid > (int64_t)num_sms * 4)
        grid = (int64_t)num_sms * 4;
    mv_reset_kernel<<<(unsigned)grid, kMvThreads, 0, s>>>(r);
}

// ------------------------------------------------------------------------------------------------
// scan: totals[view][block] -> exclusive prefixes (+ the view total in the extra column), then CSR offsets
// ------------------------------------------------------------------------------------------------

__global__ void __launch_bounds__(kMvThreads)
mv_scan_kernel(const __grid_constant__ MvArgs a) {
    chain_enter();
    __shared__ uint32_t s_warp[kMvThreads / 32];
    __shared__ long long s_warp64[kMvThreads / 32];
    __shared__ bool s_last;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nb = a.num_blocks;
    uint32_t *row = a.totals + (int64_t)blockIdx.x * (nb + 1);
    uint32_t carry = 0;
    for (int b0 = 0; b0 < nb; b0 += kMvThreads) {
        const int i = b0 + tid;
        const uint32_t c = i < nb ? row[i] : 0u;
        uint32_t total;
        const uint32_t rank = block_exclusive_scan(c, s_warp, &total);
        if (i < nb)
            row[i] = carry + rank;
        carry += total;
    }
    if (tid == 0) {
        row[nb] = carry;
        __threadfence();
        s_last = atomicAdd(a.state + 1, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!s_last)
        return;
    // the last CTA to finish: view totals -> CSR offsets (the pass starts where csr_offsets[0] says)
    __threadfence();
    long long running = a.csr_offsets[0];
    for (int v0 = 0; v0 < a.n_views; v0 += kMvThreads) {
        const int v = v0 + tid;
        const long long c = v < a.n_views ? (long long)__ldcg(a.totals + (int64_t)v * (nb + 1) + nb) : 0ll;
        long long incl = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            long long up = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o)
                incl += up;
        }
        __syncthreads();
        if (lane == 31)
            s_warp64[warp] = incl;
        __syncthreads();
        long long before = 0, sum = 0;
        for (int k = 0; k < kMvThreads / 32; k++) {
            if (k < warp)
                before += s_warp64[k];
            sum += s_warp64[k];
        }
        if (v < a.n_views) {
            a.csr_offsets[v + 1] = running + before + incl;
            if (a.host_offsets)                                 // page-locked host memory: the call's read-back without a copy
                a.host_offsets[v + 1] = running + before + incl;
        }
        running += sum;
    }
    if (a.host_offsets && tid < 8) {
        if (tid == 0)
            a.host_offsets[0] = a.csr_offsets[0];
        a.host_state[tid] = tid == 1 ? 0u : __ldcg(a.state + tid);
    }
    if (tid == 0)
        a.state[1] = 0u;
}

// ------------------------------------------------------------------------------------------------
// sparse pass
// ------------------------------------------------------------------------------------------------

__global__ void __launch_bounds__(kMvThreads)
mv_bin_lights_kernel(const __grid_constant__ MvArgs a) {
    chain_enter();
    const int g = blockIdx.x;                                  // one CTA per group: a large light covers hundreds of cells
    const MvGroup grp = a.groups[g];
    const int x0 = grp.cell_lo & 255, y0 = (grp.cell_lo >> 8) & 255, z0 = (grp.cell_lo >> 16) & 255;
    const int dx = (int)(grp.cell_hi & 255) - x0 + 1, dy = (int)((grp.cell_hi >> 8) & 255) - y0 + 1,
              dz = (int)((grp.cell_hi >> 16) & 255) - z0 + 1;
    const int cells = dx * dy * dz;
    for (int i = threadIdx.x; i < cells; i += kMvThreads) {
        const int x = x0 + i % dx, y = y0 + (i / dx) % dy, z = z0 + i / (dx * dy);
        const uint32_t node = atomicAdd(a.state + 2, 1u);
        if (node >= a.node_capacity) {
            atomicOr(a.state, kMvOverflowNodes);
            continue;
        }
        const int32_t next = atomicExch(a.heads + ((int64_t)z * a.grid.dim[1] + y) * a.grid.dim[0] + x, (int32_t)node);
        a.nodes[2 * node] = make_uint4((uint32_t)g | ((uint32_t)x0 << 12) | ((uint32_t)y0 << 18) | ((uint32_t)z0 << 24), (uint32_t)next,
                                       __float_as_uint(grp.lo[0]), __float_as_uint(grp.lo[1]));
        a.nodes[2 * node + 1] = make_uint4(__float_as_uint(grp.lo[2]), __float_as_uint(grp.hi[0]), __float_as_uint(grp.hi[1]),
                                           __float_as_uint(grp.hi[2]));
        atomicOr(a.bitmap + z * a.grid.dim[1] + y, 1ull << x);
    }
}

struct CellRange {
    int x0, x1, y0, y1, z0, z1;
};

// cell index of a coordinate: the same monotone function on the host (for the groups: fmaf, floor, clamp in float, convert)
// and here (for the entities: fma, convert to unsigned with round-down and saturation - negatives and NaN give 0 -, clamp),
// so two intervals that overlap in space always share a cell
__device__ __forceinline__ int cell_of(float v, float inv, float bias, int dim) {
    return (int)min(__float2uint_rd(__fmaf_rn(v, inv, bias)), (uint32_t)(dim - 1));
}

// K6 pre-filter, entity side: the box inflated by Re = 1.004 * largest extent + 1.002e-3 * G, G = the largest |coordinate|
// of the grid's bounds. (A box that some face of some cube group accepts overlaps that group's cube, which lies inside the
// grid's bounds, so none of its coordinates exceeds G + 2 * extent in size: Re is at least the 1.001 * extent + 1e-3 * largest
// |coordinate| the conservativeness argument asks for - DESIGN.md, K6 - without looking at the coordinates.)
// odd: a NaN or infinite side (the sum of the sides is then NaN or infinite; so is that of a finite box wider than 1e38,
// which is merely treated with the same caution).
__device__ __forceinline__ void inflate_box(const MvGrid &g, const float mn[3], const float mx[3], float lo[3], float hi[3], bool *odd) {
    // An inverted box (min > max on some axis; the reference tests it like any other, its "P-vertex" then being the least
    // favourable corner) is filtered as the box its two corners span: whatever the reference accepts of the inverted box
    // it accepts of the spanned one, whose overlap with the cube is what the argument is about.
    const float sx = fabsf(mx[0] - mn[0]), sy = fabsf(mx[1] - mn[1]), sz = fabsf(mx[2] - mn[2]);
    const float ext = fmaxf(fmaxf(sx, sy), sz);
    *odd = !((sx + sy) + sz <= 3.0e38f);
    const float re = __fmaf_rn(ext, 1.004f, g.re_bias);
#pragma unroll
    for (int c = 0; c < 3; c++) {
        lo[c] = fminf(mn[c], mx[c]) - re;
        hi[c] = fmaxf(mn[c], mx[c]) + re;
    }
}

// Cheap first look: can the (inflated) box touch the grid's bounds at all, or is it odd? Everything else would land in the
// border cells and leave with nothing.
__device__ __forceinline__ bool box_near_grid(const MvGrid &g, const float mn[3], const float mx[3]) {
    float lo[3], hi[3];
    bool odd;
    inflate_box(g, mn, mx, lo, hi, &odd);
    const bool outside = (hi[0] < g.lo[0]) | (lo[0] > g.hi[0]) | (hi[1] < g.lo[1]) | (lo[1] > g.hi[1]) | (hi[2] < g.lo[2]) | (lo[2] > g.hi[2]);
    return odd | !outside;
}

// Returns 1: candidates are in the cells of `r`, at most 2 x 2 rows of the bitmap (a box outside the grid's bounds lands in
// the border cells, which are never occupied); 2: test every group (an odd box can never be rejected; a box that spans more
// rows is cheaper to test against every cube). Branch-free on purpose: it runs for every entity of the world.
__device__ __forceinline__ int entity_cells(const MvGrid &g, const float mn[3], const float mx[3], float lo[3], float hi[3],
                                            CellRange &r, bool *odd) {
    inflate_box(g, mn, mx, lo, hi, odd);
    r.x0 = cell_of(lo[0], g.inv_cell[0], g.bias[0], g.dim[0]); r.x1 = cell_of(hi[0], g.inv_cell[0], g.bias[0], g.dim[0]);
    r.y0 = cell_of(lo[1], g.inv_cell[1], g.bias[1], g.dim[1]); r.y1 = cell_of(hi[1], g.inv_cell[1], g.bias[1], g.dim[1]);
    r.z0 = cell_of(lo[2], g.inv_cell[2], g.bias[2], g.dim[2]); r.z1 = cell_of(hi[2], g.inv_cell[2], g.bias[2], g.dim[2]);
    const bool big = (r.y1 - r.y0 > 1) | (r.z1 - r.z0 > 1);
    return (*odd | big) ? 2 : 1;
}

__device__ __forceinline__ unsigned long long x_range_bits(int x0, int x1) {
    return ((2ull << (x1 - x0)) - 1ull) << x0;              // x1 - x0 = 63: the shift wraps to 0 and 0 - 1 is all ones
}

// candidate iff the inflated box overlaps the group's (inflated) cube: separated along some axis <=> the largest of the
// six gaps is > 0 (a NaN gap is ignored by fmaxf, an all-NaN set compares false: such boxes stay candidates)
__device__ __forceinline__ bool overlaps_cube(const float glo[3], const float ghi[3], const float lo[3], const float hi[3]) {
    const float gap = fmaxf(fmaxf(fmaxf(glo[0] - hi[0], lo[0] - ghi[0]), fmaxf(glo[1] - hi[1], lo[1] - ghi[1])),
                            fmaxf(glo[2] - hi[2], lo[2] - ghi[2]));
    return !(gap > 0.0f);
}
__device__ __forceinline__ bool overlaps_cube(const MvGroup &grp, const float lo[3], const float hi[3]) {
    return overlaps_cube(grp.lo, grp.hi, lo, hi);
}

__device__ __forceinline__ void load_box(const MvArgs &a, int64_t e, float mn[3], float mx[3]) {
#pragma unroll
    for (int c = 0; c < 3; c++) {
        mn[c] = __ldg(a.bounds + c * a.stride + e);
        mx[c] = __ldg(a.bounds + (c + 3) * a.stride + e);
    }
}

// record key: group (12 bits) | entity inside the block - at first: record index - (13 bits) | face mask (6 bits); 0xffffffff = no record
__device__ __forceinline__ void append_pair(uint32_t *s_pairs, uint32_t *s_npairs, uint32_t g, uint32_t el) {
    const uint32_t i = atomicAdd(s_npairs, 1u);
    if (i < (uint32_t)kMvPairCapacity)
        s_pairs[i] = ((g << 13) | el) << 6;                    // the low six bits collect the faces that accept the box
}

struct SegCount {                 // per-face survivor counts of a (group, block) run, 16 bits each, and the segment flag
    uint32_t p0, p1, p2;
    bool cut;
};
__device__ __forceinline__ void seg_add(SegCount &s, uint32_t fm) {
    s.p0 += (fm & 1u) | (((fm >> 1) & 1u) << 16);
    s.p1 += ((fm >> 2) & 1u) | (((fm >> 3) & 1u) << 16);
    s.p2 += ((fm >> 4) & 1u) | (((fm >> 5) & 1u) << 16);
}
__device__ __forceinline__ uint32_t seg_get(const SegCount &s, int f) {
    const uint32_t p = f < 2 ? s.p0 : (f < 4 ? s.p1 : s.p2);
    return (p >> (16 * (f & 1))) & 0xffffu;
}
// `later` follows `earlier`: a cut in `later` discards what came before it
__device__ __forceinline__ SegCount seg_join(const SegCount &earlier, const SegCount &later) {
    SegCount r = later;
    if (!later.cut) {
        r.p0 += earlier.p0;
        r.p1 += earlier.p1;
        r.p2 += earlier.p2;
    }
    r.cut = earlier.cut || later.cut;
    return r;
}

// K6, entity side, streaming half: every entity of the world is classified against the occupancy bitmap - bit in `hit`:
// one of its cells is occupied, its candidates are found through the cells' lists; bit in `big`: it is tested against every
// group. One pass over the bounds at memory speed (the walk of K3: a warp takes 128 entities, 4 per lane, six 128-bit loads),
// nothing else happens here, so the loads never wait for the pair work of mv_sparse_kernel (round 2's first version did both
// in one CTA per block and streamed at 1.8 TB/s: profiles/r02_k4_config4_v2_plain_runs_sparse_ncu.csv).
__global__ void __launch_bounds__(kMvThreads, 4)
mv_classify_kernel(const __grid_constant__ MvArgs a) {
    chain_enter();
    extern __shared__ __align__(16) unsigned char s_mv[];
    unsigned long long *s_bitmap = reinterpret_cast<unsigned long long *>(s_mv);
    __shared__ float s_cbox[kMvThreads / 32][6][32];           // per warp: the boxes handed out one per lane (see below)
    __shared__ uint32_t s_cel[kMvThreads / 32][32];
    const int rows = a.grid.dim[1] * a.grid.dim[2];
    for (int i = threadIdx.x; i < rows; i += kMvThreads)
        s_bitmap[i] = a.bitmap[i];
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const long long warps_total = (long long)gridDim.x * (kMvThreads / 32);
    const float4 *__restrict__ bounds = reinterpret_cast<const float4 *>(a.bounds);
    const int64_t vstride = a.stride >> 2;
    const int dim_y = a.grid.dim[1];
    int straight_left = 0;
    for (long long group = (long long)blockIdx.x * (kMvThreads / 32) + (threadIdx.x >> 5); group < a.num_groups128; group += warps_total) {
        const int64_t v = group * 32 + lane;
        const int64_t e0 = v << 2;
        const float4 ax = __ldg(bounds + 0 * vstride + v), ay = __ldg(bounds + 1 * vstride + v), az = __ldg(bounds + 2 * vstride + v);
        const float4 bx = __ldg(bounds + 3 * vstride + v), by = __ldg(bounds + 4 * vstride + v), bz = __ldg(bounds + 5 * vstride + v);
        uint32_t alive = candidate_bits<false>(e0, a.n, 0u);
        if (a.alive_mask)
            alive &= (__ldg(a.alive_mask + (e0 >> 5)) >> (e0 & 31)) & 0xfu;
        const float mnv[4][3] = { { ax.x, ay.x, az.x }, { ax.y, ay.y, az.y }, { ax.z, ay.z, az.z }, { ax.w, ay.w, az.w } };
        const float mxv[4][3] = { { bx.x, by.x, bz.x }, { bx.y, by.y, bz.y }, { bx.z, by.z, bz.z }, { bx.w, by.w, bz.w } };
        // When the lights sit in a corner of the world (a camera's visible lights) most entities cannot touch the grid's
        // bounds. A cheap test finds those that can; if the warp's 128 hold at most 32 of them they are handed out one per
        // lane through shared memory and the cell work below runs once instead of four times.
        // (Where the grid spans the world that first look is wasted: a warp that found more than 32 skips it for its next
        // seven groups.)
        const int b = (int)(group / kMaskBlockGroups);
        float4 *__restrict__ list = a.hit_records + (int64_t)b * (2 * a.hit_capacity);
        const uint32_t el0 = (uint32_t)(e0 & (kMaskBlockEntities - 1));
        uint32_t near = 0, n0 = 0, n1 = 0, n2 = 0, n3 = 0, total_near = 128u;
        if (straight_left > 0) {
            straight_left--;
        } else {
#pragma unroll
            for (int j = 0; j < 4; j++)
                near |= (box_near_grid(a.grid, mnv[j], mxv[j]) ? 1u : 0u) << j;
            near &= alive;
            n0 = __ballot_sync(0xffffffffu, near & 1u); n1 = __ballot_sync(0xffffffffu, near & 2u);
            n2 = __ballot_sync(0xffffffffu, near & 4u); n3 = __ballot_sync(0xffffffffu, near & 8u);
            total_near = __popc(n0) + __popc(n1) + __popc(n2) + __popc(n3);
            if (total_near == 0u)
                continue;
            if (total_near > 32u)
                straight_left = 7;
        }
        if (total_near <= 32u) {
            const int warp = threadIdx.x >> 5;
            const uint32_t lt = (1u << lane) - 1u;
            const uint32_t ballots[4] = { n0, n1, n2, n3 };
            uint32_t first = 0;
#pragma unroll
            for (int j = 0; j < 4; j++) {
                if ((near >> j) & 1u) {
                    const uint32_t slot = first + __popc(ballots[j] & lt);
#pragma unroll
                    for (int c = 0; c < 3; c++) {
                        s_cbox[warp][c][slot] = mnv[j][c];
                        s_cbox[warp][c + 3][slot] = mxv[j][c];
                    }
                    s_cel[warp][slot] = el0 + j;
                }
                first += __popc(ballots[j]);
            }
            __syncwarp();
            bool mine = false, is_big = false;
            float mn[3] = { 0.f, 0.f, 0.f }, mx[3] = { 0.f, 0.f, 0.f };
            uint32_t el = 0;
            if ((uint32_t)lane < total_near) {
#pragma unroll
                for (int c = 0; c < 3; c++) {
                    mn[c] = s_cbox[warp][c][lane];
                    mx[c] = s_cbox[warp][c + 3][lane];
                }
                el = s_cel[warp][lane];
                float lo[3], hi[3];
                CellRange r;
                bool odd;
                const int kind = entity_cells(a.grid, mn, mx, lo, hi, r, &odd);
                const unsigned long long seen = (s_bitmap[r.z0 * dim_y + r.y0] | s_bitmap[r.z0 * dim_y + r.y1]) |
                                                (s_bitmap[r.z1 * dim_y + r.y0] | s_bitmap[r.z1 * dim_y + r.y1]);
                const bool occupied = ((seen >> r.x0) << (63 - (r.x1 - r.x0))) != 0ull;
                is_big = kind == 2;
                mine = is_big | ((kind == 1) & occupied);
            }
            const uint32_t who = __ballot_sync(0xffffffffu, mine);
            if (who) {
                uint32_t base = 0;
                if (lane == 0)
                    base = atomicAdd(a.block_work + b, (uint32_t)__popc(who));
                base = __shfl_sync(0xffffffffu, base, 0) + __popc(who & lt);
                if (mine && base < (uint32_t)a.hit_capacity) {
                    list[2 * base] = make_float4(mn[0], mn[1], mn[2], mx[0]);
                    list[2 * base + 1] = make_float4(mx[1], mx[2], __uint_as_float(el | ((is_big ? 1u : 0u) << 31)), 0.0f);
                }
            }
            __syncwarp();                                        // the warp's slots are rewritten by its next group
            continue;
        }
        uint32_t hit = 0, big = 0;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            float lo[3], hi[3];
            CellRange r;
            bool odd;
            const int kind = entity_cells(a.grid, mnv[j], mxv[j], lo, hi, r, &odd);
            // kind 1 spans at most 2 x 2 rows: look at the four corner rows (duplicates when it spans fewer)
            const unsigned long long seen = (s_bitmap[r.z0 * dim_y + r.y0] | s_bitmap[r.z0 * dim_y + r.y1]) |
                                            (s_bitmap[r.z1 * dim_y + r.y0] | s_bitmap[r.z1 * dim_y + r.y1]);
            const bool occupied = ((seen >> r.x0) << (63 - (r.x1 - r.x0))) != 0ull;       // any of bits x0 .. x1
            big |= (kind == 2 ? 1u : 0u) << j;
            hit |= ((kind == 1) & occupied ? 1u : 0u) << j;
        }
        // the classified entities leave as records (box + index inside the block) appended to their block's list: what
        // mv_sparse_kernel needs of them arrives there in one coalesced read (reading the boxes back out of the SoA arrays
        // cost it six sectors per entity, as much traffic as this whole pass). One atomic per group of 128 reserves the slots
        // (one per entity-of-the-lane round was tried: four dependent round trips per group, the pass went from 77 to 93 us).
        // Order inside a list is whatever the atomics give; the records are sorted further down the line.
        const uint32_t mine = (hit | big) & alive;
        if (__any_sync(0xffffffffu, mine != 0u)) {
            const uint32_t count = __popc(mine);
            uint32_t incl = count;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t up = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o)
                    incl += up;
            }
            uint32_t base = 0;
            if (lane == 31)
                base = atomicAdd(a.block_work + b, incl);
            base = __shfl_sync(0xffffffffu, base, 31) + incl - count;
#pragma unroll
            for (int j = 0; j < 4; j++)
                if ((mine >> j) & 1u) {
                    if (base < (uint32_t)a.hit_capacity) {
                        const uint32_t tag = (el0 + j) | (((big >> j) & 1u) << 31);
                        list[2 * base] = make_float4(mnv[j][0], mnv[j][1], mnv[j][2], mxv[j][0]);
                        list[2 * base + 1] = make_float4(mxv[j][1], mxv[j][2], __uint_as_float(tag), 0.0f);
                    }
                    base++;
                }
        }
    }
}

// K6, entity side, pair half: one CTA per block of 8192 entities that has classified entities.
__global__ void __launch_bounds__(kSpThreads, 6)
mv_sparse_kernel(const __grid_constant__ MvArgs a) {
    chain_enter();
    __shared__ __align__(16) uint32_t s_pairs[kMvPairCapacity];
    __shared__ uint32_t s_big[kMvBigSlots];                    // records tested against every group (huge or odd boxes)
    __shared__ uint32_t s_npairs, s_nrec, s_item_base, s_nbig;
    __shared__ uint32_t s_warp[kSpThreads / 32];
    __shared__ SegCount s_seg[kSpThreads / 32];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int b = blockIdx.x;
    const int64_t block_e0 = (int64_t)b * kMaskBlockEntities;
    const uint32_t n_hits = a.block_work[b];
    if (n_hits == 0u || n_hits > (uint32_t)a.hit_capacity) {
        if (tid == 0) {
            if (n_hits) {                                      // the block's list was too short: the host enlarges the lists and reruns
                atomicOr(a.state, kMvOverflowHits);
                atomicMax(a.state + 4, n_hits);
            }
            a.segments[b] = make_uint2(0u, 0u);
        }
        return;
    }
    if (tid == 0) {
        s_npairs = 0;
        s_nrec = 0;
        s_nbig = 0;
    }
    __syncthreads();
    const unsigned long long *__restrict__ bitmap = a.bitmap;
    const float4 *__restrict__ list = a.hit_records + (int64_t)b * (2 * a.hit_capacity);

    // ---- (2) candidate pairs: a thread takes every 256th record of the block
    float4 n0 = make_float4(0.f, 0.f, 0.f, 0.f), n1 = n0;       // the next record is requested before this one's lists are walked
    if ((uint32_t)tid < n_hits) {
        n0 = __ldg(list + 2 * tid);
        n1 = __ldg(list + 2 * tid + 1);
    }
    for (uint32_t h = tid; h < n_hits; h += kSpThreads) {
        const float4 r0 = n0, r1 = n1;
        if (h + kSpThreads < n_hits) {
            n0 = __ldg(list + 2 * (h + kSpThreads));
            n1 = __ldg(list + 2 * (h + kSpThreads) + 1);
        }
        const float mn[3] = { r0.x, r0.y, r0.z }, mx[3] = { r0.w, r1.x, r1.y };
        const uint32_t tag = __float_as_uint(r1.z);
        float lo[3], hi[3];
        CellRange r;
        bool odd;
        entity_cells(a.grid, mn, mx, lo, hi, r, &odd);
        if (tag >> 31) {
            const uint32_t slot = atomicAdd(&s_nbig, 1u);
            if (slot < (uint32_t)kMvBigSlots) {
                s_big[slot] = h;
            } else {                                           // (more such boxes in one block than slots: this thread alone)
                for (int g = 0; g < a.n_groups; g++)
                    if (odd || overlaps_cube(a.groups[g], lo, hi))
                        append_pair(s_pairs, &s_npairs, (uint32_t)g, h);
            }
            continue;
        }
        const unsigned long long xm = x_range_bits(r.x0, r.x1);
        for (int z = r.z0; z <= r.z1; z++)
            for (int y = r.y0; y <= r.y1; y++) {
                unsigned long long occupied = __ldg(bitmap + z * a.grid.dim[1] + y) & xm;
                while (occupied) {
                    const int x = __ffsll((long long)occupied) - 1;
                    occupied &= occupied - 1;
                    for (int32_t node = a.heads[((int64_t)z * a.grid.dim[1] + y) * a.grid.dim[0] + x]; node >= 0;) {
                        const uint4 q0 = __ldg(a.nodes + 2 * node), q1 = __ldg(a.nodes + 2 * node + 1);
                        node = (int32_t)q0.y;
                        // a group shared by several of the entity's cells is taken in the first of them only
                        const int gx = max(r.x0, (int)((q0.x >> 12) & 63u)), gy = max(r.y0, (int)((q0.x >> 18) & 63u)),
                                  gz = max(r.z0, (int)((q0.x >> 24) & 63u));
                        if (x != gx || y != gy || z != gz)
                            continue;
                        const float glo[3] = { __uint_as_float(q0.z), __uint_as_float(q0.w), __uint_as_float(q1.x) };
                        const float ghi[3] = { __uint_as_float(q1.y), __uint_as_float(q1.z), __uint_as_float(q1.w) };
                        if (overlaps_cube(glo, ghi, lo, hi))
                            append_pair(s_pairs, &s_npairs, q0.x & 4095u, h);
                    }
                }
            }
    }
    __syncthreads();
    // records tested against every group: a warp takes a record, its lanes every 32nd group (the whole CTA on one record at
    // a time, as at first, spends eight warps on the 37 groups of a camera's visible lights: 9 M instructions for configs[2])
    for (uint32_t k = warp, nbig = min(s_nbig, (uint32_t)kMvBigSlots); k < nbig; k += kSpThreads / 32) {
        const uint32_t h = s_big[k];
        const float4 r0 = __ldg(list + 2 * h), r1 = __ldg(list + 2 * h + 1);
        const float mn[3] = { r0.x, r0.y, r0.z }, mx[3] = { r0.w, r1.x, r1.y };
        float lo[3], hi[3];
        bool odd;
        inflate_box(a.grid, mn, mx, lo, hi, &odd);
        for (int g = lane; g < a.n_groups; g += 32)
            if (odd || overlaps_cube(a.groups[g], lo, hi))
                append_pair(s_pairs, &s_npairs, (uint32_t)g, h);
    }
    __syncthreads();
    const uint32_t npairs = s_npairs;
    if (npairs > (uint32_t)kMvPairCapacity) {                   // too dense for this pass: the host reruns these views densely
        if (tid == 0) {
            atomicOr(a.state, kMvOverflowPairs);
            a.segments[b] = make_uint2(0u, 0u);
        }
        return;
    }
    if (npairs == 0) {
        if (tid == 0)
            a.segments[b] = make_uint2(0u, 0u);
        return;
    }

    // ---- (3) exact tests (Math::intersect, intersection.cc:42-57): one thread per (pair, face), the six planes of the face
    // requested together - a thread per pair walking its faces plane by plane was up to 36 dependent trips to L2, the
    // longest chain of a block with few pairs. A face that accepts the box sets its bit in the pair's word.
    for (uint32_t t = tid; t < npairs * (uint32_t)kMvMaxGroupViews; t += kSpThreads) {
        const uint32_t i = t / (uint32_t)kMvMaxGroupViews, f = t - i * (uint32_t)kMvMaxGroupViews;
        const uint32_t pair = s_pairs[i];
        const uint32_t g = pair >> 19, h = (pair >> 6) & 8191u;   // (group, record): the box comes from the record, one sector
        const int2 views = __ldg(reinterpret_cast<const int2 *>(&a.groups[g].first_view));     // (first_view, n_views)
        if ((int)f >= views.y)
            continue;
        const int view = views.x + (int)f;
        const float4 r0 = __ldg(list + 2 * h), r1 = __ldg(list + 2 * h + 1);
        const float4 *pl = a.planes + 6 * view;
        const float4 q0 = __ldg(pl), q1 = __ldg(pl + 1), q2 = __ldg(pl + 2), q3 = __ldg(pl + 3), q4 = __ldg(pl + 4), q5 = __ldg(pl + 5);
        if (__ldg(a.view_meta + view) >> 31) {                  // shadow casters only
            const int64_t e = block_e0 + (__float_as_uint(r1.z) & (kMaskBlockEntities - 1));
            if (a.caster_mask && ((__ldg(a.caster_mask + (e >> 5)) >> (e & 31)) & 1u) == 0u)
                continue;
        }
        const float mn[3] = { r0.x, r0.y, r0.z }, mx[3] = { r0.w, r1.x, r1.y };
        const bool outside = outside_plane(q0.x, q0.y, q0.z, q0.w, mn[0], mn[1], mn[2], mx[0], mx[1], mx[2]) |
                             outside_plane(q1.x, q1.y, q1.z, q1.w, mn[0], mn[1], mn[2], mx[0], mx[1], mx[2]) |
                             outside_plane(q2.x, q2.y, q2.z, q2.w, mn[0], mn[1], mn[2], mx[0], mx[1], mx[2]) |
                             outside_plane(q3.x, q3.y, q3.z, q3.w, mn[0], mn[1], mn[2], mx[0], mx[1], mx[2]) |
                             outside_plane(q4.x, q4.y, q4.z, q4.w, mn[0], mn[1], mn[2], mx[0], mx[1], mx[2]) |
                             outside_plane(q5.x, q5.y, q5.z, q5.w, mn[0], mn[1], mn[2], mx[0], mx[1], mx[2]);
        if (!outside)
            atomicOr(&s_pairs[i], 1u << f);
    }
    __syncthreads();
    // the pairs some face accepted become records (group, entity, faces), the others are dropped
    uint32_t n_valid = 0;
    for (uint32_t i = tid; i < npairs; i += kSpThreads) {
        const uint32_t pair = s_pairs[i], fm = pair & 63u;
        uint32_t key = 0xffffffffu;
        if (fm) {
            const uint32_t el = __float_as_uint(__ldg(&list[2 * ((pair >> 6) & 8191u) + 1].z)) & (kMaskBlockEntities - 1);
            key = ((((pair >> 19) << 13) | el) << 6) | fm;
            n_valid++;
        }
        s_pairs[i] = key;
    }
    n_valid = __reduce_add_sync(0xffffffffu, n_valid);
    if (lane == 0 && n_valid)
        atomicAdd(&s_nrec, n_valid);

    // ---- (4) sort the surviving records by (group, entity), in shared memory. A few hundred pairs (the usual block) are
    // ranked by counting: every thread counts the keys below its own - one barrier instead of the 36 of a bitonic network
    // over 256 keys (ncu: those barriers were a sixth of the kernel). Keys are unique (a pair is appended once).
    __syncthreads();
    const uint32_t nrec = s_nrec;
    if (nrec == 0) {
        if (tid == 0)
            a.segments[b] = make_uint2(0u, 0u);
        return;
    }
    const uint32_t *s_rec = s_pairs;
    if (npairs <= (uint32_t)kMvRankSortBelow) {
        uint32_t *s_out = s_pairs + kMvPairCapacity / 2;
        const uint32_t padded = (npairs + 3u) & ~3u;
        for (uint32_t i = npairs + tid; i < padded; i += kSpThreads)
            s_pairs[i] = 0xffffffffu;
        __syncthreads();
        for (uint32_t i = tid; i < npairs; i += kSpThreads) {
            const uint32_t key = s_pairs[i];
            if (key == 0xffffffffu)
                continue;
            uint32_t rank = 0;
            for (uint32_t j = 0; j < padded; j += 4) {
                const uint4 q = *reinterpret_cast<const uint4 *>(s_pairs + j);
                rank += (q.x < key ? 1u : 0u) + (q.y < key ? 1u : 0u) + (q.z < key ? 1u : 0u) + (q.w < key ? 1u : 0u);
            }
            s_out[rank] = key;
        }
        __syncthreads();
        s_rec = s_out;
    } else {
        uint32_t n_sort = 32;
        while (n_sort < npairs)
            n_sort <<= 1;
        for (uint32_t i = npairs + tid; i < n_sort; i += kSpThreads)
            s_pairs[i] = 0xffffffffu;
        __syncthreads();
        for (uint32_t k = 2; k <= n_sort; k <<= 1)
            for (uint32_t j = k >> 1; j > 0; j >>= 1) {
                for (uint32_t i = tid; i < n_sort; i += kSpThreads) {
                    const uint32_t partner = i ^ j;
                    if (partner > i) {
                        const uint32_t x = s_pairs[i], y = s_pairs[partner];
                        const bool ascending = (i & k) == 0;
                        if ((x > y) == ascending) {
                            s_pairs[i] = y;
                            s_pairs[partner] = x;
                        }
                    }
                }
                __syncthreads();
            }
    }

    // ---- (5) rank inside (view, block), totals, items. Thread t owns records [t * per, (t + 1) * per).
    const uint32_t per = (nrec + kSpThreads - 1) / kSpThreads;
    const uint32_t i0 = min(nrec, (uint32_t)tid * per), i1 = min(nrec, i0 + per);
    SegCount mine = { 0u, 0u, 0u, false };
    uint32_t my_items = 0;
    {
        uint32_t prev_g = i0 > 0 ? s_rec[i0 - 1] >> 19 : 0xffffffffu;
        for (uint32_t i = i0; i < i1; i++) {
            const uint32_t key = s_rec[i], g = key >> 19, fm = key & 63u;
            if (g != prev_g) {
                mine = { 0u, 0u, 0u, true };
                prev_g = g;
            }
            seg_add(mine, fm);
            my_items += __popc(fm);
        }
    }
    // segmented exclusive scan of `mine` over the threads (a cut = a new group starts inside the thread's records)
    SegCount incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        SegCount up;
        up.p0 = __shfl_up_sync(0xffffffffu, incl.p0, o);
        up.p1 = __shfl_up_sync(0xffffffffu, incl.p1, o);
        up.p2 = __shfl_up_sync(0xffffffffu, incl.p2, o);
        up.cut = __shfl_up_sync(0xffffffffu, (int)incl.cut, o) != 0;
        if (lane >= o)
            incl = seg_join(up, incl);
    }
    SegCount carry;                                          // what precedes this thread inside its warp
    carry.p0 = __shfl_up_sync(0xffffffffu, incl.p0, 1);
    carry.p1 = __shfl_up_sync(0xffffffffu, incl.p1, 1);
    carry.p2 = __shfl_up_sync(0xffffffffu, incl.p2, 1);
    carry.cut = __shfl_up_sync(0xffffffffu, (int)incl.cut, 1) != 0;
    if (lane == 0)
        carry = { 0u, 0u, 0u, false };
    if (lane == 31)
        s_seg[warp] = incl;
    uint32_t items_total;
    const uint32_t item_rank = block_exclusive_scan<kSpThreads>(my_items, s_warp, &items_total);   // (contains the barriers s_seg needs)
    {
        SegCount before = { 0u, 0u, 0u, false };
        for (int k = 0; k < warp; k++)
            before = seg_join(before, s_seg[k]);
        carry = seg_join(before, carry);
    }
    if (tid == 0) {
        s_item_base = atomicAdd(a.state + 3, items_total);
        a.segments[b] = make_uint2(s_item_base, items_total);
    }
    __syncthreads();
    const uint32_t item_base = s_item_base;
    const bool fits = (unsigned long long)item_base + items_total <= (unsigned long long)a.item_capacity;
    if (!fits && tid == 0) {
        atomicOr(a.state, kMvOverflowItems);
        a.segments[b] = make_uint2(0u, 0u);
    }
    {
        SegCount run = carry;
        uint32_t prev_g = i0 > 0 ? s_rec[i0 - 1] >> 19 : 0xffffffffu;
        uint32_t slot = item_base + item_rank;
        int first_view = 0;
        const int row = a.num_blocks + 1;
        for (uint32_t i = i0; i < i1; i++) {
            const uint32_t key = s_rec[i], g = key >> 19, el = (key >> 6) & 8191u;
            uint32_t fm = key & 63u;
            if (g != prev_g || i == i0) {
                if (g != prev_g)
                    run = { 0u, 0u, 0u, false };
                prev_g = g;
                first_view = a.groups[g].first_view;
            }
            const uint32_t id = (uint32_t)(block_e0 + el) + a.id_base;
            for (uint32_t bits = fm; bits; bits &= bits - 1) {
                const int f = __ffs(bits) - 1;
                if (fits)
                    a.items[slot] = { (uint32_t)(first_view + f), seg_get(run, f), id };
                slot++;
            }
            seg_add(run, fm);
            const uint32_t next_g = i + 1 < nrec ? s_rec[i + 1] >> 19 : 0xffffffffu;
            if (next_g != g) {                                  // the group's run ends here: its totals for this block
#pragma unroll
                for (int f = 0; f < kMvMaxGroupViews; f++) {
                    const uint32_t c = seg_get(run, f);
                    if (c)
                        a.totals[(int64_t)(first_view + f) * row + b] = c;
                }
            }
        }
    }
}

__global__ void __launch_bounds__(kMvThreads)
mv_expand_items_kernel(const __grid_constant__ MvArgs a) {
    chain_enter();
    const int row = a.num_blocks + 1;
    for (int b = blockIdx.x; b < a.num_blocks; b += gridDim.x) {
        const uint2 seg = a.segments[b];
        for (uint32_t i = threadIdx.x; i < seg.y; i += kMvThreads) {
            const MvItem it = a.items[seg.x + i];
            const long long pos = a.csr_offsets[it.view] + (long long)a.totals[(int64_t)it.view * row + b] + it.rank;
            if (pos < a.capacity)
                a.out[pos] = it.id;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// launchers
// ------------------------------------------------------------------------------------------------

void launch_mv_plain_test(const MvArgs &a, int num_sms, cudaStream_t s) {
    if (a.n_plain <= 0 || a.num_groups128 <= 0)
        return;
    const int n_chunks = (a.n_plain + a.views_per_warp - 1) / a.views_per_warp;
    const long long items = (long long)a.num_groups128 * n_chunks;
    static int per_sm = 0;
    if (!per_sm) {
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, mv_plain_test_kernel, kPlainThreads, 0);
        if (per_sm < 1)
            per_sm = 1;
    }
    long long grid = (items + kPlainThreads / 32 - 1) / (kPlainThreads / 32);
    if (grid > (long long)num_sms * per_sm)
        grid = (long long)num_sms * per_sm;
    launch_chained(mv_plain_test_kernel, dim3((unsigned)grid), 0, s, a, kPlainThreads);
}

void launch_mv_expand_masks(const MvArgs &a, cudaStream_t s) {
    if (a.n_plain <= 0 || a.num_blocks <= 0)
        return;
    launch_chained(mv_expand_masks_kernel, dim3((unsigned)a.num_blocks, (unsigned)a.n_plain), 0, s, a);
}

void launch_mv_scan(const MvArgs &a, cudaStream_t s) {
    if (a.n_views > 0)
        launch_chained(mv_scan_kernel, dim3((unsigned)a.n_views), 0, s, a);
}

void launch_mv_bin_lights(const MvArgs &a, cudaStream_t s) {
    if (a.n_groups > 0)
        launch_chained(mv_bin_lights_kernel, dim3((unsigned)a.n_groups), 0, s, a);
}

void launch_mv_sparse(const MvArgs &a, int num_sms, cudaStream_t s) {
    if (a.n_groups <= 0 || a.num_blocks <= 0)
        return;
    const size_t smem = sizeof(unsigned long long) * (size_t)(a.grid.dim[1] * a.grid.dim[2]);     // <= 32 KB
    long long grid = ((long long)a.num_groups128 + kMvThreads / 32 - 1) / (kMvThreads / 32);
    if (grid > (long long)num_sms * 4)
        grid = (long long)num_sms * 4;
    launch_chained(mv_classify_kernel, dim3((unsigned)grid), smem, s, a);
    launch_chained(mv_sparse_kernel, dim3((unsigned)a.num_blocks), 0, s, a, kSpThreads);
}

void launch_mv_expand_items(const MvArgs &a, cudaStream_t s) {
    if (a.n_groups <= 0 || a.num_blocks <= 0)
        return;
    launch_chained(mv_expand_items_kernel, dim3((unsigned)(a.num_blocks < 1184 ? a.num_blocks : 1184)), 0, s, a);
}

}  // namespace ovis
