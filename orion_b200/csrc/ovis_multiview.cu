// ovis_multiview.cu - K4: many views in one pass over the bounds, with K6's conservative pre-filter.
//
// Replaces the reference's "one full world cull per shadow view" (render/src/deferred_render_pipeline.cc:259-284:
// 1 + sum of shadow views full scans per frame) and serves the main camera + cascade views the same way.
//
// Views come in GROUPS: the views of one group share a pre-filter ball-cube (a point light's six cube faces share the
// cube of half-side `range` about the light); a group without a cube has an infinite radius. Per pass (at most
// kMultiMaxViews views):
//
//   A  multi_pass_kernel<false>  every warp walks "supertiles" of 1024 consecutive entities (4 x 256): loads the bounds
//                                 once, and for every group runs the cube pre-filter (12 instructions per entity) and,
//                                 only where a candidate exists in the warp, the exact six-plane tests of its views with
//                                 warp-uniform early-out. It counts survivors per (supertile, view): one coalesced row of
//                                 counts per supertile plus sparse integer atomicAdds into per-(64-supertile block, view)
//                                 totals.
//   B1 multi_scan_blocks_kernel  per view: exclusive scan of the block totals, view totals, CSR offsets of the pass.
//   B2 multi_scan_tiles_kernel   per block: exclusive scan over its 64 supertiles, in place (counts -> positions).
//   C  multi_pass_kernel<true>   the same walk again, now writing each survivor id at its final position.
//
// Positions are sums of counts, so every list is in slot (= insertion) order and identical from run to run. The exact
// tests are the same instructions as K3 (ovis_device.cuh: outside_plane), so lists are bit-identical to brute force;
// the pre-filter only ever skips (entity, group) pairs that every one of the group's views would reject
// (derivation in DESIGN.md section 4, K6).
#include "ovis_kernels.h"

#include "ovis_device.cuh"

namespace ovis {

constexpr int kSuperTile = 1024;                 // entities per supertile (one warp, 4 sub-tiles of 256)
constexpr int kSubTiles = 4;
constexpr int kMultiWarps = 8;                   // warps per CTA

template <bool WRITE>
__global__ void __launch_bounds__(kMultiWarps * 32, 2)
multi_pass_kernel(const __grid_constant__ MultiArgs a) {
    extern __shared__ uint32_t s_counters[];     // kMultiWarps x n_views: survivors so far in this supertile, per view
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t *counters = s_counters + warp * a.n_views;
    const int warps_total = gridDim.x * kMultiWarps;
    const float4 *__restrict__ bounds = reinterpret_cast<const float4 *>(a.bounds);
    const int64_t vstride = a.stride >> 2;

    for (int tile = blockIdx.x * kMultiWarps + warp; tile < a.num_tiles; tile += warps_total) {
        for (int v = lane; v < a.n_views; v += 32)
            counters[v] = 0u;
        __syncwarp();
        const uint32_t *row = WRITE ? a.counts + (int64_t)tile * a.n_views : nullptr;   // positions after pass B2

        for (int sub = 0; sub < kSubTiles; sub++) {
            // ---- load 8 entities per lane (two float4 groups), as K3 does
            float mn[3][8], mx[3][8];
            uint32_t alive = 0xffu, caster = 0xffu;
            uint32_t ebase[2];
#pragma unroll
            for (int g = 0; g < 2; g++) {
                const int64_t v = (int64_t)tile * (kSuperTile / 4) + sub * 64 + g * 32 + lane;    // float4 index
                const int64_t e0 = v << 2;
                ebase[g] = (uint32_t)e0;
                const float4 ax = __ldg(bounds + 0 * vstride + v), ay = __ldg(bounds + 1 * vstride + v), az = __ldg(bounds + 2 * vstride + v);
                const float4 bx = __ldg(bounds + 3 * vstride + v), by = __ldg(bounds + 4 * vstride + v), bz = __ldg(bounds + 5 * vstride + v);
                mn[0][4 * g + 0] = ax.x; mn[0][4 * g + 1] = ax.y; mn[0][4 * g + 2] = ax.z; mn[0][4 * g + 3] = ax.w;
                mn[1][4 * g + 0] = ay.x; mn[1][4 * g + 1] = ay.y; mn[1][4 * g + 2] = ay.z; mn[1][4 * g + 3] = ay.w;
                mn[2][4 * g + 0] = az.x; mn[2][4 * g + 1] = az.y; mn[2][4 * g + 2] = az.z; mn[2][4 * g + 3] = az.w;
                mx[0][4 * g + 0] = bx.x; mx[0][4 * g + 1] = bx.y; mx[0][4 * g + 2] = bx.z; mx[0][4 * g + 3] = bx.w;
                mx[1][4 * g + 0] = by.x; mx[1][4 * g + 1] = by.y; mx[1][4 * g + 2] = by.z; mx[1][4 * g + 3] = by.w;
                mx[2][4 * g + 0] = bz.x; mx[2][4 * g + 1] = bz.y; mx[2][4 * g + 2] = bz.z; mx[2][4 * g + 3] = bz.w;
                uint32_t valid = 0xfu;
                if (e0 + 3 >= a.n) {
                    int64_t left = a.n - e0;
                    valid = left <= 0 ? 0u : ((1u << left) - 1u);
                }
                uint32_t al = valid, ca = valid;
                if (a.alive_mask)
                    al &= (__ldg(a.alive_mask + (e0 >> 5)) >> (e0 & 31)) & 0xfu;
                if (a.caster_mask)
                    ca &= (__ldg(a.caster_mask + (e0 >> 5)) >> (e0 & 31)) & 0xfu;       // caster mask already implies alive
                else
                    ca = al;
                alive = (alive & ~(0xfu << (4 * g))) | (al << (4 * g));
                caster = (caster & ~(0xfu << (4 * g))) | (ca << (4 * g));
            }
            if (!__any_sync(0xffffffffu, alive))
                continue;

            // ---- per-entity terms of the pre-filter: largest extent and largest coordinate magnitude
            float ext[8], mag[8];
#pragma unroll
            for (int j = 0; j < 8; j++) {
                const float sx = mx[0][j] - mn[0][j], sy = mx[1][j] - mn[1][j], sz = mx[2][j] - mn[2][j];
                ext[j] = fmaxf(fmaxf(fmaxf(sx, sy), sz), 0.0f);
                mag[j] = fmaxf(fmaxf(fmaxf(fabsf(mn[0][j]), fabsf(mx[0][j])), fmaxf(fabsf(mn[1][j]), fabsf(mx[1][j]))),
                               fmaxf(fabsf(mn[2][j]), fabsf(mx[2][j])));
                // fmaxf drops NaNs: a box with a NaN or non-finite side must never be filtered out
                const bool odd = sx != sx || sy != sy || sz != sz || !(ext[j] <= 3.0e38f) || !(mag[j] <= 3.0e38f);
                if (odd)
                    ext[j] = __int_as_float(0x7f800000);     // +inf: the cube test below can then never reject
            }

            for (int gi = 0; gi < a.n_groups; gi++) {
                const ViewGroupDev grp = a.groups[gi];
                // ---- K6 pre-filter: candidate iff the box overlaps the cube of half-side (range + extent + margin)
                uint32_t cand = 0;
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    const float R = (grp.r + ext[j]) * 1.001f + (grp.pmax + mag[j]) * 1.0e-3f + 1.0e-2f;
                    const bool out = (mx[0][j] < grp.px - R) | (mn[0][j] > grp.px + R) | (mx[1][j] < grp.py - R) |
                                     (mn[1][j] > grp.py + R) | (mx[2][j] < grp.pz - R) | (mn[2][j] > grp.pz + R);
                    cand |= out ? 0u : (1u << j);
                }
                cand &= alive;
                if (!__any_sync(0xffffffffu, cand))
                    continue;

                for (int vi = 0; vi < grp.n_views; vi++) {
                    const int view = grp.first_view + vi;
                    uint32_t keep = a.view_require[view] ? (cand & caster) : cand;
                    const float4 *pl = a.planes + 6 * view;
#pragma unroll 1
                    for (int p = 0; p < 6; p++) {
                        if (!__any_sync(0xffffffffu, keep))
                            break;
                        const float4 q = __ldg(pl + p);
#pragma unroll
                        for (int j = 0; j < 8; j++)
                            if (outside_plane(q.x, q.y, q.z, q.w, mn[0][j], mn[1][j], mn[2][j], mx[0][j], mx[1][j], mx[2][j]))
                                keep &= ~(1u << j);
                    }
                    if (!__any_sync(0xffffffffu, keep))
                        continue;
                    // ---- rank in slot order: group 0 of all lanes comes before group 1 of all lanes
                    const uint32_t c0 = __popc(keep & 0xfu), c1 = __popc(keep >> 4);
                    uint32_t i0 = c0, i1 = c1;
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) {
                        uint32_t u0 = __shfl_up_sync(0xffffffffu, i0, o), u1 = __shfl_up_sync(0xffffffffu, i1, o);
                        if (lane >= o) { i0 += u0; i1 += u1; }
                    }
                    const uint32_t t0 = __shfl_sync(0xffffffffu, i0, 31), t1 = __shfl_sync(0xffffffffu, i1, 31);
                    if (WRITE) {
                        const uint32_t before = counters[view];
                        const long long base = a.csr_offsets[view] + (long long)row[view] + before;
                        long long pos = base + (i0 - c0);
#pragma unroll
                        for (int j = 0; j < 4; j++)
                            if (keep & (1u << j)) {
                                if (pos < a.capacity) a.out[pos] = ebase[0] + j + a.id_base;
                                pos++;
                            }
                        pos = base + t0 + (i1 - c1);
#pragma unroll
                        for (int j = 0; j < 4; j++)
                            if (keep & (16u << j)) {
                                if (pos < a.capacity) a.out[pos] = ebase[1] + j + a.id_base;
                                pos++;
                            }
                    }
                    __syncwarp();
                    if (lane == 0)
                        counters[view] += t0 + t1;
                    __syncwarp();
                }
            }
        }

        if (!WRITE) {
            // one coalesced row of counts per supertile, and sparse adds into the block totals
            uint32_t *out_row = a.counts + (int64_t)tile * a.n_views;
            uint32_t *totals = a.block_totals + (int64_t)(tile / kGatherTiles) * a.n_views;
            __syncwarp();
            for (int v = lane; v < a.n_views; v += 32) {
                const uint32_t c = counters[v];
                out_row[v] = c;
                if (c)
                    atomicAdd(totals + v, c);
            }
        }
        __syncwarp();
    }
}

// B1: one CTA, thread per view. block_totals[b][v] becomes the number of survivors of view v in blocks < b;
// csr_offsets[first_view + v + 1] receives the end of view v's list (csr_offsets[first_view] is the start of the pass).
__global__ void __launch_bounds__(1024)
multi_scan_blocks_kernel(uint32_t *block_totals, int num_blocks, int n_views, long long *csr_offsets) {
    __shared__ long long s_warp[32];
    const int v = threadIdx.x, lane = v & 31, warp = v >> 5;
    uint32_t running = 0;
    if (v < n_views) {
        for (int b = 0; b < num_blocks; b++) {
            uint32_t *p = block_totals + (int64_t)b * n_views + v;
            uint32_t c = *p;
            *p = running;
            running += c;
        }
    }
    // inclusive scan of the view totals across the CTA
    long long incl = running;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        long long up = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o)
            incl += up;
    }
    if (lane == 31)
        s_warp[warp] = incl;
    __syncthreads();
    long long before = 0;
    for (int w = 0; w < warp; w++)
        before += s_warp[w];
    const long long start = csr_offsets[0];
    __syncthreads();
    if (v < n_views)
        csr_offsets[v + 1] = start + before + incl;
}

// B2: one CTA per block of kGatherTiles supertiles, thread per view: counts[tile][v] becomes the number of survivors
// of view v in all earlier supertiles.
__global__ void __launch_bounds__(1024)
multi_scan_tiles_kernel(uint32_t *counts, const uint32_t *block_totals, int num_tiles, int n_views) {
    const int v = threadIdx.x;
    if (v >= n_views)
        return;
    const int b = blockIdx.x;
    uint32_t running = block_totals[(int64_t)b * n_views + v];
    const int t1 = min(num_tiles, (b + 1) * kGatherTiles);
    for (int t = b * kGatherTiles; t < t1; t++) {
        uint32_t *p = counts + (int64_t)t * n_views + v;
        uint32_t c = *p;
        *p = running;
        running += c;
    }
}

int multi_max_blocks_per_sm(int n_views) {
    int blocks = 0;
    size_t smem = sizeof(uint32_t) * kMultiWarps * (size_t)n_views;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks, multi_pass_kernel<false>, kMultiWarps * 32, smem);
    return blocks;
}

void launch_multi_pass(const MultiArgs &a, bool write, int grid, cudaStream_t s) {
    size_t smem = sizeof(uint32_t) * kMultiWarps * (size_t)a.n_views;
    if (write)
        multi_pass_kernel<true><<<grid, kMultiWarps * 32, smem, s>>>(a);
    else
        multi_pass_kernel<false><<<grid, kMultiWarps * 32, smem, s>>>(a);
}

void launch_multi_scans(const MultiArgs &a, long long *csr_offsets_of_pass, cudaStream_t s) {
    int num_blocks = (a.num_tiles + kGatherTiles - 1) / kGatherTiles;
    multi_scan_blocks_kernel<<<1, 1024, 0, s>>>(a.block_totals, num_blocks, a.n_views, csr_offsets_of_pass);
    multi_scan_tiles_kernel<<<num_blocks, 1024, 0, s>>>(a.counts, a.block_totals, a.num_tiles, a.n_views);
}

}  // namespace ovis
