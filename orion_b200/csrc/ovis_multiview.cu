// ovis_multiview.cu - K4: many views in one pass over the bounds, with K6's conservative pre-filter.
//
// Replaces the reference's "one full world cull per shadow view" (render/src/deferred_render_pipeline.cc:259-284:
// 1 + sum of shadow views full scans per frame) and serves the main camera + cascade views the same way.
//
// Views come in GROUPS: the views of one group share a pre-filter ball-cube (a point light's six cube faces share the
// cube of half-side `range` about the light); a group without a cube has an infinite radius. Per pass (at most
// kMultiMaxViews views):
//
//   A  multi_pass_kernel<false>  every warp walks "supertiles" of 1024 consecutive entities (8 x 128): loads the bounds
//                                 once, and for every group runs the cube pre-filter (6 compares per entity) and,
//                                 only where a candidate exists in the warp, the exact six-plane tests of its views with
//                                 warp-uniform early-out. It counts survivors per (supertile, view): one coalesced row of
//                                 counts per supertile plus sparse integer atomicAdds into per-(64-supertile block, view)
//                                 totals, and a bit per (supertile, group) that had any candidate.
//   B1 multi_scan_blocks_kernel  per view: exclusive scan of the block totals, view totals, CSR offsets of the pass.
//   B2 multi_scan_tiles_kernel   per block: exclusive scan over its 64 supertiles, in place (counts -> positions).
//   C  multi_pass_kernel<true>   the same walk again over the groups that had candidates, now writing each survivor id
//                                 at its final position.
//
// Positions are sums of counts, so every list is in slot (= insertion) order and identical from run to run. The exact
// tests are the same instructions as K3 (ovis_device.cuh: outside_plane), so lists are bit-identical to brute force;
// the pre-filter only ever skips (entity, group) pairs that every one of the group's views would reject
// (derivation in DESIGN.md section 4, K6).
#include "ovis_kernels.h"

#include "ovis_device.cuh"

namespace ovis {

constexpr int kSuperTile = 1024;                 // entities per supertile (one warp, 8 sub-tiles of 128)
constexpr int kSubTiles = 8;
constexpr int kMultiWarps = 8;                   // warps per CTA

// Shared memory per warp: n_views survivor counters + ceil(n_groups / 32) "this group had a candidate" words.
__host__ __device__ inline int multi_words_per_warp(int n_views, int n_groups) { return n_views + (n_groups + 31) / 32; }

template <bool WRITE, bool CUBES>
__global__ void __launch_bounds__(kMultiWarps * 32, 3)
multi_pass_kernel(const __grid_constant__ MultiArgs a) {
    extern __shared__ uint32_t s_words[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int hit_words = (a.n_groups + 31) / 32;
    uint32_t *counters = s_words + warp * multi_words_per_warp(a.n_views, a.n_groups);   // survivors so far, per view
    uint32_t *hits = counters + a.n_views;                                                // groups with a candidate
    const int warps_total = gridDim.x * kMultiWarps;
    const float4 *__restrict__ bounds = reinterpret_cast<const float4 *>(a.bounds);
    const int64_t vstride = a.stride >> 2;

    for (int tile = blockIdx.x * kMultiWarps + warp; tile < a.num_tiles; tile += warps_total) {
        // counters[v]: bits 0-15 survivors of view v so far in this supertile; bits 16-23 (pass A: collected, pass C:
        // loaded) which of the 8 sub-tiles hold a survivor of view v - pass C repeats the exact tests only there
        for (int v = lane; v < a.n_views; v += 32)
            counters[v] = WRITE ? (uint32_t)a.sub_masks[(int64_t)tile * a.n_views + v] << 16 : 0u;
        // pass A finds which groups have any candidate in this supertile; pass C only revisits those
        for (int k = lane; k < hit_words; k += 32)
            hits[k] = WRITE ? a.group_hits[(int64_t)tile * hit_words + k] : 0u;
        __syncwarp();
        const uint32_t *row = WRITE ? a.counts + (int64_t)tile * a.n_views : nullptr;   // positions after pass B2

        for (int sub = 0; sub < kSubTiles; sub++) {
            // ---- 4 consecutive entities per lane: six 128-bit loads
            const int64_t v = (int64_t)tile * (kSuperTile / 4) + sub * 32 + lane;        // float4 index
            const int64_t e0 = v << 2;
            const float4 ax = __ldg(bounds + 0 * vstride + v), ay = __ldg(bounds + 1 * vstride + v), az = __ldg(bounds + 2 * vstride + v);
            const float4 bx = __ldg(bounds + 3 * vstride + v), by = __ldg(bounds + 4 * vstride + v), bz = __ldg(bounds + 5 * vstride + v);
            const float mn[3][4] = { { ax.x, ax.y, ax.z, ax.w }, { ay.x, ay.y, ay.z, ay.w }, { az.x, az.y, az.z, az.w } };
            const float mx[3][4] = { { bx.x, bx.y, bx.z, bx.w }, { by.x, by.y, by.z, by.w }, { bz.x, bz.y, bz.z, bz.w } };
            uint32_t alive = 0xfu;
            if (e0 + 3 >= a.n) {
                int64_t left = a.n - e0;
                alive = left <= 0 ? 0u : ((1u << left) - 1u);
            }
            uint32_t caster = alive;
            if (a.alive_mask)
                alive &= (__ldg(a.alive_mask + (e0 >> 5)) >> (e0 & 31)) & 0xfu;
            if (a.caster_mask)
                caster &= (__ldg(a.caster_mask + (e0 >> 5)) >> (e0 & 31)) & 0xfu;          // already implies alive
            else
                caster = alive;
            if (!__any_sync(0xffffffffu, alive))
                continue;

            // ---- K6 pre-filter, entity side: the box inflated by Re = 1.001 * largest extent + 1e-3 * largest
            //      |coordinate|. A box with a NaN or non-finite side gets Re = +inf and can never be rejected.
            float lo[3][4], hi[3][4];
            bool lane_odd = false;
            constexpr bool cubes = CUBES;                        // plain views (camera, cascades, spot lights): nothing to pre-filter
#pragma unroll
            for (int j = 0; j < 4; j++) {
                if (!cubes)
                    break;
                const float sx = mx[0][j] - mn[0][j], sy = mx[1][j] - mn[1][j], sz = mx[2][j] - mn[2][j];
                const float ext = fmaxf(fmaxf(fmaxf(sx, sy), sz), 0.0f);
                const float mag = fmaxf(fmaxf(fmaxf(fabsf(mn[0][j]), fabsf(mx[0][j])), fmaxf(fabsf(mn[1][j]), fabsf(mx[1][j]))),
                                        fmaxf(fabsf(mn[2][j]), fabsf(mx[2][j])));
                float re = ext * 1.001f + mag * 1.0e-3f;
                const bool odd = sx != sx || sy != sy || sz != sz || !(ext <= 3.0e38f) || !(mag <= 3.0e38f);   // fmaxf drops NaNs
                if (odd)
                    re = __int_as_float(0x7f800000);
                lane_odd |= odd;
#pragma unroll
                for (int c = 0; c < 3; c++) {
                    lo[c][j] = mn[c][j] - re;
                    hi[c][j] = mx[c][j] + re;
                }
            }
            // ---- union of the warp's 128 inflated boxes: a group whose cube misses it has no candidate in this
            //      sub-tile (scenes laid out by their hierarchy keep neighbours in space next to each other in memory)
            float ulo[3], uhi[3];
#pragma unroll
            for (int c = 0; c < 3; c++) {
                if (!cubes)
                    break;
                ulo[c] = fminf(fminf(lo[c][0], lo[c][1]), fminf(lo[c][2], lo[c][3]));
                uhi[c] = fmaxf(fmaxf(hi[c][0], hi[c][1]), fmaxf(hi[c][2], hi[c][3]));
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    ulo[c] = fminf(ulo[c], __shfl_xor_sync(0xffffffffu, ulo[c], o));
                    uhi[c] = fmaxf(uhi[c], __shfl_xor_sync(0xffffffffu, uhi[c], o));
                }
            }
            const bool warp_odd = !cubes || __any_sync(0xffffffffu, lane_odd);   // fminf/fmaxf drop NaNs: then test every group

            for (int gw = 0; gw < hit_words; gw++) {
                // lane-parallel: lane L tests group gw*32 + L against the union box
                bool near = false;
                {
                    const int gl = gw * 32 + lane;
                    if (gl < a.n_groups) {
                        const ViewGroupDev gq = a.groups[gl];
                        near = warp_odd || !((uhi[0] < gq.lo[0]) | (ulo[0] > gq.hi[0]) | (uhi[1] < gq.lo[1]) |
                                             (ulo[1] > gq.hi[1]) | (uhi[2] < gq.lo[2]) | (ulo[2] > gq.hi[2]));
                    }
                }
                uint32_t todo = __ballot_sync(0xffffffffu, near);
                if (WRITE)
                    todo &= hits[gw];
                uint32_t found = 0;
                while (todo) {
                    const int bit = __ffs(todo) - 1;
                    todo &= todo - 1;
                    const int gi = gw * 32 + bit;
                    const ViewGroupDev grp = a.groups[gi];
                    // ---- group side: candidate iff the inflated box overlaps the group's (inflated) cube
                    uint32_t cand = cubes ? 0u : 0xfu;
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        if (!cubes)
                            break;
                        // separated along some axis <=> the largest of the six gaps is > 0. Subtractions run on the FMA
                        // pipe and the maxima fold three at a time (FMNMX3), which leaves ONE compare on the half-rate ALU
                        // pipe instead of six plus their combination; a NaN gap is ignored by fmaxf and an all-NaN set
                        // compares false, so odd boxes stay candidates exactly as with the six compares.
                        const float gap = fmaxf(fmaxf(fmaxf(grp.lo[0] - hi[0][j], lo[0][j] - grp.hi[0]),
                                                      fmaxf(grp.lo[1] - hi[1][j], lo[1][j] - grp.hi[1])),
                                                fmaxf(grp.lo[2] - hi[2][j], lo[2][j] - grp.hi[2]));
                        cand |= gap > 0.0f ? 0u : (1u << j);
                    }
                    cand &= alive;
                    if (!__any_sync(0xffffffffu, cand))
                        continue;
                    found |= 1u << bit;

                    for (int vi = 0; vi < grp.n_views; vi++) {
                        const int view = grp.first_view + vi;
                        if (WRITE && !((counters[view] >> (16 + sub)) & 1u))              // pass A found nobody here
                            continue;
                        uint32_t keep = a.view_require[view] ? (cand & caster) : cand;
                        const float4 *pl = a.planes + 6 * view;
                        const uint32_t codes = __ldg(a.view_codes + view);
                        // Two planes at a time: the P-vertex corner is picked by a warp-uniform switch on the plane's sign
                        // bits (no select per entity), products are packed FMUL2, and "either distance < 0" is one compare
                        // of the smaller distance (fminf ignores a NaN operand, as a NaN distance rejects nothing). The
                        // warp-uniform early-outs sit where they were: after the side planes and after top / bottom.
#pragma unroll
                        for (int pp = 0; pp < 3; pp++) {
                            if (pp > 0 && !__any_sync(0xffffffffu, keep))
                                break;
                            const float4 d0 = plane_dist4_coded(codes >> (6 * pp), __ldg(pl + 2 * pp), ax, ay, az, bx, by, bz);
                            const float4 d1 = plane_dist4_coded(codes >> (6 * pp + 3), __ldg(pl + 2 * pp + 1), ax, ay, az, bx, by, bz);
                            if (fminf(d0.x, d1.x) < 0.0f) keep &= ~1u;
                            if (fminf(d0.y, d1.y) < 0.0f) keep &= ~2u;
                            if (fminf(d0.z, d1.z) < 0.0f) keep &= ~4u;
                            if (fminf(d0.w, d1.w) < 0.0f) keep &= ~8u;
                        }
                        if (!__any_sync(0xffffffffu, keep))
                            continue;
                        // ---- rank in slot order inside the sub-tile
                        const uint32_t c = __popc(keep);
                        uint32_t incl = c;
#pragma unroll
                        for (int o = 1; o < 32; o <<= 1) {
                            uint32_t up = __shfl_up_sync(0xffffffffu, incl, o);
                            if (lane >= o)
                                incl += up;
                        }
                        const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
                        if (WRITE) {
                            long long pos = a.csr_offsets[view] + (long long)row[view] + (counters[view] & 0xffffu) + (incl - c);
#pragma unroll
                            for (int j = 0; j < 4; j++)
                                if (keep & (1u << j)) {
                                    if (pos < a.capacity) a.out[pos] = (uint32_t)e0 + j + a.id_base;
                                    pos++;
                                }
                        }
                        __syncwarp();
                        if (lane == 0)
                            counters[view] += WRITE ? total : (total | (0x10000u << sub));    // each (sub, view) comes by once
                        __syncwarp();
                    }
                }
                if (!WRITE && found) {
                    if (lane == 0)
                        hits[gw] |= found;
                    __syncwarp();
                }
            }
        }

        if (!WRITE) {
            // one coalesced row of counts per supertile, sparse adds into the block totals, and the group hits
            uint32_t *out_row = a.counts + (int64_t)tile * a.n_views;
            uint32_t *totals = a.block_totals + (int64_t)(tile / kGatherTiles) * a.n_views;
            __syncwarp();
            for (int v = lane; v < a.n_views; v += 32) {
                const uint32_t c = counters[v] & 0xffffu;
                out_row[v] = c;
                a.sub_masks[(int64_t)tile * a.n_views + v] = (uint8_t)(counters[v] >> 16);
                if (c)
                    atomicAdd(totals + v, c);
            }
            for (int k = lane; k < hit_words; k += 32)
                a.group_hits[(int64_t)tile * hit_words + k] = hits[k];
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
        for (int b0 = 0; b0 < num_blocks; b0 += 8) {            // 8 independent loads in flight, then the prefix
            uint32_t c[8];
#pragma unroll
            for (int k = 0; k < 8; k++)
                c[k] = (b0 + k < num_blocks) ? block_totals[(int64_t)(b0 + k) * n_views + v] : 0u;
#pragma unroll
            for (int k = 0; k < 8; k++) {
                if (b0 + k < num_blocks)
                    block_totals[(int64_t)(b0 + k) * n_views + v] = running;
                running += c[k];
            }
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
    for (int t0 = b * kGatherTiles; t0 < t1; t0 += 8) {
        uint32_t c[8];
#pragma unroll
        for (int k = 0; k < 8; k++)
            c[k] = (t0 + k < t1) ? counts[(int64_t)(t0 + k) * n_views + v] : 0u;
#pragma unroll
        for (int k = 0; k < 8; k++) {
            if (t0 + k < t1)
                counts[(int64_t)(t0 + k) * n_views + v] = running;
            running += c[k];
        }
    }
}

int multi_max_blocks_per_sm(int n_views, int n_groups) {
    int blocks = 0;
    size_t smem = sizeof(uint32_t) * kMultiWarps * (size_t)multi_words_per_warp(n_views, n_groups);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks, multi_pass_kernel<false, true>, kMultiWarps * 32, smem);
    return blocks;
}

void launch_multi_pass(const MultiArgs &a, bool write, int grid, cudaStream_t s) {
    size_t smem = sizeof(uint32_t) * kMultiWarps * (size_t)multi_words_per_warp(a.n_views, a.n_groups);
    if (a.any_cube) {
        if (write)
            multi_pass_kernel<true, true><<<grid, kMultiWarps * 32, smem, s>>>(a);
        else
            multi_pass_kernel<false, true><<<grid, kMultiWarps * 32, smem, s>>>(a);
    } else {
        if (write)
            multi_pass_kernel<true, false><<<grid, kMultiWarps * 32, smem, s>>>(a);
        else
            multi_pass_kernel<false, false><<<grid, kMultiWarps * 32, smem, s>>>(a);
    }
}

void launch_multi_scans(const MultiArgs &a, long long *csr_offsets_of_pass, cudaStream_t s) {
    int num_blocks = (a.num_tiles + kGatherTiles - 1) / kGatherTiles;
    multi_scan_blocks_kernel<<<1, 1024, 0, s>>>(a.block_totals, num_blocks, a.n_views, csr_offsets_of_pass);
    multi_scan_tiles_kernel<<<num_blocks, 1024, 0, s>>>(a.counts, a.block_totals, a.num_tiles, a.n_views);
}

}  // namespace ovis
