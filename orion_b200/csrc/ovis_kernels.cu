// ovis_kernels.cu - sm_100a kernels of the scene-visibility path and their launchers.
//
// K1+K2  update_level_kernel     Entity::transformed + Transform::matrix + BoundingBox::transform
// K3     cull_frustum_kernel     SimpleRenderWorld::cull entity loop, one view, stable compaction
// K5     cull_lights_kernel      RenderLight::cull
// plus layout movers (AoS host layout <-> SoA device layout).
//
// The path is HBM-bound byte/flop-light work: no tensor cores, no GEMM shaping. Bounds live in HBM as six
// SoA float arrays; a thread owns four consecutive entities and reads them with six 128-bit loads, so a
// warp covers 128 consecutive entities with fully coalesced 512 B requests per array.
#include "ovis_kernels.h"

#include <cstdlib>

#include <cooperative_groups.h>

#include "ovis_device.cuh"

namespace ovis {

// ------------------------------------------------------------------------------------------------
// Layout movers
// ------------------------------------------------------------------------------------------------

template <int K>
__global__ void scatter_aos_kernel(const float *__restrict__ src, const uint32_t *__restrict__ ids, uint32_t first,
                                   int64_t n, float *dst, int64_t stride) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
        return;
    int64_t slot = ids ? (int64_t)ids[i] : (int64_t)first + i;
#pragma unroll
    for (int k = 0; k < K; k++)
        dst[k * stride + slot] = src[i * K + k];
}

template <int K>
__global__ void gather_aos_kernel(const float *__restrict__ src, int64_t stride, uint32_t first, int64_t n, float *dst) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
        return;
#pragma unroll
    for (int k = 0; k < K; k++)
        dst[i * K + k] = src[k * stride + first + i];
}

__global__ void scatter_u32_kernel(const uint32_t *__restrict__ src, const uint32_t *__restrict__ ids, int64_t n,
                                   uint32_t *dst) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n)
        dst[ids[i]] = src[i];
}

__global__ void fill_u32_kernel(uint32_t *dst, int64_t n, uint32_t value) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n)
        dst[i] = value;
}

static inline unsigned blocks_for(int64_t n, int threads) { return (unsigned)((n + threads - 1) / threads); }

void launch_scatter_aos(int k, const float *src, const uint32_t *ids, uint32_t first, int64_t n, float *dst,
                        int64_t stride, cudaStream_t s) {
    if (n <= 0)
        return;
    unsigned g = blocks_for(n, 256);
    switch (k) {
    case 6: scatter_aos_kernel<6><<<g, 256, 0, s>>>(src, ids, first, n, dst, stride); break;
    case 10: scatter_aos_kernel<10><<<g, 256, 0, s>>>(src, ids, first, n, dst, stride); break;
    }
}

void launch_gather_aos(int k, const float *src, int64_t stride, uint32_t first, int64_t n, float *dst, cudaStream_t s) {
    if (n <= 0)
        return;
    unsigned g = blocks_for(n, 256);
    switch (k) {
    case 6: gather_aos_kernel<6><<<g, 256, 0, s>>>(src, stride, first, n, dst); break;
    case 10: gather_aos_kernel<10><<<g, 256, 0, s>>>(src, stride, first, n, dst); break;
    }
}

void launch_scatter_u32(const uint32_t *src, const uint32_t *ids, int64_t n, uint32_t *dst, cudaStream_t s) {
    if (n > 0)
        scatter_u32_kernel<<<blocks_for(n, 256), 256, 0, s>>>(src, ids, n, dst);
}

void launch_fill_u32(uint32_t *dst, int64_t n, uint32_t value, cudaStream_t s) {
    if (n > 0)
        fill_u32_kernel<<<blocks_for(n, 256), 256, 0, s>>>(dst, n, value);
}

// ------------------------------------------------------------------------------------------------
// K2 alone: world AABB of entities that have no node (identity transform), used when entities are added
// or their local box changes. world = local.transform(Transform().matrix()), render_entity.cc:90-93.
// ------------------------------------------------------------------------------------------------

__device__ __forceinline__ DTrs identity_trs() { return { { 0.0f, 0.0f, 0.0f }, { 1.0f, 0.0f, 0.0f, 0.0f }, { 1.0f, 1.0f, 1.0f } }; }

__global__ void identity_bounds_kernel(const uint32_t *__restrict__ ids, uint32_t first, int64_t n,
                                       const float *__restrict__ local, float *world, int64_t stride) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
        return;
    int64_t slot = ids ? (int64_t)ids[i] : (int64_t)first + i;
    DAabb b;
#pragma unroll
    for (int k = 0; k < 3; k++) {
        b.mn[k] = local[k * stride + slot];
        b.mx[k] = local[(k + 3) * stride + slot];
    }
    DAabb w = transform_aabb(b, trs_matrix(identity_trs()));
#pragma unroll
    for (int k = 0; k < 3; k++) {
        world[k * stride + slot] = w.mn[k];
        world[(k + 3) * stride + slot] = w.mx[k];
    }
}

void launch_identity_bounds(const uint32_t *ids, uint32_t first, int64_t n, const float *local, float *world,
                            int64_t stride, cudaStream_t s) {
    if (n > 0)
        identity_bounds_kernel<<<blocks_for(n, 256), 256, 0, s>>>(ids, first, n, local, world, stride);
}

// ------------------------------------------------------------------------------------------------
// K1 + K2 fused, one launch per tree level (the TRS composition is not associative in fp32, so levels
// cannot be merged into a prefix scan). Per node: read local TRS (40 B) + parent index (4 B) + bound
// entity (4 B), gather the parent's world TRS (siblings share it: cache hits), write world TRS (40 B);
// if an entity is bound: read its local AABB (24 B), write its world AABB (24 B).
// ------------------------------------------------------------------------------------------------

__device__ __forceinline__ DTrs load_trs(const float *base, int64_t stride, int64_t i) {
    DTrs t;
    t.p = { base[0 * stride + i], base[1 * stride + i], base[2 * stride + i] };
    t.q = { base[3 * stride + i], base[4 * stride + i], base[5 * stride + i], base[6 * stride + i] };
    t.s = { base[7 * stride + i], base[8 * stride + i], base[9 * stride + i] };
    return t;
}

__device__ __forceinline__ void store_trs(float *base, int64_t stride, int64_t i, const DTrs &t) {
    base[0 * stride + i] = t.p.x; base[1 * stride + i] = t.p.y; base[2 * stride + i] = t.p.z;
    base[3 * stride + i] = t.q.w; base[4 * stride + i] = t.q.x; base[5 * stride + i] = t.q.y; base[6 * stride + i] = t.q.z;
    base[7 * stride + i] = t.s.x; base[8 * stride + i] = t.s.y; base[9 * stride + i] = t.s.z;
}

__device__ __forceinline__ void entity_bounds_from_trs(const DTrs &w, uint32_t entity, const float *local, float *world,
                                                       int64_t estride) {
    DAabb b;
#pragma unroll
    for (int k = 0; k < 3; k++) {
        b.mn[k] = local[k * estride + entity];
        b.mx[k] = local[(k + 3) * estride + entity];
    }
    DAabb o = transform_aabb(b, trs_matrix(w));
#pragma unroll
    for (int k = 0; k < 3; k++) {
        world[k * estride + entity] = o.mn[k];
        world[(k + 3) * estride + entity] = o.mx[k];
    }
}

__global__ void __launch_bounds__(256)
update_level_kernel(LevelArgs a) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.count)
        return;
    int64_t node = a.list ? (int64_t)a.list[i] : a.first + i;
    int32_t parent = a.parent[node];
    if (a.dirty) {
        // Entity::transformed only runs below a node that was moved (entity.cc:309-363 -> :382-411): a node is
        // recomputed if it changed itself or its parent was recomputed (parents are in an earlier launch).
        bool moved = a.dirty[node] != 0 || (parent >= 0 && a.dirty[parent] != 0);
        if (!moved)
            return;
        a.dirty[node] = 1;
    }
    DTrs w = load_trs(a.local_trs, a.nstride, node);
    if (parent >= 0)
        w = compose_trs(load_trs(a.world_trs, a.nstride, parent), w);
    store_trs(a.world_trs, a.nstride, node, w);
    uint32_t entity = a.node_entity[node];
    if (entity != kNone)
        entity_bounds_from_trs(w, entity, a.local_aabb, a.world_aabb, a.estride);
}

void launch_update_level(const LevelArgs &a, cudaStream_t s) {
    if (a.count > 0)
        update_level_kernel<<<blocks_for(a.count, 256), 256, 0, s>>>(a);
}

// Entities beyond the first bound to a node (a MeshRenderer owns one RenderEntity per sub-mesh,
// graphics/src/mesh_renderer.cc:180-187): (entity, node) pairs, run after all levels.
__global__ void extra_bounds_kernel(const uint32_t *__restrict__ entities, const uint32_t *__restrict__ nodes, int64_t n,
                                    const float *__restrict__ world_trs, int64_t nstride, const float *__restrict__ local,
                                    float *world, int64_t estride, const uint8_t *__restrict__ dirty) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
        return;
    if (dirty && !dirty[nodes[i]])
        return;
    entity_bounds_from_trs(load_trs(world_trs, nstride, nodes[i]), entities[i], local, world, estride);
}

void launch_extra_bounds(const uint32_t *entities, const uint32_t *nodes, int64_t n, const float *world_trs,
                         int64_t nstride, const float *local, float *world, int64_t estride, const uint8_t *dirty,
                         cudaStream_t s) {
    if (n > 0)
        extra_bounds_kernel<<<blocks_for(n, 256), 256, 0, s>>>(entities, nodes, n, world_trs, nstride, local, world, estride, dirty);
}

__global__ void mark_dirty_kernel(const uint32_t *__restrict__ ids, int64_t n, uint8_t *dirty) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n)
        dirty[ids[i]] = 1;
}

void launch_mark_dirty(const uint32_t *ids, int64_t n, uint8_t *dirty, cudaStream_t s) {
    if (n > 0)
        mark_dirty_kernel<<<blocks_for(n, 256), 256, 0, s>>>(ids, n, dirty);
}

// ------------------------------------------------------------------------------------------------
// K3: one view over all entity slots with ORDER-PRESERVING compaction, in two launches.
//
//   K3a cull_test_kernel     pure streaming, no barrier and no inter-CTA dependency: every WARP owns
//                            "warp tiles" of 128 x GROUPS consecutive entities, tests 4 x GROUPS entities
//                            per lane (six 128-bit loads per group, the next tile's loads issued before the
//                            current tile's ranking), ranks survivors with shuffle scans and stores their ids
//                            at the start of the tile's fixed slot in a staging array, plus one count per tile
//                            and (integer atomicAdd, order-independent) one total per block of 64 tiles.
//   K3b cull_gather_kernel   one CTA per block of 64 tiles: base = sum of the earlier block totals, exclusive
//                            scan of its 64 counts, then moves the staged ids to their final, dense position.
//
// Output order = slot order = insertion order and is identical from run to run: positions come from sums
// of counts only. An earlier single-launch version (decoupled look-back over tile status words) was limited
// to ~2.1 TB/s by CTA-wide barrier stalls behind the look-back chain; see profiles/r01_cull_notes.md.
// ------------------------------------------------------------------------------------------------

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

// The six plane tests of one warp tile, specialised: see outside-plane notes in ovis_device.cuh (plane_dist4_signed).
template <int GROUPS>
__device__ __forceinline__ void test_planes_packed(const CullArgs &a, const float4 (&mnx)[GROUPS], const float4 (&mny)[GROUPS],
                                                   const float4 (&mnz)[GROUPS], const float4 (&mxx)[GROUPS], const float4 (&mxy)[GROUPS],
                                                   const float4 (&mxz)[GROUPS], uint32_t (&keep)[GROUPS]) {
    // one warp-uniform switch per plane with all GROUPS x 4 entities of the lane inside it (independent chains to
    // overlap); distances are folded into a running minimum, two planes per FMNMX3, one compare per entity at the end
    const float inf = __int_as_float(0x7f800000);
    float4 lowest[GROUPS], held[GROUPS];
#pragma unroll
    for (int g = 0; g < GROUPS; g++)
        lowest[g] = make_float4(inf, inf, inf, inf);
#pragma unroll
    for (int p = 0; p < 6; p++) {
        const float nx = a.planes[p].x, ny = a.planes[p].y, nz = a.planes[p].z, d = a.planes[p].w;
        const int code = (nx >= 0.0f ? 1 : 0) | (ny >= 0.0f ? 2 : 0) | (nz >= 0.0f ? 4 : 0);
#define OVIS_PLANE_CASE(C, SX, SY, SZ)                                                                                         \
    case C:                                                                                                                    \
_Pragma("unroll") for (int g = 0; g < GROUPS; g++) {                                                                   \
    const float4 dd = plane_dist4_signed<SX, SY, SZ>(nx, ny, nz, d, mnx[g], mny[g], mnz[g], mxx[g], mxy[g], mxz[g]);   \
    if (p & 1)                                                                                                         \
        lowest[g] = min3_4(lowest[g], held[g], dd);                                                                    \
    else                                                                                                               \
        held[g] = dd;                                                                                                  \
}                                                                                                                      \
break;
        switch (code) {
            OVIS_PLANE_CASE(0, false, false, false)
            OVIS_PLANE_CASE(1, true, false, false)
            OVIS_PLANE_CASE(2, false, true, false)
            OVIS_PLANE_CASE(3, true, true, false)
            OVIS_PLANE_CASE(4, false, false, true)
            OVIS_PLANE_CASE(5, true, false, true)
            OVIS_PLANE_CASE(6, false, true, true)
            OVIS_PLANE_CASE(7, true, true, true)
        }
#undef OVIS_PLANE_CASE
    }
#pragma unroll
    for (int g = 0; g < GROUPS; g++) {
        uint32_t out = 0;
        if (lowest[g].x < 0.0f) out |= 1u;
        if (lowest[g].y < 0.0f) out |= 2u;
        if (lowest[g].z < 0.0f) out |= 4u;
        if (lowest[g].w < 0.0f) out |= 8u;
        keep[g] &= ~out;
    }
}

template <int GROUPS, bool MASKED, bool PACKED = false>
__device__ __forceinline__ void cull_test_body(const CullArgs &a) {
    constexpr int WT = 128 * GROUPS;                       // entities per warp tile
    const int lane = threadIdx.x & 31;
    const int warps_total = gridDim.x * (kCullThreads / 32);
    const int gwarp = blockIdx.x * (kCullThreads / 32) + (threadIdx.x >> 5);
    const float4 *__restrict__ bounds = reinterpret_cast<const float4 *>(a.bounds);
    const int64_t vstride = a.stride >> 2;                 // component stride in float4 units

    // re-arm the block totals the NEXT launch will accumulate into (this launch uses the other set)
    for (int i = blockIdx.x * kCullThreads + threadIdx.x; i < a.num_blocks; i += gridDim.x * kCullThreads)
        a.block_totals_next[i] = 0u;

    float4 mnx[GROUPS], mny[GROUPS], mnz[GROUPS], mxx[GROUPS], mxy[GROUPS], mxz[GROUPS];
    uint32_t mask_bits[GROUPS];
    auto issue_loads = [&](int wt) {
        if (wt >= a.num_tiles)
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
    for (; wt < a.num_tiles; wt += warps_total) {
        uint32_t keep[GROUPS];
#pragma unroll
        for (int g = 0; g < GROUPS; g++) {
            const int64_t e0 = ((int64_t)wt * (32 * GROUPS) + g * 32 + lane) << 2;   // entities e0 .. e0+3
            uint32_t k = 0xfu;                                                // bit j: entity e0 + j still inside
            if (MASKED)
                k = (mask_bits[g] >> (e0 & 31)) & 0xfu;
            if (e0 + 3 >= a.n) {                                              // ragged tail of the last tile
                int64_t valid = a.n - e0;
                k &= valid <= 0 ? 0u : ((1u << valid) - 1u);
            }
#pragma unroll
            for (int p = 0; p < 6; p++) {                                     // Left Right Top Bottom Near Far
                const float nx = a.planes[p].x, ny = a.planes[p].y, nz = a.planes[p].z, d = a.planes[p].w;
                if (PACKED)
                    break;                                                    // handled below, plane by plane for all groups
                if (outside_plane(nx, ny, nz, d, mnx[g].x, mny[g].x, mnz[g].x, mxx[g].x, mxy[g].x, mxz[g].x)) k &= ~1u;
                if (outside_plane(nx, ny, nz, d, mnx[g].y, mny[g].y, mnz[g].y, mxx[g].y, mxy[g].y, mxz[g].y)) k &= ~2u;
                if (outside_plane(nx, ny, nz, d, mnx[g].z, mny[g].z, mnz[g].z, mxx[g].z, mxy[g].z, mxz[g].z)) k &= ~4u;
                if (outside_plane(nx, ny, nz, d, mnx[g].w, mny[g].w, mnz[g].w, mxx[g].w, mxy[g].w, mxz[g].w)) k &= ~8u;
            }
            keep[g] = k;
        }
        if (PACKED)
            test_planes_packed<GROUPS>(a, mnx, mny, mnz, mxx, mxy, mxz, keep);
        issue_loads(wt + warps_total);                                        // this tile's registers are dead now

        // rank inside the warp tile: per group an inclusive shuffle scan, groups in index order
        uint32_t base = 0;
        uint32_t *slot = a.staging + (int64_t)wt * WT;
#pragma unroll
        for (int g = 0; g < GROUPS; g++) {
            uint32_t sum = __popc(keep[g]);
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                uint32_t up = __shfl_up_sync(0xffffffffu, sum, o);
                if (lane >= o)
                    sum += up;
            }
            if (keep[g]) {
                uint32_t rank = base + sum - __popc(keep[g]);
                const uint32_t e0 = (uint32_t)(((int64_t)wt * (32 * GROUPS) + g * 32 + lane) << 2) + a.id_base;
#pragma unroll
                for (int j = 0; j < 4; j++)
                    if (keep[g] & (1u << j))
                        slot[rank++] = e0 + j;
            }
            base += __shfl_sync(0xffffffffu, sum, 31);
        }
        if (lane == 0) {
            a.tile_counts[wt] = base;
            if (base)
                atomicAdd(a.block_totals + wt / kGatherTiles, base);
        }
    }
}

// ---- software-pipelined form of the same pass ----------------------------------------------------------------------
// In cull_test_body a warp has loads in flight only while it ranks the previous tile (its registers are reused), i.e. about
// a quarter of the time; tools/read_probe.cu shows that the same bytes stream in 45 us when every warp always has loads
// outstanding. Here a warp owns TWO register sets and requests tile t + 1 before it tests tile t.
template <int GROUPS>
struct TileRegs {
    float4 mnx[GROUPS], mny[GROUPS], mnz[GROUPS], mxx[GROUPS], mxy[GROUPS], mxz[GROUPS];
    uint32_t mask_bits[GROUPS];
};

template <int GROUPS, bool MASKED>
__device__ __forceinline__ void tile_load(TileRegs<GROUPS> &r, const CullArgs &a, const float4 *__restrict__ bounds, int64_t vstride,
                                          int wt, int lane) {
    if (wt >= a.num_tiles)
        return;
#pragma unroll
    for (int g = 0; g < GROUPS; g++) {
        const int64_t v = (int64_t)wt * (32 * GROUPS) + g * 32 + lane;
        r.mnx[g] = __ldg(bounds + 0 * vstride + v); r.mny[g] = __ldg(bounds + 1 * vstride + v); r.mnz[g] = __ldg(bounds + 2 * vstride + v);
        r.mxx[g] = __ldg(bounds + 3 * vstride + v); r.mxy[g] = __ldg(bounds + 4 * vstride + v); r.mxz[g] = __ldg(bounds + 5 * vstride + v);
        if (MASKED)
            r.mask_bits[g] = __ldg(a.mask + (v >> 3));
    }
}

template <int GROUPS, bool MASKED, bool PACKED = false>
__device__ __forceinline__ void tile_process(const TileRegs<GROUPS> &r, const CullArgs &a, int wt, int lane) {
    constexpr int WT = 128 * GROUPS;
    uint32_t keep[GROUPS];
#pragma unroll
    for (int g = 0; g < GROUPS; g++) {
        const int64_t e0 = ((int64_t)wt * (32 * GROUPS) + g * 32 + lane) << 2;
        uint32_t k = 0xfu;
        if (MASKED)
            k = (r.mask_bits[g] >> (e0 & 31)) & 0xfu;
        if (e0 + 3 >= a.n) {
            int64_t valid = a.n - e0;
            k &= valid <= 0 ? 0u : ((1u << valid) - 1u);
        }
#pragma unroll
        for (int p = 0; p < 6; p++) {
            if (PACKED)
                break;
            const float nx = a.planes[p].x, ny = a.planes[p].y, nz = a.planes[p].z, d = a.planes[p].w;
            if (outside_plane(nx, ny, nz, d, r.mnx[g].x, r.mny[g].x, r.mnz[g].x, r.mxx[g].x, r.mxy[g].x, r.mxz[g].x)) k &= ~1u;
            if (outside_plane(nx, ny, nz, d, r.mnx[g].y, r.mny[g].y, r.mnz[g].y, r.mxx[g].y, r.mxy[g].y, r.mxz[g].y)) k &= ~2u;
            if (outside_plane(nx, ny, nz, d, r.mnx[g].z, r.mny[g].z, r.mnz[g].z, r.mxx[g].z, r.mxy[g].z, r.mxz[g].z)) k &= ~4u;
            if (outside_plane(nx, ny, nz, d, r.mnx[g].w, r.mny[g].w, r.mnz[g].w, r.mxx[g].w, r.mxy[g].w, r.mxz[g].w)) k &= ~8u;
        }
        keep[g] = k;
    }
    if (PACKED)
        test_planes_packed<GROUPS>(a, r.mnx, r.mny, r.mnz, r.mxx, r.mxy, r.mxz, keep);
    uint32_t base = 0;
    uint32_t *slot = a.staging + (int64_t)wt * WT;
#pragma unroll
    for (int g = 0; g < GROUPS; g++) {
        uint32_t sum = __popc(keep[g]);
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t up = __shfl_up_sync(0xffffffffu, sum, o);
            if (lane >= o)
                sum += up;
        }
        if (keep[g]) {
            uint32_t rank = base + sum - __popc(keep[g]);
            const uint32_t e0 = (uint32_t)(((int64_t)wt * (32 * GROUPS) + g * 32 + lane) << 2) + a.id_base;
#pragma unroll
            for (int j = 0; j < 4; j++)
                if (keep[g] & (1u << j))
                    slot[rank++] = e0 + j;
        }
        base += __shfl_sync(0xffffffffu, sum, 31);
    }
    if (lane == 0) {
        a.tile_counts[wt] = base;
        if (base)
            atomicAdd(a.block_totals + wt / kGatherTiles, base);
    }
}

template <int GROUPS, bool MASKED, bool PACKED = false>
__global__ void __launch_bounds__(kCullThreads, GROUPS == 1 ? 2 : 1)
cull_test_pipelined_kernel(const __grid_constant__ CullArgs a) {
    if (a.n_wait && blockIdx.x == 0 && threadIdx.x < 32)
        wait_flags_warp(a.wait_flags, a.n_wait, a.wait_value, a.error_flag);
    const int lane = threadIdx.x & 31;
    const int warps_total = gridDim.x * (kCullThreads / 32);
    const float4 *__restrict__ bounds = reinterpret_cast<const float4 *>(a.bounds);
    const int64_t vstride = a.stride >> 2;
    for (int i = blockIdx.x * kCullThreads + threadIdx.x; i < a.num_blocks; i += gridDim.x * kCullThreads)
        a.block_totals_next[i] = 0u;
    TileRegs<GROUPS> ra, rb;
    int wt = blockIdx.x * (kCullThreads / 32) + (threadIdx.x >> 5);
    tile_load<GROUPS, MASKED>(ra, a, bounds, vstride, wt, lane);
    while (wt < a.num_tiles) {
        tile_load<GROUPS, MASKED>(rb, a, bounds, vstride, wt + warps_total, lane);
        tile_process<GROUPS, MASKED, PACKED>(ra, a, wt, lane);
        wt += warps_total;
        if (wt >= a.num_tiles)
            break;
        tile_load<GROUPS, MASKED>(ra, a, bounds, vstride, wt + warps_total, lane);
        tile_process<GROUPS, MASKED, PACKED>(rb, a, wt, lane);
        wt += warps_total;
    }
}

// ---- asynchronous-copy form: the bounds stream into a per-warp shared-memory ring --------------------------------------
// Every lane copies the 16 bytes it will itself test (cp.async, global -> shared, no registers held while in flight) STAGES
// - 1 tiles ahead, so each warp keeps (STAGES - 1) x 3 KB outstanding all the time however long its arithmetic takes; a
// lane only ever reads back what it copied itself, so there is no barrier and no cross-lane visibility to arrange.
template <int STAGES>
struct AsyncRing {
    static constexpr int kStageBytes = 6 * 32 * 16 + 32 * 4;          // six float4 rows + one mask word per lane
    static constexpr int kWarpBytes = STAGES * kStageBytes;
    static constexpr int kCtaBytes = (kCullThreads / 32) * kWarpBytes;
};

__device__ __forceinline__ void cp_async_16(void *smem_dst, const void *gmem_src) {
    const uint32_t dst = (uint32_t)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_4(void *smem_dst, const void *gmem_src) {
    const uint32_t dst = (uint32_t)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <bool MASKED, int STAGES, bool PACKED = false>
__global__ void __launch_bounds__(kCullThreads, STAGES == 2 ? 4 : 2)
cull_test_async_kernel(const __grid_constant__ CullArgs a) {
    extern __shared__ __align__(16) unsigned char s_ring_bytes[];
    if (a.n_wait && blockIdx.x == 0 && threadIdx.x < 32)
        wait_flags_warp(a.wait_flags, a.n_wait, a.wait_value, a.error_flag);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int warps_total = gridDim.x * (kCullThreads / 32);
    const float4 *__restrict__ bounds = reinterpret_cast<const float4 *>(a.bounds);
    const int64_t vstride = a.stride >> 2;
    for (int i = blockIdx.x * kCullThreads + threadIdx.x; i < a.num_blocks; i += gridDim.x * kCullThreads)
        a.block_totals_next[i] = 0u;
    unsigned char *ring = s_ring_bytes + warp * AsyncRing<STAGES>::kWarpBytes;

    auto issue = [&](int wt, int stage) {
        if (wt < a.num_tiles) {
            float4 *rows = reinterpret_cast<float4 *>(ring + stage * AsyncRing<STAGES>::kStageBytes);
            const int64_t v = (int64_t)wt * 32 + lane;                         // float4 index: entities 4v .. 4v+3
#pragma unroll
            for (int c = 0; c < 6; c++)
                cp_async_16(rows + c * 32 + lane, bounds + c * vstride + v);
            if (MASKED)
                cp_async_4(reinterpret_cast<uint32_t *>(rows + 6 * 32) + lane, a.mask + (v >> 3));
        }
        cp_async_commit();                                                     // an empty group keeps the count uniform
    };

    int wt = blockIdx.x * (kCullThreads / 32) + warp;
#pragma unroll
    for (int st = 0; st < STAGES - 1; st++)
        issue(wt + st * warps_total, st);
    int stage = 0;
    for (; wt < a.num_tiles; wt += warps_total) {
        int ahead = stage + STAGES - 1;
        if (ahead >= STAGES)
            ahead -= STAGES;
        issue(wt + (STAGES - 1) * warps_total, ahead);
        cp_async_wait<STAGES - 1>();                                           // this lane's copies of tile wt have landed
        const float4 *rows = reinterpret_cast<const float4 *>(ring + stage * AsyncRing<STAGES>::kStageBytes);
        TileRegs<1> r;
        r.mnx[0] = rows[0 * 32 + lane]; r.mny[0] = rows[1 * 32 + lane]; r.mnz[0] = rows[2 * 32 + lane];
        r.mxx[0] = rows[3 * 32 + lane]; r.mxy[0] = rows[4 * 32 + lane]; r.mxz[0] = rows[5 * 32 + lane];
        if (MASKED)
            r.mask_bits[0] = reinterpret_cast<const uint32_t *>(rows + 6 * 32)[lane];
        tile_process<1, MASKED, PACKED>(r, a, wt, lane);
        stage = stage + 1 == STAGES ? 0 : stage + 1;
    }
    cp_async_wait<0>();
}

template <int GROUPS, bool MASKED, bool PACKED = false>
__global__ void __launch_bounds__(kCullThreads, 4 / GROUPS)
cull_test_kernel(const __grid_constant__ CullArgs a) {
    // Sharded exchange: the segment buffer this frame's gather kernel will refill must have been released by every peer.
    // One warp of the whole grid checks that here, hidden behind the streaming pass (the gather kernel starts only after
    // this kernel has finished), instead of every gather block paying a round of system-scope loads before its stores.
    if (a.n_wait && blockIdx.x == 0 && threadIdx.x < 32)
        wait_flags_warp(a.wait_flags, a.n_wait, a.wait_value, a.error_flag);
    cull_test_body<GROUPS, MASKED, PACKED>(a);
}

// Gather of one block of kGatherTiles warp tiles by one 256-thread CTA; `num_gather_blocks` = blocks in the launch.
__device__ __forceinline__ void cull_gather_body(const GatherArgs &a, int block, int num_gather_blocks) {
    __shared__ uint32_t s_scan[kGatherTiles + 1];   // exclusive scan of this block's tile counts, then the block sum
    __shared__ uint32_t s_warp[8];
    __shared__ uint32_t s_base;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    // A block without survivors has nothing to place (most blocks of a large world under a narrow frustum); only the last
    // block must go on, it publishes the grand total.
    if (block != num_gather_blocks - 1 && a.block_totals[block] == 0u)
        return;

    // base = survivors in all earlier blocks
    uint32_t part = 0;
    for (int i = tid; i < block; i += 256)
        part += a.block_totals[i];
    part = __reduce_add_sync(0xffffffffu, part);
    if (lane == 0)
        s_warp[warp] = part;

    // exclusive scan of the block's tile counts (kGatherTiles = 64: two warps)
    uint32_t c = 0, incl = 0;
    if (tid < kGatherTiles) {
        const int tile = block * kGatherTiles + tid;
        c = tile < a.num_tiles ? a.tile_counts[tile] : 0u;
        incl = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t up = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o)
                incl += up;
        }
        if (tid == 31)
            s_scan[kGatherTiles] = incl;                    // total of the first warp, fixed up below
    }
    __syncthreads();
    if (tid < kGatherTiles) {
        const uint32_t first = s_scan[kGatherTiles];
        s_scan[tid] = (warp ? first : 0u) + incl - c;
    }
    if (tid == 0) {
        uint32_t b = 0;
#pragma unroll
        for (int w = 0; w < 8; w++)
            b += s_warp[w];
        s_base = b;
    }
    __syncthreads();
    if (tid == kGatherTiles - 1)
        s_scan[kGatherTiles] = s_scan[tid] + c;             // block sum
    __syncthreads();

    const uint32_t total = s_scan[kGatherTiles];
    const long long out_base = a.d_base ? *a.d_base : 0ll;
    const long long dst0 = out_base + (long long)s_base;
    if (block == num_gather_blocks - 1 && tid == 0) {       // the last block knows the grand total
        uint32_t grand = s_base + total;
        if (a.d_count) *a.d_count = grand;
        if (a.d_next) *a.d_next = out_base + (long long)grand;
        for (int x = 0; x < a.n_extra; x++)
            *a.extra_count[x] = grand;
    }
    // output position k of this block comes from the tile t with s_scan[t] <= k < s_scan[t+1];
    // four independent positions per pass keep several staging loads in flight per thread
    const int64_t tile0 = (int64_t)block * kGatherTiles;
    for (uint32_t k0 = tid; k0 < total; k0 += 4 * 256) {
        uint32_t val[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const uint32_t k = k0 + u * 256;
            if (k < total) {
                int lo = 0;
#pragma unroll
                for (int step = kGatherTiles / 2; step > 0; step >>= 1)      // last t with s_scan[t] <= k
                    if (s_scan[lo + step] <= k)
                        lo += step;
                val[u] = a.staging[(tile0 + lo) * a.tile_entities + (k - s_scan[lo])];
            }
        }
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const uint32_t k = k0 + u * 256;
            if (k < total) {
                if (dst0 + k < a.capacity)
                    a.out[dst0 + k] = val[u];
                if (dst0 + k < a.extra_capacity)
                    for (int x = 0; x < a.n_extra; x++)        // peers' copies of this shard's segment / the host's copy
                        a.extra_out[x][dst0 + k] = val[u];
            }
        }
    }
}

__global__ void __launch_bounds__(256)
cull_gather_kernel(const __grid_constant__ GatherArgs a) {
    // Let the next cull's test kernel start now (if it was launched to allow that): it shares nothing with this kernel.
    asm volatile("griddepcontrol.launch_dependents;");
    if (a.n_wait) {                                           // peers must be done with the buffer we are about to refill
        if (threadIdx.x == 0)
            wait_flags_at_least(a.wait_flags, a.n_wait, a.wait_value, a.error_flag);
        __syncthreads();
    }
    // a CTA takes every gridDim.x-th block of 64 tiles (blocks without survivors return at once inside the body)
    const int num_gather_blocks = (a.num_tiles + kGatherTiles - 1) / kGatherTiles;
    for (int b = blockIdx.x; b < num_gather_blocks; b += gridDim.x) {
        cull_gather_body(a, b, num_gather_blocks);
        __syncthreads();                                      // the body's shared arrays are reused by the next block
    }
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
    cull_test_body<GROUPS, MASKED>(a);
    cooperative_groups::this_grid().sync();
    const int num_gather_blocks = (g.num_tiles + kGatherTiles - 1) / kGatherTiles;
    for (int b = blockIdx.x; b < num_gather_blocks; b += gridDim.x) {
        cull_gather_body(g, b, num_gather_blocks);
        __syncthreads();                                    // shared scan arrays are reused by the next block
    }
}

// Launch shapes selectable at run time (OVIS_CULL_VARIANT, tuning only). Variant 0 is the default.
static const CullVariant kVariants[] = { { 2, false, false }, { 1, false, false }, { 4, false, false }, { 1, true, false },
                                         { 2, true, false },  { 2, false, true },  { 1, false, true },
                                         { 1, false, false, 4 }, { 1, false, false, 3 }, { 1, false, false, 2 },
                                         { 1, true, true, 0 },   { 2, true, true, 0 },   { 1, false, true, 4 },  { 1, false, true, 2 } };

int cull_num_variants() { return (int)(sizeof(kVariants) / sizeof(kVariants[0])); }
int cull_variant_smem(int v) {
    const int st = kVariants[v].async_stages;
    return st == 4 ? AsyncRing<4>::kCtaBytes : st == 3 ? AsyncRing<3>::kCtaBytes : st == 2 ? AsyncRing<2>::kCtaBytes : 0;
}
CullVariant cull_variant(int v) { return kVariants[v]; }

template <typename F>
static void with_variant(int variant, bool masked, F &&f) {
    const CullVariant &cv = kVariants[variant];
    if (cv.packed && cv.async_stages == 4) { if (masked) f(cull_test_async_kernel<true, 4, true>); else f(cull_test_async_kernel<false, 4, true>); return; }
    if (cv.packed && cv.async_stages == 2) { if (masked) f(cull_test_async_kernel<true, 2, true>); else f(cull_test_async_kernel<false, 2, true>); return; }
    if (cv.packed && cv.pipelined) {
        if (cv.groups == 1) { if (masked) f(cull_test_pipelined_kernel<1, true, true>); else f(cull_test_pipelined_kernel<1, false, true>); }
        else { if (masked) f(cull_test_pipelined_kernel<2, true, true>); else f(cull_test_pipelined_kernel<2, false, true>); }
        return;
    }
    if (kVariants[variant].async_stages == 4) { if (masked) f(cull_test_async_kernel<true, 4>); else f(cull_test_async_kernel<false, 4>); return; }
    if (kVariants[variant].async_stages == 3) { if (masked) f(cull_test_async_kernel<true, 3>); else f(cull_test_async_kernel<false, 3>); return; }
    if (kVariants[variant].async_stages == 2) { if (masked) f(cull_test_async_kernel<true, 2>); else f(cull_test_async_kernel<false, 2>); return; }
    if (kVariants[variant].pipelined) {
        if (kVariants[variant].groups == 1) { if (masked) f(cull_test_pipelined_kernel<1, true>); else f(cull_test_pipelined_kernel<1, false>); }
        else { if (masked) f(cull_test_pipelined_kernel<2, true>); else f(cull_test_pipelined_kernel<2, false>); }
        return;
    }
    if (kVariants[variant].packed) {
        if (kVariants[variant].groups == 1) { if (masked) f(cull_test_kernel<1, true, true>); else f(cull_test_kernel<1, false, true>); }
        else { if (masked) f(cull_test_kernel<2, true, true>); else f(cull_test_kernel<2, false, true>); }
        return;
    }
    switch (kVariants[variant].groups) {
    case 1: if (masked) f(cull_test_kernel<1, true>); else f(cull_test_kernel<1, false>); break;
    case 2: if (masked) f(cull_test_kernel<2, true>); else f(cull_test_kernel<2, false>); break;
    case 4: if (masked) f(cull_test_kernel<4, true>); else f(cull_test_kernel<4, false>); break;
    }
}

// overlap_previous_gather: launch with programmatic stream serialization, so this test kernel may start while the
// PREVIOUS cull's gather kernel (which signals its dependents at once) is still running. Only legal when the two use
// different scratch sets; every other predecessor still completes first.
void launch_cull_test(const CullArgs &a, bool masked, int variant, int grid, cudaStream_t s, bool overlap_previous_gather) {
    with_variant(variant, masked, [&](auto kernel) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)grid);
        cfg.blockDim = dim3(kCullThreads);
        cfg.dynamicSmemBytes = (size_t)cull_variant_smem(variant);
        if (cfg.dynamicSmemBytes)
            cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cfg.dynamicSmemBytes);
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
    switch (kVariants[variant].groups) {
    case 1: if (masked) q(cull_fused_kernel<1, true>); else q(cull_fused_kernel<1, false>); break;
    case 2: if (masked) q(cull_fused_kernel<2, true>); else q(cull_fused_kernel<2, false>); break;
    }
    return blocks;
}

void launch_cull_gather(const GatherArgs &a, cudaStream_t s) {
    int blocks = (a.num_tiles + kGatherTiles - 1) / kGatherTiles;
    static const int cap = getenv("OVIS_GATHER_CTAS") ? atoi(getenv("OVIS_GATHER_CTAS")) : 0;   // tuning
    if (cap > 0 && blocks > cap)
        blocks = cap;
    cull_gather_kernel<<<blocks, 256, 0, s>>>(a);
}

int cull_test_max_blocks_per_sm(bool masked, int variant) {
    int blocks = 0;
    with_variant(variant, masked, [&](auto kernel) {
        const int smem = cull_variant_smem(variant);
        if (smem)
            cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks, kernel, kCullThreads, (size_t)smem);
    });
    return blocks;
}

// n == 0 never launches the tile loop; this writes the empty result instead.
__global__ void empty_result_kernel(const long long *d_base, long long *d_next, uint32_t *d_count) {
    if (d_count) *d_count = 0u;
    if (d_next) *d_next = d_base ? *d_base : 0ll;
}

void launch_empty_result(const long long *d_base, long long *d_next, uint32_t *d_count, cudaStream_t s) {
    empty_result_kernel<<<1, 1, 0, s>>>(d_base, d_next, d_count);
}

// ------------------------------------------------------------------------------------------------
// K5: RenderLight::cull for every light slot, ordered output. Light counts are small (<= a few
// thousand), so one CTA walks the slots in 256-wide chunks with a running base.
// ------------------------------------------------------------------------------------------------

__global__ void __launch_bounds__(256)
cull_lights_kernel(const __grid_constant__ LightCullArgs a) {
    __shared__ uint32_t s_warp_total[8];
    __shared__ uint32_t s_base;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0)
        s_base = 0;
    __syncthreads();
    for (int chunk = 0; chunk < a.n; chunk += 256) {
        int i = chunk + tid;
        bool visible = false;
        if (i < a.n) {
            DeviceLight l = a.lights[i];
            if (l.kind == kLightAlways) {
                visible = true;                                     // ambient / directional, render_light.cc:193-195
            } else if (l.kind == kLightSphere) {                    // point: Sphere(position, range), :197-201
                visible = true;
#pragma unroll
                for (int p = 0; p < 6; p++) {                       // intersection.cc:28-36
                    float dist = sub(add(add(mul(a.planes[p].x, l.a.x), mul(a.planes[p].y, l.a.y)), mul(a.planes[p].z, l.a.z)),
                                     a.planes[p].w);
                    if (dist < -l.a.w)
                        visible = false;
                }
            } else if (l.kind == kLightBox) {                       // spot: cone-fitted AABB, :203-204
                visible = true;
#pragma unroll
                for (int p = 0; p < 6; p++)
                    if (outside_plane(a.planes[p].x, a.planes[p].y, a.planes[p].z, a.planes[p].w, l.a.x, l.a.y, l.a.z,
                                      l.b.x, l.b.y, l.b.z))
                        visible = false;
            }
        }
        uint32_t ballot = __ballot_sync(0xffffffffu, visible);
        if (lane == 0)
            s_warp_total[warp] = __popc(ballot);
        __syncthreads();
        uint32_t rank = s_base + __popc(ballot & ((1u << lane) - 1u));
        uint32_t chunk_total = 0;
#pragma unroll
        for (int w = 0; w < 8; w++) {
            if (w < warp)
                rank += s_warp_total[w];
            chunk_total += s_warp_total[w];
        }
        if (visible)
            a.out[rank] = (uint32_t)i;
        __syncthreads();
        if (tid == 0)
            s_base += chunk_total;
        __syncthreads();
    }
    if (tid == 0)
        *a.d_count = s_base;
}

void launch_cull_lights(const LightCullArgs &a, cudaStream_t s) { cull_lights_kernel<<<1, 256, 0, s>>>(a); }

}  // namespace ovis
