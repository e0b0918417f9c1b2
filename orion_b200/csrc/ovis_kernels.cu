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
    DTrs w = load_trs(a.local_trs, a.nstride, node);
    int32_t parent = a.parent[node];
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
                                    float *world, int64_t estride) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
        return;
    entity_bounds_from_trs(load_trs(world_trs, nstride, nodes[i]), entities[i], local, world, estride);
}

void launch_extra_bounds(const uint32_t *entities, const uint32_t *nodes, int64_t n, const float *world_trs,
                         int64_t nstride, const float *local, float *world, int64_t estride, cudaStream_t s) {
    if (n > 0)
        extra_bounds_kernel<<<blocks_for(n, 256), 256, 0, s>>>(entities, nodes, n, world_trs, nstride, local, world, estride);
}

// ------------------------------------------------------------------------------------------------
// K3: one view over all entity slots with ORDER-PRESERVING compaction in a single pass.
//
// Persistent CTAs take 1024-entity tiles from an atomic ticket (so a tile's predecessors are always
// running or finished), test 4 entities per thread, rank survivors with a warp shuffle scan + a
// block scan, and obtain the tile's global base with a decoupled look-back over 64-bit tile status
// words {epoch, state, value}. Output order = slot order = insertion order, identical from run to
// run (no atomics decide a position). The epoch tag makes stale words from the previous launch read as
// "not ready", so the status array never needs clearing; the last CTA to leave resets the ticket.
// ------------------------------------------------------------------------------------------------

constexpr uint32_t kStateAggregate = 1u, kStatePrefix = 2u;

__device__ __forceinline__ uint64_t pack_status(uint32_t epoch, uint32_t state, uint32_t value) {
    return ((uint64_t)((epoch << 2) | state) << 32) | value;
}

__device__ __forceinline__ uint64_t load_status(const uint64_t *p) {
    uint64_t v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ void store_status(uint64_t *p, uint64_t v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// Called by all 32 lanes of warp 0. Returns the number of survivors in all earlier tiles.
__device__ __forceinline__ uint32_t tile_exclusive_prefix(uint64_t *status, int tile, uint32_t aggregate, uint32_t epoch,
                                                          int lane) {
    if (tile == 0) {
        if (lane == 0)
            store_status(status, pack_status(epoch, kStatePrefix, aggregate));
        return 0;
    }
    if (lane == 0)
        store_status(status + tile, pack_status(epoch, kStateAggregate, aggregate));
    uint32_t exclusive = 0;
    int look = tile - 1;
    while (true) {
        int idx = look - lane;
        uint32_t state = kStatePrefix, value = 0;      // tiles before tile 0: prefix 0
        if (idx >= 0) {
            uint64_t s = load_status(status + idx);
            uint32_t hi = (uint32_t)(s >> 32);
            state = ((hi >> 2) == epoch) ? (hi & 3u) : 0u;
            value = (uint32_t)s;
        }
        uint32_t not_ready = __ballot_sync(0xffffffffu, state == 0u);
        uint32_t prefixes = __ballot_sync(0xffffffffu, state == kStatePrefix);
        // lanes [0, stop) are usable when none of them is "not ready"; stop at the first prefix (inclusive)
        int first_prefix = prefixes ? (__ffs(prefixes) - 1) : 32;
        int first_bad = not_ready ? (__ffs(not_ready) - 1) : 32;
        if (first_bad < first_prefix && first_bad < 32) {
            if (first_bad == 0)
                continue;                                // nothing usable yet: poll again
            // consume the ready aggregates in front of the first unready tile, then poll from there
            uint32_t part = (lane < first_bad) ? value : 0u;
            exclusive += __reduce_add_sync(0xffffffffu, part);
            look -= first_bad;
            continue;
        }
        int take = first_prefix < 32 ? first_prefix + 1 : 32;
        uint32_t part = (lane < take) ? value : 0u;
        exclusive += __reduce_add_sync(0xffffffffu, part);
        if (first_prefix < 32)
            break;
        look -= 32;
    }
    if (lane == 0)
        store_status(status + tile, pack_status(epoch, kStatePrefix, exclusive + aggregate));
    return exclusive;
}

template <bool MASKED>
__global__ void __launch_bounds__(kCullThreads)
cull_frustum_kernel(const __grid_constant__ CullArgs a) {
    __shared__ uint32_t s_warp_total[kCullThreads / 32];
    __shared__ uint32_t s_tile_base;
    __shared__ int s_tile;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int num_tiles = a.num_tiles;
    const float4 *__restrict__ bx0 = reinterpret_cast<const float4 *>(a.bounds);
    const int64_t vstride = a.stride >> 2;            // stride in float4 units

    // output base of this view inside a CSR buffer (0 for a stand-alone list)
    const long long out_base = a.d_base ? *a.d_base : 0ll;

    while (true) {
        if (tid == 0)
            s_tile = (int)atomicAdd(a.ticket, 1u);
        __syncthreads();
        const int tile = s_tile;
        if (tile >= num_tiles)
            break;

        const int64_t v = (int64_t)tile * kCullThreads + tid;      // float4 index: entities 4v .. 4v+3
        const int64_t e0 = v << 2;
        const float4 mnx = __ldg(bx0 + 0 * vstride + v), mny = __ldg(bx0 + 1 * vstride + v), mnz = __ldg(bx0 + 2 * vstride + v);
        const float4 mxx = __ldg(bx0 + 3 * vstride + v), mxy = __ldg(bx0 + 4 * vstride + v), mxz = __ldg(bx0 + 5 * vstride + v);

        uint32_t keep = 0xfu;                                        // bit j: entity e0 + j still inside
        if (MASKED)
            keep = (__ldg(a.mask + (e0 >> 5)) >> (e0 & 31)) & 0xfu;
        if (e0 + 3 >= a.n) {                                         // ragged tail of the last tile
            int valid = (int)(a.n - e0);
            keep &= valid <= 0 ? 0u : ((1u << valid) - 1u);
        }
#pragma unroll
        for (int p = 0; p < 6; p++) {                                // Left Right Top Bottom Near Far
            const float nx = a.planes[p].x, ny = a.planes[p].y, nz = a.planes[p].z, d = a.planes[p].w;
            if (outside_plane(nx, ny, nz, d, mnx.x, mny.x, mnz.x, mxx.x, mxy.x, mxz.x)) keep &= ~1u;
            if (outside_plane(nx, ny, nz, d, mnx.y, mny.y, mnz.y, mxx.y, mxy.y, mxz.y)) keep &= ~2u;
            if (outside_plane(nx, ny, nz, d, mnx.z, mny.z, mnz.z, mxx.z, mxy.z, mxz.z)) keep &= ~4u;
            if (outside_plane(nx, ny, nz, d, mnx.w, mny.w, mnz.w, mxx.w, mxy.w, mxz.w)) keep &= ~8u;
        }

        // rank inside the warp: inclusive shuffle scan of the per-thread survivor counts
        const uint32_t mine = __popc(keep);
        uint32_t incl = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t up = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o)
                incl += up;
        }
        if (lane == 31)
            s_warp_total[warp] = incl;
        __syncthreads();

        if (warp == 0) {
            uint32_t wt = (lane < kCullThreads / 32) ? s_warp_total[lane] : 0u;
            uint32_t aggregate = __reduce_add_sync(0xffffffffu, wt);
            uint32_t excl = tile_exclusive_prefix(a.status, tile, aggregate, a.epoch, lane);
            if (lane == 0) {
                s_tile_base = excl;
                if (tile == num_tiles - 1) {                         // the last tile knows the total
                    uint32_t total = excl + aggregate;
                    if (a.d_count) *a.d_count = total;
                    if (a.d_next) *a.d_next = out_base + (long long)total;
                }
            }
        }
        __syncthreads();

        uint32_t rank = s_tile_base + incl - mine;
#pragma unroll
        for (int w = 0; w < kCullThreads / 32; w++)
            if (w < warp)
                rank += s_warp_total[w];
        if (keep) {
            long long pos = out_base + (long long)rank;
#pragma unroll
            for (int j = 0; j < 4; j++) {
                if (keep & (1u << j)) {
                    if (pos < a.capacity)
                        a.out[pos] = (uint32_t)(e0 + j) + a.id_base;
                    pos++;
                }
            }
        }
        // No trailing barrier: s_tile is rewritten by thread 0 only after every thread has passed the two
        // barriers above (so has read it), and s_warp_total / s_tile_base are rewritten only after the
        // next iteration's first barrier.
    }

    // the last CTA out re-arms the ticket for the next launch
    if (tid == 0) {
        __threadfence();
        unsigned int done = atomicAdd(a.done, 1u);
        if (done == gridDim.x - 1) {
            *a.ticket = 0u;
            *a.done = 0u;
            __threadfence();
        }
    }
}

void launch_cull_frustum(const CullArgs &a, bool masked, int grid, cudaStream_t s) {
    if (masked)
        cull_frustum_kernel<true><<<grid, kCullThreads, 0, s>>>(a);
    else
        cull_frustum_kernel<false><<<grid, kCullThreads, 0, s>>>(a);
}

int cull_frustum_max_blocks_per_sm(bool masked) {
    int blocks = 0;
    if (masked)
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks, cull_frustum_kernel<true>, kCullThreads, 0);
    else
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks, cull_frustum_kernel<false>, kCullThreads, 0);
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
