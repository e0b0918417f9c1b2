// ovis_kernels.cu - sm_100a kernels of the scene-visibility path and their launchers.
//
// K1+K2  update_level_kernel     Entity::transformed + Transform::matrix + BoundingBox::transform
// (K3, the entity cull, lives in ovis_cull.cu; K4 / K6 in ovis_multiview.cu)
// K5     cull_lights_kernel      RenderLight::cull
// plus layout movers (AoS host layout <-> SoA device layout).
//
// The path is HBM-bound byte/flop-light work: no tensor cores, no GEMM shaping. Bounds live in HBM as six
// SoA float arrays; a thread owns four consecutive entities and reads them with six 128-bit loads, so a
// warp covers 128 consecutive entities with fully coalesced 512 B requests per array.
#include "ovis_kernels.h"

#include <cstdlib>

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

// Stable compaction of per-entity arrays: item i of the new array is item ids[i] of the old one (ids ascending).
template <int K>
__global__ void compact_soa_kernel(const float *__restrict__ src, const uint32_t *__restrict__ ids, int64_t n, int64_t stride, float *dst) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
        return;
    const int64_t from = ids[i];
#pragma unroll
    for (int k = 0; k < K; k++)
        dst[k * stride + i] = src[k * stride + from];
}

__global__ void compact_u32_kernel(const uint32_t *__restrict__ src, const uint32_t *__restrict__ ids, int64_t n, uint32_t *dst) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n)
        dst[i] = src[ids[i]];
}

void launch_compact_soa6(const float *src, const uint32_t *ids, int64_t n, int64_t stride, float *dst, cudaStream_t s) {
    if (n > 0)
        compact_soa_kernel<6><<<blocks_for(n, 256), 256, 0, s>>>(src, ids, n, stride, dst);
}

void launch_compact_u32(const uint32_t *src, const uint32_t *ids, int64_t n, uint32_t *dst, cudaStream_t s) {
    if (n > 0)
        compact_u32_kernel<<<blocks_for(n, 256), 256, 0, s>>>(src, ids, n, dst);
}

// world TRS of a list of nodes, AoS (10 floats each): what lights bound to nodes read back after an update
__global__ void gather_trs_kernel(const float *__restrict__ src, int64_t stride, const uint32_t *__restrict__ ids, int64_t n, float *dst) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
        return;
    const int64_t node = ids[i];
#pragma unroll
    for (int k = 0; k < 10; k++)
        dst[i * 10 + k] = src[k * stride + node];
}

void launch_gather_trs(const float *src, int64_t stride, const uint32_t *ids, int64_t n, float *dst, cudaStream_t s) {
    if (n > 0)
        gather_trs_kernel<<<blocks_for(n, 256), 256, 0, s>>>(src, stride, ids, n, dst);
}

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
// K5: RenderLight::cull for every light slot, ordered output. One thread per light, 256 per CTA: every CTA leaves its
// visibility bits (one ballot word per warp); the CTA that finishes last turns the bit words into the ordered id list
// with a block scan of their popcounts (positions are sums of counts: insertion order, the same from run to run).
// (Round 1 walked all lights with ONE CTA, 16 chunks x 3 barriers for 4096 lights: 0.3 ms per cull with lights.)
// ------------------------------------------------------------------------------------------------

__global__ void __launch_bounds__(256)
cull_lights_kernel(const __grid_constant__ LightCullArgs a) {
    __shared__ uint32_t s_warp[8];
    __shared__ bool s_last;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int i = blockIdx.x * 256 + tid;
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
    const uint32_t ballot = __ballot_sync(0xffffffffu, visible);
    if (lane == 0)
        a.bits[blockIdx.x * 8 + warp] = ballot;
    __syncthreads();
    if (tid == 0) {
        __threadfence();
        s_last = atomicAdd(a.counter, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!s_last)
        return;
    __threadfence();
    const int words = gridDim.x * 8;
    uint32_t base = 0;
    for (int w0 = 0; w0 < words; w0 += 256) {
        uint32_t word = w0 + tid < words ? __ldcg(a.bits + w0 + tid) : 0u;
        const uint32_t c = __popc(word);
        uint32_t incl = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t up = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o)
                incl += up;
        }
        __syncthreads();                                       // s_warp is reused by every chunk
        if (lane == 31)
            s_warp[warp] = incl;
        __syncthreads();
        uint32_t before = 0, total = 0;
#pragma unroll
        for (int k = 0; k < 8; k++) {
            if (k < warp)
                before += s_warp[k];
            total += s_warp[k];
        }
        uint32_t rank = base + before + incl - c;
        const uint32_t id0 = (uint32_t)(w0 + tid) * 32u;
        while (word) {
            const int bit = __ffs(word) - 1;
            word &= word - 1;
            a.out[rank++] = id0 + bit;
        }
        base += total;
    }
    if (tid == 0) {
        *a.d_count = base;
        *a.counter = 0u;
    }
}

void launch_cull_lights(const LightCullArgs &a, cudaStream_t s) {
    if (a.n > 0)
        cull_lights_kernel<<<(unsigned)((a.n + 255) / 256), 256, 0, s>>>(a);
    else
        cudaMemsetAsync(a.d_count, 0, sizeof(uint32_t), s);
}

}  // namespace ovis
