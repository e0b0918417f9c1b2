// ovis_drawlist.cu - consumers of the visible lists, on the device:
//
//   draw lists     DeferredRenderPipeline::prepareEntities (render/src/deferred_render_pipeline.cc:318-330) and the
//                  shadow-caster loop of prepareLights (:272-283) with DrawList::add (render/src/draw_list.cc:35-45):
//                  every visible entity expands to one draw per shader pass of the chosen pass type, in list order.
//   uniforms       EntityUniforms{mat4 transform; vec3 position} (render/include/render_entity.h:34-37) as
//                  RenderEntity::setTransform fills them (render/src/render_entity.cc:67-69), for visible entities only.
//
// The reference hands these to GL/Vulkan command lists and uniform buffers; here they land in headless device buffers.
// Expansion is an ordered scan over the (short) visible list: per 1024-entry tile the draw counts are summed, each tile
// then adds the totals of the tiles before it (L2 reads) and writes its draws. Positions are sums of counts: the draw
// order is the list order, as in the reference's std::deque.
#include "ovis_kernels.h"

#include "ovis_device.cuh"

#include <algorithm>

namespace ovis {

constexpr int kDrawTile = 1024;       // visible-list entries per CTA (256 threads x 4)
constexpr int kDrawSingleLaunchTiles = 2;    // measured: re-counting more than ~2K entries costs more than the second launch

__device__ __forceinline__ void draw_counts(const DrawArgs &a, uint32_t id, uint32_t &c0, uint32_t &c1) {
    const uint32_t slot = id - a.id_base;
    c0 = c1 = 0;
    if (slot >= a.n_slots)                                   // not an id of this world (stale after a compaction, another shard's)
        return;
    const uint32_t pc = a.pass_counts[slot];
    if (a.mode == 0) {
        const uint32_t deferred = pc & 0xffu, basic = (pc >> 8) & 0xffu;
        if (deferred > 0) c0 = deferred;                     // deferred_render_pipeline.cc:322-323
        else c1 = basic;                                      // :324-325 (0 passes: nothing drawn, :326-327)
    } else {
        const bool caster = (a.caster_mask[slot >> 5] >> (slot & 31)) & 1u;   // entity->castsShadow(), :273
        if (caster) c0 = (pc >> 16) & 0xffu;
    }
}

// Phase 1: totals per tile.
__global__ void __launch_bounds__(256)
draw_count_kernel(const __grid_constant__ DrawArgs a) {
    __shared__ uint32_t s0[8], s1[8];
    // never beyond the host's bound that sized the grid: a longer device list is cut there (and reported: see draw_write_kernel)
    const uint32_t m = min(*a.d_visible_count, (uint32_t)min((long long)gridDim.x * kDrawTile, (long long)0xffffffffll));
    const int64_t base = (int64_t)blockIdx.x * kDrawTile;
    if (base >= m)
        return;
    uint32_t t0 = 0, t1 = 0;
    for (int k = 0; k < 4; k++) {
        int64_t i = base + k * 256 + threadIdx.x;
        if (i < m) {
            uint32_t c0, c1;
            draw_counts(a, a.d_visible[i], c0, c1);
            t0 += c0;
            t1 += c1;
        }
    }
    t0 = __reduce_add_sync(0xffffffffu, t0);
    t1 = __reduce_add_sync(0xffffffffu, t1);
    if ((threadIdx.x & 31) == 0) {
        s0[threadIdx.x >> 5] = t0;
        s1[threadIdx.x >> 5] = t1;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t a0 = 0, a1 = 0;
        for (int w = 0; w < 8; w++) { a0 += s0[w]; a1 += s1[w]; }
        a.tile_totals[2 * blockIdx.x] = a0;
        a.tile_totals[2 * blockIdx.x + 1] = a1;
    }
}

// Phase 2: base of the tile = totals of the tiles before it; then ordered expansion.
__global__ void __launch_bounds__(256)
draw_write_kernel(const __grid_constant__ DrawArgs a) {
    __shared__ uint32_t s_w0[8], s_w1[8];
    __shared__ uint32_t s_base[2];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // never beyond the host's bound that sized the grid: a longer device list is cut there (and reported: see draw_write_kernel)
    const uint32_t m = min(*a.d_visible_count, (uint32_t)min((long long)gridDim.x * kDrawTile, (long long)0xffffffffll));
    const int64_t base = (int64_t)blockIdx.x * kDrawTile;
    const int tiles = (int)((m + kDrawTile - 1) / kDrawTile);
    if (base >= m) {
        if (blockIdx.x == 0 && tid == 0) {                   // empty list
            a.d_draw_counts[0] = 0;
            a.d_draw_counts[1] = 0;
        }
        return;
    }
    uint32_t p0 = 0, p1 = 0;
    if (a.tile_totals) {
        for (int t = tid; t < (int)blockIdx.x; t += 256) {
            p0 += a.tile_totals[2 * t];
            p1 += a.tile_totals[2 * t + 1];
        }
    } else {                                                 // short list, single launch: count the earlier entries here
        for (int64_t i = tid; i < base; i += 256) {
            uint32_t c0, c1;
            draw_counts(a, a.d_visible[i], c0, c1);
            p0 += c0;
            p1 += c1;
        }
    }
    p0 = __reduce_add_sync(0xffffffffu, p0);
    p1 = __reduce_add_sync(0xffffffffu, p1);
    if (lane == 0) { s_w0[warp] = p0; s_w1[warp] = p1; }
    __syncthreads();
    if (tid == 0) {
        uint32_t b0 = 0, b1 = 0;
        for (int w = 0; w < 8; w++) { b0 += s_w0[w]; b1 += s_w1[w]; }
        s_base[0] = b0;
        s_base[1] = b1;
    }
    __syncthreads();

    // each thread owns 4 CONSECUTIVE entries so that a thread-level scan keeps list order
    uint32_t ids[4], c0[4], c1[4];
    uint32_t t0 = 0, t1 = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        int64_t i = base + tid * 4 + k;
        ids[k] = 0; c0[k] = 0; c1[k] = 0;
        if (i < m) {
            ids[k] = a.d_visible[i];
            draw_counts(a, ids[k], c0[k], c1[k]);
        }
        t0 += c0[k];
        t1 += c1[k];
    }
    uint32_t i0 = t0, i1 = t1;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t u0 = __shfl_up_sync(0xffffffffu, i0, o), u1 = __shfl_up_sync(0xffffffffu, i1, o);
        if (lane >= o) { i0 += u0; i1 += u1; }
    }
    if (lane == 31) { s_w0[warp] = i0; s_w1[warp] = i1; }
    __syncthreads();
    uint32_t b0 = s_base[0], b1 = s_base[1], tile0 = 0, tile1 = 0;
#pragma unroll
    for (int w = 0; w < 8; w++) {
        if (w < warp) { b0 += s_w0[w]; b1 += s_w1[w]; }
        tile0 += s_w0[w];
        tile1 += s_w1[w];
    }
    long long pos0 = (long long)b0 + (i0 - t0), pos1 = (long long)b1 + (i1 - t1);
#pragma unroll
    for (int k = 0; k < 4; k++) {
        for (uint32_t p = 0; p < c0[k]; p++, pos0++)          // DrawList::add: one draw per pass, pass index ascending
            if (pos0 < a.cap0) a.d_draws0[pos0] = make_uint2(ids[k], p);
        for (uint32_t p = 0; p < c1[k]; p++, pos1++)
            if (pos1 < a.cap1) a.d_draws1[pos1] = make_uint2(ids[k], p);
    }
    if ((int)blockIdx.x == tiles - 1 && tid == 0) {          // the last tile knows the totals
        // a visible list longer than the caller's bound cannot be expanded completely: say so instead of reporting the
        // draws of its first part as if they were all (0xffffffff in both counts = "max_visible was too small")
        const bool cut = *a.d_visible_count > m;
        a.d_draw_counts[0] = cut ? 0xffffffffu : s_base[0] + tile0;
        a.d_draw_counts[1] = cut ? 0xffffffffu : s_base[1] + tile1;
    }
}

int launch_draw_lists(const DrawArgs &a, int64_t max_visible, cudaStream_t s) {
    int tiles = (int)((max_visible + kDrawTile - 1) / kDrawTile);
    if (tiles < 1)
        tiles = 1;
    if (tiles <= kDrawSingleLaunchTiles) {                   // very short list: re-counting one tile beats a second launch
        DrawArgs b = a;
        b.tile_totals = nullptr;
        draw_write_kernel<<<tiles, 256, 0, s>>>(b);
        return 1;
    }
    draw_count_kernel<<<tiles, 256, 0, s>>>(a);
    draw_write_kernel<<<tiles, 256, 0, s>>>(a);
    return 2;
}

// EntityUniforms for the visible entities: 20 floats each (mat4 column-major, position, pad) = std140's 80 bytes.
__global__ void __launch_bounds__(256)
entity_uniforms_kernel(const __grid_constant__ UniformArgs a) {
    const uint32_t m = *a.d_visible_count;
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m || i >= a.capacity)
        return;
    const uint32_t slot = a.d_visible[i] - a.id_base;
    float4 *out = reinterpret_cast<float4 *>(a.d_out + 20 * i);
    if (slot >= a.n_slots) {                                 // not an id of this world: a zero block, no table is read
        out[0] = out[1] = out[2] = out[3] = out[4] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        return;
    }
    const uint32_t node = a.entity_node[slot];
    DTrs t = { { 0.0f, 0.0f, 0.0f }, { 1.0f, 0.0f, 0.0f, 0.0f }, { 1.0f, 1.0f, 1.0f } };   // Transform() of an entity never moved
    if (node != kNone) {
        const float *b = a.world_trs;
        const int64_t st = a.nstride;
        t.p = { b[0 * st + node], b[1 * st + node], b[2 * st + node] };
        t.q = { b[3 * st + node], b[4 * st + node], b[5 * st + node], b[6 * st + node] };
        t.s = { b[7 * st + node], b[8 * st + node], b[9 * st + node] };
    }
    const DMat4 mtx = trs_matrix(t);                           // Transform::matrix(), render_entity.cc:68
    out[0] = make_float4(mtx.c[0].x, mtx.c[0].y, mtx.c[0].z, mtx.c[0].w);
    out[1] = make_float4(mtx.c[1].x, mtx.c[1].y, mtx.c[1].z, mtx.c[1].w);
    out[2] = make_float4(mtx.c[2].x, mtx.c[2].y, mtx.c[2].z, mtx.c[2].w);
    out[3] = make_float4(mtx.c[3].x, mtx.c[3].y, mtx.c[3].z, mtx.c[3].w);
    out[4] = make_float4(t.p.x, t.p.y, t.p.z, 0.0f);           // uniforms->position = m_transform.position(), :69
}

void launch_entity_uniforms(const UniformArgs &a, int64_t max_visible, cudaStream_t s) {
    if (max_visible > 0)
        entity_uniforms_kernel<<<(unsigned)((max_visible + 255) / 256), 256, 0, s>>>(a);
}

// LightUniforms of the visible lights, in list order: 208-byte blocks gathered from the per-light table (13 float4 each).
__global__ void __launch_bounds__(256)
light_uniforms_kernel(const uint32_t *visible, const uint32_t *visible_count, const float4 *table, uint32_t n_lights, float4 *out,
                      int64_t capacity) {
    const uint32_t m = *visible_count;
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t i = t / 13;
    const int q = (int)(t - i * 13);
    if (i >= m || i >= capacity)
        return;
    const uint32_t id = visible[i];
    out[i * 13 + q] = id < n_lights ? table[(int64_t)id * 13 + q] : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
}

void launch_light_uniforms(const uint32_t *d_visible, const uint32_t *d_visible_count, const float *d_table, uint32_t n_lights,
                           float *d_out, int64_t max_visible, int64_t capacity, cudaStream_t s) {
    const int64_t threads = std::min(max_visible, capacity) * 13;
    if (threads > 0)
        light_uniforms_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, s>>>(
            d_visible, d_visible_count, reinterpret_cast<const float4 *>(d_table), n_lights, reinterpret_cast<float4 *>(d_out), capacity);
}

}  // namespace ovis
