// ovis_raster.cu - headless rasteriser: the G-buffer pass and the shadow-map passes of the reference's deferred pipeline
// as CUDA kernels over plain device buffers.
//
// What it replaces: DeferredRenderPipeline::renderShadowMaps / renderDeferredGBuffer
// (render/src/deferred_render_pipeline.cc:335-370, 384-412) drawing DrawList entries (render/src/draw_list.cc:52-60)
// with the shaders engine/shaders/lit_vtx.glsl + lit_deferred_frag.glsl and shadow_vtx.glsl + shadow_frag.glsl, under
// the default pipeline state (gpu/include/state.h:105-145: depth test LessOrEqual, depth write on, cull back faces,
// counter-clockwise = front; clear depth 1, clear colours 0). The reference hands that to OpenGL / Vulkan; here the
// fixed-function part is written out:
//
//   raster_setup_kernel    one CTA per drawn instance (an entity of a view's list): MVP = viewProjection * transform
//                          (GLM's mat4 * mat4), every triangle's clip-space vertices (GLM's mat4 * vec4, as
//                          "view.viewProjection * entity.transform * vec4(attribPosition, 1.0)" evaluates), back-face and
//                          trivial-reject tests, homogeneous edge equations (Olano & Greer: no clipping is ever needed,
//                          a triangle crossing the near plane is rasterised as it is), screen bounding box, one work item
//                          per 16 x 16-pixel tile the box touches.
//   raster_fill_kernel     a warp per work item, 8 pixels per lane: coverage at pixel centres (top-left rule on exact
//                          zeros), perspective-correct depth z/w, near / far discard, then ONE 64-bit atomicMin per
//                          covered pixel into a visibility buffer: depth bits in the high word, ~triangle id in the low
//                          word - equal depths resolve in favour of the LATER triangle, which is what LessOrEqual does
//                          with primitives in submission order. Depth-only targets (shadow maps) take a 32-bit atomicMin
//                          on the depth bits (depths are in [0, 1]: their bit patterns order like the values).
//   raster_resolve_kernel  G-buffer only: per pixel, the winning triangle's attributes are interpolated with the same
//                          barycentrics and lit_deferred_frag.glsl runs once per pixel (a visibility buffer: no
//                          overdraw of the fragment work).
//
// Draw order = list order (DrawList keeps the cull's order, draw_list.cc:35-45); triangle ids are
// first_triangle[instance] + index, a scan over the instances, so the image is identical from run to run.
// Textured / normal-mapped shader variants and the skybox are not built (flat material parameters: the shader variants
// without TEXTURED / NORMALMAP); MSAA does not exist in the reference.
#include "orion_vis.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include <cuda_runtime.h>

namespace ovis {
int report_error(int code, const char *message);
}

namespace {

int fail(int code, const std::string &m) { return ovis::report_error(code, m.c_str()); }

#define RS_CUDA(expr)                                                                                                  \
    do {                                                                                                               \
        cudaError_t err__ = (expr);                                                                                    \
        if (err__ != cudaSuccess)                                                                                      \
            return fail(OVIS_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(err__));                         \
    } while (0)
#define RS_TRY(expr)                                                                                                   \
    do {                                                                                                               \
        int rc__ = (expr);                                                                                             \
        if (rc__ != OVIS_OK)                                                                                           \
            return rc__;                                                                                               \
    } while (0)

constexpr int kTile = 16;                       // work item = one 16 x 16-pixel tile of one triangle
constexpr int kRasterThreads = 256;

struct RsView {                                 // one target of a call
    float vp[16];                               // viewProjection, column-major
    int width, height;
    long long first_instance;                   // its instances are [first_instance, first_instance + n_instances)
    long long n_instances;
    unsigned long long *vis;                    // G-buffer target: width x height keys; nullptr for depth-only
    unsigned int *depth_bits;                   // depth-only target: width x height float bit patterns
};

struct RsMeshRef {
    long long first_vertex, first_index;
    int n_triangles;
    int pad;
};

struct RsTriangle {                             // 80 bytes: what fill and resolve need of a set-up triangle
    float a[3], b[3], c[3];                     // edge functions E_i(x, y) = a_i x + b_i y + c_i, oriented so that inside is >= 0
    float z[3], w[3];                           // clip-space z and w of the vertices
    unsigned int instance;                      // index into the call's instance list
    unsigned int view;
    unsigned short x0, y0, x1, y1;              // pixel bounding box, inclusive, clamped to the target
    unsigned int pad;
};
struct RsNormals {                              // G-buffer passes only: world-space normals of the three vertices
    float n[9];                                 // vec3(transform * vec4(normal, 0))
    float pad[3];
};

struct RsTexAttribs {                           // G-buffer passes of worlds with textures: what TEXTURED / NORMALMAP interpolate
    float uv[6];                                // vtxTexcoord of the three vertices
    float t[9], b[9];                           // vtxTangent, vtxBitangent (lit_vtx.glsl:48-55)
};
struct RsTexture {
    const float4 *texels;                       // decoded to linear float at upload
    int width, height;
    int wrap;                                   // repeat addressing, else clamp to edge
    int pad;
};

struct RsItem {
    unsigned int triangle;                      // id = first_triangle[instance] + index
    unsigned short tx, ty;                      // tile coordinates
};

struct RsArgs {
    const RsView *views;
    int n_views;
    long long n_instances;
    const unsigned int *instance_entity;        // entity id of every instance (the CSR entity list of the call)
    const float *instance_uniforms;             // 20 floats per instance: transform (column-major), position, pad
    const unsigned int *entity_mesh;            // per entity: mesh id, 0xffffffff = none
    long long n_entities_bound;
    unsigned int id_base;
    const RsMeshRef *meshes;
    const float *vertices;                      // 8 floats per vertex: position, normal, texcoord
    const unsigned int *indices;
    long long *first_triangle;                  // [n_instances + 1]: scan of the instances' triangle counts
    RsTriangle *triangles;
    RsNormals *normals;                         // nullptr for depth-only calls
    RsTexAttribs *tex_attribs;                  // nullptr unless the raster holds textures (and this is a G-buffer pass)
    const float *tangents;                      // 4 floats per vertex (zeros for meshes that were given none)
    long long triangle_capacity;
    RsItem *items;
    unsigned int item_capacity;
    unsigned int *counters;                     // [0] items appended (may exceed the capacity), [1] overflow flag
};

// ---- GLM 0.9.8 arithmetic, written out (no FMA: the library is built -fmad=false and these use the rn intrinsics) ----
__device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }

// mat4 * mat4 (detail/type_mat4x4.inl:601-620): Res[j] = A0 * B[j][0] + A1 * B[j][1] + A2 * B[j][2] + A3 * B[j][3], left to right
__device__ void mat4_mul(const float *A, const float *B, float *R) {
#pragma unroll
    for (int j = 0; j < 4; j++)
#pragma unroll
        for (int r = 0; r < 4; r++)
            R[4 * j + r] = add(add(add(mul(A[r], B[4 * j]), mul(A[4 + r], B[4 * j + 1])), mul(A[8 + r], B[4 * j + 2])), mul(A[12 + r], B[4 * j + 3]));
}

// mat4 * vec4 (detail/type_mat4x4.inl:541-566): (m[0] * x + m[1] * y) + (m[2] * z + m[3] * w)
__device__ __forceinline__ void mat4_vec4(const float *M, float x, float y, float z, float w, float out[4]) {
#pragma unroll
    for (int r = 0; r < 4; r++)
        out[r] = add(add(mul(M[r], x), mul(M[4 + r], y)), add(mul(M[8 + r], z), mul(M[12 + r], w)));
}

// ---- instance -> triangles: counts and their scan --------------------------------------------------------------------
__device__ __forceinline__ int instance_triangles(const RsArgs &a, long long k) {
    const unsigned int e = a.instance_entity[k] - a.id_base;
    if ((long long)e >= a.n_entities_bound)
        return 0;
    const unsigned int m = a.entity_mesh[e];
    return m == 0xffffffffu ? 0 : a.meshes[m].n_triangles;
}

// one CTA: exclusive scan of the instances' triangle counts (a call draws thousands of instances, not millions)
__global__ void __launch_bounds__(1024)
raster_scan_kernel(const RsArgs a) {
    __shared__ long long s_warp[32];
    __shared__ long long s_carry;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0)
        s_carry = 0;
    __syncthreads();
    for (long long k0 = 0; k0 < a.n_instances; k0 += 1024) {
        const long long k = k0 + tid;
        const long long c = k < a.n_instances ? instance_triangles(a, k) : 0;
        long long incl = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const long long up = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o)
                incl += up;
        }
        if (lane == 31)
            s_warp[warp] = incl;
        __syncthreads();
        long long before = s_carry, sum = 0;
        for (int q = 0; q < 32; q++) {
            if (q < warp)
                before += s_warp[q];
            sum += s_warp[q];
        }
        if (k < a.n_instances)
            a.first_triangle[k] = before + incl - c;
        __syncthreads();
        if (tid == 0)
            s_carry += sum;
        __syncthreads();
    }
    if (tid == 0)
        a.first_triangle[a.n_instances] = s_carry;
}

__device__ __forceinline__ int view_of_instance(const RsArgs &a, long long k) {
    int lo = 0, hi = a.n_views - 1;                          // views hold consecutive, ascending instance ranges
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (a.views[mid].first_instance <= k)
            lo = mid;
        else
            hi = mid - 1;
    }
    return lo;
}

// ---- set-up ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kRasterThreads)
raster_setup_kernel(const RsArgs a) {
    __shared__ float s_mvp[16], s_model[16];
    __shared__ int s_view;
    const long long k = blockIdx.x;
    const int n_tri = instance_triangles(a, k);
    if (n_tri == 0)
        return;
    if (threadIdx.x == 0) {
        const int v = view_of_instance(a, k);
        s_view = v;
        const float *M = a.instance_uniforms + 20 * k;
        float mvp[16];
        mat4_mul(a.views[v].vp, M, mvp);                     // "view.viewProjection * entity.transform" (lit_vtx.glsl:57)
        for (int i = 0; i < 16; i++) {
            s_mvp[i] = mvp[i];
            s_model[i] = M[i];
        }
    }
    __syncthreads();
    const RsView &view = a.views[s_view];
    const RsMeshRef mesh = a.meshes[a.entity_mesh[a.instance_entity[k] - a.id_base]];
    const long long first = a.first_triangle[k];
    const float half_w = 0.5f * (float)view.width, half_h = 0.5f * (float)view.height;
    for (int t = threadIdx.x; t < n_tri; t += kRasterThreads) {
        const long long id = first + t;
        if (id >= a.triangle_capacity) {
            a.counters[1] = 1u;
            continue;
        }
        RsTriangle tri;
        RsNormals nrms;
        RsTexAttribs ta;
        float X[3], Y[3];
#pragma unroll
        for (int i = 0; i < 3; i++) {
            const float *vtx = a.vertices + 8 * (mesh.first_vertex + a.indices[mesh.first_index + 3 * t + i]);
            float clip[4], nrm[4];
            mat4_vec4(s_mvp, vtx[0], vtx[1], vtx[2], 1.0f, clip);
            mat4_vec4(s_model, vtx[3], vtx[4], vtx[5], 0.0f, nrm);       // vtxNormal = vec3(entity.transform * vec4(attribNormal, 0.0))
            X[i] = mul(add(clip[0], clip[3]), half_w);         // X / w = window x, Y / w = window y (origin bottom left)
            Y[i] = mul(add(clip[1], clip[3]), half_h);
            tri.z[i] = clip[2];
            tri.w[i] = clip[3];
            nrms.n[3 * i] = nrm[0]; nrms.n[3 * i + 1] = nrm[1]; nrms.n[3 * i + 2] = nrm[2];
            if (a.tex_attribs) {
                const float *tg = a.tangents + 4 * (mesh.first_vertex + a.indices[mesh.first_index + 3 * t + i]);
                float tw[4];
                mat4_vec4(s_model, tg[0], tg[1], tg[2], 0.0f, tw);        // vtxTangent = vec3(entity.transform * vec4(tangent, 0.0))
                ta.uv[2 * i] = vtx[6]; ta.uv[2 * i + 1] = vtx[7];
                ta.t[3 * i] = tw[0]; ta.t[3 * i + 1] = tw[1]; ta.t[3 * i + 2] = tw[2];
                // vtxBitangent = cross(vtxNormal, vtxTangent) * attribTangent.w (GLM cross: x = a.y b.z - b.y a.z, ...)
                ta.b[3 * i] = mul(sub(mul(nrm[1], tw[2]), mul(tw[1], nrm[2])), tg[3]);
                ta.b[3 * i + 1] = mul(sub(mul(nrm[2], tw[0]), mul(tw[2], nrm[0])), tg[3]);
                ta.b[3 * i + 2] = mul(sub(mul(nrm[0], tw[1]), mul(tw[0], nrm[1])), tg[3]);
            }
        }
        // edge functions: rows of the adjugate of [X; Y; w] (2-D homogeneous rasterisation)
#pragma unroll
        for (int i = 0; i < 3; i++) {
            const int i1 = (i + 1) % 3, i2 = (i + 2) % 3;
            tri.a[i] = sub(mul(Y[i1], tri.w[i2]), mul(Y[i2], tri.w[i1]));
            tri.b[i] = sub(mul(X[i2], tri.w[i1]), mul(X[i1], tri.w[i2]));
            tri.c[i] = sub(mul(X[i1], Y[i2]), mul(X[i2], Y[i1]));
        }
        const float det = add(add(mul(X[0], tri.a[0]), mul(Y[0], tri.b[0])), mul(tri.w[0], tri.c[0]));
        tri.instance = (unsigned int)k;
        tri.view = (unsigned int)s_view;
        bool draw = det > 0.0f;                                // counter-clockwise = front, back faces culled; NaN: not drawn
        // trivial rejects: all three vertices beyond one clip plane
        draw = draw && !(tri.z[0] < 0.0f && tri.z[1] < 0.0f && tri.z[2] < 0.0f) &&
               !(tri.z[0] > tri.w[0] && tri.z[1] > tri.w[1] && tri.z[2] > tri.w[2]);
        int x0 = 0, x1 = view.width - 1, y0 = 0, y1 = view.height - 1;
        if (draw && tri.w[0] > 0.0f && tri.w[1] > 0.0f && tri.w[2] > 0.0f) {          // all in front: the projected box
            const float px[3] = { X[0] / tri.w[0], X[1] / tri.w[1], X[2] / tri.w[2] };
            const float py[3] = { Y[0] / tri.w[0], Y[1] / tri.w[1], Y[2] / tri.w[2] };
            const float lx = fminf(fminf(px[0], px[1]), px[2]), hx = fmaxf(fmaxf(px[0], px[1]), px[2]);
            const float ly = fminf(fminf(py[0], py[1]), py[2]), hy = fmaxf(fmaxf(py[0], py[1]), py[2]);
            if (!(hx >= 0.0f) || !(lx <= (float)view.width) || !(hy >= 0.0f) || !(ly <= (float)view.height)) {
                draw = false;
            } else {                                           // pixel centres at +0.5; one pixel of slack for the rounding of px, py
                x0 = max(x0, (int)floorf(lx - 0.5f));
                x1 = min(x1, (int)ceilf(hx - 0.5f));
                y0 = max(y0, (int)floorf(ly - 0.5f));
                y1 = min(y1, (int)ceilf(hy - 0.5f));
            }
        }
        if (!draw || x1 < x0 || y1 < y0)
            continue;                                          // (its record is never referenced: not written)
        tri.x0 = (unsigned short)x0; tri.y0 = (unsigned short)y0; tri.x1 = (unsigned short)x1; tri.y1 = (unsigned short)y1;
        tri.pad = 0u;
        a.triangles[id] = tri;
        if (a.normals)
            a.normals[id] = nrms;
        if (a.tex_attribs)
            a.tex_attribs[id] = ta;
        const int tx0 = x0 / kTile, tx1 = x1 / kTile, ty0 = y0 / kTile, ty1 = y1 / kTile;
        const unsigned int n_items = (unsigned int)((tx1 - tx0 + 1) * (ty1 - ty0 + 1));
        const unsigned int at = atomicAdd(a.counters, n_items);
        unsigned int i = 0;
        for (int ty = ty0; ty <= ty1; ty++)
            for (int tx = tx0; tx <= tx1; tx++, i++)
                if (at + i < a.item_capacity)
                    a.items[at + i] = { (unsigned int)id, (unsigned short)tx, (unsigned short)ty };
    }
}

// ---- coverage and depth of one pixel centre: shared by fill and resolve so that both see the same numbers --------------
struct RsSample {
    float e[3];
    float depth;
    bool covered;
};

template <bool EARLY_OUT>
__device__ __forceinline__ RsSample sample_pixel(const RsTriangle &t, int px, int py) {
    RsSample s;
    const float fx = (float)px + 0.5f, fy = (float)py + 0.5f;
    bool inside = true;
#pragma unroll
    for (int i = 0; i < 3; i++) {
        s.e[i] = add(add(mul(t.a[i], fx), mul(t.b[i], fy)), t.c[i]);
        // a centre exactly on an edge belongs to the triangle whose edge is a "top" or "left" one
        inside = inside && (s.e[i] > 0.0f || (s.e[i] == 0.0f && (t.a[i] > 0.0f || (t.a[i] == 0.0f && t.b[i] > 0.0f))));
    }
    if (EARLY_OUT && !inside) {                                 // (the fill kernel: most centres of a bounding box are outside)
        s.depth = 2.0f;
        s.covered = false;
        return s;
    }
    const float zn = add(add(mul(s.e[0], t.z[0]), mul(s.e[1], t.z[1])), mul(s.e[2], t.z[2]));
    const float wn = add(add(mul(s.e[0], t.w[0]), mul(s.e[1], t.w[1])), mul(s.e[2], t.w[2]));
    s.depth = __fdiv_rn(zn, wn);                               // z / w, perspective-correct (the normalisation of the weights cancels)
    s.covered = inside && wn > 0.0f && s.depth >= 0.0f && s.depth <= 1.0f;       // near and far clip per pixel
    return s;
}

__global__ void __launch_bounds__(kRasterThreads)
raster_fill_kernel(const RsArgs a) {
    const unsigned int n_items = min(a.counters[0], a.item_capacity);
    const int lane = threadIdx.x & 31;
    const unsigned int warps = gridDim.x * (kRasterThreads / 32);
    for (unsigned int i = blockIdx.x * (kRasterThreads / 32) + (threadIdx.x >> 5); i < n_items; i += warps) {
        const RsItem it = a.items[i];
        const RsTriangle tri = a.triangles[it.triangle];
        const RsView &view = a.views[tri.view];
        // the part of the tile inside the triangle's bounding box, walked 32 pixels at a time (a tessellated mesh's triangles
        // cover a handful of pixels: testing all 256 of the tile for each was 385 instructions per item)
        const int rx0 = max((int)tri.x0, it.tx * kTile), rx1 = min((int)tri.x1, it.tx * kTile + kTile - 1);
        const int ry0 = max((int)tri.y0, it.ty * kTile), ry1 = min((int)tri.y1, it.ty * kTile + kTile - 1);
        const int rw = rx1 - rx0 + 1, count = rw * (ry1 - ry0 + 1);
        const unsigned int inv_rw = 65536u / (unsigned int)rw + 1u;      // p / rw for p < 256, rw <= 16: (p * inv_rw) >> 16
        for (int p = lane; p < count; p += 32) {
            const int row = (int)(((unsigned int)p * inv_rw) >> 16);
            const int px = rx0 + (p - row * rw), py = ry0 + row;
            const RsSample s = sample_pixel<true>(tri, px, py);
            if (!s.covered)
                continue;
            const long long at = (long long)py * view.width + px;
            if (view.vis)
                atomicMin(view.vis + at, ((unsigned long long)__float_as_uint(s.depth) << 32) | (unsigned long long)(~it.triangle));
            else
                atomicMin(view.depth_bits + at, __float_as_uint(s.depth));
        }
    }
}

// texture(sampler2D, uv) as this library defines it (orion_vis.h, ovis_raster_add_texture): bilinear on the base level
__device__ __forceinline__ int tex_coord(int i, int size, int wrap) {
    if (wrap) {
        i %= size;
        return i < 0 ? i + size : i;
    }
    return min(max(i, 0), size - 1);
}

__device__ float4 sample_texture(const RsTexture &t, float u, float v) {
    float x = sub(mul(u, (float)t.width), 0.5f), y = sub(mul(v, (float)t.height), 0.5f);
    if (!(fabsf(x) < 1.0e9f)) x = 0.0f;                        // NaN or absurd coordinates: texel 0
    if (!(fabsf(y) < 1.0e9f)) y = 0.0f;
    const float fx0 = floorf(x), fy0 = floorf(y);
    const float fx = sub(x, fx0), fy = sub(y, fy0);
    const int i0 = tex_coord((int)fx0, t.width, t.wrap), i1 = tex_coord((int)fx0 + 1, t.width, t.wrap);
    const int j0 = tex_coord((int)fy0, t.height, t.wrap), j1 = tex_coord((int)fy0 + 1, t.height, t.wrap);
    const float4 t00 = t.texels[(long long)j0 * t.width + i0], t10 = t.texels[(long long)j0 * t.width + i1];
    const float4 t01 = t.texels[(long long)j1 * t.width + i0], t11 = t.texels[(long long)j1 * t.width + i1];
    const float gx = sub(1.0f, fx), gy = sub(1.0f, fy);
    float4 r;
    r.x = add(mul(add(mul(t00.x, gx), mul(t10.x, fx)), gy), mul(add(mul(t01.x, gx), mul(t11.x, fx)), fy));
    r.y = add(mul(add(mul(t00.y, gx), mul(t10.y, fx)), gy), mul(add(mul(t01.y, gx), mul(t11.y, fx)), fy));
    r.z = add(mul(add(mul(t00.z, gx), mul(t10.z, fx)), gy), mul(add(mul(t01.z, gx), mul(t11.z, fx)), fy));
    r.w = add(mul(add(mul(t00.w, gx), mul(t10.w, fx)), gy), mul(add(mul(t01.w, gx), mul(t11.w, fx)), fy));
    return r;
}

__device__ __forceinline__ void normalize3(float v[3]) {          // v * inversesqrt(dot(v, v))
    const float inv = __fdiv_rn(1.0f, __fsqrt_rn(add(add(mul(v[0], v[0]), mul(v[1], v[1])), mul(v[2], v[2]))));
    v[0] = mul(v[0], inv); v[1] = mul(v[1], inv); v[2] = mul(v[2], inv);
}

struct RsResolveArgs {
    const unsigned long long *vis;
    const RsTriangle *triangles;
    const RsNormals *normals;
    const RsTexAttribs *tex_attribs;
    const RsTexture *textures;
    const unsigned int *instance_entity;
    const unsigned int *entity_material;
    const OvisMaterial *materials;
    unsigned int id_base;
    int width, height;
    float4 *buffer_a, *buffer_b, *buffer_c;
    float *depth;
};

// lit_deferred_frag.glsl:45-74 without TEXTURED / NORMALMAP, once per pixel for the triangle that won the depth test
__global__ void __launch_bounds__(kRasterThreads)
raster_resolve_kernel(const RsResolveArgs a) {
    const long long p = (long long)blockIdx.x * kRasterThreads + threadIdx.x;
    if (p >= (long long)a.width * a.height)
        return;
    const unsigned long long key = a.vis[p];
    if (key == ~0ull) {                                        // nothing drawn: the clear values (deferred_render_pipeline.cc:391-394)
        a.buffer_a[p] = a.buffer_b[p] = a.buffer_c[p] = make_float4(0.f, 0.f, 0.f, 0.f);
        a.depth[p] = 1.0f;
        return;
    }
    const RsTriangle t = a.triangles[~(unsigned int)key];
    const RsNormals tn = a.normals[~(unsigned int)key];
    const int px = (int)(p % a.width), py = (int)(p / a.width);
    const RsSample s = sample_pixel<false>(t, px, py);
    const float sum = add(add(s.e[0], s.e[1]), s.e[2]);
    const float b0 = __fdiv_rn(s.e[0], sum), b1 = __fdiv_rn(s.e[1], sum), b2 = __fdiv_rn(s.e[2], sum);
    float n[3];
#pragma unroll
    for (int c = 0; c < 3; c++)
        n[c] = add(add(mul(b0, tn.n[c]), mul(b1, tn.n[3 + c])), mul(b2, tn.n[6 + c]));
    const OvisMaterial m = a.materials[a.entity_material[a.instance_entity[t.instance] - a.id_base]];
    float diffuse[3] = { m.diffuse[0], m.diffuse[1], m.diffuse[2] };
    if ((m.flags & (OVIS_MATERIAL_TEXTURED | OVIS_MATERIAL_NORMALMAP)) && a.tex_attribs) {
        const RsTexAttribs ta = a.tex_attribs[~(unsigned int)key];
        const float u = add(add(mul(b0, ta.uv[0]), mul(b1, ta.uv[2])), mul(b2, ta.uv[4]));
        const float v = add(add(mul(b0, ta.uv[1]), mul(b1, ta.uv[3])), mul(b2, ta.uv[5]));
        if (m.flags & OVIS_MATERIAL_TEXTURED) {                 // diffuse = texture(diffuseTexture, vtxTexcoord)
            const float4 d = sample_texture(a.textures[m.diffuse_texture], u, v);
            diffuse[0] = d.x; diffuse[1] = d.y; diffuse[2] = d.z;
        }
        if (m.flags & OVIS_MATERIAL_NORMALMAP) {                // lit_deferred_frag.glsl:52-63
            const float4 sn4 = sample_texture(a.textures[m.normal_texture], u, v);
            float sn[3] = { sub(mul(sn4.x, 2.0f), 1.0f), sub(mul(sn4.y, 2.0f), 1.0f), sub(mul(sn4.z, 2.0f), 1.0f) };
            normalize3(sn);
            float T[3], B[3], N[3] = { n[0], n[1], n[2] };
#pragma unroll
            for (int c = 0; c < 3; c++) {
                T[c] = add(add(mul(b0, ta.t[c]), mul(b1, ta.t[3 + c])), mul(b2, ta.t[6 + c]));
                B[c] = add(add(mul(b0, ta.b[c]), mul(b1, ta.b[3 + c])), mul(b2, ta.b[6 + c]));
            }
            normalize3(T);
            normalize3(B);
            normalize3(N);
#pragma unroll
            for (int c = 0; c < 3; c++)                          // mat3(T, B, N) * sampledNormal (GLM: m0 * x + m1 * y + m2 * z)
                n[c] = add(add(mul(T[c], sn[0]), mul(B[c], sn[1])), mul(N[c], sn[2]));
        }
    }
    // normalize(normal): v * inversesqrt(dot(v, v)) (GLM func_geometric.inl:95-104, func_exponential.inl: 1 / sqrt)
    const float inv = __fdiv_rn(1.0f, __fsqrt_rn(add(add(mul(n[0], n[0]), mul(n[1], n[1])), mul(n[2], n[2]))));
    a.buffer_a[p] = make_float4(__fdiv_rn(add(mul(n[0], inv), 1.0f), 2.0f), __fdiv_rn(add(mul(n[1], inv), 1.0f), 2.0f),
                                __fdiv_rn(add(mul(n[2], inv), 1.0f), 2.0f), 1.0f);
    a.buffer_b[p] = make_float4(diffuse[0], diffuse[1], diffuse[2], 1.0f);
    a.buffer_c[p] = (m.flags & OVIS_MATERIAL_SPECULAR) ? make_float4(m.specular[0], m.specular[1], m.specular[2], m.shininess)
                                                       : make_float4(0.f, 0.f, 0.f, 1.0f);
    a.depth[p] = __uint_as_float((unsigned int)(key >> 32));
}

}  // namespace

// ---------------------------------------------------------------------------------------------------------------------
// Host side
// ---------------------------------------------------------------------------------------------------------------------

struct OvisRaster {
    OvisWorld *world = nullptr;
    int device = 0;
    cudaStream_t stream = nullptr;
    uint32_t id_base = 0;
    // geometry and bindings (host mirrors are only what the API needs to validate and append)
    std::vector<RsMeshRef> meshes;
    int64_t n_vertices = 0, n_indices = 0, n_materials = 0;
    float *d_vertices = nullptr;
    int64_t cap_vertices = 0;
    uint32_t *d_indices = nullptr;
    int64_t cap_indices = 0;
    RsMeshRef *d_meshes = nullptr;
    int64_t cap_meshes = 0;
    OvisMaterial *d_materials = nullptr;
    int64_t cap_materials = 0;
    float *d_tangents = nullptr;                    // 4 floats per vertex, parallel to d_vertices
    int64_t cap_tangents = 0;
    std::vector<RsTexture> textures;                // host mirror of the table (texel pointers are device pointers)
    RsTexture *d_textures = nullptr;
    int64_t cap_textures = 0;
    RsTexAttribs *d_tex_attribs = nullptr;
    int64_t cap_tex_attribs = 0;
    uint32_t *d_entity_mesh = nullptr, *d_entity_material = nullptr;
    int64_t n_bound = 0;                            // entity slots the two tables cover
    // scratch of a call
    RsView *d_views = nullptr;
    int64_t cap_views = 0;
    float *d_uniforms = nullptr;
    int64_t cap_uniforms = 0;
    long long *d_first_triangle = nullptr;
    int64_t cap_first = 0;
    RsTriangle *d_triangles = nullptr;
    int64_t cap_triangles = 0;
    RsNormals *d_normals = nullptr;
    int64_t cap_normals = 0;
    RsItem *d_items = nullptr;
    int64_t cap_items = 0;
    unsigned int *d_counters = nullptr;             // [0] items, [1] overflow, [2] instance count of the call
    unsigned long long *d_vis = nullptr;
    int64_t cap_vis = 0;
    bool want_normals = true;                       // false while the Basic pass draws (it shades with the material colour alone)
    uint32_t *d_dense_ids = nullptr;                // entity ids of a strided (draw-list) input, densified
    int64_t cap_dense_ids = 0;
};

namespace {

__global__ void fill_words_kernel(unsigned int *dst, long long n, unsigned int value) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        dst[i] = value;
}

// later entries win when an entity is named twice in one call only by accident of scheduling: callers name each once
__global__ void bind_scatter_kernel(const unsigned int *stage, long long n, unsigned int *entity_mesh, unsigned int *entity_material) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        entity_mesh[stage[i]] = stage[n + i];
        entity_material[stage[i]] = stage[2 * n + i];
    }
}

// a draw list is (entity, pass index) pairs (ovis_build_draw_lists_device): every `stride`-th word is an entity id
__global__ void gather_ids_kernel(const unsigned int *src, int stride, long long n, unsigned int *dst) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        dst[i] = src[i * stride];
}

// Basic pass: the visibility buffer starts from the depth buffer the G-buffer pass left (no triangle: low word all ones)
__global__ void vis_from_depth_kernel(const float *depth, long long n, unsigned long long *vis) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        vis[i] = ((unsigned long long)__float_as_uint(depth[i]) << 32) | 0xffffffffull;
}

struct RsBasicArgs {
    const unsigned long long *vis;
    const RsTriangle *triangles;
    const unsigned int *instance_entity;
    const unsigned int *entity_material;
    const OvisMaterial *materials;
    unsigned int id_base;
    long long pixels;
    float4 *rgba;
    float *depth;
};

// unlit_frag.glsl without TEXTURED: fragColour = vec4(diffuseColour, 1.0), blending off, for the pixels a triangle of this
// pass won (depth test LessOrEqual against what the G-buffer pass left)
__global__ void __launch_bounds__(kRasterThreads)
raster_resolve_basic_kernel(const RsBasicArgs a) {
    const long long p = (long long)blockIdx.x * kRasterThreads + threadIdx.x;
    if (p >= a.pixels)
        return;
    const unsigned long long key = a.vis[p];
    if ((unsigned int)key == 0xffffffffu)
        return;
    const RsTriangle &t = a.triangles[~(unsigned int)key];
    const OvisMaterial m = a.materials[a.entity_material[a.instance_entity[t.instance] - a.id_base]];
    a.rgba[p] = make_float4(m.diffuse[0], m.diffuse[1], m.diffuse[2], 1.0f);
    a.depth[p] = __uint_as_float((unsigned int)(key >> 32));
}

int fill_words(unsigned int *dst, int64_t n, unsigned int value, cudaStream_t s) {
    if (n <= 0)
        return OVIS_OK;
    fill_words_kernel<<<(unsigned)std::min<int64_t>((n + 255) / 256, 1184), 256, 0, s>>>(dst, (long long)n, value);
    RS_CUDA(cudaGetLastError());
    return OVIS_OK;
}

template <typename T>
int ensure_device(T *&ptr, int64_t &cap, int64_t want, int64_t keep, cudaStream_t s) {
    if (want <= cap)
        return OVIS_OK;
    const int64_t fresh = std::max<int64_t>(want, cap + cap / 2);
    T *p = nullptr;
    RS_CUDA(cudaMalloc(&p, sizeof(T) * (size_t)fresh));
    if (ptr && keep > 0)
        RS_CUDA(cudaMemcpyAsync(p, ptr, sizeof(T) * (size_t)keep, cudaMemcpyDeviceToDevice, s));
    if (ptr) {
        RS_CUDA(cudaStreamSynchronize(s));
        RS_CUDA(cudaFree(ptr));
    }
    ptr = p;
    cap = fresh;
    return OVIS_OK;
}

int use(OvisRaster *r) {
    RS_CUDA(cudaSetDevice(r->device));
    void *s = nullptr;
    RS_TRY(ovis_world_get_stream(r->world, &s, nullptr));      // the world's stream may have been replaced since
    r->stream = (cudaStream_t)s;
    RS_TRY(ovis_world_get_id_base(r->world, &r->id_base));
    return OVIS_OK;
}

// The kernels of one call over `n_instances` instances spread over the views in h_views; item and triangle buffers are
// enlarged and the call repeated when they overflow.
int draw(OvisRaster *r, std::vector<RsView> &h_views, int64_t n_instances, const uint32_t *d_instance_entity) {
    if (n_instances == 0 || r->meshes.empty())
        return OVIS_OK;
    cudaStream_t s = r->stream;
    RS_TRY(ensure_device(r->d_views, r->cap_views, (int64_t)h_views.size(), 0, s));
    RS_TRY(ensure_device(r->d_uniforms, r->cap_uniforms, 20 * n_instances, 0, s));
    RS_TRY(ensure_device(r->d_first_triangle, r->cap_first, n_instances + 1, 0, s));
    if (!r->d_counters)
        RS_CUDA(cudaMalloc(&r->d_counters, 4 * sizeof(unsigned int)));
    // EntityUniforms of the instances, in list order (render_entity.cc:64-72): transform = Transform::matrix()
    const unsigned int count_word = (unsigned int)n_instances;
    RS_CUDA(cudaMemcpyAsync(r->d_counters + 2, &count_word, sizeof count_word, cudaMemcpyHostToDevice, s));
    RS_TRY(ovis_pack_entity_uniforms_device(r->world, d_instance_entity, r->d_counters + 2, n_instances, r->d_uniforms, n_instances));
    RS_CUDA(cudaMemcpyAsync(r->d_views, h_views.data(), sizeof(RsView) * h_views.size(), cudaMemcpyHostToDevice, s));
    RsArgs a{};
    a.views = r->d_views;
    a.n_views = (int)h_views.size();
    a.n_instances = n_instances;
    a.instance_entity = d_instance_entity;
    a.instance_uniforms = r->d_uniforms;
    a.entity_mesh = r->d_entity_mesh;
    a.n_entities_bound = r->n_bound;
    a.id_base = r->id_base;
    a.meshes = r->d_meshes;
    a.vertices = r->d_vertices;
    a.indices = r->d_indices;
    a.first_triangle = r->d_first_triangle;
    a.counters = r->d_counters;
    raster_scan_kernel<<<1, 1024, 0, s>>>(a);
    long long total_triangles = 0;
    RS_CUDA(cudaMemcpyAsync(&total_triangles, r->d_first_triangle + n_instances, sizeof total_triangles, cudaMemcpyDeviceToHost, s));
    RS_CUDA(cudaStreamSynchronize(s));
    if (total_triangles == 0)
        return OVIS_OK;
    if (total_triangles > 0xfffffffell)
        return fail(OVIS_ERR_CAPACITY, "ovis_raster: more than 2^32 triangles in one call");
    RS_TRY(ensure_device(r->d_triangles, r->cap_triangles, total_triangles, 0, s));
    const bool shaded = h_views[0].vis != nullptr && r->want_normals;
    const bool textured = shaded && !r->textures.empty();
    if (shaded)
        RS_TRY(ensure_device(r->d_normals, r->cap_normals, total_triangles, 0, s));
    if (textured)
        RS_TRY(ensure_device(r->d_tex_attribs, r->cap_tex_attribs, total_triangles, 0, s));
    if (r->cap_items == 0)
        RS_TRY(ensure_device(r->d_items, r->cap_items, (int64_t)1 << 20, 0, s));
    int sms = 0;
    RS_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, r->device));
    for (int attempt = 0; attempt < 6; attempt++) {
        a.triangles = r->d_triangles;
        a.normals = shaded ? r->d_normals : nullptr;
        a.tex_attribs = textured ? r->d_tex_attribs : nullptr;
        a.tangents = r->d_tangents;
        a.triangle_capacity = r->cap_triangles;
        a.items = r->d_items;
        a.item_capacity = (unsigned int)std::min<int64_t>(r->cap_items, 0xffffffffll);
        RS_CUDA(cudaMemsetAsync(r->d_counters, 0, 2 * sizeof(unsigned int), s));
        raster_setup_kernel<<<(unsigned)n_instances, kRasterThreads, 0, s>>>(a);
        unsigned int counters[2] = { 0, 0 };
        RS_CUDA(cudaMemcpyAsync(counters, r->d_counters, sizeof counters, cudaMemcpyDeviceToHost, s));
        RS_CUDA(cudaStreamSynchronize(s));
        if (counters[1])
            return fail(OVIS_ERR_CUDA, "ovis_raster: triangle buffer overflow (internal error)");
        if ((int64_t)counters[0] > r->cap_items) {             // more tiles than the item list holds: enlarge it, set up again
            RS_TRY(ensure_device(r->d_items, r->cap_items, (int64_t)counters[0] + counters[0] / 4, 0, s));
            continue;
        }
        if (counters[0])
            raster_fill_kernel<<<(unsigned)(sms * 8), kRasterThreads, 0, s>>>(a);
        RS_CUDA(cudaGetLastError());
        return OVIS_OK;
    }
    return fail(OVIS_ERR_NOMEM, "ovis_raster: the work-item list did not fit after regrowing");
}

}  // namespace

extern "C" {

int ovis_raster_create(OvisWorld *world, OvisRaster **out_raster) {
    if (!world || !out_raster)
        return fail(OVIS_ERR_INVALID, "ovis_raster_create: null argument");
    *out_raster = nullptr;
    OvisRaster *r = new (std::nothrow) OvisRaster;
    if (!r)
        return fail(OVIS_ERR_NOMEM, "out of host memory");
    r->world = world;
    void *s = nullptr;
    int32_t device = 0;
    int rc = ovis_world_get_stream(world, &s, &device);
    if (rc != OVIS_OK) {
        delete r;
        return rc;
    }
    r->device = device;
    r->stream = (cudaStream_t)s;
    *out_raster = r;
    return OVIS_OK;
}

int ovis_raster_destroy(OvisRaster *r) {
    if (!r)
        return OVIS_OK;
    cudaSetDevice(r->device);
    cudaDeviceSynchronize();
    void *bufs[] = { r->d_vertices, r->d_indices, r->d_meshes, r->d_materials, r->d_entity_mesh, r->d_entity_material, r->d_views,
                     r->d_uniforms, r->d_first_triangle, r->d_triangles, r->d_normals, r->d_items, r->d_counters, r->d_vis, r->d_dense_ids, r->d_tangents, r->d_textures, r->d_tex_attribs };
    for (const RsTexture &t : r->textures)
        cudaFree(const_cast<float4 *>(t.texels));
    for (void *b : bufs)
        if (b)
            cudaFree(b);
    delete r;
    return OVIS_OK;
}

int ovis_raster_add_mesh(OvisRaster *r, int64_t n_vertices, const float *positions, const float *normals, const float *texcoords,
                         int64_t n_indices, const uint32_t *indices, uint32_t *out_mesh) {
    if (!r || n_vertices < 1 || !positions || !normals || n_indices < 0 || n_indices % 3 != 0 || (n_indices > 0 && !indices) ||
        n_indices / 3 > 0x7fffffff)
        return fail(OVIS_ERR_INVALID, "ovis_raster_add_mesh: bad arguments (positions and normals are required, indices come in threes)");
    for (int64_t i = 0; i < n_indices; i++)
        if ((int64_t)indices[i] >= n_vertices)
            return fail(OVIS_ERR_INVALID, "ovis_raster_add_mesh: index " + std::to_string(i) + " is out of range");
    RS_TRY(use(r));
    std::vector<float> packed((size_t)n_vertices * 8);
    for (int64_t v = 0; v < n_vertices; v++) {
        float *o = packed.data() + 8 * v;
        o[0] = positions[3 * v]; o[1] = positions[3 * v + 1]; o[2] = positions[3 * v + 2];
        o[3] = normals[3 * v]; o[4] = normals[3 * v + 1]; o[5] = normals[3 * v + 2];
        o[6] = texcoords ? texcoords[2 * v] : 0.0f;
        o[7] = texcoords ? texcoords[2 * v + 1] : 0.0f;
    }
    RS_TRY(ensure_device(r->d_vertices, r->cap_vertices, 8 * (r->n_vertices + n_vertices), 8 * r->n_vertices, r->stream));
    RS_TRY(ensure_device(r->d_indices, r->cap_indices, std::max<int64_t>(r->n_indices + n_indices, 1), r->n_indices, r->stream));
    RS_TRY(ensure_device(r->d_tangents, r->cap_tangents, 4 * (r->n_vertices + n_vertices), 4 * r->n_vertices, r->stream));
    RS_CUDA(cudaMemsetAsync(r->d_tangents + 4 * r->n_vertices, 0, sizeof(float) * 4 * (size_t)n_vertices, r->stream));
    RS_TRY(ensure_device(r->d_meshes, r->cap_meshes, (int64_t)r->meshes.size() + 1, (int64_t)r->meshes.size(), r->stream));
    RsMeshRef ref{ r->n_vertices, r->n_indices, (int)(n_indices / 3), 0 };
    RS_CUDA(cudaMemcpyAsync(r->d_vertices + 8 * r->n_vertices, packed.data(), sizeof(float) * packed.size(), cudaMemcpyHostToDevice, r->stream));
    if (n_indices)
        RS_CUDA(cudaMemcpyAsync(r->d_indices + r->n_indices, indices, sizeof(uint32_t) * (size_t)n_indices, cudaMemcpyHostToDevice, r->stream));
    RS_CUDA(cudaMemcpyAsync(r->d_meshes + r->meshes.size(), &ref, sizeof ref, cudaMemcpyHostToDevice, r->stream));
    RS_CUDA(cudaStreamSynchronize(r->stream));                 // (the staging vector goes out of scope)
    if (out_mesh)
        *out_mesh = (uint32_t)r->meshes.size();
    r->meshes.push_back(ref);
    r->n_vertices += n_vertices;
    r->n_indices += n_indices;
    return OVIS_OK;
}

int ovis_raster_set_mesh_tangents(OvisRaster *r, uint32_t mesh, int64_t n_vertices, const float *tangents) {
    if (!r || mesh >= r->meshes.size() || !tangents)
        return fail(OVIS_ERR_INVALID, "ovis_raster_set_mesh_tangents: unknown mesh or null tangents");
    const int64_t first = r->meshes[mesh].first_vertex;
    const int64_t count = (mesh + 1 < r->meshes.size() ? r->meshes[mesh + 1].first_vertex : r->n_vertices) - first;
    if (n_vertices != count)
        return fail(OVIS_ERR_INVALID, "ovis_raster_set_mesh_tangents: the mesh has " + std::to_string(count) + " vertices");
    RS_TRY(use(r));
    RS_CUDA(cudaMemcpyAsync(r->d_tangents + 4 * first, tangents, sizeof(float) * 4 * (size_t)n_vertices, cudaMemcpyHostToDevice, r->stream));
    RS_CUDA(cudaStreamSynchronize(r->stream));
    return OVIS_OK;
}

int ovis_raster_add_texture(OvisRaster *r, int32_t width, int32_t height, const uint8_t *rgba, int32_t srgb, int32_t wrap, uint32_t *out_texture) {
    if (!r || width < 1 || height < 1 || width > 32768 || height > 32768 || !rgba)
        return fail(OVIS_ERR_INVALID, "ovis_raster_add_texture: bad arguments (1 .. 32768 texels a side)");
    RS_TRY(use(r));
    // texels are decoded once, on the host: c / 255, and for the colour channels of an sRGB texture the transfer function of
    // the GL / Vulkan specifications (c <= 0.04045 ? c / 12.92 : ((c + 0.055) / 1.055) ^ 2.4) - a 256-entry table
    float linear[256], decoded[256];
    for (int i = 0; i < 256; i++) {
        linear[i] = (float)i / 255.0f;
        decoded[i] = linear[i] <= 0.04045f ? linear[i] / 12.92f : powf((linear[i] + 0.055f) / 1.055f, 2.4f);
    }
    const size_t texels = (size_t)width * (size_t)height;
    std::vector<float> host(4 * texels);
    for (size_t i = 0; i < texels; i++) {
        const float *table = srgb ? decoded : linear;
        host[4 * i] = table[rgba[4 * i]];
        host[4 * i + 1] = table[rgba[4 * i + 1]];
        host[4 * i + 2] = table[rgba[4 * i + 2]];
        host[4 * i + 3] = linear[rgba[4 * i + 3]];
    }
    float4 *d_texels = nullptr;
    RS_CUDA(cudaMalloc(&d_texels, sizeof(float4) * texels));
    cudaError_t err = cudaMemcpyAsync(d_texels, host.data(), sizeof(float4) * texels, cudaMemcpyHostToDevice, r->stream);
    if (err == cudaSuccess)
        err = cudaStreamSynchronize(r->stream);
    if (err != cudaSuccess) {
        cudaFree(d_texels);
        return fail(OVIS_ERR_CUDA, std::string("ovis_raster_add_texture: ") + cudaGetErrorString(err));
    }
    RsTexture t{ d_texels, width, height, wrap ? 1 : 0, 0 };
    RS_TRY(ensure_device(r->d_textures, r->cap_textures, (int64_t)r->textures.size() + 1, (int64_t)r->textures.size(), r->stream));
    RS_CUDA(cudaMemcpyAsync(r->d_textures + r->textures.size(), &t, sizeof t, cudaMemcpyHostToDevice, r->stream));
    RS_CUDA(cudaStreamSynchronize(r->stream));
    if (out_texture)
        *out_texture = (uint32_t)r->textures.size();
    r->textures.push_back(t);
    return OVIS_OK;
}

int ovis_raster_add_material(OvisRaster *r, const OvisMaterial *material, uint32_t *out_material) {
    if (!r || !material)
        return fail(OVIS_ERR_INVALID, "ovis_raster_add_material: null argument");
    if (((material->flags & OVIS_MATERIAL_TEXTURED) && material->diffuse_texture >= r->textures.size()) ||
        ((material->flags & OVIS_MATERIAL_NORMALMAP) && material->normal_texture >= r->textures.size()))
        return fail(OVIS_ERR_INVALID, "ovis_raster_add_material: the material names a texture that was not added");
    RS_TRY(use(r));
    RS_TRY(ensure_device(r->d_materials, r->cap_materials, r->n_materials + 1, r->n_materials, r->stream));
    RS_CUDA(cudaMemcpyAsync(r->d_materials + r->n_materials, material, sizeof *material, cudaMemcpyHostToDevice, r->stream));
    RS_CUDA(cudaStreamSynchronize(r->stream));
    if (out_material)
        *out_material = (uint32_t)r->n_materials;
    r->n_materials++;
    return OVIS_OK;
}

int ovis_raster_bind_entities(OvisRaster *r, int64_t n, const uint32_t *entity_ids, const uint32_t *mesh_ids, const uint32_t *material_ids) {
    if (!r || n < 0 || (n > 0 && (!entity_ids || !mesh_ids || !material_ids)))
        return fail(OVIS_ERR_INVALID, "ovis_raster_bind_entities: bad arguments");
    RS_TRY(use(r));
    int64_t n_entities = 0;
    RS_TRY(ovis_world_counts(r->world, &n_entities, nullptr, nullptr));
    for (int64_t i = 0; i < n; i++) {
        if ((int64_t)(entity_ids[i] - r->id_base) >= n_entities)
            return fail(OVIS_ERR_INVALID, "ovis_raster_bind_entities: unknown entity " + std::to_string(entity_ids[i]));
        if (mesh_ids[i] >= r->meshes.size() || (int64_t)material_ids[i] >= r->n_materials)
            return fail(OVIS_ERR_INVALID, "ovis_raster_bind_entities: unknown mesh or material for entity " + std::to_string(entity_ids[i]));
    }
    if (n_entities > r->n_bound) {                             // the tables follow the world's entity count; new slots are unbound
        uint32_t *mesh = nullptr, *mat = nullptr;
        RS_CUDA(cudaMalloc(&mesh, sizeof(uint32_t) * (size_t)n_entities));
        RS_CUDA(cudaMalloc(&mat, sizeof(uint32_t) * (size_t)n_entities));
        RS_CUDA(cudaMemsetAsync(mesh, 0xff, sizeof(uint32_t) * (size_t)n_entities, r->stream));
        RS_CUDA(cudaMemsetAsync(mat, 0, sizeof(uint32_t) * (size_t)n_entities, r->stream));
        if (r->n_bound) {
            RS_CUDA(cudaMemcpyAsync(mesh, r->d_entity_mesh, sizeof(uint32_t) * (size_t)r->n_bound, cudaMemcpyDeviceToDevice, r->stream));
            RS_CUDA(cudaMemcpyAsync(mat, r->d_entity_material, sizeof(uint32_t) * (size_t)r->n_bound, cudaMemcpyDeviceToDevice, r->stream));
        }
        RS_CUDA(cudaStreamSynchronize(r->stream));
        cudaFree(r->d_entity_mesh);
        cudaFree(r->d_entity_material);
        r->d_entity_mesh = mesh;
        r->d_entity_material = mat;
        r->n_bound = n_entities;
    }
    if (n == 0)
        return OVIS_OK;
    // ids, meshes and materials go up in one staging block; a small kernel scatters them into the two tables
    uint32_t *d_stage = nullptr;
    RS_CUDA(cudaMalloc(&d_stage, sizeof(uint32_t) * 3 * (size_t)n));
    std::vector<uint32_t> h_stage(3 * (size_t)n);
    for (int64_t i = 0; i < n; i++) {
        h_stage[(size_t)i] = entity_ids[i] - r->id_base;
        h_stage[(size_t)(n + i)] = mesh_ids[i];
        h_stage[(size_t)(2 * n + i)] = material_ids[i];
    }
    cudaError_t err = cudaMemcpyAsync(d_stage, h_stage.data(), sizeof(uint32_t) * h_stage.size(), cudaMemcpyHostToDevice, r->stream);
    if (err == cudaSuccess) {
        bind_scatter_kernel<<<(unsigned)std::min<int64_t>((n + 255) / 256, 65535), 256, 0, r->stream>>>(d_stage, (long long)n, r->d_entity_mesh, r->d_entity_material);
        err = cudaStreamSynchronize(r->stream);
    }
    cudaFree(d_stage);
    if (err != cudaSuccess)
        return fail(OVIS_ERR_CUDA, std::string("ovis_raster_bind_entities: ") + cudaGetErrorString(err));
    return OVIS_OK;
}

}  // extern "C"

namespace {

// The instance list of a single-view pass: `count` ids read from every `stride`-th word of d_ids (1 = a visible list,
// 2 = a draw list of (entity, pass) pairs); returns a dense device list.
int instance_list(OvisRaster *r, const uint32_t *d_ids, int stride, const uint32_t *d_count, int64_t max_count, const uint32_t **out_dense, int64_t *out_n) {
    unsigned int count = 0;
    RS_CUDA(cudaMemcpyAsync(&count, d_count, sizeof count, cudaMemcpyDeviceToHost, r->stream));
    RS_CUDA(cudaStreamSynchronize(r->stream));
    int64_t n = count;
    if (max_count >= 0)
        n = std::min<int64_t>(n, max_count);
    *out_n = n;
    *out_dense = d_ids;
    if (stride != 1 && n > 0) {
        RS_TRY(ensure_device(r->d_dense_ids, r->cap_dense_ids, n, 0, r->stream));
        gather_ids_kernel<<<(unsigned)std::min<int64_t>((n + 255) / 256, 1184), 256, 0, r->stream>>>(d_ids, stride, (long long)n, r->d_dense_ids);
        RS_CUDA(cudaGetLastError());
        *out_dense = r->d_dense_ids;
    }
    return OVIS_OK;
}

int check_view_and_size(const char *who, const OvisView *view, int32_t width, int32_t height) {
    if (width < 1 || height < 1 || width > 65535 || height > 65535)
        return fail(OVIS_ERR_INVALID, std::string(who) + ": the target must be 1 .. 65535 pixels a side");
    if (view->vp_width != width || view->vp_height != height)
        return fail(OVIS_ERR_INVALID, std::string(who) + ": the view's viewport is not the target's size");
    return OVIS_OK;
}

int one_view(const OvisView *view, int32_t width, int32_t height, int64_t n, unsigned long long *vis, std::vector<RsView> &views) {
    float mats[64];
    RS_TRY(ovis_view_matrices(view, mats, nullptr));
    views.resize(1);
    std::memcpy(views[0].vp, mats + 32, sizeof views[0].vp);    // viewProjection
    views[0].width = width;
    views[0].height = height;
    views[0].first_instance = 0;
    views[0].n_instances = n;
    views[0].vis = vis;
    views[0].depth_bits = nullptr;
    return OVIS_OK;
}

int gbuffer_pass(OvisRaster *r, const OvisView *view, const uint32_t *d_ids, int stride, const uint32_t *d_count, int64_t max_count,
                 const OvisGBufferTarget *target) {
    if (!r || !view || !d_ids || !d_count || !target || !target->d_buffer_a || !target->d_buffer_b || !target->d_buffer_c || !target->d_depth)
        return fail(OVIS_ERR_INVALID, "ovis_raster_gbuffer: null argument");
    if ((((uintptr_t)target->d_buffer_a | (uintptr_t)target->d_buffer_b | (uintptr_t)target->d_buffer_c) & 15u) || ((uintptr_t)target->d_depth & 3u))
        return fail(OVIS_ERR_INVALID, "ovis_raster_gbuffer: the colour buffers must be 16-byte aligned");
    RS_TRY(check_view_and_size("ovis_raster_gbuffer", view, target->width, target->height));
    RS_TRY(use(r));
    const uint32_t *d_list = nullptr;
    int64_t n = 0;
    RS_TRY(instance_list(r, d_ids, stride, d_count, max_count, &d_list, &n));
    const int64_t pixels = (int64_t)target->width * target->height;
    RS_TRY(ensure_device(r->d_vis, r->cap_vis, pixels, 0, r->stream));
    RS_CUDA(cudaMemsetAsync(r->d_vis, 0xff, sizeof(unsigned long long) * (size_t)pixels, r->stream));
    std::vector<RsView> views;
    RS_TRY(one_view(view, target->width, target->height, n, r->d_vis, views));
    RS_TRY(draw(r, views, n, d_list));
    RsResolveArgs ra{};
    ra.vis = r->d_vis;
    ra.triangles = r->d_triangles;
    ra.normals = r->d_normals;
    ra.tex_attribs = r->textures.empty() ? nullptr : r->d_tex_attribs;
    ra.textures = r->d_textures;
    ra.instance_entity = d_list;
    ra.entity_material = r->d_entity_material;
    ra.materials = r->d_materials;
    ra.id_base = r->id_base;
    ra.width = target->width;
    ra.height = target->height;
    ra.buffer_a = reinterpret_cast<float4 *>(target->d_buffer_a);
    ra.buffer_b = reinterpret_cast<float4 *>(target->d_buffer_b);
    ra.buffer_c = reinterpret_cast<float4 *>(target->d_buffer_c);
    ra.depth = target->d_depth;
    raster_resolve_kernel<<<(unsigned)((pixels + kRasterThreads - 1) / kRasterThreads), kRasterThreads, 0, r->stream>>>(ra);
    RS_CUDA(cudaGetLastError());
    return OVIS_OK;
}

int basic_pass(OvisRaster *r, const OvisView *view, const uint32_t *d_ids, int stride, const uint32_t *d_count, int64_t max_count,
               int32_t width, int32_t height, float *d_rgba, float *d_depth) {
    if (!r || !view || !d_ids || !d_count || !d_rgba || !d_depth)
        return fail(OVIS_ERR_INVALID, "ovis_raster_basic: null argument");
    if (((uintptr_t)d_rgba & 15u) || ((uintptr_t)d_depth & 3u))
        return fail(OVIS_ERR_INVALID, "ovis_raster_basic: the colour buffer must be 16-byte aligned");
    RS_TRY(check_view_and_size("ovis_raster_basic", view, width, height));
    RS_TRY(use(r));
    const uint32_t *d_list = nullptr;
    int64_t n = 0;
    RS_TRY(instance_list(r, d_ids, stride, d_count, max_count, &d_list, &n));
    if (n == 0)
        return OVIS_OK;
    const int64_t pixels = (int64_t)width * height;
    RS_TRY(ensure_device(r->d_vis, r->cap_vis, pixels, 0, r->stream));
    vis_from_depth_kernel<<<(unsigned)std::min<int64_t>((pixels + 255) / 256, 4736), 256, 0, r->stream>>>(d_depth, (long long)pixels, r->d_vis);
    std::vector<RsView> views;
    RS_TRY(one_view(view, width, height, n, r->d_vis, views));
    r->want_normals = false;
    const int rc_draw = draw(r, views, n, d_list);
    r->want_normals = true;
    RS_TRY(rc_draw);
    RsBasicArgs ba{};
    ba.vis = r->d_vis;
    ba.triangles = r->d_triangles;
    ba.instance_entity = d_list;
    ba.entity_material = r->d_entity_material;
    ba.materials = r->d_materials;
    ba.id_base = r->id_base;
    ba.pixels = pixels;
    ba.rgba = reinterpret_cast<float4 *>(d_rgba);
    ba.depth = d_depth;
    raster_resolve_basic_kernel<<<(unsigned)((pixels + kRasterThreads - 1) / kRasterThreads), kRasterThreads, 0, r->stream>>>(ba);
    RS_CUDA(cudaGetLastError());
    return OVIS_OK;
}

}  // namespace

extern "C" {

int ovis_raster_gbuffer_device(OvisRaster *r, const OvisView *view, const uint32_t *d_visible, const uint32_t *d_visible_count,
                               int64_t max_visible, const OvisGBufferTarget *target) {
    return gbuffer_pass(r, view, d_visible, 1, d_visible_count, max_visible, target);
}

int ovis_raster_gbuffer_draws_device(OvisRaster *r, const OvisView *view, const uint32_t *d_draws, const uint32_t *d_draw_count,
                                     int64_t max_draws, const OvisGBufferTarget *target) {
    return gbuffer_pass(r, view, d_draws, 2, d_draw_count, max_draws, target);
}

int ovis_raster_basic_device(OvisRaster *r, const OvisView *view, const uint32_t *d_visible, const uint32_t *d_visible_count,
                             int64_t max_visible, int32_t width, int32_t height, float *d_rgba, float *d_depth) {
    return basic_pass(r, view, d_visible, 1, d_visible_count, max_visible, width, height, d_rgba, d_depth);
}

int ovis_raster_basic_draws_device(OvisRaster *r, const OvisView *view, const uint32_t *d_draws, const uint32_t *d_draw_count,
                                   int64_t max_draws, int32_t width, int32_t height, float *d_rgba, float *d_depth) {
    return basic_pass(r, view, d_draws, 2, d_draw_count, max_draws, width, height, d_rgba, d_depth);
}

int ovis_raster_shadow_maps_device(OvisRaster *r, int32_t n_views, const OvisView *views, const int64_t *host_offsets,
                                   const uint32_t *d_entities, int32_t shadow_size, float *const *d_maps) {
    if (!r || n_views < 0 || (n_views > 0 && (!views || !host_offsets || !d_maps)) || shadow_size < 1 || shadow_size > 65535)
        return fail(OVIS_ERR_INVALID, "ovis_raster_shadow_maps_device: bad arguments");
    if (n_views == 0)
        return OVIS_OK;
    RS_TRY(use(r));
    const int64_t total = host_offsets[n_views] - host_offsets[0];
    if (total < 0 || (total > 0 && !d_entities))
        return fail(OVIS_ERR_INVALID, "ovis_raster_shadow_maps_device: bad offsets or null entity list");
    std::vector<RsView> h((size_t)n_views);
    const float one = 1.0f;
    unsigned int one_bits;
    std::memcpy(&one_bits, &one, sizeof one_bits);
    for (int32_t k = 0; k < n_views; k++) {
        if (!d_maps[k] || host_offsets[k + 1] < host_offsets[k])
            return fail(OVIS_ERR_INVALID, "ovis_raster_shadow_maps_device: view " + std::to_string(k) + ": null map or descending offsets");
        float mats[64];
        RS_TRY(ovis_view_matrices(&views[k], mats, nullptr));
        std::memcpy(h[(size_t)k].vp, mats + 32, sizeof h[(size_t)k].vp);
        h[(size_t)k].width = h[(size_t)k].height = shadow_size;
        h[(size_t)k].first_instance = host_offsets[k] - host_offsets[0];
        h[(size_t)k].n_instances = host_offsets[k + 1] - host_offsets[k];
        h[(size_t)k].vis = nullptr;
        h[(size_t)k].depth_bits = reinterpret_cast<unsigned int *>(d_maps[k]);
    }
    for (int32_t k = 0; k < n_views; k++)                       // clear depth 1 (deferred_render_pipeline.cc:354)
        RS_TRY(fill_words(reinterpret_cast<unsigned int *>(d_maps[k]), (int64_t)shadow_size * shadow_size, one_bits, r->stream));
    return draw(r, h, total, d_entities + host_offsets[0]);
}

}  // extern "C"
