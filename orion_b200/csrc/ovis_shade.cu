// ovis_shade.cu - the deferred light pass and the colour post effects over headless buffers.
//
//   shade_deferred_kernel  engine/shaders/deferred_light_frag.glsl:30-81 (G-buffer decode, position from depth) and
//                          engine/shaders/lighting.h:46-211 (3x3 PCF shadows, Blinn-Phong, attenuation, spot cone) for
//                          EVERY visible light of a pixel in one pass: the reference rasterises one additive pass per
//                          light (render/src/deferred_render_pipeline.cc:424-470), re-reading the G-buffer and
//                          read-modify-writing the HDR target each time; here a pixel's G-buffer sample is decoded once,
//                          the lights' contributions are summed in registers in list order (the order the blend unit
//                          would add them) and the pixel is written once.
//   post_effects_kernel    tonemap_effect_frag.glsl:30-71 (Uncharted 2) and gamma_correction_effect_frag.glsl:26-31,
//                          optionally fused (the reference runs them as two full-screen passes).
//
// Arithmetic: fp32, operations in the shader's source order, IEEE division and square root, no FMA (the library is built
// with -fmad=false), GLM's association for mat4 * vec4 - i.e. the definition oracle/ovis_shading.c states, which is
// bit-identical to the reference's shader source compiled against GLM. The one deviation is powf (CUDA's is accurate to
// a few ulp, not correctly rounded): specular and gamma terms carry that, everything else is exact.
// The light loop is uniform across a CTA (every thread shades the same light at the same time), so the per-type branches
// of the reference's shader variants do not diverge; light blocks are staged through shared memory in chunks.
#include "ovis_kernels.h"

#include "orion_vis.h"

#include <cstdlib>

namespace ovis {

namespace {

struct F3 {
    float x, y, z;
};
__device__ __forceinline__ F3 f3(float x, float y, float z) { return { x, y, z }; }
__device__ __forceinline__ F3 operator+(F3 a, F3 b) { return { a.x + b.x, a.y + b.y, a.z + b.z }; }
__device__ __forceinline__ F3 operator-(F3 a, F3 b) { return { a.x - b.x, a.y - b.y, a.z - b.z }; }
__device__ __forceinline__ F3 operator*(F3 a, F3 b) { return { a.x * b.x, a.y * b.y, a.z * b.z }; }
__device__ __forceinline__ F3 operator*(F3 a, float s) { return { a.x * s, a.y * s, a.z * s }; }
__device__ __forceinline__ float dot3(F3 a, F3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ float length3(F3 a) { return sqrtf(dot3(a, a)); }
__device__ __forceinline__ F3 normalize3(F3 a) { return a * (1.0f / length3(a)); }
__device__ __forceinline__ float max_glsl(float a, float b) { return a > b ? a : b; }
__device__ __forceinline__ float clamp_glsl(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }

// (c0 v0 + c1 v1) + (c2 v2 + c3 v3), row r of a column-major mat4
__device__ __forceinline__ float mat_row(const float *m, int r, float v0, float v1, float v2, float v3) {
    return (m[0 + r] * v0 + m[4 + r] * v1) + (m[8 + r] * v2 + m[12 + r] * v3);
}

enum { L_POSITION = 0, L_INTENSITY = 3, L_DIRECTION = 4, L_COS_CUTOFF = 7, L_COLOUR = 8, L_RANGE = 11, L_SHADOW_SPACE = 28,
       L_SHADOW_ZNEAR = 44, L_ATT_CONSTANT = 45, L_ATT_LINEAR = 46, L_ATT_EXP = 47, L_SHADOW_BIAS = 48, L_FLOATS = 52 };
enum { V_INVERSE_VP = 48, V_VIEWPORT_POS = 64, V_VIEWPORT_SIZE = 66, V_POSITION = 68, V_FLOATS = 72 };

// texture(sampler2DShadow, vec3(u, v, ref)): bilinear, compare Less, clamp to edge
__device__ __forceinline__ float shadow_lookup_2d(const float *__restrict__ map, int size, float u, float v, float ref) {
    ref = clamp_glsl(ref, 0.0f, 1.0f);
    const float x = u * (float)size - 0.5f, y = v * (float)size - 0.5f;
    const float fx0 = floorf(x), fy0 = floorf(y);
    const float wx = x - fx0, wy = y - fy0;
    // clamp to edge, in float BEFORE the conversion: a coordinate beyond int's range (a light at 1e30, a projective divide
    // by ~0) converts to INT_MAX, and x0 + 1 is then signed overflow - which the compiler is free to, and did, turn into an
    // index outside the map. NaN takes the lower edge (fmaxf returns the other operand); its weights are NaN anyway.
    int x0 = (int)fminf(fmaxf(fx0, -1.0f), (float)size), y0 = (int)fminf(fmaxf(fy0, -1.0f), (float)size), x1 = x0 + 1, y1 = y0 + 1;
    x0 = min(max(x0, 0), size - 1); x1 = min(max(x1, 0), size - 1);
    y0 = min(max(y0, 0), size - 1); y1 = min(max(y1, 0), size - 1);
    const float c00 = ref < __ldg(map + (size_t)y0 * size + x0) ? 1.0f : 0.0f, c10 = ref < __ldg(map + (size_t)y0 * size + x1) ? 1.0f : 0.0f;
    const float c01 = ref < __ldg(map + (size_t)y1 * size + x0) ? 1.0f : 0.0f, c11 = ref < __ldg(map + (size_t)y1 * size + x1) ? 1.0f : 0.0f;
    const float top = c00 * (1.0f - wx) + c10 * wx, bottom = c01 * (1.0f - wx) + c11 * wx;
    return top * (1.0f - wy) + bottom * wy;
}

// texture(samplerCubeShadow, vec4(d, ref)): GL major-axis face selection, faces +x -x +y -y +z -z
__device__ __forceinline__ float shadow_lookup_cube(const float *__restrict__ map, int size, F3 d, float ref) {
    const float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
    float sc, tc, ma;
    int face;
    if (ax >= ay && ax >= az) { face = d.x >= 0.0f ? 0 : 1; ma = ax; sc = d.x >= 0.0f ? -d.z : d.z; tc = -d.y; }
    else if (ay >= az)         { face = d.y >= 0.0f ? 2 : 3; ma = ay; sc = d.x; tc = d.y >= 0.0f ? d.z : -d.z; }
    else                       { face = d.z >= 0.0f ? 4 : 5; ma = az; sc = d.z >= 0.0f ? d.x : -d.x; tc = -d.y; }
    const float s = (sc / ma + 1.0f) * 0.5f, t = (tc / ma + 1.0f) * 0.5f;
    return shadow_lookup_2d(map + (size_t)face * size * size, size, s, t, ref);
}

struct Lighting {
    F3 normal, position, diffuse, specular;
    float shininess;
};

// lighting.h:46-143
__device__ __forceinline__ float calc_shadow(const Lighting &d, const float *light, int type, const float *map, int size) {
    if (!map)
        return 1.0f;
    if (type == OVIS_LIGHT_SPOT) {
        const float *m = light + L_SHADOW_SPACE;
        const float sx = mat_row(m, 0, d.position.x, d.position.y, d.position.z, 1.0f);
        const float sy = mat_row(m, 1, d.position.x, d.position.y, d.position.z, 1.0f);
        const float sz = mat_row(m, 2, d.position.x, d.position.y, d.position.z, 1.0f);
        const float sw = mat_row(m, 3, d.position.x, d.position.y, d.position.z, 1.0f);
        const float u = sx / sw, v = sy / sw;
        float depth = sz / sw;
        depth += light[L_SHADOW_BIAS];
        float shadow = 0.0f;
        const float texel = 1.0f / (float)size;
        for (int x = -1; x <= 1; x++)
            for (int y = -1; y <= 1; y++)
                shadow += shadow_lookup_2d(map, size, u + (float)x * texel, v + (float)y * texel, depth);
        return shadow / 9.0f;
    }
    if (type == OVIS_LIGHT_POINT) {
        const F3 direction = d.position - f3(light[0], light[1], light[2]);
        const float ax = fabsf(direction.x), ay = fabsf(direction.y), az = fabsf(direction.z);
        float local_depth;
        F3 oa, ob;
        if (ax >= ay && ax >= az) { local_depth = ax; oa = f3(0, 1, 0); ob = f3(0, 0, 1); }
        else if (ay >= ax && ay >= az) { local_depth = ay; oa = f3(1, 0, 0); ob = f3(0, 0, 1); }
        else { local_depth = az; oa = f3(1, 0, 0); ob = f3(0, 1, 0); }
        const float far = light[L_RANGE], near = light[L_SHADOW_ZNEAR];
        const float depth = -(far / (near - far)) - ((far * near) / (far - near) / local_depth);
        const float texel = 1.0f / (float)size;
        oa = (oa * texel) * 2.0f;
        ob = (ob * texel) * 2.0f;
        const F3 nd = normalize3(direction);
        const float ref = depth + light[L_SHADOW_BIAS];
        float shadow = 0.0f;
        for (int a = -1; a <= 1; a++)
            for (int b = -1; b <= 1; b++)
                shadow += shadow_lookup_cube(map, size, (nd + oa * (float)a) + ob * (float)b, ref);
        return shadow / 9.0f;
    }
    return 1.0f;
}

// lighting.h:149-168
__device__ __forceinline__ F3 blinn_phong(const Lighting &d, F3 direction, const float *light, F3 to_view) {
    const F3 light_colour = f3(light[L_COLOUR], light[L_COLOUR + 1], light[L_COLOUR + 2]);
    const F3 to_light = f3(-direction.x, -direction.y, -direction.z);
    const float angle = max_glsl(dot3(d.normal, to_light), 0.0f);
    F3 colour = (d.diffuse * light_colour) * (light[L_INTENSITY] * angle);
    const F3 half_vector = normalize3(to_light + to_view);
    const float specular_angle = max_glsl(dot3(d.normal, half_vector), 0.0f);
    colour = colour + (d.specular * light_colour) * (light[L_INTENSITY] * powf(specular_angle, d.shininess));
    return colour;
}

// lighting.h:173-211. `may_skip`: light and pixel are known to be finite and in bounds (see kLightSkippable), so that an
// attenuation of exactly 0 means a contribution of exactly +-0 - which the caller's sum would not notice.
__device__ __forceinline__ bool calc_light(const Lighting &d, const float *light, int type, F3 to_view, const float *map, int size,
                                           bool may_skip, F3 &out) {
    if (type == OVIS_LIGHT_AMBIENT) {
        out = (d.diffuse * f3(light[L_COLOUR], light[L_COLOUR + 1], light[L_COLOUR + 2])) * light[L_INTENSITY];
        return true;
    }
    F3 direction;
    float attenuation = 1.0f;
    const F3 light_direction = f3(light[L_DIRECTION], light[L_DIRECTION + 1], light[L_DIRECTION + 2]);
    if (type == OVIS_LIGHT_DIRECTIONAL) {
        direction = light_direction;
    } else {
        const F3 to_fragment = d.position - f3(light[0], light[1], light[2]);
        const float distance = length3(to_fragment);
        attenuation = clamp_glsl(floorf(light[L_RANGE] / distance), 0.0f, 1.0f);
        if (may_skip && attenuation == 0.0f)
            return false;
        direction = to_fragment * (1.0f / distance);             // normalize(): the same sqrt, computed once
        if (type == OVIS_LIGHT_SPOT) {
            const float spot = max_glsl(dot3(direction, light_direction), light[L_COS_CUTOFF]);
            attenuation *= 1.0f - ((1.0f - spot) / (1.0f - light[L_COS_CUTOFF]));
        }
        attenuation /= light[L_ATT_CONSTANT] + (light[L_ATT_LINEAR] * distance) + (light[L_ATT_EXP] * distance * distance);
    }
    attenuation *= calc_shadow(d, light, type, map, size);
    out = blinn_phong(d, direction, light, to_view) * attenuation;
    return true;
}

constexpr int kShadeChunk = 16;    // lights staged per round: 16 x 208 B of shared memory
constexpr uint32_t kShadeTileRejectMinLights = 8;

template <bool TILED>
__global__ void __launch_bounds__(256)
shade_deferred_kernel(const __grid_constant__ ShadeArgs a) {
    __shared__ float s_light[kShadeChunk * L_FLOATS];
    __shared__ float s_view[V_FLOATS];
    __shared__ int s_type[kShadeChunk];
    __shared__ const float *s_map[kShadeChunk];
    const int tid = threadIdx.x;
    if (tid < V_FLOATS)
        s_view[tid] = a.view_uniforms[tid];
    __syncthreads();
    // a CTA shades a 16 x 16 pixel tile: neighbours on screen are neighbours in space, which is what makes the
    // tile-level light rejection below (and warp-uniform out-of-range skips) effective
    int px, py;
    bool active;
    if (TILED) {
        const int tiles_x = (a.width + 15) >> 4;
        px = (int)(blockIdx.x % tiles_x) * 16 + (tid & 15);
        py = (int)(blockIdx.x / tiles_x) * 16 + (tid >> 4);
        active = px < a.width && py < a.height;
    } else {                                                  // OVIS_SHADE_LINEAR=1 (tuning): 256 consecutive pixels per CTA
        const int64_t i = (int64_t)blockIdx.x * 256 + tid;
        active = i < (int64_t)a.width * a.height;
        px = (int)(i % a.width);
        py = (int)(i / a.width);
    }
    const int64_t pixel = (int64_t)py * a.width + px;
    Lighting d = {};
    bool pixel_clean = false;
    if (active) {
        const float4 sa = __ldg(reinterpret_cast<const float4 *>(a.buffer_a) + pixel);
        const float4 sb = __ldg(reinterpret_cast<const float4 *>(a.buffer_b) + pixel);
        const float4 sc = __ldg(reinterpret_cast<const float4 *>(a.buffer_c) + pixel);
        const float depth = __ldg(a.depth + pixel);
        d.normal = f3(sa.x * 2.0f - 1.0f, sa.y * 2.0f - 1.0f, sa.z * 2.0f - 1.0f);
        d.diffuse = f3(sb.x, sb.y, sb.z);
        d.specular = f3(sc.x, sc.y, sc.z);
        d.shininess = sc.w;
        const int *vi = reinterpret_cast<const int *>(s_view);
        const float nx = (((float)px - (float)vi[V_VIEWPORT_POS]) / (float)vi[V_VIEWPORT_SIZE]) * 2.0f - 1.0f;
        const float ny = (((float)py - (float)vi[V_VIEWPORT_POS + 1]) / (float)vi[V_VIEWPORT_SIZE + 1]) * 2.0f - 1.0f;
        const float *inv = s_view + V_INVERSE_VP;
        const float hx = mat_row(inv, 0, nx, ny, depth, 1.0f), hy = mat_row(inv, 1, nx, ny, depth, 1.0f);
        const float hz = mat_row(inv, 2, nx, ny, depth, 1.0f), hw = mat_row(inv, 3, nx, ny, depth, 1.0f);
        d.position = f3(hx / hw, hy / hw, hz / hw);
        // comparisons are false for NaN, so one chain covers "finite and within bounds"
        pixel_clean = dot3(d.normal, d.normal) <= 1.001f && d.shininess >= 0.0f && d.shininess <= 4096.0f &&
                      fabsf(d.diffuse.x) <= 1e6f && fabsf(d.diffuse.y) <= 1e6f && fabsf(d.diffuse.z) <= 1e6f &&
                      fabsf(d.specular.x) <= 1e6f && fabsf(d.specular.y) <= 1e6f && fabsf(d.specular.z) <= 1e6f &&
                      fabsf(d.position.x) <= 1e18f && fabsf(d.position.y) <= 1e18f && fabsf(d.position.z) <= 1e18f;
    }
    // ---- bounds of the tile's positions; usable only if every pixel of the tile is clean
    __shared__ float s_box[8][6];
    __shared__ int s_tile_clean;
    __shared__ unsigned char s_skip[kShadeChunk];
    uint32_t m = *a.d_visible_count;
    if ((int64_t)m > a.max_visible)
        m = (uint32_t)a.max_visible;
    if (m <= kShadeTileRejectMinLights) {                     // too few lights for the bounds to pay for themselves
        if (tid == 0)
            s_tile_clean = 0;
    } else {
        const float inf = __int_as_float(0x7f800000);
        const bool use = active && pixel_clean;
        float lo[3] = { use ? d.position.x : inf, use ? d.position.y : inf, use ? d.position.z : inf };
        float hi[3] = { use ? d.position.x : -inf, use ? d.position.y : -inf, use ? d.position.z : -inf };
#pragma unroll
        for (int c = 0; c < 3; c++)
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                lo[c] = fminf(lo[c], __shfl_xor_sync(0xffffffffu, lo[c], o));
                hi[c] = fmaxf(hi[c], __shfl_xor_sync(0xffffffffu, hi[c], o));
            }
        const int all_clean = __syncthreads_and(!active || pixel_clean);
        if ((tid & 31) == 0)
            for (int c = 0; c < 3; c++) {
                s_box[tid >> 5][c] = lo[c];
                s_box[tid >> 5][3 + c] = hi[c];
            }
        __syncthreads();
        if (tid < 3) {
            float l = s_box[0][tid], h = s_box[0][3 + tid];
            for (int wv = 1; wv < 8; wv++) {
                l = fminf(l, s_box[wv][tid]);
                h = fmaxf(h, s_box[wv][3 + tid]);
            }
            s_box[0][tid] = l;
            s_box[0][3 + tid] = h;
        }
        if (tid == 0)
            s_tile_clean = all_clean;
        __syncthreads();
    }
    // normalize(view.position - data.position) does not depend on the light: the shader recomputes it per light, here it
    // is computed once per pixel (same operations, same bits)
    const F3 to_view = normalize3(f3(s_view[V_POSITION], s_view[V_POSITION + 1], s_view[V_POSITION + 2]) - d.position);
    F3 sum = f3(0.0f, 0.0f, 0.0f);
    for (uint32_t base = 0; base < m; base += kShadeChunk) {
        const int n = (int)min((uint32_t)kShadeChunk, m - base);
        __syncthreads();                                          // the previous chunk has been consumed
        for (int i = tid; i < n * L_FLOATS; i += blockDim.x) {
            const uint32_t id = min(a.d_visible[base + i / L_FLOATS], a.n_lights - 1u);     // (an id beyond the table: skipped below)
            s_light[i] = __ldg(a.light_table + (size_t)id * L_FLOATS + i % L_FLOATS);
        }
        if (tid < n) {
            const uint32_t listed = a.d_visible[base + tid], id = min(listed, a.n_lights - 1u);
            // every load is issued before any of them is used: one memory round trip, not a chain of them
            const float *light = a.light_table + (size_t)id * L_FLOATS;
            const int packed_type = __ldg(a.light_types + id);
            const float *map = a.shadow_maps ? a.shadow_maps[id] : nullptr;
            const float r = __ldg(light + L_RANGE);
            const float p0 = __ldg(light + 0), p1 = __ldg(light + 1), p2 = __ldg(light + 2);
            s_type[tid] = packed_type;
            s_map[tid] = map;
            // Tile-level rejection: every pixel of the tile is clean, the light is skippable (see kLightSkippable) and not
            // a shadowed spot light, and the tile's bounds are farther from the light than its range with a margin far
            // above the rounding of the shader's own distance: then every pixel's attenuation is exactly 0.
            const float g0 = fmaxf(fmaxf(s_box[0][0] - p0, p0 - s_box[0][3]), 0.0f);
            const float g1 = fmaxf(fmaxf(s_box[0][1] - p1, p1 - s_box[0][4]), 0.0f);
            const float g2 = fmaxf(fmaxf(s_box[0][2] - p2, p2 - s_box[0][5]), 0.0f);
            const float d2 = g0 * g0 + g1 * g1 + g2 * g2;
            const bool allowed = s_tile_clean && (packed_type & kLightSkippable) && !((packed_type & 0xff) == OVIS_LIGHT_SPOT && map);
            s_skip[tid] = (listed >= a.n_lights || (allowed && (r < 0.0f || d2 > r * r * 1.001f))) ? 1 : 0;   // not a light of this world: nothing
        }
        __syncthreads();
        if (active)
            for (int l = 0; l < n; l++) {
                if (s_skip[l])
                    continue;
                const float *light = s_light + l * L_FLOATS;
                const int type = s_type[l] & 0xff;
                // Out of the light's range the shader multiplies a finite colour by an attenuation of exactly 0 and the
                // blend adds +-0, which leaves the sum as it is (the sum is never -0: it starts at +0). Where every
                // operand is known to be finite and in bounds (light checked on the host, pixel checked above), that
                // work is skipped; a shadowed spot light is not (its projective divide can produce NaN off-range).
                const bool may_skip = (s_type[l] & kLightSkippable) && pixel_clean && !(type == OVIS_LIGHT_SPOT && s_map[l]);
                F3 colour;
                if (calc_light(d, light, type, to_view, s_map[l], a.shadow_size, may_skip, colour))
                    sum = sum + colour;
            }
    }
    if (active)
        reinterpret_cast<float4 *>(a.out_rgba)[pixel] = make_float4(sum.x, sum.y, sum.z, 1.0f);
}

// tonemap_effect_frag.glsl:35-44
__device__ __forceinline__ float uncharted2(float x) {
    const float A = 0.15f, B = 0.50f, C = 0.10f, D = 0.20f, E = 0.02f, F = 0.30f;
    return ((x * (A * x + C * B) + D * E) / (x * (A * x + B) + D * F)) - E / F;
}

__global__ void __launch_bounds__(256)
post_effects_kernel(int64_t n, const float4 *__restrict__ in, uint32_t effects, float exposure, float white_point, float gamma,
                    float4 *__restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
        return;
    float4 c = __ldg(in + i);
    if (effects & OVIS_POST_TONEMAP) {
        const float white = uncharted2(white_point);
        c.x = uncharted2((c.x * exposure) * 2.0f) / white;
        c.y = uncharted2((c.y * exposure) * 2.0f) / white;
        c.z = uncharted2((c.z * exposure) * 2.0f) / white;
    }
    if (effects & OVIS_POST_GAMMA) {
        const float e = 1.0f / gamma;
        c.x = powf(c.x, e);
        c.y = powf(c.y, e);
        c.z = powf(c.z, e);
    }
    out[i] = c;
}

}  // namespace

void launch_shade_deferred(const ShadeArgs &a, cudaStream_t s) {
    static const bool linear = getenv("OVIS_SHADE_LINEAR") && atoi(getenv("OVIS_SHADE_LINEAR")) != 0;
    const int64_t tiles = (int64_t)((a.width + 15) / 16) * ((a.height + 15) / 16);
    const int64_t pixels = (int64_t)a.width * a.height;
    if (pixels <= 0)
        return;
    if (linear)
        shade_deferred_kernel<false><<<(unsigned)((pixels + 255) / 256), 256, 0, s>>>(a);
    else
        shade_deferred_kernel<true><<<(unsigned)tiles, 256, 0, s>>>(a);
}

void launch_post_effects(int64_t n_pixels, const float *in, uint32_t effects, float exposure, float white_point, float gamma, float *out,
                         cudaStream_t s) {
    if (n_pixels > 0)
        post_effects_kernel<<<(unsigned)((n_pixels + 255) / 256), 256, 0, s>>>(
            n_pixels, reinterpret_cast<const float4 *>(in), effects, exposure, white_point, gamma, reinterpret_cast<float4 *>(out));
}

}  // namespace ovis
