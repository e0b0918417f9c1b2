/*
 * oracle/ovis_shading.c - TEST INFRASTRUCTURE. CPU restatement of the reference's deferred light shader and its two
 * colour post effects. Nothing in the product may call this; only tests/ and the CPU
 * legs of the benchmarks do.
 *
 * Restated from the reference's GLSL, statement for statement, in float arithmetic (a GLSL literal is a float):
 *   engine/shaders/deferred_light_frag.glsl:30-81   decodeGBuffer, main
 *   engine/shaders/lighting.h:46-143                calcShadow (3x3 PCF, spot and point)
 *   engine/shaders/lighting.h:149-211               calcLightBlinnPhong, calcLight
 *   engine/shaders/tonemap_effect_frag.glsl:30-71   Uncharted 2 operator
 *   engine/shaders/gamma_correction_effect_frag.glsl:26-31
 * and the fixed-function steps around them as the pipeline configures them: additive blending of one pass per light in
 * list order into a target cleared to (0, 0, 0, 1) with alpha = source alpha (render/src/deferred_render_pipeline.cc:
 * 424-446), shadow sampler = bilinear, compare Less, clamp addressing (:250-254).
 *
 * PARITY UNPINNED by reference code: GLSL has no CPU implementation in the reference and a graphics API is unavailable
 * on this box, so there is nothing to execute the shaders with. GLSL also leaves the precision of pow / normalize /
 * division and the exact filtering of shadow samplers to the implementation. What this file fixes as the definition:
 *   - IEEE float, operations in source order, libm powf / sqrtf, no contraction (-ffp-contract=off);
 *   - inputs in fp32 (the reference's R10G10B10A2 / fp16 / Depth32 targets are storage formats of the rasteriser);
 *   - sampler2DShadow: texel centres at (i + 0.5) / size, clamp to edge, 4 taps each compared (ref < texel, ref clamped to
 *     [0, 1]) and blended with the bilinear weights - what every current GPU does for a linear-filtered shadow sampler;
 *   - samplerCubeShadow: major-axis face selection of the GL specification (ties: x, then y), face order +x -x +y -y +z -z,
 *     then the same 2D lookup inside the face (no filtering across face edges).
 * The CUDA path is compared against it with the tolerance written in tests/test_shading_gpu.py.
 */
#include "ovis_oracle.h"

#include <math.h>
#include <string.h>

typedef struct { float x, y, z; } s3;

static s3 S3(float x, float y, float z) { s3 r = { x, y, z }; return r; }
static s3 s3_add(s3 a, s3 b) { return S3(a.x + b.x, a.y + b.y, a.z + b.z); }
static s3 s3_sub(s3 a, s3 b) { return S3(a.x - b.x, a.y - b.y, a.z - b.z); }
static s3 s3_mul(s3 a, s3 b) { return S3(a.x * b.x, a.y * b.y, a.z * b.z); }
static s3 s3_scale(s3 a, float s) { return S3(a.x * s, a.y * s, a.z * s); }
static float s3_dot(s3 a, s3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static float s3_length(s3 a) { return sqrtf(s3_dot(a, a)); }
static s3 s3_normalize(s3 a) { return s3_scale(a, 1.0f / s3_length(a)); }
static float maxf(float a, float b) { return a > b ? a : b; }      /* GLSL max: y if x < y else x; same for non-NaN */
static float clampf(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }

/* column-major mat4 * vec4. GLSL does not say how the four products are summed; this takes GLM's association
 * (glm/detail/type_mat4x4.inl:545-571: (c0 v0 + c1 v1) + (c2 v2 + c3 v3)) so that this file and the reference's shader source
 * compiled against GLM (oracle/glsl_harness.cc) can be compared for equality. */
static void mat4_mul_vec4(const float *m, const float v[4], float out[4]) {
    for (int r = 0; r < 4; r++)
        out[r] = (m[0 + r] * v[0] + m[4 + r] * v[1]) + (m[8 + r] * v[2] + m[12 + r] * v[3]);
}

/* LightUniforms block, float indices (render/include/render_light.h:33-47) */
enum { L_POSITION = 0, L_INTENSITY = 3, L_DIRECTION = 4, L_COS_CUTOFF = 7, L_COLOUR = 8, L_RANGE = 11, L_SHADOW_SPACE = 28,
       L_SHADOW_ZNEAR = 44, L_ATT_CONSTANT = 45, L_ATT_LINEAR = 46, L_ATT_EXP = 47, L_SHADOW_BIAS = 48, L_FLOATS = 52 };
/* ViewUniforms block (render/include/render_view.h:28-36) */
enum { V_INVERSE_VP = 48, V_VIEWPORT_POS = 64, V_VIEWPORT_SIZE = 66, V_POSITION = 68 };

/* texture(sampler2DShadow, vec3(u, v, ref)) on one size x size depth image */
static float shadow_lookup_2d(const float *map, int size, float u, float v, float ref) {
    ref = clampf(ref, 0.0f, 1.0f);
    float x = u * (float)size - 0.5f, y = v * (float)size - 0.5f;
    float fx0 = floorf(x), fy0 = floorf(y);
    float wx = x - fx0, wy = y - fy0;
    /* clamp to edge, in float before the conversion (a float beyond int's range converts to anything in C, and x0 + 1
     * could overflow): far to the right is the right edge, NaN the left one (its weights are NaN anyway) */
    int x0 = (int)fminf(fmaxf(fx0, -1.0f), (float)size), y0 = (int)fminf(fmaxf(fy0, -1.0f), (float)size), x1 = x0 + 1, y1 = y0 + 1;
    if (x0 < 0) x0 = 0; if (x0 > size - 1) x0 = size - 1;
    if (x1 < 0) x1 = 0; if (x1 > size - 1) x1 = size - 1;
    if (y0 < 0) y0 = 0; if (y0 > size - 1) y0 = size - 1;
    if (y1 < 0) y1 = 0; if (y1 > size - 1) y1 = size - 1;
    float c00 = ref < map[(size_t)y0 * size + x0] ? 1.0f : 0.0f, c10 = ref < map[(size_t)y0 * size + x1] ? 1.0f : 0.0f;
    float c01 = ref < map[(size_t)y1 * size + x0] ? 1.0f : 0.0f, c11 = ref < map[(size_t)y1 * size + x1] ? 1.0f : 0.0f;
    float top = c00 * (1.0f - wx) + c10 * wx, bottom = c01 * (1.0f - wx) + c11 * wx;
    return top * (1.0f - wy) + bottom * wy;
}

/* texture(samplerCubeShadow, vec4(direction, ref)) on six size x size faces */
static float shadow_lookup_cube(const float *map, int size, s3 d, float ref) {
    float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z), sc, tc, ma;
    int face;
    if (ax >= ay && ax >= az) { face = d.x >= 0.0f ? 0 : 1; ma = ax; sc = d.x >= 0.0f ? -d.z : d.z; tc = -d.y; }
    else if (ay >= az)         { face = d.y >= 0.0f ? 2 : 3; ma = ay; sc = d.x; tc = d.y >= 0.0f ? d.z : -d.z; }
    else                       { face = d.z >= 0.0f ? 4 : 5; ma = az; sc = d.z >= 0.0f ? d.x : -d.x; tc = -d.y; }
    float s = (sc / ma + 1.0f) * 0.5f, t = (tc / ma + 1.0f) * 0.5f;
    return shadow_lookup_2d(map + (size_t)face * size * size, size, s, t, ref);
}

/* the sampler definitions, exported so that the GLSL-compiled tier (oracle/glsl_harness.cc) uses the very same ones */
float orc_shadow_lookup_2d(const float *map, int size, float u, float v, float ref) {
    return shadow_lookup_2d(map, size, u, v, ref);
}
float orc_shadow_lookup_cube(const float *map, int size, float x, float y, float z, float ref) {
    return shadow_lookup_cube(map, size, S3(x, y, z), ref);
}

typedef struct { s3 normal, position, diffuse, specular; float shininess; } lighting_data;

/* lighting.h:46-143 */
static float calc_shadow(const lighting_data *data, const float *light, int type, const float *map, int size) {
    if (!map)
        return 1.0f;
    if (type == ORC_LIGHT_SPOT) {
        float pos[4] = { data->position.x, data->position.y, data->position.z, 1.0f }, sp[4];
        mat4_mul_vec4(light + L_SHADOW_SPACE, pos, sp);                         /* :52 */
        float u = sp[0] / sp[3], v = sp[1] / sp[3], depth = sp[2] / sp[3];       /* :56 */
        depth += light[L_SHADOW_BIAS];                                          /* :59 */
        float shadow = 0.0f;
        float texel = 1.0f / (float)size;                                       /* :63 */
        for (int x = -1; x <= 1; x++)
            for (int y = -1; y <= 1; y++)
                shadow += shadow_lookup_2d(map, size, u + (float)x * texel, v + (float)y * texel, depth);   /* :66-72 */
        return shadow / 9.0f;                                                   /* :76 */
    }
    if (type == ORC_LIGHT_POINT) {
        s3 direction = s3_sub(data->position, S3(light[0], light[1], light[2]));  /* :84 */
        float ax = fabsf(direction.x), ay = fabsf(direction.y), az = fabsf(direction.z), local_depth;
        s3 oa, ob;
        if (ax >= ay && ax >= az) { local_depth = ax; oa = S3(0, 1, 0); ob = S3(0, 0, 1); }        /* :90-102 */
        else if (ay >= ax && ay >= az) { local_depth = ay; oa = S3(1, 0, 0); ob = S3(0, 0, 1); }
        else { local_depth = az; oa = S3(1, 0, 0); ob = S3(0, 1, 0); }
        float far = light[L_RANGE], near = light[L_SHADOW_ZNEAR];
        float depth = -(far / (near - far)) - ((far * near) / (far - near) / local_depth);           /* :107 */
        float texel = 1.0f / (float)size;                                       /* :112 */
        oa = s3_scale(s3_scale(oa, texel), 2.0f);                               /* :113-114 */
        ob = s3_scale(s3_scale(ob, texel), 2.0f);
        s3 nd = s3_normalize(direction);                                        /* :119 */
        float ref = depth + light[L_SHADOW_BIAS];
        float shadow = 0.0f;
        for (int a = -1; a <= 1; a++)
            for (int b = -1; b <= 1; b++) {
                s3 p = s3_add(s3_add(nd, s3_scale(oa, (float)a)), s3_scale(ob, (float)b));          /* :125-127 */
                shadow += shadow_lookup_cube(map, size, p, ref);
            }
        return shadow / 9.0f;
    }
    return 1.0f;
}

/* lighting.h:149-168 */
static s3 blinn_phong(const lighting_data *data, s3 direction, const float *light, s3 view_position) {
    s3 to_light = S3(-direction.x, -direction.y, -direction.z);
    s3 to_view = s3_normalize(s3_sub(view_position, data->position));
    float angle = maxf(s3_dot(data->normal, to_light), 0.0f);
    s3 light_colour = S3(light[L_COLOUR], light[L_COLOUR + 1], light[L_COLOUR + 2]);
    s3 colour = s3_scale(s3_mul(data->diffuse, light_colour), light[L_INTENSITY] * angle);
    s3 half_vector = s3_normalize(s3_add(to_light, to_view));
    float specular_angle = maxf(s3_dot(data->normal, half_vector), 0.0f);
    colour = s3_add(colour, s3_scale(s3_mul(data->specular, light_colour),
                                     light[L_INTENSITY] * powf(specular_angle, data->shininess)));
    return colour;
}

/* lighting.h:173-211 */
static s3 calc_light(const lighting_data *data, const float *light, int type, s3 view_position, const float *map, int size) {
    s3 light_colour = S3(light[L_COLOUR], light[L_COLOUR + 1], light[L_COLOUR + 2]);
    if (type == ORC_LIGHT_AMBIENT)
        return s3_scale(s3_mul(data->diffuse, light_colour), light[L_INTENSITY]);                    /* :175 */
    s3 direction;
    float attenuation = 1.0f;
    s3 light_direction = S3(light[L_DIRECTION], light[L_DIRECTION + 1], light[L_DIRECTION + 2]);
    if (type == ORC_LIGHT_DIRECTIONAL) {
        direction = light_direction;                                                                 /* :182 */
    } else {
        s3 to_fragment = s3_sub(data->position, S3(light[0], light[1], light[2]));                   /* :185 */
        float distance = s3_length(to_fragment);
        direction = s3_normalize(to_fragment);
        attenuation = clampf(floorf(light[L_RANGE] / distance), 0.0f, 1.0f);                          /* :192 */
        if (type == ORC_LIGHT_SPOT) {
            float spot = maxf(s3_dot(direction, light_direction), light[L_COS_CUTOFF]);              /* :196 */
            attenuation *= 1.0f - ((1.0f - spot) / (1.0f - light[L_COS_CUTOFF]));                     /* :197 */
        }
        attenuation /= light[L_ATT_CONSTANT] + (light[L_ATT_LINEAR] * distance) +
                       (light[L_ATT_EXP] * distance * distance);                                      /* :201-204 */
    }
    attenuation *= calc_shadow(data, light, type, map, size);                                        /* :208 */
    return s3_scale(blinn_phong(data, direction, light, view_position), attenuation);                /* :210 */
}

void orc_shade_deferred(int32_t width, int32_t height, const float *buffer_a, const float *buffer_b, const float *buffer_c,
                        const float *depth, const void *view_uniforms, int32_t n_lights, const float *light_uniforms,
                        const int32_t *light_types, const float *const *shadow_maps, int32_t shadow_size, float *out_rgba) {
    const float *view = (const float *)view_uniforms;
    const int32_t *view_i = (const int32_t *)view_uniforms;
    s3 view_position = S3(view[V_POSITION], view[V_POSITION + 1], view[V_POSITION + 2]);
    for (int32_t py = 0; py < height; py++)
        for (int32_t px = 0; px < width; px++) {
            size_t i = (size_t)py * width + px;
            lighting_data data;                                                 /* deferred_light_frag.glsl:32-70 */
            data.normal = S3(buffer_a[4 * i] * 2.0f - 1.0f, buffer_a[4 * i + 1] * 2.0f - 1.0f, buffer_a[4 * i + 2] * 2.0f - 1.0f);
            data.diffuse = S3(buffer_b[4 * i], buffer_b[4 * i + 1], buffer_b[4 * i + 2]);
            data.specular = S3(buffer_c[4 * i], buffer_c[4 * i + 1], buffer_c[4 * i + 2]);
            data.shininess = buffer_c[4 * i + 3];
            /* gl_FragCoord with pixel_center_integer: the integer pixel coordinate within the target (:24, :65) */
            float ndc[4] = { (((float)px - (float)view_i[V_VIEWPORT_POS]) / (float)view_i[V_VIEWPORT_SIZE]) * 2.0f - 1.0f,
                             (((float)py - (float)view_i[V_VIEWPORT_POS + 1]) / (float)view_i[V_VIEWPORT_SIZE + 1]) * 2.0f - 1.0f,
                             depth[i], 1.0f };
            float hp[4];
            mat4_mul_vec4(view + V_INVERSE_VP, ndc, hp);                        /* :68 */
            data.position = S3(hp[0] / hp[3], hp[1] / hp[3], hp[2] / hp[3]);    /* :69 */
            s3 sum = S3(0.0f, 0.0f, 0.0f);                                      /* clear colour, pipeline :428 */
            for (int32_t l = 0; l < n_lights; l++)                              /* one additive pass per light, in list order */
                sum = s3_add(sum, calc_light(&data, light_uniforms + (size_t)l * L_FLOATS, light_types[l], view_position,
                                             shadow_maps ? shadow_maps[l] : 0, shadow_size));
            out_rgba[4 * i] = sum.x; out_rgba[4 * i + 1] = sum.y; out_rgba[4 * i + 2] = sum.z;
            out_rgba[4 * i + 3] = 1.0f;                                         /* source alpha x 1 + dest alpha x 0 */
        }
}

/* tonemap_effect_frag.glsl:35-44 */
static float uncharted2(float x) {
    const float A = 0.15f, B = 0.50f, C = 0.10f, D = 0.20f, E = 0.02f, F = 0.30f;
    return ((x * (A * x + C * B) + D * E) / (x * (A * x + B) + D * F)) - E / F;
}

void orc_tonemap(int64_t n_pixels, const float *rgba, float exposure, float white_point, float *out_rgba) {
    for (int64_t i = 0; i < n_pixels; i++) {
        for (int c = 0; c < 3; c++) {
            float colour = rgba[4 * i + c] * exposure;                          /* :61 */
            colour *= 2.0f;                                                     /* exposureBias, :53-54 */
            out_rgba[4 * i + c] = uncharted2(colour) / uncharted2(white_point); /* :56 */
        }
        out_rgba[4 * i + 3] = rgba[4 * i + 3];                                  /* :71 */
    }
}

/* gamma_correction_effect_frag.glsl:26-31 */
void orc_gamma(int64_t n_pixels, const float *rgba, float gamma, float *out_rgba) {
    for (int64_t i = 0; i < n_pixels; i++) {
        for (int c = 0; c < 3; c++)
            out_rgba[4 * i + c] = powf(rgba[4 * i + c], 1.0f / gamma);
        out_rgba[4 * i + 3] = rgba[4 * i + 3];
    }
}
