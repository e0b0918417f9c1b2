/*
 * oracle/glsl_harness.cc - TEST INFRASTRUCTURE. Runs the reference's own shader SOURCE on the CPU: the GLSL files under
 * /root/reference/engine/shaders are (after the mechanical rewrite below) valid C++ against GLM, the vector library the
 * reference vendors, so the arithmetic of the deferred light pass and the two colour post effects is the reference's
 * text, not a restatement. Only built into oracle/_ref/liborion_ref.so, where /root/reference exists.
 *
 * The Makefile writes rewritten copies of lighting.h, deferred_light_frag.glsl, tonemap_effect_frag.glsl and
 * gamma_correction_effect_frag.glsl into oracle/_ref/glsl/ (git-ignored build output; the sources are read in place):
 *   1.25        -> 1.25f               a GLSL literal is a float, a C++ one a double
 *   v.xyz / .rgb / .xy -> vec3(v) / vec2(v)   GLM's truncating constructors instead of swizzles
 *   (out LightingData data) -> (LightingData &data)
 *   the `in vec4 gl_FragCoord;` redeclaration is dropped (the built-in lives in this file)
 * Everything else GLSL-specific is supplied here: the uniform blocks, samplers and built-ins the shaders name. What the
 * shaders leave to fixed-function hardware - texel fetch, the shadow samplers' compare-and-filter - is NOT reference code:
 * it calls the same definitions oracle/ovis_shading.c states (and exports for that purpose).
 */
#include <glm/glm.hpp>

#include <cstdint>
#include <cstring>

#include "ovis_oracle.h"

extern "C" {
/* the sampler definitions shared with the C restatement (oracle/ovis_shading.c) */
float orc_shadow_lookup_2d(const float *map, int size, float u, float v, float ref);
float orc_shadow_lookup_cube(const float *map, int size, float x, float y, float z, float ref);
}

namespace glsl {

using namespace glm;

/* GLSL lets an int meet a float vector; GLM's templates do not deduce that */
inline vec3 operator*(const vec3 &v, int s) { return v * (float)s; }

struct LightBlock {                    /* render/include/render_light.h:33-47 */
    vec3 position; float intensity;
    vec3 direction; float cosCutoff;
    vec3 colour; float range;
    mat4 volumeTransform;
    mat4 shadowSpace;
    float shadowZNear, attenuationConstant, attenuationLinear, attenuationExp, shadowBiasConstant;
};
struct ViewBlock {                     /* render/include/render_view.h:28-36; the ivec2 members as the float GLSL converts them to */
    mat4 view, projection, viewProjection, inverseViewProjection;
    vec2 viewportPosition, viewportSize;
    vec3 position;
};

static thread_local LightBlock light;
static thread_local ViewBlock view;
static thread_local vec4 gl_FragCoord;

struct sampler2D { const float *data; int width, height; };
struct sampler2DShadow { const float *map; int size; };
struct samplerCubeShadow { const float *map; int size; };

inline vec4 texelFetch(const sampler2D &s, ivec2 uv, int) {
    const float *p = s.data + 4 * ((size_t)uv.y * s.width + uv.x);
    return vec4(p[0], p[1], p[2], p[3]);
}
inline vec2 textureSize(const sampler2DShadow &s, int) { return vec2((float)s.size, (float)s.size); }
inline vec2 textureSize(const samplerCubeShadow &s, int) { return vec2((float)s.size, (float)s.size); }
inline vec2 textureSize(const sampler2D &s, int) { return vec2((float)s.width, (float)s.height); }
inline float texture(const sampler2DShadow &s, vec3 p) { return s.map ? orc_shadow_lookup_2d(s.map, s.size, p.x, p.y, p.z) : 1.0f; }
inline float texture(const samplerCubeShadow &s, vec4 p) { return s.map ? orc_shadow_lookup_cube(s.map, s.size, p.x, p.y, p.z, p.w) : 1.0f; }
/* post effects sample the source at the pixel's own texel (nearest): gl_FragCoord.xy / size * size */
inline vec4 texture(const sampler2D &s, vec2 uv) {
    ivec2 t((int)(uv.x * (float)s.width), (int)(uv.y * (float)s.height));
    return texelFetch(s, t, 0);
}

static thread_local sampler2D deferredBufferA, deferredBufferB, deferredBufferC, deferredBufferD, sourceTexture;
static thread_local float exposure, whitePoint, gamma;

#define layout(...)
#define uniform
#define in
#define out
#define main shader_main

/* one namespace per shader variant (the engine compiles one per light type, with and without SHADOW) */
namespace ambient {
#define AMBIENT_LIGHT
#include "_ref/glsl/deferred_light_frag.glsl"
#undef AMBIENT_LIGHT
#undef __LIGHTING_H
}
namespace directional {
#define DIRECTIONAL_LIGHT
#include "_ref/glsl/deferred_light_frag.glsl"
#undef DIRECTIONAL_LIGHT
#undef __LIGHTING_H
}
namespace point {
#define POINT_LIGHT
#include "_ref/glsl/deferred_light_frag.glsl"
#undef __LIGHTING_H
}
namespace point_shadow {
#define SHADOW
#include "_ref/glsl/deferred_light_frag.glsl"
#undef SHADOW
#undef POINT_LIGHT
#undef __LIGHTING_H
}
namespace spot {
#define SPOT_LIGHT
#include "_ref/glsl/deferred_light_frag.glsl"
#undef __LIGHTING_H
}
namespace spot_shadow {
#define SHADOW
#include "_ref/glsl/deferred_light_frag.glsl"
#undef SHADOW
#undef SPOT_LIGHT
#undef __LIGHTING_H
}
namespace tonemap {
#include "_ref/glsl/tonemap_effect_frag.glsl"
#undef __POST_EFFECT_H
}
namespace gamma_correction {
#include "_ref/glsl/gamma_correction_effect_frag.glsl"
#undef __POST_EFFECT_H
}

/* ---- the shaders the rasteriser runs (oracle/ovis_raster.c restates them): lit_vtx.glsl, lit_deferred_frag.glsl without
 * TEXTURED / NORMALMAP, shadow_vtx.glsl. The uniform blocks and material parameters are injected by the engine's shader
 * compiler, here by hand; `X.rgb = e;` was rewritten to set_rgb(X, e) (a swizzle on the left of an assignment). */
struct EntityBlock { mat4 transform; vec3 position; };          /* render/include/render_entity.h:34-37 */
static thread_local EntityBlock entity;
static thread_local vec4 gl_Position;
static thread_local vec3 diffuseColour, specularColour;
static thread_local float shininess;
inline void set_rgb(vec4 &dst, const vec3 &v) { dst.x = v.x; dst.y = v.y; dst.z = v.z; }

/* material textures: sampled with the definition oracle/ovis_raster.c states (orc_texture_sample) */
struct texture_ref { const void *t; };
static thread_local texture_ref diffuseTexture, normalTexture;
inline vec4 texture(const texture_ref &s, vec2 uv) {
    float o[4];
    orc_texture_sample(s.t, uv.x, uv.y, o);
    return vec4(o[0], o[1], o[2], o[3]);
}

namespace lit_vtx {
#include "_ref/glsl/lit_vtx.glsl"
}
namespace lit_vtx_normalmap {
#define NORMALMAP
#include "_ref/glsl/lit_vtx.glsl"
#undef NORMALMAP
}
/* (the eight keyword combinations, each in its own namespace) */
namespace lit_frag_t {
#define TEXTURED
#include "_ref/glsl/lit_deferred_frag.glsl"
#undef TEXTURED
}
namespace lit_frag_n {
#define NORMALMAP
#include "_ref/glsl/lit_deferred_frag.glsl"
#undef NORMALMAP
}
namespace lit_frag_tn {
#define TEXTURED
#define NORMALMAP
#include "_ref/glsl/lit_deferred_frag.glsl"
#undef TEXTURED
#undef NORMALMAP
}
namespace lit_frag_st {
#define SPECULAR
#define TEXTURED
#include "_ref/glsl/lit_deferred_frag.glsl"
#undef TEXTURED
#undef SPECULAR
}
namespace lit_frag_sn {
#define SPECULAR
#define NORMALMAP
#include "_ref/glsl/lit_deferred_frag.glsl"
#undef NORMALMAP
#undef SPECULAR
}
namespace lit_frag_stn {
#define SPECULAR
#define TEXTURED
#define NORMALMAP
#include "_ref/glsl/lit_deferred_frag.glsl"
#undef TEXTURED
#undef NORMALMAP
#undef SPECULAR
}

namespace shadow_vtx {
#include "_ref/glsl/shadow_vtx.glsl"
}
namespace unlit_vtx {
#include "_ref/glsl/unlit_vtx.glsl"
}
namespace unlit_frag {
#include "_ref/glsl/unlit_frag.glsl"
}
namespace lit_frag {
#include "_ref/glsl/lit_deferred_frag.glsl"
}
namespace lit_frag_specular {
#define SPECULAR
#include "_ref/glsl/lit_deferred_frag.glsl"
#undef SPECULAR
}

#undef layout
#undef uniform
#undef in
#undef out
#undef main

}  // namespace glsl

extern "C" {

/* Same contract as orc_shade_deferred; every pixel runs the reference's main() of the variant for each light. */
void orc_glsl_shade_deferred(int32_t width, int32_t height, const float *buffer_a, const float *buffer_b, const float *buffer_c,
                             const float *depth, const void *view_uniforms, int32_t n_lights, const float *light_uniforms,
                             const int32_t *light_types, const float *const *shadow_maps, int32_t shadow_size, float *out_rgba) {
    using namespace glsl;
    const float *vf = (const float *)view_uniforms;
    const int32_t *vi = (const int32_t *)view_uniforms;
    memcpy(&view.view, vf, 256);
    view.viewportPosition = vec2((float)vi[64], (float)vi[65]);
    view.viewportSize = vec2((float)vi[66], (float)vi[67]);
    view.position = vec3(vf[68], vf[69], vf[70]);
    /* the depth copy is a one-channel image: present it as the r channel of a 4-channel fetch */
    float *depth4 = new float[(size_t)width * height * 4]();
    for (size_t i = 0; i < (size_t)width * height; i++)
        depth4[4 * i] = depth[i];
    deferredBufferA = { buffer_a, width, height };
    deferredBufferB = { buffer_b, width, height };
    deferredBufferC = { buffer_c, width, height };
    deferredBufferD = { depth4, width, height };
    for (int32_t py = 0; py < height; py++)
        for (int32_t px = 0; px < width; px++) {
            gl_FragCoord = vec4((float)px, (float)py, 0.0f, 1.0f);     /* pixel_center_integer */
            vec3 sum(0.0f);
            for (int32_t l = 0; l < n_lights; l++) {
                const float *lf = light_uniforms + 52 * (size_t)l;
                memcpy(&light.position, lf, 12); light.intensity = lf[3];
                memcpy(&light.direction, lf + 4, 12); light.cosCutoff = lf[7];
                memcpy(&light.colour, lf + 8, 12); light.range = lf[11];
                memcpy(&light.volumeTransform, lf + 12, 64);
                memcpy(&light.shadowSpace, lf + 28, 64);
                light.shadowZNear = lf[44]; light.attenuationConstant = lf[45]; light.attenuationLinear = lf[46];
                light.attenuationExp = lf[47]; light.shadowBiasConstant = lf[48];
                const float *map = shadow_maps ? shadow_maps[l] : nullptr;
                vec4 colour;
                switch (light_types[l]) {
                case ORC_LIGHT_AMBIENT: ambient::shader_main(); colour = ambient::fragColour; break;
                case ORC_LIGHT_DIRECTIONAL: directional::shader_main(); colour = directional::fragColour; break;
                case ORC_LIGHT_POINT:
                    if (map) { point_shadow::shadowMap = { map, shadow_size }; point_shadow::shader_main(); colour = point_shadow::fragColour; }
                    else { point::shader_main(); colour = point::fragColour; }
                    break;
                default:
                    if (map) { spot_shadow::shadowMap = { map, shadow_size }; spot_shadow::shader_main(); colour = spot_shadow::fragColour; }
                    else { spot::shader_main(); colour = spot::fragColour; }
                    break;
                }
                sum = sum + vec3(colour);                              /* additive blend, one pass per light */
            }
            float *o = out_rgba + 4 * ((size_t)py * width + px);
            o[0] = sum.x; o[1] = sum.y; o[2] = sum.z; o[3] = 1.0f;
        }
    delete[] depth4;
}

/* lit_vtx.glsl's main() for one vertex: clip position and world-space normal (the two outputs the rasteriser consumes) */
void orc_glsl_lit_vertex(const float *transform, const float *view_projection, const float *position, const float *normal,
                         float *out_clip, float *out_normal) {
    using namespace glsl;
    memcpy(&entity.transform, transform, 64);
    memcpy(&view.viewProjection, view_projection, 64);
    lit_vtx::attribPosition = vec3(position[0], position[1], position[2]);
    lit_vtx::attribNormal = vec3(normal[0], normal[1], normal[2]);
    lit_vtx::attribTexcoord = vec2(0.0f);
    lit_vtx::shader_main();
    out_clip[0] = gl_Position.x; out_clip[1] = gl_Position.y; out_clip[2] = gl_Position.z; out_clip[3] = gl_Position.w;
    out_normal[0] = lit_vtx::vtxNormal.x; out_normal[1] = lit_vtx::vtxNormal.y; out_normal[2] = lit_vtx::vtxNormal.z;
}

void orc_glsl_shadow_vertex(const float *transform, const float *view_projection, const float *position, float *out_clip) {
    using namespace glsl;
    memcpy(&entity.transform, transform, 64);
    memcpy(&view.viewProjection, view_projection, 64);
    shadow_vtx::attribPosition = vec3(position[0], position[1], position[2]);
    shadow_vtx::shader_main();
    out_clip[0] = gl_Position.x; out_clip[1] = gl_Position.y; out_clip[2] = gl_Position.z; out_clip[3] = gl_Position.w;
}

/* lit_vtx.glsl with NORMALMAP for one vertex: world normal, tangent, bitangent */
void orc_glsl_lit_vertex_normalmap(const float *transform, const float *view_projection, const float *position, const float *normal,
                                   const float *tangent4, float *out_normal, float *out_tangent, float *out_bitangent) {
    using namespace glsl;
    memcpy(&entity.transform, transform, 64);
    memcpy(&view.viewProjection, view_projection, 64);
    lit_vtx_normalmap::attribPosition = vec3(position[0], position[1], position[2]);
    lit_vtx_normalmap::attribNormal = vec3(normal[0], normal[1], normal[2]);
    lit_vtx_normalmap::attribTexcoord = vec2(0.0f);
    lit_vtx_normalmap::attribTangent = vec4(tangent4[0], tangent4[1], tangent4[2], tangent4[3]);
    lit_vtx_normalmap::shader_main();
    memcpy(out_normal, &lit_vtx_normalmap::vtxNormal, 12);
    memcpy(out_tangent, &lit_vtx_normalmap::vtxTangent, 12);
    memcpy(out_bitangent, &lit_vtx_normalmap::vtxBitangent, 12);
}

/* lit_deferred_frag.glsl for one fragment with any combination of SPECULAR (1) / TEXTURED (2) / NORMALMAP (4) in material[7];
 * same arguments as orc_raster_fragment_textured */
#define OVIS_RUN_FRAG(NS)                                                                                                  \
    do {                                                                                                                   \
        NS::vtxNormal = n; NS::vtxTexcoord = tc;                                                                           \
        NS::shader_main();                                                                                                 \
        a = NS::deferredBufferA; b = NS::deferredBufferB; c = NS::deferredBufferC;                                         \
    } while (0)
#define OVIS_RUN_FRAG_N(NS)                                                                                                \
    do {                                                                                                                   \
        NS::vtxTangent = tg; NS::vtxBitangent = bt;                                                                        \
        OVIS_RUN_FRAG(NS);                                                                                                 \
    } while (0)
void orc_glsl_lit_fragment_textured(const float *normal, const float *uv, const float *tangent, const float *bitangent,
                                    const float *material, const void *textures, float *out_a, float *out_b, float *out_c) {
    using namespace glsl;
    struct tex_record { const uint8_t *rgba; int32_t width, height, srgb, wrap; };
    const tex_record *tex = (const tex_record *)textures;
    diffuseColour = vec3(material[0], material[1], material[2]);
    specularColour = vec3(material[3], material[4], material[5]);
    shininess = material[6];
    const int flags = (int)material[7];
    if (flags & 2) diffuseTexture.t = tex + (int)material[8];
    if (flags & 4) normalTexture.t = tex + (int)material[9];
    const vec3 n(normal[0], normal[1], normal[2]), tg(tangent[0], tangent[1], tangent[2]), bt(bitangent[0], bitangent[1], bitangent[2]);
    const vec2 tc(uv[0], uv[1]);
    vec4 a, b, c;
    switch (flags & 7) {
    case 0: OVIS_RUN_FRAG(lit_frag); break;
    case 1: OVIS_RUN_FRAG(lit_frag_specular); break;
    case 2: OVIS_RUN_FRAG(lit_frag_t); break;
    case 3: OVIS_RUN_FRAG(lit_frag_st); break;
    case 4: OVIS_RUN_FRAG_N(lit_frag_n); break;
    case 5: OVIS_RUN_FRAG_N(lit_frag_sn); break;
    case 6: OVIS_RUN_FRAG_N(lit_frag_tn); break;
    default: OVIS_RUN_FRAG_N(lit_frag_stn); break;
    }
    memcpy(out_a, &a, 16); memcpy(out_b, &b, 16); memcpy(out_c, &c, 16);
}
#undef OVIS_RUN_FRAG
#undef OVIS_RUN_FRAG_N

/* unlit_vtx.glsl + unlit_frag.glsl (the Basic pass) for one vertex / fragment */
void orc_glsl_unlit(const float *transform, const float *view_projection, const float *position, const float *diffuse,
                    float *out_clip, float *out_colour) {
    using namespace glsl;
    memcpy(&entity.transform, transform, 64);
    memcpy(&view.viewProjection, view_projection, 64);
    unlit_vtx::attribPosition = vec3(position[0], position[1], position[2]);
    unlit_vtx::attribTexcoord = vec2(0.0f);
    unlit_vtx::shader_main();
    out_clip[0] = gl_Position.x; out_clip[1] = gl_Position.y; out_clip[2] = gl_Position.z; out_clip[3] = gl_Position.w;
    diffuseColour = vec3(diffuse[0], diffuse[1], diffuse[2]);
    unlit_frag::shader_main();
    memcpy(out_colour, &unlit_frag::fragColour, 16);
}

/* lit_deferred_frag.glsl's main() for one fragment; material = diffuse rgb, specular rgb, shininess, SPECULAR keyword (0 / 1) */
void orc_glsl_lit_fragment(const float *normal, const float *material, float *out_a, float *out_b, float *out_c) {
    using namespace glsl;
    diffuseColour = vec3(material[0], material[1], material[2]);
    specularColour = vec3(material[3], material[4], material[5]);
    shininess = material[6];
    vec4 a, b, c;
    if (material[7] != 0.0f) {
        lit_frag_specular::vtxNormal = vec3(normal[0], normal[1], normal[2]);
        lit_frag_specular::shader_main();
        a = lit_frag_specular::deferredBufferA; b = lit_frag_specular::deferredBufferB; c = lit_frag_specular::deferredBufferC;
    } else {
        lit_frag::vtxNormal = vec3(normal[0], normal[1], normal[2]);
        lit_frag::shader_main();
        a = lit_frag::deferredBufferA; b = lit_frag::deferredBufferB; c = lit_frag::deferredBufferC;
    }
    memcpy(out_a, &a, 16); memcpy(out_b, &b, 16); memcpy(out_c, &c, 16);
}

void orc_glsl_tonemap(int64_t n_pixels, const float *rgba, float exposure_, float white_point, float *out_rgba) {
    using namespace glsl;
    glsl::exposure = exposure_;
    glsl::whitePoint = white_point;
    sourceTexture = { rgba, (int)n_pixels, 1 };
    for (int64_t i = 0; i < n_pixels; i++) {
        gl_FragCoord = vec4((float)i + 0.5f, 0.5f, 0.0f, 1.0f);       /* default pixel centres */
        tonemap::shader_main();
        memcpy(out_rgba + 4 * i, &tonemap::fragColour, 16);
    }
}

void orc_glsl_gamma(int64_t n_pixels, const float *rgba, float gamma_, float *out_rgba) {
    using namespace glsl;
    glsl::gamma = gamma_;
    sourceTexture = { rgba, (int)n_pixels, 1 };
    for (int64_t i = 0; i < n_pixels; i++) {
        gl_FragCoord = vec4((float)i + 0.5f, 0.5f, 0.0f, 1.0f);
        gamma_correction::shader_main();
        memcpy(out_rgba + 4 * i, &gamma_correction::fragColour, 16);
    }
}

}  /* extern "C" */
