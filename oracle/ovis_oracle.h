/*
 * oracle/ovis_oracle.h - the flat "checker" API shared by BOTH CPU oracles.
 *
 * TEST INFRASTRUCTURE ONLY. Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may load
 * the libraries that export these symbols; nothing under orion_b200/ does.
 *
 *   oracle/libovis_oracle.so      (ovis_oracle.c)   plain-C restatement, prefix orc_
 *   oracle/_ref/liborion_ref.so   (ref_harness.cc)  the same orc_ entry points implemented by calling
 *                                                   the reference's own compiled classes, plus the
 *                                                   object-world oref_ entry points (tier A)
 *
 * Conventions (identical in include/orion_vis.h):
 *   TRS    = 10 floats: position xyz, orientation w x y z, scale xyz
 *   AABB   = 6 floats:  minimum xyz, maximum xyz
 *   planes = 24 floats: 6 x (normal xyz, distance), order Left Right Top Bottom Near Far
 *            (engine/runtime/core/include/math/frustum.h:39-47)
 *   mat4   = 16 floats column-major (GLM: m[col][row])
 */
#ifndef OVIS_ORACLE_H
#define OVIS_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* RenderLight::Type, engine/runtime/render/include/render_light.h:52-58 */
enum { ORC_LIGHT_AMBIENT = 0, ORC_LIGHT_DIRECTIONAL = 1, ORC_LIGHT_POINT = 2, ORC_LIGHT_SPOT = 3 };
/* RenderLight::kCastsShadows (render_light.h:61-64), RenderEntity::kCastsShadow (render_entity.h:50-53) */
enum { ORC_CASTS_SHADOW = 1u };
/* RenderWorld::CullFlags::kCullLights, engine/runtime/render/include/render_world.h:43-46 */
enum { ORC_CULL_LIGHTS = 1u };

typedef struct OrcLight {
    int32_t  type;
    uint32_t flags;
    float    position[3];
    float    direction[3];   /* normalised by the light, as RenderLight::setDirection does */
    float    colour[3];
    float    intensity;
    float    range;
    float    cutoff;         /* degrees, spot only, <= 45 */
} OrcLight;                   /* 56 bytes */

typedef struct OrcView {
    float   position[3];
    float   orientation[4];  /* w x y z */
    float   hfov_deg;
    float   znear;
    float   zfar;
    int32_t vp_width;        /* aspect = (float)vp_width / (float)vp_height, render_view.cc:84 */
    int32_t vp_height;
} OrcView;                    /* 48 bytes */

/* ---- view / frustum (render_view.cc:102-135, frustum.cc:27-85) ---- */
void orc_view_planes(const OrcView *view, float out_planes[24]);
/* out_mats = view, projection, viewProjection, inverseViewProjection (4 x 16); corners = 8 x 3 */
void orc_view_matrices(const OrcView *view, float out_mats[64], float out_corners[24]);

/* ---- transform propagation (entity.cc:382-411) over depth-sorted nodes: parent[i] < i, -1 = root child of identity ---- */
void orc_world_trs(int64_t n, const int32_t *parent, const float *local_trs, float *out_world_trs);
/* Transform::matrix (transform.cc:26-35) */
void orc_trs_matrix(int64_t n, const float *trs, float *out_mat16);
/* RenderEntity::updateWorld (render_entity.cc:90-93) = BoundingBox::transform(Transform::matrix()) */
void orc_world_aabb(int64_t n, const float *world_trs, const float *local_aabb, float *out_aabb);

/* ---- entity cull over flat AABBs (simple_render_world.cc:38-41 + intersection.cc:42-57) ----
 * flags may be NULL. An entity is emitted iff it intersects and (flags[i] & require) == require.
 * Returns the number of indices written to out_idx (capacity n). */
int64_t orc_cull_aabb(int64_t n, const float *aabb, const float planes[24],
                      const uint32_t *flags, uint32_t require, uint32_t *out_idx);
/* Same result, index range split over nthreads host threads (CPU baseline "all cores"). */
int64_t orc_cull_aabb_mt(int64_t n, const float *aabb, const float planes[24],
                         const uint32_t *flags, uint32_t require, uint32_t *out_idx, int nthreads);

/* ---- lights ---- */
/* RenderLight::updateVolumeTransform, spot branch (render_light.cc:227-249): cone-fitted AABB */
void orc_light_spot_aabb(const OrcLight *light, float out_aabb[6]);
/* RenderLight::cull (render_light.cc:187-209): writes indices of lights that are NOT culled */
int32_t orc_light_cull(int32_t n, const OrcLight *lights, const float planes[24], uint32_t *out_idx);
/* RenderLight::updateShadowViews (render_light.cc:258-332) with the square shadow-map viewport of
 * deferred_render_pipeline.cc:261-265. Returns the number of shadow views (0, 1 or 6). */
int32_t orc_light_shadow_views(const OrcLight *light, OrcView out_views[6]);

/* ---- per-light shadow-caster lists (deferred_render_pipeline.cc:207,247,257-284) ----
 * For every light NOT culled by main_planes that casts shadows, for each of its shadow views:
 * the ordered list of entities that intersect that view and have ORC_CASTS_SHADOW.
 * out_list_light[k], out_list_view[k] identify list k; out_offsets has n_lists + 1 entries.
 * Returns n_lists, or -1 if a capacity is too small. */
int32_t orc_shadow_lists(int64_t n, const float *aabb, const uint32_t *flags,
                         int32_t n_lights, const OrcLight *lights, const float main_planes[24],
                         int32_t cap_lists, int32_t *out_list_light, int32_t *out_list_view,
                         int64_t *out_offsets, int64_t cap_entities, uint32_t *out_entities);

/* ---- draw lists ----
 * pass_counts[i] packs Shader::numPasses of entity i's material: bits 0-7 deferred, 8-15 basic, 16-23 shadow caster.
 * mode 0 = DeferredRenderPipeline::prepareEntities (deferred_render_pipeline.cc:318-330): for each visible entity in
 *          order, numPasses(deferred) draws into list 0, else numPasses(basic) draws into list 1 (DrawList::add,
 *          draw_list.cc:35-45: one draw per pass, pass index ascending);
 * mode 1 = the shadow-caster loop of prepareLights (:272-283): castsShadow() entities get numPasses(shadow caster)
 *          draws into list 0.
 * A draw is (entity id, pass index). out_draws0/1 hold 2 uint32 per draw; returns counts through out_counts[2]. */
void orc_draw_lists(int64_t n_visible, const uint32_t *visible, const uint32_t *pass_counts, const uint32_t *flags,
                    int32_t mode, uint32_t *out_draws0, uint32_t *out_draws1, int64_t out_counts[2]);
/* ---- per-entity uniforms (section 8(f) row 2): EntityUniforms{mat4 transform; vec3 position} (render_entity.h:34-37),
 * filled as RenderEntity::setTransform does (render_entity.cc:67-69). 20 floats per entity (std140: 80 bytes). */
void orc_entity_uniforms(int64_t n, const float *world_trs, float *out20);
/* The same blocks read back from the uniform buffer of a RenderEntity after setTransform + getResources (the reference
 * tier goes through UniformBuffer::flush into the headless GPU buffer; the port fills the bytes itself). 80 B each. */
void orc_entity_uniform_blocks(int64_t n, const float *world_trs, void *out80);
/* glm quat (w x y z) * vec3, glm/gtc/quaternion.inl:313-321 - how a Light gets its world direction from its entity's
 * world orientation (graphics/src/light.cc:105). */
void orc_quat_rotate(const float q[4], const float v[3], float out[3]);
/* sizeof(EntityUniforms), sizeof(ViewUniforms), sizeof(LightUniforms): 80, 288, 208. */
void orc_uniform_sizes(int32_t out[3]);
/* ViewUniforms (render/include/render_view.h:28-36) of a view whose viewport starts at viewport_pos: 288 bytes =
 * view, projection, viewProjection, inverseViewProjection (4 x mat4), ivec2 viewportPosition, ivec2 viewportSize,
 * vec3 position, 4 bytes padding (render_view.cc:51-58,74-90,102-135). */
void orc_view_uniforms(const OrcView *view, const int32_t viewport_pos[2], void *out288);
/* LightUniforms (render/include/render_light.h:33-47), 208 bytes, in the state the graphics Light component leaves a
 * RenderLight in (graphics/src/light.cc:35-45,58-74,100-108): floats [0..2] position, [3] intensity, [4..6] normalised
 * direction, [7] cosCutoff (spot), [8..10] colour, [11] range, [12..27] volumeTransform (point, spot: render_light.cc:
 * 219-242), [28..43] shadowSpace (shadow-casting spot: :284-290), [44] shadowZNear (shadow-casting point: :324),
 * [45..47] attenuation (point, spot), [48] shadowBiasConstant, [49..51] padding. Fields a type never writes are 0. */
void orc_light_uniforms(const OrcLight *light, const float attenuation[3], float shadow_bias, void *out208);

const char *orc_name(void);   /* "port" or "reference" */

/* ---- deferred light shading and colour post effects (section 8(f) row 4) - oracle/ovis_shading.c, which states what is
 * restated from the GLSL and what this oracle has to define itself. Buffers are row-major, width x height:
 *   buffer_a/b/c  4 floats per pixel: encoded normal (n * 0.5 + 0.5) | diffuse rgb | specular rgb + shininess
 *   depth         1 float per pixel (0..1)
 *   view_uniforms the 288-byte ViewUniforms block; light_uniforms n_lights x 52 floats (LightUniforms blocks);
 *   light_types   ORC_LIGHT_* per light (the reference compiles one shader variant per type);
 *   shadow_maps   NULL, or per light NULL / a depth image: shadow_size^2 floats (spot) or 6 x shadow_size^2 (point).
 * out_rgba: 4 floats per pixel - the sum of the lights' contributions in list order, alpha 1. */
void orc_shade_deferred(int32_t width, int32_t height, const float *buffer_a, const float *buffer_b, const float *buffer_c,
                        const float *depth, const void *view_uniforms, int32_t n_lights, const float *light_uniforms,
                        const int32_t *light_types, const float *const *shadow_maps, int32_t shadow_size, float *out_rgba);
void orc_tonemap(int64_t n_pixels, const float *rgba, float exposure, float white_point, float *out_rgba);
void orc_gamma(int64_t n_pixels, const float *rgba, float gamma, float *out_rgba);

void orc_raster_vertex(const float *transform, const float *view_projection, const float *position, const float *normal,
                       float *out_clip, float *out_normal);
void orc_raster_fragment(const float *normal, const float *material, float *out_a, float *out_b, float *out_c);
/* ovis_raster.c: the G-buffer pass / a shadow-map pass over the given instances (see the file's header) */
void orc_raster(int64_t n_instances, const float *matrix, const int64_t *mesh, const float *material, const float *vertices,
                const uint32_t *indices, const float *vp, int32_t width, int32_t height, float *out_a, float *out_b, float *out_c,
                float *out_depth);

void orc_raster_basic(int64_t n_instances, const float *matrix, const int64_t *mesh, const float *material, const float *vertices,
                      const uint32_t *indices, const float *vp, int32_t width, int32_t height, float *rgba, float *depth_buffer);

void orc_raster_bitangent(const float *world_normal, const float *world_tangent, float handedness, float *out);
/* textures of the rasteriser's restatement: records of { const uint8_t *rgba; int32_t width, height, srgb, wrap; } */
void orc_texture_sample(const void *texture, float u, float v, float out[4]);
void orc_raster_fragment_textured(const float *normal, const float *uv, const float *tangent, const float *bitangent, const float *material,
                                  const void *textures, float *out_a, float *out_b, float *out_c);
void orc_raster_textured(int64_t n_instances, const float *matrix, const int64_t *mesh, const float *material, const float *vertices,
                         const float *tangents, const uint32_t *indices, const void *textures, const float *vp, int32_t width,
                         int32_t height, float *out_a, float *out_b, float *out_c, float *out_depth);

#ifdef __cplusplus
}
#endif

#endif
