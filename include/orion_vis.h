/*
 * orion_vis.h - C ABI of the B200 scene-visibility library (liborion_vis.so).
 *
 * This is the drop-in boundary for the one hot path this project accelerates: Orion's per-frame
 * scene visibility (world-transform propagation + world AABB recompute, frustum culling of render
 * entities and lights per view, per-light shadow-caster lists). Each entry point names the reference
 * interface it replaces; paths are relative to /root/reference/engine/runtime/ unless they start
 * with engine/ or apps/. INTEGRATION.md shows the C++ adapter (CudaRenderWorld : RenderWorld) a
 * maintainer drops in at graphics/src/graphics_system.cc:33-36.
 *
 * Rules of the ABI
 *   - plain C types only; every function returns an OvisStatus (0 = ok) and never throws or exits;
 *     ovis_last_error() returns a thread-local description of the last failure.
 *   - one host thread per world (the engine is single-threaded).
 *   - there is NO CPU fallback: without a CUDA device every call that needs one fails with
 *     OVIS_ERR_CUDA.
 *   - entity / node / light ids are stable uint32 handles, assigned in insertion order. Every result
 *     list is emitted in insertion order of the ids it holds, which is the order the reference's
 *     std::list walk produces (render/src/simple_render_world.cc:38-41,53-55,65-67).
 *
 * Layouts (all little-endian fp32 / uint32)
 *   TRS    10 floats: position xyz, orientation w x y z, scale xyz       (core/include/math/transform.h:55-61)
 *   AABB    6 floats: minimum xyz, maximum xyz                            (core/include/math/bounding_box.h:28-31)
 *   planes 24 floats: 6 x (normal xyz, distance), Left Right Top Bottom Near Far (core/include/math/frustum.h:39-47)
 */
#ifndef ORION_VIS_H
#define ORION_VIS_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct OvisWorld OvisWorld;

typedef enum OvisStatus {
    OVIS_OK = 0,
    OVIS_ERR_INVALID = 1,     /* bad argument (null pointer, unknown id, unsupported flag mask ...) */
    OVIS_ERR_CUDA = 2,        /* CUDA runtime error or no device */
    OVIS_ERR_CAPACITY = 3,    /* caller-provided output buffer too small; required size is reported */
    OVIS_ERR_NOMEM = 4
} OvisStatus;

/* RenderLight::Type, render/include/render_light.h:52-58 */
enum { OVIS_LIGHT_AMBIENT = 0, OVIS_LIGHT_DIRECTIONAL = 1, OVIS_LIGHT_POINT = 2, OVIS_LIGHT_SPOT = 3 };
/* RenderEntity::kCastsShadow (render/include/render_entity.h:50-53), RenderLight::kCastsShadows (render_light.h:61-64) */
enum { OVIS_CASTS_SHADOW = 1u };
/* RenderWorld::CullFlags::kCullLights, render/include/render_world.h:43-46 */
enum { OVIS_CULL_LIGHTS = 1u };
#define OVIS_NO_PARENT (-1)

/* Parameters a RenderLight is configured with through its setters (render/src/render_light.cc:67-142). */
typedef struct OvisLight {
    int32_t  type;
    uint32_t flags;
    float    position[3];
    float    direction[3];    /* normalised on use, as RenderLight::setDirection does (:78-85) */
    float    colour[3];
    float    intensity;
    float    range;
    float    cutoff;          /* degrees, spot lights, <= 45 (:104) */
} OvisLight;                   /* 56 bytes */

/* Parameters a RenderView is configured with: setTransform / perspective / setViewport
 * (render/src/render_view.cc:51-90). aspect = (float)vp_width / (float)vp_height (:84). */
typedef struct OvisView {
    float   position[3];
    float   orientation[4];   /* w x y z */
    float   hfov_deg;
    float   znear;
    float   zfar;
    int32_t vp_width;
    int32_t vp_height;
} OvisView;                    /* 48 bytes */

const char *ovis_last_error(void);
/* Library build description, e.g. "orion_vis sm_100a fmad=off". */
const char *ovis_version(void);

/* ------------------------------------------------------------------------------------------------
 * World lifetime - replaces `new SimpleRenderWorld` (graphics/src/graphics_system.cc:33-36,
 * render/include/simple_render_world.h:29-53). All device memory of the world lives on `device`.
 * ---------------------------------------------------------------------------------------------- */
int ovis_world_create(int device, OvisWorld **out_world);
int ovis_world_destroy(OvisWorld *world);
/* Pre-size device storage (optional; storage otherwise grows geometrically). */
int ovis_world_reserve(OvisWorld *world, int64_t n_entities, int64_t n_nodes);
/* All kernels and copies of the world are issued on this cudaStream_t (default: the legacy stream). */
int ovis_world_set_stream(OvisWorld *world, void *cuda_stream);
/* The stream the world's work is ordered on and the device it lives on (either output may be NULL). */
int ovis_world_get_stream(OvisWorld *world, void **out_cuda_stream, int32_t *out_device);
/* Sharded worlds: this world holds the contiguous id range starting at id_base
 * of a larger scene; ids reported by the entity culls are local id + id_base, so the per-shard lists
 * concatenated in rank order are the single-world list of simple_render_world.cc:38-41. Ids passed
 * INTO the API stay local. Default 0. */
int ovis_world_set_id_base(OvisWorld *world, uint32_t id_base);
int ovis_world_get_id_base(OvisWorld *world, uint32_t *out_id_base);
int ovis_world_counts(OvisWorld *world, int64_t *n_entities, int64_t *n_nodes, int32_t *n_lights);
/* Live entities vs entity slots: removed entities keep their slot (ids are stable handles), so a world with spawn /
 * despawn churn grows, and as long as one slot is dead every cull reads the alive mask. */
int ovis_world_live_counts(OvisWorld *world, int64_t *out_live_entities, int64_t *out_entity_slots);
/* Reclaims the dead slots: the live entities move together, IN ORDER (the list order of simple_render_world.cc:38-41 is
 * insertion order and stays it), and get the ids 0 .. new_slots - 1. out_remap[old id] = new id, or 0xffffffff for a
 * removed entity (cap_remap >= the old slot count; NULL with cap 0 if the caller keeps no ids). Every id held outside the
 * library must be translated. The reference's std::list shrinks on remove (simple_render_world.cc:65-67); this is the
 * same effect, taken when the caller decides (e.g. more than a quarter of the slots dead). O(entities) on the device. */
int ovis_world_compact(OvisWorld *world, int64_t cap_remap, uint32_t *out_remap, int64_t *out_old_slots, int64_t *out_new_slots);

/* ------------------------------------------------------------------------------------------------
 * Render entities - RenderWorld::addEntity / updateEntity / removeEntity
 * (render/include/render_world.h:69-83; callers render/src/render_entity.cc:52-97).
 * ---------------------------------------------------------------------------------------------- */
/* RenderEntity::setWorld -> addEntity (render_entity.cc:52-60) for n entities; local AABBs as
 * RenderEntity::setBoundingBox (:76-80), flags as setFlags (render_entity.h:64). Ids are
 * first_id .. first_id + n - 1. An entity has no world bounds until it is given a transform
 * (ovis_entity_set_transforms / ovis_node_bind_entities) or explicit bounds (ovis_entity_set_world_aabbs);
 * until then its world AABB is its local AABB (identity transform, Transform::Transform() transform.h:64-70). */
int ovis_entity_add(OvisWorld *world, int64_t n, const float *local_aabb, const uint32_t *flags, uint32_t *out_first_id);
/* The batched setters below (and ovis_node_set_local_trs) check EVERY id before they change anything: one id that names
 * no live entity refuses the whole call. A batch is a parallel scatter on the device: each id should occur once in it; if
 * an id is repeated, one of its values is stored, unspecified which (the reference's one-setter-call-at-a-time order has
 * no counterpart inside a batch - callers that coalesce a frame's changes, like the adapter, keep the last one). */
/* RenderEntity::~RenderEntity -> setWorld(nullptr) -> removeEntity (render_entity.cc:46-60;
 * simple_render_world.cc:65-67): survivors keep their relative order. */
int ovis_entity_remove(OvisWorld *world, int64_t n, const uint32_t *ids);
/* RenderEntity::setBoundingBox (render_entity.cc:76-80) */
int ovis_entity_set_local_aabbs(OvisWorld *world, int64_t n, const uint32_t *ids, const float *local_aabb);
/* RenderEntity::setFlags (render_entity.h:64) */
int ovis_entity_set_flags(OvisWorld *world, int64_t n, const uint32_t *ids, const uint32_t *flags);
/* RenderEntity::setTransform with a WORLD transform (render_entity.cc:64-72; caller
 * graphics/src/renderer.cc:62-66). World AABBs are recomputed on the device by the next
 * ovis_update_transforms (or implicitly by the next cull). */
int ovis_entity_set_transforms(OvisWorld *world, int64_t n, const uint32_t *ids, const float *world_trs);
/* Static geometry: give n consecutive entities their WORLD AABBs directly (what updateWorld,
 * render_entity.cc:90-93, would have produced), bypassing the transform stage. */
int ovis_entity_set_world_aabbs(OvisWorld *world, uint32_t first_id, int64_t n, const float *world_aabb);
/* Same for scattered ids: what a RenderWorld::updateEntity implementation forwards when the RenderEntity has
 * already recomputed worldBoundingBox() on the host (render_entity.cc:90-97). */
int ovis_entity_set_world_aabbs_by_id(OvisWorld *world, int64_t n, const uint32_t *ids, const float *world_aabb);

/* ------------------------------------------------------------------------------------------------
 * Transform tree - Entity local/world transforms (engine/include/entity.h:45-190,
 * engine/src/entity.cc:309-363 setters, :382-411 transformed()).
 * ---------------------------------------------------------------------------------------------- */
/* Add n nodes. parent[i] is an existing node id, an id assigned earlier in this same call, or
 * OVIS_NO_PARENT for a child of the (identity) root. Ids are first_id .. first_id + n - 1. */
int ovis_node_add(OvisWorld *world, int64_t n, const int32_t *parent, const float *local_trs, uint32_t *out_first_id);
/* Entity::setPosition / setOrientation / setScale, batched (entity.cc:309-363). */
int ovis_node_set_local_trs(OvisWorld *world, int64_t n, const uint32_t *ids, const float *local_trs);
/* Renderer component on an Entity: its RenderEntities follow the node's world transform
 * (graphics/src/renderer.cc:62-66). */
int ovis_node_bind_entities(OvisWorld *world, int64_t n, const uint32_t *node_ids, const uint32_t *entity_ids);
/* Entity::transformed over everything that changed since the last call (entity.cc:382-411), then
 * Transform::matrix + BoundingBox::transform for every affected entity (core/src/math/transform.cc:26-35,
 * core/src/math/bounding_box.cc:27-40, render_entity.cc:90-93). Asynchronous on the world's stream. */
/* Entity::addChild / re-parenting (engine/src/entity.cc:210-236): `node` becomes a child of `new_parent` (OVIS_NO_PARENT =
 * a root). A node may not become a descendant of itself. Everything below is recomputed at the next update. */
int ovis_node_set_parent(OvisWorld *world, uint32_t node, int32_t new_parent);
/* Entity::destroy (engine/src/entity.cc:58-90): the node and its whole subtree leave the tree, and every render entity
 * bound to one of them is removed as by ovis_entity_remove (destroying an Entity destroys its components, whose
 * RenderEntities leave the world: graphics/src/renderer.cc:87-94). Lights bound to a removed node stay where they are and
 * follow nothing. Node ids are not reused. Cost: O(nodes) on the host. */
int ovis_node_remove(OvisWorld *world, uint32_t node, int64_t *out_removed_nodes, int64_t *out_removed_entities);
/* A light follows a node as graphics/src/light.cc:100-108 (Light::transformed) does: after every transform update its
 * position is the node's world position and its direction the node's world orientation applied to (0, 0, -1).
 * OVIS_NO_PARENT unbinds. Costs one small device-to-host copy per update while any light is bound. */
int ovis_light_bind_node(OvisWorld *world, uint32_t light, int32_t node);
int ovis_update_transforms(OvisWorld *world);

/* ------------------------------------------------------------------------------------------------
 * Lights - RenderWorld::addLight / updateLight / removeLight (render_world.h:85-95).
 * ---------------------------------------------------------------------------------------------- */
int ovis_light_add(OvisWorld *world, int32_t n, const OvisLight *lights, uint32_t *out_first_id);
int ovis_light_update(OvisWorld *world, uint32_t id, const OvisLight *light);
int ovis_light_remove(OvisWorld *world, uint32_t id);
/* RenderLight::setAttenuation / setShadowBias (render/src/render_light.cc:130-159): the LightUniforms fields that
 * OvisLight does not carry. A new light has its component's defaults (graphics/src/light.cc:44,63,72): bias -0.001,
 * attenuation (1, 0.045, 0.0075) for point and (1, 0.09, 0.032) for spot lights. */
int ovis_light_set_shading(OvisWorld *world, uint32_t light_id, const float attenuation[3], float shadow_bias_constant);

/* ------------------------------------------------------------------------------------------------
 * Host-side view / light set-up (O(views); kept on the host so the planes are built by the same
 * libm and the same operation order as the reference).
 * ---------------------------------------------------------------------------------------------- */
/* RenderView::updateMatrices + Frustum::update (render/src/render_view.cc:102-135, core/src/math/frustum.cc:27-67) */
int ovis_view_planes(const OvisView *view, float out_planes[24]);
/* Same, also returning view, projection, viewProjection, inverseViewProjection (4 x 16 floats,
 * column-major) and the 8 frustum corners (frustum.cc:70-84). Either output may be NULL. */
int ovis_view_matrices(const OvisView *view, float out_mats[64], float out_corners[24]);
/* RenderLight::updateVolumeTransform, spot branch (render/src/render_light.cc:227-249) */
int ovis_light_spot_aabb(const OvisLight *light, float out_aabb[6]);
/* RenderLight::updateShadowViews (render_light.cc:258-332) with the square shadow-map viewport of
 * render/src/deferred_render_pipeline.cc:261-265. *out_count = 0, 1 (spot) or 6 (point). */
int ovis_light_shadow_views(const OvisLight *light, int32_t shadow_map_resolution, OvisView out_views[6], int32_t *out_count);
/* The uniform blocks the reference uploads per view and per light, computed on the host
 * with the reference's operation order and libm calls (O(views), O(lights) work):
 *   ViewUniforms  (render/include/render_view.h:28-36), 288 bytes: mat4 view, projection, viewProjection,
 *                 inverseViewProjection; ivec2 viewportPosition, viewportSize; vec3 position; 4 bytes padding - filled
 *                 by render_view.cc:55,79-80,113-133. viewport_pos NULL = (0, 0).
 *   LightUniforms (render/include/render_light.h:33-47), 208 bytes, as floats: [0..2] position, [3] intensity, [4..6]
 *                 normalised direction, [7] cosCutoff (spot), [8..10] colour, [11] range, [12..27] volumeTransform
 *                 (point, spot: render_light.cc:219-242), [28..43] shadowSpace = bias * projection * view of the shadow
 *                 view (shadow-casting spot: :284-290), [44] shadowZNear (shadow-casting point: :324), [45..47]
 *                 attenuation (point, spot: :136-138), [48] shadowBiasConstant (:158), [49..51] padding - the state the
 *                 graphics Light component leaves a RenderLight in (graphics/src/light.cc:35-45,58-74,100-108). The
 *                 block starts zeroed (render_core/src/uniform_buffer.cc:136-137); what a type never writes stays 0.
 *                 `range` is copied for every type (the reference copies an uninitialised member for ambient and
 *                 directional lights, render_light.cc:336). */
int ovis_view_uniforms(const OvisView *view, const int32_t viewport_pos[2], void *out288);
int ovis_light_uniforms(const OvisLight *light, const float attenuation[3], float shadow_bias_constant, void *out208);

/* ------------------------------------------------------------------------------------------------
 * Culling - RenderWorld::cull (render/include/render_world.h:69-77; render/src/simple_render_world.cc:37-49).
 * Pending transform updates are flushed first.
 * ---------------------------------------------------------------------------------------------- */
/* cull(view, results, flags) into HOST buffers. out_entities receives ids of entities whose world
 * AABB intersects the view frustum (core/src/math/intersection.cc:42-57); when flags has
 * OVIS_CULL_LIGHTS, out_lights receives ids of lights for which RenderLight::cull is false
 * (render_light.cc:187-209). Counts are always written; if a capacity is too small the call returns
 * OVIS_ERR_CAPACITY and the content of that list is unspecified. Nothing beyond a buffer's capacity is ever
 * written; entries between the count and the capacity are unspecified (a small list comes over in one
 * speculative copy together with its count). out_lights may be NULL when OVIS_CULL_LIGHTS is not set. */
int ovis_cull(OvisWorld *world, const OvisView *view, uint32_t flags,
              uint32_t *out_entities, int64_t cap_entities, int64_t *out_n_entities,
              uint32_t *out_lights, int32_t cap_lights, int32_t *out_n_lights);
/* Entity cull against explicit planes; require = 0 or OVIS_CASTS_SHADOW (the castsShadow() filter of
 * deferred_render_pipeline.cc:273 folded into the pass). HOST output. */
int ovis_cull_planes(OvisWorld *world, const float planes[24], uint32_t require,
                     uint32_t *out_entities, int64_t cap_entities, int64_t *out_n_entities);
/* Same, DEVICE output, asynchronous on the world's stream: d_out_entities must hold one uint32 per
 * live entity slot (ovis_world_counts), d_out_count is one uint32 in device memory. */
int ovis_cull_planes_device(OvisWorld *world, const float planes[24], uint32_t require,
                            uint32_t *d_out_entities, uint32_t *d_out_count);

/* Sharded worlds: the same cull with the exchange fused in. The list (at most `capacity`
 * ids, clipped beyond) and its true count are stored to n_dst destinations (1..9): destination 0 is local, the others
 * are typically the peers' segments for this shard, i.e. peer-GPU memory mapped over NVLink (CUDA IPC / symmetric
 * memory), written by the gather kernel itself with plain stores, so no separate all-gather of the visible lists is
 * needed - only a barrier before the peers read. No counterpart in the reference (single process, single device). */
int ovis_cull_planes_device_multi(OvisWorld *world, const float planes[24], uint32_t require, int64_t capacity,
                                  int32_t n_dst, uint32_t *const *d_out_entities, uint32_t *const *d_out_count);

/* The same with one-way ready flags instead of a barrier, so that nothing but the two cull kernels is enqueued per
 * frame (and back-to-back frames keep overlapping). Before the gather kernel may store, each of the n_wait (<= 32) words
 * at d_wait_flags (local memory that the peers write) must be >= wait_value - i.e. the peers have finished with the
 * buffer being refilled; one warp of the test kernel checks that, hidden behind the streaming pass. After every gather
 * block has stored, signal_value is written to each of the n_signal words d_signal_flags[x] (typically this shard's slot
 * in every rank's flag array, peers' over NVLink). A reader calls ovis_wait_flags (one tiny kernel on the world's stream
 * that spins until all n words are >= value) or ovis_read_segments_host before it touches the segments. Waits give up
 * after ~2 s and raise the world's exchange error (ovis_world_exchange_error) rather than hang. Frame numbers compare
 * modulo 2^32 (signed difference). No counterpart in the reference. */
int ovis_cull_planes_device_exchange(OvisWorld *world, const float planes[24], uint32_t require, int64_t capacity,
                                     int32_t n_dst, uint32_t *const *d_out_entities, uint32_t *const *d_out_count,
                                     int32_t n_wait, const uint32_t *d_wait_flags, uint32_t wait_value,
                                     int32_t n_signal, uint32_t *const *d_signal_flags, uint32_t signal_value);
int ovis_wait_flags(OvisWorld *world, const uint32_t *d_flags, int32_t n, uint32_t value);
/* The host as the reader of a sharded exchange: waits for each segment's flag (d_flags[r] >= value; NULL = no wait) and
 * delivers the n_segments segments - [count | ids...] blocks of segment_words uint32 each, the layout the exchange
 * writes - into host_out with the same layout, then synchronises the world's stream. With page-locked host memory
 * (ovis_host_alloc, cudaHostAlloc) this is ONE kernel that stores each segment's count word and exactly `count` ids
 * into host memory as soon as that segment has landed; with pageable memory it is a wait kernel and a copy of the
 * whole buffer. Words of host_out beyond 1 + count of a segment are not written. No counterpart in the reference. */
int ovis_read_segments_host(OvisWorld *world, const uint32_t *d_flags, uint32_t value, const uint32_t *d_segments,
                            int64_t segment_words, int32_t n_segments, uint32_t *host_out);
int ovis_world_exchange_error(OvisWorld *world, int32_t *out_error);

/* ---- one world over several GPUs, one host process (SURVEY.md section 8(b) "ovis_world_create(device_ids[], n, ...)",
 * section 8(e)) ------------------------------------------------------------------------------------------------------
 * The reference is one process and one thread (engine/src/engine.cc:147-195): its RenderWorld replacement cannot ask
 * for a process per GPU. An OvisShardedWorld owns one OvisWorld per listed device plus one worker thread per device, so
 * the caller's single thread still starts all shards' culls side by side. Entities are split into contiguous id ranges
 * [r N / G, (r + 1) N / G); shard r reports global ids, so the shards' lists read in shard order are the reference's
 * list (insertion order, simple_render_world.cc:38-41). Listing a device twice is allowed (two shards on one GPU).
 * ovis_sharded_shard gives the shard's ordinary world for everything else (flags, removal, lights, transform trees whose
 * nodes live on that shard). No counterpart in the reference. */
typedef struct OvisShardedWorld OvisShardedWorld;
int ovis_sharded_create(const int *device_ids, int32_t n_devices, OvisShardedWorld **out_world);
int ovis_sharded_destroy(OvisShardedWorld *world);
int ovis_sharded_count(OvisShardedWorld *world, int32_t *out_shards, int64_t *out_entities);
int ovis_sharded_shard(OvisShardedWorld *world, int32_t shard, OvisWorld **out_world, int64_t *out_id_base);
/* The whole entity set at once (world-space AABBs n x 6, flags n or NULL): split, uploaded, id bases set. Once per world. */
int ovis_sharded_add_entities(OvisShardedWorld *world, int64_t n, const float *world_aabb, const uint32_t *flags);
/* SimpleRenderWorld::cull over all shards for a HOST consumer: every device's gather kernel stores its segment straight
 * into one page-locked buffer all devices map (no GPU-to-GPU exchange, no copy kernel); out_entities receives the merged
 * list, dense and in id order; out_shard_counts (optional, one per shard) the per-shard counts. OVIS_ERR_CAPACITY (with
 * *out_n_entities = required) if cap_entities is too small. */
int ovis_sharded_cull(OvisShardedWorld *world, const OvisView *view, uint32_t require, uint32_t *out_entities,
                      int64_t cap_entities, int64_t *out_n_entities, int64_t *out_shard_counts);
int ovis_sharded_cull_planes(OvisShardedWorld *world, const float planes[24], uint32_t require, uint32_t *out_entities,
                             int64_t cap_entities, int64_t *out_n_entities, int64_t *out_shard_counts);
/* Lights are replicated: every shard's world receives the same lights under the same ids (render_world.h:69-72). */
int ovis_sharded_add_lights(OvisShardedWorld *world, int32_t n, const OvisLight *lights, uint32_t *out_first_id);
/* ovis_cull_views / ovis_build_shadow_lists over all shards (BASELINE configs[4]: "then 5 views"): every shard runs the
 * many-view pass on its entities into its own page-locked CSR, the caller's thread splices list v of shard 0, 1, ... one
 * after the other - per view that is the single-world list, ids global and ascending. Same arguments, outputs and
 * OVIS_ERR_CAPACITY behaviour as the single-world calls. */
int ovis_sharded_cull_views(OvisShardedWorld *world, int32_t n_views, const float *planes, const uint32_t *require,
                            int64_t *out_offsets, int64_t cap_entities, uint32_t *out_entities, int64_t *out_n_entities);
int ovis_sharded_build_shadow_lists(OvisShardedWorld *world, const OvisView *main_view, int32_t shadow_map_resolution,
                                    int32_t cap_lists, int32_t *out_list_light, int32_t *out_list_view, int64_t *out_offsets,
                                    int32_t *out_n_lists, int64_t cap_entities, uint32_t *out_entities, int64_t *out_n_entities);
/* The page-locked segments of the last ovis_sharded_cull*: shard r's [count | ids...] at segments + r * segment_words. */
int ovis_sharded_segments(OvisShardedWorld *world, const uint32_t **out_segments, int64_t *out_segment_words);
/* For DEVICE consumers: the local culls, then one ncclAllGather per device of fixed-size [count | ids...] blocks of
 * 1 + capacity words (the collective north_star names; NCCL is loaded at run time). On return (*out_d_blocks)[d] is, on
 * device d, the n_shards blocks in shard order. A count above `capacity` means that block was clipped: call again with
 * more. Needs distinct devices (one NCCL communicator). */
int ovis_sharded_allgather_device(OvisShardedWorld *world, const float planes[24], uint32_t require, int64_t capacity,
                                  const uint32_t *const **out_d_blocks, int64_t *out_shard_counts);

/* Entity cull for callers that cannot allocate GPU memory - a render thread in the middle of a frame, or an
 * engine that forbids cudaMalloc/cudaFree after start-up because they synchronise the device, can fail under
 * a fixed VRAM budget, and bypass its own pool. No counterpart in the reference. Same result as
 * ovis_cull_planes_device (ids ascending, id_base added, *d_out_count = number of visible entities even if
 * `capacity` clipped the writes), but it needs no OvisWorld and allocates nothing: the caller passes the
 * world AABBs as device memory in the library's SoA layout (component k of entity i at
 * d_world_aabb[k * stride + i], k = min xyz then max xyz; stride a multiple of 4 and >= n_entities rounded
 * up to 256; d_world_aabb 16-byte aligned, because the kernels read 128-bit vectors - a misaligned pointer is
 * refused with OVIS_ERR_INVALID rather than left to fault on the device) and the device outputs, and the kernels'
 * scratch is a fixed ~4.2 MB array inside the library
 * image (present on every GPU that loads the library, until exit). Limits and rules: at most
 * OVIS_NOALLOC_MAX_ENTITIES entities per call (more -> OVIS_ERR_CAPACITY); all calls on a GPU share that one
 * scratch, so calls whose kernels could overlap on the GPU (different streams or threads) race - keep them on
 * one stream or serialise them; the first call on each device resolves the scratch address and launch
 * geometry, so make that first call before starting any stream capture. Asynchronous on `cuda_stream`. */
#define OVIS_NOALLOC_MAX_ENTITIES 1048576
int ovis_cull_planes_noalloc(int device, void *cuda_stream, const float *d_world_aabb, int64_t stride,
                             int64_t n_entities, const float planes[24], uint32_t id_base,
                             uint32_t *d_out_entities, int64_t capacity, uint32_t *d_out_count);

/* Many views in one pass over the bounds (main camera + cascades + every shadow view:
 * deferred_render_pipeline.cc:147,259-284). planes = n_views x 24 floats, require = n_views masks (NULL = 0).
 * Results stay on the device in CSR form and are valid until the next call on this world:
 * *d_offsets -> n_views + 1 int64, *d_entities -> offsets[n_views] uint32 (insertion order per view).
 * host_offsets (optional, n_views + 1) receives a synchronous copy of the offsets. */
int ovis_cull_views_device(OvisWorld *world, int32_t n_views, const float *planes, const uint32_t *require,
                           const int64_t **d_offsets, const uint32_t **d_entities, int64_t *host_offsets);

/* The same pass with HOST output: out_offsets receives n_views + 1 entries, out_entities the lists back to back.
 * OVIS_ERR_CAPACITY (with *out_n_entities = required) if cap_entities is too small. */
int ovis_cull_views(OvisWorld *world, int32_t n_views, const float *planes, const uint32_t *require, int64_t *out_offsets,
                    int64_t cap_entities, uint32_t *out_entities, int64_t *out_n_entities);

/* The shadow-view culls of prepareLights (deferred_render_pipeline.cc:259-284) for an explicit set of lights,
 * regardless of their visibility: every shadow view (RenderLight::updateShadowViews, render_light.cc:258-332; six
 * cube faces per point light, one per spot light, none otherwise or without OVIS_CASTS_SHADOW) is culled for
 * castsShadow() entities in one multi-view pass set. List k belongs to light out_list_light[k], view
 * out_list_view[k]; results stay on the device in CSR form as for ovis_cull_views_device (host_offsets optional,
 * n_lists + 1 entries). OVIS_ERR_CAPACITY (with *out_n_lists = required) if cap_lists is too small. */
int ovis_cull_shadow_views_device(OvisWorld *world, int32_t n_lights, const uint32_t *light_ids, int32_t shadow_map_resolution,
                                  int32_t cap_lists, int32_t *out_list_light, int32_t *out_list_view, int32_t *out_n_lists,
                                  const int64_t **d_offsets, const uint32_t **d_entities, int64_t *host_offsets);

/* The same culls with HOST output and a choice of filter, for callers that serve the reference's per-view
 * RenderWorld::cull(shadowView, results, 0) calls from one batched pass (adapter/cuda_render_world.cc): `require` = 0 keeps
 * every entity that intersects the view, as SimpleRenderWorld::cull does (the pipeline filters castsShadow() itself,
 * deferred_render_pipeline.cc:272-283); OVIS_CASTS_SHADOW keeps the casters only. OVIS_ERR_CAPACITY (with *out_n_lists /
 * *out_n_entities = required) if a buffer is too small. */
int ovis_cull_shadow_views(OvisWorld *world, int32_t n_lights, const uint32_t *light_ids, int32_t shadow_map_resolution, uint32_t require,
                           int32_t cap_lists, int32_t *out_list_light, int32_t *out_list_view, int64_t *out_offsets, int32_t *out_n_lists,
                           int64_t cap_entities, uint32_t *out_entities, int64_t *out_n_entities);

/* DeferredRenderPipeline::prepareLights' cull+filter loop (deferred_render_pipeline.cc:207,247,257-284):
 * for every light not culled by main_view that casts shadows (spot, point), one list per shadow view
 * holding the ids of the castsShadow() entities intersecting it. HOST output:
 * out_list_light[k] / out_list_view[k] name list k, out_offsets has n_lists + 1 entries.
 * Returns OVIS_ERR_CAPACITY (with *out_n_lists / *out_n_entities = required sizes) if a buffer is too small. */
int ovis_build_shadow_lists(OvisWorld *world, const OvisView *main_view, int32_t shadow_map_resolution,
                            int32_t cap_lists, int32_t *out_list_light, int32_t *out_list_view,
                            int64_t *out_offsets, int32_t *out_n_lists,
                            int64_t cap_entities, uint32_t *out_entities, int64_t *out_n_entities);

/* ------------------------------------------------------------------------------------------------
 * Read-back for parity checks (RenderEntity::worldBoundingBox / Entity::worldTransform /
 * RenderEntity::transform().matrix()). HOST output, synchronous.
 * ---------------------------------------------------------------------------------------------- */
int ovis_get_world_aabbs(OvisWorld *world, uint32_t first_entity, int64_t n, float *out_aabb);
int ovis_get_node_world_trs(OvisWorld *world, uint32_t first_node, int64_t n, float *out_trs);

/* ------------------------------------------------------------------------------------------------
 * Consumers of the visible lists. The reference runs these on
 * the CPU right after culling and hands the results to GL/Vulkan; here the lists and the uniform
 * block stay in headless device buffers (graphics API: unavailable on this box).
 * ---------------------------------------------------------------------------------------------- */
/* One draw call of a draw list: DrawList::add (render/src/draw_list.cc:35-45) appends, for a given
 * entity and pass type, one EntityDrawCall per shader pass of that type (passIndex 0, 1, ...). */
typedef struct OvisDraw {
    uint32_t entity;      /* entity id (global id for a sharded world) */
    uint32_t pass_index;  /* index of the shader pass within its pass type */
} OvisDraw;

/* Shader-pass counts of an entity's material, which the reference looks up per draw through
 * entity->shader()->numPasses(type) (draw_list.cc:41, deferred_render_pipeline.cc:322-325).
 * Packed word: numPasses(kDeferred) | numPasses(kBasic) << 8 | numPasses(kShadowCaster) << 16
 * (each 0..255; a new entity has 0 of everything and is never drawn). */
#define OVIS_PASS_COUNTS(deferred, basic, shadow_caster) \
    ((uint32_t)(deferred) | ((uint32_t)(basic) << 8) | ((uint32_t)(shadow_caster) << 16))
int ovis_entity_set_pass_counts(OvisWorld *world, int64_t n, const uint32_t *ids, const uint32_t *pass_counts);

#define OVIS_DRAWS_MAIN 0   /* prepareEntities: list 0 = deferred-pass draws, list 1 = basic-pass draws */
#define OVIS_DRAWS_SHADOW 1 /* shadow-caster loop: list 0 = shadow-caster draws of castsShadow() entities */
/* Expand a visible list into draw lists, in list order, on the world's stream (one or two small kernels).
 *   OVIS_DRAWS_MAIN    DeferredRenderPipeline::prepareEntities (render/src/deferred_render_pipeline.cc:
 *                      318-330): an entity whose shader has deferred passes goes to list 0 with one
 *                      draw per deferred pass, otherwise one draw per basic pass goes to list 1
 *                      (no pass of either type: not drawn).
 *   OVIS_DRAWS_SHADOW  the per-shadow-view loop of prepareLights (:272-283): castsShadow() entities
 *                      get one draw per shadow-caster pass in list 0; list 1 stays empty.
 * d_visible/d_visible_count: a device list of entity ids and its device count, e.g. what
 * ovis_cull_planes_device wrote, or one view's slice of ovis_cull_views_device's CSR; max_visible is
 * an upper bound of the count known to the host (sizes the grids; < 0 = the world's entity count).
 * Both NULL = the entity list of this world's last ovis_cull / ovis_cull_planes.
 * Draws beyond a list's capacity are dropped, but d_draw_counts[0..1] (device; NULL = an internal
 * pair, readable through ovis_get_draw_counts) always receive the full draw counts, so a caller can
 * detect the overflow and retry. If the device count turns out larger than max_visible (the bound was wrong) only the
 * first max_visible entries, rounded up to 1024, are expanded and both counts read 0xffffffff. Never allocates after the
 * world's arrays exist.
 * Ids in a caller's device list that do not name a slot of this world (stale after ovis_world_compact, another shard's)
 * index nothing: here they expand to no draw, ovis_pack_entity_uniforms_device / ovis_pack_light_uniforms_device write
 * a zero block for them, ovis_shade_deferred_device skips such a light, the rasteriser such an entity. */
int ovis_build_draw_lists_device(OvisWorld *world, const uint32_t *d_visible, const uint32_t *d_visible_count, int64_t max_visible,
                                 int32_t mode, OvisDraw *d_draws0, int64_t capacity0, OvisDraw *d_draws1, int64_t capacity1,
                                 uint32_t *d_draw_counts);
/* Host copy of the internal draw-count pair (synchronises the world's stream). */
int ovis_get_draw_counts(OvisWorld *world, uint32_t out_counts[2]);

/* EntityUniforms (render/include/render_entity.h:34-37: mat4 transform; vec3 position) of every
 * entity of a visible list, in list order, 20 floats (80 bytes, the std140 size) per entity:
 * [0..15] Transform::matrix() column-major, [16..18] position, [19] zero padding - what
 * RenderEntity::setTransform (render/src/render_entity.cc:64-72) writes into each entity's uniform
 * buffer. The reference rewrites the uniforms of every entity that moved; packing only the visible
 * ones from the world TRS already resident on the device keeps the traffic proportional to what is
 * drawn. Entities that were never given a transform get the identity (Transform's default).
 * d_uniforms must be 16-byte aligned with room for `capacity` entries; entries beyond it are dropped.
 * Pending transform updates are flushed first. List arguments as for ovis_build_draw_lists_device. */
int ovis_pack_entity_uniforms_device(OvisWorld *world, const uint32_t *d_visible, const uint32_t *d_visible_count, int64_t max_visible,
                                     float *d_uniforms, int64_t capacity);
/* LightUniforms blocks (208 bytes each, layout under ovis_light_uniforms) of the lights of a visible-light list, in list
 * order, gathered on the device from a per-light table the world keeps current (one small kernel). d_visible NULL = the
 * light list of this world's last ovis_cull with OVIS_CULL_LIGHTS. d_uniforms: 16-byte aligned, `capacity` blocks. */
int ovis_pack_light_uniforms_device(OvisWorld *world, const uint32_t *d_visible, const uint32_t *d_visible_count, int64_t max_visible,
                                    float *d_uniforms, int64_t capacity);
/* ViewUniforms blocks (288 bytes each) of n_views views into device memory: host math, one asynchronous upload on the
 * world's stream. viewport_pos: 2 ints per view, or NULL for (0, 0). */
int ovis_pack_view_uniforms_device(OvisWorld *world, int32_t n_views, const OvisView *views, const int32_t *viewport_pos,
                                   void *d_uniforms);

/* ------------------------------------------------------------------------------------------------
 * Deferred light pass and colour post effects over headless device buffers:
 * the arithmetic of the reference's shaders as CUDA kernels. The G-buffer and the shadow maps are plain device
 * arrays: filled by the headless rasteriser below (ovis_raster_*), or by the caller.
 * ---------------------------------------------------------------------------------------------- */
typedef struct OvisGBuffer {            /* row-major, width x height, device pointers, 16-byte aligned */
    int32_t width, height;
    const float *d_buffer_a;            /* 4 floats per pixel: world normal encoded n * 0.5 + 0.5 (lit_deferred_frag.glsl), a unused */
    const float *d_buffer_b;            /* 4 floats per pixel: diffuse rgb, a unused */
    const float *d_buffer_c;            /* 4 floats per pixel: specular rgb, shininess */
    const float *d_depth;               /* 1 float per pixel: depth buffer value 0..1 */
} OvisGBuffer;                          /* the reference stores these as R10G10B10A2 / fp16 / Depth32 (deferred_render_pipeline.h:51-54) */
/* engine/shaders/deferred_light_frag.glsl:30-81 + engine/shaders/lighting.h:46-211 for every light of a visible-light
 * list, summed per pixel in list order as the additive light passes of DeferredRenderPipeline::renderDeferredLights would
 * (render/src/deferred_render_pipeline.cc:424-470; target cleared to (0,0,0,1), alpha = source alpha = 1): one kernel, one
 * G-buffer read and one write per pixel whatever the number of lights. Light parameters come from the world's per-light
 * LightUniforms table (ovis_light_uniforms), the light type selects the shader variant. d_view_uniforms: the 288-byte
 * ViewUniforms block of the view the G-buffer was rendered from, in device memory (ovis_pack_view_uniforms_device).
 * d_visible_lights / count / max as for ovis_pack_light_uniforms_device (NULL = the last ovis_cull's light list).
 * d_shadow_maps: NULL (no light is shadowed: the shader variants without SHADOW), or a device array indexed by light id
 * of device pointers to that light's shadow map - shadow_size^2 floats for a spot light, 6 x shadow_size^2 (faces +x -x
 * +y -y +z -z) for a point light, NULL for an unshadowed light; sampled like the pipeline's sampler (bilinear, compare
 * Less, clamp: render/src/deferred_render_pipeline.cc:250-254) with the 3x3 PCF of lighting.h. d_out_rgba: 4 floats per
 * pixel. Asynchronous on the world's stream. */
int ovis_shade_deferred_device(OvisWorld *world, const OvisGBuffer *gbuffer, const void *d_view_uniforms,
                               const uint32_t *d_visible_lights, const uint32_t *d_visible_count, int64_t max_visible,
                               const float *const *d_shadow_maps, int32_t shadow_size, float *d_out_rgba);
enum { OVIS_POST_TONEMAP = 1u, OVIS_POST_GAMMA = 2u };
/* engine/shaders/tonemap_effect_frag.glsl:30-71 (Uncharted 2 operator; parameters exposure, whitePoint) and / or
 * engine/shaders/gamma_correction_effect_frag.glsl:26-31 (pow(rgb, 1 / gamma)) over n_pixels rgba float4 values; with
 * both bits set the two effects run fused in one pass, tonemap first (the order of the cubes demo's effect chain).
 * d_out_rgba may equal d_rgba. */
int ovis_post_effects_device(OvisWorld *world, int64_t n_pixels, const float *d_rgba, uint32_t effects, float exposure,
                             float white_point, float gamma, float *d_out_rgba);

/* ------------------------------------------------------------------------------------------------
 * Headless rasteriser: what fills the G-buffer and the shadow maps the light pass above reads, as CUDA kernels -
 * DeferredRenderPipeline::renderDeferredGBuffer and ::renderShadowMaps (render/src/deferred_render_pipeline.cc:384-412,
 * 335-370) drawing the entities of a visible list with engine/shaders/lit_vtx.glsl + lit_deferred_frag.glsl, resp.
 * shadow_vtx.glsl + shadow_frag.glsl, under the reference's default pipeline state (gpu/include/state.h:105-145: depth
 * test LessOrEqual with depth write, back faces culled, counter-clockwise = front; clear depth 1, colours 0). Pixel
 * centres at +0.5, row 0 = the bottom row (GL window coordinates, as ovis_shade_deferred_device reads them), top-left
 * rule, perspective-correct interpolation, near / far clip per pixel. Materials: flat parameters or base-level bilinear
 * textures (TEXTURED, NORMALMAP). An OvisRaster belongs to one world and is driven from the same host thread; its work is
 * ordered on the world's stream.
 * ---------------------------------------------------------------------------------------------- */
typedef struct OvisRaster OvisRaster;
enum { OVIS_MATERIAL_SPECULAR = 1u,     /* the shader keywords of lit_deferred_frag.glsl: SPECULAR (:68-73), */
       OVIS_MATERIAL_TEXTURED = 2u,     /* TEXTURED (:46-50: diffuse from diffuse_texture instead of diffuseColour), */
       OVIS_MATERIAL_NORMALMAP = 4u };  /* NORMALMAP (:52-63: normal from normal_texture in tangent space; the mesh needs tangents) */
typedef struct OvisMaterial {
    float    diffuse[3];                /* diffuseColour */
    float    specular[3];               /* specularColour */
    float    shininess;
    uint32_t flags;
    uint32_t diffuse_texture;           /* ids from ovis_raster_add_texture, read when the flag is set */
    uint32_t normal_texture;
} OvisMaterial;                         /* 40 bytes */
typedef struct OvisGBufferTarget {      /* the writable form of OvisGBuffer: row-major, 16-byte aligned device pointers */
    int32_t width, height;
    float *d_buffer_a, *d_buffer_b, *d_buffer_c;   /* 4 floats per pixel each */
    float *d_depth;                     /* 1 float per pixel */
} OvisGBufferTarget;
int ovis_raster_create(OvisWorld *world, OvisRaster **out_raster);
int ovis_raster_destroy(OvisRaster *raster);
/* A triangle mesh (Mesh / SubMesh vertex data as the OBJ loader builds it, engine/src/loaders/obj_loader.cc): positions and
 * normals 3 floats per vertex, texcoords 2 per vertex or NULL, indices 3 per triangle into this mesh's vertices. */
int ovis_raster_add_mesh(OvisRaster *raster, int64_t n_vertices, const float *positions, const float *normals,
                         const float *texcoords, int64_t n_indices, const uint32_t *indices, uint32_t *out_mesh);
/* Tangents of a mesh added before (4 floats per vertex: xyz + handedness, e.g. from ovis_mesh_tangents): the kTangentSemantic
 * attribute lit_vtx.glsl reads under NORMALMAP. Meshes without them have zero tangents. */
int ovis_raster_set_mesh_tangents(OvisRaster *raster, uint32_t mesh, int64_t n_vertices, const float *tangents);
/* A 2-D texture from R, G, B, A bytes (rows bottom-up as GL stores them, e.g. ovis_tga_load's output). srgb != 0: the colour
 * channels are sRGB-encoded and decoded to linear per texel before filtering (TGALoader's default, sRGB = true). wrap != 0:
 * repeat addressing (kWrap), else clamp to edge. Sampling is bilinear on the base level with texel centres at
 * (i + 0.5) / size - the reference asks the graphics API for mip-mapped 8x anisotropic filtering (engine/src/texture.cc:41-42),
 * whose result the API leaves to the implementation; this is the part of it every implementation agrees on up close. */
int ovis_raster_add_texture(OvisRaster *raster, int32_t width, int32_t height, const uint8_t *rgba, int32_t srgb, int32_t wrap,
                            uint32_t *out_texture);
int ovis_raster_add_material(OvisRaster *raster, const OvisMaterial *material, uint32_t *out_material);
/* RenderEntity(subMesh, material) (graphics/src/mesh_renderer.cc:180-187): what entity ids[i] is drawn with. */
int ovis_raster_bind_entities(OvisRaster *raster, int64_t n, const uint32_t *entity_ids, const uint32_t *mesh_ids,
                              const uint32_t *material_ids);
/* The G-buffer pass for the entities of a visible list (device ids + device count, e.g. from ovis_cull_planes_device),
 * drawn in list order from `view` (its viewport must equal the target's size). Entities without a bound mesh are skipped.
 * Synchronises the world's stream (it sizes its scratch from the list). */
int ovis_raster_gbuffer_device(OvisRaster *raster, const OvisView *view, const uint32_t *d_visible, const uint32_t *d_visible_count,
                               int64_t max_visible, const OvisGBufferTarget *target);
/* The same pass fed by a draw list: d_draws holds (entity id, pass index) pairs as ovis_build_draw_lists_device writes them
 * (the deferred list), one draw per pair, in list order. */
int ovis_raster_gbuffer_draws_device(OvisRaster *raster, const OvisView *view, const uint32_t *d_draws, const uint32_t *d_draw_count,
                                     int64_t max_draws, const OvisGBufferTarget *target);
/* DeferredRenderPipeline::renderBasic (render/src/deferred_render_pipeline.cc:506-522): the entities of the basic draw list
 * drawn with engine/shaders/unlit_vtx.glsl + unlit_frag.glsl (without TEXTURED: fragColour = vec4(diffuseColour, 1)) into
 * the colour target the light pass wrote, depth-tested (LessOrEqual, with write) against the depth buffer the G-buffer pass
 * left: d_rgba (4 floats per pixel) and d_depth are read and updated in place. Either an entity list or a draw list. */
int ovis_raster_basic_device(OvisRaster *raster, const OvisView *view, const uint32_t *d_visible, const uint32_t *d_visible_count,
                             int64_t max_visible, int32_t width, int32_t height, float *d_rgba, float *d_depth);
int ovis_raster_basic_draws_device(OvisRaster *raster, const OvisView *view, const uint32_t *d_draws, const uint32_t *d_draw_count,
                                   int64_t max_draws, int32_t width, int32_t height, float *d_rgba, float *d_depth);
/* The shadow-map passes: view k draws the entities d_entities[host_offsets[k] .. host_offsets[k + 1]) (the CSR result of
 * ovis_cull_shadow_views_device, shadow casters only) depth-only into d_maps[k]: shadow_size x shadow_size floats, cleared
 * to 1. d_maps is a HOST array of n_views device pointers (a point light's six faces are six consecutive views: point
 * d_maps[k + f] at map + f * shadow_size^2 to build the layout ovis_shade_deferred_device reads). */
int ovis_raster_shadow_maps_device(OvisRaster *raster, int32_t n_views, const OvisView *views, const int64_t *host_offsets,
                                   const uint32_t *d_entities, int32_t shadow_size, float *const *d_maps);

/* ------------------------------------------------------------------------------------------------
 * Scene import: the reference's on-disk formats, read into a world.
 * Host-side parsing; everything it creates goes through the calls above.
 * ---------------------------------------------------------------------------------------------- */
typedef struct OvisSceneInfo {
    uint32_t first_node, first_entity, first_light;   /* ids of what the import created (consecutive from these) */
    int32_t  n_nodes;                                 /* one per Entity that is active in the world */
    int64_t  n_entities;                              /* one per sub-mesh of every active MeshRenderer, one per Skybox */
    int32_t  n_lights;
    int32_t  n_cameras;
    int32_t  shadow_map_resolution;                   /* of the first camera's DeferredRenderPipeline, 0 if none */
} OvisSceneInfo;
/* Load a serialised world (`.object`: the JSON array engine/src/json_serialiser.cc writes; World -> root Entity,
 * engine/src/world.cc:56-71; Entity properties, components, children, engine/src/entity.cc:106-148) and create, in the
 * reference's activation order (entity.cc:414-421: an entity's components, then its children, depth first; inactive
 * entities and components are skipped as they never reach the render world):
 *   - one transform node per Entity (local position / orientation / scale; the root keeps the identity, entity.cc:127-130);
 *   - one render entity per sub-mesh of each MeshRenderer (graphics/src/mesh_renderer.cc:50-59,129-134; local AABB from
 *     the .obj, see ovis_obj_submesh_aabbs) and one per Skybox (+-FLT_MAX, graphics/src/skybox.cc:46-54), bound to the
 *     Entity's node, flags from castsShadow (graphics/src/renderer.cc:70-85);
 *   - one light per Ambient/Directional/Point/SpotLight with the component defaults of graphics/src/light.cc:35-74, world
 *     position and direction from the Entity's world transform (light.cc:100-108); the light is bound to the Entity's
 *     node (ovis_light_bind_node) and follows it from then on; attenuation and shadow bias via ovis_light_set_shading;
 *   - one OvisView per Camera (graphics/src/camera.cc:46-57,71-79) with the given viewport size, returned in
 *     out_cameras (up to cap_cameras; OVIS_ERR_CAPACITY if there are more, the world is still complete).
 * Mesh assets "game/<path>" resolve to <game_asset_dir>/<path>.obj. Physics, controllers and post effects are ignored.
 * The file is read as the reference's JSONSerialiser reads it (json_serialiser.cc:232-244, 486-626; entity.cc:127-147):
 * {"objectID": n} is the n-th object of the document; a property of the wrong JSON type keeps its default (a float has to be
 * written with a fraction or an exponent: 50.0, not 50); a component / child list ends at the first unreadable reference;
 * files the reference would assert on (an entity attached twice, a null reference in a list, no World first) are refused
 * with OVIS_ERR_INVALID. */
int ovis_scene_import(OvisWorld *world, const char *world_path, const char *game_asset_dir, int32_t vp_width,
                      int32_t vp_height, int32_t cap_cameras, OvisView *out_cameras, OvisSceneInfo *out_info);
/* The same in two steps, for hosts that want the parsed arrays (or have no device yet): ovis_scene_load parses into a
 * host-side OvisScene (needs no GPU), ovis_scene_info / ovis_scene_arrays read it (node parent indices and entity->node
 * indices are relative to the scene; light_shading = attenuation[3] + shadow bias per light; any output may be NULL; sizes
 * from ovis_scene_info), ovis_scene_instantiate creates its content in a world, ovis_scene_free releases it. */
typedef struct OvisScene OvisScene;
int ovis_scene_load(const char *world_path, const char *game_asset_dir, int32_t vp_width, int32_t vp_height, OvisScene **out_scene);
int ovis_scene_info(const OvisScene *scene, OvisSceneInfo *out_info);
int ovis_scene_arrays(const OvisScene *scene, int32_t *out_parent, float *out_local_trs, float *out_local_aabb, uint32_t *out_flags,
                      uint32_t *out_entity_node, OvisLight *out_lights, float *out_light_shading, OvisView *out_cameras);
int ovis_scene_instantiate(const OvisScene *scene, OvisWorld *world, OvisSceneInfo *out_info);
/* What render entity `entity` (0-based, in the scene's order) draws: the .obj file of its MeshRenderer's mesh asset and the
 * sub-mesh index (mesh_renderer.cc:129-134); an empty path and -1 for a Skybox. Feeds ovis_obj_mesh / ovis_raster_add_mesh. */
int ovis_scene_entity_mesh(const OvisScene *scene, int64_t entity, char *out_obj_path, int32_t capacity, int32_t *out_submesh);
int ovis_scene_free(OvisScene *scene);
/* Local bounding box of every sub-mesh of a Wavefront .obj file, as OBJLoader + MeshLoader::calculateBoundingBox produce
 * them (engine/src/loaders/obj_loader.cc:94-279: a sub-mesh starts at the first face after a `usemtl` that changes the
 * material or after any `g`; values parsed with strtof; engine/src/loaders/mesh_loader.cc:187-199: min/max over the
 * positions the sub-mesh's faces reference). 6 floats per sub-mesh. */
int ovis_obj_submesh_aabbs(const char *obj_path, int32_t capacity, float *out_aabb, int32_t *out_count);
/* The vertex data of the same file as OBJLoader builds it (obj_loader.cc:245-276): one vertex per distinct v/vt/vn triple
 * in order of first use, shared by all sub-meshes; a quad becomes the triangles (0, 1, 2), (2, 3, 0). positions / normals
 * 3 floats per vertex, texcoords 2; sub-mesh k owns indices [out_first_index[k], out_first_index[k + 1]) (cap_submeshes + 1
 * entries). The sizes are always returned; OVIS_ERR_CAPACITY if a buffer is too small. Feeds ovis_raster_add_mesh. */
int ovis_obj_mesh(const char *obj_path, int64_t cap_vertices, float *out_positions, float *out_normals, float *out_texcoords,
                  int64_t cap_indices, uint32_t *out_indices, int32_t cap_submeshes, int64_t *out_first_index,
                  int64_t *out_n_vertices, int64_t *out_n_indices, int32_t *out_n_submeshes);
/* MeshLoader::calculateTangents (engine/src/loaders/mesh_loader.cc:201-289; what an OBJLoader with generateTangents
 * adds): 4 floats per vertex, xyz = normalize(t - n dot(n, t)) with t the sum of the triangles' s-directions, w = the
 * handedness of the bitangent. Vertex data as ovis_obj_mesh returns it (the indices of all sub-meshes). */
int ovis_mesh_tangents(int64_t n_vertices, const float *positions, const float *normals, const float *texcoords,
                       int64_t n_indices, const uint32_t *indices, float *out_tangents);
/* TGALoader::loadData (engine/src/loaders/tga_loader.cc:60-117): uncompressed 24 / 32-bit true-colour TGA -> R, G, B, A
 * bytes, rows in file order, alpha 255 for 24-bit files. The size is always returned; OVIS_ERR_CAPACITY if the buffer
 * (capacity_bytes) is smaller than 4 x width x height. */
int ovis_tga_load(const char *path, int64_t capacity_bytes, uint8_t *out_rgba, int32_t *out_width, int32_t *out_height);
/* Diagnostic: the canonical text of a JSON document as the scene loader's own reader sees it (strings byte for byte,
 * members in file order, numbers as the float the reference's JSONSerialiser reads - GetFloat() = (float)double,
 * engine/src/json_serialiser.cc:540 - plus the integer when there is one). The test suite compares it with the same dump
 * made through the reference's own JSON dependency (the vendored rapidjson, engine/3rdparty/rapidjson), which pins the
 * JSON layer of ovis_scene_load. OVIS_ERR_CAPACITY (with *out_length = required) if `capacity` is too small. */
int ovis_json_canonical(const char *text, int64_t length, char *out, int64_t capacity, int64_t *out_length);
/* Read a light back: its OvisLight, attenuation[3] and shadow bias (any output may be NULL). */
int ovis_light_get(OvisWorld *world, uint32_t light_id, OvisLight *out_light, float out_attenuation[3], float *out_shadow_bias);

/* Copy `bytes` from a device pointer this library handed out (e.g. the CSR arrays of
 * ovis_cull_views_device) to host memory, after all work queued on the world's stream; for ABI users
 * that bind no CUDA runtime of their own (cgo, ctypes). No counterpart in the reference (its results are
 * host std::lists, render_world.h:49-55). */
int ovis_read_device(OvisWorld *world, const void *d_src, int64_t bytes, void *host_dst);

/* Pinned host memory for result buffers (plain malloc'd buffers work too, only slower to copy into). */
int ovis_host_alloc(int64_t bytes, void **out_ptr);
int ovis_host_free(void *ptr);
/* Page-locks memory the caller already owns (a shared-memory segment several processes map, an engine's own result
 * arena) so that ovis_cull / ovis_cull_planes store into it straight from the gather kernel, as they do for
 * ovis_host_alloc memory. Processes that each hold one shard of a world can share one result buffer this way: every
 * rank's kernel writes its own segment, no GPU-to-GPU exchange is needed for a host consumer. No counterpart in the
 * reference. */
int ovis_host_register(void *ptr, int64_t bytes);
/* Spin until every one of n words in HOST memory is at least `value` (frame counters other processes or GPUs store: the
 * hand-over of per-rank segments in one shared block). OVIS_ERR_CAPACITY after timeout_us microseconds (< 0: never). */
int ovis_host_flags_wait(const uint32_t *flags, int32_t n, uint32_t value, int64_t timeout_us);
/* Restrict the calling thread (and the threads it starts afterwards) to the CPUs next to `device`, as the kernel's PCI
 * topology lists them (/sys/bus/pci/devices/.../local_cpulist), within the CPUs it may already use. A frame is a launch,
 * a few stores across PCIe and a poll of host memory: from the far socket each of those is slower. out_cpus (optional)
 * receives the number of CPUs pinned to, 0 when nothing was changed (no topology files, a cpuset that excludes them).
 * The worker threads of ovis_sharded_create do this for their devices (OVIS_PIN_WORKERS=0 turns it off); a process that
 * holds one shard calls it once, before it creates its world. No counterpart in the reference. */
int ovis_host_pin_thread_near_device(int32_t device, int32_t *out_cpus);
int ovis_host_unregister(void *ptr);

/* Profiling hook (no counterpart in the reference, whose cull path has no timers - SURVEY.md section 5):
 * while enabled, every entity cull records CUDA events on the world's stream before the test kernel,
 * between the test and the gather kernel, and after the gather kernel. ovis_world_read_timing synchronises
 * the stream, writes the per-cull device times (milliseconds) of the culls since the last read / enable
 * (at most `capacity` entries) and resets the record. */
int ovis_world_set_timing(OvisWorld *world, int enabled);
int ovis_world_read_timing(OvisWorld *world, int32_t capacity, float *test_ms, float *gather_ms, int32_t *out_count);

/* Number of kernels this library has launched on behalf of `world` since creation (bench.py's gpu_launches). */
int ovis_world_launch_count(OvisWorld *world, int64_t *out_launches);

#ifdef __cplusplus
}
#endif
#endif /* ORION_VIS_H */
