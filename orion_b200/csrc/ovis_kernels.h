// ovis_kernels.h - launch interface between the host world (ovis_world.cu) and the kernels.
#pragma once

#include <cstdint>
#include <cuda_runtime.h>

namespace ovis {

constexpr uint32_t kNone = 0xffffffffu;
constexpr int kCullThreads = 256;                       // threads per CTA of the cull test kernel
constexpr int kMaxPeers = 8;                            // extra destinations of the gather kernel (one NVSwitch box)
constexpr int kMaskBlockEntities = 8192;                // entity slots per block of the gather kernel ...
constexpr int kMaskBlockWords = kMaskBlockEntities / 32;   // ... = 256 words of the survivor bitmap, one per thread
constexpr int kMaskBlockGroups = kMaskBlockEntities / 128; // ... = 64 groups of 128 entities (one warp, 4 per lane)
constexpr int kCullTile = 1024;                         // storage padding: a multiple of every warp-tile size

struct CullVariant {
    int groups;                  // register-fed kernel: float4 groups (4 consecutive entities each) per lane and tile
    int ctas;                    // register-fed kernel: CTAs per SM it is compiled for
    int ring_stages;             // > 0: the TMA ring kernel with that many stages of 1024 entities per CTA
    int ring_sets;               // ... and that many sets of eight consumer warps taking the stages in turn
};

// SoA storage convention: component k of item i lives at base[k * stride + i]; stride is a multiple of
// kCullTile so every component array is 4 KiB aligned and float4-loadable.

struct LevelArgs {
    const uint32_t *list;        // node ids of this level, or nullptr when the level is the range [first, first+count)
    int64_t first, count;
    const int32_t *parent;       // per node, -1 = child of the identity root
    const float *local_trs;      // 10 x nstride
    float *world_trs;            // 10 x nstride
    int64_t nstride;
    uint8_t *dirty;              // per node: local TRS (or a bound entity's box) changed; nullptr = recompute everything.
                                 // The kernel sets it for every node it recomputes, so children inherit it.
    const uint32_t *node_entity; // first entity bound to the node, or kNone
    const float *local_aabb;     // 6 x estride
    float *world_aabb;           // 6 x estride
    int64_t estride;
};

struct CullArgs {                 // K3a
    float4 planes[6];            // normal xyz, distance
    uint32_t sign_codes;         // plane_sign_codes(planes): which corner is the P-vertex, 3 bits per plane
    const float *bounds;         // world AABBs, 6 x stride
    int64_t stride;
    int64_t n;                   // entity slots
    int num_groups;              // groups of 128 entities: ceil(n / 128)
    int num_blocks;              // blocks of kMaskBlockEntities slots
    const uint32_t *mask;        // 1 bit per slot (alive [& caster]); only read by the MASKED kernels
    uint32_t *survivors;         // scratch: 1 bit per slot, 4 words per group; every group's words are written
    uint32_t *block_totals;      // scratch: survivors per block (zero on entry)
    uint32_t *block_totals_next; // scratch: the set the next launch accumulates into; zeroed here
    // sharded exchange (optional): wait until wait_flags[i] >= wait_value for i < n_wait (<= 32) before finishing
    int n_wait;
    const uint32_t *wait_flags;
    uint32_t wait_value;
    uint32_t *error_flag;
};

struct GatherArgs {               // K3b
    const uint32_t *survivors;
    const uint32_t *block_totals;
    int64_t num_words;           // bitmap words: 4 x num_groups
    int num_blocks;
    uint32_t id_base;            // added to every emitted slot id (global id of a shard's first entity)
    uint32_t *out;               // visible slot ids, ascending
    long long capacity;          // writable entries of out
    const long long *d_base;     // optional: start offset inside out (CSR chaining), device memory
    long long *d_next;           // optional: receives base + count
    uint32_t *d_count;           // optional: receives count
    // Fused exchange of a sharded world: the same list and count are also stored to these buffers, which may be
    // peer GPUs' memory mapped over NVLink (plain st.global to peer addresses; no CSR chaining on this path).
    // The extra destinations share one capacity of their own (extra_capacity). A destination may also be pinned host
    // memory mapped into the device's address space: the ids then reach the host without any copy after the kernel.
    int n_extra;
    long long extra_capacity;
    uint32_t *extra_out[kMaxPeers];
    uint32_t *extra_count[kMaxPeers];
    // One-way flags instead of a barrier (sharded worlds, all optional):
    //   before storing, wait until every wait_flags[i] >= wait_value  (peers are done reading the buffer being reused);
    //   after ALL blocks have stored, write signal_value to every signal_flags[x] (this shard's segment is complete).
    int n_wait;
    const uint32_t *wait_flags;
    uint32_t wait_value;
    int n_signal;
    uint32_t *signal_flags[kMaxPeers + 1];
    uint32_t signal_value;
    int block_fence_sys;          // OVIS_EXCHANGE_BLOCK_FENCE=sys: system-scope fence in every block (diagnostic)
    unsigned int *done_counter;   // scratch, zero between launches
    uint32_t *error_flag;         // set to 1 if a wait gave up (a peer never arrived)
};

// ---- K4 multi-view + K6 light binning (ovis_multiview.cu) -------------------------------------------------------
// Views are numbered 0 .. n_views - 1 in the caller's order; every view owns a row of per-block totals and a CSR list.
// "Plain" views (camera, cascades, spot lights: no pre-filter) go through a dense pass that leaves one survivor bit
// per (view, entity); the six faces of a point light form a "cube group" and go through the sparse pass, which finds
// the few (entity, light) candidate pairs through a uniform grid over the lights.
constexpr int kMvMaxGroupViews = 6;           // views of one cube group (a point light's faces)
constexpr int kMvMaxGroups = 4096;            // cube groups per call (12 bits of the record key)
constexpr int kMvGridMax = 64;                // cells per axis of the light grid (bitmap rows are 64-bit words)
constexpr int kMvPairCapacity = 8192;         // candidate pairs one block of 8192 entities may produce before falling back

struct MvGroup {                  // one cube group
    float lo[3], hi[3];          // the pre-filter cube, already inflated by its rounding margin
    int first_view, n_views;     // its views are consecutive
    uint32_t cell_lo, cell_hi;   // grid cells it covers: x | y << 8 | z << 16, inclusive (host-computed)
};

struct MvGrid {
    float lo[3], hi[3];          // exact bounds of the union of the groups' cubes
    float inv_cell[3];           // cells per unit length
    float bias[3];               // cell of v = floor(fma(v, inv_cell, bias)), the same expression on host and device; lo maps to
                                 // 1.5: cell 0 and cell dim - 1 are a border no cube reaches, where everything outside lands
    float re_bias;               // 1.002e-3 x the largest |coordinate| of the grid's bounds (entity inflation, see inflate_box)
    int dim[3];                  // cells per axis, border included (<= kMvGridMax)
};

struct MvItem {                   // one survivor of the sparse pass, placed by mv_expand_items_kernel
    uint32_t view, rank, id;     // rank inside (view, block); id already includes id_base
};

struct MvArgs {
    const float *bounds;         // world AABBs, 6 x stride
    int64_t stride;
    int64_t n;
    int num_groups128;           // groups of 128 entities
    int num_blocks;              // blocks of kMaskBlockEntities
    const uint32_t *alive_mask;  // nullptr when every slot is live
    const uint32_t *caster_mask; // alive & castsShadow; nullptr when no view asks for casters
    const float4 *planes;        // 6 per view
    const uint32_t *view_meta;   // per view: bits 0-17 P-vertex sign codes (3 per plane), bit 31 = shadow casters only
    int n_views;
    uint32_t *totals;            // [n_views][num_blocks + 1]: counts, then (after the scan) exclusive prefixes + the total
    long long *csr_offsets;      // [n_views + 1]
    uint32_t *out;
    long long capacity;
    uint32_t id_base;
    uint32_t *state;             // [0] overflow bits (kMvOverflow*), [1] scan ticket, [2] light-grid nodes used, [3] items used,
                                 // [4] most records any block wanted (when kMvOverflowHits)
    // dense pass
    const uint32_t *plain_views; // view ids handled by the dense pass
    const uint2 *plain_meta;     // per plain view: x = sign codes | equal codes from here on (1..4) << 18 | casters only << 31, y = view id
    const float4 *plain_planes;  // per plain view 12: plane p = (nx, nx, ny, ny), (nz, nz, d, d)
    int n_plain;
    int views_per_warp;          // plain views one warp handles for its 128 entities (small worlds split the views)
    uint32_t *masks;             // [n_plain][num_groups128 * 4]
    // sparse pass
    const MvGroup *groups;
    int n_groups;
    MvGrid grid;
    unsigned long long *bitmap;  // [dim z][dim y]: bit x = some cube covers the cell
    int32_t *heads;              // per cell: first node of its group list, -1 = none
    uint4 *nodes;                // per node two: (group | first cell x << 12 | y << 18 | z << 24, next node, cube lo x, lo y),
                                 // (lo z, hi x, hi y, hi z): all a cell's visitor needs of the group, in one sector
    uint32_t node_capacity;
    MvItem *items;
    uint32_t item_capacity;
    uint2 *segments;             // per block: (first item, items)
    int hit_capacity;            // records per block the lists hold (the host enlarges them when a block had more)
    float4 *hit_records;         // per block of 8192 entities hit_capacity records of two: (min xyz, max x), (max yz, tag, -);
                                 // tag = index inside the block | tested against every group << 31 (mv_classify_kernel -> mv_sparse_kernel)
    uint32_t *block_work;        // per block: records appended (may exceed the capacity: overflow), zeroed per call
    long long *host_offsets;     // page-locked host copy of csr_offsets (this round's slice) and of state[0..7], written by
    uint32_t *host_state;        // the scan kernel's last CTA; nullptr = none
};
struct MvResetArgs {              // up to four arrays filled with a word (sizes in 16-byte units, pointers 16-byte aligned)
    void *ptr[4];
    int64_t quads[4];
    uint32_t value[4];
    uint32_t *state;             // first_round: state[0] = 0 and csr_offsets[0] = 0 as well
    long long *csr_offsets;
    int first_round;
    const uint4 *copy_src;       // the round's staging (planes, view records, groups) in page-locked host memory -> copy_dst on
    uint4 *copy_dst;             // the device: fetched by this kernel, which saves the call a cudaMemcpyAsync on its critical path
    int64_t copy_quads;
};
void launch_mv_reset(const MvResetArgs &r, int num_sms, cudaStream_t s);
constexpr uint32_t kMvOverflowPairs = 1u, kMvOverflowItems = 2u, kMvOverflowNodes = 4u, kMvOverflowHits = 8u;
// plain_meta.x: P-vertex codes (18 bits) | equal codes that follow, capped at 4 (3 bits) | how many views starting here
// have bit-identical side planes (left, right, top, bottom), capped at 7 (3 bits; 0 or 1: none) | ... | the view wants
// shadow casters only (bit 31)
constexpr int kPlainSharedShift = 21;

void launch_mv_plain_test(const MvArgs &a, int num_sms, cudaStream_t s);
void launch_mv_bin_lights(const MvArgs &a, cudaStream_t s);
void launch_mv_sparse(const MvArgs &a, int num_sms, cudaStream_t s);
void launch_mv_scan(const MvArgs &a, cudaStream_t s);
void launch_mv_expand_masks(const MvArgs &a, cudaStream_t s);
void launch_mv_expand_items(const MvArgs &a, cudaStream_t s);

// ---- consumers of the visible lists (section 8(f)) -----------------------------------------------------------------
struct DrawArgs {
    const uint32_t *d_visible;       // ids (slot + id_base), list order
    const uint32_t *d_visible_count; // device count
    uint32_t id_base;
    uint32_t n_slots;                // ids outside [id_base, id_base + n_slots) expand to nothing (a stale id must not index the tables)
    const uint32_t *pass_counts;     // per slot: numPasses deferred | basic << 8 | shadow caster << 16
    const uint32_t *caster_mask;     // alive & castsShadow bits (mode 1)
    int mode;                        // 0 = prepareEntities (two lists), 1 = shadow-caster list
    uint32_t *tile_totals;           // scratch: 2 per 1024-entry tile (NULL: the write kernel counts for itself)
    uint2 *d_draws0, *d_draws1;      // (entity id, pass index)
    long long cap0, cap1;
    uint32_t *d_draw_counts;         // [2]
};

struct UniformArgs {
    const uint32_t *d_visible;
    const uint32_t *d_visible_count;
    uint32_t id_base;
    uint32_t n_slots;                // an id outside [id_base, id_base + n_slots) gets a zero block
    const uint32_t *entity_node;     // per slot: node the entity follows, or kNone
    const float *world_trs;          // 10 x nstride
    int64_t nstride;
    float *d_out;                    // 20 floats per visible entity
    int64_t capacity;                // entries d_out can hold
};

int launch_draw_lists(const DrawArgs &a, int64_t max_visible, cudaStream_t s);   // returns the launches made
void launch_entity_uniforms(const UniformArgs &a, int64_t max_visible, cudaStream_t s);
void launch_light_uniforms(const uint32_t *d_visible, const uint32_t *d_visible_count, const float *d_table, uint32_t n_lights,
                           float *d_out, int64_t max_visible, int64_t capacity, cudaStream_t s);

// ---- deferred shading over headless buffers (section 8(f) row 4) ----------------------------------------------------
struct ShadeArgs {
    int width, height;
    const float *buffer_a, *buffer_b, *buffer_c;   // float4 per pixel
    const float *depth;                            // float per pixel
    const float *view_uniforms;                    // 72 floats (ViewUniforms block), device
    const uint32_t *d_visible;                     // light ids, list order
    const uint32_t *d_visible_count;
    int64_t max_visible;
    uint32_t n_lights;                             // light slots: an id beyond them is skipped
    const float *light_table;                      // 52 floats per light slot
    const int32_t *light_types;                    // OVIS_LIGHT_* per light slot, | kLightSkippable
    const float *const *shadow_maps;               // per light slot: depth image or NULL; NULL = no shadows at all
    int shadow_size;
    float *out_rgba;                               // float4 per pixel
};
// Set by the host for a point / spot light whose block is finite and bounded, with attenuation constant > 0, linear and
// exponential terms >= 0 and (spot) cosCutoff < 1: beyond its range such a light contributes exactly +-0 to a clean pixel.
constexpr int32_t kLightSkippable = 1 << 8;
void launch_shade_deferred(const ShadeArgs &a, cudaStream_t s);
void launch_post_effects(int64_t n_pixels, const float *in, uint32_t effects, float exposure, float white_point, float gamma, float *out,
                         cudaStream_t s);

enum : uint32_t { kLightNever = 0, kLightAlways = 1, kLightSphere = 2, kLightBox = 3 };

struct DeviceLight {
    float4 a;                    // sphere: centre xyz + radius; box: minimum xyz
    float4 b;                    // box: maximum xyz
    uint32_t kind;
    uint32_t pad[3];
};

struct LightCullArgs {
    float4 planes[6];
    const DeviceLight *lights;
    int n;
    uint32_t *out;
    uint32_t *d_count;
    uint32_t *bits;              // scratch: one ballot word per 32 light slots (8 per CTA)
    uint32_t *counter;           // scratch: CTAs finished, zero between launches
};

void launch_scatter_aos(int k, const float *src, const uint32_t *ids, uint32_t first, int64_t n, float *dst,
                        int64_t stride, cudaStream_t s);
void launch_gather_aos(int k, const float *src, int64_t stride, uint32_t first, int64_t n, float *dst, cudaStream_t s);
void launch_compact_soa6(const float *src, const uint32_t *ids, int64_t n, int64_t stride, float *dst, cudaStream_t s);
void launch_compact_u32(const uint32_t *src, const uint32_t *ids, int64_t n, uint32_t *dst, cudaStream_t s);
void launch_gather_trs(const float *src, int64_t stride, const uint32_t *ids, int64_t n, float *dst, cudaStream_t s);
void launch_scatter_u32(const uint32_t *src, const uint32_t *ids, int64_t n, uint32_t *dst, cudaStream_t s);
void launch_fill_u32(uint32_t *dst, int64_t n, uint32_t value, cudaStream_t s);
void launch_identity_bounds(const uint32_t *ids, uint32_t first, int64_t n, const float *local, float *world,
                            int64_t stride, cudaStream_t s);
void launch_update_level(const LevelArgs &a, cudaStream_t s);
void launch_extra_bounds(const uint32_t *entities, const uint32_t *nodes, int64_t n, const float *world_trs,
                         int64_t nstride, const float *local, float *world, int64_t estride, const uint8_t *dirty,
                         cudaStream_t s);
void launch_mark_dirty(const uint32_t *ids, int64_t n, uint8_t *dirty, cudaStream_t s);
int cull_num_variants();
CullVariant cull_variant(int v);
// The variant used when none is forced: small sets are latency-bound (register-fed, many small warp tiles); large sets
// stream through the TMA ring.
inline int cull_default_variant(int64_t n_entities) { return n_entities < (4 << 20) ? 1 : 2; }
int cull_test_grid(int variant, int num_groups, int resident_ctas);   // CTAs to launch
uint32_t plane_sign_codes(const float planes[24]);
void launch_cull_test(const CullArgs &a, bool masked, int variant, int grid, cudaStream_t s, bool overlap_previous_gather);
void launch_cull_gather(const GatherArgs &a, cudaStream_t s);
void launch_segments_to_host(const uint32_t *flags, uint32_t value, const uint32_t *segments, long long segment_words,
                             int n_segments, uint32_t *host, uint32_t *error_flag, cudaStream_t s);
void launch_wait_flags(const uint32_t *flags, int n, uint32_t value, uint32_t *error_flag, cudaStream_t s);
int cull_fused_max_blocks_per_sm(bool masked, int variant);
cudaError_t launch_cull_fused(const CullArgs &a, const GatherArgs &g, bool masked, int variant, int grid, cudaStream_t s);
int cull_test_max_blocks_per_sm(bool masked, int variant);
void launch_empty_result(const long long *d_base, long long *d_next, uint32_t *d_count, cudaStream_t s);
void launch_cull_lights(const LightCullArgs &a, cudaStream_t s);

}  // namespace ovis
