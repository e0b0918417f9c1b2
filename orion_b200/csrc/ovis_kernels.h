// ovis_kernels.h - launch interface between the host world (ovis_world.cu) and the kernels.
#pragma once

#include <cstdint>
#include <cuda_runtime.h>

namespace ovis {

constexpr uint32_t kNone = 0xffffffffu;
constexpr int kCullThreads = 256;                       // threads per CTA of the cull test kernel
constexpr int kMaxPeers = 8;                            // extra destinations of the gather kernel (one NVSwitch box)
constexpr int kGatherTiles = 64;                        // warp tiles per block of the gather kernel
constexpr int kCullTile = 1024;                         // storage padding: a multiple of every warp-tile size

struct CullVariant {
    int groups;                  // float4 groups (4 consecutive entities each) per lane
    bool pipelined;              // two register sets per warp: tile t + 1 is requested before tile t is tested
    bool packed;                 // sign-specialised plane code with FMUL2 products (ovis_device.cuh: outside_plane4)
    int async_stages;            // > 0: cp.async ring of that many stages per warp in shared memory (groups must be 1)
    int tile() const { return 128 * groups; }   // entities per warp tile
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
    const float *bounds;         // world AABBs, 6 x stride
    int64_t stride;
    int64_t n;                   // entity slots
    int num_tiles;               // warp tiles: ceil(n / tile of the launched variant)
    int num_blocks;              // blocks of kGatherTiles warp tiles
    const uint32_t *mask;        // 1 bit per slot (alive [& caster]); only read by the MASKED variant
    uint32_t *staging;           // scratch: num_tiles x tile ids, each tile's survivors at the start of its slot
    uint32_t *tile_counts;       // scratch: survivors per warp tile
    uint32_t *block_totals;      // scratch: survivors per block of kGatherTiles tiles (zero on entry)
    uint32_t *block_totals_next; // scratch: the set the next launch accumulates into; zeroed here
    uint32_t id_base;            // added to every emitted slot id (global id of a shard's first entity)
    // sharded exchange (optional): wait until wait_flags[i] >= wait_value for i < n_wait (<= 32) before finishing
    int n_wait;
    const uint32_t *wait_flags;
    uint32_t wait_value;
    uint32_t *error_flag;
};

struct GatherArgs {               // K3b
    const uint32_t *staging;
    const uint32_t *tile_counts;
    const uint32_t *block_totals;
    int num_tiles;
    int tile_entities;
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

// ---- K4 multi-view ------------------------------------------------------------------------------
constexpr int kMultiMaxViews = 1024;          // views per pass (one CTA of 1024 threads scans them)

struct ViewGroupDev {
    float lo[3], hi[3];          // the group's pre-filter cube, already inflated by its rounding margin
                                 // (centre -/+ (1.001 * range + 1e-3 * max|centre| + 0.01)); -inf/+inf = no pre-filter
    int first_view, n_views;     // consecutive views of the pass that share this cube
};

struct MultiArgs {
    const float *bounds;         // world AABBs, 6 x stride
    int64_t stride;
    int64_t n;
    int num_tiles;               // supertiles of 1024 entities
    const uint32_t *alive_mask;  // nullptr when every slot is live
    const uint32_t *caster_mask; // alive & castsShadow
    const ViewGroupDev *groups;
    int n_groups;
    int any_cube;                // 0: no group of this pass has a finite pre-filter cube (plain views): the K6 work is skipped
    const float4 *planes;        // 6 per view
    const uint32_t *view_require;// per view: non-zero = shadow casters only
    const uint32_t *view_codes;  // per view: 3 bits per plane (bit c = normal component c >= 0), plane p at bits 3p..3p+2
    int n_views;
    uint32_t *counts;            // [num_tiles][n_views]: counts after pass A, positions after the scans
    uint8_t *sub_masks;          // [num_tiles][n_views]: bit s = sub-tile s holds a survivor of the view (A writes, C reads)
    uint32_t *block_totals;      // [num_blocks][n_views], zero before pass A
    uint32_t *group_hits;        // [num_tiles][ceil(n_groups / 32)]: bit = group had a candidate (A writes, C reads)
    const long long *csr_offsets;// [n_views + 1] start of each view's list (valid after the scans)
    uint32_t *out;
    long long capacity;
    uint32_t id_base;
};

int multi_max_blocks_per_sm(int n_views, int n_groups);
void launch_multi_pass(const MultiArgs &a, bool write, int grid, cudaStream_t s);
void launch_multi_scans(const MultiArgs &a, long long *csr_offsets_of_pass, cudaStream_t s);

// ---- consumers of the visible lists (section 8(f)) -----------------------------------------------------------------
struct DrawArgs {
    const uint32_t *d_visible;       // ids (slot + id_base), list order
    const uint32_t *d_visible_count; // device count
    uint32_t id_base;
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
    const uint32_t *entity_node;     // per slot: node the entity follows, or kNone
    const float *world_trs;          // 10 x nstride
    int64_t nstride;
    float *d_out;                    // 20 floats per visible entity
    int64_t capacity;                // entries d_out can hold
};

int launch_draw_lists(const DrawArgs &a, int64_t max_visible, cudaStream_t s);   // returns the launches made
void launch_entity_uniforms(const UniformArgs &a, int64_t max_visible, cudaStream_t s);
void launch_light_uniforms(const uint32_t *d_visible, const uint32_t *d_visible_count, const float *d_table, float *d_out,
                           int64_t max_visible, int64_t capacity, cudaStream_t s);

// ---- deferred shading over headless buffers (section 8(f) row 4) ----------------------------------------------------
struct ShadeArgs {
    int width, height;
    const float *buffer_a, *buffer_b, *buffer_c;   // float4 per pixel
    const float *depth;                            // float per pixel
    const float *view_uniforms;                    // 72 floats (ViewUniforms block), device
    const uint32_t *d_visible;                     // light ids, list order
    const uint32_t *d_visible_count;
    int64_t max_visible;
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
};

void launch_scatter_aos(int k, const float *src, const uint32_t *ids, uint32_t first, int64_t n, float *dst,
                        int64_t stride, cudaStream_t s);
void launch_gather_aos(int k, const float *src, int64_t stride, uint32_t first, int64_t n, float *dst, cudaStream_t s);
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
int cull_variant_smem(int v);       // dynamic shared memory per CTA of the variant's test kernel
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
