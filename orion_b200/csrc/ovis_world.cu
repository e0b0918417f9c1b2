// ovis_world.cu - the host side of liborion_vis: device-resident world store and the C ABI of
// include/orion_vis.h. It mirrors the reference's RenderWorld contract (render/include/render_world.h:40-98):
// the world keeps borrowed entity/light state in insertion order and answers cull() calls; here the
// state lives in HBM as SoA arrays and every cull is one kernel launch.
//
// There is deliberately no CPU implementation of any stage in this file: if CUDA is unavailable every
// entry point that needs the device returns OVIS_ERR_CUDA.
#include "orion_vis.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cctype>
#include <sched.h>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <new>
#include <string>
#include <vector>

#include <cuda_runtime.h>

#include "ovis_kernels.h"
#include "ovis_math.hpp"

namespace {

thread_local std::string g_error;

int fail(int code, const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_error = buf;
    return code;
}

#define OVIS_CUDA(call)                                                                                   \
    do {                                                                                                  \
        cudaError_t err__ = (call);                                                                       \
        if (err__ != cudaSuccess)                                                                         \
            return fail(OVIS_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(err__), __FILE__, __LINE__); \
    } while (0)

#define OVIS_TRY(call)                 \
    do {                               \
        int rc__ = (call);             \
        if (rc__ != OVIS_OK)           \
            return rc__;               \
    } while (0)

inline int64_t round_up(int64_t v, int64_t m) { return (v + m - 1) / m * m; }

// OVIS_MV_TRACE=1: host-side time stamps of a many-view call on stderr (where the microseconds between "kernels" and the
// sum of the kernel times go)
struct MvTrace {
    bool on;
    std::chrono::steady_clock::time_point t0, last;
    MvTrace() {
        static const bool enabled = getenv("OVIS_MV_TRACE") && atoi(getenv("OVIS_MV_TRACE")) != 0;
        on = enabled;
        if (on)
            t0 = last = std::chrono::steady_clock::now();
    }
    void mark(const char *what) {
        if (!on)
            return;
        const auto now = std::chrono::steady_clock::now();
        fprintf(stderr, "[mv] %-22s +%7.1f us  (%7.1f)\n", what, std::chrono::duration<double, std::micro>(now - last).count(),
                std::chrono::duration<double, std::micro>(now - t0).count());
        last = now;
    }
};

struct Level {
    int64_t count = 0;
    int64_t first = 0;          // smallest node id of the level
    bool contiguous = true;     // ids are exactly [first, first + count)
    int64_t list_offset = 0;    // into d_level_lists when not contiguous
};

}  // namespace

struct OvisWorld {
    int device = 0;
    cudaStream_t stream = nullptr;
    int num_sms = 0;
    int forced_variant = -1;              // OVIS_CULL_VARIANT (tuning); -1 = pick by size
    bool fused_small = true;              // OVIS_CULL_FUSED=0 (tuning) keeps small worlds on the two-launch path
    bool mv_shared_sides = true;          // OVIS_MV_SHARED_SIDES=0 (tuning): views with identical side planes are tested like any others
    int64_t fused_below = 512 << 10;      // OVIS_FUSED_BELOW (tuning): worlds with fewer entities take the single cooperative launch
    int blocks_per_sm[16][2] = {};         // cull test kernel occupancy per variant: [unmasked, masked]
    int fused_blocks_per_sm[16][2] = {};   // same for the fused (cooperative) kernel; 0 = not available
    int64_t launches = 0;
    int64_t launches_after_gather = -1;   // value of `launches` right after the last gather kernel went out
    uint32_t id_base = 0;                 // global id of slot 0 (sharded worlds)

    // ---- entities (slot == id; slots are never reused, so slot order is insertion order)
    int64_t n_ent = 0, cap_ent = 0, n_live = 0;
    float *d_local = nullptr;             // 6 x cap_ent
    float *d_world = nullptr;             // 6 x cap_ent
    uint32_t *d_alive_bits = nullptr;     // cap_ent / 32 words: slot holds a live entity
    uint32_t *d_caster_bits = nullptr;    // alive & castsShadow
    std::vector<uint32_t> h_alive, h_caster_raw;   // host mirrors of the bit masks (caster before masking by alive)
    std::vector<uint32_t> h_entity_node;  // node the entity follows, or kNone
    uint32_t *d_entity_node = nullptr;    // device copy (entity uniforms)
    uint32_t *d_pass_counts = nullptr;    // per slot: shader passes deferred | basic << 8 | shadow caster << 16
    uint32_t *d_draw_totals = nullptr;    // draw-list scratch: 2 per 1024 visible entries
    uint32_t *d_draw_counts = nullptr;    // [2] draws written by the last ovis_build_draw_lists_device
    bool masks_dirty = false;
    int64_t masks_dirty_lo = 0, masks_dirty_hi = 0;     // 32-entity mask words changed since the last upload: [lo, hi)
    bool any_dead = false;

    // ---- transform tree
    int64_t n_nodes = 0, cap_nodes = 0;
    int32_t *d_parent = nullptr;
    float *d_local_trs = nullptr;         // 10 x cap_nodes
    float *d_world_trs = nullptr;         // 10 x cap_nodes
    uint32_t *d_node_entity = nullptr;
    std::vector<int32_t> h_parent, h_level;
    std::vector<uint8_t> h_node_dead;     // empty until the first ovis_node_remove; a dead node is in no level
    std::vector<int32_t> light_node;      // per light: the node it follows (ovis_light_bind_node), or OVIS_NO_PARENT
    int64_t n_bound_lights = 0;
    bool light_bindings_dirty = false;
    std::vector<uint32_t> h_subtree;      // nodes in the subtree rooted at each node (itself included)
    std::vector<uint32_t> h_node_entity;
    std::vector<uint32_t> h_extra_entity, h_extra_node;   // 2nd.. entity on a node
    uint32_t *d_extra_entity = nullptr, *d_extra_node = nullptr;
    int64_t cap_extra = 0;
    bool extra_dirty = false;
    std::vector<Level> levels;
    uint32_t *d_level_lists = nullptr;
    int64_t cap_level_lists = 0;
    bool levels_dirty = false;
    bool transforms_dirty = false;
    uint8_t *d_dirty = nullptr;           // per node: changed since the last update (cap_nodes bytes)
    bool all_dirty = false;               // recompute every node (topology changed, or most nodes moved)
    bool marks_present = false;           // d_dirty holds non-zero bytes

    // ---- lights
    std::vector<OvisLight> lights;
    std::vector<uint8_t> light_alive;
    // Shadow views and their planes are host math with libm calls (render_light.cc:258-332, render_view.cc:102-135,
    // frustum.cc:27-67): ~150 ns per view. They only change when the light does, so they are kept per light
    // (4096 lights x 6 faces every frame would otherwise be 4 ms of host time in front of a 1 ms GPU pass).
    struct LightViews { int resolution = 0, count = 0; float planes[6 * 24]; };
    std::vector<LightViews> light_views;
    struct LightShading { float attenuation[3]; float shadow_bias; };
    std::vector<LightShading> light_shading;   // what the LightUniforms block needs beyond OvisLight
    float *d_light_uniforms = nullptr;    // 52 floats (208 B) per light slot
    int32_t *d_light_types = nullptr;     // OVIS_LIGHT_* per light slot
    int64_t last_visible_lights = 0;      // light count of the previous ovis_cull with OVIS_CULL_LIGHTS
    ovis::DeviceLight *d_lights = nullptr;
    uint32_t *d_light_out = nullptr;
    uint32_t *d_light_bits = nullptr;     // K5 scratch: visibility bits per 32 lights, and the finished-CTA counter behind them
    int64_t cap_lights = 0;
    bool lights_dirty = false;
    int64_t lights_dirty_lo = 0, lights_dirty_hi = 0;   // light slots changed since the last upload: [lo, hi)

    // ---- cull scratch and results
    // Two survivor bitmaps and three block-total sets rotate from cull to cull, so that a cull's test kernel can
    // run while the previous cull's gather kernel is still reading its own set.
    uint32_t *d_survivors = nullptr;      // 2 x cap_ent / 32 words: one bit per slot, written by the test kernel
    uint32_t *d_block_totals = nullptr;   // 3 x cap_blocks totals (the idle sets are zero)
    int64_t cap_blocks = 0;
    uint64_t cull_serial = 0;             // culls launched so far (selects the scratch sets)
    int64_t last_visible = 0;             // entity count of the previous host-output cull (sizes the speculative copy)
    uint32_t *d_result = nullptr;         // cap_ent visible ids
    uint32_t *d_counts = nullptr;         // [0] entities, [1] lights
    unsigned int *d_exchange = nullptr;   // [0] gather done counter, [1] exchange error flag
    uint32_t *h_counts = nullptr;         // pinned mirror
    uint32_t *h_counts_dev = nullptr;     // the same words as the device addresses them (mapped pinned memory)
    uint32_t host_frame = 0;              // value the gather kernel's last block stores into h_counts[2] for the current cull
    bool host_flag = true;                // OVIS_HOST_FLAG=0: wait with cudaStreamSynchronize instead of watching that word
    bool direct_host = true;              // OVIS_DIRECT_HOST=0 (tuning) forces the copy path
    bool block_fence_sys = false;         // OVIS_EXCHANGE_BLOCK_FENCE=sys (diagnostic): system fence in every gather block
    uint32_t *d_csr_entities = nullptr;
    int64_t cap_csr = 0;
    long long *d_csr_offsets = nullptr;
    int64_t cap_csr_views = 0;

    // ---- multi-view (K4 / K6) scratch, see MvArgs
    void *h_mv_stage = nullptr;           // pinned: planes | view meta | plain view ids | cube groups of one round
    size_t cap_h_mv_stage = 0;
    unsigned char *d_mv_stage = nullptr;  // the device copy
    size_t cap_d_mv_stage = 0;
    void *h_mv_result = nullptr;          // pinned: CSR offsets + state words read back at the end of a call
    size_t cap_h_mv_result = 0;
    unsigned char *d_mv_state = nullptr;  // [overflow, scan ticket, nodes used, items used | pad | occupancy bitmap]
    int32_t *d_mv_heads = nullptr;        // per grid cell: first node of its light list
    uint32_t *d_mv_totals = nullptr;      // [views of a round][blocks + 1]
    int64_t cap_mv_totals = 0;
    uint32_t *d_mv_masks = nullptr;       // [plain views of a round][words]
    int64_t cap_mv_masks = 0;
    float4 *d_mv_hits = nullptr;          // sparse pass: per block of 8192 entities the records of those near a light
    int64_t cap_mv_hits = 0;
    int mv_hit_cap = 1024;                // records per block (grown on demand, at most one per entity of the block)
    uint32_t *d_mv_work = nullptr;        // ... and how many
    int64_t cap_mv_work = 0;
    uint4 *d_mv_nodes = nullptr;          // light grid: two per node (ovis_kernels.h, MvArgs::nodes)
    int64_t cap_mv_nodes = 0;
    ovis::MvItem *d_mv_items = nullptr;
    int64_t cap_mv_items = 0;
    uint2 *d_mv_segments = nullptr;
    int64_t cap_mv_segments = 0;
    bool mv_force_dense = false;          // OVIS_MV_DENSE=1 (testing): cube groups take the dense pass too
    bool mv_force_sparse = false;         // OVIS_MV_DENSE=0 (testing): ... and never by the size rule
    int mv_round_views = 0;               // OVIS_MV_ROUND_VIEWS (testing): upper bound on the views of one round, 0 = by memory
    // Host outputs of a multi-view call ride in the same round trip as the offsets: the lists change little from frame to
    // frame, so a little more than the previous total is copied speculatively; only a frame that grew pays a second copy.
    uint32_t *mv_spec_out = nullptr;
    int64_t mv_spec_cap = 0, mv_spec_copied = 0, mv_last_total = 0;
    // The light list of the last host-output cull with lights: ovis_build_shadow_lists of the same view reuses it instead
    // of culling the lights again and waiting for them (the pipeline culls the main view first: deferred_render_pipeline.cc:147).
    std::vector<uint32_t> h_visible_lights;
    float h_light_cull_planes[24] = {};
    bool light_cull_valid = false;

    // ---- optional per-kernel timing of the entity cull (bench / profiling hook)
    bool timing = false;
    std::vector<cudaEvent_t> timing_events;   // triples: before test, after test, after gather
    size_t timing_used = 0;

    // ---- upload staging
    void *d_stage[2] = { nullptr, nullptr };
    size_t stage_bytes[2] = { 0, 0 };
};

namespace {

using ovis::kCullTile;
using ovis::kNone;

constexpr int64_t kUploadChunk = 1 << 23;   // items per staged upload
constexpr int64_t kMvDenseBelowTests = 100000000;   // entity-view tests below which cube groups take the dense pass
constexpr size_t kMvBitmapOffset = 64;      // bytes: four state words, padding, then the light grid's occupancy bitmap
constexpr size_t kMvStateBytes = kMvBitmapOffset + sizeof(unsigned long long) * ovis::kMvGridMax * ovis::kMvGridMax;

int use_device(OvisWorld *w) {
    OVIS_CUDA(cudaSetDevice(w->device));
    return OVIS_OK;
}

template <typename T>
int grow(T *&ptr, int64_t old_items, int64_t new_items, cudaStream_t s, bool zero_new = true) {
    T *fresh = nullptr;
    OVIS_CUDA(cudaMalloc(&fresh, sizeof(T) * (size_t)new_items));
    if (zero_new)
        OVIS_CUDA(cudaMemsetAsync(fresh, 0, sizeof(T) * (size_t)new_items, s));
    if (ptr && old_items > 0)
        OVIS_CUDA(cudaMemcpyAsync(fresh, ptr, sizeof(T) * (size_t)old_items, cudaMemcpyDeviceToDevice, s));
    if (ptr) {
        OVIS_CUDA(cudaStreamSynchronize(s));
        OVIS_CUDA(cudaFree(ptr));
    }
    ptr = fresh;
    return OVIS_OK;
}

// SoA block of `comps` component arrays: copy each component to its new stride.
int grow_soa(float *&ptr, int comps, int64_t old_cap, int64_t new_cap, int64_t used, cudaStream_t s) {
    float *fresh = nullptr;
    OVIS_CUDA(cudaMalloc(&fresh, sizeof(float) * (size_t)(comps * new_cap)));
    OVIS_CUDA(cudaMemsetAsync(fresh, 0, sizeof(float) * (size_t)(comps * new_cap), s));
    if (ptr && used > 0)
        for (int k = 0; k < comps; k++)
            OVIS_CUDA(cudaMemcpyAsync(fresh + k * new_cap, ptr + k * old_cap, sizeof(float) * (size_t)used,
                                      cudaMemcpyDeviceToDevice, s));
    if (ptr) {
        OVIS_CUDA(cudaStreamSynchronize(s));
        OVIS_CUDA(cudaFree(ptr));
    }
    ptr = fresh;
    return OVIS_OK;
}

int stage(OvisWorld *w, int which, const void *host, size_t bytes, void **out) {
    if (w->stage_bytes[which] < bytes) {
        if (w->d_stage[which]) {
            OVIS_CUDA(cudaStreamSynchronize(w->stream));
            OVIS_CUDA(cudaFree(w->d_stage[which]));
            w->d_stage[which] = nullptr;
            w->stage_bytes[which] = 0;
        }
        size_t want = std::max(bytes, (size_t)1 << 20);
        OVIS_CUDA(cudaMalloc(&w->d_stage[which], want));
        w->stage_bytes[which] = want;
    }
    OVIS_CUDA(cudaMemcpyAsync(w->d_stage[which], host, bytes, cudaMemcpyHostToDevice, w->stream));
    *out = w->d_stage[which];
    return OVIS_OK;
}

// A set of buffers that grow together: every new block is allocated (and filled from the old one) first; only when all of
// them exist are the old ones released and the pointers swapped, so an out-of-memory half way leaves the world as it was.
struct GrowSet {
    struct Item { void **slot; void *fresh; };
    std::vector<Item> items;
    cudaStream_t stream;
    bool failed = false;

    explicit GrowSet(cudaStream_t s) : stream(s) {}
    template <typename T>
    void add(T *&ptr, int64_t old_items, int64_t new_items, bool zero_new = true, int fill_byte_new = 0) {
        if (failed)
            return;
        T *fresh = nullptr;
        if (cudaMalloc(&fresh, sizeof(T) * (size_t)new_items) != cudaSuccess) {
            cudaGetLastError();
            failed = true;
            return;
        }
        items.push_back({ reinterpret_cast<void **>(&ptr), fresh });
        if (zero_new || fill_byte_new)
            cudaMemsetAsync(fresh, fill_byte_new, sizeof(T) * (size_t)new_items, stream);
        if (ptr && old_items > 0)
            cudaMemcpyAsync(fresh, ptr, sizeof(T) * (size_t)old_items, cudaMemcpyDeviceToDevice, stream);
    }
    void add_soa(float *&ptr, int comps, int64_t old_cap, int64_t new_cap, int64_t used) {
        if (failed)
            return;
        float *fresh = nullptr;
        if (cudaMalloc(&fresh, sizeof(float) * (size_t)(comps * new_cap)) != cudaSuccess) {
            cudaGetLastError();
            failed = true;
            return;
        }
        items.push_back({ reinterpret_cast<void **>(&ptr), fresh });
        cudaMemsetAsync(fresh, 0, sizeof(float) * (size_t)(comps * new_cap), stream);
        if (ptr && used > 0)
            for (int k = 0; k < comps; k++)
                cudaMemcpyAsync(fresh + k * new_cap, ptr + k * old_cap, sizeof(float) * (size_t)used, cudaMemcpyDeviceToDevice, stream);
    }
    int commit() {
        cudaError_t err = cudaStreamSynchronize(stream);
        if (failed || err != cudaSuccess) {
            for (Item &i : items)
                cudaFree(i.fresh);
            return fail(OVIS_ERR_NOMEM, "out of device memory while growing the world (nothing was changed)");
        }
        for (Item &i : items) {
            if (*i.slot)
                cudaFree(*i.slot);
            *i.slot = i.fresh;
        }
        return OVIS_OK;
    }
};

int reserve_entities(OvisWorld *w, int64_t want) {
    if (want <= w->cap_ent)
        return OVIS_OK;
    const int64_t cap = round_up(std::max<int64_t>(want, w->cap_ent + w->cap_ent / 2), kCullTile);
    const int64_t blocks = cap / ovis::kMaskBlockEntities + 1;
    GrowSet g(w->stream);
    g.add_soa(w->d_local, 6, w->cap_ent, cap, w->n_ent);
    g.add_soa(w->d_world, 6, w->cap_ent, cap, w->n_ent);
    g.add(w->d_alive_bits, w->cap_ent / 32, cap / 32);
    g.add(w->d_caster_bits, w->cap_ent / 32, cap / 32);
    g.add(w->d_pass_counts, w->n_ent, cap);
    g.add(w->d_entity_node, w->n_ent, cap, false, 0xff);           // kNone everywhere, then the old values
    g.add(w->d_draw_totals, 0, 2 * (cap / 1024 + 1), false);
    if (!w->d_draw_counts)
        g.add(w->d_draw_counts, 0, 2);
    g.add(w->d_result, 0, cap, false);
    g.add(w->d_survivors, 0, 2 * (cap / 32), false);
    g.add(w->d_block_totals, 0, 3 * blocks);                         // all sets start zeroed
    OVIS_TRY(g.commit());
    w->cap_blocks = blocks;
    w->cull_serial = 0;
    w->cap_ent = cap;
    w->h_alive.resize((size_t)(cap / 32), 0u);
    w->h_caster_raw.resize((size_t)(cap / 32), 0u);
    return OVIS_OK;
}

int reserve_nodes(OvisWorld *w, int64_t want) {
    if (want <= w->cap_nodes)
        return OVIS_OK;
    int64_t cap = round_up(std::max<int64_t>(want, w->cap_nodes + w->cap_nodes / 2), kCullTile);
    OVIS_TRY(grow_soa(w->d_local_trs, 10, w->cap_nodes, cap, w->n_nodes, w->stream));
    OVIS_TRY(grow_soa(w->d_world_trs, 10, w->cap_nodes, cap, w->n_nodes, w->stream));
    OVIS_TRY(grow(w->d_parent, w->n_nodes, cap, w->stream));
    OVIS_TRY(grow(w->d_node_entity, w->n_nodes, cap, w->stream));
    OVIS_TRY(grow(w->d_dirty, w->n_nodes, cap, w->stream));
    w->cap_nodes = cap;
    return OVIS_OK;
}

void mark_masks(OvisWorld *w, int64_t first_id, int64_t last_id) {
    const int64_t lo = first_id >> 5, hi = (last_id >> 5) + 1;
    if (!w->masks_dirty) {
        w->masks_dirty_lo = lo;
        w->masks_dirty_hi = hi;
    } else {
        w->masks_dirty_lo = std::min(w->masks_dirty_lo, lo);
        w->masks_dirty_hi = std::max(w->masks_dirty_hi, hi);
    }
    w->masks_dirty = true;
}

// Only the range of mask words that changed goes up (100 M entities are 12.5 MB per mask: a single removal must not
// resend that).
int upload_masks(OvisWorld *w) {
    if (!w->masks_dirty)
        return OVIS_OK;
    const int64_t all = round_up(w->n_ent, 32) / 32;
    const int64_t lo = std::max<int64_t>(0, w->masks_dirty_lo), hi = std::min<int64_t>(all, w->masks_dirty_hi);
    if (hi > lo) {
        const size_t words = (size_t)(hi - lo);
        std::vector<uint32_t> caster(words);
        for (size_t i = 0; i < words; i++)
            caster[i] = w->h_caster_raw[(size_t)lo + i] & w->h_alive[(size_t)lo + i];
        OVIS_CUDA(cudaMemcpyAsync(w->d_alive_bits + lo, w->h_alive.data() + lo, words * 4, cudaMemcpyHostToDevice, w->stream));
        OVIS_CUDA(cudaMemcpyAsync(w->d_caster_bits + lo, caster.data(), words * 4, cudaMemcpyHostToDevice, w->stream));
        OVIS_CUDA(cudaStreamSynchronize(w->stream));   // `caster` is a local
    }
    w->masks_dirty = false;
    return OVIS_OK;
}

int check_entity_ids(OvisWorld *w, int64_t n, const uint32_t *ids, bool need_alive = true) {
    if (n < 0 || (n > 0 && !ids))
        return fail(OVIS_ERR_INVALID, "null id array");
    for (int64_t i = 0; i < n; i++) {
        uint32_t id = ids[i];
        if ((int64_t)id >= w->n_ent || (need_alive && !((w->h_alive[id >> 5] >> (id & 31)) & 1u)))
            return fail(OVIS_ERR_INVALID, "entity id %u does not name a live entity", id);
    }
    return OVIS_OK;
}

int add_nodes(OvisWorld *w, int64_t n, const int32_t *parent, const float *local_trs, uint32_t *out_first) {
    if (n < 0 || (n > 0 && (!parent || !local_trs)))
        return fail(OVIS_ERR_INVALID, "ovis_node_add: null input");
    int64_t first = w->n_nodes;
    if (first + n > (int64_t)0x7fffffff)
        return fail(OVIS_ERR_INVALID, "too many nodes");
    for (int64_t i = 0; i < n; i++)
        if (parent[i] != OVIS_NO_PARENT && (parent[i] < 0 || (int64_t)parent[i] >= first + i))
            return fail(OVIS_ERR_INVALID, "node %lld: parent %d is not an earlier node", (long long)(first + i), parent[i]);
    OVIS_TRY(reserve_nodes(w, first + n));
    w->h_parent.resize((size_t)(first + n));
    w->h_level.resize((size_t)(first + n));
    w->h_node_entity.resize((size_t)(first + n), kNone);
    w->h_subtree.resize((size_t)(first + n), 1u);
    if (!w->h_node_dead.empty())
        w->h_node_dead.resize((size_t)(first + n), 0);
    for (int64_t i = 0; i < n; i++) {
        int32_t p = parent[i];
        w->h_parent[(size_t)(first + i)] = p;
        w->h_level[(size_t)(first + i)] = p < 0 ? 0 : w->h_level[(size_t)p] + 1;
        int hops = 0;                                                 // an estimate is enough: stop after 64 ancestors
        for (int32_t a = p; a >= 0 && hops < 64; a = w->h_parent[(size_t)a], hops++)
            w->h_subtree[(size_t)a]++;
    }
    for (int64_t s = 0; s < n; s += kUploadChunk) {
        int64_t c = std::min(kUploadChunk, n - s);
        void *d = nullptr;
        OVIS_TRY(stage(w, 0, local_trs + 10 * s, sizeof(float) * 10 * (size_t)c, &d));
        ovis::launch_scatter_aos(10, (const float *)d, nullptr, (uint32_t)(first + s), c, w->d_local_trs, w->cap_nodes, w->stream);
        w->launches++;
    }
    OVIS_CUDA(cudaMemcpyAsync(w->d_parent + first, parent, sizeof(int32_t) * (size_t)n, cudaMemcpyHostToDevice, w->stream));
    ovis::launch_fill_u32(w->d_node_entity + first, n, kNone, w->stream);
    w->launches++;
    OVIS_CUDA(cudaGetLastError());
    w->n_nodes = first + n;
    w->levels_dirty = true;
    w->transforms_dirty = true;
    w->all_dirty = true;
    if (out_first)
        *out_first = (uint32_t)first;
    return OVIS_OK;
}

int rebuild_levels(OvisWorld *w) {
    if (!w->levels_dirty)
        return OVIS_OK;
    auto dead = [&](int64_t i) { return !w->h_node_dead.empty() && w->h_node_dead[(size_t)i] != 0; };
    int32_t depth = 0;
    for (int64_t i = 0; i < w->n_nodes; i++)
        if (!dead(i))
            depth = std::max(depth, w->h_level[(size_t)i] + 1);
    w->levels.assign((size_t)depth, Level());
    std::vector<int64_t> last((size_t)depth, -1);
    for (int64_t i = 0; i < w->n_nodes; i++) {
        if (dead(i))
            continue;
        Level &l = w->levels[(size_t)w->h_level[(size_t)i]];
        if (l.count == 0)
            l.first = i;
        l.count++;
        last[(size_t)w->h_level[(size_t)i]] = i;
    }
    int64_t list_items = 0;
    for (int32_t d = 0; d < depth; d++) {
        Level &l = w->levels[(size_t)d];
        l.contiguous = (last[(size_t)d] - l.first + 1 == l.count);
        if (!l.contiguous) {
            l.list_offset = list_items;
            list_items += l.count;
        }
    }
    if (list_items > 0) {
        std::vector<uint32_t> lists((size_t)list_items);
        std::vector<int64_t> cursor((size_t)depth);
        for (int32_t d = 0; d < depth; d++)
            cursor[(size_t)d] = w->levels[(size_t)d].list_offset;
        for (int64_t i = 0; i < w->n_nodes; i++) {
            if (dead(i))
                continue;
            int32_t d = w->h_level[(size_t)i];
            if (!w->levels[(size_t)d].contiguous)
                lists[(size_t)cursor[(size_t)d]++] = (uint32_t)i;
        }
        if (list_items > w->cap_level_lists) {
            OVIS_TRY(grow(w->d_level_lists, 0, list_items, w->stream, false));
            w->cap_level_lists = list_items;
        }
        OVIS_CUDA(cudaMemcpyAsync(w->d_level_lists, lists.data(), sizeof(uint32_t) * (size_t)list_items,
                                  cudaMemcpyHostToDevice, w->stream));
        OVIS_CUDA(cudaStreamSynchronize(w->stream));
    }
    w->levels_dirty = false;
    return OVIS_OK;
}

void mark_lights(OvisWorld *w, int64_t lo, int64_t hi);

// Lights bound to nodes (ovis_light_bind_node) take their position and direction from the node's world transform, as
// Light::transformed does (graphics/src/light.cc:100-108): position = worldPosition, direction = worldOrientation *
// (0, 0, -1). Lights are host-side state (their planes, cone boxes and uniform blocks are host math), so the bound
// nodes' world TRS comes back in one small copy after the tree update.
int follow_nodes(OvisWorld *w) {
    w->light_bindings_dirty = false;
    if (w->n_bound_lights == 0)
        return OVIS_OK;
    std::vector<uint32_t> nodes, lights;
    for (size_t i = 0; i < w->light_node.size(); i++)
        if (w->light_node[i] >= 0 && w->light_alive[i]) {
            nodes.push_back((uint32_t)w->light_node[i]);
            lights.push_back((uint32_t)i);
        }
    if (nodes.empty())
        return OVIS_OK;
    const size_t n = nodes.size();
    void *d_ids = nullptr;
    OVIS_TRY(stage(w, 1, nodes.data(), sizeof(uint32_t) * n, &d_ids));
    if (w->stage_bytes[0] < sizeof(float) * 10 * n) {
        if (w->d_stage[0]) {
            OVIS_CUDA(cudaStreamSynchronize(w->stream));
            OVIS_CUDA(cudaFree(w->d_stage[0]));
            w->d_stage[0] = nullptr;
            w->stage_bytes[0] = 0;
        }
        size_t want = std::max(sizeof(float) * 10 * n, (size_t)1 << 20);
        OVIS_CUDA(cudaMalloc(&w->d_stage[0], want));
        w->stage_bytes[0] = want;
    }
    ovis::launch_gather_trs(w->d_world_trs, w->cap_nodes, (const uint32_t *)d_ids, (int64_t)n, (float *)w->d_stage[0], w->stream);
    w->launches++;
    std::vector<float> trs(10 * n);
    OVIS_CUDA(cudaMemcpyAsync(trs.data(), w->d_stage[0], sizeof(float) * 10 * n, cudaMemcpyDeviceToHost, w->stream));
    OVIS_CUDA(cudaStreamSynchronize(w->stream));
    for (size_t k = 0; k < n; k++) {
        const float *t = trs.data() + 10 * k;
        OvisLight &l = w->lights[lights[k]];
        const ovis::hm::Vec3 d = ovis::hm::rotate(ovis::hm::Quat{ t[3], t[4], t[5], t[6] }, ovis::hm::Vec3{ 0.0f, 0.0f, -1.0f });
        const float pos[3] = { t[0], t[1], t[2] }, dir[3] = { d.x, d.y, d.z };
        if (std::memcmp(l.position, pos, sizeof pos) != 0 || std::memcmp(l.direction, dir, sizeof dir) != 0) {
            std::memcpy(l.position, pos, sizeof pos);
            std::memcpy(l.direction, dir, sizeof dir);
            mark_lights(w, lights[k], (int64_t)lights[k] + 1);
        }
    }
    return OVIS_OK;
}

int flush_transforms(OvisWorld *w) {
    if (!w->transforms_dirty)
        return w->light_bindings_dirty ? follow_nodes(w) : OVIS_OK;
    OVIS_TRY(rebuild_levels(w));
    if (w->extra_dirty) {
        int64_t n = (int64_t)w->h_extra_entity.size();
        if (n > w->cap_extra) {
            OVIS_TRY(grow(w->d_extra_entity, 0, n, w->stream, false));
            OVIS_TRY(grow(w->d_extra_node, 0, n, w->stream, false));
            w->cap_extra = n;
        }
        if (n > 0) {
            OVIS_CUDA(cudaMemcpyAsync(w->d_extra_entity, w->h_extra_entity.data(), 4 * (size_t)n, cudaMemcpyHostToDevice, w->stream));
            OVIS_CUDA(cudaMemcpyAsync(w->d_extra_node, w->h_extra_node.data(), 4 * (size_t)n, cudaMemcpyHostToDevice, w->stream));
        }
        w->extra_dirty = false;
    }
    for (const Level &l : w->levels) {
        ovis::LevelArgs a;
        a.list = l.contiguous ? nullptr : w->d_level_lists + l.list_offset;
        a.first = l.first;
        a.count = l.count;
        a.parent = w->d_parent;
        a.local_trs = w->d_local_trs;
        a.world_trs = w->d_world_trs;
        a.nstride = w->cap_nodes;
        a.dirty = w->all_dirty ? nullptr : w->d_dirty;
        a.node_entity = w->d_node_entity;
        a.local_aabb = w->d_local;
        a.world_aabb = w->d_world;
        a.estride = w->cap_ent;
        if (l.count <= 0)
            continue;
        ovis::launch_update_level(a, w->stream);
        w->launches++;
    }
    if (!w->h_extra_entity.empty()) {
        ovis::launch_extra_bounds(w->d_extra_entity, w->d_extra_node, (int64_t)w->h_extra_entity.size(), w->d_world_trs,
                                  w->cap_nodes, w->d_local, w->d_world, w->cap_ent, w->all_dirty ? nullptr : w->d_dirty,
                                  w->stream);
        w->launches++;
    }
    OVIS_CUDA(cudaGetLastError());
    if (w->marks_present)
        OVIS_CUDA(cudaMemsetAsync(w->d_dirty, 0, (size_t)w->n_nodes, w->stream));
    w->marks_present = false;
    w->all_dirty = false;
    w->transforms_dirty = false;
    return follow_nodes(w);
}

int bind_entities(OvisWorld *w, int64_t n, const uint32_t *nodes, const uint32_t *entities) {
    std::vector<uint32_t> first_nodes, first_entities;
    for (int64_t i = 0; i < n; i++) {
        uint32_t node = nodes[i], ent = entities[i];
        if (w->h_entity_node[ent] != kNone)
            return fail(OVIS_ERR_INVALID, "entity %u is already bound to node %u", ent, w->h_entity_node[ent]);
        w->h_entity_node[ent] = node;
        if (w->h_node_entity[node] == kNone) {
            w->h_node_entity[node] = ent;
            first_nodes.push_back(node);
            first_entities.push_back(ent);
        } else {
            w->h_extra_entity.push_back(ent);
            w->h_extra_node.push_back(node);
            w->extra_dirty = true;
        }
    }
    for (size_t s = 0; s < first_nodes.size(); s += (size_t)kUploadChunk) {
        size_t c = std::min((size_t)kUploadChunk, first_nodes.size() - s);
        void *dn = nullptr, *de = nullptr;
        OVIS_TRY(stage(w, 0, first_entities.data() + s, 4 * c, &de));
        OVIS_TRY(stage(w, 1, first_nodes.data() + s, 4 * c, &dn));
        ovis::launch_scatter_u32((const uint32_t *)de, (const uint32_t *)dn, (int64_t)c, w->d_node_entity, w->stream);
        w->launches++;
        OVIS_CUDA(cudaStreamSynchronize(w->stream));   // vectors are locals
    }
    for (int64_t s = 0; s < n; s += kUploadChunk) {              // every entity, first on its node or not
        int64_t c = std::min(kUploadChunk, n - s);
        void *dn = nullptr, *de = nullptr;
        OVIS_TRY(stage(w, 0, nodes + s, 4 * (size_t)c, &dn));
        OVIS_TRY(stage(w, 1, entities + s, 4 * (size_t)c, &de));
        ovis::launch_scatter_u32((const uint32_t *)dn, (const uint32_t *)de, c, w->d_entity_node, w->stream);
        w->launches++;
        OVIS_CUDA(cudaStreamSynchronize(w->stream));   // the staging buffers are reused by the next chunk
    }
    w->transforms_dirty = true;
    w->all_dirty = true;
    return OVIS_OK;
}

void load_planes(float4 dst[6], const float planes[24]) {
    for (int i = 0; i < 6; i++)
        dst[i] = make_float4(planes[4 * i], planes[4 * i + 1], planes[4 * i + 2], planes[4 * i + 3]);
}

// Enqueue one entity cull (K3). Output goes to `out` (device) starting at *d_base (or 0).
struct FlagOpts {            // one-way ready flags of a sharded exchange (ovis_cull_planes_device_exchange)
    int n_wait = 0;
    const uint32_t *wait_flags = nullptr;
    uint32_t wait_value = 0;
    int n_signal = 0;
    uint32_t *const *signal_flags = nullptr;
    uint32_t signal_value = 0;
};

int enqueue_cull(OvisWorld *w, const float planes[24], uint32_t require, uint32_t *out, long long capacity,
                 const long long *d_base, long long *d_next, uint32_t *d_count, int n_extra = 0,
                 uint32_t *const *extra_out = nullptr, uint32_t *const *extra_count = nullptr,
                 const FlagOpts *flags = nullptr, long long extra_capacity = -1) {
    if (require & ~(uint32_t)OVIS_CASTS_SHADOW)
        return fail(OVIS_ERR_INVALID, "require mask 0x%x: only 0 or OVIS_CASTS_SHADOW is supported", require);
    if (w->n_ent == 0) {
        ovis::launch_empty_result(d_base, d_next, d_count, w->stream);
        w->launches++;
        for (int x = 0; x < n_extra; x++) {
            ovis::launch_empty_result(nullptr, nullptr, extra_count[x], w->stream);
            w->launches++;
        }
        return OVIS_OK;
    }
    OVIS_TRY(upload_masks(w));
    // small worlds are latency-bound: smaller warp tiles spread them over more warps
    // large worlds: 256-entity tiles with the sign-specialised, FMUL2-packed plane code (profiles/r01_cull_notes.md)
    const int variant = w->forced_variant >= 0 ? w->forced_variant : ovis::cull_default_variant(w->n_ent);
    const bool masked = require != 0 || w->any_dead;
    ovis::CullArgs a{};
    load_planes(a.planes, planes);
    a.sign_codes = ovis::plane_sign_codes(planes);
    a.bounds = w->d_world;
    a.stride = w->cap_ent;
    a.n = w->n_ent;
    a.num_groups = (int)((w->n_ent + 127) / 128);
    a.num_blocks = (int)((w->n_ent + ovis::kMaskBlockEntities - 1) / ovis::kMaskBlockEntities);
    a.mask = require ? w->d_caster_bits : w->d_alive_bits;
    const int64_t set2 = (int64_t)(w->cull_serial & 1), set3 = (int64_t)(w->cull_serial % 3);
    a.survivors = w->d_survivors + set2 * (w->cap_ent / 32);
    a.block_totals = w->d_block_totals + set3 * w->cap_blocks;
    // zeroed by this test kernel for the next cull; last read by the gather kernel two culls ago, which has finished:
    // the previous test kernel ended with griddepcontrol.wait, i.e. not before that gather kernel had completed
    a.block_totals_next = w->d_block_totals + ((set3 + 1) % 3) * w->cap_blocks;
    a.n_wait = 0;
    a.wait_flags = nullptr;
    a.wait_value = 0;
    a.error_flag = w->d_exchange + 1;
    if (flags && flags->n_wait > 0) {                            // checked by the test kernel, see cull_test_kernel
        a.n_wait = flags->n_wait;
        a.wait_flags = flags->wait_flags;
        a.wait_value = flags->wait_value;
    }
    int grid = ovis::cull_test_grid(variant, a.num_groups, w->num_sms * w->blocks_per_sm[variant][masked ? 1 : 0]);
    cudaEvent_t *ev = nullptr;
    if (w->timing) {
        if (w->timing_used + 3 > w->timing_events.size()) {
            size_t old = w->timing_events.size();
            w->timing_events.resize(old + 3 * 256);
            for (size_t i = old; i < w->timing_events.size(); i++)
                OVIS_CUDA(cudaEventCreate(&w->timing_events[i]));
        }
        ev = &w->timing_events[w->timing_used];
        w->timing_used += 3;
        OVIS_CUDA(cudaEventRecord(ev[0], w->stream));
    }
    ovis::GatherArgs g{};
    g.survivors = a.survivors;
    g.block_totals = a.block_totals;
    g.num_words = (int64_t)a.num_groups * 4;
    g.num_blocks = a.num_blocks;
    g.id_base = w->id_base;
    g.out = out;
    g.capacity = capacity;
    g.d_base = d_base;
    g.d_next = d_next;
    g.d_count = d_count;
    g.n_extra = n_extra;
    g.extra_capacity = extra_capacity >= 0 ? extra_capacity : capacity;
    for (int x = 0; x < n_extra; x++) {
        g.extra_out[x] = extra_out[x];
        g.extra_count[x] = extra_count[x];
    }
    g.n_wait = 0;
    g.n_signal = 0;
    g.block_fence_sys = w->block_fence_sys ? 1 : 0;
    g.done_counter = w->d_exchange;
    g.error_flag = w->d_exchange + 1;
    if (flags) {
        g.n_wait = 0;                                            // the test kernel has waited already
        g.n_signal = flags->n_signal;
        for (int x = 0; x < flags->n_signal; x++)
            g.signal_flags[x] = flags->signal_flags[x];
        g.signal_value = flags->signal_value;
    }
    // Below ~half a million entities the cull is pure latency: one cooperative launch runs both phases with a grid
    // barrier in between. Larger worlds use two launches, the second overlapped with the next cull's first.
    const int fused_per_sm = w->fused_blocks_per_sm[variant][masked ? 1 : 0];
    const bool fused = w->fused_small && w->n_ent < w->fused_below && fused_per_sm > 0 && (!flags || flags->n_wait == 0);
    if (fused) {
        grid = ovis::cull_test_grid(variant, a.num_groups, w->num_sms * fused_per_sm);   // every CTA must be co-resident
        OVIS_CUDA(ovis::launch_cull_fused(a, g, masked, variant, grid, w->stream));
        if (ev) {
            OVIS_CUDA(cudaEventRecord(ev[1], w->stream));
            OVIS_CUDA(cudaEventRecord(ev[2], w->stream));
        }
        w->launches += 1;
    } else {
        // The test kernel may overlap the gather kernel in front of it only if that is really what is in front of it: any
        // other kernel this world enqueued since (transform levels, scatters, the light cull ...) may have written what
        // the test kernel reads, and must have completed first.
        const bool after_own_gather = w->launches == w->launches_after_gather;
        ovis::launch_cull_test(a, masked, variant, grid, w->stream, /*overlap_previous_gather=*/after_own_gather);
        if (ev)
            OVIS_CUDA(cudaEventRecord(ev[1], w->stream));
        ovis::launch_cull_gather(g, w->stream);
        if (ev)
            OVIS_CUDA(cudaEventRecord(ev[2], w->stream));
        w->launches += 2;
        w->launches_after_gather = w->launches;
    }
    w->cull_serial++;
    OVIS_CUDA(cudaGetLastError());
    return OVIS_OK;
}

void mark_lights(OvisWorld *w, int64_t lo, int64_t hi) {
    w->light_cull_valid = false;
    for (int64_t i = std::max<int64_t>(lo, 0); i < hi && i < (int64_t)w->light_views.size(); i++)
        w->light_views[(size_t)i].resolution = 0;             // the cached shadow views of these lights are stale
    if (!w->lights_dirty) {
        w->lights_dirty_lo = lo;
        w->lights_dirty_hi = hi;
    } else {
        w->lights_dirty_lo = std::min(w->lights_dirty_lo, lo);
        w->lights_dirty_hi = std::max(w->lights_dirty_hi, hi);
    }
    w->lights_dirty = true;
}

// Device copies of the lights that changed: cull record, LightUniforms block, type. Only the dirty slot range is
// recomputed and uploaded (a light that moves every frame does not make the other 4095 pay).
int upload_lights(OvisWorld *w) {
    if (!w->lights_dirty)
        return OVIS_OK;
    int64_t n = (int64_t)w->lights.size();
    if (n > w->cap_lights) {
        int64_t cap = std::max<int64_t>(n, 64);
        OVIS_TRY(grow(w->d_lights, 0, cap, w->stream, false));
        OVIS_TRY(grow(w->d_light_out, 0, cap, w->stream, false));
        OVIS_TRY(grow(w->d_light_bits, 0, cap / 32 + 16, w->stream));          // ballot words + (last word) the finished-CTA counter
        OVIS_TRY(grow(w->d_light_uniforms, 0, 52 * cap, w->stream, false));
        OVIS_TRY(grow(w->d_light_types, 0, cap, w->stream, false));
        w->cap_lights = cap;
        w->lights_dirty_lo = 0;                                  // fresh buffers: everything goes up again
        w->lights_dirty_hi = n;
    }
    const int64_t lo = std::max<int64_t>(0, w->lights_dirty_lo), hi = std::min<int64_t>(n, w->lights_dirty_hi);
    const int64_t m = std::max<int64_t>(0, hi - lo);
    std::vector<ovis::DeviceLight> dl((size_t)m);
    for (int64_t i = lo; i < hi; i++) {
        const OvisLight &l = w->lights[(size_t)i];
        ovis::DeviceLight &d = dl[(size_t)(i - lo)];
        std::memset(&d, 0, sizeof d);
        if (!w->light_alive[(size_t)i] || ovis::hm::lightHasNoEffect(l)) {          // render_light.cc:189-190
            d.kind = ovis::kLightNever;
        } else if (l.type == OVIS_LIGHT_AMBIENT || l.type == OVIS_LIGHT_DIRECTIONAL) {
            d.kind = ovis::kLightAlways;
        } else if (l.type == OVIS_LIGHT_POINT) {
            d.kind = ovis::kLightSphere;
            d.a = make_float4(l.position[0], l.position[1], l.position[2], l.range);
        } else if (l.type == OVIS_LIGHT_SPOT) {
            float box[6];
            ovis::hm::spotAabb(l, box);
            d.kind = ovis::kLightBox;
            d.a = make_float4(box[0], box[1], box[2], 0.0f);
            d.b = make_float4(box[3], box[4], box[5], 0.0f);
        } else {
            d.kind = ovis::kLightNever;                                              // render_light.cc:206-207
        }
    }
    std::vector<float> blocks(52 * (size_t)m);
    std::vector<int32_t> types((size_t)m);
    for (int64_t i = lo; i < hi; i++)
        types[(size_t)(i - lo)] = w->lights[(size_t)i].type;
    for (int64_t i = lo; i < hi; i++)
        ovis::hm::lightUniforms(w->lights[(size_t)i], w->light_shading[(size_t)i].attenuation, w->light_shading[(size_t)i].shadow_bias,
                                blocks.data() + 52 * (size_t)(i - lo));
    for (int64_t i = 0; i < m; i++) {                            // see kLightSkippable
        const float *b = blocks.data() + 52 * (size_t)i;
        const int32_t type = types[(size_t)i];
        if (type != OVIS_LIGHT_POINT && type != OVIS_LIGHT_SPOT)
            continue;
        bool ok = true;
        for (int k = 0; k < 52; k++)
            ok = ok && std::isfinite(b[k]);
        for (int k = 0; k < 3; k++)
            ok = ok && std::fabs(b[k]) <= 1e18f && std::fabs(b[8 + k]) <= 1e6f;
        ok = ok && std::fabs(b[3]) <= 1e6f && b[45] > 0.0f && b[46] >= 0.0f && b[47] >= 0.0f;
        if (type == OVIS_LIGHT_SPOT)
            ok = ok && b[7] < 1.0f;
        if (ok)
            types[(size_t)i] |= ovis::kLightSkippable;
    }
    if (m > 0) {
        OVIS_CUDA(cudaMemcpyAsync(w->d_lights + lo, dl.data(), sizeof(ovis::DeviceLight) * (size_t)m, cudaMemcpyHostToDevice, w->stream));
        OVIS_CUDA(cudaMemcpyAsync(w->d_light_uniforms + 52 * lo, blocks.data(), sizeof(float) * blocks.size(), cudaMemcpyHostToDevice,
                                  w->stream));
        OVIS_CUDA(cudaMemcpyAsync(w->d_light_types + lo, types.data(), sizeof(int32_t) * types.size(), cudaMemcpyHostToDevice, w->stream));
        OVIS_CUDA(cudaStreamSynchronize(w->stream));           // the vectors are locals
    }
    w->lights_dirty = false;
    return OVIS_OK;
}

int enqueue_light_cull(OvisWorld *w, const float planes[24]) {
    OVIS_TRY(upload_lights(w));
    ovis::LightCullArgs a;
    load_planes(a.planes, planes);
    a.lights = w->d_lights;
    a.n = (int)w->lights.size();
    a.out = w->d_light_out;
    a.d_count = w->d_counts + 1;
    a.bits = w->d_light_bits;
    a.counter = w->d_light_bits ? w->d_light_bits + (w->cap_lights / 32 + 15) : nullptr;
    ovis::launch_cull_lights(a, w->stream);
    w->launches++;
    OVIS_CUDA(cudaGetLastError());
    return OVIS_OK;
}

// A group of consecutive views that share a pre-filter cube (a point light's six faces), or a single plain view.
struct HostGroup {
    float lo[3], hi[3];          // the cube, already inflated by its rounding margin; meaningless when !cube
    int first_view, n_views;
    bool cube;
};

// Pre-filter cube of a view group (nullptr centre = none). The entity side adds its own extent and margin in the kernel.
HostGroup make_group(const float *centre, float range, int first_view, int n_views) {
    HostGroup g{};
    g.cube = false;
    if (centre) {
        float pmax = std::max(std::fabs(centre[0]), std::max(std::fabs(centre[1]), std::fabs(centre[2])));
        float rg = range * 1.001f + pmax * 1.0e-3f + 1.0e-2f;
        g.cube = std::isfinite(rg) && std::isfinite(pmax);                    // NaN / infinite light: no filtering
        for (int c = 0; c < 3; c++) {
            g.lo[c] = centre[c] - rg;
            g.hi[c] = centre[c] + rg;
            g.cube = g.cube && std::isfinite(g.lo[c]) && std::isfinite(g.hi[c]);
        }
    }
    g.first_view = first_view;
    g.n_views = n_views;
    return g;
}

template <typename T>
int ensure(T *&ptr, int64_t &cap, int64_t want, cudaStream_t s) {
    if (want <= cap)
        return OVIS_OK;
    int64_t fresh = std::max<int64_t>(want, cap + cap / 2);
    OVIS_TRY(grow(ptr, 0, fresh, s, false));
    cap = fresh;
    return OVIS_OK;
}

int ensure_pinned(void *&ptr, size_t &cap, size_t want) {
    if (want <= cap)
        return OVIS_OK;
    if (ptr)
        OVIS_CUDA(cudaFreeHost(ptr));
    ptr = nullptr;
    cap = 0;
    size_t fresh = std::max(want + want / 2, (size_t)1 << 16);
    OVIS_CUDA(cudaMallocHost(&ptr, fresh));
    cap = fresh;
    return OVIS_OK;
}

// Uniform grid over the union of the cubes of groups[ids]: cell size = the larger of (largest extent / 64) and a quarter
// of the smallest cube side, coarsened until the per-cell lists need at most `node_budget` nodes. Fills the groups' cell
// ranges with the same monotone cell function the kernel applies to the entities (mv cell_of).
int64_t build_light_grid(std::vector<ovis::MvGroup> &groups, ovis::MvGrid &grid, int64_t node_budget) {
    const float inf = std::numeric_limits<float>::infinity();
    float min_side = inf;
    for (int c = 0; c < 3; c++) {
        grid.lo[c] = inf;
        grid.hi[c] = -inf;
    }
    for (const ovis::MvGroup &g : groups)
        for (int c = 0; c < 3; c++) {
            grid.lo[c] = std::min(grid.lo[c], g.lo[c]);
            grid.hi[c] = std::max(grid.hi[c], g.hi[c]);
            min_side = std::min(min_side, g.hi[c] - g.lo[c]);
        }
    float extent[3], largest = 0.0f;
    for (int c = 0; c < 3; c++) {
        extent[c] = grid.hi[c] - grid.lo[c];
        largest = std::max(largest, extent[c]);
    }
    // Three of the kMvGridMax cells per axis are border: the grid's lower bound maps to 1.5, so the cubes occupy cells
    // 1 .. dim - 2 whatever the rounding, and an entity outside the bounds clamps into cell 0 or dim - 1, which no cube marks:
    // the kernel needs no "outside the grid" test.
    const int usable = ovis::kMvGridMax - 3;
    float cell = std::max(largest / (float)usable, min_side * 0.25f);
    if (!(cell > 0.0f) || !std::isfinite(cell))
        cell = 1.0f;
    float reach = 0.0f;
    for (int c = 0; c < 3; c++)
        reach = std::max(reach, std::max(std::fabs(grid.lo[c]), std::fabs(grid.hi[c])));
    grid.re_bias = reach * 1.002e-3f;
    for (;;) {
        bool single = true;
        for (int c = 0; c < 3; c++) {
            float cells = std::ceil(extent[c] / cell);
            const int inner = (int)std::min(std::max(cells, 1.0f), (float)usable);
            single = single && inner == 1;
            grid.dim[c] = inner + 3;
            grid.inv_cell[c] = 1.0f / cell;
            grid.bias[c] = 1.5f - grid.lo[c] * grid.inv_cell[c];
        }
        int64_t nodes = 0;
        for (ovis::MvGroup &g : groups) {
            int lo[3], hi[3];
            for (int c = 0; c < 3; c++) {
                float a = std::floor(std::fmaf(g.lo[c], grid.inv_cell[c], grid.bias[c]));
                float b = std::floor(std::fmaf(g.hi[c], grid.inv_cell[c], grid.bias[c]));
                lo[c] = (int)std::min(std::max(a, 0.0f), (float)(grid.dim[c] - 1));
                hi[c] = (int)std::min(std::max(b, 0.0f), (float)(grid.dim[c] - 1));
            }
            g.cell_lo = (uint32_t)lo[0] | ((uint32_t)lo[1] << 8) | ((uint32_t)lo[2] << 16);
            g.cell_hi = (uint32_t)hi[0] | ((uint32_t)hi[1] << 8) | ((uint32_t)hi[2] << 16);
            nodes += (int64_t)(hi[0] - lo[0] + 1) * (hi[1] - lo[1] + 1) * (hi[2] - lo[2] + 1);
        }
        if (nodes <= node_budget || single)
            return nodes;
        cell *= 2.0f;
    }
}

// K4: all views of `groups` (consecutive, covering views 0 .. n_views - 1 in order); result in w->d_csr_* (CSR, device).
// One upload, the kernels, one read-back of the offsets and the overflow word: a single synchronisation per call. Views
// are processed in rounds only when the scratch of one round would exceed its budget (dense bitmaps of very many views).
int cull_views_grouped(OvisWorld *w, int32_t n_views, const float *planes, const uint32_t *require,
                       const std::vector<HostGroup> &groups, int64_t *host_offsets_out,
                       std::vector<long long> &host_offsets) {
    MvTrace trace;
    OVIS_TRY(flush_transforms(w));
    host_offsets.assign((size_t)n_views + 1, 0);
    OVIS_TRY(ensure(w->d_csr_offsets, w->cap_csr_views, std::max<int64_t>(n_views + 1, 64), w->stream));
    if (w->cap_csr == 0)
        OVIS_TRY(ensure(w->d_csr_entities, w->cap_csr, std::max<int64_t>(w->cap_ent, 1 << 20), w->stream));
    if (n_views == 0 || w->n_ent == 0) {
        OVIS_CUDA(cudaMemsetAsync(w->d_csr_offsets, 0, sizeof(long long) * ((size_t)n_views + 1), w->stream));
        OVIS_CUDA(cudaStreamSynchronize(w->stream));
        if (host_offsets_out)
            for (int32_t v = 0; v <= n_views; v++)
                host_offsets_out[v] = 0;
        return OVIS_OK;
    }
    OVIS_TRY(upload_masks(w));
    bool any_require = false;
    for (int32_t v = 0; v < n_views; v++) {
        uint32_t r = require ? require[v] : 0u;
        if (r & ~(uint32_t)OVIS_CASTS_SHADOW)
            return fail(OVIS_ERR_INVALID, "require mask 0x%x: only 0 or OVIS_CASTS_SHADOW is supported", r);
        any_require |= r != 0;
    }
    const int num_groups128 = (int)((w->n_ent + 127) / 128);
    const int num_blocks = (int)((w->n_ent + ovis::kMaskBlockEntities - 1) / ovis::kMaskBlockEntities);
    const int64_t num_words = (int64_t)num_groups128 * 4;
    const int64_t row = num_blocks + 1;
    // scratch budgets of one round
    const int64_t max_plain = std::max<int64_t>(1, ((int64_t)1 << 28) / num_words);          // 1 GiB of survivor bitmaps
    int64_t max_views = std::max<int64_t>(ovis::kMvMaxGroupViews, ((int64_t)1 << 28) / row);   // 1 GiB of block totals
    if (w->mv_round_views > 0)
        max_views = std::min<int64_t>(max_views, w->mv_round_views);
    if (!w->d_mv_state) {
        OVIS_CUDA(cudaMalloc(&w->d_mv_state, kMvStateBytes));
        OVIS_CUDA(cudaMemsetAsync(w->d_mv_state, 0, kMvStateBytes, w->stream));
        OVIS_CUDA(cudaMalloc(&w->d_mv_heads, sizeof(int32_t) * ovis::kMvGridMax * ovis::kMvGridMax * ovis::kMvGridMax));
    }
    bool force_dense = w->mv_force_dense;
    // The light grid pays when there are many (entity, view) pairs to avoid. A small world is tested against every view
    // faster than its entities can be looked up in the grid (20 000 entities x 72 faces: 18 us densely, and only three
    // blocks of 8192 entities for the sparse pass to spread over the GPU).
    int64_t cube_views = 0;
    for (const HostGroup &hg : groups)
        if (hg.cube)
            cube_views += hg.n_views;
    if (!w->mv_force_sparse && (int64_t)w->n_ent * cube_views <= kMvDenseBelowTests)
        force_dense = true;

    for (int attempt = 0;; attempt++) {
        if (attempt == 8)
            return fail(OVIS_ERR_NOMEM, "multi-view result did not fit after regrowing");
        OVIS_TRY(ensure_pinned(w->h_mv_result, w->cap_h_mv_result, sizeof(long long) * ((size_t)n_views + 5)));
        long long *hr = static_cast<long long *>(w->h_mv_result);
        size_t g0 = 0;
        while (g0 < groups.size()) {
            // ---- one round: groups [g0, g1), views [v0, v1)
            const int v0 = groups[g0].first_view;
            std::vector<uint32_t> plain;                 // view ids relative to v0
            std::vector<ovis::MvGroup> cubes;
            size_t g1 = g0;
            int v1 = v0;
            for (; g1 < groups.size(); g1++) {
                const HostGroup &hg = groups[g1];
                const bool sparse = hg.cube && !force_dense && hg.n_views <= ovis::kMvMaxGroupViews &&
                                    (int64_t)cubes.size() < ovis::kMvMaxGroups;
                if (g1 > g0 && ((v1 - v0) + hg.n_views > max_views || (!sparse && (int64_t)plain.size() + hg.n_views > max_plain)))
                    break;
                if (sparse) {
                    ovis::MvGroup mg{};
                    std::memcpy(mg.lo, hg.lo, sizeof mg.lo);
                    std::memcpy(mg.hi, hg.hi, sizeof mg.hi);
                    mg.first_view = hg.first_view - v0;
                    mg.n_views = hg.n_views;
                    cubes.push_back(mg);
                } else {
                    for (int k = 0; k < hg.n_views; k++)
                        plain.push_back((uint32_t)(hg.first_view - v0 + k));
                }
                v1 += hg.n_views;
            }
            const int nv = v1 - v0;
            ovis::MvArgs a{};
            int64_t nodes = 0;
            if (!cubes.empty())
                nodes = build_light_grid(cubes, a.grid, (int64_t)1 << 24);
            // ---- staging: planes | view meta | plain view ids | cube groups, one copy
            const size_t off_meta = sizeof(float) * 24 * (size_t)nv;
            const size_t off_plain = off_meta + sizeof(uint32_t) * (size_t)nv;
            const size_t off_groups = (off_plain + sizeof(uint32_t) * plain.size() + 15) & ~(size_t)15;
            const size_t off_pmeta = (off_groups + sizeof(ovis::MvGroup) * cubes.size() + 15) & ~(size_t)15;
            const size_t off_pplanes = (off_pmeta + sizeof(uint2) * plain.size() + 15) & ~(size_t)15;
            const size_t bytes = off_pplanes + sizeof(float) * 48 * plain.size();
            OVIS_TRY(ensure_pinned(w->h_mv_stage, w->cap_h_mv_stage, bytes + 16));
            {
                int64_t cap = (int64_t)w->cap_d_mv_stage;
                OVIS_TRY(ensure(w->d_mv_stage, cap, (int64_t)bytes + 16, w->stream));
                w->cap_d_mv_stage = (size_t)cap;
            }
            unsigned char *hs = static_cast<unsigned char *>(w->h_mv_stage);
            std::memcpy(hs, planes + 24 * (size_t)v0, sizeof(float) * 24 * (size_t)nv);
            uint32_t *meta = reinterpret_cast<uint32_t *>(hs + off_meta);
            for (int v = 0; v < nv; v++)
                meta[v] = ovis::plane_sign_codes(planes + 24 * (size_t)(v0 + v)) | ((require && require[v0 + v]) ? 0x80000000u : 0u);
            // views with equal P-vertex sign codes next to each other: the kernel tests them in runs (a view's bitmap row
            // is its position in this list, its totals row and CSR list are found through the view id, so any order works)
            // ... and among those, views with bit-identical side planes (the depth slices of one frustum) next to each other:
            // the kernel computes the four side distances of such a run once
            auto sides = [&](uint32_t x) { return planes + 24 * (size_t)(v0 + x); };          // planes 0-3: 16 floats
            std::stable_sort(plain.begin(), plain.end(), [&](uint32_t x, uint32_t y) {
                const uint32_t cx = meta[x] & 0x3ffffu, cy = meta[y] & 0x3ffffu;
                return cx != cy ? cx < cy : std::memcmp(sides(x), sides(y), 16 * sizeof(float)) < 0;
            });
            if (!plain.empty())
                std::memcpy(hs + off_plain, plain.data(), sizeof(uint32_t) * plain.size());
            if (!cubes.empty())
                std::memcpy(hs + off_groups, cubes.data(), sizeof(ovis::MvGroup) * cubes.size());
            {
                // what the dense kernel reads per plain view, in list order: codes + how many equal codes follow, and
                // the planes with every component doubled (operands of the packed multiplies as loaded)
                uint32_t *pm = reinterpret_cast<uint32_t *>(hs + off_pmeta);
                float *pp = reinterpret_cast<float *>(hs + off_pplanes);
                const size_t np = plain.size();
                for (size_t k = np; k-- > 0;) {
                    const uint32_t m = meta[plain[k]];
                    uint32_t run = 1;
                    if (k + 1 < np && (meta[plain[k + 1]] & 0x3ffffu) == (m & 0x3ffffu))
                        run = std::min<uint32_t>(4u, ((pm[2 * (k + 1)] >> 18) & 7u) + 1u);
                    // how many views from here on (same codes) have the side planes of this one, capped at 7
                    uint32_t shared = 1;
                    if (w->mv_shared_sides && k + 1 < np && (meta[plain[k + 1]] & 0x3ffffu) == (m & 0x3ffffu) &&
                        std::memcmp(sides(plain[k]), sides(plain[k + 1]), 16 * sizeof(float)) == 0)
                        shared = std::min<uint32_t>(7u, std::max<uint32_t>(1u, (pm[2 * (k + 1)] >> ovis::kPlainSharedShift) & 7u) + 1u);
                    pm[2 * k] = (m & 0x8003ffffu) | (run << 18) | (shared << ovis::kPlainSharedShift);
                    pm[2 * k + 1] = plain[k];
                    const float *src = planes + 24 * (size_t)(v0 + plain[k]);
                    for (int c = 0; c < 24; c++)
                        pp[48 * k + 2 * c] = pp[48 * k + 2 * c + 1] = src[c];
                }
            }
            trace.mark("groups, grid, staging");
            // ---- scratch
            OVIS_TRY(ensure(w->d_mv_totals, w->cap_mv_totals, round_up((int64_t)nv * row, 4), w->stream));
            OVIS_TRY(ensure(w->d_mv_masks, w->cap_mv_masks, std::max<int64_t>(1, (int64_t)plain.size() * num_words), w->stream));
            if (!cubes.empty()) {
                if (nodes > w->cap_mv_nodes) {
                    const int64_t fresh = std::max<int64_t>(nodes + nodes / 2, 1 << 16);
                    OVIS_TRY(grow(w->d_mv_nodes, 0, 2 * fresh, w->stream, false));
                    w->cap_mv_nodes = fresh;
                }
                OVIS_TRY(ensure(w->d_mv_items, w->cap_mv_items, std::max<int64_t>((int64_t)1 << 20, w->n_ent / 8), w->stream));
                OVIS_TRY(ensure(w->d_mv_segments, w->cap_mv_segments, num_blocks, w->stream));
                OVIS_TRY(ensure(w->d_mv_hits, w->cap_mv_hits, (int64_t)num_blocks * 2 * w->mv_hit_cap, w->stream));
                OVIS_TRY(ensure(w->d_mv_work, w->cap_mv_work, round_up(num_blocks, 4), w->stream));
            }
            const unsigned char *ds = static_cast<const unsigned char *>(w->d_mv_stage);
            a.bounds = w->d_world;
            a.stride = w->cap_ent;
            a.n = w->n_ent;
            a.num_groups128 = num_groups128;
            a.num_blocks = num_blocks;
            a.alive_mask = w->any_dead ? w->d_alive_bits : nullptr;
            a.caster_mask = any_require ? w->d_caster_bits : nullptr;
            a.planes = reinterpret_cast<const float4 *>(ds);
            a.view_meta = reinterpret_cast<const uint32_t *>(ds + off_meta);
            a.n_views = nv;
            a.totals = w->d_mv_totals;
            a.csr_offsets = w->d_csr_offsets + v0;
            a.out = w->d_csr_entities;
            a.capacity = w->cap_csr;
            a.id_base = w->id_base;
            a.state = reinterpret_cast<uint32_t *>(w->d_mv_state);
            a.plain_views = reinterpret_cast<const uint32_t *>(ds + off_plain);
            a.n_plain = (int)plain.size();
            a.plain_meta = reinterpret_cast<const uint2 *>(ds + off_pmeta);
            a.plain_planes = reinterpret_cast<const float4 *>(ds + off_pplanes);
            {
                // small worlds: split the views over more warps so that every SM has work
                const int64_t want_warps = (int64_t)w->num_sms * 48;
                int64_t vpw = num_groups128 >= want_warps ? (int64_t)plain.size()
                                                          : ((int64_t)plain.size() * num_groups128 + want_warps - 1) / want_warps;
                a.views_per_warp = (int)std::max<int64_t>(1, vpw);
            }
            a.masks = w->d_mv_masks;
            a.groups = reinterpret_cast<const ovis::MvGroup *>(ds + off_groups);
            a.n_groups = (int)cubes.size();
            a.bitmap = reinterpret_cast<unsigned long long *>(w->d_mv_state + kMvBitmapOffset);
            a.heads = w->d_mv_heads;
            a.nodes = w->d_mv_nodes;
            a.node_capacity = (uint32_t)std::min<int64_t>(w->cap_mv_nodes, 0xffffffffll);
            a.items = w->d_mv_items;
            a.item_capacity = (uint32_t)std::min<int64_t>(w->cap_mv_items, 0xffffffffll);
            a.segments = w->d_mv_segments;
            if (!cubes.empty()) {
                a.hit_records = w->d_mv_hits;
                a.hit_capacity = w->mv_hit_cap;
                a.block_work = w->d_mv_work;
            }
            a.host_offsets = hr + v0;
            a.host_state = reinterpret_cast<uint32_t *>(hr + n_views + 1);
            {
                // ---- reset: per-block totals, the round's counters + occupancy bitmap, the cells' list heads, the blocks'
                // record counts - one launch
                ovis::MvResetArgs r{};
                r.ptr[0] = w->d_mv_totals;
                r.quads[0] = round_up((int64_t)nv * row, 4) / 4;
                r.ptr[1] = w->d_mv_state + 16;
                r.quads[1] = (int64_t)(kMvStateBytes - 16) / 16;
                if (!cubes.empty()) {
                    r.ptr[2] = w->d_mv_heads;
                    r.quads[2] = round_up((int64_t)a.grid.dim[0] * a.grid.dim[1] * a.grid.dim[2], 4) / 4;
                    r.value[2] = 0xffffffffu;
                    r.ptr[3] = w->d_mv_work;
                    r.quads[3] = round_up(num_blocks, 4) / 4;
                }
                r.state = reinterpret_cast<uint32_t *>(w->d_mv_state);
                r.csr_offsets = w->d_csr_offsets;
                r.first_round = g0 == 0 ? 1 : 0;
                // the staging goes up inside this launch (a cudaMemcpyAsync of these few KB cost the call 5-20 us of host time
                // before the first kernel could be queued)
                r.copy_src = reinterpret_cast<const uint4 *>(hs);
                r.copy_dst = reinterpret_cast<uint4 *>(w->d_mv_stage);
                r.copy_quads = (int64_t)((bytes + 15) / 16);
                ovis::launch_mv_reset(r, w->num_sms, w->stream);
            }
            // ---- kernels
            ovis::launch_mv_bin_lights(a, w->stream);
            ovis::launch_mv_plain_test(a, w->num_sms, w->stream);
            ovis::launch_mv_sparse(a, w->num_sms, w->stream);
            ovis::launch_mv_scan(a, w->stream);
            ovis::launch_mv_expand_masks(a, w->stream);
            ovis::launch_mv_expand_items(a, w->stream);
            w->launches += 2 + (plain.empty() ? 0 : 2) + (cubes.empty() ? 0 : 4);
            OVIS_CUDA(cudaGetLastError());
            trace.mark("kernels queued");
            if (g1 < groups.size())
                OVIS_CUDA(cudaStreamSynchronize(w->stream));     // the pinned staging buffer is refilled by the next round
            g0 = g1;
        }
        // ---- read-back: the scan kernel of every round has stored its offsets and the state words in page-locked host
        // memory (hr), nothing is copied
        w->mv_spec_copied = 0;
        if (w->mv_spec_out && w->mv_spec_cap > 0) {
            const int64_t guess = std::min(std::min(w->mv_spec_cap, w->cap_csr), w->mv_last_total + w->mv_last_total / 4 + 1024);
            OVIS_CUDA(cudaMemcpyAsync(w->mv_spec_out, w->d_csr_entities, sizeof(uint32_t) * (size_t)guess, cudaMemcpyDeviceToHost, w->stream));
            w->mv_spec_copied = guess;
        }
        OVIS_CUDA(cudaStreamSynchronize(w->stream));
        trace.mark("synchronised");
        const uint32_t *state = reinterpret_cast<const uint32_t *>(hr + n_views + 1);
        const uint32_t overflow = state[0];
        if (overflow & ovis::kMvOverflowPairs) {         // some block was too dense for the sparse pass: everything densely
            force_dense = true;
            continue;
        }
        if (overflow & ovis::kMvOverflowHits) {          // more entities near lights in some block than its list holds
            const int64_t want = (int64_t)state[4] + state[4] / 4;
            w->mv_hit_cap = (int)std::min<int64_t>(ovis::kMaskBlockEntities, std::max<int64_t>(2 * (int64_t)w->mv_hit_cap, (want + 255) / 256 * 256));
            continue;
        }
        if (overflow & (ovis::kMvOverflowItems | ovis::kMvOverflowNodes)) {
            if (overflow & ovis::kMvOverflowItems)
                OVIS_TRY(ensure(w->d_mv_items, w->cap_mv_items, std::max<int64_t>(w->cap_mv_items * 2, (int64_t)state[3] + 1024), w->stream));
            if (overflow & ovis::kMvOverflowNodes)
                return fail(OVIS_ERR_CUDA, "light grid: node list overflow (internal error)");
            continue;
        }
        if (hr[n_views] > w->cap_csr) {                  // counts are exact even when writes were clipped
            OVIS_TRY(grow(w->d_csr_entities, 0, round_up(hr[n_views], 1 << 20), w->stream, false));
            w->cap_csr = round_up(hr[n_views], 1 << 20);
            continue;
        }
        for (int32_t v = 0; v <= n_views; v++)
            host_offsets[(size_t)v] = hr[v];
        w->mv_last_total = hr[n_views];
        break;
    }
    if (host_offsets_out)
        for (int32_t v = 0; v <= n_views; v++)
            host_offsets_out[v] = host_offsets[(size_t)v];
    return OVIS_OK;
}

// The CSR entities of the last cull_views_grouped into a host buffer: whatever the speculative copy did not cover.
int finish_host_entities(OvisWorld *w, uint32_t *out_entities, int64_t total) {
    const int64_t have = w->mv_spec_out == out_entities ? std::min(w->mv_spec_copied, total) : 0;
    w->mv_spec_out = nullptr;
    if (total > have) {
        OVIS_CUDA(cudaMemcpyAsync(out_entities + have, w->d_csr_entities + have, sizeof(uint32_t) * (size_t)(total - have),
                                  cudaMemcpyDeviceToHost, w->stream));
        OVIS_CUDA(cudaStreamSynchronize(w->stream));
    }
    return OVIS_OK;
}

// Shadow views (render_light.cc:258-332) of the given lights, culled for shadow casters (deferred_render_pipeline.cc:
// 270-283) with K4; only spot and point lights have any (allocateShadowMap is fatal for the rest, :291-303).
int shadow_views_device(OvisWorld *w, const std::vector<uint32_t> &light_ids, int32_t shadow_map_resolution,
                        std::vector<int32_t> &list_light, std::vector<int32_t> &list_view, std::vector<long long> &offsets,
                        uint32_t require_mask = OVIS_CASTS_SHADOW) {
    std::vector<float> planes;
    std::vector<HostGroup> groups;
    list_light.clear();
    list_view.clear();
    w->light_views.resize(w->lights.size());
    planes.reserve(light_ids.size() * 6 * 24);
    for (uint32_t id : light_ids) {
        const OvisLight &l = w->lights[id];
        OvisWorld::LightViews &lv = w->light_views[id];
        if (lv.resolution != shadow_map_resolution) {
            OvisView views[6];
            lv.count = ovis::hm::shadowViews(l, shadow_map_resolution, views);
            for (int i = 0; i < lv.count; i++)
                ovis::hm::viewPlanes(views[i], lv.planes + 24 * i);
            lv.resolution = shadow_map_resolution;
        }
        if (lv.count == 0)
            continue;
        // A point light's six face frusta (fov 90, far = range, apex at the light) only accept boxes that overlap
        // the cube of half-side range + box extent about the light; a spot light's single frustum gets no pre-filter.
        // That holds while far > near: a range at or below the shadow views' near plane (0.1) turns the far plane
        // round, the frusta open outwards beyond the range (the reference then lists entities any distance away), and
        // the light is tested without a cube. A NaN range fails the comparison too.
        bool bounded = l.type == OVIS_LIGHT_POINT && l.range > ovis::hm::kShadowViewNear;
        // How far the faces really reach is what the reference's own far planes say, not the range: they come out of
        // far / (near - far) + 1 in float, which for range / near ~ 1e6 keeps three or four bits (a range of 90 000 puts the
        // far plane anywhere between 84 000 and 100 000), and for larger ratios is 0: no far plane at all. The cube is sized
        // by the largest distance from the light to one of its six far planes; no cube if one of them is degenerate.
        float reach = l.range;
        for (int i = 0; bounded && i < lv.count; i++) {
            const float *fp = lv.planes + 24 * i + 20;
            const float far_distance = (fp[0] * l.position[0] + fp[1] * l.position[1] + fp[2] * l.position[2]) - fp[3];
            if (!(far_distance > 0.0f) || !(far_distance <= 8.0f * l.range))
                bounded = false;                                   // NaN, behind the light, or nowhere near the range
            else
                reach = std::max(reach, far_distance);
        }
        groups.push_back(make_group(bounded ? l.position : nullptr, reach, (int)list_light.size(), lv.count));
        planes.insert(planes.end(), lv.planes, lv.planes + 24 * lv.count);
        for (int i = 0; i < lv.count; i++) {
            list_light.push_back((int32_t)id);
            list_view.push_back(i);
        }
    }
    int32_t n_lists = (int32_t)list_light.size();
    std::vector<uint32_t> require((size_t)n_lists, require_mask);
    return cull_views_grouped(w, n_lists, planes.data(), require.data(), groups, nullptr, offsets);
}

// Views without a shared pre-filter: one group per view, infinite cube.
int cull_views_device(OvisWorld *w, int32_t n_views, const float *planes, const uint32_t *require, int64_t *host_offsets_out,
                      std::vector<long long> &host_offsets) {
    std::vector<HostGroup> groups((size_t)n_views);
    for (int32_t v = 0; v < n_views; v++) {
        groups[(size_t)v] = make_group(nullptr, 0.0f, v, 1);
    }
    return cull_views_grouped(w, n_views, planes, require, groups, host_offsets_out, host_offsets);
}

}  // namespace

namespace ovis {
// Lets the library's other .cu files report through ovis_last_error().
int report_error(int code, const char *message) { return fail(code, "%s", message); }
}  // namespace ovis

// ================================================================================================
// C ABI
// ================================================================================================

extern "C" {

const char *ovis_last_error(void) { return g_error.c_str(); }

const char *ovis_version(void) { return "orion_vis 0.1 sm_100a fmad=off"; }

int ovis_world_create(int device, OvisWorld **out_world) {
    if (!out_world)
        return fail(OVIS_ERR_INVALID, "out_world is null");
    *out_world = nullptr;
    int count = 0;
    cudaError_t err = cudaGetDeviceCount(&count);
    if (err != cudaSuccess || count == 0)
        return fail(OVIS_ERR_CUDA, "no CUDA device available (%s); liborion_vis has no CPU path",
                    err == cudaSuccess ? "device count is 0" : cudaGetErrorString(err));
    if (device < 0 || device >= count)
        return fail(OVIS_ERR_INVALID, "device %d out of range (0..%d)", device, count - 1);
    OVIS_CUDA(cudaSetDevice(device));
    OvisWorld *w = new (std::nothrow) OvisWorld;
    if (!w)
        return fail(OVIS_ERR_NOMEM, "out of host memory");
    w->device = device;
    cudaDeviceProp prop;
    OVIS_CUDA(cudaGetDeviceProperties(&prop, device));
    w->num_sms = prop.multiProcessorCount;
    if (const char *v = getenv("OVIS_CULL_VARIANT")) {   // tuning knob, not part of the ABI
        int k = atoi(v);
        if (k >= 0 && k < ovis::cull_num_variants())
            w->forced_variant = k;
    }
    if (const char *v = getenv("OVIS_CULL_FUSED"))
        w->fused_small = atoi(v) != 0;
    for (int v = 0; v < ovis::cull_num_variants() && v < 16; v++) {
        w->blocks_per_sm[v][0] = std::max(1, ovis::cull_test_max_blocks_per_sm(false, v));
        w->blocks_per_sm[v][1] = std::max(1, ovis::cull_test_max_blocks_per_sm(true, v));
        w->fused_blocks_per_sm[v][0] = ovis::cull_fused_max_blocks_per_sm(false, v);
        w->fused_blocks_per_sm[v][1] = ovis::cull_fused_max_blocks_per_sm(true, v);
    }
    OVIS_CUDA(cudaMalloc(&w->d_exchange, 2 * sizeof(unsigned int)));
    OVIS_CUDA(cudaMemset(w->d_exchange, 0, 2 * sizeof(unsigned int)));
    OVIS_CUDA(cudaMalloc(&w->d_counts, 2 * sizeof(uint32_t)));
    OVIS_CUDA(cudaMemset(w->d_counts, 0, 2 * sizeof(uint32_t)));
    OVIS_CUDA(cudaMallocHost(&w->h_counts, 4 * sizeof(uint32_t)));     // entity count, light count, frame flag, unused
    w->h_counts[0] = w->h_counts[1] = w->h_counts[2] = w->h_counts[3] = 0;
    if (cudaHostGetDevicePointer((void **)&w->h_counts_dev, w->h_counts, 0) != cudaSuccess) {
        cudaGetLastError();
        w->h_counts_dev = nullptr;
    }
    if (const char *v = getenv("OVIS_EXCHANGE_BLOCK_FENCE"))
        w->block_fence_sys = std::strcmp(v, "sys") == 0;
    if (const char *v = getenv("OVIS_DIRECT_HOST"))
        w->direct_host = atoi(v) != 0;
    if (const char *v = getenv("OVIS_FUSED_BELOW"))
        w->fused_below = atoll(v);
    if (const char *v = getenv("OVIS_MV_SHARED_SIDES"))
        w->mv_shared_sides = atoi(v) != 0;
    if (const char *v = getenv("OVIS_HOST_FLAG"))
        w->host_flag = atoi(v) != 0;
    if (const char *v = getenv("OVIS_MV_DENSE")) {
        w->mv_force_dense = atoi(v) != 0;
        w->mv_force_sparse = atoi(v) == 0;
    }
    if (const char *v = getenv("OVIS_MV_ROUND_VIEWS"))          // (testing) views per round, as if the scratch budget were that small
        w->mv_round_views = std::max(atoi(v), ovis::kMvMaxGroupViews);
    *out_world = w;
    return OVIS_OK;
}

int ovis_world_destroy(OvisWorld *w) {
    if (!w)
        return OVIS_OK;
    cudaSetDevice(w->device);
    cudaStreamSynchronize(w->stream);
    void *bufs[] = { w->d_local, w->d_world, w->d_alive_bits, w->d_caster_bits, w->d_parent, w->d_local_trs, w->d_world_trs,
                     w->d_node_entity, w->d_dirty, w->d_light_uniforms, w->d_light_types, w->d_entity_node, w->d_pass_counts, w->d_draw_totals, w->d_draw_counts,
                     w->d_extra_entity, w->d_extra_node, w->d_level_lists, w->d_lights, w->d_light_out, w->d_light_bits,
                     w->d_survivors, w->d_block_totals, w->d_exchange, w->d_mv_stage, w->d_mv_state, w->d_mv_heads, w->d_mv_totals,
                     w->d_mv_masks, w->d_mv_hits, w->d_mv_work, w->d_mv_nodes, w->d_mv_items, w->d_mv_segments, w->d_result, w->d_counts, w->d_csr_entities, w->d_csr_offsets,
                     w->d_stage[0], w->d_stage[1] };
    for (void *b : bufs)
        if (b)
            cudaFree(b);
    if (w->h_counts)
        cudaFreeHost(w->h_counts);
    if (w->h_mv_stage)
        cudaFreeHost(w->h_mv_stage);
    if (w->h_mv_result)
        cudaFreeHost(w->h_mv_result);
    for (cudaEvent_t e : w->timing_events)
        cudaEventDestroy(e);
    delete w;
    return OVIS_OK;
}

int ovis_world_reserve(OvisWorld *w, int64_t n_entities, int64_t n_nodes) {
    if (!w)
        return fail(OVIS_ERR_INVALID, "world is null");
    OVIS_TRY(use_device(w));
    if (n_entities > 0)
        OVIS_TRY(reserve_entities(w, n_entities));
    if (n_nodes > 0)
        OVIS_TRY(reserve_nodes(w, n_nodes));
    return OVIS_OK;
}

int ovis_world_set_stream(OvisWorld *w, void *cuda_stream) {
    if (!w)
        return fail(OVIS_ERR_INVALID, "world is null");
    OVIS_TRY(use_device(w));
    OVIS_CUDA(cudaStreamSynchronize(w->stream));
    w->stream = (cudaStream_t)cuda_stream;
    w->launches_after_gather = -1;                 // nothing of ours is in front on the new stream
    return OVIS_OK;
}

int ovis_world_get_stream(OvisWorld *w, void **out_cuda_stream, int32_t *out_device) {
    if (!w)
        return fail(OVIS_ERR_INVALID, "world is null");
    if (out_cuda_stream) *out_cuda_stream = (void *)w->stream;
    if (out_device) *out_device = w->device;
    return OVIS_OK;
}

int ovis_world_get_id_base(OvisWorld *w, uint32_t *out_id_base) {
    if (!w || !out_id_base)
        return fail(OVIS_ERR_INVALID, "ovis_world_get_id_base: null argument");
    *out_id_base = w->id_base;
    return OVIS_OK;
}

int ovis_world_set_id_base(OvisWorld *w, uint32_t id_base) {
    if (!w)
        return fail(OVIS_ERR_INVALID, "world is null");
    w->id_base = id_base;
    return OVIS_OK;
}

int ovis_world_counts(OvisWorld *w, int64_t *n_entities, int64_t *n_nodes, int32_t *n_lights) {
    if (!w)
        return fail(OVIS_ERR_INVALID, "world is null");
    if (n_entities) *n_entities = w->n_ent;
    if (n_nodes) *n_nodes = w->n_nodes;
    if (n_lights) *n_lights = (int32_t)w->lights.size();
    return OVIS_OK;
}

int ovis_world_set_timing(OvisWorld *w, int enabled) {
    if (!w)
        return fail(OVIS_ERR_INVALID, "world is null");
    w->timing = enabled != 0;
    w->timing_used = 0;
    return OVIS_OK;
}

int ovis_world_read_timing(OvisWorld *w, int32_t capacity, float *test_ms, float *gather_ms, int32_t *out_count) {
    if (!w || !out_count || capacity < 0)
        return fail(OVIS_ERR_INVALID, "ovis_world_read_timing: bad arguments");
    OVIS_TRY(use_device(w));
    OVIS_CUDA(cudaStreamSynchronize(w->stream));
    int32_t n = (int32_t)(w->timing_used / 3);
    *out_count = n;
    for (int32_t i = 0; i < n && i < capacity; i++) {
        cudaEvent_t *ev = &w->timing_events[3 * (size_t)i];
        if (test_ms)
            OVIS_CUDA(cudaEventElapsedTime(&test_ms[i], ev[0], ev[1]));
        if (gather_ms)
            OVIS_CUDA(cudaEventElapsedTime(&gather_ms[i], ev[1], ev[2]));
    }
    w->timing_used = 0;
    return OVIS_OK;
}

int ovis_world_launch_count(OvisWorld *w, int64_t *out_launches) {
    if (!w || !out_launches)
        return fail(OVIS_ERR_INVALID, "null argument");
    *out_launches = w->launches;
    return OVIS_OK;
}

// ---- entities ----------------------------------------------------------------------------------

int ovis_entity_add(OvisWorld *w, int64_t n, const float *local_aabb, const uint32_t *flags, uint32_t *out_first_id) {
    if (!w || n < 0)
        return fail(OVIS_ERR_INVALID, "ovis_entity_add: bad arguments");
    OVIS_TRY(use_device(w));
    int64_t first = w->n_ent;
    if (first + n > (int64_t)0xfffffffe)
        return fail(OVIS_ERR_INVALID, "too many entities");
    OVIS_TRY(reserve_entities(w, first + n));
    for (int64_t s = 0; s < n && local_aabb; s += kUploadChunk) {
        int64_t c = std::min(kUploadChunk, n - s);
        void *d = nullptr;
        OVIS_TRY(stage(w, 0, local_aabb + 6 * s, sizeof(float) * 6 * (size_t)c, &d));
        ovis::launch_scatter_aos(6, (const float *)d, nullptr, (uint32_t)(first + s), c, w->d_local, w->cap_ent, w->stream);
        w->launches++;
    }
    // BoundingBox() is all zeros (bounding_box.h:33-36): the SoA blocks are zero-filled on allocation.
    ovis::launch_identity_bounds(nullptr, (uint32_t)first, n, w->d_local, w->d_world, w->cap_ent, w->stream);
    w->launches++;
    OVIS_CUDA(cudaGetLastError());
    w->h_entity_node.resize((size_t)(first + n), kNone);
    for (int64_t i = 0; i < n; i++) {
        int64_t id = first + i;
        w->h_alive[(size_t)(id >> 5)] |= 1u << (id & 31);
        if (flags && (flags[i] & OVIS_CASTS_SHADOW))
            w->h_caster_raw[(size_t)(id >> 5)] |= 1u << (id & 31);
    }
    w->n_ent = first + n;
    w->n_live += n;
    if (n > 0)
        mark_masks(w, first, first + n - 1);
    if (out_first_id)
        *out_first_id = (uint32_t)first;
    return OVIS_OK;
}

int ovis_entity_remove(OvisWorld *w, int64_t n, const uint32_t *ids) {
    if (!w)
        return fail(OVIS_ERR_INVALID, "world is null");
    OVIS_TRY(check_entity_ids(w, n, ids));
    for (int64_t i = 0; i < n; i++) {
        uint32_t id = ids[i];
        if (!((w->h_alive[id >> 5] >> (id & 31)) & 1u))
            continue;                                   // listed twice
        w->h_alive[id >> 5] &= ~(1u << (id & 31));
        w->n_live--;
        mark_masks(w, id, id);
    }
    if (n > 0)
        w->any_dead = true;
    return OVIS_OK;
}

int ovis_entity_set_local_aabbs(OvisWorld *w, int64_t n, const uint32_t *ids, const float *local_aabb) {
    if (!w || (n > 0 && !local_aabb))
        return fail(OVIS_ERR_INVALID, "ovis_entity_set_local_aabbs: null input");
    OVIS_TRY(use_device(w));
    OVIS_TRY(check_entity_ids(w, n, ids));
    for (int64_t s = 0; s < n; s += kUploadChunk) {
        int64_t c = std::min(kUploadChunk, n - s);
        void *d = nullptr, *di = nullptr;
        OVIS_TRY(stage(w, 0, local_aabb + 6 * s, sizeof(float) * 6 * (size_t)c, &d));
        OVIS_TRY(stage(w, 1, ids + s, sizeof(uint32_t) * (size_t)c, &di));
        ovis::launch_scatter_aos(6, (const float *)d, (const uint32_t *)di, 0, c, w->d_local, w->cap_ent, w->stream);
        // entities that follow no node keep the identity transform: refresh them here; the rest are
        // refreshed by the transform pass below
        ovis::launch_identity_bounds((const uint32_t *)di, 0, c, w->d_local, w->d_world, w->cap_ent, w->stream);
        w->launches += 2;
    }
    OVIS_CUDA(cudaGetLastError());
    std::vector<uint32_t> nodes;
    for (int64_t i = 0; i < n; i++)
        if (w->h_entity_node[ids[i]] != kNone)
            nodes.push_back(w->h_entity_node[ids[i]]);
    if (!nodes.empty()) {
        w->transforms_dirty = true;
        if (!w->all_dirty) {
            void *dn = nullptr;
            OVIS_TRY(stage(w, 1, nodes.data(), sizeof(uint32_t) * nodes.size(), &dn));
            ovis::launch_mark_dirty((const uint32_t *)dn, (int64_t)nodes.size(), w->d_dirty, w->stream);
            w->launches++;
            w->marks_present = true;
            OVIS_CUDA(cudaStreamSynchronize(w->stream));   // `nodes` is a local
        }
    }
    return OVIS_OK;
}

int ovis_entity_set_flags(OvisWorld *w, int64_t n, const uint32_t *ids, const uint32_t *flags) {
    if (!w || (n > 0 && !flags))
        return fail(OVIS_ERR_INVALID, "ovis_entity_set_flags: null input");
    OVIS_TRY(check_entity_ids(w, n, ids));
    for (int64_t i = 0; i < n; i++) {
        uint32_t id = ids[i], bit = 1u << (id & 31);
        if (flags[i] & OVIS_CASTS_SHADOW)
            w->h_caster_raw[id >> 5] |= bit;
        else
            w->h_caster_raw[id >> 5] &= ~bit;
        mark_masks(w, id, id);
    }
    return OVIS_OK;
}

int ovis_entity_set_transforms(OvisWorld *w, int64_t n, const uint32_t *ids, const float *world_trs) {
    if (!w || (n > 0 && !world_trs))
        return fail(OVIS_ERR_INVALID, "ovis_entity_set_transforms: null input");
    OVIS_TRY(use_device(w));
    OVIS_TRY(check_entity_ids(w, n, ids));
    // An entity driven directly (RenderEntity::setTransform) follows a private root-level node whose
    // local TRS is the given world TRS; the same kernels then serve both ways of moving an entity.
    std::vector<uint32_t> fresh_entities, node_ids((size_t)n);
    std::vector<float> fresh_trs;
    for (int64_t i = 0; i < n; i++) {
        uint32_t node = w->h_entity_node[ids[i]];
        if (node == kNone) {
            bool dup = false;                           // same id twice in one call: the last TRS wins
            for (size_t k = 0; k < fresh_entities.size() && !dup; k++)
                dup = fresh_entities[k] == ids[i];
            if (!dup)
                fresh_entities.push_back(ids[i]);
        } else if (w->h_parent[node] != OVIS_NO_PARENT || w->h_node_entity[node] != ids[i]) {
            return fail(OVIS_ERR_INVALID, "entity %u follows tree node %u; move the node instead", ids[i], node);
        }
    }
    if (!fresh_entities.empty()) {
        size_t m = fresh_entities.size();
        std::vector<int32_t> parents(m, OVIS_NO_PARENT);
        fresh_trs.assign(10 * m, 0.0f);
        for (size_t k = 0; k < m; k++) {
            fresh_trs[10 * k + 3] = 1.0f;
            fresh_trs[10 * k + 7] = fresh_trs[10 * k + 8] = fresh_trs[10 * k + 9] = 1.0f;
        }
        uint32_t first_node = 0;
        OVIS_TRY(add_nodes(w, (int64_t)m, parents.data(), fresh_trs.data(), &first_node));
        OVIS_CUDA(cudaStreamSynchronize(w->stream));
        std::vector<uint32_t> nodes(m);
        for (size_t k = 0; k < m; k++)
            nodes[k] = first_node + (uint32_t)k;
        OVIS_TRY(bind_entities(w, (int64_t)m, nodes.data(), fresh_entities.data()));
    }
    for (int64_t i = 0; i < n; i++)
        node_ids[(size_t)i] = w->h_entity_node[ids[i]];
    return ovis_node_set_local_trs(w, n, node_ids.data(), world_trs);
}

int ovis_entity_set_world_aabbs(OvisWorld *w, uint32_t first_id, int64_t n, const float *world_aabb) {
    if (!w || n < 0 || (n > 0 && !world_aabb))
        return fail(OVIS_ERR_INVALID, "ovis_entity_set_world_aabbs: bad arguments");
    if ((int64_t)first_id + n > w->n_ent)
        return fail(OVIS_ERR_INVALID, "entity range %u..%lld exceeds %lld entities", first_id, (long long)first_id + n - 1, (long long)w->n_ent);
    OVIS_TRY(use_device(w));
    for (int64_t s = 0; s < n; s += kUploadChunk) {
        int64_t c = std::min(kUploadChunk, n - s);
        void *d = nullptr;
        OVIS_TRY(stage(w, 0, world_aabb + 6 * s, sizeof(float) * 6 * (size_t)c, &d));
        ovis::launch_scatter_aos(6, (const float *)d, nullptr, (uint32_t)(first_id + s), c, w->d_world, w->cap_ent, w->stream);
        w->launches++;
    }
    OVIS_CUDA(cudaGetLastError());
    return OVIS_OK;
}

int ovis_entity_set_world_aabbs_by_id(OvisWorld *w, int64_t n, const uint32_t *ids, const float *world_aabb) {
    if (!w || n < 0 || (n > 0 && !world_aabb))
        return fail(OVIS_ERR_INVALID, "ovis_entity_set_world_aabbs_by_id: bad arguments");
    OVIS_TRY(use_device(w));
    OVIS_TRY(check_entity_ids(w, n, ids));
    for (int64_t s = 0; s < n; s += kUploadChunk) {
        int64_t c = std::min(kUploadChunk, n - s);
        void *d = nullptr, *di = nullptr;
        OVIS_TRY(stage(w, 0, world_aabb + 6 * s, sizeof(float) * 6 * (size_t)c, &d));
        OVIS_TRY(stage(w, 1, ids + s, sizeof(uint32_t) * (size_t)c, &di));
        ovis::launch_scatter_aos(6, (const float *)d, (const uint32_t *)di, 0, c, w->d_world, w->cap_ent, w->stream);
        w->launches++;
    }
    OVIS_CUDA(cudaGetLastError());
    return OVIS_OK;
}

// ---- transform tree ----------------------------------------------------------------------------

int ovis_node_add(OvisWorld *w, int64_t n, const int32_t *parent, const float *local_trs, uint32_t *out_first_id) {
    if (!w)
        return fail(OVIS_ERR_INVALID, "world is null");
    OVIS_TRY(use_device(w));
    return add_nodes(w, n, parent, local_trs, out_first_id);
}

int ovis_node_set_local_trs(OvisWorld *w, int64_t n, const uint32_t *ids, const float *local_trs) {
    if (!w || n < 0 || (n > 0 && (!ids || !local_trs)))
        return fail(OVIS_ERR_INVALID, "ovis_node_set_local_trs: null input");
    OVIS_TRY(use_device(w));
    for (int64_t i = 0; i < n; i++)
        if ((int64_t)ids[i] >= w->n_nodes)
            return fail(OVIS_ERR_INVALID, "node id %u out of range", ids[i]);
    for (int64_t s = 0; s < n; s += kUploadChunk) {
        int64_t c = std::min(kUploadChunk, n - s);
        void *d = nullptr, *di = nullptr;
        OVIS_TRY(stage(w, 0, local_trs + 10 * s, sizeof(float) * 10 * (size_t)c, &d));
        OVIS_TRY(stage(w, 1, ids + s, sizeof(uint32_t) * (size_t)c, &di));
        ovis::launch_scatter_aos(10, (const float *)d, (const uint32_t *)di, 0, c, w->d_local_trs, w->cap_nodes, w->stream);
        w->launches++;
        if (!w->all_dirty) {
            ovis::launch_mark_dirty((const uint32_t *)di, c, w->d_dirty, w->stream);
            w->launches++;
            w->marks_present = true;
        }
    }
    OVIS_CUDA(cudaGetLastError());
    if (n > 0)
        w->transforms_dirty = true;
    if (!w->all_dirty) {
        // estimate how much of the tree sits below the moved nodes; when it is most of it, recompute everything
        // and skip the per-node flag traffic
        int64_t below = 0;
        for (int64_t i = 0; i < n && below < w->n_nodes; i++)
            below += w->h_subtree[ids[i]];
        if (below >= w->n_nodes / 3)
            w->all_dirty = true;
    }
    return OVIS_OK;
}

int ovis_node_bind_entities(OvisWorld *w, int64_t n, const uint32_t *node_ids, const uint32_t *entity_ids) {
    if (!w || n < 0 || (n > 0 && (!node_ids || !entity_ids)))
        return fail(OVIS_ERR_INVALID, "ovis_node_bind_entities: null input");
    OVIS_TRY(use_device(w));
    OVIS_TRY(check_entity_ids(w, n, entity_ids));
    for (int64_t i = 0; i < n; i++)
        if ((int64_t)node_ids[i] >= w->n_nodes)
            return fail(OVIS_ERR_INVALID, "node id %u out of range", node_ids[i]);
    return bind_entities(w, n, node_ids, entity_ids);
}

int ovis_update_transforms(OvisWorld *w) {
    if (!w)
        return fail(OVIS_ERR_INVALID, "world is null");
    OVIS_TRY(use_device(w));
    return flush_transforms(w);
}

int ovis_node_set_parent(OvisWorld *w, uint32_t node, int32_t new_parent) {
    if (!w || (int64_t)node >= w->n_nodes || (new_parent != OVIS_NO_PARENT && (new_parent < 0 || (int64_t)new_parent >= w->n_nodes)))
        return fail(OVIS_ERR_INVALID, "ovis_node_set_parent: unknown node");
    auto dead = [&](int64_t i) { return !w->h_node_dead.empty() && w->h_node_dead[(size_t)i] != 0; };
    if (dead(node) || (new_parent >= 0 && dead(new_parent)))
        return fail(OVIS_ERR_INVALID, "ovis_node_set_parent: the node was removed");
    for (int32_t a = new_parent; a >= 0; a = w->h_parent[(size_t)a])     // Entity::addChild cannot create a cycle either
        if ((uint32_t)a == node)
            return fail(OVIS_ERR_INVALID, "ovis_node_set_parent: node %u is an ancestor of node %d", node, new_parent);
    OVIS_TRY(use_device(w));
    // the subtree-size estimates (they only steer the all-or-dirty choice of an update): move this subtree's share
    const uint32_t moved = w->h_subtree[node];
    int hops = 0;
    for (int32_t a = w->h_parent[node]; a >= 0 && hops < 64; a = w->h_parent[(size_t)a], hops++)
        w->h_subtree[(size_t)a] -= std::min(w->h_subtree[(size_t)a] - 1u, moved);
    hops = 0;
    for (int32_t a = new_parent; a >= 0 && hops < 64; a = w->h_parent[(size_t)a], hops++)
        w->h_subtree[(size_t)a] += moved;
    w->h_parent[node] = new_parent;
    OVIS_CUDA(cudaMemcpyAsync(w->d_parent + node, &w->h_parent[node], sizeof(int32_t), cudaMemcpyHostToDevice, w->stream));
    OVIS_CUDA(cudaStreamSynchronize(w->stream));
    // levels of every node again: after re-parenting a parent may have a larger id than its child, so walk up with a memo
    std::vector<int32_t> &level = w->h_level;
    std::fill(level.begin(), level.end(), -1);
    std::vector<int32_t> path;
    for (int64_t i = 0; i < w->n_nodes; i++) {
        int32_t a = (int32_t)i;
        path.clear();
        while (a >= 0 && level[(size_t)a] < 0) {
            path.push_back(a);
            a = w->h_parent[(size_t)a];
        }
        int32_t l = a >= 0 ? level[(size_t)a] : -1;
        for (size_t k = path.size(); k-- > 0;)
            level[(size_t)path[k]] = ++l;
    }
    w->levels_dirty = true;
    w->transforms_dirty = true;
    w->all_dirty = true;
    return OVIS_OK;
}

int ovis_node_remove(OvisWorld *w, uint32_t node, int64_t *out_removed_nodes, int64_t *out_removed_entities) {
    if (!w || (int64_t)node >= w->n_nodes)
        return fail(OVIS_ERR_INVALID, "ovis_node_remove: unknown node");
    if (w->h_node_dead.empty())
        w->h_node_dead.assign((size_t)w->n_nodes, 0);
    w->h_node_dead.resize((size_t)w->n_nodes, 0);
    if (w->h_node_dead[node])
        return fail(OVIS_ERR_INVALID, "ovis_node_remove: node %u was removed already", node);
    // Entity::destroy (engine/src/entity.cc:58-90) takes the whole subtree and every component with it
    w->h_node_dead[node] = 1;
    int64_t removed = 1;
    for (bool again = true; again;) {                       // to a fixed point: two passes when parents precede their children
        again = false;
        for (int64_t i = 0; i < w->n_nodes; i++) {
            const int32_t p = w->h_parent[(size_t)i];
            if (!w->h_node_dead[(size_t)i] && p >= 0 && w->h_node_dead[(size_t)p]) {
                w->h_node_dead[(size_t)i] = 1;
                removed++;
                again = true;
            }
        }
    }
    std::vector<uint32_t> gone;
    for (int64_t e = 0; e < w->n_ent; e++) {
        const uint32_t n = w->h_entity_node[(size_t)e];
        if (n != kNone && w->h_node_dead[n] && ((w->h_alive[(size_t)e >> 5] >> (e & 31)) & 1u))
            gone.push_back((uint32_t)e);
    }
    if (!gone.empty())
        OVIS_TRY(ovis_entity_remove(w, (int64_t)gone.size(), gone.data()));
    for (size_t l = 0; l < w->light_node.size(); l++)
        if (w->light_node[l] >= 0 && w->h_node_dead[(size_t)w->light_node[l]]) {
            w->light_node[l] = OVIS_NO_PARENT;                // the light stays where the node left it
            w->n_bound_lights--;
        }
    w->levels_dirty = true;
    w->transforms_dirty = true;
    if (out_removed_nodes) *out_removed_nodes = removed;
    if (out_removed_entities) *out_removed_entities = (int64_t)gone.size();
    return OVIS_OK;
}

int ovis_world_live_counts(OvisWorld *w, int64_t *out_live_entities, int64_t *out_entity_slots) {
    if (!w)
        return fail(OVIS_ERR_INVALID, "world is null");
    if (out_live_entities) *out_live_entities = w->n_live;
    if (out_entity_slots) *out_entity_slots = w->n_ent;
    return OVIS_OK;
}

int ovis_world_compact(OvisWorld *w, int64_t cap_remap, uint32_t *out_remap, int64_t *out_old_slots, int64_t *out_new_slots) {
    if (!w || cap_remap < 0 || (cap_remap > 0 && !out_remap))
        return fail(OVIS_ERR_INVALID, "ovis_world_compact: bad arguments");
    const int64_t n_old = w->n_ent;
    if (out_old_slots) *out_old_slots = n_old;
    if (out_remap && cap_remap < n_old)
        return fail(OVIS_ERR_CAPACITY, "ovis_world_compact: the remap needs %lld entries", (long long)n_old);
    OVIS_TRY(use_device(w));
    OVIS_TRY(flush_transforms(w));
    std::vector<uint32_t> keep;                              // old ids of the live entities, ascending: order is preserved
    keep.reserve((size_t)w->n_live);
    for (int64_t e = 0; e < n_old; e++)
        if ((w->h_alive[(size_t)e >> 5] >> (e & 31)) & 1u)
            keep.push_back((uint32_t)e);
    const int64_t n_new = (int64_t)keep.size();
    if (out_new_slots) *out_new_slots = n_new;
    if (out_remap) {
        std::fill(out_remap, out_remap + n_old, kNone);
        for (int64_t i = 0; i < n_new; i++)
            out_remap[keep[(size_t)i]] = (uint32_t)i;
    }
    if (n_new == n_old) {
        w->any_dead = false;
        return OVIS_OK;
    }
    // ---- device: fresh arrays of the same capacity, filled by stable gathers, swapped in together
    float *local = nullptr, *world = nullptr;
    uint32_t *passes = nullptr, *ent_node = nullptr, *d_keep = nullptr;
    const size_t soa = sizeof(float) * 6 * (size_t)w->cap_ent;
    auto release = [&]() { cudaFree(local); cudaFree(world); cudaFree(passes); cudaFree(ent_node); cudaFree(d_keep); };
    if (cudaMalloc(&local, soa) != cudaSuccess || cudaMalloc(&world, soa) != cudaSuccess ||
        cudaMalloc(&passes, sizeof(uint32_t) * (size_t)w->cap_ent) != cudaSuccess ||
        cudaMalloc(&ent_node, sizeof(uint32_t) * (size_t)w->cap_ent) != cudaSuccess ||
        cudaMalloc(&d_keep, sizeof(uint32_t) * (size_t)std::max<int64_t>(n_new, 1)) != cudaSuccess) {
        cudaGetLastError();
        release();
        return fail(OVIS_ERR_NOMEM, "ovis_world_compact: out of device memory (nothing was changed)");
    }
    OVIS_CUDA(cudaMemsetAsync(local, 0, soa, w->stream));
    OVIS_CUDA(cudaMemsetAsync(world, 0, soa, w->stream));
    OVIS_CUDA(cudaMemsetAsync(passes, 0, sizeof(uint32_t) * (size_t)w->cap_ent, w->stream));
    OVIS_CUDA(cudaMemsetAsync(ent_node, 0xff, sizeof(uint32_t) * (size_t)w->cap_ent, w->stream));
    OVIS_CUDA(cudaMemcpyAsync(d_keep, keep.data(), sizeof(uint32_t) * (size_t)n_new, cudaMemcpyHostToDevice, w->stream));
    ovis::launch_compact_soa6(w->d_local, d_keep, n_new, w->cap_ent, local, w->stream);
    ovis::launch_compact_soa6(w->d_world, d_keep, n_new, w->cap_ent, world, w->stream);
    ovis::launch_compact_u32(w->d_pass_counts, d_keep, n_new, passes, w->stream);
    ovis::launch_compact_u32(w->d_entity_node, d_keep, n_new, ent_node, w->stream);
    w->launches += 4;
    OVIS_CUDA(cudaGetLastError());
    OVIS_CUDA(cudaStreamSynchronize(w->stream));
    cudaFree(w->d_local); cudaFree(w->d_world); cudaFree(w->d_pass_counts); cudaFree(w->d_entity_node); cudaFree(d_keep);
    w->d_local = local; w->d_world = world; w->d_pass_counts = passes; w->d_entity_node = ent_node;
    // ---- host mirrors
    std::vector<uint32_t> node_of((size_t)n_new), caster((size_t)(w->cap_ent / 32), 0u);
    for (int64_t i = 0; i < n_new; i++) {
        const uint32_t e = keep[(size_t)i];
        node_of[(size_t)i] = w->h_entity_node[e];
        if ((w->h_caster_raw[e >> 5] >> (e & 31)) & 1u)
            caster[(size_t)i >> 5] |= 1u << (i & 31);
    }
    w->h_entity_node = node_of;
    w->h_caster_raw = caster;
    std::fill(w->h_alive.begin(), w->h_alive.end(), 0u);
    for (int64_t i = 0; i < n_new; i++)
        w->h_alive[(size_t)i >> 5] |= 1u << (i & 31);
    w->n_ent = n_new;
    w->n_live = n_new;
    w->any_dead = false;
    mark_masks(w, 0, std::max<int64_t>(n_old, 1) - 1);
    // ---- which entity a node carries: first bound entity per node in id order, the rest as (entity, node) pairs
    if (w->n_nodes > 0) {
        std::fill(w->h_node_entity.begin(), w->h_node_entity.end(), kNone);
        w->h_extra_entity.clear();
        w->h_extra_node.clear();
        for (int64_t i = 0; i < n_new; i++) {
            const uint32_t n = w->h_entity_node[(size_t)i];
            if (n == kNone)
                continue;
            if (w->h_node_entity[n] == kNone) {
                w->h_node_entity[n] = (uint32_t)i;
            } else {
                w->h_extra_entity.push_back((uint32_t)i);
                w->h_extra_node.push_back(n);
            }
        }
        OVIS_CUDA(cudaMemcpyAsync(w->d_node_entity, w->h_node_entity.data(), sizeof(uint32_t) * (size_t)w->n_nodes, cudaMemcpyHostToDevice, w->stream));
        OVIS_CUDA(cudaStreamSynchronize(w->stream));
        w->extra_dirty = true;
    }
    w->last_visible = 0;                                     // the device result list holds old ids
    return OVIS_OK;
}

int ovis_light_bind_node(OvisWorld *w, uint32_t light, int32_t node) {
    if (!w || light >= w->lights.size() || !w->light_alive[light])
        return fail(OVIS_ERR_INVALID, "ovis_light_bind_node: unknown light %u", light);
    if (node != OVIS_NO_PARENT && (node < 0 || (int64_t)node >= w->n_nodes || (!w->h_node_dead.empty() && w->h_node_dead[(size_t)node])))
        return fail(OVIS_ERR_INVALID, "ovis_light_bind_node: unknown node %d", node);
    w->light_node.resize(w->lights.size(), OVIS_NO_PARENT);
    if (w->light_node[light] >= 0)
        w->n_bound_lights--;
    w->light_node[light] = node;
    if (node >= 0)
        w->n_bound_lights++;
    w->light_bindings_dirty = true;
    return OVIS_OK;
}

// ---- lights ------------------------------------------------------------------------------------

int ovis_light_add(OvisWorld *w, int32_t n, const OvisLight *lights, uint32_t *out_first_id) {
    if (!w || n < 0 || (n > 0 && !lights))
        return fail(OVIS_ERR_INVALID, "ovis_light_add: bad arguments");
    for (int32_t i = 0; i < n; i++)
        if (lights[i].type == OVIS_LIGHT_SPOT && lights[i].cutoff > 45.0f)
            return fail(OVIS_ERR_INVALID, "spot light cutoff must be <= 45 degrees");   // render_light.cc:104
    const int64_t first = (int64_t)w->lights.size();
    if (out_first_id)
        *out_first_id = (uint32_t)first;
    for (int32_t i = 0; i < n; i++) {
        w->lights.push_back(lights[i]);
        w->light_alive.push_back(1);
        // the defaults of the graphics Light components (graphics/src/light.cc:44,63,72)
        OvisWorld::LightShading sh = { { 0.0f, 0.0f, 0.0f }, -0.001f };
        if (lights[i].type == OVIS_LIGHT_POINT)
            sh = { { 1.0f, 0.045f, 0.0075f }, -0.001f };
        else if (lights[i].type == OVIS_LIGHT_SPOT)
            sh = { { 1.0f, 0.09f, 0.032f }, -0.001f };
        w->light_shading.push_back(sh);
    }
    if (n > 0)
        mark_lights(w, first, first + n);
    return OVIS_OK;
}

int ovis_light_update(OvisWorld *w, uint32_t id, const OvisLight *light) {
    if (!w || !light || id >= w->lights.size() || !w->light_alive[id])
        return fail(OVIS_ERR_INVALID, "ovis_light_update: unknown light %u", id);
    if (light->type == OVIS_LIGHT_SPOT && light->cutoff > 45.0f)
        return fail(OVIS_ERR_INVALID, "spot light cutoff must be <= 45 degrees");
    if (std::memcmp(&w->lights[id], light, sizeof(OvisLight)) == 0)
        return OVIS_OK;                                          // unchanged (callers that re-describe every light per cull)
    w->lights[id] = *light;
    mark_lights(w, id, (int64_t)id + 1);
    return OVIS_OK;
}

int ovis_light_set_shading(OvisWorld *w, uint32_t id, const float attenuation[3], float shadow_bias_constant) {
    if (!w || !attenuation || id >= w->lights.size() || !w->light_alive[id])
        return fail(OVIS_ERR_INVALID, "ovis_light_set_shading: unknown light %u", id);
    std::memcpy(w->light_shading[id].attenuation, attenuation, 12);
    w->light_shading[id].shadow_bias = shadow_bias_constant;
    mark_lights(w, id, (int64_t)id + 1);
    return OVIS_OK;
}

int ovis_light_get(OvisWorld *w, uint32_t id, OvisLight *out_light, float out_attenuation[3], float *out_shadow_bias) {
    if (!w || id >= w->lights.size() || !w->light_alive[id])
        return fail(OVIS_ERR_INVALID, "ovis_light_get: unknown light %u", id);
    if (out_light)
        *out_light = w->lights[id];
    if (out_attenuation)
        std::memcpy(out_attenuation, w->light_shading[id].attenuation, 12);
    if (out_shadow_bias)
        *out_shadow_bias = w->light_shading[id].shadow_bias;
    return OVIS_OK;
}

int ovis_light_remove(OvisWorld *w, uint32_t id) {
    if (!w || id >= w->lights.size() || !w->light_alive[id])
        return fail(OVIS_ERR_INVALID, "ovis_light_remove: unknown light %u", id);
    w->light_alive[id] = 0;
    if (id < w->light_node.size() && w->light_node[id] >= 0) {
        w->light_node[id] = OVIS_NO_PARENT;
        w->n_bound_lights--;
    }
    mark_lights(w, id, (int64_t)id + 1);
    return OVIS_OK;
}

// ---- host-side set-up --------------------------------------------------------------------------

int ovis_view_planes(const OvisView *view, float out_planes[24]) {
    if (!view || !out_planes)
        return fail(OVIS_ERR_INVALID, "ovis_view_planes: null argument");
    ovis::hm::viewPlanes(*view, out_planes);
    return OVIS_OK;
}

int ovis_view_matrices(const OvisView *view, float out_mats[64], float out_corners[24]) {
    if (!view)
        return fail(OVIS_ERR_INVALID, "ovis_view_matrices: null view");
    ovis::hm::ViewMatrices m = ovis::hm::viewMatrices(*view);
    if (out_mats) {
        std::memcpy(out_mats, &m.view, 64);
        std::memcpy(out_mats + 16, &m.projection, 64);
        std::memcpy(out_mats + 32, &m.viewProjection, 64);
        std::memcpy(out_mats + 48, &m.inverseViewProjection, 64);
    }
    if (out_corners)
        ovis::hm::frustumCorners(m.inverseViewProjection, out_corners);
    return OVIS_OK;
}

int ovis_view_uniforms(const OvisView *view, const int32_t viewport_pos[2], void *out288) {
    if (!view || !out288)
        return fail(OVIS_ERR_INVALID, "ovis_view_uniforms: null argument");
    const int32_t origin[2] = { 0, 0 };
    ovis::hm::viewUniforms(*view, viewport_pos ? viewport_pos : origin, out288);
    return OVIS_OK;
}

int ovis_light_uniforms(const OvisLight *light, const float attenuation[3], float shadow_bias_constant, void *out208) {
    if (!light || !attenuation || !out208)
        return fail(OVIS_ERR_INVALID, "ovis_light_uniforms: null argument");
    if (light->type == OVIS_LIGHT_SPOT && light->cutoff > 45.0f)
        return fail(OVIS_ERR_INVALID, "spot light cutoff must be <= 45 degrees");
    ovis::hm::lightUniforms(*light, attenuation, shadow_bias_constant, out208);
    return OVIS_OK;
}

int ovis_light_spot_aabb(const OvisLight *light, float out_aabb[6]) {
    if (!light || !out_aabb)
        return fail(OVIS_ERR_INVALID, "ovis_light_spot_aabb: null argument");
    ovis::hm::spotAabb(*light, out_aabb);
    return OVIS_OK;
}

int ovis_light_shadow_views(const OvisLight *light, int32_t shadow_map_resolution, OvisView out_views[6], int32_t *out_count) {
    if (!light || !out_views || !out_count || shadow_map_resolution <= 0)
        return fail(OVIS_ERR_INVALID, "ovis_light_shadow_views: bad argument");
    *out_count = ovis::hm::shadowViews(*light, shadow_map_resolution, out_views);
    return OVIS_OK;
}

// ---- culling -----------------------------------------------------------------------------------

int ovis_cull_planes_device(OvisWorld *w, const float planes[24], uint32_t require, uint32_t *d_out_entities, uint32_t *d_out_count) {
    if (!w || !planes || !d_out_entities || !d_out_count)
        return fail(OVIS_ERR_INVALID, "ovis_cull_planes_device: null argument");
    OVIS_TRY(use_device(w));
    OVIS_TRY(flush_transforms(w));
    return enqueue_cull(w, planes, require, d_out_entities, (long long)w->n_ent, nullptr, nullptr, d_out_count);
}

int ovis_cull_planes_device_multi(OvisWorld *w, const float planes[24], uint32_t require, int64_t capacity,
                                  int32_t n_dst, uint32_t *const *d_out_entities, uint32_t *const *d_out_count) {
    if (!w || !planes || n_dst < 1 || n_dst > ovis::kMaxPeers + 1 || !d_out_entities || !d_out_count || capacity < 0)
        return fail(OVIS_ERR_INVALID, "ovis_cull_planes_device_multi: bad arguments (1..%d destinations)", ovis::kMaxPeers + 1);
    for (int32_t i = 0; i < n_dst; i++)
        if (!d_out_entities[i] || !d_out_count[i])
            return fail(OVIS_ERR_INVALID, "ovis_cull_planes_device_multi: destination %d is null", i);
    OVIS_TRY(use_device(w));
    OVIS_TRY(flush_transforms(w));
    if (w->n_ent == 0) {
        for (int32_t i = 0; i < n_dst; i++) {
            ovis::launch_empty_result(nullptr, nullptr, d_out_count[i], w->stream);
            w->launches++;
        }
        return OVIS_OK;
    }
    return enqueue_cull(w, planes, require, d_out_entities[0], capacity, nullptr, nullptr, d_out_count[0], n_dst - 1,
                        d_out_entities + 1, d_out_count + 1);
}

int ovis_cull_planes_device_exchange(OvisWorld *w, const float planes[24], uint32_t require, int64_t capacity,
                                     int32_t n_dst, uint32_t *const *d_out_entities, uint32_t *const *d_out_count,
                                     int32_t n_wait, const uint32_t *d_wait_flags, uint32_t wait_value,
                                     int32_t n_signal, uint32_t *const *d_signal_flags, uint32_t signal_value) {
    if (!w || !planes || n_dst < 1 || n_dst > ovis::kMaxPeers + 1 || !d_out_entities || !d_out_count || capacity < 0 ||
        n_wait < 0 || n_wait > 32 || (n_wait > 0 && !d_wait_flags) || n_signal < 0 || n_signal > ovis::kMaxPeers + 1 ||
        (n_signal > 0 && !d_signal_flags))
        return fail(OVIS_ERR_INVALID, "ovis_cull_planes_device_exchange: bad arguments");
    OVIS_TRY(use_device(w));
    OVIS_TRY(flush_transforms(w));
    if (w->n_ent == 0)
        return fail(OVIS_ERR_INVALID, "ovis_cull_planes_device_exchange: the shard holds no entities");
    FlagOpts f;
    f.n_wait = n_wait;
    f.wait_flags = d_wait_flags;
    f.wait_value = wait_value;
    f.n_signal = n_signal;
    f.signal_flags = d_signal_flags;
    f.signal_value = signal_value;
    return enqueue_cull(w, planes, require, d_out_entities[0], capacity, nullptr, nullptr, d_out_count[0], n_dst - 1,
                        d_out_entities + 1, d_out_count + 1, &f);
}

int ovis_wait_flags(OvisWorld *w, const uint32_t *d_flags, int32_t n, uint32_t value) {
    if (!w || !d_flags || n < 1)
        return fail(OVIS_ERR_INVALID, "ovis_wait_flags: bad arguments");
    OVIS_TRY(use_device(w));
    ovis::launch_wait_flags(d_flags, n, value, w->d_exchange + 1, w->stream);
    w->launches++;
    OVIS_CUDA(cudaGetLastError());
    return OVIS_OK;
}

static uint32_t *host_alias(OvisWorld *w, const uint32_t *host);

int ovis_read_segments_host(OvisWorld *w, const uint32_t *d_flags, uint32_t value, const uint32_t *d_segments,
                            int64_t segment_words, int32_t n_segments, uint32_t *host_out) {
    if (!w || !d_segments || !host_out || segment_words < 1 || n_segments < 1 || n_segments > 65535)
        return fail(OVIS_ERR_INVALID, "ovis_read_segments_host: bad arguments");
    OVIS_TRY(use_device(w));
    uint32_t *alias = host_alias(w, host_out);
    if (alias) {
        ovis::launch_segments_to_host(d_flags, value, d_segments, segment_words, n_segments, alias, w->d_exchange + 1, w->stream);
        w->launches++;
    } else {                                                    // pageable destination: wait, then copy every segment whole
        if (d_flags) {
            ovis::launch_wait_flags(d_flags, n_segments, value, w->d_exchange + 1, w->stream);
            w->launches++;
        }
        OVIS_CUDA(cudaMemcpyAsync(host_out, d_segments, sizeof(uint32_t) * (size_t)(segment_words * n_segments),
                                  cudaMemcpyDeviceToHost, w->stream));
    }
    OVIS_CUDA(cudaGetLastError());
    OVIS_CUDA(cudaStreamSynchronize(w->stream));
    return OVIS_OK;
}

int ovis_world_exchange_error(OvisWorld *w, int32_t *out_error) {
    if (!w || !out_error)
        return fail(OVIS_ERR_INVALID, "ovis_world_exchange_error: null argument");
    OVIS_TRY(use_device(w));
    unsigned int v = 0;
    OVIS_CUDA(cudaMemcpyAsync(&v, w->d_exchange + 1, sizeof v, cudaMemcpyDeviceToHost, w->stream));
    OVIS_CUDA(cudaStreamSynchronize(w->stream));
    *out_error = (int32_t)v;
    return OVIS_OK;
}

// Device alias of a caller's host buffer when the GPU can store to it directly: pinned memory from ovis_host_alloc,
// cudaHostAlloc / cudaMallocHost or cudaHostRegister(..., mapped). NULL for pageable memory. Asked of the driver on every
// call (well under a microsecond): a cached answer could outlive the allocation it was given for.
static uint32_t *host_alias(OvisWorld *w, const uint32_t *host) {
    if (!host || !w->direct_host || !w->h_counts_dev)
        return nullptr;
    cudaPointerAttributes attr{};
    uint32_t *dev = nullptr;
    if (cudaPointerGetAttributes(&attr, host) == cudaSuccess) {
        if (attr.type == cudaMemoryTypeHost && attr.devicePointer)
            dev = static_cast<uint32_t *>(attr.devicePointer);
    } else {
        cudaGetLastError();
    }
    return dev;
}

static int cull_to_host(OvisWorld *w, const float planes[24], uint32_t require, bool with_lights,
                        uint32_t *out_entities, int64_t cap_entities, int64_t *out_n_entities,
                        uint32_t *out_lights, int32_t cap_lights, int32_t *out_n_lights) {
    MvTrace trace;                                             // OVIS_MV_TRACE=1: host-side time stamps on stderr
    OVIS_TRY(flush_transforms(w));
    // Pinned output buffer: the gather kernel stores the ids (and the count) straight into host memory as its second
    // destination, so the frame is two kernels and one synchronise - no copy is enqueued at all.
    uint32_t *alias = cap_entities > 0 ? host_alias(w, out_entities) : nullptr;
    // With the ids going straight to host memory the only thing left to wait for is the kernel's end. The gather kernel's
    // last block can say so itself: after a system-scope fence it stores the frame number into a page-locked word, and the
    // host watches that word instead of asking the driver (cudaStreamSynchronize wakes up a few microseconds later).
    const bool watch = alias && !with_lights && w->host_flag && w->n_ent > 0;      // (an empty world launches no gather)
    if (alias) {
        uint32_t *extra_out[1] = { alias }, *extra_count[1] = { w->h_counts_dev };
        FlagOpts f;
        uint32_t *flag_ptr[1] = { w->h_counts_dev + 2 };
        if (watch) {
            f.n_signal = 1;
            f.signal_flags = flag_ptr;
            f.signal_value = ++w->host_frame;
        }
        OVIS_TRY(enqueue_cull(w, planes, require, w->d_result, (long long)w->cap_ent, nullptr, nullptr, w->d_counts, 1,
                              extra_out, extra_count, watch ? &f : nullptr, (long long)cap_entities));
    } else {
        OVIS_TRY(enqueue_cull(w, planes, require, w->d_result, (long long)w->cap_ent, nullptr, nullptr, w->d_counts));
    }
    if (with_lights)
        OVIS_TRY(enqueue_light_cull(w, planes));
    trace.mark("cull: kernels queued");
    // Pageable output buffer: one round trip in the common case. Visible sets change little from cull to cull, so the
    // ids are copied speculatively (a little more than the previous call needed) together with the counts; only a call
    // whose list turns out longer pays a second copy for the remainder.
    int64_t spec_e = (out_entities && !alias) ? std::min<int64_t>(cap_entities, std::min<int64_t>(w->n_ent, w->last_visible + w->last_visible / 4 + 1024)) : 0;
    int32_t spec_l = (with_lights && out_lights) ? (int32_t)std::min<int64_t>(cap_lights, (int64_t)w->lights.size()) : 0;
    if (!alias)
        OVIS_CUDA(cudaMemcpyAsync(w->h_counts, w->d_counts, 2 * sizeof(uint32_t), cudaMemcpyDeviceToHost, w->stream));
    else if (with_lights)
        OVIS_CUDA(cudaMemcpyAsync(w->h_counts + 1, w->d_counts + 1, sizeof(uint32_t), cudaMemcpyDeviceToHost, w->stream));
    if (spec_e > 0)
        OVIS_CUDA(cudaMemcpyAsync(out_entities, w->d_result, sizeof(uint32_t) * (size_t)spec_e, cudaMemcpyDeviceToHost, w->stream));
    if (spec_l > 0)
        OVIS_CUDA(cudaMemcpyAsync(out_lights, w->d_light_out, sizeof(uint32_t) * (size_t)spec_l, cudaMemcpyDeviceToHost, w->stream));
    bool arrived = false;
    if (watch) {
        const volatile uint32_t *flag = w->h_counts + 2;
        const auto t0 = std::chrono::steady_clock::now();
        for (unsigned spins = 0; !arrived; spins++) {
            arrived = *flag == w->host_frame;
            if (!arrived && (spins & 1023u) == 1023u &&
                std::chrono::steady_clock::now() - t0 > std::chrono::milliseconds(50))
                break;                                         // something is wrong (or very slow): let the driver tell
        }
        std::atomic_thread_fence(std::memory_order_acquire);
    }
    if (!arrived)
        OVIS_CUDA(cudaStreamSynchronize(w->stream));
    trace.mark("cull: synchronised");
    int64_t ne = w->h_counts[0];
    int32_t nl = with_lights ? (int32_t)w->h_counts[1] : 0;
    w->last_visible = ne;
    if (with_lights)
        w->last_visible_lights = nl;
    w->light_cull_valid = false;
    if (out_n_entities) *out_n_entities = ne;
    if (out_n_lights) *out_n_lights = nl;
    if (ne > cap_entities || (ne > 0 && !out_entities))
        return fail(OVIS_ERR_CAPACITY, "entity result needs %lld entries, buffer holds %lld", (long long)ne, (long long)cap_entities);
    if (nl > cap_lights || (nl > 0 && !out_lights))
        return fail(OVIS_ERR_CAPACITY, "light result needs %d entries, buffer holds %d", nl, cap_lights);
    bool again = false;
    if (!alias && ne > spec_e) {
        OVIS_CUDA(cudaMemcpyAsync(out_entities + spec_e, w->d_result + spec_e, sizeof(uint32_t) * (size_t)(ne - spec_e),
                                  cudaMemcpyDeviceToHost, w->stream));
        again = true;
    }
    if (nl > spec_l) {
        OVIS_CUDA(cudaMemcpyAsync(out_lights + spec_l, w->d_light_out + spec_l, sizeof(uint32_t) * (size_t)(nl - spec_l),
                                  cudaMemcpyDeviceToHost, w->stream));
        again = true;
    }
    if (again)
        OVIS_CUDA(cudaStreamSynchronize(w->stream));
    if (with_lights && out_lights) {                       // kept for ovis_build_shadow_lists of the same view
        w->h_visible_lights.assign(out_lights, out_lights + nl);
        std::memcpy(w->h_light_cull_planes, planes, sizeof w->h_light_cull_planes);
        w->light_cull_valid = true;
    }
    return OVIS_OK;
}

int ovis_cull_planes(OvisWorld *w, const float planes[24], uint32_t require,
                     uint32_t *out_entities, int64_t cap_entities, int64_t *out_n_entities) {
    if (!w || !planes)
        return fail(OVIS_ERR_INVALID, "ovis_cull_planes: null argument");
    OVIS_TRY(use_device(w));
    return cull_to_host(w, planes, require, false, out_entities, cap_entities, out_n_entities, nullptr, 0, nullptr);
}

int ovis_cull(OvisWorld *w, const OvisView *view, uint32_t flags,
              uint32_t *out_entities, int64_t cap_entities, int64_t *out_n_entities,
              uint32_t *out_lights, int32_t cap_lights, int32_t *out_n_lights) {
    if (!w || !view)
        return fail(OVIS_ERR_INVALID, "ovis_cull: null argument");
    OVIS_TRY(use_device(w));
    float planes[24];
    ovis::hm::viewPlanes(*view, planes);
    return cull_to_host(w, planes, 0, (flags & OVIS_CULL_LIGHTS) != 0, out_entities, cap_entities, out_n_entities,
                        out_lights, cap_lights, out_n_lights);
}

int ovis_cull_views_device(OvisWorld *w, int32_t n_views, const float *planes, const uint32_t *require,
                           const int64_t **d_offsets, const uint32_t **d_entities, int64_t *host_offsets) {
    if (!w || n_views < 0 || (n_views > 0 && !planes))
        return fail(OVIS_ERR_INVALID, "ovis_cull_views_device: bad arguments");
    OVIS_TRY(use_device(w));
    std::vector<long long> offsets;
    OVIS_TRY(cull_views_device(w, n_views, planes, require, host_offsets, offsets));
    if (d_offsets) *d_offsets = (const int64_t *)w->d_csr_offsets;
    if (d_entities) *d_entities = w->d_csr_entities;
    return OVIS_OK;
}

int ovis_cull_views(OvisWorld *w, int32_t n_views, const float *planes, const uint32_t *require, int64_t *out_offsets,
                    int64_t cap_entities, uint32_t *out_entities, int64_t *out_n_entities) {
    if (!w || n_views < 0 || (n_views > 0 && !planes) || !out_offsets || !out_n_entities || cap_entities < 0 ||
        (cap_entities > 0 && !out_entities))
        return fail(OVIS_ERR_INVALID, "ovis_cull_views: bad arguments");
    OVIS_TRY(use_device(w));
    std::vector<long long> offsets;
    w->mv_spec_out = out_entities;
    w->mv_spec_cap = out_entities ? cap_entities : 0;
    {
        const int rc = cull_views_device(w, n_views, planes, require, out_offsets, offsets);
        if (rc != OVIS_OK) {
            w->mv_spec_out = nullptr;
            return rc;
        }
    }
    const int64_t total = offsets.empty() ? 0 : offsets[(size_t)n_views];
    *out_n_entities = total;
    if (total > cap_entities) {
        w->mv_spec_out = nullptr;
        return fail(OVIS_ERR_CAPACITY, "ovis_cull_views: the lists need %lld entries, buffer holds %lld", (long long)total, (long long)cap_entities);
    }
    return finish_host_entities(w, out_entities, total);
}

int ovis_cull_shadow_views_device(OvisWorld *w, int32_t n_lights, const uint32_t *light_ids, int32_t shadow_map_resolution,
                                  int32_t cap_lists, int32_t *out_list_light, int32_t *out_list_view, int32_t *out_n_lists,
                                  const int64_t **d_offsets, const uint32_t **d_entities, int64_t *host_offsets) {
    if (!w || n_lights < 0 || (n_lights > 0 && !light_ids) || !out_n_lists || shadow_map_resolution <= 0)
        return fail(OVIS_ERR_INVALID, "ovis_cull_shadow_views_device: bad arguments");
    OVIS_TRY(use_device(w));
    std::vector<uint32_t> ids(light_ids, light_ids + n_lights);
    for (uint32_t id : ids)
        if (id >= w->lights.size() || !w->light_alive[id])
            return fail(OVIS_ERR_INVALID, "ovis_cull_shadow_views_device: unknown light %u", id);
    std::vector<int32_t> list_light, list_view;
    std::vector<long long> offsets;
    OVIS_TRY(shadow_views_device(w, ids, shadow_map_resolution, list_light, list_view, offsets));
    int32_t n_lists = (int32_t)list_light.size();
    *out_n_lists = n_lists;
    if (n_lists > cap_lists)
        return fail(OVIS_ERR_CAPACITY, "shadow views need %d list slots, %d given", n_lists, cap_lists);
    for (int32_t k = 0; k < n_lists; k++) {
        if (out_list_light) out_list_light[k] = list_light[(size_t)k];
        if (out_list_view) out_list_view[k] = list_view[(size_t)k];
    }
    if (host_offsets)
        for (int32_t k = 0; k <= n_lists; k++)
            host_offsets[k] = offsets[(size_t)k];
    if (d_offsets) *d_offsets = (const int64_t *)w->d_csr_offsets;
    if (d_entities) *d_entities = w->d_csr_entities;
    return OVIS_OK;
}

int ovis_cull_shadow_views(OvisWorld *w, int32_t n_lights, const uint32_t *light_ids, int32_t shadow_map_resolution, uint32_t require,
                           int32_t cap_lists, int32_t *out_list_light, int32_t *out_list_view, int64_t *out_offsets, int32_t *out_n_lists,
                           int64_t cap_entities, uint32_t *out_entities, int64_t *out_n_entities) {
    if (!w || n_lights < 0 || (n_lights > 0 && !light_ids) || !out_n_lists || !out_n_entities || shadow_map_resolution <= 0 ||
        (require & ~(uint32_t)OVIS_CASTS_SHADOW))
        return fail(OVIS_ERR_INVALID, "ovis_cull_shadow_views: bad arguments");
    OVIS_TRY(use_device(w));
    std::vector<uint32_t> ids(light_ids, light_ids + n_lights);
    for (uint32_t id : ids)
        if (id >= w->lights.size() || !w->light_alive[id])
            return fail(OVIS_ERR_INVALID, "ovis_cull_shadow_views: unknown light %u", id);
    std::vector<int32_t> list_light, list_view;
    std::vector<long long> offsets;
    w->mv_spec_out = out_entities;
    w->mv_spec_cap = out_entities ? cap_entities : 0;
    {
        const int rc = shadow_views_device(w, ids, shadow_map_resolution, list_light, list_view, offsets, require);
        if (rc != OVIS_OK) {
            w->mv_spec_out = nullptr;
            return rc;
        }
    }
    const int32_t n_lists = (int32_t)list_light.size();
    const int64_t total = offsets.empty() ? 0 : offsets[(size_t)n_lists];
    *out_n_lists = n_lists;
    *out_n_entities = total;
    if (n_lists > cap_lists || total > cap_entities)
        return fail(OVIS_ERR_CAPACITY, "shadow views need %d lists / %lld entities", n_lists, (long long)total);
    if ((n_lists > 0 && (!out_list_light || !out_list_view)) || !out_offsets || (total > 0 && !out_entities))
        return fail(OVIS_ERR_INVALID, "ovis_cull_shadow_views: null output");
    for (int32_t k = 0; k < n_lists; k++) {
        out_list_light[k] = list_light[(size_t)k];
        out_list_view[k] = list_view[(size_t)k];
    }
    for (int32_t k = 0; k <= n_lists; k++)
        out_offsets[k] = offsets.empty() ? 0 : offsets[(size_t)k];
    return finish_host_entities(w, out_entities, total);
}

int ovis_build_shadow_lists(OvisWorld *w, const OvisView *main_view, int32_t shadow_map_resolution,
                            int32_t cap_lists, int32_t *out_list_light, int32_t *out_list_view,
                            int64_t *out_offsets, int32_t *out_n_lists,
                            int64_t cap_entities, uint32_t *out_entities, int64_t *out_n_entities) {
    if (!w || !main_view || !out_n_lists || !out_n_entities || shadow_map_resolution <= 0)
        return fail(OVIS_ERR_INVALID, "ovis_build_shadow_lists: bad arguments");
    OVIS_TRY(use_device(w));
    OVIS_TRY(flush_transforms(w));
    // 1. lights visible from the main view, in insertion order (deferred_render_pipeline.cc:147,207)
    float main_planes[24];
    ovis::hm::viewPlanes(*main_view, main_planes);
    std::vector<uint32_t> visible;
    if (w->light_cull_valid && !w->lights_dirty && std::memcmp(main_planes, w->h_light_cull_planes, sizeof main_planes) == 0) {
        visible = w->h_visible_lights;                  // this frame's ovis_cull(main_view, OVIS_CULL_LIGHTS) found them already
    } else {
        OVIS_TRY(enqueue_light_cull(w, main_planes));
        OVIS_CUDA(cudaMemcpyAsync(w->h_counts + 1, w->d_counts + 1, sizeof(uint32_t), cudaMemcpyDeviceToHost, w->stream));
        OVIS_CUDA(cudaStreamSynchronize(w->stream));
        w->last_visible_lights = w->h_counts[1];      // the device light list is now this view's (pack / shade read it)
        visible.resize(w->h_counts[1]);
        if (!visible.empty()) {
            OVIS_CUDA(cudaMemcpyAsync(visible.data(), w->d_light_out, 4 * visible.size(), cudaMemcpyDeviceToHost, w->stream));
            OVIS_CUDA(cudaStreamSynchronize(w->stream));
        }
        w->h_visible_lights = visible;
        std::memcpy(w->h_light_cull_planes, main_planes, sizeof main_planes);
        w->light_cull_valid = true;
    }
    // 2.+3. their shadow views, culled for shadow casters in one multi-view pass set
    std::vector<int32_t> list_light, list_view;
    std::vector<long long> offsets;
    w->mv_spec_out = out_entities;
    w->mv_spec_cap = out_entities ? cap_entities : 0;
    {
        const int rc = shadow_views_device(w, visible, shadow_map_resolution, list_light, list_view, offsets);
        if (rc != OVIS_OK) {
            w->mv_spec_out = nullptr;
            return rc;
        }
    }
    int32_t n_lists = (int32_t)list_light.size();
    int64_t total = offsets.empty() ? 0 : offsets[(size_t)n_lists];
    *out_n_lists = n_lists;
    *out_n_entities = total;
    if (n_lists > cap_lists || total > cap_entities)
        return fail(OVIS_ERR_CAPACITY, "shadow lists need %d lists / %lld entities", n_lists, (long long)total);
    if ((n_lists > 0 && (!out_list_light || !out_list_view)) || !out_offsets || (total > 0 && !out_entities))
        return fail(OVIS_ERR_INVALID, "ovis_build_shadow_lists: null output");
    for (int32_t k = 0; k < n_lists; k++) {
        out_list_light[k] = list_light[(size_t)k];
        out_list_view[k] = list_view[(size_t)k];
    }
    for (int32_t k = 0; k <= n_lists; k++)
        out_offsets[k] = offsets.empty() ? 0 : offsets[(size_t)k];
    return finish_host_entities(w, out_entities, total);
}

// ---- read-back ---------------------------------------------------------------------------------

static int read_back(OvisWorld *w, int comps, const float *soa, int64_t stride, uint32_t first, int64_t n, float *out) {
    for (int64_t s = 0; s < n; s += kUploadChunk) {
        int64_t c = std::min(kUploadChunk, n - s);
        size_t bytes = sizeof(float) * (size_t)comps * (size_t)c;
        if (w->stage_bytes[0] < bytes) {
            if (w->d_stage[0]) {
                OVIS_CUDA(cudaStreamSynchronize(w->stream));
                OVIS_CUDA(cudaFree(w->d_stage[0]));
                w->d_stage[0] = nullptr;
                w->stage_bytes[0] = 0;
            }
            OVIS_CUDA(cudaMalloc(&w->d_stage[0], bytes));
            w->stage_bytes[0] = bytes;
        }
        ovis::launch_gather_aos(comps, soa, stride, (uint32_t)(first + s), c, (float *)w->d_stage[0], w->stream);
        w->launches++;
        OVIS_CUDA(cudaMemcpyAsync(out + comps * s, w->d_stage[0], bytes, cudaMemcpyDeviceToHost, w->stream));
        OVIS_CUDA(cudaStreamSynchronize(w->stream));
    }
    return OVIS_OK;
}

int ovis_get_world_aabbs(OvisWorld *w, uint32_t first_entity, int64_t n, float *out_aabb) {
    if (!w || n < 0 || (n > 0 && !out_aabb) || (int64_t)first_entity + n > w->n_ent)
        return fail(OVIS_ERR_INVALID, "ovis_get_world_aabbs: bad range");
    OVIS_TRY(use_device(w));
    OVIS_TRY(flush_transforms(w));
    return read_back(w, 6, w->d_world, w->cap_ent, first_entity, n, out_aabb);
}

int ovis_get_node_world_trs(OvisWorld *w, uint32_t first_node, int64_t n, float *out_trs) {
    if (!w || n < 0 || (n > 0 && !out_trs) || (int64_t)first_node + n > w->n_nodes)
        return fail(OVIS_ERR_INVALID, "ovis_get_node_world_trs: bad range");
    OVIS_TRY(use_device(w));
    OVIS_TRY(flush_transforms(w));
    return read_back(w, 10, w->d_world_trs, w->cap_nodes, first_node, n, out_trs);
}

int ovis_entity_set_pass_counts(OvisWorld *w, int64_t n, const uint32_t *ids, const uint32_t *pass_counts) {
    if (!w || (n > 0 && !pass_counts))
        return fail(OVIS_ERR_INVALID, "ovis_entity_set_pass_counts: null input");
    OVIS_TRY(use_device(w));
    OVIS_TRY(check_entity_ids(w, n, ids));
    for (int64_t s = 0; s < n; s += kUploadChunk) {
        int64_t c = std::min(kUploadChunk, n - s);
        void *dv = nullptr, *di = nullptr;
        OVIS_TRY(stage(w, 0, pass_counts + s, 4 * (size_t)c, &dv));
        OVIS_TRY(stage(w, 1, ids + s, 4 * (size_t)c, &di));
        ovis::launch_scatter_u32((const uint32_t *)dv, (const uint32_t *)di, c, w->d_pass_counts, w->stream);
        w->launches++;
        if (s + kUploadChunk < n)
            OVIS_CUDA(cudaStreamSynchronize(w->stream));   // the staging buffers are reused by the next chunk
    }
    OVIS_CUDA(cudaGetLastError());
    return OVIS_OK;
}

int ovis_build_draw_lists_device(OvisWorld *w, const uint32_t *d_visible, const uint32_t *d_visible_count, int64_t max_visible,
                                 int32_t mode, OvisDraw *d_draws0, int64_t capacity0, OvisDraw *d_draws1, int64_t capacity1,
                                 uint32_t *d_draw_counts) {
    if (!w || (mode != OVIS_DRAWS_MAIN && mode != OVIS_DRAWS_SHADOW) || capacity0 < 0 || capacity1 < 0 ||
        (capacity0 > 0 && !d_draws0) || (capacity1 > 0 && !d_draws1))
        return fail(OVIS_ERR_INVALID, "ovis_build_draw_lists_device: bad arguments");
    if ((d_visible == nullptr) != (d_visible_count == nullptr))
        return fail(OVIS_ERR_INVALID, "ovis_build_draw_lists_device: pass both the list and its count, or neither");
    OVIS_TRY(use_device(w));
    if (w->n_ent == 0)
        OVIS_TRY(reserve_entities(w, 1));                       // the count words exist even for an empty world
    if (!d_visible) {                                           // the entity list of the last ovis_cull*
        d_visible = w->d_result;
        d_visible_count = w->d_counts;
        max_visible = w->last_visible;                          // that cull returned its count to the host
    }
    if (max_visible < 0 || max_visible > w->n_ent)
        max_visible = w->n_ent;                                 // a list of distinct entities is no longer than the world
    OVIS_TRY(upload_masks(w));
    ovis::DrawArgs a{};
    a.d_visible = d_visible;
    a.d_visible_count = d_visible_count;
    a.id_base = w->id_base;
    a.n_slots = (uint32_t)w->n_ent;
    a.pass_counts = w->d_pass_counts;
    a.caster_mask = w->d_caster_bits;
    a.mode = mode;
    a.tile_totals = w->d_draw_totals;
    a.d_draws0 = reinterpret_cast<uint2 *>(d_draws0);
    a.d_draws1 = reinterpret_cast<uint2 *>(d_draws1);
    a.cap0 = capacity0;
    a.cap1 = capacity1;
    a.d_draw_counts = d_draw_counts ? d_draw_counts : w->d_draw_counts;
    w->launches += ovis::launch_draw_lists(a, max_visible, w->stream);
    OVIS_CUDA(cudaGetLastError());
    return OVIS_OK;
}

int ovis_pack_light_uniforms_device(OvisWorld *w, const uint32_t *d_visible, const uint32_t *d_visible_count, int64_t max_visible,
                                    float *d_uniforms, int64_t capacity) {
    if (!w || capacity < 0 || (capacity > 0 && !d_uniforms) || ((uintptr_t)d_uniforms & 15u))
        return fail(OVIS_ERR_INVALID, "ovis_pack_light_uniforms_device: bad arguments (the output must be 16-byte aligned)");
    if ((d_visible == nullptr) != (d_visible_count == nullptr))
        return fail(OVIS_ERR_INVALID, "ovis_pack_light_uniforms_device: pass both the list and its count, or neither");
    OVIS_TRY(use_device(w));
    OVIS_TRY(upload_lights(w));
    const int64_t n_lights = (int64_t)w->lights.size();
    if (!d_visible) {                                           // the light list of the last ovis_cull
        d_visible = w->d_light_out;
        d_visible_count = w->d_counts + 1;
        max_visible = w->last_visible_lights;
    }
    if (max_visible < 0 || max_visible > n_lights)
        max_visible = n_lights;
    if (n_lights > 0 && std::min(max_visible, capacity) > 0) {
        ovis::launch_light_uniforms(d_visible, d_visible_count, w->d_light_uniforms, (uint32_t)n_lights, d_uniforms, max_visible, capacity,
                                    w->stream);
        w->launches++;
    }
    OVIS_CUDA(cudaGetLastError());
    return OVIS_OK;
}

int ovis_shade_deferred_device(OvisWorld *w, const OvisGBuffer *g, const void *d_view_uniforms, const uint32_t *d_visible,
                               const uint32_t *d_visible_count, int64_t max_visible, const float *const *d_shadow_maps,
                               int32_t shadow_size, float *d_out_rgba) {
    if (!w || !g || !d_view_uniforms || !d_out_rgba || g->width < 0 || g->height < 0)
        return fail(OVIS_ERR_INVALID, "ovis_shade_deferred_device: null argument or negative size");
    if ((int64_t)g->width * g->height > 0 && (!g->d_buffer_a || !g->d_buffer_b || !g->d_buffer_c || !g->d_depth))
        return fail(OVIS_ERR_INVALID, "ovis_shade_deferred_device: incomplete G-buffer");
    if ((((uintptr_t)g->d_buffer_a | (uintptr_t)g->d_buffer_b | (uintptr_t)g->d_buffer_c | (uintptr_t)d_out_rgba) & 15u) ||
        ((uintptr_t)d_view_uniforms & 3u))
        return fail(OVIS_ERR_INVALID, "ovis_shade_deferred_device: the colour buffers must be 16-byte aligned");
    if (d_shadow_maps && shadow_size < 1)
        return fail(OVIS_ERR_INVALID, "ovis_shade_deferred_device: shadow maps need a size");
    if ((d_visible == nullptr) != (d_visible_count == nullptr))
        return fail(OVIS_ERR_INVALID, "ovis_shade_deferred_device: pass both the light list and its count, or neither");
    OVIS_TRY(use_device(w));
    OVIS_TRY(upload_lights(w));
    const int64_t n_lights = (int64_t)w->lights.size();
    if (!d_visible) {
        d_visible = w->d_light_out;
        d_visible_count = w->d_counts + 1;
        max_visible = w->last_visible_lights;
    }
    if (max_visible < 0 || max_visible > n_lights)
        max_visible = n_lights;
    ovis::ShadeArgs a{};
    a.width = g->width;
    a.height = g->height;
    a.buffer_a = g->d_buffer_a;
    a.buffer_b = g->d_buffer_b;
    a.buffer_c = g->d_buffer_c;
    a.depth = g->d_depth;
    a.view_uniforms = static_cast<const float *>(d_view_uniforms);
    a.d_visible = d_visible;
    a.d_visible_count = d_visible_count;
    a.max_visible = n_lights > 0 ? max_visible : 0;              // no lights: the count word is 0 anyway
    a.n_lights = (uint32_t)n_lights;
    a.light_table = w->d_light_uniforms;
    a.light_types = w->d_light_types;
    a.shadow_maps = d_shadow_maps;
    a.shadow_size = shadow_size;
    a.out_rgba = d_out_rgba;
    if ((int64_t)g->width * g->height > 0) {
        ovis::launch_shade_deferred(a, w->stream);
        w->launches++;
    }
    OVIS_CUDA(cudaGetLastError());
    return OVIS_OK;
}

int ovis_post_effects_device(OvisWorld *w, int64_t n_pixels, const float *d_rgba, uint32_t effects, float exposure, float white_point,
                             float gamma, float *d_out_rgba) {
    if (!w || n_pixels < 0 || (n_pixels > 0 && (!d_rgba || !d_out_rgba)) || (effects & ~(uint32_t)(OVIS_POST_TONEMAP | OVIS_POST_GAMMA)) ||
        (((uintptr_t)d_rgba | (uintptr_t)d_out_rgba) & 15u))
        return fail(OVIS_ERR_INVALID, "ovis_post_effects_device: bad arguments (buffers must be 16-byte aligned)");
    OVIS_TRY(use_device(w));
    if (n_pixels > 0) {
        ovis::launch_post_effects(n_pixels, d_rgba, effects, exposure, white_point, gamma, d_out_rgba, w->stream);
        w->launches++;
    }
    OVIS_CUDA(cudaGetLastError());
    return OVIS_OK;
}

int ovis_pack_view_uniforms_device(OvisWorld *w, int32_t n_views, const OvisView *views, const int32_t *viewport_pos, void *d_uniforms) {
    if (!w || n_views < 0 || (n_views > 0 && (!views || !d_uniforms)))
        return fail(OVIS_ERR_INVALID, "ovis_pack_view_uniforms_device: bad arguments");
    OVIS_TRY(use_device(w));
    if (n_views == 0)
        return OVIS_OK;
    std::vector<uint8_t> blocks(288 * (size_t)n_views);
    const int32_t origin[2] = { 0, 0 };
    for (int32_t i = 0; i < n_views; i++)
        ovis::hm::viewUniforms(views[i], viewport_pos ? viewport_pos + 2 * i : origin, blocks.data() + 288 * (size_t)i);
    // pageable source: the runtime has taken its copy of `blocks` when this returns
    OVIS_CUDA(cudaMemcpyAsync(d_uniforms, blocks.data(), blocks.size(), cudaMemcpyHostToDevice, w->stream));
    return OVIS_OK;
}

int ovis_get_draw_counts(OvisWorld *w, uint32_t out_counts[2]) {
    if (!w || !out_counts)
        return fail(OVIS_ERR_INVALID, "ovis_get_draw_counts: null argument");
    OVIS_TRY(use_device(w));
    out_counts[0] = out_counts[1] = 0;
    if (!w->d_draw_counts)
        return OVIS_OK;
    OVIS_CUDA(cudaMemcpyAsync(out_counts, w->d_draw_counts, 8, cudaMemcpyDeviceToHost, w->stream));
    OVIS_CUDA(cudaStreamSynchronize(w->stream));
    return OVIS_OK;
}

int ovis_pack_entity_uniforms_device(OvisWorld *w, const uint32_t *d_visible, const uint32_t *d_visible_count, int64_t max_visible,
                                     float *d_uniforms, int64_t capacity) {
    if (!w || capacity < 0 || (capacity > 0 && !d_uniforms) || ((uintptr_t)d_uniforms & 15u))
        return fail(OVIS_ERR_INVALID, "ovis_pack_entity_uniforms_device: bad arguments (the output must be 16-byte aligned)");
    if ((d_visible == nullptr) != (d_visible_count == nullptr))
        return fail(OVIS_ERR_INVALID, "ovis_pack_entity_uniforms_device: pass both the list and its count, or neither");
    OVIS_TRY(use_device(w));
    if (w->n_ent == 0)
        OVIS_TRY(reserve_entities(w, 1));
    if (!d_visible) {
        d_visible = w->d_result;
        d_visible_count = w->d_counts;
        max_visible = w->last_visible;                          // that cull returned its count to the host
    }
    if (max_visible < 0 || max_visible > w->n_ent)
        max_visible = w->n_ent;
    OVIS_TRY(flush_transforms(w));
    ovis::UniformArgs a{};
    a.d_visible = d_visible;
    a.d_visible_count = d_visible_count;
    a.id_base = w->id_base;
    a.n_slots = (uint32_t)w->n_ent;
    a.entity_node = w->d_entity_node;
    a.world_trs = w->d_world_trs;
    a.nstride = w->cap_nodes;
    a.d_out = d_uniforms;
    a.capacity = capacity;
    ovis::launch_entity_uniforms(a, std::min(max_visible, capacity), w->stream);
    w->launches += std::min(max_visible, capacity) > 0 ? 1 : 0;
    OVIS_CUDA(cudaGetLastError());
    return OVIS_OK;
}

int ovis_read_device(OvisWorld *w, const void *d_src, int64_t bytes, void *host_dst) {
    if (!w || bytes < 0 || (bytes > 0 && (!d_src || !host_dst)))
        return fail(OVIS_ERR_INVALID, "ovis_read_device: bad arguments");
    OVIS_TRY(use_device(w));
    if (bytes > 0)
        OVIS_CUDA(cudaMemcpyAsync(host_dst, d_src, (size_t)bytes, cudaMemcpyDeviceToHost, w->stream));
    OVIS_CUDA(cudaStreamSynchronize(w->stream));
    return OVIS_OK;
}

int ovis_host_alloc(int64_t bytes, void **out_ptr) {
    if (!out_ptr || bytes < 0)
        return fail(OVIS_ERR_INVALID, "ovis_host_alloc: bad arguments");
    OVIS_CUDA(cudaMallocHost(out_ptr, (size_t)std::max<int64_t>(bytes, 1)));
    return OVIS_OK;
}

int ovis_host_free(void *ptr) {
    if (ptr)
        OVIS_CUDA(cudaFreeHost(ptr));
    return OVIS_OK;
}

int ovis_host_flags_wait(const uint32_t *flags, int32_t n, uint32_t value, int64_t timeout_us) {
    if (!flags || n < 0)
        return fail(OVIS_ERR_INVALID, "ovis_host_flags_wait: bad arguments");
    const volatile uint32_t *f = flags;
    const auto t0 = std::chrono::steady_clock::now();
    for (unsigned spins = 0;; spins++) {
        bool all = true;
        for (int32_t i = 0; i < n; i++)
            all = all && (int32_t)(f[i] - value) >= 0;           // "at least", robust to wrap-around
        if (all)
            break;
        if ((spins & 255u) == 255u && timeout_us >= 0 &&
            std::chrono::steady_clock::now() - t0 > std::chrono::microseconds(timeout_us))
            return fail(OVIS_ERR_CAPACITY, "ovis_host_flags_wait: timed out");
    }
    std::atomic_thread_fence(std::memory_order_acquire);
    return OVIS_OK;
}

// The CPUs next to a GPU, from the kernel's own topology files: /sys/bus/pci/devices/<domain:bus:device.function>/local_cpulist
// ("0-31,64-95"). Only CPUs the calling thread is already allowed to run on are kept (containers, cpusets).
int ovis_host_pin_thread_near_device(int32_t device, int32_t *out_cpus) {
    if (out_cpus)
        *out_cpus = 0;
    char bus[32] = {};
    OVIS_CUDA(cudaDeviceGetPCIBusId(bus, (int)sizeof bus, device));
    for (char *c = bus; *c; c++)
        *c = (char)std::tolower((unsigned char)*c);
    const std::string path = std::string("/sys/bus/pci/devices/") + bus + "/local_cpulist";
    FILE *f = std::fopen(path.c_str(), "r");
    if (!f)
        return OVIS_OK;                                          // no topology information: leave the thread where it is
    char line[4096] = {};
    const bool got = std::fgets(line, (int)sizeof line, f) != nullptr;
    std::fclose(f);
    if (!got)
        return OVIS_OK;
    cpu_set_t allowed, wanted;
    CPU_ZERO(&allowed);
    CPU_ZERO(&wanted);
    if (sched_getaffinity(0, sizeof allowed, &allowed) != 0)
        return OVIS_OK;
    int n = 0;
    for (const char *p = line; *p;) {
        char *end = nullptr;
        const long a = std::strtol(p, &end, 10);
        if (end == p)
            break;
        long b = a;
        p = end;
        if (*p == '-') {
            b = std::strtol(p + 1, &end, 10);
            p = end;
        }
        for (long c = a; c <= b && c < CPU_SETSIZE; c++)
            if (c >= 0 && CPU_ISSET((int)c, &allowed)) {
                CPU_SET((int)c, &wanted);
                n++;
            }
        while (*p == ',' || *p == ' ' || *p == '\n')
            p++;
    }
    if (n > 0 && sched_setaffinity(0, sizeof wanted, &wanted) == 0 && out_cpus)
        *out_cpus = n;
    return OVIS_OK;
}

int ovis_host_register(void *ptr, int64_t bytes) {
    if (!ptr || bytes <= 0)
        return fail(OVIS_ERR_INVALID, "ovis_host_register: bad arguments");
    OVIS_CUDA(cudaHostRegister(ptr, (size_t)bytes, cudaHostRegisterPortable | cudaHostRegisterMapped));
    return OVIS_OK;
}

int ovis_host_unregister(void *ptr) {
    if (ptr)
        OVIS_CUDA(cudaHostUnregister(ptr));
    return OVIS_OK;
}

}  // extern "C"
