// ovis_sharded.cu - one world sharded over several GPUs of one box, driven by ONE host process through the C ABI
// (ovis_world_create(device_ids[], n, ...); section 8(e)).
//
// The reference engine is one process and one thread (engine/src/engine.cc:147-195), so a drop-in RenderWorld cannot ask
// for a process per GPU. Here the caller keeps its single thread; the library owns one worker thread per device, each
// bound to its device for life, so the n culls of a frame are launched side by side instead of one after the other
// (two launches per device cost ~8 us of host time; queued from one thread that is 64 us at 8 GPUs before the last
// device even starts, more than the cull itself).
//
// Entities are sharded by contiguous id range; shard r's world reports global ids (ovis_world_set_id_base). For a HOST
// consumer nothing is exchanged between GPUs: every device's gather kernel stores its segment [count | ids...] straight
// into ONE page-locked buffer that all devices map (cudaHostAllocPortable | Mapped) - the same zero-copy path ovis_cull
// takes on one GPU - and reading the segments in shard order is the reference's list order
// (render/src/simple_render_world.cc:38-41: insertion order). For DEVICE consumers ovis_sharded_allgather_device leaves
// every segment on every GPU with one ncclAllGather per device (the collective BASELINE.json's north_star names);
// NCCL is loaded at run time (dlopen), so the library has no link-time dependency on it.
#include "orion_vis.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <thread>
#include <vector>

#include <dlfcn.h>

#include <cuda_runtime.h>

namespace ovis {
int report_error(int code, const char *message);
}

namespace {

int fail(int code, const std::string &m) { return ovis::report_error(code, m.c_str()); }

// ---- the few NCCL entry points used, resolved at run time
struct Nccl {
    void *lib = nullptr;
    int (*CommInitAll)(void **comms, int n, const int *devs) = nullptr;
    int (*CommDestroy)(void *comm) = nullptr;
    int (*AllGather)(const void *send, void *recv, size_t count, int dtype, void *comm, cudaStream_t s) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
    bool load() {
        if (lib)
            return true;
        for (const char *name : { "libnccl.so.2", "libnccl.so" }) {
            lib = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
            if (lib)
                break;
        }
        if (!lib)
            return false;
        CommInitAll = reinterpret_cast<decltype(CommInitAll)>(dlsym(lib, "ncclCommInitAll"));
        CommDestroy = reinterpret_cast<decltype(CommDestroy)>(dlsym(lib, "ncclCommDestroy"));
        AllGather = reinterpret_cast<decltype(AllGather)>(dlsym(lib, "ncclAllGather"));
        GroupStart = reinterpret_cast<decltype(GroupStart)>(dlsym(lib, "ncclGroupStart"));
        GroupEnd = reinterpret_cast<decltype(GroupEnd)>(dlsym(lib, "ncclGroupEnd"));
        GetErrorString = reinterpret_cast<decltype(GetErrorString)>(dlsym(lib, "ncclGetErrorString"));
        return CommInitAll && CommDestroy && AllGather && GroupStart && GroupEnd;
    }
};
constexpr int kNcclUint32 = 3;                      // ncclUint32 (nccl.h: int8 0, uint8 1, int32 2, uint32 3)

struct Worker {
    std::thread thread;
    std::atomic<uint64_t> posted{ 0 };             // last job number handed to this worker
    std::atomic<uint64_t> finished{ 0 };           // last job number it completed
    int rc = OVIS_OK;
    std::string error;
    int64_t count = 0;
    // many-view jobs: this shard's CSR (page-locked entity buffer, grown by the worker itself)
    std::vector<int64_t> offsets;
    std::vector<int32_t> list_light, list_view;
    int32_t n_lists = 0;
    uint32_t *h_entities = nullptr;
    int64_t cap_entities = 0;
};
enum JobKind { kJobCull = 0, kJobViews = 1, kJobShadowLists = 2 };

}  // namespace

struct OvisShardedWorld {
    int n = 0;
    std::vector<int> devices;
    std::vector<OvisWorld *> shards;
    std::vector<int64_t> base;                      // global id of each shard's first entity, n + 1 entries
    std::vector<Worker> workers;
    std::mutex mutex;
    std::condition_variable wake;
    std::atomic<bool> quit{ false };
    uint64_t job = 0;
    // the job: one cull of every shard into the shared host segments
    float planes[24];
    uint32_t require = 0;
    int kind = kJobCull;
    // ... or one many-view pass of every shard into its worker's CSR
    int32_t n_views = 0;
    const float *view_planes = nullptr;
    const uint32_t *view_require = nullptr;
    OvisView main_view;
    int32_t shadow_res = 0;
    uint32_t *h_segments = nullptr;                 // portable pinned: n x (1 + seg_capacity) words
    int64_t seg_capacity = 0;
    // device consumers (NCCL)
    Nccl nccl;
    std::vector<void *> comms;
    std::vector<cudaStream_t> streams;
    std::vector<uint32_t *> d_send, d_recv;         // per device: [1 + cap] and n x [1 + cap]
    int64_t dev_capacity = 0;
};

namespace {

// One many-view pass of shard r into the worker's own buffers; a result that outgrows them is repeated once with more room.
void many_views_job(OvisShardedWorld *sw, int r) {
    Worker &me = sw->workers[r];
    for (int attempt = 0; attempt < 3; attempt++) {
        int64_t total = 0;
        if (sw->kind == kJobViews) {
            me.offsets.assign((size_t)sw->n_views + 1, 0);
            me.n_lists = sw->n_views;
            me.rc = ovis_cull_views(sw->shards[r], sw->n_views, sw->view_planes, sw->view_require, me.offsets.data(), me.cap_entities,
                                    me.h_entities, &total);
        } else {
            int32_t n_lights = 0;
            ovis_world_counts(sw->shards[r], nullptr, nullptr, &n_lights);
            const size_t cap_lists = 6 * (size_t)std::max(n_lights, 1);
            me.list_light.resize(cap_lists);
            me.list_view.resize(cap_lists);
            me.offsets.assign(cap_lists + 1, 0);
            me.rc = ovis_build_shadow_lists(sw->shards[r], &sw->main_view, sw->shadow_res, (int32_t)cap_lists, me.list_light.data(),
                                            me.list_view.data(), me.offsets.data(), &me.n_lists, me.cap_entities, me.h_entities, &total);
        }
        me.count = total;
        if (me.rc != OVIS_ERR_CAPACITY || total <= me.cap_entities)
            break;
        if (me.h_entities)
            ovis_host_free(me.h_entities);
        me.h_entities = nullptr;
        me.cap_entities = 0;
        void *p = nullptr;
        const int64_t want = total + total / 4 + 4096;
        if (ovis_host_alloc((int64_t)sizeof(uint32_t) * want, &p) != OVIS_OK) {
            me.rc = OVIS_ERR_NOMEM;
            break;
        }
        me.h_entities = static_cast<uint32_t *>(p);
        me.cap_entities = want;
    }
    if (me.rc != OVIS_OK)
        me.error = ovis_last_error();
}

void worker_loop(OvisShardedWorld *sw, int r) {
    cudaSetDevice(sw->devices[r]);
    const char *pin = getenv("OVIS_PIN_WORKERS");
    if (!pin || atoi(pin) != 0)
        ovis_host_pin_thread_near_device(sw->devices[r], nullptr);   // launches, PCIe stores and the host poll stay on the GPU's socket
    Worker &me = sw->workers[r];
    uint64_t seen = 0;
    while (true) {
        // a frame loop posts jobs every few tens of microseconds: spin briefly, then sleep on the condition variable
        uint64_t posted = me.posted.load(std::memory_order_acquire);
        if (posted == seen) {
            const auto t0 = std::chrono::steady_clock::now();
            while ((posted = me.posted.load(std::memory_order_acquire)) == seen && !sw->quit.load(std::memory_order_relaxed)) {
                if (std::chrono::steady_clock::now() - t0 > std::chrono::microseconds(300)) {
                    std::unique_lock<std::mutex> lock(sw->mutex);
                    sw->wake.wait_for(lock, std::chrono::milliseconds(2), [&] {
                        return me.posted.load(std::memory_order_acquire) != seen || sw->quit.load(std::memory_order_relaxed);
                    });
                }
            }
        }
        if (sw->quit.load(std::memory_order_relaxed))
            return;
        if (sw->kind == kJobCull) {
            uint32_t *seg = sw->h_segments + (int64_t)r * (1 + sw->seg_capacity);
            int64_t count = 0;
            me.rc = ovis_cull_planes(sw->shards[r], sw->planes, sw->require, seg + 1, sw->seg_capacity, &count);
            if (me.rc != OVIS_OK)
                me.error = ovis_last_error();
            me.count = count;
            seg[0] = (uint32_t)count;
        } else {
            many_views_job(sw, r);
        }
        seen = posted;
        me.finished.store(posted, std::memory_order_release);
    }
}

int ensure_segments(OvisShardedWorld *sw, int64_t capacity) {
    if (capacity <= sw->seg_capacity)
        return OVIS_OK;
    if (sw->h_segments)
        cudaFreeHost(sw->h_segments);
    sw->h_segments = nullptr;
    sw->seg_capacity = 0;
    void *p = nullptr;
    cudaError_t err = cudaHostAlloc(&p, sizeof(uint32_t) * (size_t)sw->n * (size_t)(1 + capacity), cudaHostAllocPortable | cudaHostAllocMapped);
    if (err != cudaSuccess)
        return fail(OVIS_ERR_NOMEM, std::string("ovis_sharded: cudaHostAlloc failed: ") + cudaGetErrorString(err));
    sw->h_segments = static_cast<uint32_t *>(p);
    sw->seg_capacity = capacity;
    return OVIS_OK;
}

// one cull of every shard; on return the segments hold [count | ids] and counts[r] the true counts
int run_frame(OvisShardedWorld *sw) {
    sw->job++;
    for (Worker &w : sw->workers)
        w.posted.store(sw->job, std::memory_order_release);
    sw->wake.notify_all();
    for (Worker &w : sw->workers)
        while (w.finished.load(std::memory_order_acquire) != sw->job)
            std::this_thread::yield();
    for (int r = 0; r < sw->n; r++)
        if (sw->workers[r].rc != OVIS_OK && sw->workers[r].rc != OVIS_ERR_CAPACITY)
            return fail(sw->workers[r].rc, "shard " + std::to_string(r) + ": " + sw->workers[r].error);
    return OVIS_OK;
}

}  // namespace

extern "C" {

int ovis_sharded_create(const int *device_ids, int32_t n_devices, OvisShardedWorld **out_world) {
    if (!device_ids || n_devices < 1 || n_devices > 64 || !out_world)
        return fail(OVIS_ERR_INVALID, "ovis_sharded_create: bad arguments (1..64 devices)");
    *out_world = nullptr;
    OvisShardedWorld *sw = new (std::nothrow) OvisShardedWorld;
    if (!sw)
        return fail(OVIS_ERR_NOMEM, "out of host memory");
    sw->n = n_devices;
    sw->devices.assign(device_ids, device_ids + n_devices);
    sw->base.assign((size_t)n_devices + 1, 0);
    sw->workers = std::vector<Worker>((size_t)n_devices);
    for (int r = 0; r < n_devices; r++) {
        OvisWorld *w = nullptr;
        int rc = ovis_world_create(device_ids[r], &w);
        if (rc == OVIS_OK) {
            // its own non-blocking stream: culls of two shards that share a device (tests) must not serialise on stream 0
            cudaStream_t s = nullptr;
            cudaSetDevice(device_ids[r]);
            if (cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking) == cudaSuccess)
                rc = ovis_world_set_stream(w, s);
            sw->streams.push_back(s);
        }
        if (rc != OVIS_OK) {
            std::string why = ovis_last_error();
            for (OvisWorld *made : sw->shards)
                ovis_world_destroy(made);
            delete sw;
            return fail(rc, "ovis_sharded_create: device " + std::to_string(device_ids[r]) + ": " + why);
        }
        sw->shards.push_back(w);
    }
    for (int r = 0; r < n_devices; r++)
        sw->workers[r].thread = std::thread(worker_loop, sw, r);
    *out_world = sw;
    return OVIS_OK;
}

int ovis_sharded_destroy(OvisShardedWorld *sw) {
    if (!sw)
        return OVIS_OK;
    sw->quit.store(true);
    sw->wake.notify_all();
    for (Worker &w : sw->workers)
        if (w.thread.joinable())
            w.thread.join();
    for (size_t r = 0; r < sw->comms.size(); r++)
        if (sw->comms[r])
            sw->nccl.CommDestroy(sw->comms[r]);
    for (int r = 0; r < sw->n; r++) {
        cudaSetDevice(sw->devices[r]);
        if (r < (int)sw->d_send.size() && sw->d_send[r]) cudaFree(sw->d_send[r]);
        if (r < (int)sw->d_recv.size() && sw->d_recv[r]) cudaFree(sw->d_recv[r]);
        ovis_world_destroy(sw->shards[r]);
        if (sw->streams[r])
            cudaStreamDestroy(sw->streams[r]);
    }
    if (sw->h_segments)
        cudaFreeHost(sw->h_segments);
    for (Worker &w : sw->workers)
        if (w.h_entities)
            ovis_host_free(w.h_entities);
    delete sw;
    return OVIS_OK;
}

int ovis_sharded_count(OvisShardedWorld *sw, int32_t *out_shards, int64_t *out_entities) {
    if (!sw)
        return fail(OVIS_ERR_INVALID, "sharded world is null");
    if (out_shards) *out_shards = sw->n;
    if (out_entities) *out_entities = sw->base[(size_t)sw->n];
    return OVIS_OK;
}

int ovis_sharded_shard(OvisShardedWorld *sw, int32_t shard, OvisWorld **out_world, int64_t *out_id_base) {
    if (!sw || shard < 0 || shard >= sw->n || !out_world)
        return fail(OVIS_ERR_INVALID, "ovis_sharded_shard: bad arguments");
    *out_world = sw->shards[(size_t)shard];
    if (out_id_base) *out_id_base = sw->base[(size_t)shard];
    return OVIS_OK;
}

int ovis_sharded_add_entities(OvisShardedWorld *sw, int64_t n, const float *world_aabb, const uint32_t *flags) {
    if (!sw || n < 0 || (n > 0 && !world_aabb))
        return fail(OVIS_ERR_INVALID, "ovis_sharded_add_entities: bad arguments");
    if (sw->base[(size_t)sw->n] != 0)
        return fail(OVIS_ERR_INVALID, "ovis_sharded_add_entities: the world already holds entities (the split into contiguous "
                                      "shards is made once; add to a shard's world directly for more)");
    if (n > (int64_t)0xfffffffe)
        return fail(OVIS_ERR_INVALID, "too many entities");
    for (int r = 0; r < sw->n; r++) {
        const int64_t lo = n * r / sw->n, hi = n * (r + 1) / sw->n;        // [r N / G, (r + 1) N / G): SURVEY.md 8(e)
        sw->base[(size_t)r] = lo;
        sw->base[(size_t)r + 1] = hi;
        uint32_t first = 0;
        int rc = ovis_entity_add(sw->shards[r], hi - lo, world_aabb + 6 * lo, flags ? flags + lo : nullptr, &first);
        if (rc == OVIS_OK)
            rc = ovis_entity_set_world_aabbs(sw->shards[r], first, hi - lo, world_aabb + 6 * lo);
        if (rc == OVIS_OK)
            rc = ovis_world_set_id_base(sw->shards[r], (uint32_t)lo);
        if (rc != OVIS_OK)
            return rc;
    }
    return OVIS_OK;
}

int ovis_sharded_cull_planes(OvisShardedWorld *sw, const float planes[24], uint32_t require, uint32_t *out_entities,
                             int64_t cap_entities, int64_t *out_n_entities, int64_t *out_shard_counts) {
    if (!sw || !planes || !out_n_entities || cap_entities < 0 || (cap_entities > 0 && !out_entities))
        return fail(OVIS_ERR_INVALID, "ovis_sharded_cull_planes: bad arguments");
    int64_t largest = 0;
    for (int r = 0; r < sw->n; r++)
        largest = std::max(largest, sw->base[(size_t)r + 1] - sw->base[(size_t)r]);
    if (sw->seg_capacity == 0) {
        int rc = ensure_segments(sw, std::max<int64_t>(1024, largest / 8));
        if (rc != OVIS_OK)
            return rc;
    }
    std::memcpy(sw->planes, planes, sizeof sw->planes);
    sw->require = require;
    for (int attempt = 0; attempt < 3; attempt++) {
        int rc = run_frame(sw);
        if (rc != OVIS_OK)
            return rc;
        int64_t top = 0;
        for (const Worker &w : sw->workers)
            top = std::max(top, w.count);
        if (top <= sw->seg_capacity)
            break;
        rc = ensure_segments(sw, std::min<int64_t>(largest, top + top / 4 + 1024));   // a list outgrew its segment: redo the frame
        if (rc != OVIS_OK)
            return rc;
    }
    int64_t total = 0;
    for (int r = 0; r < sw->n; r++) {
        if (out_shard_counts)
            out_shard_counts[r] = sw->workers[r].count;
        total += sw->workers[r].count;
    }
    *out_n_entities = total;
    if (total > cap_entities)
        return fail(OVIS_ERR_CAPACITY, "entity result needs " + std::to_string(total) + " entries, buffer holds " + std::to_string(cap_entities));
    int64_t at = 0;
    for (int r = 0; r < sw->n; r++) {                            // shard order = id order = the reference's list order
        std::memcpy(out_entities + at, sw->h_segments + (int64_t)r * (1 + sw->seg_capacity) + 1, sizeof(uint32_t) * (size_t)sw->workers[r].count);
        at += sw->workers[r].count;
    }
    return OVIS_OK;
}

int ovis_sharded_cull(OvisShardedWorld *sw, const OvisView *view, uint32_t require, uint32_t *out_entities, int64_t cap_entities,
                      int64_t *out_n_entities, int64_t *out_shard_counts) {
    if (!view)
        return fail(OVIS_ERR_INVALID, "ovis_sharded_cull: null view");
    float planes[24];
    int rc = ovis_view_planes(view, planes);
    if (rc != OVIS_OK)
        return rc;
    return ovis_sharded_cull_planes(sw, planes, require, out_entities, cap_entities, out_n_entities, out_shard_counts);
}

int ovis_sharded_add_lights(OvisShardedWorld *sw, int32_t n, const OvisLight *lights, uint32_t *out_first_id) {
    if (!sw || n < 0 || (n > 0 && !lights))
        return fail(OVIS_ERR_INVALID, "ovis_sharded_add_lights: bad arguments");
    uint32_t first0 = 0;
    for (int r = 0; r < sw->n; r++) {
        uint32_t first = 0;
        int rc = ovis_light_add(sw->shards[r], n, lights, &first);
        if (rc != OVIS_OK)
            return rc;
        if (r == 0)
            first0 = first;
        else if (first != first0)
            return fail(OVIS_ERR_INVALID, "ovis_sharded_add_lights: the shards' light ids have diverged (lights were added to a shard directly)");
    }
    if (out_first_id)
        *out_first_id = first0;
    return OVIS_OK;
}

// list v of the merged result = list v of shard 0, 1, ... one after the other (contiguous shards, global ascending ids)
static int splice_lists(OvisShardedWorld *sw, const char *who, int32_t n_lists, int64_t *out_offsets, int64_t cap_entities,
                        uint32_t *out_entities, int64_t *out_n_entities) {
    int64_t total = 0;
    for (const Worker &w : sw->workers)
        total += w.count;
    *out_n_entities = total;
    if (total > cap_entities)
        return fail(OVIS_ERR_CAPACITY, std::string(who) + ": the lists need " + std::to_string(total) + " entries, buffer holds " + std::to_string(cap_entities));
    int64_t at = 0;
    for (int32_t v = 0; v < n_lists; v++) {
        out_offsets[v] = at;
        for (const Worker &w : sw->workers) {
            const int64_t lo = w.offsets[(size_t)v], hi = w.offsets[(size_t)v + 1];
            if (hi > lo)
                std::memcpy(out_entities + at, w.h_entities + lo, sizeof(uint32_t) * (size_t)(hi - lo));
            at += hi - lo;
        }
    }
    out_offsets[n_lists] = at;
    return OVIS_OK;
}

int ovis_sharded_cull_views(OvisShardedWorld *sw, int32_t n_views, const float *planes, const uint32_t *require,
                            int64_t *out_offsets, int64_t cap_entities, uint32_t *out_entities, int64_t *out_n_entities) {
    if (!sw || n_views < 0 || (n_views > 0 && !planes) || !out_offsets || !out_n_entities || cap_entities < 0 ||
        (cap_entities > 0 && !out_entities))
        return fail(OVIS_ERR_INVALID, "ovis_sharded_cull_views: bad arguments");
    sw->kind = kJobViews;
    sw->n_views = n_views;
    sw->view_planes = planes;
    sw->view_require = require;
    int rc = run_frame(sw);
    sw->kind = kJobCull;
    if (rc != OVIS_OK)
        return rc;
    return splice_lists(sw, "ovis_sharded_cull_views", n_views, out_offsets, cap_entities, out_entities, out_n_entities);
}

int ovis_sharded_build_shadow_lists(OvisShardedWorld *sw, const OvisView *main_view, int32_t shadow_map_resolution,
                                    int32_t cap_lists, int32_t *out_list_light, int32_t *out_list_view, int64_t *out_offsets,
                                    int32_t *out_n_lists, int64_t cap_entities, uint32_t *out_entities, int64_t *out_n_entities) {
    if (!sw || !main_view || !out_n_lists || !out_n_entities || shadow_map_resolution <= 0 || cap_entities < 0 ||
        (cap_entities > 0 && !out_entities))
        return fail(OVIS_ERR_INVALID, "ovis_sharded_build_shadow_lists: bad arguments");
    sw->kind = kJobShadowLists;
    sw->main_view = *main_view;
    sw->shadow_res = shadow_map_resolution;
    int rc = run_frame(sw);
    sw->kind = kJobCull;
    if (rc != OVIS_OK)
        return rc;
    // the lights are replicated, so every shard names the same lists in the same order
    const Worker &w0 = sw->workers[0];
    for (const Worker &w : sw->workers)
        if (w.n_lists != w0.n_lists || !std::equal(w.list_light.begin(), w.list_light.begin() + w.n_lists, w0.list_light.begin()) ||
            !std::equal(w.list_view.begin(), w.list_view.begin() + w.n_lists, w0.list_view.begin()))
            return fail(OVIS_ERR_INVALID, "ovis_sharded_build_shadow_lists: the shards disagree on the lights (they must hold the same lights: "
                                          "ovis_sharded_add_lights)");
    *out_n_lists = w0.n_lists;
    int64_t total = 0;
    for (const Worker &w : sw->workers)
        total += w.count;
    *out_n_entities = total;
    if (w0.n_lists > cap_lists)
        return fail(OVIS_ERR_CAPACITY, "ovis_sharded_build_shadow_lists: " + std::to_string(w0.n_lists) + " lists, room for " + std::to_string(cap_lists));
    if (!out_offsets || (w0.n_lists > 0 && (!out_list_light || !out_list_view)))
        return fail(OVIS_ERR_INVALID, "ovis_sharded_build_shadow_lists: null output");
    for (int32_t k = 0; k < w0.n_lists; k++) {
        out_list_light[k] = w0.list_light[(size_t)k];
        out_list_view[k] = w0.list_view[(size_t)k];
    }
    return splice_lists(sw, "ovis_sharded_build_shadow_lists", w0.n_lists, out_offsets, cap_entities, out_entities, out_n_entities);
}

int ovis_sharded_segments(OvisShardedWorld *sw, const uint32_t **out_segments, int64_t *out_segment_words) {
    if (!sw || !out_segments || !out_segment_words)
        return fail(OVIS_ERR_INVALID, "ovis_sharded_segments: null argument");
    *out_segments = sw->h_segments;
    *out_segment_words = 1 + sw->seg_capacity;
    return OVIS_OK;
}

int ovis_sharded_allgather_device(OvisShardedWorld *sw, const float planes[24], uint32_t require, int64_t capacity,
                                  const uint32_t *const **out_d_blocks, int64_t *out_shard_counts) {
    if (!sw || !planes || capacity < 1 || !out_d_blocks)
        return fail(OVIS_ERR_INVALID, "ovis_sharded_allgather_device: bad arguments");
    if (!sw->nccl.load())
        return fail(OVIS_ERR_CUDA, "ovis_sharded_allgather_device: libnccl.so.2 could not be loaded");
    if (sw->comms.empty()) {
        sw->comms.assign((size_t)sw->n, nullptr);
        int rc = sw->nccl.CommInitAll(sw->comms.data(), sw->n, sw->devices.data());
        if (rc != 0) {
            sw->comms.clear();
            return fail(OVIS_ERR_CUDA, std::string("ncclCommInitAll failed: ") + (sw->nccl.GetErrorString ? sw->nccl.GetErrorString(rc) : "?") +
                                           " (one communicator needs distinct devices)");
        }
    }
    if (capacity > sw->dev_capacity) {
        sw->d_send.resize((size_t)sw->n, nullptr);
        sw->d_recv.resize((size_t)sw->n, nullptr);
        for (int r = 0; r < sw->n; r++) {
            cudaSetDevice(sw->devices[r]);
            cudaStreamSynchronize(sw->streams[r]);
            if (sw->d_send[r]) cudaFree(sw->d_send[r]);
            if (sw->d_recv[r]) cudaFree(sw->d_recv[r]);
            if (cudaMalloc(&sw->d_send[r], sizeof(uint32_t) * (size_t)(1 + capacity)) != cudaSuccess ||
                cudaMalloc(&sw->d_recv[r], sizeof(uint32_t) * (size_t)sw->n * (size_t)(1 + capacity)) != cudaSuccess)
                return fail(OVIS_ERR_NOMEM, "ovis_sharded_allgather_device: out of device memory");
        }
        sw->dev_capacity = capacity;
    }
    const int64_t words = 1 + sw->dev_capacity;
    // local culls into [count | ids] blocks (the kernel bounds its writes by `capacity`; the count is always the true one)
    for (int r = 0; r < sw->n; r++) {
        uint32_t *dst[1] = { sw->d_send[r] + 1 }, *cnt[1] = { sw->d_send[r] };
        int rc = ovis_cull_planes_device_multi(sw->shards[r], planes, require, sw->dev_capacity, 1, dst, cnt);
        if (rc != OVIS_OK)
            return rc;
    }
    sw->nccl.GroupStart();
    for (int r = 0; r < sw->n; r++)
        sw->nccl.AllGather(sw->d_send[r], sw->d_recv[r], (size_t)words, kNcclUint32, sw->comms[r], sw->streams[r]);
    int rc = sw->nccl.GroupEnd();
    if (rc != 0)
        return fail(OVIS_ERR_CUDA, std::string("ncclAllGather failed: ") + (sw->nccl.GetErrorString ? sw->nccl.GetErrorString(rc) : "?"));
    for (int r = 0; r < sw->n; r++) {
        cudaSetDevice(sw->devices[r]);
        cudaError_t err = cudaStreamSynchronize(sw->streams[r]);
        if (err != cudaSuccess)
            return fail(OVIS_ERR_CUDA, std::string("ovis_sharded_allgather_device: ") + cudaGetErrorString(err));
    }
    if (out_shard_counts) {
        cudaSetDevice(sw->devices[0]);
        for (int r = 0; r < sw->n; r++) {
            uint32_t c = 0;
            cudaMemcpy(&c, sw->d_recv[0] + (int64_t)r * words, sizeof c, cudaMemcpyDeviceToHost);
            out_shard_counts[r] = c;
        }
    }
    *out_d_blocks = sw->d_recv.data();
    return OVIS_OK;
}

}  // extern "C"
