// ovis_noalloc.cu - ovis_cull_planes_noalloc: the single-view entity cull for callers that must not
// allocate GPU memory.
//
// Who needs this: a renderer calling the cull from inside its frame loop. There, cudaMalloc / cudaFree
// are off limits - they synchronise the whole device (a stall in the middle of a frame), they can fail
// when the graphics API has already taken the VRAM budget, and many engines simply forbid driver
// allocations after start-up and hand out all device memory from their own pool. Such a caller owns its
// bounds and result buffers already; the only thing the cull needs beyond them is scratch space for the
// survivor bitmap (one bit per entity) and the per-block totals. Here that scratch is a fixed array that is part of the library
// image, so the call path contains no cudaMalloc, cudaFree, or any allocating runtime/library routine.
//
// The price is a hard size limit (OVIS_NOALLOC_MAX_ENTITIES) and one shared scratch per GPU:
//   - every GPU that loads liborion_vis.so carries the ~130 KB array until the process exits;
//   - all calls on a GPU use the same array. It still holds the previous call's data, so each call clears
//     the one part that has to start at zero (the block totals); the rest is overwritten before it is read;
//   - two calls whose kernels overlap on the same GPU (different streams, or different host threads) would
//     write the same scratch and race. Issue them on one stream, or serialise them yourself.
#include "orion_vis.h"

#include <cstdio>
#include <mutex>

#include <cuda_runtime.h>

#include "ovis_kernels.h"

namespace ovis {
int report_error(int code, const char *message);
}

namespace {

constexpr int64_t kMaxEntities = OVIS_NOALLOC_MAX_ENTITIES;
constexpr int64_t kStrideTile = 256;                            // the documented stride rule: whole 256-entity tiles
constexpr int64_t kMaxBlocks = kMaxEntities / ovis::kMaskBlockEntities + 1;

// Scratch layout inside the array, in bytes: the survivor bitmap (one bit per entity slot) and two sets of block totals.
constexpr size_t kBitmapBytes = sizeof(uint32_t) * (kMaxEntities / 32);
constexpr size_t kTotalsBytes = sizeof(uint32_t) * 2 * kMaxBlocks;
constexpr size_t kScratchBytes = kBitmapBytes + ((kTotalsBytes + 255) / 256) * 256;

// The scratch itself. In host code this name is only a handle for the runtime, NOT a device address: it must
// never be handed to a kernel. The address kernels may use comes from cudaGetSymbolAddress (below).
static __device__ __align__(256) unsigned char g_noalloc_scratch[kScratchBytes];

struct DeviceSetup {
    bool ready = false;
    unsigned char *scratch = nullptr;     // device address of g_noalloc_scratch on this GPU
    int num_sms = 0;
    int variant = 0;                      // what a world holding kMaxEntities or fewer would use
    int blocks_per_sm = 0;
};

constexpr int kMaxDevices = 64;
DeviceSetup g_setup[kMaxDevices];
std::mutex g_setup_mutex;

int fail_cuda(const char *what, cudaError_t err) {
    char buf[256];
    snprintf(buf, sizeof buf, "ovis_cull_planes_noalloc: %s failed: %s", what, cudaGetErrorString(err));
    return ovis::report_error(OVIS_ERR_CUDA, buf);
}

// One-time set-up for `device`, which must be the current device: resolve the scratch array's device address
// and cache the launch geometry. None of these calls allocates.
int setup_device(int device, DeviceSetup **out) {
    std::lock_guard<std::mutex> lock(g_setup_mutex);
    DeviceSetup &s = g_setup[device];
    if (!s.ready) {
        void *ptr = nullptr;
        cudaError_t err = cudaGetSymbolAddress(&ptr, g_noalloc_scratch);
        if (err != cudaSuccess)
            return fail_cuda("cudaGetSymbolAddress(scratch)", err);
        s.scratch = static_cast<unsigned char *>(ptr);
        err = cudaDeviceGetAttribute(&s.num_sms, cudaDevAttrMultiProcessorCount, device);
        if (err != cudaSuccess)
            return fail_cuda("cudaDeviceGetAttribute", err);
        s.variant = 1;                                // register-fed, 128-entity warp tiles: what a world of this size uses
        s.blocks_per_sm = ovis::cull_test_max_blocks_per_sm(false, s.variant);
        if (s.blocks_per_sm < 1)
            s.blocks_per_sm = 1;
        s.ready = true;
    }
    *out = &s;
    return OVIS_OK;
}

}  // namespace

extern "C" int ovis_cull_planes_noalloc(int device, void *cuda_stream, const float *d_world_aabb, int64_t stride,
                                        int64_t n_entities, const float planes[24], uint32_t id_base,
                                        uint32_t *d_out_entities, int64_t capacity, uint32_t *d_out_count) {
    if (!planes || !d_out_count || n_entities < 0 || capacity < 0 || (n_entities > 0 && (!d_world_aabb || !d_out_entities)))
        return ovis::report_error(OVIS_ERR_INVALID, "ovis_cull_planes_noalloc: null or negative argument");
    if (n_entities > kMaxEntities) {
        char buf[200];
        snprintf(buf, sizeof buf, "ovis_cull_planes_noalloc: %lld entities exceed the fixed scratch (at most %lld per call); "
                                  "use a world (ovis_cull_planes_device) for larger sets",
                 (long long)n_entities, (long long)kMaxEntities);
        return ovis::report_error(OVIS_ERR_CAPACITY, buf);
    }
    const int64_t padded = (n_entities + kStrideTile - 1) / kStrideTile * kStrideTile;
    if (stride % 4 != 0 || stride < padded)
        return ovis::report_error(OVIS_ERR_INVALID, "ovis_cull_planes_noalloc: stride must be a multiple of 4 and at least "
                                                    "n_entities rounded up to 256 (whole warp tiles are loaded)");
    // The kernels read the bounds as 128-bit vectors: a pointer that is not 16-byte aligned would fault on the device and
    // take the whole CUDA context with it, so it is refused here.
    if (reinterpret_cast<uintptr_t>(d_world_aabb) & 15u)
        return ovis::report_error(OVIS_ERR_INVALID, "ovis_cull_planes_noalloc: d_world_aabb must be 16-byte aligned");
    if ((reinterpret_cast<uintptr_t>(d_out_entities) | reinterpret_cast<uintptr_t>(d_out_count)) & 3u)
        return ovis::report_error(OVIS_ERR_INVALID, "ovis_cull_planes_noalloc: the outputs must be 4-byte aligned");
    if (device < 0 || device >= kMaxDevices)
        return ovis::report_error(OVIS_ERR_INVALID, "ovis_cull_planes_noalloc: device out of range");
    cudaError_t err = cudaSetDevice(device);
    if (err != cudaSuccess)
        return fail_cuda("cudaSetDevice", err);
    DeviceSetup *setup = nullptr;
    int rc = setup_device(device, &setup);
    if (rc != OVIS_OK)
        return rc;
    cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);

    if (n_entities == 0) {
        ovis::launch_empty_result(nullptr, nullptr, d_out_count, stream);
        err = cudaGetLastError();
        return err == cudaSuccess ? OVIS_OK : fail_cuda("empty-result launch", err);
    }

    uint32_t *survivors = reinterpret_cast<uint32_t *>(setup->scratch);
    uint32_t *totals = reinterpret_cast<uint32_t *>(setup->scratch + kBitmapBytes);

    ovis::CullArgs a{};
    for (int i = 0; i < 6; i++)
        a.planes[i] = make_float4(planes[4 * i], planes[4 * i + 1], planes[4 * i + 2], planes[4 * i + 3]);
    a.bounds = d_world_aabb;
    a.stride = stride;
    a.n = n_entities;
    a.sign_codes = ovis::plane_sign_codes(planes);
    a.num_groups = (int)((n_entities + 127) / 128);
    a.num_blocks = (int)((n_entities + ovis::kMaskBlockEntities - 1) / ovis::kMaskBlockEntities);
    a.mask = nullptr;
    a.survivors = survivors;
    a.block_totals = totals;                       // accumulated into by this call: must start at zero
    a.block_totals_next = totals + kMaxBlocks;     // the kernel's "re-arm the other set" write lands here, unused

    // The array keeps whatever the previous call left. Only the block totals are added to rather than
    // overwritten, so they are the only part that has to be cleared.
    err = cudaMemsetAsync(totals, 0, sizeof(uint32_t) * (size_t)a.num_blocks, stream);
    if (err != cudaSuccess)
        return fail_cuda("cudaMemsetAsync(block totals)", err);

    const int grid = ovis::cull_test_grid(setup->variant, a.num_groups, setup->num_sms * setup->blocks_per_sm);
    // one shared scratch: this call's test kernel must not overlap the previous call's gather kernel
    ovis::launch_cull_test(a, false, setup->variant, grid, stream, /*overlap_previous_gather=*/false);

    ovis::GatherArgs g{};
    g.survivors = survivors;
    g.block_totals = totals;
    g.num_words = (int64_t)a.num_groups * 4;
    g.num_blocks = a.num_blocks;
    g.id_base = id_base;
    g.out = d_out_entities;
    g.capacity = capacity;
    g.d_base = nullptr;
    g.d_next = nullptr;
    g.d_count = d_out_count;
    g.n_extra = 0;
    g.n_wait = 0;
    g.n_signal = 0;
    g.done_counter = nullptr;
    g.error_flag = nullptr;
    ovis::launch_cull_gather(g, stream);
    err = cudaGetLastError();
    if (err != cudaSuccess)
        return fail_cuda("kernel launch", err);
    return OVIS_OK;
}
