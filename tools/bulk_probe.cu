// bulk_probe.cu - how fast can cp.async.bulk (TMA, non-tensor form) stream the cull's input into shared memory?
//
// The entity cull reads six SoA float arrays once. tools/read_probe.cu gives the ceiling for plain 128-bit loads (45 us for
// 12.5 M entities). This probe streams the same bytes with bulk copies into shared-memory rings and folds them into a
// checksum from shared memory, to find the ring shape (copy size, stages, who issues) that reaches that ceiling:
//   warp_ring<ROW, STAGES>   every warp owns a ring; lane 0 issues six ROW-byte copies per tile (ROW / 4 entities)
//   cta_ring<TILE, STAGES>   one producer warp per CTA issues six copies of TILE * 4 bytes per stage; 8 consumer warps
//                            each read their 1/8 of the stage; full / empty mbarriers, no __syncthreads
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/bulk_probe tools/bulk_probe.cu
// Run:   tools/bulk_probe [entities]
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

__device__ __forceinline__ uint32_t smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t n) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(n) : "memory"); }
__device__ __forceinline__ void mbar_expect(uint32_t bar, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_arrive(uint32_t bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    } while (!done);
}
__device__ __forceinline__ void bulk(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ float fold(float4 v) { return (v.x + v.y) + (v.z + v.w); }

template <int ROW, int STAGES>
__global__ void __launch_bounds__(256) warp_ring(const float *__restrict__ p, long long stride, int num_tiles, float *out) {
    extern __shared__ __align__(128) unsigned char smem[];
    constexpr int kStage = 6 * ROW, kWarp = STAGES * kStage, kEnt = ROW / 4;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int warps_total = gridDim.x * 8;
    unsigned char *ring = smem + warp * kWarp;
    const uint32_t ring_a = smem_addr(ring), bar0 = smem_addr(smem + 8 * kWarp) + warp * STAGES * 8;
    if (lane == 0) {
        for (int s = 0; s < STAGES; s++) mbar_init(bar0 + 8 * s, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    auto request = [&](int t, int s) {
        if (t >= num_tiles) return;
        mbar_expect(bar0 + 8 * s, kStage);
#pragma unroll
        for (int c = 0; c < 6; c++) bulk(ring_a + s * kStage + c * ROW, p + c * stride + (long long)t * kEnt, ROW, bar0 + 8 * s);
    };
    int t = blockIdx.x * 8 + warp;
    if (lane == 0)
        for (int s = 0; s < STAGES; s++) request(t + s * warps_total, s);
    int stage = 0; uint32_t par = 0; float acc = 0.0f;
    for (; t < num_tiles; t += warps_total) {
        mbar_wait(bar0 + 8 * stage, par);
        const float4 *rows = reinterpret_cast<const float4 *>(ring + stage * kStage);
#pragma unroll
        for (int i = lane; i < kStage / 16; i += 32) acc += fold(rows[i]);
        __syncwarp();
        if (lane == 0) request(t + STAGES * warps_total, stage);
        if (++stage == STAGES) { stage = 0; par ^= 1; }
    }
    if (acc == 123.456f) *out = acc;
}

// TILE entities per stage (multiple of 1024): 8 consumer warps x TILE / 8 entities each, warp 8 = producer
template <int TILE, int STAGES>
__global__ void __launch_bounds__(288) cta_ring(const float *__restrict__ p, long long stride, int num_tiles, float *out, int delay, unsigned *bits) {
    extern __shared__ __align__(128) unsigned char smem[];
    constexpr int kRow = TILE * 4, kStage = 6 * kRow;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t ring_a = smem_addr(smem), full0 = smem_addr(smem + STAGES * kStage), empty0 = full0 + 8 * STAGES;
    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; s++) { mbar_init(full0 + 8 * s, 1); mbar_init(empty0 + 8 * s, 8); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    float acc = 0.0f;
    if (warp == 8) {
        if (lane == 0) {
            int stage = 0; uint32_t par = 0;
            for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
                mbar_wait(empty0 + 8 * stage, par ^ 1u);                       // first round passes at once
                mbar_expect(full0 + 8 * stage, kStage);
#pragma unroll
                for (int c = 0; c < 6; c++) bulk(ring_a + stage * kStage + c * kRow, p + c * stride + (long long)t * TILE, kRow, full0 + 8 * stage);
                if (++stage == STAGES) { stage = 0; par ^= 1; }
            }
        }
    } else {
        int stage = 0; uint32_t par = 0;
        for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
            mbar_wait(full0 + 8 * stage, par);
            const unsigned char *st = smem + stage * kStage;
#pragma unroll
            for (int c = 0; c < 6; c++) {
                const float4 *row = reinterpret_cast<const float4 *>(st + c * kRow) + warp * (TILE / 32);
#pragma unroll
                for (int i = lane; i < TILE / 32; i += 32) acc += fold(row[i]);
            }
            if (delay) {                                      // stand-in for the plane tests: hold the stage this long
                const long long t0 = clock64();
                while (clock64() - t0 < delay) {}
            }
            if (bits && (lane & 7) == 0)                      // the cull's output: 16 bytes per 128 entities
                bits[((long long)t * (TILE / 128) + warp) * 4 + (lane >> 3)] = __float_as_uint(acc);
            __syncwarp();
            if (lane == 0) mbar_arrive(empty0 + 8 * stage);
            if (++stage == STAGES) { stage = 0; par ^= 1; }
        }
    }
    if (acc == 123.456f) *out = acc;
}

template <typename F>
static float time_us(F launch, int iters) {
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    for (int i = 0; i < 3; i++) launch();
    cudaEventRecord(a);
    for (int i = 0; i < iters; i++) launch();
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms = 0.0f;
    cudaEventElapsedTime(&ms, a, b);
    return ms * 1000.0f / iters;
}

static float *g_d, *g_out;
static long long g_stride, g_n;
static int g_sms;

template <int ROW, int STAGES>
static void run_warp(int ctas_per_sm) {
    const int smem = 8 * STAGES * 6 * ROW + 8 * STAGES * 8;
    if (smem * ctas_per_sm > 225 * 1024) return;
    cudaFuncSetAttribute(warp_ring<ROW, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    const int tiles = (int)((g_n + ROW / 4 - 1) / (ROW / 4));
    float us = time_us([&] { warp_ring<ROW, STAGES><<<g_sms * ctas_per_sm, 256, smem>>>(g_d, g_stride, tiles, g_out); }, 100);
    cudaError_t e = cudaDeviceSynchronize();
    printf("warp_ring row=%4d stages=%d ctas/sm=%d smem=%6d : %7.2f us  %6.0f GB/s %s\n", ROW, STAGES, ctas_per_sm, smem, us, 24.0 * g_n / us / 1e3,
           e == cudaSuccess ? "" : cudaGetErrorString(e));
}

static unsigned *g_bits;
template <int TILE, int STAGES>
static void run_cta(int ctas_per_sm, int delay = 0, bool store = false) {
    const int smem = STAGES * 6 * TILE * 4 + 16 * STAGES;
    if (smem * ctas_per_sm > 225 * 1024) return;
    cudaFuncSetAttribute(cta_ring<TILE, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    const int tiles = (int)((g_n + TILE - 1) / TILE);
    float us = time_us([&] { cta_ring<TILE, STAGES><<<g_sms * ctas_per_sm, 288, smem>>>(g_d, g_stride, tiles, g_out, delay, store ? g_bits : nullptr); }, 100);
    cudaError_t e = cudaDeviceSynchronize();
    printf("cta_ring tile=%4d stages=%d ctas/sm=%d smem=%6d delay=%4d store=%d : %7.2f us  %6.0f GB/s %s\n", TILE, STAGES, ctas_per_sm, smem, delay, (int)store, us, 24.0 * g_n / us / 1e3,
           e == cudaSuccess ? "" : cudaGetErrorString(e));
}

int main(int argc, char **argv) {
    g_n = argc > 1 ? atoll(argv[1]) : 12500000ll;
    g_stride = (g_n + 4095) / 4096 * 4096;
    CK(cudaMalloc(&g_d, sizeof(float) * 6 * g_stride));
    CK(cudaMalloc(&g_out, 4));
    CK(cudaMemset(g_d, 0, sizeof(float) * 6 * g_stride));
    CK(cudaDeviceGetAttribute(&g_sms, cudaDevAttrMultiProcessorCount, 0));
    CK(cudaMalloc(&g_bits, g_stride / 8 + 4096));
    run_warp<512, 2>(4); run_warp<512, 3>(3); run_warp<512, 4>(2); run_warp<512, 4>(1);
    run_warp<1024, 2>(2); run_warp<1024, 3>(1); run_warp<1024, 4>(1);
    run_warp<2048, 2>(1); run_warp<2048, 2>(2); run_warp<4096, 2>(1);
    run_cta<1024, 2>(4); run_cta<1024, 3>(3); run_cta<1024, 4>(2); run_cta<1024, 8>(1);
    run_cta<2048, 2>(2); run_cta<2048, 3>(1); run_cta<2048, 4>(1); run_cta<4096, 2>(1);
    // the cull's shape (1024-entity stages, 4 stages, 2 CTAs / SM) with the consumers holding a stage for `delay` cycles
    for (int delay : { 0, 500, 1000, 1500, 2000, 3000 }) run_cta<1024, 4>(2, delay, false);
    run_cta<1024, 4>(2, 0, true); run_cta<1024, 4>(2, 1000, true);
    run_cta<1024, 8>(1, 1000, false); run_cta<1024, 8>(1, 2000, false); run_cta<1024, 3>(3, 1000, false); run_cta<1024, 3>(3, 2000, false);
    CK(cudaDeviceSynchronize());
    return 0;
}
