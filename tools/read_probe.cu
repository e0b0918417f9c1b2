// read_probe.cu - what does a bare streaming READ of the cull's input cost on this GPU? (evidence for DESIGN.md section 4)
//
// The entity cull (K3) reads 6 component arrays of N floats once (24 B per entity) and writes almost nothing. Its roofline
// denominator (MEASURED_PEAKS.json) is a copy bandwidth. This probe times kernels that do nothing but read the same bytes,
// with the same access shapes, and fold them into a checksum so that the loads cannot be dropped:
//   one_stream      one array of 6N floats, float4 per lane, grid-stride over a persistent grid
//   six_streams     the K3 shape: 6 arrays, per warp tile of 256 entities every lane loads 2 float4 from each array
//   six_streams_np  the same with one warp tile per warp (non-persistent grid)
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/read_probe tools/read_probe.cu
// Run:   tools/read_probe [entities]     (prints one JSON line)
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

__device__ __forceinline__ float fold(float4 v) { return (v.x + v.y) + (v.z + v.w); }

__global__ void __launch_bounds__(256, 2) one_stream(const float4 *__restrict__ p, long long n4, float *out) {
    float acc = 0.0f;
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n4; i += (long long)gridDim.x * 256)
        acc += fold(__ldg(p + i));
    if (acc == 123.456f)
        *out = acc;
}

// K3's walk: warp tiles of 256 entities (64 float4 per array), lane L takes float4 L and L + 32 of the tile, 6 arrays
__global__ void __launch_bounds__(256, 2) six_streams(const float4 *__restrict__ p, long long stride4, int num_tiles, float *out) {
    const int lane = threadIdx.x & 31;
    const int warps_total = gridDim.x * 8, gwarp = blockIdx.x * 8 + (threadIdx.x >> 5);
    float acc = 0.0f;
    for (int t = gwarp; t < num_tiles; t += warps_total) {
        float4 v[12];
#pragma unroll
        for (int g = 0; g < 2; g++)
#pragma unroll
            for (int c = 0; c < 6; c++)
                v[g * 6 + c] = __ldg(p + c * stride4 + (long long)t * 64 + g * 32 + lane);
#pragma unroll
        for (int k = 0; k < 12; k++)
            acc += fold(v[k]);
    }
    if (acc == 123.456f)
        *out = acc;
}

__global__ void __launch_bounds__(256) six_streams_np(const float4 *__restrict__ p, long long stride4, int num_tiles, float *out) {
    const int lane = threadIdx.x & 31;
    const int t = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (t >= num_tiles)
        return;
    float acc = 0.0f;
#pragma unroll
    for (int g = 0; g < 2; g++)
#pragma unroll
        for (int c = 0; c < 6; c++)
            acc += fold(__ldg(p + c * stride4 + (long long)t * 64 + g * 32 + lane));
    if (acc == 123.456f)
        *out = acc;
}

template <typename F>
static float time_us(F launch, int iters) {
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    for (int i = 0; i < 3; i++)
        launch();
    cudaEventRecord(a);
    for (int i = 0; i < iters; i++)
        launch();
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms = 0.0f;
    cudaEventElapsedTime(&ms, a, b);
    return ms * 1000.0f / iters;
}

int main(int argc, char **argv) {
    const long long n = argc > 1 ? atoll(argv[1]) : 12500000ll;
    const long long stride = (n + 1023) / 1024 * 1024;
    float *d = nullptr, *out = nullptr;
    CK(cudaMalloc(&d, sizeof(float) * 6 * stride));
    CK(cudaMalloc(&out, 4));
    CK(cudaMemset(d, 0, sizeof(float) * 6 * stride));
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    const int tiles = (int)((n + 255) / 256);
    const double bytes = 24.0 * n;
    const int iters = 200;
    float t1 = time_us([&] { one_stream<<<sms * 2, 256>>>((const float4 *)d, 6 * stride / 4, out); }, iters);
    float t1b = time_us([&] { one_stream<<<sms * 8, 256>>>((const float4 *)d, 6 * stride / 4, out); }, iters);
    float t6 = time_us([&] { six_streams<<<sms * 2, 256>>>((const float4 *)d, stride / 4, tiles, out); }, iters);
    float t6n = time_us([&] { six_streams_np<<<(tiles + 7) / 8, 256>>>((const float4 *)d, stride / 4, tiles, out); }, iters);
    CK(cudaDeviceSynchronize());
    printf("{\"entities\": %lld, \"bytes\": %.0f, \"one_stream_2cta_us\": %.2f, \"one_stream_8cta_us\": %.2f, \"six_streams_persistent_us\": %.2f, "
           "\"six_streams_one_tile_per_warp_us\": %.2f, \"gbs\": [%.0f, %.0f, %.0f, %.0f]}\n",
           n, bytes, t1, t1b, t6, t6n, bytes / t1 / 1e3, bytes / t1b / 1e3, bytes / t6 / 1e3, bytes / t6n / 1e3);
    return 0;
}
