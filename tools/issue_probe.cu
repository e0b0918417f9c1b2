// issue_probe.cu - the fp32 issue-rate peak WITHOUT fused multiply-add, the roofline denominator of the many-view culls.
//
// The plane tests may not use FMA (bit-exactness with the reference's x86-64 build), so their ceiling is not the
// advertised FMA rate but how many FADD / FMUL / FMNMX / compare warp-instructions the SMs can issue. This probe times
// long unrolled runs of independent chains, one kernel per instruction mix:
//   fadd_fmul    FADD and FMUL alternating (the FMA pipe)
//   fmul2_fadd   packed FMUL2 + 2 scalar FADD (the cull's products and sums)
//   fmnmx_cmp    FMNMX + FSETP/SEL (the half-rate ALU pipe)
//   cull_mix     per 4 entities and plane: 6 FMUL2, 12 FADD, 2 FMNMX3-equivalents (the K3 / K4 inner loop's mix)
// Output: one JSON line with warp-instructions per second for each mix and the nominal 148 SMs x 4 schedulers x clock.
// Build: nvcc -O3 -fmad=false -gencode arch=compute_100a,code=sm_100a -o tools/issue_probe tools/issue_probe.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

constexpr int kChains = 8;
constexpr int kIters = 2048;

__device__ __forceinline__ void mul2(float a0, float a1, float n, float &r0, float &r1) {
    unsigned long long a, b, d;
    asm volatile("mov.b64 %0, {%1, %2};" : "=l"(a) : "f"(a0), "f"(a1));
    asm volatile("mov.b64 %0, {%1, %2};" : "=l"(b) : "f"(n), "f"(n));
    asm volatile("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    asm volatile("mov.b64 {%0, %1}, %2;" : "=f"(r0), "=f"(r1) : "l"(d));
}

__global__ void __launch_bounds__(256) fadd_fmul(float *out, float s) {
    float v[kChains];
#pragma unroll
    for (int k = 0; k < kChains; k++) v[k] = threadIdx.x * 0.001f + k;
    for (int i = 0; i < kIters; i++) {
#pragma unroll
        for (int k = 0; k < kChains; k++) v[k] = __fadd_rn(v[k], s);
#pragma unroll
        for (int k = 0; k < kChains; k++) v[k] = __fmul_rn(v[k], s);
    }
    float acc = 0;
#pragma unroll
    for (int k = 0; k < kChains; k++) acc += v[k];
    if (acc == 123.456f) *out = acc;
}

__global__ void __launch_bounds__(256) fmul2_fadd(float *out, float s) {
    float v[kChains];
#pragma unroll
    for (int k = 0; k < kChains; k++) v[k] = threadIdx.x * 0.001f + k;
    for (int i = 0; i < kIters; i++) {
#pragma unroll
        for (int k = 0; k < kChains; k += 2) mul2(v[k], v[k + 1], s, v[k], v[k + 1]);
#pragma unroll
        for (int k = 0; k < kChains; k++) v[k] = __fadd_rn(v[k], s);
    }
    float acc = 0;
#pragma unroll
    for (int k = 0; k < kChains; k++) acc += v[k];
    if (acc == 123.456f) *out = acc;
}

__global__ void __launch_bounds__(256) fmnmx_cmp(float *out, float s) {
    float v[kChains];
    unsigned bits = 0;
#pragma unroll
    for (int k = 0; k < kChains; k++) v[k] = threadIdx.x * 0.001f + k;
    for (int i = 0; i < kIters; i++) {
#pragma unroll
        for (int k = 0; k < kChains; k++) v[k] = fminf(v[k], s + i);
#pragma unroll
        for (int k = 0; k < kChains; k++) bits += v[k] < (float)i ? 1u : 0u;
    }
    float acc = (float)bits;
#pragma unroll
    for (int k = 0; k < kChains; k++) acc += v[k];
    if (acc == 123.456f) *out = acc;
}

// 4 entities x 1 plane per step: 6 FMUL2 + 12 FADD + 4 fminf(fminf()) (folded to FMNMX3 by ptxas)
__global__ void __launch_bounds__(256) cull_mix(float *out, float nx, float ny, float nz, float d) {
    float px[4], py[4], pz[4], low[4];
#pragma unroll
    for (int k = 0; k < 4; k++) { px[k] = threadIdx.x + k; py[k] = k * 0.5f; pz[k] = -k; low[k] = 1e30f; }
    for (int i = 0; i < kIters; i++) {
        float x[4], y[4], z[4], dist[4];
        mul2(px[0], px[1], nx, x[0], x[1]); mul2(px[2], px[3], nx, x[2], x[3]);
        mul2(py[0], py[1], ny, y[0], y[1]); mul2(py[2], py[3], ny, y[2], y[3]);
        mul2(pz[0], pz[1], nz, z[0], z[1]); mul2(pz[2], pz[3], nz, z[2], z[3]);
#pragma unroll
        for (int k = 0; k < 4; k++) {
            dist[k] = __fsub_rn(__fadd_rn(__fadd_rn(x[k], y[k]), z[k]), d);
            low[k] = fminf(fminf(low[k], dist[k]), px[k]);
            px[k] = __fadd_rn(px[k], 1.0f);                    // next step's operands differ (1 more FADD per entity)
        }
    }
    float acc = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) acc += low[k];
    if (acc == 123.456f) *out = acc;
}

template <typename F>
static double time_s(F launch) {
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    launch();
    cudaEventRecord(a);
    for (int i = 0; i < 5; i++) launch();
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    return ms * 1e-3 / 5;
}

int main() {
    int sms = 0, khz = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    CK(cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0));
    float *out = nullptr;
    CK(cudaMalloc(&out, 4));
    const int grid = sms * 8, threads = 256;
    const double warps = (double)grid * threads / 32;
    const double t1 = time_s([&] { fadd_fmul<<<grid, threads>>>(out, 1.0001f); });
    const double t2 = time_s([&] { fmul2_fadd<<<grid, threads>>>(out, 1.0001f); });
    const double t3 = time_s([&] { fmnmx_cmp<<<grid, threads>>>(out, 0.5f); });
    const double t4 = time_s([&] { cull_mix<<<grid, threads>>>(out, 0.3f, -0.5f, 0.8f, 2.0f); });
    CK(cudaDeviceSynchronize());
    const double i1 = warps * kIters * 2.0 * kChains;                    // FADD + FMUL per chain
    const double i2 = warps * kIters * (kChains / 2 + kChains);          // FMUL2 per pair + FADD per chain
    const double i3 = warps * kIters * 3.0 * kChains;                    // FMNMX + FSETP + SEL/IADD per chain (approx.)
    const double i4 = warps * kIters * (6 + 12 + 4 + 4);                 // FMUL2, FADD (incl. the sub), FMNMX3, FADD
    printf("{\"sms\": %d, \"clock_mhz\": %.0f, \"nominal_warp_inst_per_s\": %.4g, \"fadd_fmul_warp_inst_per_s\": %.4g, "
           "\"fmul2_fadd_warp_inst_per_s\": %.4g, \"fmnmx_cmp_warp_inst_per_s\": %.4g, \"cull_mix_warp_inst_per_s\": %.4g, "
           "\"cull_mix_plane_tests_per_s\": %.4g, \"note\": \"no FMA issued (-fmad=false, rn intrinsics); cull_mix = 6 FMUL2 + 16 FADD + 4 FMNMX3 "
           "per 4 plane tests\"}\n",
           sms, khz / 1e3, sms * 4.0 * khz * 1e3, i1 / t1, i2 / t2, i3 / t3, i4 / t4, warps * 32 * kIters * 4.0 / t4);
    return 0;
}
