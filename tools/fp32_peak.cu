// tools/fp32_peak.cu — measures the FP32 pipe of the GPU it runs on: scalar FFMA, packed FFMA2
// (fma.rn.f32x2), FFMA2 with one LDS.64 per 4 (k_convolve's mix), and F2F.F64.F32 + DADD beside FFMA2.
// Diagnostic only (sizes the k_convolve roofline in DESIGN.md / profiles/r02_conv_arith.md).
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o build/fp32_peak tools/fp32_peak.cu
#include <cstdio>
#include <cuda_runtime.h>

typedef unsigned long long u64;
__device__ __forceinline__ u64 pack2(float a, float b) { u64 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void unpack2(u64 v, float &a, float &b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c) { u64 d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }

#define NACC 16
__global__ void k_ffma(float *out, int iters, float a, float b) {
    float acc[NACC];
#pragma unroll
    for (int i = 0; i < NACC; i++) acc[i] = threadIdx.x + i;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < NACC; i++) acc[i] = fmaf(acc[i], a, b);
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < NACC; i++) s += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_ffma2(float *out, int iters, float a, float b) {
    u64 acc[NACC];
    const u64 pa = pack2(a, a * 0.5f), pb = pack2(b, b * 0.25f);
#pragma unroll
    for (int i = 0; i < NACC; i++) acc[i] = pack2(threadIdx.x + i, i);
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < NACC; i++) acc[i] = ffma2(acc[i], pa, pb);
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < NACC; i++) { float x, y; unpack2(acc[i], x, y); s += x + y; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// 28 FFMA2 per 7 LDS.64: the ratio of k_convolve's inner loop (147 : 34)
__global__ void k_ffma2_lds(float *out, int iters) {
    __shared__ u64 sm[2048];
    for (int i = threadIdx.x; i < 2048; i += blockDim.x) sm[i] = pack2(1.0f + i * 1e-6f, 1e-3f);
    __syncthreads();
    u64 acc[NACC];
#pragma unroll
    for (int i = 0; i < NACC; i++) acc[i] = pack2(threadIdx.x + i, i);
    int idx = threadIdx.x;
    for (int it = 0; it < iters; it++) {
        u64 w[7];
#pragma unroll
        for (int q = 0; q < 7; q++) w[q] = sm[(idx + q) & 2047];
#pragma unroll
        for (int q = 0; q < 28; q++) acc[q & 15] = ffma2(w[q % 7], w[(q + 3) % 7], acc[q & 15]);
        idx += 33;
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < NACC; i++) { float x, y; unpack2(acc[i], x, y); s += x + y; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// 21 FFMA2 + 1 F2F.F64.F32 + 1 DADD + 1 FADD: one kernel row of k_convolve with its fold
__global__ void k_ffma2_fold(float *out, int iters, float a, float b) {
    u64 acc[NACC]; double tot[NACC]; float tv[NACC];
    const u64 pa = pack2(a, a * 0.5f), pb = pack2(b, b * 0.25f);
#pragma unroll
    for (int i = 0; i < NACC; i++) { acc[i] = pack2(threadIdx.x + i, i); tot[i] = 0; tv[i] = 0; }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < 21; r++)
#pragma unroll
            for (int i = 0; i < NACC; i++) acc[i] = ffma2(acc[i], pa, pb);
#pragma unroll
        for (int i = 0; i < NACC; i++) { float x, y; unpack2(acc[i], x, y); tot[i] += (double)x; tv[i] += y; }
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < NACC; i++) s += tot[i] + tv[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = (float)s;
}

template <typename F> static float time_ms(F f) {
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    f(); cudaDeviceSynchronize();
    cudaEventRecord(a); f(); cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b); return ms;
}

int main() {
    cudaDeviceProp pr; cudaGetDeviceProperties(&pr, 0);
    const int sms = pr.multiProcessorCount; int khz = 0; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const double clk = khz * 1e3;
    printf("%s: %d SMs, max clock %.0f MHz\n", pr.name, sms, clk / 1e6);
    float *out; cudaMalloc(&out, sizeof(float) * sms * 8 * 1024);
    const int iters = 20000;
    for (int wps = 4; wps <= 32; wps *= 2) {
        const int thr = wps * 32 > 1024 ? 1024 : wps * 32, blocks = sms * (wps * 32 / thr);
        float ms = time_ms([&] { k_ffma<<<blocks, thr>>>(out, iters, 1.0001f, 0.5f); });
        double n = (double)blocks * thr * iters * NACC;
        printf("FFMA        %2d warps/SM: %.2f ms  %.2f T FMA/s = %.1f /clk/SM\n", wps, ms, n / ms / 1e9, n / (ms * 1e-3) / sms / clk);
        ms = time_ms([&] { k_ffma2<<<blocks, thr>>>(out, iters, 1.0001f, 0.5f); });
        printf("FFMA2       %2d warps/SM: %.2f ms  %.2f T FMA/s = %.1f /clk/SM (2 per instruction lane)\n", wps, ms, 2 * n / ms / 1e9, 2 * n / (ms * 1e-3) / sms / clk);
        ms = time_ms([&] { k_ffma2_lds<<<blocks, thr>>>(out, iters); });
        n = (double)blocks * thr * iters * 28;
        printf("FFMA2+LDS64 %2d warps/SM: %.2f ms  %.2f T FMA/s = %.1f /clk/SM (28 FFMA2 : 7 LDS.64)\n", wps, ms, 2 * n / ms / 1e9, 2 * n / (ms * 1e-3) / sms / clk);
        ms = time_ms([&] { k_ffma2_fold<<<blocks, thr>>>(out, iters / 20, 1.0001f, 0.5f); });
        n = (double)blocks * thr * (iters / 20) * NACC * 21;
        printf("FFMA2+fold  %2d warps/SM: %.2f ms  %.2f T FMA/s = %.1f /clk/SM (21 FFMA2 : F2F + DADD + FADD)\n", wps, ms, 2 * n / ms / 1e9, 2 * n / (ms * 1e-3) / sms / clk);
    }
    cudaFree(out);
    return 0;
}
