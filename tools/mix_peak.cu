// tools/mix_peak.cu — can the FP64 and FP32 pipes of a B200 SM run side by side?  DFMA alone, DFMA + FFMA alternating
// (k_convolve's image / variance mix), DFMA + FFMA2 (two variance products per instruction), at 8 and 12 warps per SM.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o build/mix_peak tools/mix_peak.cu
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pack2(float a, float b) { u64 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void unpack2(u64 v, float &a, float &b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c) { u64 d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
#define NA 21
template <int MODE>   // 0: DFMA only, 1: DFMA + FFMA, 2: DFMA + FFMA2 for every second DFMA
__global__ void k_mix(double *out, int iters, double a, float b)
{
    double acc[NA]; float fa[NA]; u64 pa[NA / 2 + 1];
#pragma unroll
    for (int i = 0; i < NA; i++) { acc[i] = threadIdx.x + i; fa[i] = i; }
#pragma unroll
    for (int i = 0; i < NA / 2 + 1; i++) pa[i] = pack2(i, i + 1);
    const u64 pb = pack2(b, b * 0.5f);
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < NA; i++) {
            acc[i] = fma(acc[i], a, 0.25);
            if (MODE == 1) fa[i] = fmaf(fa[i], b, 0.5f);
            if (MODE == 2 && (i & 1) == 0) pa[i / 2] = ffma2(pa[i / 2], pb, pb);
        }
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < NA; i++) s += acc[i] + fa[i];
#pragma unroll
    for (int i = 0; i < NA / 2 + 1; i++) { float x, y; unpack2(pa[i], x, y); s += x + y; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
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
    double *out; cudaMalloc(&out, sizeof(double) * sms * 1024);
    const int iters = 4000;
    for (int wps : {4, 8, 12, 16}) {
        const int thr = wps * 32;
        const double n = (double)sms * thr * iters * NA;
        float m0 = time_ms([&] { k_mix<0><<<sms, thr>>>(out, iters, 1.0000001, 1.0001f); });
        float m1 = time_ms([&] { k_mix<1><<<sms, thr>>>(out, iters, 1.0000001, 1.0001f); });
        float m2 = time_ms([&] { k_mix<2><<<sms, thr>>>(out, iters, 1.0000001, 1.0001f); });
        printf("%2d warps/SM: DFMA only %.1f /clk/SM | DFMA+FFMA %.1f DFMA/clk/SM | DFMA+FFMA2(1:2) %.1f DFMA/clk/SM\n", wps,
               n / (m0 * 1e-3) / sms / clk, n / (m1 * 1e-3) / sms / clk, n / (m2 * 1e-3) / sms / clk);
    }
    return 0;
}
