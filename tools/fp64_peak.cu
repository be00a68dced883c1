// tools/fp64_peak.cu — measures the FP64 pipe of the GPU it runs on: DFMA, DADD/DMUL, F2F.F64.F32 and
// LDS.128 issue rates per SM per clock.  Diagnostic only (sizes the k_convolve roofline in DESIGN.md).
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o build/fp64_peak tools/fp64_peak.cu
#include <cstdio>
#include <cuda_runtime.h>

#define NACC 16
__global__ void k_dfma(double *out, int iters, double a, double b) {
    double acc[NACC];
#pragma unroll
    for (int i = 0; i < NACC; i++) acc[i] = threadIdx.x + i;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < NACC; i++) acc[i] = fma(acc[i], a, b);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < NACC; i++) s += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_dmuladd(double *out, int iters, double a, double b) {
    double acc[NACC];
#pragma unroll
    for (int i = 0; i < NACC; i++) acc[i] = threadIdx.x + i;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < NACC; i++) acc[i] = __dadd_rn(__dmul_rn(acc[i], a), b);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < NACC; i++) s += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_cvt(double *out, int iters, float a) {
    float f[NACC]; double acc[NACC];
#pragma unroll
    for (int i = 0; i < NACC; i++) { f[i] = threadIdx.x * a + i; acc[i] = 0; }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < NACC; i++) { acc[i] = (double)f[i]; f[i] = f[i] * a + __double2float_rn(acc[i]) * 0.f; }
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < NACC; i++) s += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_lds128(double *out, int iters) {
    __shared__ double2 sm[1024];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) sm[i] = make_double2(i, 1);
    __syncthreads();
    double2 acc = make_double2(0, 0);
    int idx = threadIdx.x;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) { double2 v = sm[(idx + i * 33) & 1023]; acc.x += v.x; acc.y += v.y; }
        idx += 7;
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc.x + acc.y;
}

template <typename F> static float time_ms(F f) {
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    f(); cudaDeviceSynchronize();
    cudaEventRecord(a); f(); cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b); return ms;
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    const int sms = p.multiProcessorCount;
    printf("%s: %d SMs, max clock %d MHz\n", p.name, sms, clk / 1000);
    double *out; cudaMalloc(&out, sizeof(double) * sms * 8 * 512);
    const int iters = 20000;
    for (int warps = 4; warps <= 32; warps *= 2) {
        int threads = 256, blocks = sms * warps * 32 / threads;
        float ms = time_ms([&] { k_dfma<<<blocks, threads>>>(out, iters, 1.0000001, 1e-9); });
        double ops = (double)blocks * threads * iters * NACC;
        printf("DFMA      %2d warps/SM: %.2f ms  %.2f T DFMA/s  = %.1f /clk/SM at %d MHz\n", warps, ms, ops / ms / 1e9, ops / (ms * 1e-3) / sms / (clk * 1e3), clk / 1000);
        ms = time_ms([&] { k_dmuladd<<<blocks, threads>>>(out, iters, 1.0000001, 1e-9); });
        printf("DMUL+DADD %2d warps/SM: %.2f ms  %.2f T pairs/s = %.1f pairs/clk/SM\n", warps, ms, ops / ms / 1e9, ops / (ms * 1e-3) / sms / (clk * 1e3));
    }
    {
        int threads = 256, blocks = sms * 8;
        float ms = time_ms([&] { k_cvt<<<blocks, threads>>>(out, iters, 1.0001f); });
        double ops = (double)blocks * threads * iters * NACC;
        printf("F2F.F64.F32 (+FFMA, F2F.F32.F64): %.2f ms = %.1f cvt-pairs/clk/SM\n", ms, ops / (ms * 1e-3) / sms / (clk * 1e3));
        ms = time_ms([&] { k_lds128<<<blocks, threads>>>(out, iters); });
        ops = (double)blocks * threads * iters * 8;
        printf("LDS.128 (+2 DADD): %.2f ms = %.1f B/clk/SM\n", ms, 16 * ops / (ms * 1e-3) / sms / (clk * 1e3));
    }
    return 0;
}
