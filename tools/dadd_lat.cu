// Dependent-chain latency of FP64 add / mul-then-sub / fused multiply-add on one thread: the floor of the
// back-substitution chain of k_lu_subst (n^2/2 dependent subtractions in the reference's order) and of the panel's
// row updates.   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -fmad=false tools/dadd_lat.cu -o build/dadd_lat
#include <cstdio>
#include <cuda_runtime.h>

__global__ void k_chain(double *out, long long *cyc, const double *in, int n, int mode)
{
    double s = in[0];
    const double q0 = in[1], q1 = in[2], q2 = in[3], q3 = in[4], m = in[5];
    long long t0 = clock64();
    if (mode == 0) for (int i = 0; i < n; i += 4) { s = __dsub_rn(s, q0); s = __dsub_rn(s, q1); s = __dsub_rn(s, q2); s = __dsub_rn(s, q3); }
    else if (mode == 1) for (int i = 0; i < n; i += 4) { s = __dsub_rn(s, __dmul_rn(s, m)); s = __dsub_rn(s, __dmul_rn(s, m)); s = __dsub_rn(s, __dmul_rn(s, m)); s = __dsub_rn(s, __dmul_rn(s, m)); }
    else for (int i = 0; i < n; i += 4) { s = fma(s, m, q0); s = fma(s, m, q1); s = fma(s, m, q2); s = fma(s, m, q3); }
    long long t1 = clock64();
    out[mode] = s; cyc[mode] = t1 - t0;
}

int main()
{
    double h[6] = {1.0, 1e-9, 2e-9, 3e-9, 4e-9, 1e-7}, *d_in, *d_out; long long *d_cyc, cyc[3];
    cudaMalloc(&d_in, sizeof h); cudaMalloc(&d_out, 3 * sizeof(double)); cudaMalloc(&d_cyc, 3 * sizeof(long long));
    cudaMemcpy(d_in, h, sizeof h, cudaMemcpyHostToDevice);
    const int n = 1 << 20;
    const char *names[3] = {"DADD chain", "DMUL -> DADD chain", "DFMA chain"};
    for (int rep = 0; rep < 2; rep++)
        for (int mode = 0; mode < 3; mode++) k_chain<<<1, 1>>>(d_out, d_cyc, d_in, n, mode);
    cudaDeviceSynchronize();
    cudaMemcpy(cyc, d_cyc, sizeof cyc, cudaMemcpyDeviceToHost);
    for (int mode = 0; mode < 3; mode++) printf("%-20s %.2f cycles per dependent step\n", names[mode], (double)cyc[mode] / n);
    printf("status %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
