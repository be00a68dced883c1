// tools/tma_probe.cu — isolates the TMA tensor load k_convolve uses (cp.async.bulk.tensor.2d of a float box with
// out-of-bounds zero fill).  usage: tma_probe NX NY BOXW BOXH X0 Y0 [smem_offset]
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }

__global__ void k_probe(const __grid_constant__ CUtensorMap tm, float *out, int boxw, int boxh, int x0, int y0, int off)
{
    extern __shared__ __align__(128) unsigned char dyn[];
    float *dst = (float *)(dyn + off);
    uint64_t *bar = (uint64_t *)(dyn + off + ((boxw * boxh * 4 + 127) & ~127));
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(1) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(boxw * boxh * 4) : "memory");
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                     ::"r"(smem_u32(dst)), "l"(&tm), "r"(x0), "r"(y0), "r"(smem_u32(bar)) : "memory");
    }
    asm volatile("{\n\t.reg .pred P1;\n\tLAB_WAIT:\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t@P1 bra DONE;\n\tbra LAB_WAIT;\n\tDONE:\n\t}"
                 ::"r"(smem_u32(bar)), "r"(0) : "memory");
    for (int k = threadIdx.x; k < boxw * boxh; k += blockDim.x) out[k] = dst[k];
}

typedef CUresult (*enc_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *, const cuuint32_t *,
                           const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main(int argc, char **argv)
{
    if (argc < 7) { printf("usage\n"); return 2; }
    const int nx = atoi(argv[1]), ny = atoi(argv[2]), bw = atoi(argv[3]), bh = atoi(argv[4]), x0 = atoi(argv[5]), y0 = atoi(argv[6]);
    const int off = argc > 7 ? atoi(argv[7]) : 0;
    float *h = (float *)malloc((size_t)nx * ny * 4), *d, *o, *ho = (float *)malloc((size_t)bw * bh * 4);
    for (int i = 0; i < nx * ny; i++) h[i] = (float)i;
    cudaMalloc(&d, (size_t)nx * ny * 4); cudaMalloc(&o, (size_t)bw * bh * 4);
    cudaMemcpy(d, h, (size_t)nx * ny * 4, cudaMemcpyHostToDevice);
    void *f = nullptr; cudaDriverEntryPointQueryResult q;
    cudaError_t ce = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q);
    printf("entry point: %s q=%d f=%p\n", cudaGetErrorString(ce), (int)q, f);
    CUtensorMap tm; memset(&tm, 0, sizeof tm);
    const cuuint64_t gdim[2] = {(cuuint64_t)nx, (cuuint64_t)ny}, gstr[1] = {(cuuint64_t)nx * 4};
    const cuuint32_t box[2] = {(cuuint32_t)bw, (cuuint32_t)bh}, estr[2] = {1, 1};
    CUresult r = ((enc_fn)f)(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, d, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                             CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode: %d\n", (int)r);
    if (r != CUDA_SUCCESS) return 1;
    const size_t sm = off + ((bw * bh * 4 + 127) & ~127) + 64;
    cudaFuncSetAttribute(k_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    k_probe<<<1, 128, sm>>>(tm, o, bw, bh, x0, y0, off);
    ce = cudaDeviceSynchronize();
    printf("kernel: %s\n", cudaGetErrorString(ce));
    if (ce != cudaSuccess) return 1;
    cudaMemcpy(ho, o, (size_t)bw * bh * 4, cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int y = 0; y < bh; y++) for (int x = 0; x < bw; x++) {
        const int gx = x0 + x, gy = y0 + y;
        const float want = (gx < nx && gy < ny && gx >= 0 && gy >= 0) ? h[(size_t)gy * nx + gx] : 0.f;
        if (ho[y * bw + x] != want) bad++;
    }
    printf("box %dx%d at (%d,%d) of %dx%d, smem offset %d: %d wrong of %d\n", bw, bh, x0, y0, nx, ny, off, bad, bw * bh);
    return bad != 0;
}
