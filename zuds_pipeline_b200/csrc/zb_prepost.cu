// zb_prepost.cu — the elementwise steps zuds does around the hotpants call, on the GPU:
//   zb_mask_combine        zuds/subtraction.py:135  badpix = refmask | scimask
//                          zuds/mask.py:42-72       boolean = (badpix & BAD_SUM) > 0   (uint8 0/1, what -imi reads)
//   zb_mask_flag_sentinel  zuds/subtraction.py:167-177  submask |= 2**17 where sub == 1e-30 (exact float equality)
//   zb_rms_from_weight     zuds/image.py:179-190    rms = 1/sqrt(w); BIG_RMS where masked or >= 0.9 SATURATE
//   zb_weight_from_rms     zuds/image.py:146-149    w = 1/rms^2; 0 where masked or >= 0.9 SATURATE
// Integer / correctly-rounded float32 work, bit-identical to the numpy statements they replace.  HBM-bound
// streaming (5-13 bytes per pixel), grid-stride, 16 bytes per thread per trip where the types allow.
// Device pointers in and out, so that a caller holding the images in HBM never goes back to the host between
// these steps, zb_background_estimate and zb_subtract.
#include "zb_internal.h"

__global__ void __launch_bounds__(256)
k_mask_combine(const void *__restrict__ a, int adt, const void *__restrict__ b, int bdt, long long n, int bad_sum,
               int32_t *__restrict__ out_or, uint8_t *__restrict__ out_bool)
{
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const int v = zb_read_mask(a, adt, i) | zb_read_mask(b, bdt, i);
        if (out_or) out_or[i] = v;
        if (out_bool) out_bool[i] = (uint8_t)((v & bad_sum) > 0);
    }
}

__global__ void __launch_bounds__(256)
k_mask_flag_sentinel(int32_t *__restrict__ mask, const float *__restrict__ diff, long long n, float sentinel, int bit_value)
{
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        if (diff[i] == sentinel) mask[i] |= bit_value;
}

// mode 0: rms from weight; mode 1: weight from rms
__global__ void __launch_bounds__(256)
k_rms_weight(const float *__restrict__ in, const uint8_t *__restrict__ bad, const float *__restrict__ data, float sat_level,
             float fill, long long n, float *__restrict__ out, int mode)
{
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        float v;
        if (bad && bad[i]) v = fill;
        else {
            const float x = in[i];
            v = mode == 0 ? __fdiv_rn(1.0f, __fsqrt_rn(x)) : __fdiv_rn(1.0f, __fmul_rn(x, x));
        }
        if (data && data[i] >= sat_level) v = fill;
        out[i] = v;
    }
}

// -tp / -ip (hotpants/main.c:728-733): the ADU pedestal comes off every pixel before anything else looks at it
__global__ void __launch_bounds__(256)
k_sub_pedestal(const float *__restrict__ in, float ped, long long n, float *__restrict__ out)
{
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        out[i] = __fsub_rn(in[i], ped);
}

static int grid_for(long long n) {
    long long g = (n + 255) / 256;
    return (int)(g < 1 ? 1 : (g > 148 * 16 ? 148 * 16 : g));
}

static bool mask_dt_ok(const void *m, int dt) { return !m || dt == ZB_MASK_U8 || dt == ZB_MASK_I16 || dt == ZB_MASK_I32; }

extern "C" int zb_mask_combine(zb_ctx *ctx, const void *sci_mask, int sci_dtype, const void *ref_mask, int ref_dtype, long long n,
                               int bad_sum, int32_t *out_or, uint8_t *out_bool, void *stream)
{
    if (!ctx || n < 0 || (!out_or && !out_bool)) { if (ctx) snprintf(ctx->err, sizeof ctx->err, "null argument"); return ZB_ERR_PARAM; }
    if (!mask_dt_ok(sci_mask, sci_dtype) || !mask_dt_ok(ref_mask, ref_dtype)) { snprintf(ctx->err, sizeof ctx->err, "bad mask dtype"); return ZB_ERR_PARAM; }
    if (cudaSetDevice(ctx->device) != cudaSuccess) return zb_cuda_fail(ctx, cudaGetLastError(), "cudaSetDevice", __FILE__, __LINE__);
    cudaStream_t st = zb_pick_stream(ctx, stream);
    if (n) k_mask_combine<<<grid_for(n), 256, 0, st>>>(sci_mask, sci_dtype, ref_mask, ref_dtype, n, bad_sum, out_or, out_bool);
    ctx->launches++;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return zb_cuda_fail(ctx, e, "k_mask_combine", __FILE__, __LINE__);
    return ZB_OK;
}

extern "C" int zb_mask_flag_sentinel(zb_ctx *ctx, int32_t *mask, const float *diff, long long n, float sentinel, int bit_value, void *stream)
{
    if (!ctx || !mask || !diff || n < 0) { if (ctx) snprintf(ctx->err, sizeof ctx->err, "null argument"); return ZB_ERR_PARAM; }
    if (cudaSetDevice(ctx->device) != cudaSuccess) return zb_cuda_fail(ctx, cudaGetLastError(), "cudaSetDevice", __FILE__, __LINE__);
    cudaStream_t st = zb_pick_stream(ctx, stream);
    if (n) k_mask_flag_sentinel<<<grid_for(n), 256, 0, st>>>(mask, diff, n, sentinel, bit_value);
    ctx->launches++;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return zb_cuda_fail(ctx, e, "k_mask_flag_sentinel", __FILE__, __LINE__);
    return ZB_OK;
}

static int rms_weight(zb_ctx *ctx, const float *in, const uint8_t *bad, const float *data, float sat_level, float fill, long long n,
                      float *out, int mode, void *stream)
{
    if (!ctx || !in || !out || n < 0) { if (ctx) snprintf(ctx->err, sizeof ctx->err, "null argument"); return ZB_ERR_PARAM; }
    if (cudaSetDevice(ctx->device) != cudaSuccess) return zb_cuda_fail(ctx, cudaGetLastError(), "cudaSetDevice", __FILE__, __LINE__);
    cudaStream_t st = zb_pick_stream(ctx, stream);
    if (n) k_rms_weight<<<grid_for(n), 256, 0, st>>>(in, bad, data, sat_level, fill, n, out, mode);
    ctx->launches++;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return zb_cuda_fail(ctx, e, "k_rms_weight", __FILE__, __LINE__);
    return ZB_OK;
}

extern "C" int zb_rms_from_weight(zb_ctx *ctx, const float *weight, const uint8_t *bad, const float *data, float sat_level, float big_rms,
                                  long long n, float *rms, void *stream)
{
    return rms_weight(ctx, weight, bad, data, sat_level, big_rms, n, rms, 0, stream);
}

extern "C" int zb_weight_from_rms(zb_ctx *ctx, const float *rms, const uint8_t *bad, const float *data, float sat_level,
                                  long long n, float *weight, void *stream)
{
    return rms_weight(ctx, rms, bad, data, sat_level, 0.f, n, weight, 1, stream);
}

// Copies of the two images with their pedestals removed, in buffers the context keeps (zb_subtract works on them
// instead of the caller's arrays when -tp / -ip are given).
int zb_launch_pedestals(zb_ctx *ctx, const float *sci, const float *ref, float i_ped, float t_ped, long long n,
                        const float **sci_out, const float **ref_out, cudaStream_t st)
{
    const size_t bytes = (size_t)n * sizeof(float);
    for (int k = 0; k < 2; k++) {
        if (ctx->ped_bytes[k] >= bytes && ctx->d_ped[k]) continue;
        cudaStreamSynchronize(st);
        if (ctx->d_ped[k]) cudaFree(ctx->d_ped[k]);
        ctx->d_ped[k] = NULL; ctx->ped_bytes[k] = 0;
        if (cudaMalloc((void **)&ctx->d_ped[k], bytes) != cudaSuccess) { snprintf(ctx->err, sizeof ctx->err, "cudaMalloc of %zu bytes failed (pedestal copy)", bytes); return ZB_ERR_NOMEM; }
        ctx->ped_bytes[k] = bytes;
    }
    k_sub_pedestal<<<grid_for(n), 256, 0, st>>>(sci, i_ped, n, ctx->d_ped[0]);
    k_sub_pedestal<<<grid_for(n), 256, 0, st>>>(ref, t_ped, n, ctx->d_ped[1]);
    ctx->launches += 2;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return zb_cuda_fail(ctx, e, "k_sub_pedestal", __FILE__, __LINE__);
    *sci_out = ctx->d_ped[0]; *ref_out = ctx->d_ped[1];
    return ZB_OK;
}
