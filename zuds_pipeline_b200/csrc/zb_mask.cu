// zb_mask.cu — input-stage mask of every buffered region.
//
// Restates, for all regions of a pair at once:
//   hotpants/main.c:778-803      -imi / -tmi ingest   (v | 0x20*(v>0))
//   hotpants/functions.c:1388    makeInputMask        (value tests -> 0x81/0x82/0x84)
//   hotpants/functions.c:1353    spreadMask           (0x40 on non-bad pixels within +-width/2 of a bad one)
//   hotpants/main.c:811-824      sBorder frame        (0x100|0x400)
// Integer/byte work, HBM-bound: 2 float loads + 1-2 mask loads + one int32 store per pixel.
#include "zb_internal.h"

#define TILE 32

__global__ void __launch_bounds__(256)
k_input_mask(const ZbDev *__restrict__ D, const float *__restrict__ sci, const float *__restrict__ ref,
             const void *__restrict__ smask, int sdt, const void *__restrict__ tmask, int tdt,
             int32_t *__restrict__ m_in, int y_lo, int y_hi)
{
    const ZbRegion R = D->reg[blockIdx.z];
    const float fill = D->p.fill_val;
    const float tU = D->tU, tL = D->tL, iU = D->iU, iL = D->iL;
    int y = blockIdx.y;
    if (y >= R.rpy || y < y_lo || y >= y_hi) return;          // [y_lo, y_hi): row-tile window (whole region: 0 .. INT_MAX)
    for (int x = blockIdx.x * blockDim.x + threadIdx.x; x < R.rpx; x += gridDim.x * blockDim.x) {
        long long g = (long long)(R.by0 + y) * D->nx + (R.bx0 + x);
        float t = ref[g], i = sci[g];
        int m = 0;
        int v = zb_read_mask(smask, sdt, g);
        m |= v | (ZB_FLAG_INPUT_MASK * (v > 0));
        v = zb_read_mask(tmask, tdt, g);
        m |= v | (ZB_FLAG_INPUT_MASK * (v > 0));
        m |= (ZB_FLAG_INPUT_ISBAD | ZB_FLAG_BAD_PIXVAL) * (t == fill || i == fill);
        m |= (ZB_FLAG_INPUT_ISBAD | ZB_FLAG_SAT_PIXEL) * (t >= tU || i >= iU);
        m |= (ZB_FLAG_INPUT_ISBAD | ZB_FLAG_LOW_PIXEL) * (t <= tL || i <= iL);
        m_in[R.off + (long long)y * R.rpx + x] = m;
    }
}

// Box dilation of the 0x80 bit by +-w2 (clipped to the region) + the sBorder frame.
// One CTA = TILE x TILE output pixels; the bad flags of the tile + halo are staged in shared
// memory and OR-reduced separably (rows, then columns).
__global__ void __launch_bounds__(TILE * 8)
k_spread_mask(const ZbDev *__restrict__ D, int32_t *__restrict__ m_in, int w2, int sborder, int y_lo, int y_hi)
{
    extern __shared__ unsigned char sm[];
    const ZbRegion R = D->reg[blockIdx.z];
    const int x0 = blockIdx.x * TILE, y0 = blockIdx.y * TILE;
    if (x0 >= R.rpx || y0 >= R.rpy || y0 >= y_hi || y0 + TILE <= y_lo) return;
    const int tw = TILE + 2 * w2;
    unsigned char *bad = sm;                 // [tw][tw]
    unsigned char *rowor = sm + tw * tw;     // [tw][TILE]
    int32_t *m = m_in + R.off;
    const int tid = threadIdx.y * blockDim.x + threadIdx.x, nt = blockDim.x * blockDim.y;
    for (int k = tid; k < tw * tw; k += nt) {
        int ty = k / tw, tx = k - ty * tw;
        int x = x0 + tx - w2, y = y0 + ty - w2;
        unsigned char b = 0;
        if (x >= 0 && x < R.rpx && y >= max(0, y_lo) && y < min(R.rpy, y_hi)) b = (m[(long long)y * R.rpx + x] & ZB_FLAG_INPUT_ISBAD) ? 1 : 0;
        bad[k] = b;
    }
    __syncthreads();
    for (int k = tid; k < tw * TILE; k += nt) {
        int ty = k / TILE, tx = k - ty * TILE;
        unsigned char b = 0;
        for (int d = 0; d <= 2 * w2; d++) b |= bad[ty * tw + tx + d];
        rowor[k] = b;
    }
    __syncthreads();
    for (int k = tid; k < TILE * TILE; k += nt) {
        int ty = k / TILE, tx = k - ty * TILE;
        int x = x0 + tx, y = y0 + ty;
        if (x >= R.rpx || y >= R.rpy || y < y_lo || y >= y_hi) continue;
        unsigned char b = 0;
        for (int d = 0; d <= 2 * w2; d++) b |= rowor[(ty + d) * TILE + tx];
        int v = m[(long long)y * R.rpx + x];
        int nv = v;
        if (b && !(v & ZB_FLAG_INPUT_ISBAD)) nv |= ZB_FLAG_OK_CONV;
        if (x < sborder || x >= R.rpx - sborder || y < sborder || y >= R.rpy - sborder)
            nv |= (ZB_FLAG_T_BAD | ZB_FLAG_I_BAD);
        if (nv != v) m[(long long)y * R.rpx + x] = nv;
    }
}

int zb_launch_input_mask(zb_ctx *ctx, const float *sci, const float *ref, const void *smask, int sdt,
                         const void *tmask, int tdt, cudaStream_t st, int y_lo, int y_hi)
{
    const zb_sizes &sz = ctx->sz;
    dim3 g1((sz.max_rpix_x + 255) / 256, sz.max_rpix_y, sz.n_regions);
    k_input_mask<<<g1, 256, 0, st>>>(ctx->d_dev, sci, ref, smask, sdt, tmask, tdt, ctx->d_min, y_lo, y_hi); ZB_DBG("k_input_mask");
    ctx->launches++;
    int width = (int)(ctx->p.hw_kernel * ctx->p.kf_spread_mask1);
    int w2 = width > 0 ? width / 2 : 0;
    int tw = TILE + 2 * w2;
    size_t smem = (size_t)tw * tw + (size_t)tw * TILE;
    if (smem > ZB_SMEM_OPTIN) { snprintf(ctx->err, sizeof ctx->err, "spread width %d too large", width); return ZB_ERR_UNSUPPORTED; }
    if (smem > 48 * 1024) cudaFuncSetAttribute(k_spread_mask, cudaFuncAttributeMaxDynamicSharedMemorySize, ZB_SMEM_OPTIN);
    dim3 g2((sz.max_rpix_x + TILE - 1) / TILE, (sz.max_rpix_y + TILE - 1) / TILE, sz.n_regions);
    // width <= 0: spreadMask returns at once (functions.c:1362-1364) but the frame is still drawn
    k_spread_mask<<<g2, dim3(TILE, 8), smem, st>>>(ctx->d_dev, ctx->d_min, width > 0 ? w2 : 0, sz.s_border, y_lo, y_hi); ZB_DBG("k_spread_mask");
    ctx->launches++;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return zb_cuda_fail(ctx, e, "input mask kernels", __FILE__, __LINE__);
    return ZB_OK;
}
