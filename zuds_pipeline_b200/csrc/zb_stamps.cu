// zb_stamps.cu — stamp statistics and sub-stamp (PSF centre) search, one CTA per grid cell.
//
// Restates hotpants/functions.c:187-361 buildStamps (getCenters branch), :363-388 cutStamp,
// :763-1035 getStampStats3, :1037-1107 sigma_clip, :504-716 getPsfCenters, :445-502
// checkPsfCenter and :1676-1714 quick_sort.  Stamps of one region are spatially disjoint and
// every mask side effect of the search stays inside the stamp rectangle, so the
// cells are independent work items; the raster-order greedy semantics are kept INSIDE the CTA:
// candidates are found by a block-wide "first index that passes" search and processed one at
// a time.  The stamp (<= fwStamp^2 pixels + mask) lives in shared memory for the whole search.
//
// Compiled with -fmad=false: discrete decisions here must match the reference's unfused
// double arithmetic bit for bit.
#include "zb_internal.h"
#include "zb_hist.cuh"

#define ZB_NFAIL 96          // remembered failed sub-stamp centres per grid cell (more are simply re-checked)
// Two CTA sizes: 256 threads for survey-sized stamps (hundreds of grid cells keep the SMs busy), 1024
// threads for the few large stamps of a coadd-scale frame, where a cell's raster-greedy search is the
// critical path and evaluates 32 candidates per round instead of 8.
namespace zs256 {
#define NT 256
#define NW (NT / 32)     // candidates evaluated concurrently by the sub-stamp search
#include "zb_stamps_body.cuh"
#undef NT
#undef NW
}
namespace zs1024 {
#define NT 1024
#define NW (NT / 32)
#include "zb_stamps_body.cuh"
#undef NT
#undef NW
}

int zb_launch_stamp_search(zb_ctx *ctx, const float *conv_img, int first_cell, int n_cells, cudaStream_t st, int dir_t)
{
    if (dir_t < 0) dir_t = ctx->h_dev->conv_template;
    if (n_cells <= 0) return ZB_OK;
    size_t npx = (size_t)ctx->sz.fw_stamp * ctx->sz.fw_stamp;
    size_t smem = npx * sizeof(float) + ((npx + 1) & ~(size_t)1) * sizeof(unsigned short) + npx + 32 + ((npx + 31) / 32) * sizeof(unsigned);
    unsigned char *gs = NULL;
    size_t per = 0;
    if (smem > ZB_SMEM_OPTIN) {
        per = (smem + 255) & ~(size_t)255;
        const size_t need = per * (size_t)ctx->n_cells;
        if (ctx->search_scratch_bytes < need) {
            cudaStreamSynchronize(st);
            if (ctx->d_search_scratch) cudaFree(ctx->d_search_scratch);
            ctx->d_search_scratch = NULL; ctx->search_scratch_bytes = 0;
            if (cudaMalloc((void **)&ctx->d_search_scratch, need) != cudaSuccess) {
                snprintf(ctx->err, sizeof ctx->err, "cudaMalloc of %zu bytes of stamp-search scratch failed (stamps of %d x %d pixels)", need, ctx->sz.fw_stamp, ctx->sz.fw_stamp);
                return ZB_ERR_NOMEM;
            }
            ctx->search_scratch_bytes = need;
        }
        gs = ctx->d_search_scratch; smem = 0;
    }
    if (npx >= 160 * 160) {
        cudaFuncSetAttribute(zs1024::k_stamp_search, cudaFuncAttributeMaxDynamicSharedMemorySize, ZB_SMEM_OPTIN);
        zs1024::k_stamp_search<<<n_cells, 1024, smem, st>>>(ctx->d_dev, conv_img, ctx->d_min, ctx->d_cells, first_cell, ctx->d_ran, ctx->n_ran, gs, per, dir_t);
    } else {
        cudaFuncSetAttribute(zs256::k_stamp_search, cudaFuncAttributeMaxDynamicSharedMemorySize, ZB_SMEM_OPTIN);
        zs256::k_stamp_search<<<n_cells, 256, smem, st>>>(ctx->d_dev, conv_img, ctx->d_min, ctx->d_cells, first_cell, ctx->d_ran, ctx->n_ran, gs, per, dir_t);
    }
    ZB_DBG("k_stamp_search");
    ctx->launches++;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return zb_cuda_fail(ctx, e, "k_stamp_search", __FILE__, __LINE__);
    return ZB_OK;
}
