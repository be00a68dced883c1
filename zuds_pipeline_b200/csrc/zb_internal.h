// zb_internal.h — shared host/device definitions of the B200 subtraction path.
// Everything here is private to libzb200.so; the public surface is include/zb200.h.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include "../../include/zb200.h"

#define ZB_MAXVAL  1e10      // hotpants/defaults.h:24
#define ZB_ZEROVAL 1e-10     // hotpants/defaults.h:25
// Dynamic shared memory opt-in, the SAME constant for every kernel and every launch: contexts on
// other host threads set it concurrently, and a per-launch value would race with their launches.
#define ZB_SMEM_OPTIN (216 * 1024)
#define ZB_MAX_REGIONS 64
#define ZB_MAX_BASIS   160   // nCompKer + nBGVectors upper bound handled by the fit kernels
#define ZB_MAX_COMP    32    // (ko+1)(ko+2)/2 upper bound (ko <= 6)

// One buffered region (hotpants/main.c:606-658).  All coordinates are image pixels.
struct ZbRegion {
    int rx0, rx1, ry0, ry1;      // good-data bounds, inclusive
    int bx0, bx1, by0, by1;      // buffered bounds, inclusive
    int rpx, rpy;                // buffered size
    int own_x1, own_y1;          // last column/row this region is the LAST writer of (main.c:548-551)
    long long off;               // element offset of this region in the per-region scratch arrays
    int stamp_base;              // first stamp slot of this region
    int skip;                    // 1: no usable stamps, region left untouched (main.c:1002-1009)
};

// Device-resident constants of one call.
struct ZbDev {
    zb_params p;
    zb_sizes  sz;
    int nx, ny;
    int n_regions;
    int conv_template;           // 1: convolve template (ref), 0: convolve image (sci)
    float tU, tUK, tL, iU, iUK, iL;   // thresholds after pedestal removal (main.c:212-217)
    ZbRegion reg[ZB_MAX_REGIONS];
};

// One grid cell of the stamp search (hotpants/main.c:837-906) and its outcome.
struct ZbCell {
    int region, gk, gl;
    int sx0, sy0, spx, spy;      // stamp origin (region coords) and size
    float kft;                   // kerFitThresh in force (may be halved on retry)
    int  stat_status;            // getStampStats3 return code
    int  clk_stats, clk_search;  // SM clock cycles spent in the two halves of k_stamp_search (diagnostics)
    int  nss;
    int  xss[ZB_MAX_KSS], yss[ZB_MAX_KSS];    // region coords
    double in_mode, in_fwhm;     // slot contents before this cell ran (stale-slot quirk)
    double sum, mean, median, mode, sd, fwhm, lfwhm;
};

// A stamp that survived the search: a slot with big per-stamp arrays.
struct ZbStamp {
    int region;
    int x0, y0;
    int nss, sscnt;
    int xss[ZB_MAX_KSS], yss[ZB_MAX_KSS];
    int refill;                  // set by the clip kernel when (sscnt) changed and the slot must be refilled
    double mode, fwhm, chi2, sum;
    double sig1;                 // last getStampSig value
    int sig_flag;
    int excl;                    // -c b: left out of check_stamps' trial fit (kernel sum beyond kerSigReject sigma, alard.c:590-606)
};

// Per-region state of the fit loop.
struct ZbFit {
    int n_stamps;                // kept stamps
    int active;                  // 1 while check_again keeps asking for another pass
    int iters, singular, n_skipped, pad;
    double meansig, scatter;
};

// Region QA statistics (main.c:1479-1554): four getStampStats3 runs per region, advanced in lock
// step by whole-grid passes over the region (zb_stats.cu).
#define ZB_QA_NBLK 256          // stride of the per-CTA partial sums (max CTAs per region)
#define ZB_QA_NBLK_MIN 64
struct ZbQaStat {
    int phase;                   // 0 sigma-clip passes, 1 histogram passes, 2 done
    int status;                  // getStampStats3 return code
    int maxiter, umask, smask, use_noise;
    int ncrit, cnt_prev, tries, imax, npts_s, ticket;
    int exact, pad;                // exact: keep the double clip test (degenerate mean / sd)
    float clo, chi, llo, lhi;      // accepted float interval: all criteria so far / the last criterion
    double crit_mean[8], crit_istd[8];
    double mean, sd, bin1, binsize, samp_med;
    double sum, median, mode, fwhm, lfwhm;
    // clip iterations >= 2 from the shell list (zb_stats.cu): sums over the first criterion's interval, and the
    // float interval every later criterion must still contain for the list to hold all later rejections
    double base_s1, base_s2;
    int base_n, shell_bad;
    float core_lo, core_hi;        // core of the shell lists as collected by the last shell pass
    float ncore_lo, ncore_hi;      // core the next frame pass collects with (from the latest mean / sd)
    int bins[256];
};
#define ZB_QA_SHELL_DIV 4       // a warp's shell segment holds 1/4 of the pixels it reads in a pass
struct ZbQaPart { double s1, s2; int n, pad; };
// CTAs per region of a k_qa_pass launch: enough to put ~4 CTAs on every SM whatever the number of regions
static inline int zb_qa_nblk(int n_regions) {
    int nblk = n_regions >= 5 ? ZB_QA_NBLK_MIN : (4 * 148 + n_regions - 1) / n_regions;
    return nblk < ZB_QA_NBLK_MIN ? ZB_QA_NBLK_MIN : (nblk > ZB_QA_NBLK ? ZB_QA_NBLK : nblk);
}

#define ZB_CUDA_OK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return zb_cuda_fail(ctx, e_, #call, __FILE__, __LINE__); } while (0)

// Row-tile session state (zb_tile.cu): one oversized single-region frame split over ranks by rows.
struct ZbTile {
    int on, world, rank;
    zb_tile_info info;
    const float *sci, *ref, *sci_rms, *ref_rms;
    const void *smask, *tmask; int sdt, tdt;
    float *diff, *noise; int32_t *out_mask;
    cudaStream_t st;
    float kft; int retried;
    int cell_lo, cell_hi;          // grid cells of this rank
    int slot_lo, slot_hi;          // kept-stamp slots of this rank (contiguous: compaction keeps grid order)
    int n_kept;
    int pending_cell;              // cell the stale-slot walk wants searched again, with the values it inherits
    double pend[7];
    zb_region_info ri;             // the region record, completed by zb_tile_finish on rank 0
};

struct zb_ctx {
    int device;
    cudaStream_t stream;
    char err[512];
    long long launches;
    // cached configuration
    zb_params p; zb_sizes sz; int configured;
    ZbDev *h_dev; ZbDev *d_dev;
    // device workspace
    int32_t *d_min;              // input-stage mask per region   [sum rpx*rpy]
    int32_t *d_cmask;            // mask of the convolved image   [sum rpx*rpy]
    int32_t *d_omask;            // output-stage mask per region  [sum rpx*rpy]
    float   *d_rdiff, *d_rnoise; // per-region diff / noise before output masking
    long long region_elems;
    double  *d_ran;  int n_ran;  // ran1 stream, seed -666 (functions.c:811)
    double  *d_fx, *d_fy, *d_kvec; int *d_ren;   // basis (alard.c:134-197)
    ZbCell  *d_cells, *h_cells; int n_cells;
    unsigned char *d_search_scratch; size_t search_scratch_bytes;   // stamps larger than shared memory (zb_stamps.cu)
    ZbStamp *d_stamps, *h_stamps; int n_slots;
    double  *d_vec;              // [slot][nvec][npix]
    double  *d_kref;             // [slot][npix]
    double  *d_gram;             // [slot][nvec*(nvec+1)/2]
    double  *d_rhs;              // [slot][nvec]
    double  *d_wxy;              // [slot][n_comp]
    double  *d_mat;              // [region][n*n]
    double  *d_sol;              // [region][n_comp_total+1]
    double  *d_vv; int *d_indx;  // LU scratch
    ZbFit   *d_fit, *h_fit;
    float   *d_ped[2]; size_t ped_bytes[2];      // -tp / -ip: the images with their pedestals removed (zb_prepost.cu)
    double  *d_check, *h_check;  // -c b: [n_slots] per-stamp kernel sums + [n_regions] trial kernel sums (zb_check.cu)
    int     *d_jobs, *d_njobs, *h_njobs;
    double  *d_kaks;             // [region][blocks+1]  sum |K| per block in the reference's tap order
    void    *d_ktap;             // [region][blocks+1][fw*fwp] 16-byte taps (K double, K^2 float), reversed, rows padded: k_convolve's operand
    int     *d_tilectr;          // tile counter of k_convolve's dynamic schedule
    cudaGraphExec_t lu_exec; int lu_key[2];       // the LU launch sequence of zb_launch_fit_solve as a graph, for (mat_size, n_regions)
    unsigned char conv_geom[512]; int conv_geom_key[6]; int conv_geom_bytes;   // cached tile shape of k_convolve (zb_conv.cu)
    int      n_sm;               // multiprocessors of the device (grid of the persistent kernels)
    double  *d_kstat;            // per-region: ksum, aks
    double  *d_stats;  double *h_stats;   // region QA scratch
    ZbQaStat *d_qa;              // [region][4]
    ZbQaPart *d_qpart;           // [region][5][ZB_QA_NBLK]  (4 stats + getNoiseStats3)
    float *d_qshell; int *d_qshell_n; int qshell_cap;   // [region][4][ZB_QA_NBLK * 8 warps][cap] clip-shell values and their counts
    cudaEvent_t ev_sync;         // blocking-sync event (zb_stream_wait)
    void *d_bkg; double *h_bkg;  // background-estimate selection state (zb_bkg.cu)
    // staging for zb_subtract_host
    void *d_stage[9]; size_t stage_bytes[9];
    // zb_subtract_batch: a second staging set, copy streams and events (H2D of pair n+1 and D2H of pair n-1 overlap pair n)
    void *d_stage_b[9]; size_t stage_b_bytes[9];
    cudaStream_t st_h2d, st_d2h;
    cudaEvent_t ev_h2d[2], ev_d2h[2], ev_done[2];
    int batch_ready;
    ZbTile tile;
    // timings
    cudaEvent_t ev[ZB_NSTAGES + 1];
    float ms[ZB_NSTAGES];
};

int zb_cuda_fail(zb_ctx *ctx, cudaError_t e, const char *what, const char *file, int line);
// the `stream` argument of the C ABI: ZB_STREAM_OWN -> the context's private stream, NULL -> the legacy default stream
static inline cudaStream_t zb_pick_stream(zb_ctx *ctx, void *stream) {
    if (stream == ZB_STREAM_OWN) return ctx->stream;
    return stream ? (cudaStream_t)stream : cudaStreamLegacy;
}
// Wait for the stream: cudaStreamSynchronize (spinning; measured faster: 158 vs 143 pairs/s end to end with 24
// contexts on 16 cores), or with ZB_SYNC=block a blocking-sync event that yields the core.
cudaError_t zb_stream_wait(zb_ctx *ctx, cudaStream_t st);
// ZB_DEBUG_SYNC=1 in the environment: synchronise after every kernel so a fault names its kernel
int zb_dbg_sync(zb_ctx *ctx, cudaStream_t st, const char *what);
#define ZB_DBG(what) do { int rc_d = zb_dbg_sync(ctx, st, what); if (rc_d != ZB_OK) return rc_d; } while (0)

// Small control transfers (cell records, stamp table, fit state, job lists) between PINNED host memory and the
// device, done by a copy KERNEL through the unified address space instead of cudaMemcpyAsync: the copy engines
// are FIFOs shared by every context of the GPU, and a 200-byte read-back queued behind another pair's 38 MB image
// copy waits milliseconds (measured: fit loop 4.8 -> 10.5 ms per pair with the D2H engine busy).  dst/src: one
// device pointer and one cudaMallocHost pointer (or two device pointers), 4-byte aligned, bytes a multiple of 4.
cudaError_t zb_copy_small(zb_ctx *ctx, void *dst, const void *src, size_t bytes, cudaStream_t st);
// the same for small device memsets to zero (a memset node may also be served by a copy engine)
cudaError_t zb_zero_small(zb_ctx *ctx, void *dst, size_t bytes, cudaStream_t st);

// kernel launchers (each returns a zb_status)
int zb_launch_input_mask(zb_ctx *ctx, const float *sci, const float *ref, const void *smask, int sdt,
                         const void *tmask, int tdt, cudaStream_t st, int y_lo = 0, int y_hi = 0x7fffffff);
// dir_t: 1 the template is searched, 0 the science image, -1 whichever the call convolves (ZbDev::conv_template)
int zb_launch_stamp_search(zb_ctx *ctx, const float *conv_img, int first_cell, int n_cells, cudaStream_t st, int dir_t = -1);
int zb_launch_fill(zb_ctx *ctx, const float *conv_img, const float *ref_img, const int *d_jobs, int n_jobs, cudaStream_t st);
int zb_launch_fit_iteration(zb_ctx *ctx, const float *conv_img, const float *ref_img,
                            const float *sci, const float *ref, const float *sci_rms, const float *ref_rms, cudaStream_t st);
int zb_launch_convolve(zb_ctx *ctx, const float *sci, const float *ref, const float *sci_rms, const float *ref_rms,
                       const void *smask, int sdt, const void *tmask, int tdt,
                       float *diff, float *noise, int32_t *out_mask, cudaStream_t st,
                       int blk_row0 = 0, int blk_row1 = 0x7fffffff, int row0 = 0, int row1 = 0x7fffffff);
int zb_launch_fit_solve(zb_ctx *ctx, cudaStream_t st);
int zb_launch_stamp_sig(zb_ctx *ctx, const float *sci, const float *ref, const float *sci_rms, const float *ref_rms,
                        int slot0, int slot1, cudaStream_t st);
int zb_launch_clip(zb_ctx *ctx, cudaStream_t st);
int zb_launch_region_stats(zb_ctx *ctx, float *noise, int32_t *out_mask, cudaStream_t st);
int zb_launch_final_sig(zb_ctx *ctx, cudaStream_t st);
int zb_launch_pedestals(zb_ctx *ctx, const float *sci, const float *ref, float i_ped, float t_ped, long long n,
                        const float **sci_out, const float **ref_out, cudaStream_t st);
int zb_check_stamps(zb_ctx *ctx, const float *sci, const float *ref, const float *sci_rms, const float *ref_rms,
                    double *merit, cudaStream_t st);

// host helpers
int zb_stage_buf(zb_ctx *ctx, int i, size_t bytes);
int zb_configure(zb_ctx *ctx, const zb_params *p);
void zb_init_cells(zb_ctx *ctx, int r, float kft, int cell_base = 0);
int zb_finish_readback(zb_ctx *ctx, zb_region_info *regions, double *kernel_sol, zb_stamp_info *stamps_out, int timing, cudaStream_t st);
void zb_host_ran1_stream(double *out, int n);
void zb_host_basis(const zb_params *p, double *kvec, double *fx, double *fy, int *ren);

__host__ __device__ static inline int zb_read_mask(const void *m, int dt, long long idx) {
    if (!m) return 0;
    if (dt == ZB_MASK_U8)  return (int)((const uint8_t *)m)[idx];
    if (dt == ZB_MASK_I16) return (int)((const int16_t *)m)[idx];
    return ((const int32_t *)m)[idx];
}
