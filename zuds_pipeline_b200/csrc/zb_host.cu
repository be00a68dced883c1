// zb_host.cu — C ABI of libzb200.so: parameter handling, context/workspace management and
// the per-pair orchestration that replaces hotpants' main() region loop (hotpants/main.c:606-1867).
// There is no CPU compute path here: every pixel and every fit quantity is produced by the CUDA
// kernels in zb_mask.cu / zb_stamps.cu / zb_fit.cu / zb_conv.cu / zb_stats.cu.  The host only
// sequences launches and takes the few discrete decisions hotpants takes between stages
// (stamp compaction, the one-shot threshold retry, "does check_again want another pass").
#include "zb_internal.h"
#include <sched.h>
#include <math.h>
#include <stdlib.h>
#include <new>
#include <vector>

static inline int imin_(int a, int b) { return a < b ? a : b; }
static inline int imax_(int a, int b) { return a < b ? b : a; }

extern "C" const char *zb_version(void) { return "zb200 0.1 (hotpants 5.1.11 semantics, sm_100a)"; }

int zb_cuda_fail(zb_ctx *ctx, cudaError_t e, const char *what, const char *file, int line) {
    if (ctx) snprintf(ctx->err, sizeof ctx->err, "CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString(e), file, line, what);
    return ZB_ERR_CUDA;
}

cudaError_t zb_stream_wait(zb_ctx *ctx, cudaStream_t st) {
    // ZB_SYNC: "yield" (default: poll an event, sched_yield between polls, so that many worker threads can
    // share few cores: 24 contexts on 4 cores drive a B200 as fast as on 16), "spin" (cudaStreamSynchronize
    // busy-wait; needs a core per worker), "block" (blocking event: sleep until the driver's interrupt)
    static int mode = -1;
    if (mode < 0) { const char *e = getenv("ZB_SYNC"); mode = !e ? 2 : (e[0] == 'b' ? 1 : (e[0] == 's' ? 0 : 2)); }
    if (mode == 0 || !ctx->ev_sync) return cudaStreamSynchronize(st);
    cudaError_t e = cudaEventRecord(ctx->ev_sync, st);
    if (e != cudaSuccess) return e;
    if (mode == 1) return cudaEventSynchronize(ctx->ev_sync);
    while ((e = cudaEventQuery(ctx->ev_sync)) == cudaErrorNotReady) sched_yield();
    return e;
}

__global__ void k_copy_words(uint32_t *__restrict__ dst, const uint32_t *__restrict__ src, size_t nwords) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nwords; i += (size_t)gridDim.x * blockDim.x) dst[i] = src[i];
}

cudaError_t zb_copy_small(zb_ctx *ctx, void *dst, const void *src, size_t bytes, cudaStream_t st) {
    static int use_dma = -1;
    if (use_dma < 0) { const char *e = getenv("ZB_SMALL_COPY"); use_dma = (e && e[0] == 'd') ? 1 : 0; }     // "dma": cudaMemcpyAsync
    if (bytes == 0) return cudaSuccess;
    if (use_dma || (bytes & 3) || ((uintptr_t)dst & 3) || ((uintptr_t)src & 3))
        return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, st);
    const size_t nwords = bytes >> 2;
    const unsigned blocks = (unsigned)((nwords + 255) / 256 < 64 ? (nwords + 255) / 256 : 64);
    k_copy_words<<<blocks, 256, 0, st>>>((uint32_t *)dst, (const uint32_t *)src, nwords);
    if (ctx) ctx->launches++;
    return cudaGetLastError();
}

__global__ void k_zero_words(uint32_t *__restrict__ dst, size_t nwords) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nwords; i += (size_t)gridDim.x * blockDim.x) dst[i] = 0u;
}

cudaError_t zb_zero_small(zb_ctx *ctx, void *dst, size_t bytes, cudaStream_t st) {
    if (bytes == 0) return cudaSuccess;
    if ((bytes & 3) || ((uintptr_t)dst & 3)) return cudaMemsetAsync(dst, 0, bytes, st);
    const size_t nwords = bytes >> 2;
    const unsigned blocks = (unsigned)((nwords + 255) / 256 < 64 ? (nwords + 255) / 256 : 64);
    k_zero_words<<<blocks, 256, 0, st>>>((uint32_t *)dst, nwords);
    if (ctx) ctx->launches++;
    return cudaGetLastError();
}

int zb_dbg_sync(zb_ctx *ctx, cudaStream_t st, const char *what) {
    static int on = -1;
    if (on < 0) { const char *e = getenv("ZB_DEBUG_SYNC"); on = (e && e[0] && e[0] != '0') ? 1 : 0; }
    if (!on) return ZB_OK;
    cudaError_t e = cudaStreamSynchronize(st);
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e != cudaSuccess) return zb_cuda_fail(ctx, e, what, __FILE__, __LINE__);
    return ZB_OK;
}

// hotpants/vargs.c:55-128 + defaults.h:27-80
extern "C" void zb_params_default(zb_params *p) {
    static const double sig[3] = {0.7, 1.5, 3.0};
    static const int deg[3] = {6, 4, 2};
    memset(p, 0, sizeof *p);
    p->ngauss = 3;
    for (int i = 0; i < 3; i++) { p->deg[i] = deg[i]; p->beta[i] = (float)(1.0 / (2.0 * sig[i] * sig[i])); }
    p->t_uthresh = p->t_ukthresh = p->i_uthresh = p->i_ukthresh = 25000.f;
    p->t_gain = p->i_gain = 1.f;
    p->hw_kernel = 10;
    p->ker_fit_thresh = 20.f; p->min_frac_good_stamps = 0.1f; p->scale_fit_thresh = 0.5f;
    p->kf_spread_mask1 = 1.f; p->kf_spread_mask2 = 1.f;
    p->nreg_x = p->nreg_y = 1;
    p->nstamp_x = p->nstamp_y = 10;
    p->n_ksstamps = 3; p->hw_ksstamp = 15;
    p->ker_order = 2; p->bg_order = 1;
    p->stat_sig = 3.f; p->ker_sig_reject = 2.f; p->ker_frac_mask = 0.99f;
    p->fill_val = 1.e-30f; p->fill_val_noise = 0.f;
    p->force_convolve = 'b'; p->phot_normalize = 't'; p->fig_merit = 'v';
}

#define FAIL(code, ...) do { if (msg && msglen > 0) snprintf(msg, (size_t)msglen, __VA_ARGS__); return code; } while (0)

// hotpants/vargs.c:482-647 (the FATAL ERROR conditions), then this build's scope limits
extern "C" int zb_params_check(const zb_params *p, char *msg, int msglen) {
    if (p->nx <= 0 || p->ny <= 0) FAIL(ZB_ERR_PARAM, "image size %d x %d", p->nx, p->ny);
    if (p->t_uthresh && p->t_uthresh <= 0) FAIL(ZB_ERR_PARAM, "FATAL ERROR tu (%f) : cannot be negative (or zero)", p->t_uthresh);
    if (p->t_uthresh && p->t_lthresh && p->t_uthresh < p->t_lthresh) FAIL(ZB_ERR_PARAM, "FATAL ERROR tu (%f) : cannot be smaller than tl (%f)", p->t_uthresh, p->t_lthresh);
    if (p->t_gain && p->t_gain <= 0) FAIL(ZB_ERR_PARAM, "FATAL ERROR tg (%f) : cannot be negative (or zero)", p->t_gain);
    if (p->t_rdnoise && p->t_rdnoise <= 0) FAIL(ZB_ERR_PARAM, "FATAL ERROR tr (%f) : cannot be negative (or zero)", p->t_rdnoise);
    if (p->i_uthresh && p->i_uthresh <= 0) FAIL(ZB_ERR_PARAM, "FATAL ERROR iu (%f) : cannot be negative (or zero)", p->i_uthresh);
    if (p->i_uthresh && p->i_lthresh && p->i_uthresh < p->i_lthresh) FAIL(ZB_ERR_PARAM, "FATAL ERROR iu (%f) : cannot be smaller than il (%f)", p->i_uthresh, p->i_lthresh);
    if (p->i_gain && p->i_gain <= 0) FAIL(ZB_ERR_PARAM, "FATAL ERROR ig (%f) : cannot be negative (or zero)", p->i_gain);
    if (p->i_rdnoise && p->i_rdnoise <= 0) FAIL(ZB_ERR_PARAM, "FATAL ERROR ir (%f) : cannot be negative (or zero)", p->i_rdnoise);
    if (p->hw_kernel <= 0) FAIL(ZB_ERR_PARAM, "FATAL ERROR r (%d) : cannot be negative (or zero)", p->hw_kernel);
    if (p->kc_step < 0) FAIL(ZB_ERR_PARAM, "FATAL ERROR kcs (%d) : cannot be negative (or zero)", p->kc_step);
    if (p->ker_fit_thresh < 0) FAIL(ZB_ERR_PARAM, "FATAL ERROR ft (%f) : cannot be negative", p->ker_fit_thresh);
    if (p->scale_fit_thresh > 1 || p->scale_fit_thresh < 0) FAIL(ZB_ERR_PARAM, "FATAL ERROR sft (%f) : cannot be negative or greater than 1", p->scale_fit_thresh);
    if (p->min_frac_good_stamps > 1 || p->min_frac_good_stamps < 0) FAIL(ZB_ERR_PARAM, "FATAL ERROR nft (%f) : cannot be negative or greater than 1", p->min_frac_good_stamps);
    if (p->ker_sig_reject < 0) FAIL(ZB_ERR_PARAM, "FATAL ERROR ks (%f) : cannot be negative", p->ker_sig_reject);
    if (p->kf_spread_mask1 < 0) FAIL(ZB_ERR_PARAM, "FATAL ERROR mins (%f) : cannot be negative", p->kf_spread_mask1);
    if (p->n_ksstamps < 0) FAIL(ZB_ERR_PARAM, "FATAL ERROR nss (%d) : cannot be negative", p->n_ksstamps);
    if (p->hw_ksstamp < 0) FAIL(ZB_ERR_PARAM, "FATAL ERROR nss (%d) : cannot be negative", p->hw_ksstamp);
    if (p->hw_ksstamp < p->hw_kernel) FAIL(ZB_ERR_PARAM, "FATAL ERROR rss (%d) < r (%d) : conflicting requests", p->hw_ksstamp, p->hw_kernel);
    if (!(p->force_convolve == 'i' || p->force_convolve == 't' || p->force_convolve == 'b')) FAIL(ZB_ERR_PARAM, "FATAL ERROR c (%c) : invalid char ('i' or 't' or 'b')", p->force_convolve);
    if (!(p->phot_normalize == 'i' || p->phot_normalize == 't' || p->phot_normalize == 'u')) FAIL(ZB_ERR_PARAM, "FATAL ERROR n (%c) : invalid char ('i', 't', or 'u')", p->phot_normalize);
    if (!(p->fig_merit == 'v' || p->fig_merit == 's' || p->fig_merit == 'h')) FAIL(ZB_ERR_PARAM, "FATAL ERROR fom (%c) : invalid char ('v', 's' or 'h')", p->fig_merit);
    if (p->ker_order < 0) FAIL(ZB_ERR_PARAM, "FATAL ERROR ko (%d) : cannot be negative", p->ker_order);
    if (p->bg_order < 0) FAIL(ZB_ERR_PARAM, "FATAL ERROR bgo (%d) : cannot be negative", p->bg_order);
    if (p->nreg_x < 1 || p->nreg_y < 1 || p->nstamp_x < 1 || p->nstamp_y < 1) FAIL(ZB_ERR_PARAM, "region/stamp grid must be >= 1");
    if (p->ngauss < 1) FAIL(ZB_ERR_PARAM, "ngauss must be >= 1");
    // scope of this build
    if (p->fig_merit != 'v') FAIL(ZB_ERR_UNSUPPORTED, "-fom %c is not built; only the default 'v'", p->fig_merit);
    if (p->use_full_ss) FAIL(ZB_ERR_UNSUPPORTED, "useFullSS is not built");
    if (p->ngauss > ZB_MAX_GAUSS) FAIL(ZB_ERR_UNSUPPORTED, "-ng %d > %d", p->ngauss, ZB_MAX_GAUSS);
    if (p->n_ksstamps > ZB_MAX_KSS || p->n_ksstamps < 1) FAIL(ZB_ERR_UNSUPPORTED, "-nss %d outside 1..%d", p->n_ksstamps, ZB_MAX_KSS);
    if (p->nreg_x * p->nreg_y > ZB_MAX_REGIONS) FAIL(ZB_ERR_UNSUPPORTED, "more than %d regions", ZB_MAX_REGIONS);
    if (p->ker_order > 6) FAIL(ZB_ERR_UNSUPPORTED, "-ko %d > 6", p->ker_order);
    if (p->scale_fit_thresh >= 1.f) FAIL(ZB_ERR_UNSUPPORTED, "-sft 1 (re-search on the already claimed mask, main.c:832-998) is not built");
    // getPsfCenters marks pixels >= tUKThresh while it scans; that is unobservable only while every
    // such pixel is already masked by the -tu/-iu test, i.e. while tuk >= tu.
    if (p->force_convolve != 'i' && p->t_ukthresh < p->t_uthresh) FAIL(ZB_ERR_UNSUPPORTED, "-tuk (%f) < -tu (%f) is not built", p->t_ukthresh, p->t_uthresh);
    if (p->force_convolve != 't' && p->i_ukthresh < p->i_uthresh) FAIL(ZB_ERR_UNSUPPORTED, "-iuk (%f) < -iu (%f) is not built", p->i_ukthresh, p->i_uthresh);
    zb_sizes s; zb_sizes_of(p, &s);
    if (s.n_comp_ker + s.n_bg_vectors > ZB_MAX_BASIS) FAIL(ZB_ERR_UNSUPPORTED, "%d basis vectors > %d", s.n_comp_ker + s.n_bg_vectors, ZB_MAX_BASIS);
    if (s.nstamp_x < 1 || s.nstamp_y < 1) FAIL(ZB_ERR_PARAM, "image too small for any stamp");
    if (msg && msglen > 0) msg[0] = 0;
    return ZB_OK;
}

static void region_bounds(const zb_params *p, const zb_sizes *s, int ir, int jr, ZbRegion *R) {
    const int xMax = p->nx - 1, yMax = p->ny - 1;
    R->rx0 = ir * xMax / p->nreg_x;  R->ry0 = jr * yMax / p->nreg_y;
    R->rx1 = imin_((ir + 1) * xMax / p->nreg_x, xMax);  R->ry1 = imin_((jr + 1) * yMax / p->nreg_y, yMax);
    const int buf = s->n_regions > 1 ? s->fw_stamp / 2 : p->hw_kernel;     // main.c:623-635
    R->bx0 = imax_(0, R->rx0 - buf);  R->by0 = imax_(0, R->ry0 - buf);
    R->bx1 = imin_(xMax, R->rx1 + buf);  R->by1 = imin_(yMax, R->ry1 + buf);
    R->rpx = R->bx1 - R->bx0 + 1;  R->rpy = R->by1 - R->by0 + 1;
    R->own_x1 = R->rx1;  R->own_y1 = R->ry1;
}

// hotpants/main.c:382-428 (sizes) and :535-555 (region grid)
extern "C" int zb_sizes_of(const zb_params *p, zb_sizes *s) {
    memset(s, 0, sizeof *s);
    for (int i = 0; i < p->ngauss && i < ZB_MAX_GAUSS; i++) s->n_comp_ker += ((p->deg[i] + 1) * (p->deg[i] + 2)) / 2;
    s->n_comp = ((p->ker_order + 1) * (p->ker_order + 2)) / 2;
    s->n_bg_vectors = ((p->bg_order + 1) * (p->bg_order + 2)) / 2;
    s->n_comp_total = s->n_comp_ker * s->n_comp + s->n_bg_vectors;
    s->mat_size = (s->n_comp_ker - 1) * s->n_comp + s->n_bg_vectors + 1;
    s->fw_kernel = 2 * p->hw_kernel + 1;
    s->fw_ksstamp = 2 * p->hw_ksstamp + 1;
    s->nstamp_x = p->nstamp_x; s->nstamp_y = p->nstamp_y;
    if (p->nreg_x < 1 || p->nreg_y < 1 || s->nstamp_x < 1 || s->nstamp_y < 1) return ZB_ERR_PARAM;
    s->fw_stamp = imin_(p->nx / p->nreg_x / s->nstamp_x, p->ny / p->nreg_y / s->nstamp_y);
    s->fw_stamp -= s->fw_kernel;
    s->fw_stamp -= (s->fw_stamp % 2 == 0) ? 1 : 0;
    if (s->fw_stamp < s->fw_ksstamp) {
        s->fw_stamp = s->fw_ksstamp + s->fw_kernel;
        s->fw_stamp -= (s->fw_stamp % 2 == 0) ? 1 : 0;
        s->nstamp_x = p->nx / p->nreg_x / s->fw_stamp;
        s->nstamp_y = p->ny / p->nreg_y / s->fw_stamp;
    }
    s->kc_step = p->kc_step ? p->kc_step : s->fw_kernel;
    s->n_stamps = s->nstamp_x * s->nstamp_y;
    s->s_border = p->hw_ksstamp + p->hw_kernel;
    s->n_regions = p->nreg_x * p->nreg_y;
    for (int j = 0; j < p->nreg_y; j++)
        for (int i = 0; i < p->nreg_x; i++) {
            ZbRegion R; region_bounds(p, s, i, j, &R);
            s->max_rpix_x = imax_(s->max_rpix_x, R.rpx);
            s->max_rpix_y = imax_(s->max_rpix_y, R.rpy);
        }
    return ZB_OK;
}

// Numerical Recipes ran1 (the generator hotpants/functions.c:1632-1662 uses), reseeded with -666 by
// every getStampStats3 call, so one precomputed stream serves all calls.
void zb_host_ran1_stream(double *out, int n) {
    const long M1 = 259200, IA1 = 7141, IC1 = 54773, M2 = 134456, IA2 = 8121, IC2 = 28411, M3 = 243000, IA3 = 4561, IC3 = 51349;
    const double RM1 = 1.0 / 259200, RM2 = 1.0 / 134456;
    double r[98]; long ix1, ix2, ix3; const int idum = -666;
    ix1 = (IC1 - idum) % M1;  ix1 = (IA1 * ix1 + IC1) % M1;  ix2 = ix1 % M2;
    ix1 = (IA1 * ix1 + IC1) % M1;  ix3 = ix1 % M3;
    for (int j = 1; j <= 97; j++) { ix1 = (IA1 * ix1 + IC1) % M1; ix2 = (IA2 * ix2 + IC2) % M2; r[j] = (ix1 + ix2 * RM2) * RM1; }
    for (int k = 0; k < n; k++) {
        ix1 = (IA1 * ix1 + IC1) % M1;  ix2 = (IA2 * ix2 + IC2) % M2;  ix3 = (IA3 * ix3 + IC3) % M3;
        int j = 1 + (int)((97 * ix3) / M3);
        out[k] = r[j];
        r[j] = (ix1 + ix2 * RM2) * RM1;
    }
}

// alard.c:134-197 kernel_vector for every basis (host, once per configuration; ~50 x 441 doubles)
void zb_host_basis(const zb_params *p, double *kvec, double *fx, double *fy, int *ren) {
    const int hw = p->hw_kernel, fw = 2 * hw + 1;
    int n = 0;
    for (int g = 0; g < p->ngauss; g++)
        for (int px = 0; px <= p->deg[g]; px++)
            for (int py = 0; py <= p->deg[g] - px; py++, n++) {
                double sx = 0.0, sy = 0.0, *v = kvec + (size_t)n * fw * fw;
                const bool even = (px % 2 == 0) && (py % 2 == 0);
                for (int ix = 0; ix < fw; ix++) {
                    double x = (double)(ix - hw), qe = exp(-x * x * p->beta[g]);
                    fx[ix + n * fw] = qe * pow(x, px);
                    fy[ix + n * fw] = qe * pow(x, py);
                    sx += fx[ix + n * fw]; sy += fy[ix + n * fw];
                }
                sx = 1. / sx; sy = 1. / sy;
                if (even) for (int ix = 0; ix < fw; ix++) { fx[ix + n * fw] *= sx; fy[ix + n * fw] *= sy; }
                for (int i = 0; i < fw; i++) for (int j = 0; j < fw; j++) v[i + fw * j] = fx[i + n * fw] * fy[j + n * fw];
                ren[n] = (even && n > 0) ? 1 : 0;
                if (ren[n]) for (int i = 0; i < fw * fw; i++) v[i] -= kvec[i];
            }
}

// ------------------------------------------------------------------ context

extern "C" int zb_create(int device, zb_ctx **out) {
    if (!out) return ZB_ERR_PARAM;
    *out = NULL;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev <= 0 || device < 0 || device >= ndev) return ZB_ERR_CUDA;
    zb_ctx *ctx = new (std::nothrow) zb_ctx;
    if (!ctx) return ZB_ERR_NOMEM;
    memset(ctx, 0, sizeof *ctx);
    ctx->device = device;
    if (cudaSetDevice(device) != cudaSuccess || cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) { delete ctx; return ZB_ERR_CUDA; }
    for (int i = 0; i <= ZB_NSTAGES; i++) cudaEventCreate(&ctx->ev[i]);
    if (cudaDeviceGetAttribute(&ctx->n_sm, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || ctx->n_sm <= 0) ctx->n_sm = 148;
    if (cudaEventCreateWithFlags(&ctx->ev_sync, cudaEventBlockingSync | cudaEventDisableTiming) != cudaSuccess) ctx->ev_sync = NULL;
    *out = ctx;
    return ZB_OK;
}

static void free_workspace(zb_ctx *ctx) {
    if (ctx->lu_exec) { cudaGraphExecDestroy(ctx->lu_exec); ctx->lu_exec = NULL; }      // its nodes hold workspace pointers
    void **dptrs[] = {(void **)&ctx->d_dev, (void **)&ctx->d_min, (void **)&ctx->d_omask, (void **)&ctx->d_rdiff, (void **)&ctx->d_rnoise,
                      (void **)&ctx->d_ran, (void **)&ctx->d_fx, (void **)&ctx->d_fy, (void **)&ctx->d_kvec, (void **)&ctx->d_ren,
                      (void **)&ctx->d_cells, (void **)&ctx->d_stamps, (void **)&ctx->d_vec, (void **)&ctx->d_kref, (void **)&ctx->d_gram,
                      (void **)&ctx->d_rhs, (void **)&ctx->d_wxy, (void **)&ctx->d_mat, (void **)&ctx->d_sol, (void **)&ctx->d_vv,
                      (void **)&ctx->d_indx, (void **)&ctx->d_fit, (void **)&ctx->d_jobs, (void **)&ctx->d_njobs, (void **)&ctx->d_kaks,
                      (void **)&ctx->d_ktap, (void **)&ctx->d_tilectr,
                      (void **)&ctx->d_kstat, (void **)&ctx->d_stats, (void **)&ctx->d_qa, (void **)&ctx->d_qpart,
                      (void **)&ctx->d_qshell, (void **)&ctx->d_qshell_n, (void **)&ctx->d_check};
    for (size_t i = 0; i < sizeof dptrs / sizeof dptrs[0]; i++) { if (*dptrs[i]) cudaFree(*dptrs[i]); *dptrs[i] = NULL; }
    void **hptrs[] = {(void **)&ctx->h_dev, (void **)&ctx->h_cells, (void **)&ctx->h_stamps, (void **)&ctx->h_fit, (void **)&ctx->h_njobs, (void **)&ctx->h_stats, (void **)&ctx->h_check};
    for (size_t i = 0; i < sizeof hptrs / sizeof hptrs[0]; i++) { if (*hptrs[i]) cudaFreeHost(*hptrs[i]); *hptrs[i] = NULL; }
    ctx->configured = 0;
}

extern "C" void zb_destroy(zb_ctx *ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    free_workspace(ctx);
    for (int i = 0; i < 9; i++) if (ctx->d_stage[i]) cudaFree(ctx->d_stage[i]);
    for (int i = 0; i < 9; i++) if (ctx->d_stage_b[i]) cudaFree(ctx->d_stage_b[i]);
    if (ctx->batch_ready) {
        cudaStreamDestroy(ctx->st_h2d); cudaStreamDestroy(ctx->st_d2h);
        for (int i = 0; i < 2; i++) { cudaEventDestroy(ctx->ev_h2d[i]); cudaEventDestroy(ctx->ev_d2h[i]); cudaEventDestroy(ctx->ev_done[i]); }
    }
    for (int i = 0; i < 2; i++) if (ctx->d_ped[i]) cudaFree(ctx->d_ped[i]);
    if (ctx->d_search_scratch) cudaFree(ctx->d_search_scratch);
    if (ctx->d_bkg) cudaFree(ctx->d_bkg);
    if (ctx->h_bkg) cudaFreeHost(ctx->h_bkg);
    for (int i = 0; i <= ZB_NSTAGES; i++) cudaEventDestroy(ctx->ev[i]);
    if (ctx->ev_sync) cudaEventDestroy(ctx->ev_sync);
    cudaStreamDestroy(ctx->stream);
    delete ctx;
}

extern "C" const char *zb_last_error(const zb_ctx *ctx) { return ctx ? ctx->err : "no context"; }
extern "C" long long zb_launch_count(const zb_ctx *ctx) { return ctx ? ctx->launches : 0; }

static const char *k_stage_names[ZB_NSTAGES] = {"input_mask", "stamp_search", "stamp_resolve", "fill_stamps", "fit_loop",
                                                "make_kernels+convolve", "final_sig", "region_stats", "readback", "k_convolve", "search_stats_kclk", "search_scan_kclk"};
extern "C" const char *zb_stage_name(int i) { return (i >= 0 && i < ZB_NSTAGES) ? k_stage_names[i] : ""; }
extern "C" int zb_last_timings(const zb_ctx *ctx, float *ms, int n) {
    if (!ctx || !ms) return ZB_ERR_PARAM;
    for (int i = 0; i < n && i < ZB_NSTAGES; i++) ms[i] = ctx->ms[i];
    return ZB_OK;
}

#define DMALLOC(ptr, bytes) do { if (cudaMalloc((void **)&(ptr), (size_t)(bytes) ? (size_t)(bytes) : 16) != cudaSuccess) { \
    snprintf(ctx->err, sizeof ctx->err, "cudaMalloc of %zu bytes failed (%s)", (size_t)(bytes), #ptr); free_workspace(ctx); return ZB_ERR_NOMEM; } } while (0)
#define HMALLOC(ptr, bytes) do { if (cudaMallocHost((void **)&(ptr), (size_t)(bytes) ? (size_t)(bytes) : 16) != cudaSuccess) { \
    snprintf(ctx->err, sizeof ctx->err, "cudaMallocHost of %zu bytes failed (%s)", (size_t)(bytes), #ptr); free_workspace(ctx); return ZB_ERR_NOMEM; } } while (0)

// The parameters that size the workspace or enter the precomputed tables (basis, region grid).
// Everything else (thresholds, gains, clipping constants, directions) only travels in ZbDev.
static bool same_geometry(const zb_params *a, const zb_params *b) {
    if (a->nx != b->nx || a->ny != b->ny || a->hw_kernel != b->hw_kernel || a->hw_ksstamp != b->hw_ksstamp ||
        a->n_ksstamps != b->n_ksstamps || a->ker_order != b->ker_order || a->bg_order != b->bg_order ||
        a->nreg_x != b->nreg_x || a->nreg_y != b->nreg_y || a->nstamp_x != b->nstamp_x || a->nstamp_y != b->nstamp_y ||
        a->kc_step != b->kc_step || a->ngauss != b->ngauss || (a->force_convolve == 'b') != (b->force_convolve == 'b')) return false;
    for (int i = 0; i < a->ngauss && i < ZB_MAX_GAUSS; i++)
        if (a->deg[i] != b->deg[i] || a->beta[i] != b->beta[i]) return false;
    return true;
}

static void set_scalars(zb_ctx *ctx, const zb_params *p) {
    ZbDev *H = ctx->h_dev;
    ctx->p = *p;
    H->p = *p;
    H->conv_template = (p->force_convolve == 't');
    H->tU = p->t_uthresh - p->t_pedestal; H->tUK = p->t_ukthresh - p->t_pedestal; H->tL = p->t_lthresh - p->t_pedestal;
    H->iU = p->i_uthresh - p->i_pedestal; H->iUK = p->i_ukthresh - p->i_pedestal; H->iL = p->i_lthresh - p->i_pedestal;
}

// (Re)build the workspace for a parameter set.  Pairs of one survey share the geometry and differ
// only in thresholds, so the usual call just refreshes the scalars.
int zb_configure(zb_ctx *ctx, const zb_params *p) {
    if (ctx->configured && memcmp(&ctx->p, p, sizeof *p) == 0) return ZB_OK;
    int rc = zb_params_check(p, ctx->err, (int)sizeof ctx->err);
    if (rc != ZB_OK) return rc;
    if (ctx->configured && same_geometry(&ctx->p, p)) { set_scalars(ctx, p); return ZB_OK; }
    cudaStreamSynchronize(ctx->stream);
    free_workspace(ctx);
    ctx->p = *p;
    zb_sizes_of(p, &ctx->sz);
    const zb_sizes &sz = ctx->sz;
    HMALLOC(ctx->h_dev, sizeof(ZbDev));
    ZbDev *H = ctx->h_dev;
    memset(H, 0, sizeof *H);
    set_scalars(ctx, p);
    H->sz = sz; H->nx = p->nx; H->ny = p->ny; H->n_regions = sz.n_regions;
    long long off = 0; int nR = 0;
    for (int j = 0; j < p->nreg_y; j++)
        for (int i = 0; i < p->nreg_x; i++, nR++) {
            ZbRegion &R = H->reg[nR];
            region_bounds(p, &sz, i, j, &R);
            R.off = off; off += (long long)R.rpx * R.rpy;
            R.stamp_base = nR * sz.n_stamps;
        }
    ctx->region_elems = off;
    DMALLOC(ctx->d_dev, sizeof(ZbDev));
    DMALLOC(ctx->d_min, off * sizeof(int32_t));
    DMALLOC(ctx->d_omask, off * sizeof(int32_t));
    DMALLOC(ctx->d_rdiff, off * sizeof(float));
    DMALLOC(ctx->d_rnoise, off * sizeof(float));
    // ran1 stream: getStampStats3 draws at most 2 numbers per pixel of the area it is called on
    long long maxarea = imax_(sz.fw_stamp * sz.fw_stamp, sz.max_rpix_x * sz.max_rpix_y);
    ctx->n_ran = (int)(2 * maxarea + 2);
    {
        double *tmp = (double *)malloc((size_t)ctx->n_ran * sizeof(double));
        if (!tmp) return ZB_ERR_NOMEM;
        zb_host_ran1_stream(tmp, ctx->n_ran);
        DMALLOC(ctx->d_ran, (size_t)ctx->n_ran * sizeof(double));
        cudaMemcpyAsync(ctx->d_ran, tmp, (size_t)ctx->n_ran * sizeof(double), cudaMemcpyHostToDevice, ctx->stream);
        cudaStreamSynchronize(ctx->stream);
        free(tmp);
    }
    const int fw = sz.fw_kernel, nK = sz.n_comp_ker, nvec = nK + sz.n_bg_vectors, npix = sz.fw_ksstamp * sz.fw_ksstamp;
    {
        double *kv = (double *)malloc((size_t)nK * fw * fw * sizeof(double));
        double *fx = (double *)malloc((size_t)nK * fw * sizeof(double)), *fy = (double *)malloc((size_t)nK * fw * sizeof(double));
        int *ren = (int *)malloc((size_t)nK * sizeof(int));
        zb_host_basis(p, kv, fx, fy, ren);
        DMALLOC(ctx->d_kvec, (size_t)nK * fw * fw * sizeof(double));
        DMALLOC(ctx->d_fx, (size_t)nK * fw * sizeof(double));
        DMALLOC(ctx->d_fy, (size_t)nK * fw * sizeof(double));
        DMALLOC(ctx->d_ren, (size_t)nK * sizeof(int));
        cudaMemcpyAsync(ctx->d_kvec, kv, (size_t)nK * fw * fw * sizeof(double), cudaMemcpyHostToDevice, ctx->stream);
        cudaMemcpyAsync(ctx->d_fx, fx, (size_t)nK * fw * sizeof(double), cudaMemcpyHostToDevice, ctx->stream);
        cudaMemcpyAsync(ctx->d_fy, fy, (size_t)nK * fw * sizeof(double), cudaMemcpyHostToDevice, ctx->stream);
        cudaMemcpyAsync(ctx->d_ren, ren, (size_t)nK * sizeof(int), cudaMemcpyHostToDevice, ctx->stream);
        cudaStreamSynchronize(ctx->stream);
        free(kv); free(fx); free(fy); free(ren);
    }
    ctx->n_cells = sz.n_regions * sz.n_stamps;
    ctx->n_slots = sz.n_regions * sz.n_stamps;
    const size_t ns = (size_t)ctx->n_slots;
    // -c b searches both images: the science image's grid cells live behind the template's (cell index + n_cells)
    const size_t ncell_alloc = p->force_convolve == 'b' ? 2 * ns : ns;
    DMALLOC(ctx->d_cells, ncell_alloc * sizeof(ZbCell));   HMALLOC(ctx->h_cells, ncell_alloc * sizeof(ZbCell));
    if (p->force_convolve == 'b') {       // per-stamp kernel sums, per-region trial kernel sums, a copy of the regions' statistics blocks
        DMALLOC(ctx->d_check, (ns + (size_t)sz.n_regions * 65) * sizeof(double));
        HMALLOC(ctx->h_check, (ns + sz.n_regions) * sizeof(double));
    }
    DMALLOC(ctx->d_stamps, ns * sizeof(ZbStamp)); HMALLOC(ctx->h_stamps, ns * sizeof(ZbStamp));
    DMALLOC(ctx->d_vec, ns * nvec * npix * sizeof(double));
    DMALLOC(ctx->d_kref, ns * npix * sizeof(double));
    DMALLOC(ctx->d_gram, ns * (size_t)(nvec * (nvec + 1) / 2) * sizeof(double));
    DMALLOC(ctx->d_rhs, ns * nvec * sizeof(double));
    DMALLOC(ctx->d_wxy, ns * sz.n_comp * sizeof(double));
    DMALLOC(ctx->d_mat, (size_t)sz.n_regions * sz.mat_size * sz.mat_size * sizeof(double));
    DMALLOC(ctx->d_sol, (size_t)sz.n_regions * (sz.n_comp_total + 1) * sizeof(double));
    DMALLOC(ctx->d_vv, (size_t)sz.n_regions * sz.mat_size * sizeof(double));
    DMALLOC(ctx->d_indx, (size_t)sz.n_regions * sz.mat_size * sizeof(int));
    DMALLOC(ctx->d_fit, (size_t)sz.n_regions * sizeof(ZbFit));   HMALLOC(ctx->h_fit, (size_t)sz.n_regions * sizeof(ZbFit));
    DMALLOC(ctx->d_jobs, ns * sizeof(int));
    DMALLOC(ctx->d_njobs, sizeof(int));   HMALLOC(ctx->h_njobs, (4 + ns) * sizeof(int));     // [0]: job count; [4..]: job list scratch (zb_tile.cu)
    const int B = sz.kc_step, nbx = (sz.max_rpix_x + B - 1) / B, nby = (sz.max_rpix_y + B - 1) / B;
    DMALLOC(ctx->d_kaks, (size_t)sz.n_regions * ((size_t)nbx * nby + 1) * sizeof(double));
    DMALLOC(ctx->d_tilectr, 16);
    DMALLOC(ctx->d_ktap, (size_t)sz.n_regions * ((size_t)nbx * nby + 1) * fw * (fw + 8) * 16);
    DMALLOC(ctx->d_kstat, (size_t)sz.n_regions * 4 * sizeof(double));
    const size_t stat_bytes = (size_t)sz.n_regions * ((size_t)sz.n_comp_total + 512) * sizeof(double) + ns * 8;
    DMALLOC(ctx->d_stats, stat_bytes);   HMALLOC(ctx->h_stats, stat_bytes);
    cudaMemsetAsync(ctx->d_stats, 0, stat_bytes, ctx->stream);
    DMALLOC(ctx->d_qa, (size_t)sz.n_regions * 4 * sizeof(ZbQaStat));
    DMALLOC(ctx->d_qpart, (size_t)sz.n_regions * 5 * ZB_QA_NBLK * sizeof(ZbQaPart));
    {   // clip-shell lists (zb_stats.cu): one segment per warp of every k_qa_pass CTA and statistic
        const int nblk = zb_qa_nblk(sz.n_regions);
        const long long per_warp = ((long long)sz.max_rpix_x * sz.max_rpix_y + (long long)nblk * 8 - 1) / ((long long)nblk * 8);
        ctx->qshell_cap = (int)(per_warp / ZB_QA_SHELL_DIV) + 32;
        DMALLOC(ctx->d_qshell, (size_t)sz.n_regions * 4 * nblk * 8 * ctx->qshell_cap * sizeof(float));
        DMALLOC(ctx->d_qshell_n, (size_t)sz.n_regions * 4 * nblk * 8 * sizeof(int));
    }
    zb_copy_small(ctx, ctx->d_dev, H, sizeof(ZbDev), ctx->stream);
    cudaError_t e = cudaStreamSynchronize(ctx->stream);
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e != cudaSuccess) { free_workspace(ctx); return zb_cuda_fail(ctx, e, "configure", __FILE__, __LINE__); }
    ctx->configured = 1;
    return ZB_OK;
}

// ------------------------------------------------------------- orchestration

void zb_init_cells(zb_ctx *ctx, int r, float kft, int cell_base) {
    const zb_sizes &sz = ctx->sz;
    const ZbRegion &R = ctx->h_dev->reg[r];
    for (int gl = 0; gl < sz.nstamp_y; gl++)
        for (int gk = 0; gk < sz.nstamp_x; gk++) {
            ZbCell &C = ctx->h_cells[(size_t)cell_base + (size_t)r * sz.n_stamps + gl * sz.nstamp_x + gk];
            memset(&C, 0, sizeof C);
            // main.c:849-852, image coordinates
            int sXMin = R.bx0 + gk * R.rpx / sz.nstamp_x, sYMin = R.by0 + gl * R.rpy / sz.nstamp_y;
            int sXMax = imin_(sXMin + sz.fw_stamp - 1, R.bx1), sYMax = imin_(sYMin + sz.fw_stamp - 1, R.by1);
            C.region = r; C.gk = gk; C.gl = gl;
            C.sx0 = sXMin - R.bx0; C.sy0 = sYMin - R.by0;
            C.spx = sXMax - sXMin + 1; C.spy = sYMax - sYMin + 1;
            C.kft = kft;
        }
}

// Stale-slot quirk: a grid cell that found nothing leaves its stamp slot to the next
// cell, and getStampStats3 can return early without overwriting mode/fwhm.  Cells are searched in
// parallel assuming a fresh slot; the rare cell whose statistics bailed out while the slot was NOT
// fresh is searched again with the inherited values.  Returns the number of kept stamps.
static int resolve_region(zb_ctx *ctx, int r, const float *conv_img, cudaStream_t st, int *rc_out, int cell_base = 0, int dir_t = -1) {
    const zb_sizes &sz = ctx->sz;
    double s_mode = 0, s_fwhm = 0, s_mean = 0, s_sd = 0, s_median = 0, s_lfwhm = 0, s_sum = 0;
    int kept = 0;
    for (int c = 0; c < sz.n_stamps; c++) {
        const int ci = cell_base + r * sz.n_stamps + c;
        ZbCell *C = &ctx->h_cells[ci];
        const bool depends = (C->stat_status == 1 || C->stat_status == 4 || C->stat_status == 5);
        if (depends && (C->in_mode != s_mode || C->in_fwhm != s_fwhm)) {
            float kft = C->kft; int gk = C->gk, gl = C->gl, sx0 = C->sx0, sy0 = C->sy0, spx = C->spx, spy = C->spy;
            memset(C, 0, sizeof *C);
            C->region = r; C->gk = gk; C->gl = gl; C->sx0 = sx0; C->sy0 = sy0; C->spx = spx; C->spy = spy; C->kft = kft;
            C->in_mode = s_mode; C->in_fwhm = s_fwhm; C->mean = s_mean; C->sd = s_sd; C->median = s_median; C->lfwhm = s_lfwhm; C->sum = s_sum;
            zb_copy_small(ctx, &ctx->d_cells[ci], C, sizeof(ZbCell), st);
            int rc = zb_launch_stamp_search(ctx, conv_img, ci, 1, st, dir_t);
            if (rc != ZB_OK) { *rc_out = rc; return 0; }
            zb_copy_small(ctx, C, &ctx->d_cells[ci], sizeof(ZbCell), st);
            if (zb_stream_wait(ctx, st) != cudaSuccess) { *rc_out = zb_cuda_fail(ctx, cudaGetLastError(), "stamp re-search", __FILE__, __LINE__); return 0; }
        }
        if (C->nss > 0) { kept++; s_mode = s_fwhm = s_mean = s_sd = s_median = s_lfwhm = s_sum = 0; }
        else { s_mode = C->mode; s_fwhm = C->fwhm; s_mean = C->mean; s_sd = C->sd; s_median = C->median; s_lfwhm = C->lfwhm; s_sum = C->sum; }
    }
    *rc_out = ZB_OK;
    return kept;
}

#define STAGE(i) cudaEventRecord(ctx->ev[i], st)
#define CHECK_RC(x) do { int rc_ = (x); if (rc_ != ZB_OK) return rc_; } while (0)
#define SYNC(what) do { cudaError_t e_ = zb_stream_wait(ctx, st); if (e_ != cudaSuccess) return zb_cuda_fail(ctx, e_, what, __FILE__, __LINE__); } while (0)

// Stage 8 of a subtraction: the small device-to-host read-backs and the per-region record (main.c:1719-1953 keywords)
int zb_finish_readback(zb_ctx *ctx, zb_region_info *regions, double *kernel_sol, zb_stamp_info *stamps_out, int timing /* 0 none, 1 stages, 2 stages + k_convolve */, cudaStream_t st)
{
    const zb_sizes &sz = ctx->sz;
    ZbDev *H = ctx->h_dev;
    const int nR = sz.n_regions;
    zb_copy_small(ctx, ctx->h_stamps, ctx->d_stamps, (size_t)ctx->n_slots * sizeof(ZbStamp), st);
    zb_copy_small(ctx, ctx->h_stats, ctx->d_stats, (size_t)nR * 64 * sizeof(double), st);
    double *h_sol = (double *)((char *)ctx->h_stats + (size_t)nR * 64 * sizeof(double));
    zb_copy_small(ctx, h_sol, ctx->d_sol, (size_t)nR * (sz.n_comp_total + 1) * sizeof(double), st);
    double *h_ks = h_sol + (size_t)nR * (sz.n_comp_total + 1);
    zb_copy_small(ctx, h_ks, ctx->d_kstat, (size_t)nR * 4 * sizeof(double), st);
    STAGE(9);
    SYNC("final read-back");
    if (timing >= 1) for (int i = 0; i < 9; i++) cudaEventElapsedTime(&ctx->ms[i], ctx->ev[i], ctx->ev[i + 1]);
    if (timing >= 2) cudaEventElapsedTime(&ctx->ms[9], ctx->ev[10], ctx->ev[11]);
    for (int r = 0; r < nR; r++) {
        zb_region_info *ri = &regions[r];
        const ZbFit &F = ctx->h_fit[r];
        const double *q = ctx->h_stats + (size_t)r * 64;
        if (kernel_sol) {
            double *dst = kernel_sol + (size_t)r * (sz.n_comp_total + 1);
            if (H->reg[r].skip) memset(dst, 0, (size_t)(sz.n_comp_total + 1) * sizeof(double));
            else { memcpy(dst, h_sol + (size_t)r * (sz.n_comp_total + 1), (size_t)(sz.n_comp_total + 1) * sizeof(double)); dst[0] = 0.0;
                   for (int k = sz.mat_size + 1; k <= sz.n_comp_total; k++) dst[k] = 0.0; }
        }
        if (stamps_out) {
            for (int k = 0; k < sz.n_stamps; k++) {
                zb_stamp_info *so = &stamps_out[(size_t)r * sz.n_stamps + k];
                memset(so, 0, sizeof *so);
                if (k >= F.n_stamps) continue;
                const ZbStamp &S = ctx->h_stamps[H->reg[r].stamp_base + k];
                so->x0 = S.x0; so->y0 = S.y0; so->nss = S.nss; so->sscnt = S.sscnt;
                for (int s = 0; s < S.nss; s++) { so->xss[s] = S.xss[s]; so->yss[s] = S.yss[s]; }
                so->mode = S.mode; so->fwhm = S.fwhm; so->chi2 = S.chi2;
            }
        }
        if (H->reg[r].skip) continue;
        if (F.singular) ri->flags |= ZB_REGION_SINGULAR;
        ri->n_skipped = F.n_skipped; ri->fit_iters = F.iters;
        ri->ksum = h_ks[(size_t)r * 4]; ri->sssig = F.meansig; ri->ssscat = F.scatter;
        ri->fsig = q[0]; ri->fscat = q[1]; ri->x2norm = q[2]; ri->nx2norm = (int)q[3];
        ri->dmean = q[4]; ri->dsige = q[5]; ri->dsig = q[6]; ri->dmeano = q[7]; ri->dsigeo = q[8]; ri->dsigo = q[9]; ri->nscalo = q[10];
    }
    return ZB_OK;
}


// -c b (hotpants' own default; zuds always passes -c t).  Both images are searched for stamps (the two searches use
// their own mask bits, functions.c:195, and with tuk >= tu / iuk >= iu leave nothing behind that the other could see),
// check_stamps gives each direction a merit per region (main.c:1026-1072, alard.c:483-724, here zb_check.cu), and the
// smaller merit decides which image that region convolves (main.c:1076-1085).  The regions are then taken through
// the fit and the output pass in two groups, one per direction, by the kernels of the fixed-direction path: while a
// group is processed the other group's regions are inactive (fit) / marked skipped (output pass).
static int subtract_both(zb_ctx *ctx, const zb_params *p,
                         const float *sci, const float *ref, const float *sci_rms, const float *ref_rms,
                         const void *sci_mask, int sci_mask_dtype, const void *ref_mask, int ref_mask_dtype,
                         float *diff, float *noise, int32_t *out_mask,
                         zb_region_info *regions, double *kernel_sol, zb_stamp_info *stamps_out, cudaStream_t st)
{
    const zb_sizes &sz = ctx->sz;
    ZbDev *H = ctx->h_dev;
    const int nR = sz.n_regions, nC = ctx->n_cells, nS = sz.n_stamps;
    // direction d: 0 = convolve the template (stamps searched on ref), 1 = convolve the science image
    const float *conv_of[2] = {ref, sci}, *ref_of[2] = {sci, ref};
    const int dir_t_of[2] = {1, 0}, base_of[2] = {0, nC};
    ctx->err[0] = 0;
    memset(ctx->ms, 0, sizeof ctx->ms);

    // ---- stage 0: input masks
    STAGE(0);
    for (int r = 0; r < nR; r++) { H->reg[r].skip = 0; H->reg[r].own_x1 = H->reg[r].rx1; H->reg[r].own_y1 = H->reg[r].ry1; }
    H->conv_template = 1;
    zb_copy_small(ctx, ctx->d_dev, H, sizeof(ZbDev), st);
    CHECK_RC(zb_launch_input_mask(ctx, sci, ref, sci_mask, sci_mask_dtype, ref_mask, ref_mask_dtype, st));

    // ---- stage 1: stamp statistics + sub-stamp search on both images
    STAGE(1);
    for (int d = 0; d < 2; d++) for (int r = 0; r < nR; r++) zb_init_cells(ctx, r, p->ker_fit_thresh, base_of[d]);
    zb_copy_small(ctx, ctx->d_cells, ctx->h_cells, (size_t)2 * nC * sizeof(ZbCell), st);
    for (int d = 0; d < 2; d++) CHECK_RC(zb_launch_stamp_search(ctx, conv_of[d], base_of[d], nC, st, dir_t_of[d]));
    zb_copy_small(ctx, ctx->h_cells, ctx->d_cells, (size_t)2 * nC * sizeof(ZbCell), st);
    SYNC("stamp search");

    // ---- stage 2: per direction compaction, stale-slot fix-ups; one retry of BOTH searches when either found too few (main.c:921-995)
    STAGE(2);
    std::vector<ZbStamp> tab[2];
    std::vector<int> cnt[2], dir(nR, 0), gone(nR, 0);
    for (int d = 0; d < 2; d++) { tab[d].assign((size_t)ctx->n_slots, ZbStamp()); memset(tab[d].data(), 0, tab[d].size() * sizeof(ZbStamp)); cnt[d].assign(nR, 0); }
    for (int r = 0; r < nR; r++) {
        zb_region_info *ri = &regions[r];
        const ZbRegion &R = H->reg[r];
        memset(ri, 0, sizeof *ri);
        ri->rx_min = R.rx0; ri->rx_max = R.rx1; ri->ry_min = R.ry0; ri->ry_max = R.ry1;
        ri->bx_min = R.bx0; ri->bx_max = R.bx1; ri->by_min = R.by0; ri->by_max = R.by1;
        int kept[2], rc = ZB_OK;
        for (int d = 0; d < 2; d++) { kept[d] = resolve_region(ctx, r, conv_of[d], st, &rc, base_of[d], dir_t_of[d]); CHECK_RC(rc); }
        if (((kept[0] / (float)nS) < p->min_frac_good_stamps || (kept[1] / (float)nS) < p->min_frac_good_stamps) && p->scale_fit_thresh < 1.f) {
            float kft = p->ker_fit_thresh;
            kft *= p->scale_fit_thresh;
            for (int d = 0; d < 2; d++) {
                zb_init_cells(ctx, r, kft, base_of[d]);
                const int c0 = base_of[d] + r * nS;
                zb_copy_small(ctx, &ctx->d_cells[c0], &ctx->h_cells[c0], (size_t)nS * sizeof(ZbCell), st);
                CHECK_RC(zb_launch_stamp_search(ctx, conv_of[d], c0, nS, st, dir_t_of[d]));
                zb_copy_small(ctx, &ctx->h_cells[c0], &ctx->d_cells[c0], (size_t)nS * sizeof(ZbCell), st);
            }
            SYNC("stamp search retry");
            for (int d = 0; d < 2; d++) { kept[d] = resolve_region(ctx, r, conv_of[d], st, &rc, base_of[d], dir_t_of[d]); CHECK_RC(rc); }
            ri->flags |= ZB_REGION_RETRIED;
        }
        for (int d = 0; d < 2; d++) {           // kept stamps, in grid order, are this direction's slots
            int k = 0;
            for (int c = 0; c < nS; c++) {
                const ZbCell &C = ctx->h_cells[(size_t)base_of[d] + (size_t)r * nS + c];
                if (C.nss <= 0) continue;
                ZbStamp &S = tab[d][R.stamp_base + k];
                S.region = r; S.x0 = C.sx0; S.y0 = C.sy0; S.nss = C.nss; S.sscnt = 0;
                for (int q = 0; q < C.nss; q++) { S.xss[q] = C.xss[q]; S.yss[q] = C.yss[q]; }
                S.mode = C.mode; S.fwhm = C.fwhm;
                k++;
            }
            cnt[d][r] = k;
        }
    }

    // ---- stage 3: fill both stamp sets, merit of each direction (main.c:1026-1072), the choice (main.c:1076-1085)
    STAGE(3);
    std::vector<double> merit[2];
    int *jobs = (int *)ctx->h_stats;          // pinned scratch
    for (int d = 0; d < 2; d++) {
        merit[d].assign(nR, 666.0);           // what check_stamps returns for a direction without stamps (alard.c:693-721)
        memcpy(ctx->h_stamps, tab[d].data(), (size_t)ctx->n_slots * sizeof(ZbStamp));
        int njobs = 0;
        for (int r = 0; r < nR; r++) {
            ZbFit &F = ctx->h_fit[r];
            memset(&F, 0, sizeof F);
            F.n_stamps = cnt[d][r];
            for (int k = 0; k < F.n_stamps; k++) jobs[njobs++] = H->reg[r].stamp_base + k;
        }
        if (njobs > 0) {
            zb_copy_small(ctx, ctx->d_stamps, ctx->h_stamps, (size_t)ctx->n_slots * sizeof(ZbStamp), st);
            zb_copy_small(ctx, ctx->d_jobs, jobs, (size_t)njobs * sizeof(int), st);
            CHECK_RC(zb_launch_fill(ctx, conv_of[d], ref_of[d], ctx->d_jobs, njobs, st));
            CHECK_RC(zb_check_stamps(ctx, sci, ref, sci_rms, ref_rms, merit[d].data(), st));
        }
        SYNC("check_stamps");
    }
    bool any_skip = false, has[2] = {false, false};
    for (int r = 0; r < nR; r++) {
        zb_region_info *ri = &regions[r];
        const ZbRegion &R = H->reg[r];
        ZbFit &F = ctx->h_fit[r];
        memset(&F, 0, sizeof F);
        for (int q = 0; q < nS; q++) memset(&ctx->h_stamps[R.stamp_base + q], 0, sizeof(ZbStamp));
        if (cnt[0][r] > 0 || cnt[1][r] > 0) {
            ri->t_merit = merit[0][r]; ri->i_merit = merit[1][r];
            dir[r] = (ri->t_merit < ri->i_merit) ? 0 : 1;
        }
        const int k = cnt[dir[r]][r];
        ri->conv_template = dir[r] == 0;
        ri->n_stamps = k;
        F.n_stamps = k;
        if (k == 0) { ri->flags |= ZB_REGION_SKIPPED; gone[r] = 1; any_skip = true; continue; }     // main.c:1002-1009
        memcpy(&ctx->h_stamps[R.stamp_base], &tab[dir[r]][R.stamp_base], (size_t)k * sizeof(ZbStamp));
        has[dir[r]] = true;
    }
    // last-writer-wins ownership of the shared region edges (main.c:548-551,655-658), skipping silent regions
    for (int j = 0; j < p->nreg_y; j++)
        for (int i = 0; i < p->nreg_x; i++) {
            ZbRegion &R = H->reg[j * p->nreg_x + i];
            R.skip = gone[j * p->nreg_x + i];
            R.own_x1 = R.rx1; R.own_y1 = R.ry1;
            if (i + 1 < p->nreg_x && !gone[j * p->nreg_x + i + 1]) R.own_x1 = R.rx1 - 1;
            if (j + 1 < p->nreg_y && !gone[(j + 1) * p->nreg_x + i]) R.own_y1 = R.ry1 - 1;
        }
    if (any_skip) {
        cudaMemsetAsync(diff, 0, (size_t)p->nx * p->ny * sizeof(float), st);
        cudaMemsetAsync(noise, 0, (size_t)p->nx * p->ny * sizeof(float), st);
        if (out_mask) cudaMemsetAsync(out_mask, 0, (size_t)p->nx * p->ny * sizeof(int32_t), st);
    }
    zb_copy_small(ctx, ctx->d_dev, H, sizeof(ZbDev), st);
    zb_copy_small(ctx, ctx->d_stamps, ctx->h_stamps, (size_t)ctx->n_slots * sizeof(ZbStamp), st);
    zb_copy_small(ctx, ctx->d_fit, ctx->h_fit, (size_t)nR * sizeof(ZbFit), st);
    SYNC("direction choice");

    // ---- stage 4: fill + global fit with sigma clipping (alard.c:316-381), one group of regions per direction
    STAGE(4);
    for (int d = 0; d < 2; d++) {
        if (!has[d]) continue;
        int njobs = 0;
        for (int r = 0; r < nR; r++) {
            if (gone[r] || dir[r] != d) continue;
            ctx->h_fit[r].active = 1;
            for (int k = 0; k < ctx->h_fit[r].n_stamps; k++) jobs[njobs++] = H->reg[r].stamp_base + k;
        }
        zb_copy_small(ctx, ctx->d_fit, ctx->h_fit, (size_t)nR * sizeof(ZbFit), st);
        zb_copy_small(ctx, ctx->d_jobs, jobs, (size_t)njobs * sizeof(int), st);
        CHECK_RC(zb_launch_fill(ctx, conv_of[d], ref_of[d], ctx->d_jobs, njobs, st));
        bool any_active = true;
        int guard = 0;
        while (any_active) {
            CHECK_RC(zb_launch_fit_iteration(ctx, conv_of[d], ref_of[d], sci, ref, sci_rms, ref_rms, st));
            zb_copy_small(ctx, ctx->h_fit, ctx->d_fit, (size_t)nR * sizeof(ZbFit), st);
            zb_copy_small(ctx, ctx->h_njobs, ctx->d_njobs, sizeof(int), st);
            SYNC("fit iteration");
            any_active = false;
            for (int r = 0; r < nR; r++) any_active |= ctx->h_fit[r].active != 0;
            if (ctx->h_njobs[0] > 0) CHECK_RC(zb_launch_fill(ctx, conv_of[d], ref_of[d], ctx->d_jobs, ctx->h_njobs[0], st));
            if (++guard > 4 * nS * ZB_MAX_KSS) { snprintf(ctx->err, sizeof ctx->err, "fit loop did not converge"); return ZB_ERR_INTERNAL; }
        }
    }

    // ---- stages 5-7: the fused output pass, FSIG and the region statistics, again one group per direction.  The
    // statistics kernels clear the result block of every region when they start, so the first group's is set aside.
    STAGE(5);
    double *d_qsave = ctx->d_check + ctx->n_slots + nR;
    const int last = has[1] ? 1 : 0;
    for (int d = 0; d < 2; d++) {
        if (!has[d]) continue;
        SYNC("direction group");                 // H is read when the copy kernel runs, not when it is queued
        H->conv_template = d == 0;
        for (int r = 0; r < nR; r++) H->reg[r].skip = gone[r] || dir[r] != d;
        zb_copy_small(ctx, ctx->d_dev, H, sizeof(ZbDev), st);
        CHECK_RC(zb_launch_convolve(ctx, sci, ref, sci_rms, ref_rms, sci_mask, sci_mask_dtype, ref_mask, ref_mask_dtype, diff, noise, out_mask, st));
        if (d == last) STAGE(6);
        CHECK_RC(zb_launch_final_sig(ctx, st));
        if (d == last) STAGE(7);
        CHECK_RC(zb_launch_region_stats(ctx, noise, out_mask, st));
        if (d == 0 && has[1]) zb_copy_small(ctx, d_qsave, ctx->d_stats, (size_t)nR * 64 * sizeof(double), st);
        if (d == 1 && has[0])
            for (int r = 0; r < nR; r++)
                if (!gone[r] && dir[r] == 0) zb_copy_small(ctx, ctx->d_stats + (size_t)r * 64, d_qsave + (size_t)r * 64, 64 * sizeof(double), st);
    }
    if (!has[0] && !has[1]) { STAGE(6); STAGE(7); }
    SYNC("output pass");
    for (int r = 0; r < nR; r++) H->reg[r].skip = gone[r];

    // ---- stage 8: small read-backs
    STAGE(8);
    return zb_finish_readback(ctx, regions, kernel_sol, stamps_out, (has[0] || has[1]) ? 2 : 1, st);
}

extern "C" int zb_subtract(zb_ctx *ctx, const zb_params *p,
                           const float *sci, const float *ref, const float *sci_rms, const float *ref_rms,
                           const void *sci_mask, int sci_mask_dtype, const void *ref_mask, int ref_mask_dtype,
                           float *diff, float *noise, int32_t *out_mask,
                           zb_region_info *regions, double *kernel_sol, zb_stamp_info *stamps_out, void *stream)
{
    if (!ctx || !p || !sci || !ref || !diff || !noise || !regions) { if (ctx) snprintf(ctx->err, sizeof ctx->err, "null argument"); return ZB_ERR_PARAM; }
    if (sci_mask && !(sci_mask_dtype == ZB_MASK_U8 || sci_mask_dtype == ZB_MASK_I16 || sci_mask_dtype == ZB_MASK_I32)) { snprintf(ctx->err, sizeof ctx->err, "bad sci_mask_dtype %d", sci_mask_dtype); return ZB_ERR_PARAM; }
    if (ref_mask && !(ref_mask_dtype == ZB_MASK_U8 || ref_mask_dtype == ZB_MASK_I16 || ref_mask_dtype == ZB_MASK_I32)) { snprintf(ctx->err, sizeof ctx->err, "bad ref_mask_dtype %d", ref_mask_dtype); return ZB_ERR_PARAM; }
    if (cudaSetDevice(ctx->device) != cudaSuccess) return zb_cuda_fail(ctx, cudaGetLastError(), "cudaSetDevice", __FILE__, __LINE__);
    CHECK_RC(zb_configure(ctx, p));
    cudaStream_t st = zb_pick_stream(ctx, stream);
    if (p->t_pedestal != 0.f || p->i_pedestal != 0.f)          // main.c:728-733; the thresholds moved with them in set_scalars
        CHECK_RC(zb_launch_pedestals(ctx, sci, ref, p->i_pedestal, p->t_pedestal, (long long)p->nx * p->ny, &sci, &ref, st));
    if (p->force_convolve == 'b')
        return subtract_both(ctx, p, sci, ref, sci_rms, ref_rms, sci_mask, sci_mask_dtype, ref_mask, ref_mask_dtype,
                             diff, noise, out_mask, regions, kernel_sol, stamps_out, st);
    const zb_sizes &sz = ctx->sz;
    ZbDev *H = ctx->h_dev;
    const int nR = sz.n_regions;
    const bool ct = H->conv_template != 0;
    const float *conv_img = ct ? ref : sci, *ref_img = ct ? sci : ref;
    ctx->err[0] = 0;
    memset(ctx->ms, 0, sizeof ctx->ms);
    float cell_kclk[2] = {0, 0};

    // ---- stage 0: input masks
    STAGE(0);
    for (int r = 0; r < nR; r++) { H->reg[r].skip = 0; H->reg[r].own_x1 = H->reg[r].rx1; H->reg[r].own_y1 = H->reg[r].ry1; }
    zb_copy_small(ctx, ctx->d_dev, H, sizeof(ZbDev), st);     // scalars of THIS call
    CHECK_RC(zb_launch_input_mask(ctx, sci, ref, sci_mask, sci_mask_dtype, ref_mask, ref_mask_dtype, st));

    // ---- stage 1: stamp statistics + sub-stamp search, all grid cells of all regions at once
    STAGE(1);
    for (int r = 0; r < nR; r++) zb_init_cells(ctx, r, p->ker_fit_thresh);
    zb_copy_small(ctx, ctx->d_cells, ctx->h_cells, (size_t)ctx->n_cells * sizeof(ZbCell), st);
    CHECK_RC(zb_launch_stamp_search(ctx, conv_img, 0, ctx->n_cells, st));
    zb_copy_small(ctx, ctx->h_cells, ctx->d_cells, (size_t)ctx->n_cells * sizeof(ZbCell), st);
    SYNC("stamp search");
    {   // diagnostics: mean SM kilo-cycles per grid cell in the two halves of k_stamp_search
        double a = 0, b = 0;
        for (int c = 0; c < ctx->n_cells; c++) { a += ctx->h_cells[c].clk_stats; b += ctx->h_cells[c].clk_search; }
        cell_kclk[0] = (float)(a / ctx->n_cells / 1e3); cell_kclk[1] = (float)(b / ctx->n_cells / 1e3);
    }

    // ---- stage 2: compaction, stale-slot fix-ups, one retry at a lower threshold (main.c:909-998)
    STAGE(2);
    for (int r = 0; r < nR; r++) {
        zb_region_info *ri = &regions[r];
        const ZbRegion &R = H->reg[r];
        memset(ri, 0, sizeof *ri);
        ri->rx_min = R.rx0; ri->rx_max = R.rx1; ri->ry_min = R.ry0; ri->ry_max = R.ry1;
        ri->bx_min = R.bx0; ri->bx_max = R.bx1; ri->by_min = R.by0; ri->by_max = R.by1;
        ri->conv_template = ct;
        int rc = ZB_OK;
        int kept = resolve_region(ctx, r, conv_img, st, &rc);
        CHECK_RC(rc);
        if ((kept / (float)sz.n_stamps) < p->min_frac_good_stamps && p->scale_fit_thresh < 1.f) {
            float kft = p->ker_fit_thresh;
            kft *= p->scale_fit_thresh;
            zb_init_cells(ctx, r, kft);
            const int c0 = r * sz.n_stamps;
            zb_copy_small(ctx, &ctx->d_cells[c0], &ctx->h_cells[c0], (size_t)sz.n_stamps * sizeof(ZbCell), st);
            CHECK_RC(zb_launch_stamp_search(ctx, conv_img, c0, sz.n_stamps, st));
            zb_copy_small(ctx, &ctx->h_cells[c0], &ctx->d_cells[c0], (size_t)sz.n_stamps * sizeof(ZbCell), st);
            SYNC("stamp search retry");
            kept = resolve_region(ctx, r, conv_img, st, &rc);
            CHECK_RC(rc);
            ri->flags |= ZB_REGION_RETRIED;
        }
        // kept stamps, in grid order, become this region's slots
        int k = 0;
        for (int c = 0; c < sz.n_stamps; c++) {
            const ZbCell &C = ctx->h_cells[(size_t)r * sz.n_stamps + c];
            if (C.nss <= 0) continue;
            ZbStamp &S = ctx->h_stamps[R.stamp_base + k];
            memset(&S, 0, sizeof S);
            S.region = r; S.x0 = C.sx0; S.y0 = C.sy0; S.nss = C.nss; S.sscnt = 0;
            for (int q = 0; q < C.nss; q++) { S.xss[q] = C.xss[q]; S.yss[q] = C.yss[q]; }
            S.mode = C.mode; S.fwhm = C.fwhm;
            k++;
        }
        for (int q = k; q < sz.n_stamps; q++) memset(&ctx->h_stamps[R.stamp_base + q], 0, sizeof(ZbStamp));
        ri->n_stamps = k;
        ctx->h_fit[r].n_stamps = k; ctx->h_fit[r].active = k > 0; ctx->h_fit[r].iters = 0; ctx->h_fit[r].singular = 0;
        ctx->h_fit[r].n_skipped = 0; ctx->h_fit[r].meansig = 0; ctx->h_fit[r].scatter = 0;
        if (k == 0) { ri->flags |= ZB_REGION_SKIPPED; H->reg[r].skip = 1; }
    }
    // last-writer-wins ownership of the shared region edges (main.c:548-551,655-658), skipping silent regions
    bool any_skip = false;
    for (int j = 0; j < p->nreg_y; j++)
        for (int i = 0; i < p->nreg_x; i++) {
            ZbRegion &R = H->reg[j * p->nreg_x + i];
            any_skip |= R.skip != 0;
            R.own_x1 = R.rx1; R.own_y1 = R.ry1;
            if (i + 1 < p->nreg_x && !H->reg[j * p->nreg_x + i + 1].skip) R.own_x1 = R.rx1 - 1;
            if (j + 1 < p->nreg_y && !H->reg[(j + 1) * p->nreg_x + i].skip) R.own_y1 = R.ry1 - 1;
        }
    if (any_skip) {
        // corner pixels shared diagonally and regions that stay silent: start from hotpants' zero-filled outputs
        cudaMemsetAsync(diff, 0, (size_t)p->nx * p->ny * sizeof(float), st);
        cudaMemsetAsync(noise, 0, (size_t)p->nx * p->ny * sizeof(float), st);
        if (out_mask) cudaMemsetAsync(out_mask, 0, (size_t)p->nx * p->ny * sizeof(int32_t), st);
    }
    zb_copy_small(ctx, ctx->d_dev, H, sizeof(ZbDev), st);
    zb_copy_small(ctx, ctx->d_stamps, ctx->h_stamps, (size_t)ctx->n_slots * sizeof(ZbStamp), st);
    zb_copy_small(ctx, ctx->d_fit, ctx->h_fit, (size_t)nR * sizeof(ZbFit), st);

    // ---- stage 3: fill every kept stamp (alard.c:50-132)
    STAGE(3);
    int njobs = 0;
    {
        int *jobs = (int *)ctx->h_stats;      // pinned scratch
        for (int r = 0; r < nR; r++) for (int k = 0; k < ctx->h_fit[r].n_stamps; k++) jobs[njobs++] = H->reg[r].stamp_base + k;
        zb_copy_small(ctx, ctx->d_jobs, jobs, (size_t)njobs * sizeof(int), st);
        CHECK_RC(zb_launch_fill(ctx, conv_img, ref_img, ctx->d_jobs, njobs, st));
    }

    // ---- stage 4: global fit with sigma clipping of stamps (alard.c:316-381)
    STAGE(4);
    bool any_active = njobs > 0;
    int guard = 0;
    while (any_active) {
        CHECK_RC(zb_launch_fit_iteration(ctx, conv_img, ref_img, sci, ref, sci_rms, ref_rms, st));
        zb_copy_small(ctx, ctx->h_fit, ctx->d_fit, (size_t)nR * sizeof(ZbFit), st);
        zb_copy_small(ctx, ctx->h_njobs, ctx->d_njobs, sizeof(int), st);
        SYNC("fit iteration");
        any_active = false;
        for (int r = 0; r < nR; r++) any_active |= ctx->h_fit[r].active != 0;
        if (ctx->h_njobs[0] > 0) CHECK_RC(zb_launch_fill(ctx, conv_img, ref_img, ctx->d_jobs, ctx->h_njobs[0], st));
        if (++guard > 4 * sz.n_stamps * ZB_MAX_KSS) { snprintf(ctx->err, sizeof ctx->err, "fit loop did not converge"); return ZB_ERR_INTERNAL; }
    }

    // ---- stage 5: convolution + everything downstream of it, fused
    STAGE(5);
    CHECK_RC(zb_launch_convolve(ctx, sci, ref, sci_rms, ref_rms, sci_mask, sci_mask_dtype, ref_mask, ref_mask_dtype, diff, noise, out_mask, st));
    STAGE(6);
    CHECK_RC(zb_launch_final_sig(ctx, st));
    STAGE(7);
    CHECK_RC(zb_launch_region_stats(ctx, noise, out_mask, st));

    // ---- stage 8: small read-backs
    STAGE(8);
    ctx->ms[10] = cell_kclk[0]; ctx->ms[11] = cell_kclk[1];
    return zb_finish_readback(ctx, regions, kernel_sol, stamps_out, njobs > 0 ? 2 : 1, st);
}

// ------------------------------------------------------------ host-buffer entry

int zb_stage_buf(zb_ctx *ctx, int i, size_t bytes) {
    if (ctx->stage_bytes[i] >= bytes && ctx->d_stage[i]) return ZB_OK;
    if (ctx->d_stage[i]) cudaFree(ctx->d_stage[i]);
    ctx->d_stage[i] = NULL; ctx->stage_bytes[i] = 0;
    if (cudaMalloc(&ctx->d_stage[i], bytes) != cudaSuccess) { snprintf(ctx->err, sizeof ctx->err, "cudaMalloc of %zu staging bytes failed", bytes); return ZB_ERR_NOMEM; }
    ctx->stage_bytes[i] = bytes;
    return ZB_OK;
}

static size_t mask_elem(int dt) { return dt == ZB_MASK_U8 ? 1 : (dt == ZB_MASK_I16 ? 2 : 4); }

// What the reference delivers through FITS files and a fork/exec (zuds/subtraction.py:162): host
// arrays in, host arrays out.  Copies ride the ctx stream; pinned caller buffers make them async.
extern "C" int zb_subtract_host(zb_ctx *ctx, const zb_params *p,
                                const float *sci, const float *ref, const float *sci_rms, const float *ref_rms,
                                const void *sci_mask, int sci_mask_dtype, const void *ref_mask, int ref_mask_dtype,
                                float *diff, float *noise, int32_t *out_mask,
                                zb_region_info *regions, double *kernel_sol, zb_stamp_info *stamps)
{
    if (!ctx || !p || !sci || !ref || !diff || !noise) return ZB_ERR_PARAM;
    if (cudaSetDevice(ctx->device) != cudaSuccess) return zb_cuda_fail(ctx, cudaGetLastError(), "cudaSetDevice", __FILE__, __LINE__);
    const size_t n = (size_t)p->nx * p->ny;
    cudaStream_t st = ctx->stream;
    const void *src[6] = {sci, ref, sci_rms, ref_rms, sci_mask, ref_mask};
    const size_t esz[6] = {4, 4, 4, 4, mask_elem(sci_mask_dtype), mask_elem(ref_mask_dtype)};
    for (int i = 0; i < 6; i++) {
        if (!src[i]) continue;
        CHECK_RC(zb_stage_buf(ctx, i, n * esz[i]));
        cudaMemcpyAsync(ctx->d_stage[i], src[i], n * esz[i], cudaMemcpyHostToDevice, st);
    }
    CHECK_RC(zb_stage_buf(ctx, 6, n * 4)); CHECK_RC(zb_stage_buf(ctx, 7, n * 4));
    if (out_mask) CHECK_RC(zb_stage_buf(ctx, 8, n * 4));
    int rc = zb_subtract(ctx, p, (const float *)ctx->d_stage[0], (const float *)ctx->d_stage[1],
                         sci_rms ? (const float *)ctx->d_stage[2] : NULL, ref_rms ? (const float *)ctx->d_stage[3] : NULL,
                         sci_mask ? ctx->d_stage[4] : NULL, sci_mask_dtype, ref_mask ? ctx->d_stage[5] : NULL, ref_mask_dtype,
                         (float *)ctx->d_stage[6], (float *)ctx->d_stage[7], out_mask ? (int32_t *)ctx->d_stage[8] : NULL,
                         regions, kernel_sol, stamps, st);
    if (rc != ZB_OK) return rc;
    cudaMemcpyAsync(diff, ctx->d_stage[6], n * 4, cudaMemcpyDeviceToHost, st);
    cudaMemcpyAsync(noise, ctx->d_stage[7], n * 4, cudaMemcpyDeviceToHost, st);
    if (out_mask) cudaMemcpyAsync(out_mask, ctx->d_stage[8], n * 4, cudaMemcpyDeviceToHost, st);
    cudaError_t e = zb_stream_wait(ctx, st);
    if (e != cudaSuccess) return zb_cuda_fail(ctx, e, "zb_subtract_host copies", __FILE__, __LINE__);
    return ZB_OK;
}

// ------------------------------------------------------------ batched host-buffer entry

static int stage_buf_set(zb_ctx *ctx, int set, int i, size_t bytes) {
    void **buf = set ? &ctx->d_stage_b[i] : &ctx->d_stage[i];
    size_t *have = set ? &ctx->stage_b_bytes[i] : &ctx->stage_bytes[i];
    if (*have >= bytes && *buf) return ZB_OK;
    if (*buf) cudaFree(*buf);
    *buf = NULL; *have = 0;
    if (cudaMalloc(buf, bytes) != cudaSuccess) { snprintf(ctx->err, sizeof ctx->err, "cudaMalloc of %zu staging bytes failed", bytes); return ZB_ERR_NOMEM; }
    *have = bytes;
    return ZB_OK;
}

// npairs subtractions of one geometry, software-pipelined inside the context: while pair n is computed, the inputs
// of pair n+1 travel to the other staging set on a copy stream and the products of pair n-1 travel back on a third
// (SURVEY.md 8(b)'s zb_subtract_batch).  One cudaMemcpyAsync per array.  Per-pair status in rcs[] (may be NULL);
// returns the first non-OK status after finishing the batch.
extern "C" int zb_subtract_batch(zb_ctx *ctx, int npairs, const zb_params *params, const zb_host_pair *pairs,
                                 zb_region_info *regions, double *kernel_sol, zb_stamp_info *stamps, int *rcs)
{
    if (!ctx || npairs < 0 || (npairs && (!params || !pairs))) { if (ctx) snprintf(ctx->err, sizeof ctx->err, "null argument"); return ZB_ERR_PARAM; }
    if (cudaSetDevice(ctx->device) != cudaSuccess) return zb_cuda_fail(ctx, cudaGetLastError(), "cudaSetDevice", __FILE__, __LINE__);
    if (!ctx->batch_ready) {
        if (cudaStreamCreateWithFlags(&ctx->st_h2d, cudaStreamNonBlocking) != cudaSuccess || cudaStreamCreateWithFlags(&ctx->st_d2h, cudaStreamNonBlocking) != cudaSuccess)
            return zb_cuda_fail(ctx, cudaGetLastError(), "copy streams", __FILE__, __LINE__);
        for (int i = 0; i < 2; i++) {
            cudaEventCreateWithFlags(&ctx->ev_h2d[i], cudaEventDisableTiming); cudaEventCreateWithFlags(&ctx->ev_d2h[i], cudaEventDisableTiming);
            cudaEventCreateWithFlags(&ctx->ev_done[i], cudaEventDisableTiming);
        }
        ctx->batch_ready = 1;
    }
    cudaStream_t st = ctx->stream;
    int first_bad = ZB_OK;
    // upload of pair i into staging set i & 1 (after the set's previous products have left and its previous inputs were consumed)
    auto upload = [&](int i) -> int {
        const int set = i & 1;
        const zb_params *p = &params[i];
        const zb_host_pair &h = pairs[i];
        if (!h.sci || !h.ref || !h.diff || !h.noise) { snprintf(ctx->err, sizeof ctx->err, "pair %d: null image pointer", i); return ZB_ERR_PARAM; }
        const size_t n = (size_t)p->nx * p->ny;
        const void *src[6] = {h.sci, h.ref, h.sci_rms, h.ref_rms, h.sci_mask, h.ref_mask};
        const size_t esz[6] = {4, 4, 4, 4, mask_elem(h.sci_mask_dtype), mask_elem(h.ref_mask_dtype)};
        void **buf = set ? ctx->d_stage_b : ctx->d_stage;
        for (int k = 0; k < 6; k++) if (src[k]) CHECK_RC(stage_buf_set(ctx, set, k, n * esz[k]));
        CHECK_RC(stage_buf_set(ctx, set, 6, n * 4)); CHECK_RC(stage_buf_set(ctx, set, 7, n * 4));
        if (h.out_mask) CHECK_RC(stage_buf_set(ctx, set, 8, n * 4));
        if (i >= 2) { cudaStreamWaitEvent(ctx->st_h2d, ctx->ev_done[set], 0); cudaStreamWaitEvent(ctx->st_h2d, ctx->ev_d2h[set], 0); }
        for (int k = 0; k < 6; k++) if (src[k]) cudaMemcpyAsync(buf[k], src[k], n * esz[k], cudaMemcpyHostToDevice, ctx->st_h2d);
        cudaEventRecord(ctx->ev_h2d[set], ctx->st_h2d);
        return ZB_OK;
    };
    if (npairs > 0) { int rc = upload(0); if (rc != ZB_OK) return rc; }
    for (int i = 0; i < npairs; i++) {
        const int set = i & 1;
        const zb_params *p = &params[i];
        const zb_host_pair &h = pairs[i];
        const size_t n = (size_t)p->nx * p->ny;
        void **buf = set ? ctx->d_stage_b : ctx->d_stage;
        if (i + 1 < npairs) { int rc = upload(i + 1); if (rc != ZB_OK) { if (rcs) rcs[i + 1] = rc; if (first_bad == ZB_OK) first_bad = rc; } }
        cudaStreamWaitEvent(st, ctx->ev_h2d[set], 0);
        zb_sizes sz; zb_sizes_of(p, &sz);
        int rc = zb_subtract(ctx, p, (const float *)buf[0], (const float *)buf[1],
                             h.sci_rms ? (const float *)buf[2] : NULL, h.ref_rms ? (const float *)buf[3] : NULL,
                             h.sci_mask ? buf[4] : NULL, h.sci_mask_dtype, h.ref_mask ? buf[5] : NULL, h.ref_mask_dtype,
                             (float *)buf[6], (float *)buf[7], h.out_mask ? (int32_t *)buf[8] : NULL,
                             regions + (size_t)i * sz.n_regions, kernel_sol ? kernel_sol + (size_t)i * sz.n_regions * (sz.n_comp_total + 1) : NULL,
                             stamps ? stamps + (size_t)i * sz.n_regions * sz.n_stamps : NULL, ZB_STREAM_OWN);
        if (rcs) rcs[i] = rc;
        cudaEventRecord(ctx->ev_done[set], st);
        if (rc != ZB_OK) { if (first_bad == ZB_OK) first_bad = rc; continue; }
        // zb_subtract returns with the products complete on `st`: send them home on the third stream
        cudaStreamWaitEvent(ctx->st_d2h, ctx->ev_done[set], 0);
        cudaMemcpyAsync(h.diff, buf[6], n * 4, cudaMemcpyDeviceToHost, ctx->st_d2h);
        cudaMemcpyAsync(h.noise, buf[7], n * 4, cudaMemcpyDeviceToHost, ctx->st_d2h);
        if (h.out_mask) cudaMemcpyAsync(h.out_mask, buf[8], n * 4, cudaMemcpyDeviceToHost, ctx->st_d2h);
        cudaEventRecord(ctx->ev_d2h[set], ctx->st_d2h);
    }
    cudaError_t e = zb_stream_wait(ctx, ctx->st_d2h);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->st_h2d);
    if (e != cudaSuccess) return zb_cuda_fail(ctx, e, "zb_subtract_batch copies", __FILE__, __LINE__);
    return first_bad;
}

// ABI self-check used by the loader tests: sizeof of the public structs as this library was built.
extern "C" int zb_abi_sizeof(int which) {
    switch (which) {
    case 0: return (int)sizeof(zb_params);
    case 1: return (int)sizeof(zb_sizes);
    case 2: return (int)sizeof(zb_region_info);
    case 3: return (int)sizeof(zb_stamp_info);
    case 4: return (int)sizeof(zb_tile_info);
    case 5: return (int)sizeof(zb_host_pair);
    default: return -1;
    }
}
