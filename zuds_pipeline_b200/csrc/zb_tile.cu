// zb_tile.cu — row-tile sessions: one oversized single-region frame (BASELINE.json configs[4]) split
// over `world` GPUs by image rows.  include/zb200.h documents the protocol; this file holds the
// ownership arithmetic and the per-phase entry points.  Every phase launches the SAME kernels
// zb_subtract launches (zb_mask/zb_stamps/zb_fit/zb_conv/zb_stats.cu) on the rank's share of rows,
// grid cells, stamp slots and convolution blocks, so the merged result equals zb_subtract on the
// whole frame bit for bit:
//   * stamps are independent work items (hotpants/main.c:832-906): a rank owns whole grid-cell rows;
//   * the global normal matrix (alard.c:843-1025) is rebuilt on every rank from the merged per-stamp
//     Gram matrices in the reference's stamp order — the all-reduce only adds zeros to each term;
//   * check_again (alard.c:1143-1247) and the stale-slot walk (main.c:909-998) are replicated;
//   * the convolution blocks stay anchored to frame coordinates (alard.c:1272-1278).
#include "zb_internal.h"
#include <stdlib.h>

static inline int imin_(int a, int b) { return a < b ? a : b; }
static inline int imax_(int a, int b) { return a < b ? b : a; }

#define CHECK_RC(x) do { int rc_ = (x); if (rc_ != ZB_OK) return rc_; } while (0)
#define SYNC(what) do { cudaError_t e_ = zb_stream_wait(ctx, st); if (e_ != cudaSuccess) return zb_cuda_fail(ctx, e_, what, __FILE__, __LINE__); } while (0)
#define NEED_TILE() do { if (!ctx || !ctx->tile.on) { if (ctx) snprintf(ctx->err, sizeof ctx->err, "no row-tile session (call zb_tile_begin first)"); return ZB_ERR_PARAM; } \
                         if (cudaSetDevice(ctx->device) != cudaSuccess) return zb_cuda_fail(ctx, cudaGetLastError(), "cudaSetDevice", __FILE__, __LINE__); } while (0)

// boundary b (0..world) of the three partitions; the frame is the region (nreg 1x1: bx0 = by0 = 0, rpy = ny)
static int cell_row_of(const zb_sizes &sz, int world, int b) { return (int)((long long)b * sz.nstamp_y / world); }
static int pix_row_of(const zb_params *p, const zb_sizes &sz, int world, int b) {
    if (b <= 0) return 0;
    if (b >= world) return p->ny;
    return (int)((long long)cell_row_of(sz, world, b) * p->ny / sz.nstamp_y);       // sYMin of the first own cell row (main.c:850)
}
static int blk_row_of(const zb_params *p, const zb_sizes &sz, int world, int b) {
    const int B = sz.kc_step, nby = (p->ny + B - 1) / B;
    if (b <= 0) return 0;
    if (b >= world) return nby;
    const int r = pix_row_of(p, sz, world, b) - p->hw_kernel;
    return imin_(nby, imax_(0, (r + B - 1) / B));
}
static int out_row_of(const zb_params *p, const zb_sizes &sz, int world, int b) {
    if (b <= 0) return 0;
    if (b >= world) return p->ny;
    return imin_(p->ny, p->hw_kernel + blk_row_of(p, sz, world, b) * sz.kc_step);
}

extern "C" int zb_tile_plan(const zb_params *p, int world, int rank, zb_tile_info *out) {
    if (!p || !out || world < 1 || rank < 0 || rank >= world) return ZB_ERR_PARAM;
    if (p->nreg_x != 1 || p->nreg_y != 1) return ZB_ERR_UNSUPPORTED;      // several regions shard by region/pair instead
    if (p->force_convolve == 'b') return ZB_ERR_UNSUPPORTED;               // row tiles fix the direction (-c t | i)
    if (p->t_pedestal != 0.f || p->i_pedestal != 0.f) return ZB_ERR_UNSUPPORTED;   // the caller's rows are used as they are
    zb_sizes sz;
    int rc = zb_sizes_of(p, &sz);
    if (rc != ZB_OK) return rc;
    memset(out, 0, sizeof *out);
    out->rank = rank; out->world = world;
    out->cell_row0 = cell_row_of(sz, world, rank);  out->cell_row1 = cell_row_of(sz, world, rank + 1);
    out->row0 = pix_row_of(p, sz, world, rank);     out->row1 = pix_row_of(p, sz, world, rank + 1);
    out->blk_row0 = blk_row_of(p, sz, world, rank); out->blk_row1 = blk_row_of(p, sz, world, rank + 1);
    out->out_row0 = out_row_of(p, sz, world, rank); out->out_row1 = out_row_of(p, sz, world, rank + 1);
    // rows needed: own sub-stamps reach rss + hw beyond the stamp rows (alard.c:230-279), the convolution hw
    // beyond the output rows (alard.c:1286-1306), and the mask dilation width/2 beyond either (functions.c:1353)
    const int width = (int)(p->hw_kernel * p->kf_spread_mask1), w2 = width > 0 ? width / 2 : 0;
    int lo = out->row0, hi = out->row1;
    if (out->cell_row1 > out->cell_row0) { lo = imin_(lo, out->row0 - sz.s_border); hi = imax_(hi, out->row1 + sz.s_border); }
    if (out->out_row1 > out->out_row0) { lo = imin_(lo, out->out_row0 - p->hw_kernel); hi = imax_(hi, out->out_row1 + p->hw_kernel); }
    out->in_row0 = imax_(0, lo - w2);  out->in_row1 = imin_(p->ny, hi + w2);
    out->halo = imax_(out->row0 - out->in_row0, out->in_row1 - out->row1);
    out->n_cells = sz.n_stamps;  out->cell_words = (int)(sizeof(ZbCell) / sizeof(int32_t));
    out->n_slots = sz.n_stamps;
    const int nvec = sz.n_comp_ker + sz.n_bg_vectors;
    out->gram_len = nvec * (nvec + 1) / 2;  out->rhs_len = nvec;
    return ZB_OK;
}

static int search_own(zb_ctx *ctx) {
    ZbTile &T = ctx->tile;
    cudaStream_t st = T.st;
    const float *conv_img = ctx->h_dev->conv_template ? T.ref : T.sci;
    zb_init_cells(ctx, 0, T.kft);
    zb_copy_small(ctx, ctx->d_cells, ctx->h_cells, (size_t)ctx->n_cells * sizeof(ZbCell), st);
    if (T.cell_hi > T.cell_lo) {
        CHECK_RC(zb_launch_stamp_search(ctx, conv_img, T.cell_lo, T.cell_hi - T.cell_lo, st));
        zb_copy_small(ctx, &ctx->h_cells[T.cell_lo], &ctx->d_cells[T.cell_lo], (size_t)(T.cell_hi - T.cell_lo) * sizeof(ZbCell), st);
    }
    return ZB_OK;
}

extern "C" int zb_tile_begin(zb_ctx *ctx, const zb_params *p, int world, int rank,
                             const float *sci, const float *ref, const float *sci_rms, const float *ref_rms,
                             const void *sci_mask, int sci_mask_dtype, const void *ref_mask, int ref_mask_dtype,
                             float *diff, float *noise, int32_t *out_mask, void *stream)
{
    if (!ctx || !p || !sci || !ref || !diff || !noise) { if (ctx) snprintf(ctx->err, sizeof ctx->err, "null argument"); return ZB_ERR_PARAM; }
    if (cudaSetDevice(ctx->device) != cudaSuccess) return zb_cuda_fail(ctx, cudaGetLastError(), "cudaSetDevice", __FILE__, __LINE__);
    ZbTile &T = ctx->tile;
    T.on = 0;
    int rc = zb_tile_plan(p, world, rank, &T.info);
    if (rc != ZB_OK) { snprintf(ctx->err, sizeof ctx->err, "row tiles need nreg_x = nreg_y = 1 and 0 <= rank < world"); return rc; }
    CHECK_RC(zb_configure(ctx, p));
    T.world = world; T.rank = rank;
    T.sci = sci; T.ref = ref; T.sci_rms = sci_rms; T.ref_rms = ref_rms;
    T.smask = sci_mask; T.sdt = sci_mask_dtype; T.tmask = ref_mask; T.tdt = ref_mask_dtype;
    T.diff = diff; T.noise = noise; T.out_mask = out_mask;
    T.st = zb_pick_stream(ctx, stream);
    T.kft = p->ker_fit_thresh; T.retried = 0; T.pending_cell = -1;
    T.cell_lo = T.info.cell_row0 * ctx->sz.nstamp_x; T.cell_hi = T.info.cell_row1 * ctx->sz.nstamp_x;
    T.slot_lo = T.slot_hi = 0; T.n_kept = 0;
    cudaStream_t st = T.st;
    ZbDev *H = ctx->h_dev;
    ctx->err[0] = 0;
    memset(ctx->ms, 0, sizeof ctx->ms);
    H->reg[0].skip = 0; H->reg[0].own_x1 = H->reg[0].rx1; H->reg[0].own_y1 = H->reg[0].ry1;
    zb_copy_small(ctx, ctx->d_dev, H, sizeof(ZbDev), st);
    CHECK_RC(zb_launch_input_mask(ctx, sci, ref, sci_mask, sci_mask_dtype, ref_mask, ref_mask_dtype, st, T.info.in_row0, T.info.in_row1));
    T.on = 1;
    return search_own(ctx);
}

// main.c:948-995: one retry of the whole grid at kerFitThresh * scaleFitThresh
extern "C" int zb_tile_begin_retry(zb_ctx *ctx) {
    NEED_TILE();
    ZbTile &T = ctx->tile;
    float kft = ctx->p.ker_fit_thresh;
    kft *= ctx->p.scale_fit_thresh;
    T.kft = kft; T.retried = 1; T.pending_cell = -1;
    return search_own(ctx);
}

extern "C" int zb_tile_cells_export(zb_ctx *ctx, int32_t *words) {
    NEED_TILE();
    if (!words) return ZB_ERR_PARAM;
    ZbTile &T = ctx->tile;
    cudaStream_t st = T.st;
    SYNC("stamp search");
    memset(words, 0, (size_t)ctx->n_cells * sizeof(ZbCell));
    if (T.cell_hi > T.cell_lo)
        memcpy((char *)words + (size_t)T.cell_lo * sizeof(ZbCell), &ctx->h_cells[T.cell_lo], (size_t)(T.cell_hi - T.cell_lo) * sizeof(ZbCell));
    return ZB_OK;
}

extern "C" int zb_tile_cells_import(zb_ctx *ctx, const int32_t *words) {
    NEED_TILE();
    if (!words) return ZB_ERR_PARAM;
    memcpy(ctx->h_cells, words, (size_t)ctx->n_cells * sizeof(ZbCell));
    return ZB_OK;
}

// The stale-slot walk of zb_host.cu:resolve_region, replicated: every rank sees the same merged records
// and takes the same decisions; a cell that must be searched again is handed back to the caller.
extern "C" int zb_tile_resolve(zb_ctx *ctx, int *action, int *cell) {
    NEED_TILE();
    if (!action || !cell) return ZB_ERR_PARAM;
    ZbTile &T = ctx->tile;
    const zb_sizes &sz = ctx->sz;
    const zb_params &p = ctx->p;
    ZbDev *H = ctx->h_dev;
    cudaStream_t st = T.st;
    *action = 0; *cell = -1; T.pending_cell = -1;
    double s[7] = {0, 0, 0, 0, 0, 0, 0};      // mode fwhm mean sd median lfwhm sum left in the slot by cells that found nothing
    int kept = 0;
    for (int c = 0; c < sz.n_stamps; c++) {
        const ZbCell *C = &ctx->h_cells[c];
        const bool depends = (C->stat_status == 1 || C->stat_status == 4 || C->stat_status == 5);
        if (depends && (C->in_mode != s[0] || C->in_fwhm != s[1])) {
            T.pending_cell = c; memcpy(T.pend, s, sizeof s);
            *action = 1; *cell = c;
            return ZB_OK;
        }
        if (C->nss > 0) { kept++; memset(s, 0, sizeof s); }
        else { s[0] = C->mode; s[1] = C->fwhm; s[2] = C->mean; s[3] = C->sd; s[4] = C->median; s[5] = C->lfwhm; s[6] = C->sum; }
    }
    if ((kept / (float)sz.n_stamps) < p.min_frac_good_stamps && p.scale_fit_thresh < 1.f && !T.retried) { *action = 2; return ZB_OK; }

    // compaction (main.c:909-946): kept stamps, in grid order, become the slots; a rank's slots are contiguous
    const ZbRegion &R = H->reg[0];
    zb_region_info *ri = &T.ri;
    memset(ri, 0, sizeof *ri);
    ri->rx_min = R.rx0; ri->rx_max = R.rx1; ri->ry_min = R.ry0; ri->ry_max = R.ry1;
    ri->bx_min = R.bx0; ri->bx_max = R.bx1; ri->by_min = R.by0; ri->by_max = R.by1;
    ri->conv_template = H->conv_template != 0;
    if (T.retried) ri->flags |= ZB_REGION_RETRIED;
    int k = 0;
    T.slot_lo = T.slot_hi = 0;
    for (int c = 0; c < sz.n_stamps; c++) {
        const ZbCell &C = ctx->h_cells[c];
        if (c == T.cell_lo) T.slot_lo = k;
        if (C.nss > 0) {
            ZbStamp &S = ctx->h_stamps[R.stamp_base + k];
            memset(&S, 0, sizeof S);
            S.region = 0; S.x0 = C.sx0; S.y0 = C.sy0; S.nss = C.nss; S.sscnt = 0;
            for (int q = 0; q < C.nss; q++) { S.xss[q] = C.xss[q]; S.yss[q] = C.yss[q]; }
            S.mode = C.mode; S.fwhm = C.fwhm;
            k++;
        }
        if (c + 1 == T.cell_hi) T.slot_hi = k;
    }
    if (T.cell_hi <= T.cell_lo) T.slot_lo = T.slot_hi = 0;
    for (int q = k; q < sz.n_stamps; q++) memset(&ctx->h_stamps[R.stamp_base + q], 0, sizeof(ZbStamp));
    T.n_kept = k;
    ri->n_stamps = k;
    ZbFit &F = ctx->h_fit[0];
    memset(&F, 0, sizeof F);
    F.n_stamps = k; F.active = k > 0;
    if (k == 0) {
        // no usable stamp anywhere (main.c:1002-1009): the frame stays as hotpants' zero-filled outputs
        ri->flags |= ZB_REGION_SKIPPED; H->reg[0].skip = 1;
        const size_t row_elems = (size_t)p.nx, r0 = (size_t)T.info.out_row0, nrow = (size_t)(T.info.out_row1 - T.info.out_row0);
        cudaMemsetAsync(T.diff + r0 * row_elems, 0, nrow * row_elems * sizeof(float), st);
        cudaMemsetAsync(T.noise + r0 * row_elems, 0, nrow * row_elems * sizeof(float), st);
        if (T.out_mask) cudaMemsetAsync(T.out_mask + r0 * row_elems, 0, nrow * row_elems * sizeof(int32_t), st);
    }
    zb_copy_small(ctx, ctx->d_dev, H, sizeof(ZbDev), st);
    zb_copy_small(ctx, ctx->d_stamps, ctx->h_stamps, (size_t)ctx->n_slots * sizeof(ZbStamp), st);
    zb_copy_small(ctx, ctx->d_fit, ctx->h_fit, sizeof(ZbFit), st);
    return ZB_OK;
}

extern "C" int zb_tile_research(zb_ctx *ctx, int cell) {
    NEED_TILE();
    ZbTile &T = ctx->tile;
    cudaStream_t st = T.st;
    if (cell < 0 || cell >= ctx->n_cells || cell != T.pending_cell) { snprintf(ctx->err, sizeof ctx->err, "cell %d is not the one zb_tile_resolve asked for", cell); return ZB_ERR_PARAM; }
    if (cell < T.cell_lo || cell >= T.cell_hi) return ZB_OK;          // not the owner: the merged records bring the result
    ZbCell *C = &ctx->h_cells[cell];
    const float kft = C->kft; const int gk = C->gk, gl = C->gl, sx0 = C->sx0, sy0 = C->sy0, spx = C->spx, spy = C->spy;
    memset(C, 0, sizeof *C);
    C->region = 0; C->gk = gk; C->gl = gl; C->sx0 = sx0; C->sy0 = sy0; C->spx = spx; C->spy = spy; C->kft = kft;
    C->in_mode = T.pend[0]; C->in_fwhm = T.pend[1]; C->mean = T.pend[2]; C->sd = T.pend[3]; C->median = T.pend[4]; C->lfwhm = T.pend[5]; C->sum = T.pend[6];
    zb_copy_small(ctx, &ctx->d_cells[cell], C, sizeof(ZbCell), st);
    CHECK_RC(zb_launch_stamp_search(ctx, ctx->h_dev->conv_template ? T.ref : T.sci, cell, 1, st));
    zb_copy_small(ctx, C, &ctx->d_cells[cell], sizeof(ZbCell), st);
    SYNC("stamp re-search");
    return ZB_OK;
}

extern "C" int zb_tile_rows(zb_ctx *ctx, int which, int import, int r0, int r1, void *buf) {
    NEED_TILE();
    void *base = which == ZB_TILE_MASK_IN ? (void *)ctx->d_min : which == ZB_TILE_RDIFF ? (void *)ctx->d_rdiff :
                 which == ZB_TILE_RNOISE ? (void *)ctx->d_rnoise : which == ZB_TILE_OMASK ? (void *)ctx->d_omask : NULL;
    if (!base || !buf || r0 < 0 || r1 > ctx->p.ny || r1 < r0) { snprintf(ctx->err, sizeof ctx->err, "zb_tile_rows(%d, rows %d..%d)", which, r0, r1); return ZB_ERR_PARAM; }
    if (r1 == r0) return ZB_OK;
    // nreg 1x1: the region is the frame, region offset 0, pitch nx
    char *w = (char *)base + (size_t)r0 * ctx->p.nx * 4;
    const size_t bytes = (size_t)(r1 - r0) * ctx->p.nx * 4;
    cudaError_t e = import ? cudaMemcpyAsync(w, buf, bytes, cudaMemcpyDeviceToDevice, ctx->tile.st)
                           : cudaMemcpyAsync(buf, w, bytes, cudaMemcpyDeviceToDevice, ctx->tile.st);
    if (e != cudaSuccess) return zb_cuda_fail(ctx, e, "zb_tile_rows", __FILE__, __LINE__);
    return ZB_OK;
}

static int fill_slots(zb_ctx *ctx, const int *slots, int n) {
    ZbTile &T = ctx->tile;
    if (n <= 0) return ZB_OK;
    const bool ct = ctx->h_dev->conv_template != 0;
    int *jobs = ctx->h_njobs + 4;                         // pinned, n_slots ints (see zb_configure)
    for (int i = 0; i < n; i++) jobs[i] = slots ? slots[i] : T.slot_lo + i;
    zb_copy_small(ctx, ctx->d_jobs, jobs, (size_t)n * sizeof(int), T.st);
    return zb_launch_fill(ctx, ct ? T.ref : T.sci, ct ? T.sci : T.ref, ctx->d_jobs, n, T.st);
}

extern "C" int zb_tile_fill(zb_ctx *ctx) {
    NEED_TILE();
    cudaStream_t st = ctx->tile.st;
    SYNC("before fill");                                  // the pinned job list is reused between calls
    return fill_slots(ctx, NULL, ctx->tile.slot_hi - ctx->tile.slot_lo);
}

extern "C" int zb_tile_normal_export(zb_ctx *ctx, double *gram, double *rhs) {
    NEED_TILE();
    if (!gram || !rhs) return ZB_ERR_PARAM;
    ZbTile &T = ctx->tile;
    const size_t gl = (size_t)T.info.gram_len, rl = (size_t)T.info.rhs_len, ns = (size_t)ctx->n_slots, lo = (size_t)T.slot_lo, n = (size_t)(T.slot_hi - T.slot_lo);
    cudaMemsetAsync(gram, 0, ns * gl * sizeof(double), T.st);
    cudaMemsetAsync(rhs, 0, ns * rl * sizeof(double), T.st);
    if (n) {
        cudaMemcpyAsync(gram + lo * gl, ctx->d_gram + lo * gl, n * gl * sizeof(double), cudaMemcpyDeviceToDevice, T.st);
        cudaMemcpyAsync(rhs + lo * rl, ctx->d_rhs + lo * rl, n * rl * sizeof(double), cudaMemcpyDeviceToDevice, T.st);
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return zb_cuda_fail(ctx, e, "zb_tile_normal_export", __FILE__, __LINE__);
    return ZB_OK;
}

extern "C" int zb_tile_normal_import(zb_ctx *ctx, const double *gram, const double *rhs) {
    NEED_TILE();
    if (!gram || !rhs) return ZB_ERR_PARAM;
    ZbTile &T = ctx->tile;
    const size_t ns = (size_t)ctx->n_slots;
    cudaMemcpyAsync(ctx->d_gram, gram, ns * T.info.gram_len * sizeof(double), cudaMemcpyDeviceToDevice, T.st);
    cudaMemcpyAsync(ctx->d_rhs, rhs, ns * T.info.rhs_len * sizeof(double), cudaMemcpyDeviceToDevice, T.st);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return zb_cuda_fail(ctx, e, "zb_tile_normal_import", __FILE__, __LINE__);
    return ZB_OK;
}

extern "C" int zb_tile_solve(zb_ctx *ctx) {
    NEED_TILE();
    ZbTile &T = ctx->tile;
    CHECK_RC(zb_launch_fit_solve(ctx, T.st));
    return zb_launch_stamp_sig(ctx, T.sci, T.ref, T.sci_rms, T.ref_rms, T.slot_lo, T.slot_hi, T.st);
}

__global__ void k_sig_gather(const ZbStamp *__restrict__ stamps, int n, int lo, int hi, double *__restrict__ sig) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < n) sig[k] = (k >= lo && k < hi) ? stamps[k].sig1 : 0.0;
}
__global__ void k_sig_scatter(ZbStamp *__restrict__ stamps, int n, const double *__restrict__ sig) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < n) stamps[k].sig1 = sig[k];
}

extern "C" int zb_tile_sig_export(zb_ctx *ctx, double *sig) {
    NEED_TILE();
    if (!sig) return ZB_ERR_PARAM;
    k_sig_gather<<<(ctx->n_slots + 127) / 128, 128, 0, ctx->tile.st>>>(ctx->d_stamps, ctx->n_slots, ctx->tile.slot_lo, ctx->tile.slot_hi, sig);
    ctx->launches++;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return zb_cuda_fail(ctx, e, "k_sig_gather", __FILE__, __LINE__);
    return ZB_OK;
}

extern "C" int zb_tile_sig_import(zb_ctx *ctx, const double *sig) {
    NEED_TILE();
    if (!sig) return ZB_ERR_PARAM;
    k_sig_scatter<<<(ctx->n_slots + 127) / 128, 128, 0, ctx->tile.st>>>(ctx->d_stamps, ctx->n_slots, sig);
    ctx->launches++;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return zb_cuda_fail(ctx, e, "k_sig_scatter", __FILE__, __LINE__);
    return ZB_OK;
}

extern "C" int zb_tile_clip(zb_ctx *ctx, int *active) {
    NEED_TILE();
    if (!active) return ZB_ERR_PARAM;
    ZbTile &T = ctx->tile;
    cudaStream_t st = T.st;
    CHECK_RC(zb_launch_clip(ctx, st));
    int *h_jobs = ctx->h_njobs + 4;
    zb_copy_small(ctx, ctx->h_fit, ctx->d_fit, sizeof(ZbFit), st);
    zb_copy_small(ctx, ctx->h_njobs, ctx->d_njobs, sizeof(int), st);
    zb_copy_small(ctx, h_jobs, ctx->d_jobs, (size_t)ctx->n_slots * sizeof(int), st);
    SYNC("clip");
    *active = ctx->h_fit[0].active != 0;
    // the slots check_again replaced (alard.c:1196-1234): each owner refills its own
    int n = 0, *mine = (int *)malloc((size_t)(ctx->n_slots + 1) * sizeof(int));
    if (!mine) return ZB_ERR_NOMEM;
    for (int i = 0; i < ctx->h_njobs[0]; i++) if (h_jobs[i] >= T.slot_lo && h_jobs[i] < T.slot_hi) mine[n++] = h_jobs[i];
    int rc = fill_slots(ctx, mine, n);
    free(mine);
    return rc;
}

extern "C" int zb_tile_convolve(zb_ctx *ctx) {
    NEED_TILE();
    ZbTile &T = ctx->tile;
    return zb_launch_convolve(ctx, T.sci, T.ref, T.sci_rms, T.ref_rms, T.smask, T.sdt, T.tmask, T.tdt, T.diff, T.noise, T.out_mask, T.st,
                              T.info.blk_row0, T.info.blk_row1, T.info.out_row0, T.info.out_row1);
}

// main.c:1417-1554 on the gathered frame: final stamp sigma, the four region statistics, the record
extern "C" int zb_tile_finish(zb_ctx *ctx, zb_region_info *regions, double *kernel_sol, zb_stamp_info *stamps) {
    NEED_TILE();
    if (!regions) return ZB_ERR_PARAM;
    ZbTile &T = ctx->tile;
    *regions = T.ri;
    CHECK_RC(zb_launch_final_sig(ctx, T.st));
    CHECK_RC(zb_launch_region_stats(ctx, T.noise, T.out_mask, T.st));
    return zb_finish_readback(ctx, regions, kernel_sol, stamps, 0, T.st);
}
