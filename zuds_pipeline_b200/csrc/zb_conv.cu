// zb_conv.cu — the output pass: spatially varying convolution fused with background, noise,
// mask propagation, subtraction, photometric scaling and output masking.
//
//   k_make_kernels   alard.c:1346-1385 make_kernel, once per kcStep x kcStep block (alard.c:1272-1280)
//                    plus the region-centre kernel whose sum is KSUMnn (main.c:1131-1137)
//   k_convolve       alard.c:1249-1344 spatial_convolve  (image, variance, mask OR, uks/aks test)
//                    main.c:1126-1128 background, :1145-1163 noise, :1167-1173 unconvolved-side mask,
//                    :1399-1415 / :1439-1454 subtract + scale, :1584-1605 output masking
//   k_region_frame   the hwKernel frame spatial_convolve never touches (main.c:1089,1320-1331; alard.c:1264)
//
// Bound: the reference semantics (one dense (2hw+1)^2 kernel per block, double accumulation of
// float x double products) make this pass FP64-pipe bound, not HBM bound: 2*(2hw+1)^2 DFMA per
// pixel against 29 bytes of traffic.  Layout: a CTA owns NB horizontally adjacent blocks; the
// image/variance tile (+halo) is converted to double ONCE while it is staged in shared memory,
// each thread keeps TX adjacent outputs in registers and slides a register window along the
// kernel row so that one shared-memory load feeds up to TX DFMAs.
#include "zb_internal.h"
#include <stdlib.h>
#include <type_traits>


// ----------------------------------------------------------------------- kernels table

#define MK_NB 16      // convolution blocks per CTA of k_make_kernels
__global__ void __launch_bounds__(256)
k_make_kernels(const ZbDev *__restrict__ D, const double *__restrict__ sols, const double *__restrict__ kvec,
               double *__restrict__ kern, double *__restrict__ kernr, int fwp,
               double *__restrict__ ksum_out, double *__restrict__ kaks, int nbx_max, int nby_max)
{
    __shared__ double coef[MK_NB][ZB_MAX_BASIS];
    const int r = blockIdx.y;
    const ZbRegion R = D->reg[r];
    if (R.skip) return;
    const int hw = D->p.hw_kernel, fw = D->sz.fw_kernel, fw2 = fw * fw, kc = D->sz.kc_step;
    const int nK = D->sz.n_comp_ker, nc2 = D->sz.n_comp, ko = D->p.ker_order;
    const int nbx = (R.rpx + kc - 1) / kc, nby = (R.rpy + kc - 1) / kc, nb = nbx * nby;   // ceil(xSize/kcStep)
    const int per_region = nbx_max * nby_max + 1;
    const int b0 = blockIdx.x * MK_NB;
    if (b0 > nb) return;
    const double *sol = sols + (long long)r * (D->sz.n_comp_total + 1);
    const int tid = threadIdx.x;
    // coefficients of every basis at the block centre, terms in the reference order
    for (int it = tid; it < MK_NB * nK; it += blockDim.x) {
        int lb = it / nK, i1 = it - lb * nK, b = b0 + lb;
        if (b > nb) continue;
        int xi, yi;
        if (b == nb) { xi = R.rpx / 2; yi = R.rpy / 2; }          // KSUM kernel, main.c:1136
        else { int j1 = b / nbx, i1b = b - j1 * nbx; xi = i1b * kc + hw + hw; yi = j1 * kc + hw + hw; }
        double xf = (xi - 0.5 * R.rpx) / (0.5 * R.rpx), yf = (yi - 0.5 * R.rpy) / (0.5 * R.rpy);
        double c;
        if (i1 == 0) c = sol[1];
        else {
            int k = 2 + (i1 - 1) * nc2;
            c = 0.0; double ax = 1.0;
            for (int ix = 0; ix <= ko; ix++) {
                double ay = 1.0;
                for (int iy = 0; iy <= ko - ix; iy++) { c = __dadd_rn(c, __dmul_rn(__dmul_rn(sol[k++], ax), ay)); ay = __dmul_rn(ay, yf); }
                ax = __dmul_rn(ax, xf);
            }
        }
        coef[lb][i1] = c;
    }
    __syncthreads();
    for (int t = tid; t < fw2; t += blockDim.x) {
        const int roff = (fw - 1 - t / fw) * fwp + (fw - 1 - t % fw);      // position in the reversed, row-padded copy
        double acc[MK_NB];
#pragma unroll
        for (int lb = 0; lb < MK_NB; lb++) acc[lb] = 0.0;
        for (int n = 0; n < nK; n++) {
            double kv = kvec[(long long)n * fw2 + t];
#pragma unroll
            for (int lb = 0; lb < MK_NB; lb++) acc[lb] = __dadd_rn(acc[lb], __dmul_rn(coef[lb][n], kv));
        }
#pragma unroll
        for (int lb = 0; lb < MK_NB; lb++) {
            int b = b0 + lb;
            if (b > nb) continue;
            kern[((long long)r * per_region + b) * fw2 + t] = acc[lb];
            // the same kernel reversed (image coordinates ascending) with rows padded to fwp: what k_convolve
            // copies straight into shared memory
            kernr[((long long)r * per_region + b) * fw * fwp + roff] = acc[lb];
        }
    }
    for (int lb = 0; lb < MK_NB; lb++) {                      // zero padding of the reversed rows
        const int b = b0 + lb, npad = fwp - fw;
        if (b > nb || npad <= 0) continue;
        for (int t = tid; t < fw * npad; t += blockDim.x)
            kernr[((long long)r * per_region + b) * fw * fwp + (t / npad) * fwp + fw + t % npad] = 0.0;
    }
    __syncthreads();
    // aks = sum |K| in spatial_convolve's tap order (image coordinates ascending = kernel index descending,
    // alard.c:1289-1304): one dependent chain per block, needed only where a bad pixel lies under a footprint
    if (tid < MK_NB && b0 + tid < nb) {
        const double *o = kern + ((long long)r * per_region + b0 + tid) * fw2;
        double s = 0.0;
        for (int i = fw2 - 1; i >= 0; i--) s = __dadd_rn(s, fabs(o[i]));
        kaks[(long long)r * per_region + b0 + tid] = s;
    }
    if (b0 <= nb && nb < b0 + MK_NB) {                             // this CTA made the centre kernel
        if (tid == MK_NB) {
            const double *o = kern + ((long long)r * per_region + nb) * fw2;
            double s = 0.0;
            for (int i = 0; i < fw2; i++) s = __dadd_rn(s, o[i]);
            ksum_out[(size_t)r * 4] = s;
        }
    }
}

// --------------------------------------------------------------------------- helpers

struct PixCtx {
    const float *conv, *other, *conv_rms, *other_rms;
    const void *cmask_src; int cdt;       // user mask of the convolved image (-tmi for -c t)
    const void *omask_src; int odt;       // user mask of the other image     (-imi for -c t)
    float cU, cL, oU, oL;                 // value thresholds: convolved / other image
    float c_invgain, c_quad, o_invgain, o_quad;
};

__device__ static inline int value_bits(float v, float fill, float hi, float lo) {
    int m = 0;
    m |= (ZB_FLAG_INPUT_ISBAD | ZB_FLAG_BAD_PIXVAL) * (v == fill);
    m |= (ZB_FLAG_INPUT_ISBAD | ZB_FLAG_SAT_PIXEL) * (v >= hi);
    m |= (ZB_FLAG_INPUT_ISBAD | ZB_FLAG_LOW_PIXEL) * (v <= lo);
    return m;
}

__device__ static inline int user_bits(const void *m, int dt, long long g) {
    int v = zb_read_mask(m, dt, g);
    return v | (ZB_FLAG_INPUT_MASK * (v > 0));
}

__device__ static inline double background_at(const ZbDev *D, const ZbRegion &R, const double *sol, int xi, int yi) {
    const int ncompBG = (D->sz.n_comp_ker - 1) * D->sz.n_comp + 1, bgo = D->p.bg_order;
    double xf = (xi - 0.5 * R.rpx) / (0.5 * R.rpx), yf = (yi - 0.5 * R.rpy) / (0.5 * R.rpy);
    double bg = 0.0, ax = 1.0; int k = 1;
    for (int i = 0; i <= bgo; i++) {
        double ay = 1.0;
        for (int j = 0; j <= bgo - i; j++) { bg = __dadd_rn(bg, __dmul_rn(__dmul_rn(sol[ncompBG + k++], ax), ay)); ay = __dmul_rn(ay, yf); }
        ax = __dmul_rn(ax, xf);
    }
    return bg;
}

// everything after the convolution sum for one region pixel; the other image's value `oth`, its rms
// value `orv` (ignored without an rms map) and its user-mask bits `ub` are passed in so that callers
// can issue those global loads in batches
__device__ static inline void finish_pixel2(const ZbDev *D, const ZbRegion &R, const PixCtx &P, const double *sol,
                                            int x, int y, bool interior, float conv_out, float var_out, int mconv,
                                            double inv1, bool do_scale, float oth, float orv, int ub,
                                            float *rdiff, float *rnoise, int32_t *omask,
                                            float *diff, float *noise, int32_t *out_mask)
{
    const long long g = (long long)(R.by0 + y) * D->nx + (R.bx0 + x);
    const long long l = R.off + (long long)y * R.rpx + x;
    const float fill = D->p.fill_val;
    int m = mconv;
    int ob = value_bits(oth, fill, P.oU, P.oL);
    if (ob) m |= ob | ZB_FLAG_OUTPUT_ISBAD;
    m |= ub;
    if (ub & ZB_FLAG_INPUT_ISBAD) m |= ZB_FLAG_OUTPUT_ISBAD;
    float ovar;
    if (P.other_rms) ovar = __fmul_rn(orv, orv);
    else { double qq = __fmul_rn(P.o_quad, P.o_quad); ovar = (float)__dadd_rn(__dmul_rn(fabs((double)oth), (double)P.o_invgain), qq); }
    float nz = __fadd_rn(ovar, var_out);
    nz = (float)sqrt((double)nz);
    float d;
    if (interior) {
        float o = (float)__dadd_rn((double)conv_out, background_at(D, R, sol, x, y));
        d = __fsub_rn(o, oth);
        if (do_scale) { d = (float)__dmul_rn((double)d, inv1); nz = (float)__dmul_rn((double)nz, inv1); }
        if (D->conv_template) d = -d;
    } else {
        d = fill;                                   // oRData was pre-filled and never written here
        m |= ZB_FLAG_OUTPUT_ISBAD;                  // main.c:1320-1331
    }
    rdiff[l] = d; rnoise[l] = nz; omask[l] = m;
    if (x >= R.rx0 - R.bx0 && x <= R.own_x1 - R.bx0 && y >= R.ry0 - R.by0 && y <= R.own_y1 - R.by0) {
        float dd = d, nn = nz;
        if (interior && D->p.kf_spread_mask2 >= 0 && (m & ZB_FLAG_OUTPUT_ISBAD)) { dd = fill; nn = D->p.fill_val_noise; }
        diff[g] = dd; noise[g] = nn;
        if (out_mask) out_mask[g] = m;
    }
}

__device__ static inline void finish_pixel(const ZbDev *D, const ZbRegion &R, const PixCtx &P, const double *sol,
                                           int x, int y, bool interior, float conv_out, float var_out, int mconv,
                                           double inv1, bool do_scale,
                                           float *rdiff, float *rnoise, int32_t *omask,
                                           float *diff, float *noise, int32_t *out_mask)
{
    const long long g = (long long)(R.by0 + y) * D->nx + (R.bx0 + x);
    finish_pixel2(D, R, P, sol, x, y, interior, conv_out, var_out, mconv, inv1, do_scale,
                  P.other[g], P.other_rms ? P.other_rms[g] : 0.f, user_bits(P.omask_src, P.odt, g),
                  rdiff, rnoise, omask, diff, noise, out_mask);
}

// --------------------------------------------------------------------------- convolution
//
// Register-tiled direct convolution.  A CTA owns NBX x NBY convolution blocks (each B x B output
// pixels with its own kernel, alard.c:1272-1280).  The (image, variance) tile + halo is staged once in
// shared memory as double2 (float -> double converted at staging); the kernels K of the CTA's blocks
// are staged reversed, so all loops ascend in image coordinates.  A thread owns TX x TY outputs of one
// block and walks the tile rows s = 0 .. TY+fw-2: each tile row it loads (TX+CH-1 values per chunk of
// CH kernel columns) feeds the TY output rows whose footprint contains it, with kernel row
// dy = s - oy.  Per output the taps are visited dy-outer / dx-inner, the reference's order
// (alard.c:1289-1306).  Shared-memory traffic per DFMA falls by ~TY against a one-row register tile,
// which is what moves the pass from shared-memory-bound to FP64-pipe-bound.
//   per (s, chunk):  TX+CH-1 LDS.128 (window) + TY*CH LDS.64 (kernel, warp-broadcast)
//                    -> TY*CH*TX*2 DFMA (+ TY*CH DMUL for K^2)
// alard.c:1304-1308,1327: sum |K| over all taps (aks) and over the taps whose pixel is not INPUT_ISBAD
// (uks), both in the reference's tap order; true when uks/aks < kerFracMask.  Rare path (a bad
// pixel under the footprint), kept out of line so the output pass stays small.
__device__ __noinline__ static bool conv_kernel_fraction_low(const double *Kb, const unsigned char *f0, const unsigned char *h0,
                                                             int fw, int fwp, int W, float kfm, double aks_blk)
{
    // Fast decision: uks = aks - (the bad taps), accurate to ~1e-13 relative; the reference's
    // sequentially rounded sums are only reproduced when the ratio is within 1e-9 of the threshold.
    // h0[dy * W] bit 1: some INPUT_ISBAD pixel in the fw columns of footprint row dy.
    double bs = 0.0;
    for (int dy = 0; dy < fw; dy++) {
        if (!(h0[dy * W] & 2)) continue;
        const unsigned char *fr = f0 + dy * W;
        const double *kr = Kb + dy * fwp;
        for (int dx = 0; dx < fw; dx++) if (fr[dx] & 2) bs += fabs(kr[dx]);
    }
    const double ratio = (aks_blk - bs) / aks_blk;
    if (fabs(ratio - (double)kfm) > 1e-9) return ratio < (double)kfm;
    double aks = 0.0, uks = 0.0;
    for (int dy = 0; dy < fw; dy++)
        for (int dx = 0; dx < fw; dx++) {
            const double ak = fabs(Kb[dy * fwp + dx]);
            aks = __dadd_rn(aks, ak);
            if (!(f0[dy * W + dx] & 2)) uks = __dadd_rn(uks, ak);
        }
    return (uks / aks) < kfm;
}

__device__ static inline void cp_async4(void *smem_dst, const void *gsrc) {
    unsigned sa = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sa), "l"(gsrc) : "memory");
}
__device__ static inline void cp_async8(void *smem_dst, const void *gsrc) {
    unsigned sa = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(sa), "l"(gsrc) : "memory");
}
__device__ static inline void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

template <int TX, int TY, int CH, bool EXACT /* fw % CH == 0 */, bool KKONLY /* -convvar */>
__global__ void __launch_bounds__(256)
k_convolve(const ZbDev *__restrict__ D, PixCtx P, const double *__restrict__ sols, const double *__restrict__ kernr,
           const double *__restrict__ ksums, const double *__restrict__ kaks,
           int nbx_max, int nby_max, int NBX, int NBY, int GX, int GY, int W, int TH, int blk_row0, int blk_row1,
           float *__restrict__ rdiff, float *__restrict__ rnoise, int32_t *__restrict__ omask,
           float *__restrict__ diff, float *__restrict__ noise, int32_t *__restrict__ out_mask)
{
    extern __shared__ __align__(16) unsigned char dyn[];
    const int r = blockIdx.z;
    const ZbRegion R = D->reg[r];
    if (R.skip) return;
    const int hw = D->p.hw_kernel, fw = D->sz.fw_kernel, fw2 = fw * fw, B = D->sz.kc_step;
    const int nbx = (R.rpx + B - 1) / B, nby = min((R.rpy + B - 1) / B, blk_row1);   // block rows [blk_row0, blk_row1): row-tile window
    const int bx0 = blockIdx.x * NBX, by0 = blk_row0 + blockIdx.y * NBY;
    if (by0 >= nby || bx0 >= nbx) return;
    const int i00 = bx0 * B + hw, j00 = by0 * B + hw;         // first output pixel of the CTA (region coords)
    if (j00 >= R.rpy - hw || i00 >= R.rpx - hw) return;
    const int nch = (fw + CH - 1) / CH, fwp = nch * CH;
    const int NB = NBX * NBY;
    double2 *tile = (double2 *)dyn;                           // [TH][W]  (image, variance)
    double *Kr = (double *)(tile + (size_t)TH * W);           // [NB][fw][fwp]  K reversed
    unsigned char *flag = (unsigned char *)(Kr + (size_t)NB * fw * fwp);   // [TH][W] bit0: any mask bit, bit1: 0x80
    unsigned char *hor = flag + (size_t)TH * W;               // [TH][W] OR of flag over the fw columns to the right
    __shared__ int s_any;
    const int tid = threadIdx.x, nt = blockDim.x;
    const int per_region = nbx_max * nby_max + 1;
    const float fill = D->p.fill_val;
    if (tid == 0) s_any = 0;
    __syncthreads();
    // ---- stage the tile, origin (i00 - hw, j00 - hw) in region coords.  The raw floats travel by
    // cp.async into the low 8 bytes of their own double2 slot (all copies of the CTA in flight at
    // once), then each thread widens the slots it copied in place.
    const bool have_rms = P.conv_rms != nullptr;
    const int ntile = TH * W;
    const int lane = tid & 31, warp = tid >> 5, nw = nt >> 5;
    const int nxi = D->nx;
    const int xlim = R.rpx - (i00 - hw), ylim = R.rpy - (j00 - hw);      // tile columns / rows inside the region
    const long long gbase = (long long)(R.by0 + j00 - hw) * nxi + (R.bx0 + i00 - hw);
    for (int ty = warp; ty < TH; ty += nw) {
        const long long grow = gbase + (long long)ty * nxi;
        const bool yin = ty < ylim;
        for (int tx = lane; tx < W; tx += 32) {
            const int k = ty * W + tx;
            unsigned char f = 0;
            if (yin && tx < xlim) {
                float *slot = (float *)&tile[k];
                cp_async4(slot, P.conv + grow + tx);
                if (have_rms) cp_async4(slot + 1, P.conv_rms + grow + tx);
                if (P.cmask_src) { const int ub = user_bits(P.cmask_src, P.cdt, grow + tx); f = (ub ? 1 : 0) | ((ub & ZB_FLAG_INPUT_ISBAD) ? 2 : 0); }
            }
            flag[k] = f;
        }
    }
    {   // kernels of the CTA's blocks: k_make_kernels left them reversed and row-padded, a flat cp.async copy
        const int kn = fw * fwp;
        for (int lb = 0; lb < NB; lb++) {
            const int lby = lb / NBX, lbx = lb - lby * NBX;
            const bool ex = bx0 + lbx < nbx && by0 + lby < nby;
            const double *src = kernr + ((long long)r * per_region + (long long)(by0 + lby) * nbx + (bx0 + lbx)) * kn;
            double *dst = Kr + (size_t)lb * kn;
            if (ex) for (int e = tid; e < kn; e += nt) cp_async8(dst + e, src + e);
            else for (int e = tid; e < kn; e += nt) dst[e] = 0.0;
        }
    }
    cp_async_wait_all();
    int anyf = 0;
    const double c_qq = __fmul_rn(P.c_quad, P.c_quad);
    for (int ty = warp; ty < TH; ty += nw) {
        const bool yin = ty < ylim;
        for (int tx = lane; tx < W; tx += 32) {
            const int k = ty * W + tx;
            double2 v = make_double2(0.0, 0.0); unsigned char f = flag[k];
            if (yin && tx < xlim) {
                const float2 raw = *(const float2 *)&tile[k];
                const float pv = raw.x;
                float var;
                if (have_rms) var = __fmul_rn(raw.y, raw.y);
                else var = (float)__dadd_rn(__dmul_rn(fabs((double)pv), (double)P.c_invgain), c_qq);
                v.x = (double)pv; v.y = (double)var;
                const int cm = value_bits(pv, fill, P.cU, P.cL);
                f |= (cm ? 1 : 0) | ((cm & ZB_FLAG_INPUT_ISBAD) ? 2 : 0);
            }
            tile[k] = v; flag[k] = f; anyf |= f;
        }
    }
    if (anyf) s_any = 1;
    __syncthreads();
    const int any = s_any;
    if (any) {                                                // row-wise OR: the box dilation is separable
        for (int k = tid; k < ntile; k += nt) {
            const int ty = k / W, tx = k - ty * W;
            const int n = min(fw, W - tx);
            int h = 0;
            for (int dx = 0; dx < n; dx++) h |= flag[k + dx];
            hor[k] = (unsigned char)h;
        }
    }

    // thread -> (block of the CTA, column group, row group); row groups fastest
    const int tpb = GX * GY;
    const int lb = tid / tpb, lt = tid - lb * tpb;
    const int gx = lt / GY, gy = lt - gx * GY;
    const int lby = lb / NBX, lbx = lb - lby * NBX;
    const bool active = lb < NB && bx0 + lbx < nbx && by0 + lby < nby;
    const int x0 = lbx * B + gx * TX, y0 = lby * B + gy * TY;   // first output of the thread, relative to (i00, j00)
    double aq[TY][TX], av[TY][TX];
#pragma unroll
    for (int oy = 0; oy < TY; oy++)
#pragma unroll
        for (int t = 0; t < TX; t++) { aq[oy][t] = 0.0; av[oy][t] = 0.0; }
    const int ns = TY + fw - 1;
    if (active) {
        const double *Kb = Kr + (size_t)lb * fw * fwp;
        for (int s = 0; s < ns; s++) {
            const double2 *trow = tile + (size_t)(y0 + s) * W + x0;
            const bool steady = s >= TY - 1 && s < fw;              // every output row of the thread sees this tile row
            for (int ch = 0; ch < fwp; ch += CH) {
                double2 w[TX + CH - 1];
#pragma unroll
                for (int q = 0; q < TX + CH - 1; q++) w[q] = trow[ch + q];
#pragma unroll
                for (int oy = 0; oy < TY; oy++) {
                    const int dy = s - oy;
                    if (!steady && (dy < 0 || dy >= fw)) continue;  // warp-uniform
                    const double *krow = Kb + dy * fwp + ch;
                    if (EXACT || ch + CH <= fw) {
                        double k1[CH], k2[CH];
#pragma unroll
                        for (int dx = 0; dx < CH; dx++) k1[dx] = krow[dx];
#pragma unroll
                        for (int dx = 0; dx < CH; dx++) k2[dx] = KKONLY ? k1[dx] : __dmul_rn(k1[dx], k1[dx]);
#pragma unroll
                        for (int dx = 0; dx < CH; dx++)
#pragma unroll
                            for (int t = 0; t < TX; t++) {
                                aq[oy][t] = fma(w[t + dx].x, k1[dx], aq[oy][t]);
                                av[oy][t] = fma(w[t + dx].y, k2[dx], av[oy][t]);
                            }
                    } else {
#pragma unroll
                        for (int dx = 0; dx < CH; dx++) {
                            if (ch + dx >= fw) break;               // padded taps are never multiplied (0 x NaN)
                            const double k1 = krow[dx];
                            const double k2 = KKONLY ? k1 : __dmul_rn(k1, k1);
#pragma unroll
                            for (int t = 0; t < TX; t++) {
                                aq[oy][t] = fma(w[t + dx].x, k1, aq[oy][t]);
                                av[oy][t] = fma(w[t + dx].y, k2, av[oy][t]);
                            }
                        }
                    }
                }
            }
        }
    }
    // ---- hand the sums over through shared memory (the tile is dead) so that the output pass
    // below runs row-major over the CTA's rectangle: coalesced global traffic, loads batched
    const int OW = NBX * B, OH = NBY * B;
    float2 *res = (float2 *)dyn;                                  // [OH][OW] (conv, var)
    __syncthreads();
    if (active) {
#pragma unroll
        for (int oy = 0; oy < TY; oy++) {
            if (gy * TY + oy >= B) break;
#pragma unroll
            for (int t = 0; t < TX; t++) {
                if (gx * TX + t >= B) break;
                res[(size_t)(y0 + oy) * OW + x0 + t] = make_float2((float)aq[oy][t], (float)av[oy][t]);
            }
        }
    }
    __syncthreads();
    // ---- output pass: background, noise, masks, subtraction, scaling, output masking (main.c:1126-1173,
    // 1399-1454, 1584-1605).  A warp owns output rows, lanes own columns; EPR rows are in flight per step.
    const double *sol = sols + (long long)r * (D->sz.n_comp_total + 1);
    const double ksum = ksums[(size_t)r * 4], inv1 = 1. / ksum;
    const bool conv_t = D->conv_template != 0;
    const bool do_scale = conv_t ? (D->p.phot_normalize == 't') : (D->p.phot_normalize == 'i');
    const bool bg_const = D->p.bg_order == 0;
    const double bg0 = background_at(D, R, sol, 0, 0);             // the whole background when bgo == 0
    const bool have_orms = P.other_rms != nullptr;
    const double o_qq = __fmul_rn(P.o_quad, P.o_quad);
    const float kfm = D->p.ker_frac_mask, fill_noise = D->p.fill_val_noise;
    const bool mask_out = D->p.kf_spread_mask2 >= 0;
    const int ox_lim = min(OW, R.rpx - hw - i00), oy_lim = min(min(OH, (nby - by0) * B), R.rpy - hw - j00);
    const int own_x0 = R.rx0 - R.bx0 - i00, own_x1 = R.own_x1 - R.bx0 - i00;   // ownership window, CTA coords
    const int own_y0 = R.ry0 - R.by0 - j00, own_y1 = R.own_y1 - R.by0 - j00;
    const long long gout = (long long)(R.by0 + j00) * nxi + (R.bx0 + i00);
    const long long lout = R.off + (long long)j00 * R.rpx + i00;
    constexpr int EPB = 8;                                      // pixels per lane whose global loads are in flight together
    const int nchunk = (ox_lim + 31) >> 5;
    const int nitems = (oy_lim > warp ? (oy_lim - warp + nw - 1) / nw : 0) * nchunk;   // (row of this warp, column chunk)
    // item -> (row of this warp, column chunk) without a division: nitems * nchunk is far below 2^16
    const unsigned magic = (65536u + (unsigned)nchunk - 1u) / (unsigned)nchunk;
    // the global loads of batch n+1 (other image, its rms, its user mask) are in flight while batch n is finished
#define CONV_OUT_LOAD(IT0, OTH, ORV, UB) do { \
        _Pragma("unroll") for (int e = 0; e < EPB; e++) { \
            const int it = (IT0) + e, j = (int)(((unsigned)it * magic) >> 16), c = it - j * nchunk; \
            const int lx = lane + 32 * c, ly = warp + nw * j; \
            OTH[e] = 0.f; ORV[e] = 0.f; UB[e] = 0; \
            if (it < nitems && lx < ox_lim) { \
                const long long g = gout + (long long)ly * nxi + lx; \
                OTH[e] = P.other[g]; \
                if (have_orms) ORV[e] = P.other_rms[g]; \
                UB[e] = zb_read_mask(P.omask_src, P.odt, g); } } } while (0)   /* raw value: nothing may depend on a load here */
    float oth[EPB], orv[EPB]; int ub[EPB];
    CONV_OUT_LOAD(0, oth, orv, ub);
    for (int it0 = 0; it0 < nitems; it0 += EPB) {
        {
            float noth[EPB], norv[EPB]; int nub[EPB];
            if (it0 + EPB < nitems) CONV_OUT_LOAD(it0 + EPB, noth, norv, nub);
            int eox[EPB], eoy[EPB];
#pragma unroll
            for (int e = 0; e < EPB; e++) {
                const int it = it0 + e, j = (int)(((unsigned)it * magic) >> 16), c = it - j * nchunk;
                eox[e] = lane + 32 * c; eoy[e] = warp + nw * j;
                if (it >= nitems || eox[e] >= ox_lim) eox[e] = -1;
            }
#pragma unroll
            for (int e = 0; e < EPB; e++) {
                const int ox = eox[e], oy = eoy[e];
                if (ox < 0) continue;
                // mask of the convolved image at this pixel and under the footprint
                int m = 0;
                if (any) {
                    const unsigned char *h0 = hor + (size_t)oy * W + ox;
                    int mbit = 0;
                    for (int dy = 0; dy < fw; dy++) mbit |= h0[dy * W];
                    if (flag[(size_t)(oy + hw) * W + ox + hw]) {
                        const long long g = gout + (long long)oy * nxi + ox;
                        const int cm = user_bits(P.cmask_src, P.cdt, g) | value_bits(P.conv[g], fill, P.cU, P.cL);
                        m |= cm;
                        if (cm & ZB_FLAG_INPUT_ISBAD) m |= ZB_FLAG_OUTPUT_ISBAD;
                    }
                    if (mbit) {
                        bool low = false;                            // no bad tap: uks == aks exactly
                        if (mbit & 2) {
                            const int oby = oy / B, obx = ox / B;
                            const double aks_blk = kaks[(long long)r * per_region + (long long)(by0 + oby) * nbx + (bx0 + obx)];
                            low = conv_kernel_fraction_low(Kr + (size_t)(oby * NBX + obx) * fw * fwp, flag + (size_t)oy * W + ox, h0, fw, fwp, W, kfm,
                                                           aks_blk);
                        }
                        if (low) m |= (ZB_FLAG_OUTPUT_ISBAD | ZB_FLAG_BAD_CONV);
                        else m |= ZB_FLAG_OK_CONV;
                    }
                }
                const float2 qv = res[(size_t)oy * OW + ox];
                const float ov = oth[e];
                const int ob = value_bits(ov, fill, P.oU, P.oL);
                if (ob) m |= ob | ZB_FLAG_OUTPUT_ISBAD;
                const int ubit = ub[e] | (ZB_FLAG_INPUT_MASK * (ub[e] > 0));       // user_bits() of the prefetched raw mask value
                m |= ubit;
                if (ubit & ZB_FLAG_INPUT_ISBAD) m |= ZB_FLAG_OUTPUT_ISBAD;
                float ovar;
                if (have_orms) ovar = __fmul_rn(orv[e], orv[e]);
                else ovar = (float)__dadd_rn(__dmul_rn(fabs((double)ov), (double)P.o_invgain), o_qq);
                // (float)sqrt((double)x) == sqrtf(x) correctly rounded: 53 >= 2*24+2 makes the double rounding innocuous
                float nz = __fsqrt_rn(__fadd_rn(ovar, qv.y));
                const double bg = bg_const ? bg0 : background_at(D, R, sol, i00 + ox, j00 + oy);
                float d = __fsub_rn((float)__dadd_rn((double)qv.x, bg), ov);
                if (do_scale) { d = (float)__dmul_rn((double)d, inv1); nz = (float)__dmul_rn((double)nz, inv1); }
                if (conv_t) d = -d;
                const long long l = lout + (long long)oy * R.rpx + ox;
                rdiff[l] = d; rnoise[l] = nz; omask[l] = m;
                if (ox >= own_x0 && ox <= own_x1 && oy >= own_y0 && oy <= own_y1) {
                    if (mask_out && (m & ZB_FLAG_OUTPUT_ISBAD)) { d = fill; nz = fill_noise; }
                    const long long g = gout + (long long)oy * nxi + ox;
                    diff[g] = d; noise[g] = nz;
                    if (out_mask) out_mask[g] = m;
                }
            }
            if (it0 + EPB < nitems) {
#pragma unroll
                for (int e = 0; e < EPB; e++) { oth[e] = noth[e]; orv[e] = norv[e]; ub[e] = nub[e]; }
            }
        }
    }
#undef CONV_OUT_LOAD
}

// The hwKernel frame of every region (not touched by spatial_convolve).
__global__ void __launch_bounds__(256)
k_region_frame(const ZbDev *__restrict__ D, PixCtx P, const double *__restrict__ sols,
               float *__restrict__ rdiff, float *__restrict__ rnoise, int32_t *__restrict__ omask,
               float *__restrict__ diff, float *__restrict__ noise, int32_t *__restrict__ out_mask, int row0, int row1)
{
    const int r = blockIdx.y;
    const ZbRegion R = D->reg[r];
    if (R.skip) return;
    const int hw = D->p.hw_kernel;
    const int nrow = 2 * hw * R.rpx, ncol = 2 * hw * (R.rpy - 2 * hw);
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < nrow + ncol; k += gridDim.x * blockDim.x) {
        int x, y;
        if (k < nrow) { int q = k / R.rpx; x = k - q * R.rpx; y = q < hw ? q : R.rpy - 2 * hw + q; }
        else { int kk = k - nrow, q = kk / (2 * hw), c = kk - q * 2 * hw; y = hw + q; x = c < hw ? c : R.rpx - 2 * hw + c; }
        if (y < row0 || y >= row1) continue;                       // row-tile window (whole region: 0 .. INT_MAX)
        // vData is calloc'ed: the convolved variance is 0 on the frame (alard.c:1264)
        finish_pixel(D, R, P, sols + (long long)r * (D->sz.n_comp_total + 1), x, y, false, 0.f, 0.f, 0, 1.0, false,
                     rdiff, rnoise, omask, diff, noise, out_mask);
    }
}

int zb_launch_convolve(zb_ctx *ctx, const float *sci, const float *ref, const float *sci_rms, const float *ref_rms,
                       const void *smask, int sdt, const void *tmask, int tdt,
                       float *diff, float *noise, int32_t *out_mask, cudaStream_t st,
                       int blk_row0, int blk_row1, int row0, int row1)
{
    const zb_sizes &sz = ctx->sz;
    const zb_params &p = ctx->p;
    const ZbDev *H = ctx->h_dev;
    const int fw = sz.fw_kernel, B = sz.kc_step, nR = sz.n_regions;
    const int nbx_max = (sz.max_rpix_x + B - 1) / B, nby_max = (sz.max_rpix_y + B - 1) / B;
    PixCtx P;
    const bool ct = H->conv_template != 0;
    P.conv = ct ? ref : sci;            P.other = ct ? sci : ref;
    P.conv_rms = ct ? ref_rms : sci_rms; P.other_rms = ct ? sci_rms : ref_rms;
    P.cmask_src = ct ? tmask : smask;   P.cdt = ct ? tdt : sdt;
    P.omask_src = ct ? smask : tmask;   P.odt = ct ? sdt : tdt;
    P.cU = ct ? H->tU : H->iU; P.cL = ct ? H->tL : H->iL;
    P.oU = ct ? H->iU : H->tU; P.oL = ct ? H->iL : H->tL;
    const float tinv = (float)(1. / p.t_gain), tq = p.t_rdnoise / p.t_gain, iinv = (float)(1. / p.i_gain), iq = p.i_rdnoise / p.i_gain;
    P.c_invgain = ct ? tinv : iinv; P.c_quad = ct ? tq : iq;
    P.o_invgain = ct ? iinv : tinv; P.o_quad = ct ? iq : tq;

    // register tile per kernel-block size: (TX, TY, CH) with GX*TX >= B, GY*TY >= B
    int TXr, TYr, CHr;
    if (B <= 12)      { TXr = 4; TYr = 3; CHr = 4; }
    else if (B <= 24) { TXr = 7; TYr = 3; CHr = 7; }
    else              { TXr = 8; TYr = 3; CHr = 8; }
    const int GX = (B + TXr - 1) / TXr, GY = (B + TYr - 1) / TYr, tpb = GX * GY;
    if (tpb > 256) { snprintf(ctx->err, sizeof ctx->err, "kc_step %d too large for the convolution kernel", B); return ZB_ERR_UNSUPPORTED; }
    const int nch = (fw + CHr - 1) / CHr, fwp = nch * CHr;
    dim3 gk((nbx_max * nby_max + 1 + MK_NB - 1) / MK_NB + 1, nR);
    k_make_kernels<<<gk, 256, 0, st>>>(ctx->d_dev, ctx->d_sol, ctx->d_kvec, ctx->d_kern, ctx->d_kernr, fwp, ctx->d_kstat, ctx->d_kaks, nbx_max, nby_max); ZB_DBG("k_make_kernels");
    // blocks per CTA: ~128 threads and two CTAs per SM when shared memory allows (228 KB per SM, 1 KB
    // reserved per CTA), else one CTA with as many blocks as fit; among equals the squarer tile (less halo)
    int NBX = 1, NBY = 1, W = 0, TH = 0, threads = 0; size_t smem = 0;
    {
        static int env_nb = -1, env_ctas = 2;
        if (env_nb < 0) { const char *e = getenv("ZB_CONV_NB"); env_nb = e ? atoi(e) : 0; e = getenv("ZB_CONV_CTAS"); env_ctas = e ? atoi(e) : 2; if (env_ctas < 1) env_ctas = 2; }
        const size_t cap2 = (233472 - (size_t)env_ctas * 1024) / env_ctas;
        auto shape = [&](int bxn, int by, int *w, int *th) -> size_t {
            int ww = (bxn - 1) * B + GX * TXr + fwp - 1; while ((TYr * ww) % 8 != 1) ww++;   // conflict-free row-group stride
            *w = ww; *th = (by - 1) * B + GY * TYr + fw - 1;
            return (size_t)*th * ww * sizeof(double2) + (size_t)bxn * by * fw * fwp * sizeof(double) + 2 * (size_t)*th * ww;
        };
        const int target = env_nb > 0 ? env_nb : (128 / tpb > 0 ? 128 / tpb : 1);
        long best = -1;
        for (int pass = 0; pass < 2 && best < 0; pass++)
            for (int nb = target; nb >= 1 && best < 0; nb--)
                for (int by = 1; by * by <= nb; by++) {
                    if (nb % by) continue;
                    int w, th; const size_t sm = shape(nb / by, by, &w, &th);
                    if (sm > (pass == 0 ? cap2 : (size_t)ZB_SMEM_OPTIN)) continue;
                    best = nb; NBX = nb / by; NBY = by; W = w; TH = th; smem = sm;          // larger `by` overwrites: squarer
                }
        if (best < 0) smem = shape(1, 1, &W, &TH);
        threads = ((NBX * NBY * tpb + 31) / 32) * 32;
    }
    if (smem > ZB_SMEM_OPTIN) { snprintf(ctx->err, sizeof ctx->err, "convolution tile needs %zu bytes of shared memory", smem); return ZB_ERR_UNSUPPORTED; }
    if (blk_row1 > nby_max) blk_row1 = nby_max;
    if (blk_row0 < 0) blk_row0 = 0;
    dim3 gc((nbx_max + NBX - 1) / NBX, (blk_row1 - blk_row0 + NBY - 1) / NBY, nR);
    cudaEventRecord(ctx->ev[10], st);          // k_convolve alone: the kernel bench.py's roofline is quoted on
#define ZB_CONV_LAUNCH2(TX_, TY_, CH_, EX_, KK_) do { \
        cudaFuncSetAttribute(k_convolve<TX_, TY_, CH_, EX_, KK_>, cudaFuncAttributeMaxDynamicSharedMemorySize, ZB_SMEM_OPTIN); \
        k_convolve<TX_, TY_, CH_, EX_, KK_><<<gc, threads, smem, st>>>(ctx->d_dev, P, ctx->d_sol, ctx->d_kernr, ctx->d_kstat, ctx->d_kaks, nbx_max, nby_max, \
            NBX, NBY, GX, GY, W, TH, blk_row0, blk_row1, ctx->d_rdiff, ctx->d_rnoise, ctx->d_omask, diff, noise, out_mask); } while (0)
#define ZB_CONV_LAUNCH(TX_, TY_, CH_) do { \
        const bool ex_ = (fw % CH_) == 0, kk_ = p.convolve_variance != 0; \
        if (ex_ && !kk_) ZB_CONV_LAUNCH2(TX_, TY_, CH_, true, false); \
        else if (!ex_ && !kk_) ZB_CONV_LAUNCH2(TX_, TY_, CH_, false, false); \
        else ZB_CONV_LAUNCH2(TX_, TY_, CH_, false, true); } while (0)
    if (B <= 12)      ZB_CONV_LAUNCH(4, 3, 4);
    else if (B <= 24) ZB_CONV_LAUNCH(7, 3, 7);
    else              ZB_CONV_LAUNCH(8, 3, 8);
#undef ZB_CONV_LAUNCH2
#undef ZB_CONV_LAUNCH
    ZB_DBG("k_convolve");
    cudaEventRecord(ctx->ev[11], st);
    k_region_frame<<<dim3(64, nR), 256, 0, st>>>(ctx->d_dev, P, ctx->d_sol, ctx->d_rdiff, ctx->d_rnoise, ctx->d_omask, diff, noise, out_mask, row0, row1); ZB_DBG("k_region_frame");
    ctx->launches += 3;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return zb_cuda_fail(ctx, e, "convolution kernels", __FILE__, __LINE__);
    return ZB_OK;
}
