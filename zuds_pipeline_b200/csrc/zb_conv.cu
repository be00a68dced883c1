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
// Bound: one dense (2hw+1)^2 kernel per block, image AND variance convolved: 2*(2hw+1)^2 multiply-adds
// per pixel against 29 bytes of traffic, so the pass is arithmetic-bound for hw >= 5.  The image term runs on
// the FP64 pipe exactly as the reference accumulates it (float x double products summed in double), the
// variance term — positive terms, needed to 1e-4 relative — on the otherwise idle FP32 pipe, so the two
// halves of the work overlap; the measured decision between this split, the all-FP64 form of round 1, packed
// FP32 for both terms and a tensor-core form is in profiles/r02_conv_arith.md.
// Structure (persistent, warp-specialised, one CTA per SM):
//   warp 0        producer: TMA (cp.async.bulk.tensor.2d) of the next tile's float image / rms boxes and
//                 bulk copies of its blocks' kernels, signalled through mbarriers
//   warps 1-3     epilogue of tile n-1 (background, noise, masks, subtraction, outputs) while
//   warps 4-11    convert tile n+... to (image - sky, variance) pairs once, then run the FMA/LDS loop
#include "zb_internal.h"
#include <cuda.h>
#include <stdlib.h>
#include <type_traits>


// ----------------------------------------------------------------------- kernels table

#define MK_NB 16      // convolution blocks per CTA of k_make_kernels
struct __align__(16) KTap { double k; float k2; float pad; };   // one kernel tap as k_convolve reads it (one LDS.128)
__global__ void __launch_bounds__(256)
k_make_kernels(const ZbDev *__restrict__ D, const double *__restrict__ sols, const double *__restrict__ kvec,
               int fwp, double *__restrict__ ksum_out, double *__restrict__ kaks, int nbx_max, int nby_max,
               KTap *__restrict__ ktap, int kstride)
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
    double ak[MK_NB];                                                      // this thread's share of sum |K| per block
#pragma unroll
    for (int lb = 0; lb < MK_NB; lb++) ak[lb] = 0.0;
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
            // k_convolve's operand: the kernel reversed (image coordinates ascending), rows padded to fwp, K for the
            // image term (double) and K^2 for the variance term (float) in one 16-byte record per tap
            ak[lb] += fabs(acc[lb]);
            KTap kt; kt.k = acc[lb]; kt.k2 = (float)(acc[lb] * acc[lb]); kt.pad = 0.f;
            ktap[((long long)r * per_region + b) * kstride + roff] = kt;
        }
    }
    for (int lb = 0; lb < MK_NB; lb++) {                      // zero padding of the reversed rows
        const int b = b0 + lb, npad = fwp - fw;
        if (b > nb || npad <= 0) continue;
        for (int t = tid; t < fw * npad; t += blockDim.x) {
            KTap z; z.k = 0.0; z.k2 = 0.f; z.pad = 0.f;
            ktap[((long long)r * per_region + b) * kstride + (t / npad) * fwp + fw + t % npad] = z;
        }
    }

    // aks = sum |K| of every block: k_convolve decides uks/aks < kerFracMask from it with a 1e-9 margin and re-sums both
    // in the reference's tap order when the ratio is that close to the threshold (conv_kernel_fraction_low), so any
    // summation order does here: per-thread shares, warp shuffles, then the eight warps in order (deterministic)
    __shared__ double ak_w[8][MK_NB];
#pragma unroll
    for (int lb = 0; lb < MK_NB; lb++) {
        double v = ak[lb];
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if ((tid & 31) == 0) ak_w[tid >> 5][lb] = v;
    }
    __syncthreads();
    if (tid < MK_NB && b0 + tid < nb) {
        double sum = 0.0;
        for (int w = 0; w < 8; w++) sum += ak_w[w][tid];
        kaks[(long long)r * per_region + b0 + tid] = sum;
    }
    if (b0 <= nb && nb < b0 + MK_NB) {                             // this CTA made the centre kernel: KSUM, kernel index ascending (main.c:1136)
        if (tid == MK_NB) {
            const KTap *o = ktap + ((long long)r * per_region + nb) * kstride;
            double s = 0.0;
            for (int dy = fw - 1; dy >= 0; dy--) for (int dx = fw - 1; dx >= 0; dx--) s = __dadd_rn(s, o[dy * fwp + dx].k);
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

__device__ __forceinline__ static double background_at(const ZbDev *D, const ZbRegion &R, const double *sol, int xi, int yi) {
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

// the same, kept out of line: k_convolve's epilogue calls it per pixel only for -bgo > 0, and inlined there the compiler
// hoists its coordinate normalisation (two double divisions) into the per-row code of the common -bgo 0 case too
__device__ __noinline__ static double background_at_call(const ZbDev *D, int rpx, int rpy, const double *sol, int xi, int yi) {
    ZbRegion R; R.rpx = rpx; R.rpy = rpy;
    return background_at(D, R, sol, xi, yi);
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
// Persistent, warp-specialised direct convolution (alard.c:1249-1344).  The unit of work is a tile of
// NBX x NBY convolution blocks (each B x B output pixels with its own kernel, alard.c:1272-1280); a CTA
// takes tiles t = blockIdx.x, blockIdx.x + gridDim.x, ... (static schedule, deterministic).  12 warps:
//
//   warp 0 (producer)   one thread: TMA tensor loads (cp.async.bulk.tensor.2d) of the tile's float boxes of the
//                       convolved image and its rms map (+halo) into `raw`, bulk copies (cp.async.bulk) of the
//                       blocks' (K, K^2) float pairs into `kbuf`, completion on mbarriers.  The boxes of tile
//                       n+1 land while tile n is being multiplied.
//   warps 4-11 (FMA)    convert `raw` once into `compd` (image widened to double) and `compv` (variance, float)
//                       + flag bit rows, then the register-tiled FMA/LDS loop: a thread owns TX x TY outputs
//                       of one block and (K-split KS) a band of kernel rows; it walks the tile rows, each row's
//                       window of TX+CH-1 values feeding the TY output rows whose footprint contains it.  Per
//                       tap and output: one DFMA (image x K, the reference's float x double products summed in
//                       double, alard.c:1289-1299) and one FFMA (variance x K^2, alard.c:1300-1306) — two
//                       pipes that run side by side.  Sums go to `res` (bands add in band order).
//   warps 1-3 (epilogue) everything after the sum for tile n-1 while tile n is multiplied: sky term, background,
//                       noise, mask propagation incl. the uks/aks test, subtraction, scaling, output masking,
//                       global stores (main.c:1126-1173, 1399-1454, 1584-1605).
//
// Flags: bit rows.  R_any / R_bad hold one bit per tile pixel (any mask bit / INPUT_ISBAD), V_* their OR over
// the 2hw+1 rows of a footprint; "is there a flagged pixel under this output's footprint" is a shift and a
// mask, and the bad taps of a footprint row are the set bits of a word.
//
// Arithmetic note (profiles/r02_conv_arith.md): the image term is the reference's arithmetic (the taps of an
// output are visited in the reference's order; DFMA contracts the product and the add, and the KS bands of
// kernel rows are added at the end, which moves the last bit of the double sum at most and the float result
// in practice never).  The variance term is a sum of positive FP32 products: relative error ~1e-6 against the
// north star's 1e-4.  With -convvar (variance x K, terms of both signs, alard.c:1300-1306) it runs in FP64 too.

typedef unsigned long long u64;

#define ZB_SMEM_CONV (227 * 1024)   // dynamic shared memory opt-in of k_convolve (the same constant for every launch, see ZB_SMEM_OPTIN)
#define CONV_THREADS 384            // warp group 0: producer + 3 epilogue warps; warp groups 1, 2: FMA warps
#define CONV_FMA_THREADS 256
#define CONV_EPI_WARPS 3
#define CONV_MAX_KS 4
#define CONV_MAX_NB 64
#define CONV_REGS_LIGHT 120          // setmaxnreg of warp group 0
#define CONV_REGS_FMA 192           // setmaxnreg of the FMA warp groups: 128*LIGHT + 256*FMA = 64512 = 384 * 168, what the launch allocates

struct ConvGeom {
    int NBX, NBY, KS, GX, GY;       // blocks per tile, kernel-row split, thread grid per block
    int Wt, TH;                     // tile columns / rows actually used (outputs + halo + chunk padding)
    int WR;                         // raw row pitch in floats = TMA box width (multiple of 4)
    int Wc, Wv;                     // row pitch of the double image plane / the float variance plane
    int NW;                         // 32-bit words per flag row (incl. 2 zero words of padding)
    int fwp, kstride;               // kernel row pitch / block stride in taps (KTap records)
    int nbx_max, nby_max, blk_row0, blk_row1;
    int use_tma, have_rms;
    unsigned o_raw_img, o_raw_rms, o_compd, o_compv, o_k, o_resq, o_resv, o_rmask, o_vmask, o_pre, o_resm, o_misc;   // shared-memory offsets
    int n_tiles;
    int first[ZB_MAX_REGIONS + 1];  // tiles of region r: [first[r], first[r+1])
};

// ---- PTX wrappers -----------------------------------------------------------------------------
__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *b, int count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count) : "memory"); }
__device__ __forceinline__ void mbar_arrive(uint64_t *b) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t *b, unsigned bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t *b, unsigned parity) {
    asm volatile("{\n\t.reg .pred P1;\n\tLAB_WAIT:\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t@P1 bra DONE;\n\tbra LAB_WAIT;\n\tDONE:\n\t}"
                 ::"r"(smem_u32(b)), "r"(parity) : "memory");
}
// The waits of the roles that run ahead of the multiply (producer) or behind it (epilogue) last tens of microseconds; a
// tight try_wait loop there was a third of all instructions the CTA issued (ncu: 31 M + 28 M executions of the two
// loops against 554 M in total), on the schedulers the FMA warps need.  These sleep between polls instead.
__device__ __forceinline__ void mbar_wait_idle(uint64_t *b, unsigned parity) {
    for (;;) {
        unsigned done;
        asm volatile("{\n\t.reg .pred P1;\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n\tselp.u32 %0, 1, 0, P1;\n\t}"
                     : "=r"(done) : "r"(smem_u32(b)), "r"(parity) : "memory");
        if (done) return;
#ifndef CONV_WAIT_SPIN
        __nanosleep(400);
#endif
    }
}
__device__ __forceinline__ void tma_load_2d(void *dst, const CUtensorMap *tm, int x, int y, uint64_t *bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(smem_u32(dst)), "l"(tm), "r"(x), "r"(y), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void bulk_load(void *dst, const void *src, unsigned bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory"); }

// n (<= 64) bits of a flag row starting at bit x; rows carry two zero words of padding
__device__ __forceinline__ u64 bits_at(const unsigned *row, int x, int n) {
    const int w = x >> 5, sh = x & 31;
    u64 v = ((u64)row[w] | ((u64)row[w + 1] << 32)) >> sh;
    if (sh) v |= (u64)row[w + 2] << (64 - sh);
    return n >= 64 ? v : (v & ((1ull << n) - 1ull));
}

// alard.c:1304-1308,1327: sum |K| over all taps (aks) and over the taps whose pixel is not INPUT_ISBAD (uks), both in
// the reference's tap order; true when uks/aks < kerFracMask.  Fast decision: uks = aks - (the bad taps), the bad taps
// of a footprint row summed run by run from the row's prefix sums of |K| (a saturated core, a dead column or the frame
// border is one run per row: summing such runs tap by tap made the launch 70 % slower), accurate to ~1e-13 relative;
// the reference's sequentially rounded sums are only reproduced when the ratio is within 1e-9 of the threshold.
//   pre   [fw][fwp+1] prefix sums of |K| along each (reversed) kernel row of the output's block
//   Kb    the block's taps (reversed, rows padded to fwp)
//   rbad  R_bad row of the footprint's first row; rows: bit dy set when tile row (first row + dy) holds any bad pixel
__device__ __forceinline__ static bool conv_kernel_fraction_low(const double *__restrict__ pre, const KTap *__restrict__ Kb, const unsigned *rbad,
                                                                 int NW, int ox, u64 rows, int fw, int fwp, float kfm, double aks_blk)
{
    double bs = 0.0;
    while (rows) {
        const int dy = __ffsll((long long)rows) - 1; rows &= rows - 1;
        u64 rb = bits_at(rbad + (size_t)dy * NW, ox, fw);
        const double *pr = pre + dy * (fwp + 1);
        while (rb) {
            const int lo = __ffsll((long long)rb) - 1;
            const int len = __ffsll((long long)~(rb >> lo)) - 1;     // run of bad taps lo .. lo+len-1 (fw < 64: a zero bit always follows)
            bs += pr[lo + len] - pr[lo];
            rb &= ~(((1ull << len) - 1ull) << lo);
        }
    }
    const double ratio = (aks_blk - bs) / aks_blk;
    if (fabs(ratio - (double)kfm) > 1e-9) return ratio < (double)kfm;
    double aks = 0.0, uks = 0.0;
    for (int dy = 0; dy < fw; dy++) {
        const u64 rb = bits_at(rbad + (size_t)dy * NW, ox, fw);
        for (int dx = 0; dx < fw; dx++) {
            const double ak = fabs(Kb[dy * fwp + dx].k);
            aks = __dadd_rn(aks, ak);
            if (!((rb >> dx) & 1ull)) uks = __dadd_rn(uks, ak);
        }
    }
    return (uks / aks) < kfm;
}

// tile t -> region, first block of the tile; false when the tile has no output pixel
struct TileRef { int r, bx0, by0, nbx, nby; };
__device__ __forceinline__ bool conv_tile(const ZbDev *D, const ConvGeom &G, int t, TileRef &T, ZbRegion &R) {
    int r = 0;
    while (t >= G.first[r + 1]) r++;
    R = D->reg[r];
    const int B = D->sz.kc_step, hw = D->p.hw_kernel;
    T.r = r;
    T.nbx = (R.rpx + B - 1) / B;
    T.nby = min((R.rpy + B - 1) / B, G.blk_row1);
    const int ntx = (T.nbx + G.NBX - 1) / G.NBX;
    const int lt = t - G.first[r], ty = lt / ntx, tx = lt - ty * ntx;
    T.bx0 = tx * G.NBX; T.by0 = G.blk_row0 + ty * G.NBY;
    return !(R.skip || T.by0 >= T.nby || T.bx0 >= T.nbx || T.by0 * B + hw >= R.rpy - hw || T.bx0 * B + hw >= R.rpx - hw);
}

// One chunk of CH kernel columns for the TY output rows of a thread.  VAR64 (-convvar): the variance is convolved
// with K itself, in double.
template <int TX, int CH, typename AV>
__device__ __forceinline__ void conv_window(const double *__restrict__ trd, const float *__restrict__ trv, int ch,
                                            double (&wd)[TX + CH - 1], AV (&wv)[TX + CH - 1])
{
#pragma unroll
    for (int q = 0; q < TX + CH - 1; q++) { wd[q] = trd[ch + q]; wv[q] = (AV)trv[ch + q]; }
}

template <int TX, int TY, int CH, bool EXACT, bool VAR64, typename AV>
__device__ __forceinline__ void conv_chunk(const double (&wd)[TX + CH - 1], const AV (&wv)[TX + CH - 1], const KTap *__restrict__ Kb,
                                           int ch, int s, int fw, int fwp, int dy0, int dy1, bool steady,
                                           double (&aq)[TY][TX], AV (&av)[TY][TX])
{
    if (steady && (EXACT || ch + CH <= fw)) {
        // every output row of the thread takes this tile row: the taps of output row oy+1 are requested before the
        // multiply-adds of output row oy are issued, so their shared-memory latency hides behind ~100 instructions
        double2 kk[2][CH];                                       // one LDS.128 per tap: (K, [K^2 float | pad])
        const double2 *k0 = reinterpret_cast<const double2 *>(Kb + s * fwp + ch);
#pragma unroll
        for (int dx = 0; dx < CH; dx++) kk[0][dx] = k0[dx];
#pragma unroll
        for (int oy = 0; oy < TY; oy++) {
            if (oy + 1 < TY) {
                const double2 *kn = reinterpret_cast<const double2 *>(Kb + (s - oy - 1) * fwp + ch);
#pragma unroll
                for (int dx = 0; dx < CH; dx++) kk[(oy + 1) & 1][dx] = kn[dx];
            }
#pragma unroll
            for (int dx = 0; dx < CH; dx++) {
                const double k = kk[oy & 1][dx].x;
                const float k2 = __int_as_float(__double2loint(kk[oy & 1][dx].y));
#pragma unroll
                for (int t = 0; t < TX; t++) {
                    aq[oy][t] = fma(wd[t + dx], k, aq[oy][t]);
                    if (VAR64) av[oy][t] = (AV)fma((double)wv[t + dx], k, (double)av[oy][t]);
                    else av[oy][t] = (AV)fmaf((float)wv[t + dx], k2, (float)av[oy][t]);
                }
            }
        }
        return;
    }
#pragma unroll
    for (int oy = 0; oy < TY; oy++) {
        const int dy = s - oy;
        if (!steady && (dy < dy0 || dy >= dy1)) continue;       // warp-uniform
        const KTap *krow = Kb + dy * fwp + ch;
        if (EXACT || ch + CH <= fw) {
            double2 kk[CH];                                      // one LDS.128 per tap: (K, [K^2 float | pad])
#pragma unroll
            for (int dx = 0; dx < CH; dx++) kk[dx] = reinterpret_cast<const double2 *>(krow)[dx];
#pragma unroll
            for (int dx = 0; dx < CH; dx++) {
                const double k = kk[dx].x;
                const float k2 = __int_as_float(__double2loint(kk[dx].y));
#pragma unroll
                for (int t = 0; t < TX; t++) {
                    aq[oy][t] = fma(wd[t + dx], k, aq[oy][t]);
                    if (VAR64) av[oy][t] = (AV)fma((double)wv[t + dx], k, (double)av[oy][t]);
                    else av[oy][t] = (AV)fmaf((float)wv[t + dx], k2, (float)av[oy][t]);
                }
            }
        } else {
#pragma unroll
            for (int dx = 0; dx < CH; dx++) {
                if (ch + dx >= fw) break;                        // padded taps are never multiplied (0 x NaN)
                const KTap k = krow[dx];
#pragma unroll
                for (int t = 0; t < TX; t++) {
                    aq[oy][t] = fma(wd[t + dx], k.k, aq[oy][t]);
                    if (VAR64) av[oy][t] = (AV)fma((double)wv[t + dx], k.k, (double)av[oy][t]);
                    else av[oy][t] = (AV)fmaf((float)wv[t + dx], k.k2, (float)av[oy][t]);
                }
            }
        }
    }
}

template <int TX, int TY, int CH, bool EXACT, bool VAR64>
__global__ void __launch_bounds__(CONV_THREADS, 1)
k_convolve(const __grid_constant__ CUtensorMap tm_img, const __grid_constant__ CUtensorMap tm_rms,
           const ZbDev *__restrict__ D, const PixCtx P, const __grid_constant__ ConvGeom G,
           const double *__restrict__ sols, const KTap *__restrict__ ktap,
           const double *__restrict__ ksums, const double *__restrict__ kaks, int *__restrict__ tile_counter,
           float *__restrict__ rdiff, float *__restrict__ rnoise, int32_t *__restrict__ omask,
           float *__restrict__ diff, float *__restrict__ noise, int32_t *__restrict__ out_mask)
{
    extern __shared__ __align__(128) unsigned char dyn[];
    float *raw_img = (float *)(dyn + G.o_raw_img);             // [TH][WR]
    float *raw_rms = (float *)(dyn + G.o_raw_rms);             // [TH][WR]
    double *compd = (double *)(dyn + G.o_compd);               // [TH][Wc]  image, widened once
    float *compv = (float *)(dyn + G.o_compv);                 // [TH][Wv]  variance
    KTap *kbuf = (KTap *)(dyn + G.o_k);                        // [NB][kstride]  (K, K^2) reversed, rows padded to fwp
    double *resq = (double *)(dyn + G.o_resq);                 // [OH][OW]  convolved image (sum over the KS bands)
    float *resv = (float *)(dyn + G.o_resv);                   // [OH][OW]  convolved variance
    unsigned *rmask = (unsigned *)(dyn + G.o_rmask);           // [2 parity][2 any/bad][TH][NW]
    unsigned *vmask = (unsigned *)(dyn + G.o_vmask);           // [2 parity][2 any/bad][OH][NW]
    uint64_t *bars = (uint64_t *)(dyn + G.o_misc);             // raw_full, raw_empty, k_full, k_empty, res_full, res_empty
    int *s_any = (int *)(bars + 6);                            // [2 parity]
    double *s_pre = (double *)(dyn + G.o_pre);                 // [NB][fw][fwp+1]  prefix sums of |K| along each kernel row (uks/aks test)
    unsigned short *resm = (unsigned short *)(dyn + G.o_resm); // [OH][OW]  mask bits the convolution contributes (alard.c:1312-1332)
    unsigned *s_rowbad = (unsigned *)(bars + 8);               // [2 parity][16]  bit per tile row: the row holds a bad pixel
    int *s_tile = (int *)(s_rowbad + 32);                      // [4] ring of tile indices handed out by the producer; -1: no more work
    uint64_t *raw_full = bars, *raw_empty = bars + 1, *k_full = bars + 2, *k_empty = bars + 3, *res_full = bars + 4, *res_empty = bars + 5;
    uint64_t *v_full = bars + 7;                               // the V flag rows of the tile are complete (FMA warps, tiles with flagged pixels)

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int hw = D->p.hw_kernel, fw = D->sz.fw_kernel, B = D->sz.kc_step;
    const int NBX = G.NBX, NBY = G.NBY, NB = NBX * NBY, KS = G.KS;
    const int OW = NBX * B, OH = NBY * B;
    const int TH = G.TH, Wt = G.Wt, WR = G.WR, Wc = G.Wc, Wv = G.Wv, NW = G.NW, fwp = G.fwp;
    const int per_region = G.nbx_max * G.nby_max + 1;
    const int nxi = D->nx, nyi = D->ny;
    const float fill = D->p.fill_val;
    constexpr int NFW = CONV_FMA_THREADS / 32;

    if (tid == 0) {
        mbar_init(raw_full, 1); mbar_init(raw_empty, NFW); mbar_init(k_full, 1); mbar_init(k_empty, NFW);
        mbar_init(res_full, NFW); mbar_init(res_empty, CONV_EPI_WARPS); mbar_init(v_full, NFW);
        s_any[0] = s_any[1] = 0;
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    // zero the flag rows once: the padding words stay zero for the whole launch
    for (int k = tid; k < 2 * 2 * (TH + OH) * NW; k += CONV_THREADS) (k < 4 * TH * NW ? rmask[k] : vmask[k - 4 * TH * NW]) = 0u;
    __syncthreads();

    if (warp < 4) {
#ifndef CONV_NO_SETMAXNREG
        asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(CONV_REGS_LIGHT));
#endif
        if (warp == 0) {
            // ================================================================== producer
            // Tiles are handed out dynamically (an atomic counter): their cost varies with what lies under them
            // (flagged pixels, partial blocks at region edges), and a static split left SMs idle for 17 % of the launch.
            // The tile of iteration `it` travels to the other roles in s_tile[it & 3], published by the raw_full barrier.
            int it = 0;
            for (;;) {
                int t = 0;
                if (lane == 0) t = atomicAdd(tile_counter, 1);
                t = __shfl_sync(0xffffffffu, t, 0);
                TileRef T; ZbRegion R;
                if (t < G.n_tiles && !conv_tile(D, G, t, T, R)) continue;
                const unsigned par = it & 1;
                if (lane == 0) mbar_wait_idle(raw_empty, par ^ 1);
                __syncwarp();
                if (t >= G.n_tiles) {                                              // no more work: tell the other roles and leave
                    if (lane == 0) { s_tile[it & 3] = -1; mbar_arrive(raw_full); }
                    break;
                }
                if (lane == 0) s_tile[it & 3] = t;
                it++;
                // tile origin, image coordinates; the box starts at the 16-byte aligned column at or before it: a tensor
                // load whose innermost start is not 16-byte aligned raises an illegal-instruction trap (tools/tma_probe.cu)
                const int gx0 = (R.bx0 + T.bx0 * B) & ~3, gy0 = R.by0 + T.by0 * B;
                if (G.use_tma) {
                    if (lane == 0) {
                        const unsigned bytes = (unsigned)(WR * TH * 4) * (G.have_rms ? 2u : 1u);
                        mbar_expect_tx(raw_full, bytes);
                        tma_load_2d(raw_img, &tm_img, gx0, gy0, raw_full);
                        if (G.have_rms) tma_load_2d(raw_rms, &tm_rms, gx0, gy0, raw_full);
                    }
                } else {
                    // row pitch or base address not TMA-encodable (nx % 4 != 0): the warp copies the box itself
                    for (int ty = 0; ty < TH; ty++) {
                        const int gy = gy0 + ty;
                        for (int tx = lane; tx < Wt + 3; tx += 32) {
                            const int gx = gx0 + tx;
                            const bool in = gx < nxi && gy < nyi;
                            raw_img[ty * WR + tx] = in ? P.conv[(long long)gy * nxi + gx] : 0.f;
                            if (G.have_rms) raw_rms[ty * WR + tx] = in ? P.conv_rms[(long long)gy * nxi + gx] : 0.f;
                        }
                    }
                    __syncwarp();
                    if (lane == 0) mbar_arrive(raw_full);
                }
                if (lane == 0) {
                    mbar_wait_idle(k_empty, par ^ 1);
                    int nex = 0;
                    for (int lb = 0; lb < NB; lb++) { const int lby = lb / NBX, lbx = lb - lby * NBX; nex += (T.bx0 + lbx < T.nbx && T.by0 + lby < T.nby); }
                    mbar_expect_tx(k_full, (unsigned)(nex * G.kstride * sizeof(KTap)));
                    for (int lb = 0; lb < NB; lb++) {
                        const int lby = lb / NBX, lbx = lb - lby * NBX;
                        if (!(T.bx0 + lbx < T.nbx && T.by0 + lby < T.nby)) continue;
                        const KTap *src = ktap + ((long long)T.r * per_region + (long long)(T.by0 + lby) * T.nbx + (T.bx0 + lbx)) * G.kstride;
                        bulk_load(kbuf + (size_t)lb * G.kstride, src, (unsigned)(G.kstride * sizeof(KTap)), k_full);
                    }
                }
                __syncwarp();
            }
        } else {
            // ================================================================== epilogue (warps 1-3)
            // A warp owns output rows ew, ew+3, ...; lanes own columns.  The global loads of a row (other image, its
            // rms, its user mask: up to EPC column chunks of 32) are issued one row ahead of their use.
            const int ew = warp - 1;
            constexpr int EPC = 2;                                                 // column chunks handled per pass over a row
            const bool conv_t = D->conv_template != 0;
            const bool do_scale = conv_t ? (D->p.phot_normalize == 't') : (D->p.phot_normalize == 'i');
            const bool bg_const = D->p.bg_order == 0;
            const bool have_orms = P.other_rms != nullptr;
            const double o_qq = __fmul_rn(P.o_quad, P.o_quad);
            const float fill_noise = D->p.fill_val_noise;
            const bool mask_out = D->p.kf_spread_mask2 >= 0;
            for (int it = 0;; it++) {
                const unsigned par = it & 1;
                mbar_wait_idle(res_full, par);
                const int t = s_tile[it & 3];
                if (t < 0) break;
                TileRef T; ZbRegion R;
                conv_tile(D, G, t, T, R);
                const int r = T.r;
                const int i00 = T.bx0 * B + hw, j00 = T.by0 * B + hw;              // first output pixel, region coordinates
                const double *sol = sols + (long long)r * (D->sz.n_comp_total + 1);
                const double ksum = ksums[(size_t)r * 4], inv1 = 1. / ksum;
                const double bg0 = background_at(D, R, sol, 0, 0);                 // the whole background when bgo == 0
                const int ox_lim = min(OW, R.rpx - hw - i00), oy_lim = min(min(OH, (T.nby - T.by0) * B), R.rpy - hw - j00);
                const int own_x0 = R.rx0 - R.bx0 - i00, own_x1 = R.own_x1 - R.bx0 - i00;   // ownership window, tile coords
                const int own_y0 = R.ry0 - R.by0 - j00, own_y1 = R.own_y1 - R.by0 - j00;
                const long long gout = (long long)(R.by0 + j00) * nxi + (R.bx0 + i00);
                const long long lout = R.off + (long long)j00 * R.rpx + i00;
                const int any = s_any[par];
                const int nchunk = (ox_lim + 31) >> 5;
                for (int c0 = 0; c0 < nchunk; c0 += EPC) {
                    float oth[EPC], orv[EPC]; int ub[EPC];
                    // loads of row `oy` for the chunks of this pass
#define CONV_EPI_LOAD(OY, OTH, ORV, UB) do { \
                        const long long grow_ = gout + (long long)(OY) * nxi; \
                        _Pragma("unroll") for (int e = 0; e < EPC; e++) { \
                            const int ox_ = lane + 32 * (c0 + e); \
                            OTH[e] = 0.f; ORV[e] = 0.f; UB[e] = 0; \
                            if ((OY) < oy_lim && ox_ < ox_lim) { \
                                OTH[e] = P.other[grow_ + ox_]; \
                                if (have_orms) ORV[e] = P.other_rms[grow_ + ox_]; \
                                UB[e] = zb_read_mask(P.omask_src, P.odt, grow_ + ox_); } } } while (0)   /* raw values: nothing may depend on a load here */
                    CONV_EPI_LOAD(ew, oth, orv, ub);
                    for (int oy = ew; oy < oy_lim; oy += CONV_EPI_WARPS) {
                        float noth[EPC], norv[EPC]; int nub[EPC];
                        CONV_EPI_LOAD(oy + CONV_EPI_WARPS, noth, norv, nub);       // next row of this warp, in flight while this one is finished
                        const long long grow = gout + (long long)oy * nxi, lrow = lout + (long long)oy * R.rpx;
                        const bool own_row = oy >= own_y0 && oy <= own_y1;
                        const double *rq = resq + (size_t)oy * OW;
                        const float *rv = resv + (size_t)oy * OW;
                        const unsigned short *rm = resm + (size_t)oy * OW;
#pragma unroll 1
                        for (int e = 0; e < EPC; e++) {
                            const int ox = lane + 32 * (c0 + e);
                            const float ov = oth[0], orv0 = orv[0]; const int ub0 = ub[0];
#pragma unroll
                            for (int q = 0; q + 1 < EPC; q++) { oth[q] = oth[q + 1]; orv[q] = orv[q + 1]; ub[q] = ub[q + 1]; }
                            if (ox >= ox_lim) continue;
                            int m = any ? (int)rm[ox] : 0;                     // what the convolution contributes (FMA warps)
                            const float conv_out = (float)rq[ox];
                            const float var = rv[ox];
                            const int ob = value_bits(ov, fill, P.oU, P.oL);
                            if (ob) m |= ob | ZB_FLAG_OUTPUT_ISBAD;
                            const int ubit = ub0 | (ZB_FLAG_INPUT_MASK * (ub0 > 0));       // user_bits() of the prefetched raw mask value
                            m |= ubit;
                            if (ubit & ZB_FLAG_INPUT_ISBAD) m |= ZB_FLAG_OUTPUT_ISBAD;
                            float ovar;
                            if (have_orms) ovar = __fmul_rn(orv0, orv0);
                            else ovar = (float)__dadd_rn(__dmul_rn(fabs((double)ov), (double)P.o_invgain), o_qq);
                            // (float)sqrt((double)x) == sqrtf(x) correctly rounded: 53 >= 2*24+2 makes the double rounding innocuous
                            float nz = __fsqrt_rn(__fadd_rn(ovar, var));
                            #ifdef CONV_BG_INLINE
                            const double bg = bg_const ? bg0 : background_at(D, R, sol, i00 + ox, j00 + oy);
#else
                            const double bg = bg_const ? bg0 : background_at_call(D, R.rpx, R.rpy, sol, i00 + ox, j00 + oy);
#endif
                            float d = __fsub_rn((float)__dadd_rn((double)conv_out, bg), ov);
                            if (do_scale) { d = (float)__dmul_rn((double)d, inv1); nz = (float)__dmul_rn((double)nz, inv1); }
                            if (conv_t) d = -d;
                            rdiff[lrow + ox] = d; rnoise[lrow + ox] = nz; omask[lrow + ox] = m;
                            if (own_row && ox >= own_x0 && ox <= own_x1) {
                                if (mask_out && (m & ZB_FLAG_OUTPUT_ISBAD)) { d = fill; nz = fill_noise; }
                                diff[grow + ox] = d; noise[grow + ox] = nz;
                                if (out_mask) out_mask[grow + ox] = m;
                            }
                        }
#pragma unroll
                        for (int e = 0; e < EPC; e++) { oth[e] = noth[e]; orv[e] = norv[e]; ub[e] = nub[e]; }
                    }
#undef CONV_EPI_LOAD
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(res_empty);
            }
        }
    } else {
        // ====================================================================== FMA warps (4-11)
#ifndef CONV_NO_SETMAXNREG
        asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(CONV_REGS_FMA));
#endif
        const int ft = tid - 128, fwarp = ft >> 5;
        const int Tg = CONV_FMA_THREADS / KS;                                      // threads of one kernel-row band
        const int g = ft / Tg, u = ft - g * Tg;
        const int tpb = G.GX * G.GY;
        const int lb = u / tpb, lt = u - lb * tpb;
        const int gxi = lt / G.GY, gyi = lt - gxi * G.GY;
        const int lby = lb / NBX, lbx = lb - lby * NBX;
        const int x0 = lbx * B + gxi * TX, y0 = lby * B + gyi * TY;               // first output of the thread, relative to the tile's first output
        const int dy0 = (g * fw) / KS, dy1 = ((g + 1) * fw) / KS;                  // this band's kernel rows (ascending image rows)
        const bool have_rms = G.have_rms != 0;
        const double c_qq = __fmul_rn(P.c_quad, P.c_quad);
        int vit = 0;                                                               // tiles that had flagged pixels so far (phase of v_full)
        for (int it = 0;; it++) {
            const unsigned par = it & 1;
            mbar_wait(raw_full, par);
            const int t = s_tile[it & 3];
            if (t < 0) {                                                           // pass the end of work on to the epilogue warps
                mbar_wait(res_empty, par ^ 1);
                __syncwarp();
                if (lane == 0) mbar_arrive(res_full);
                break;
            }
            TileRef T; ZbRegion R;
            conv_tile(D, G, t, T, R);
            const int gx0 = R.bx0 + T.bx0 * B, gy0 = R.by0 + T.by0 * B;
            const int sh = gx0 & 3;                                                // the raw box starts at the aligned column gx0 - sh
            unsigned *Rany = rmask + (size_t)(par * 2 + 0) * TH * NW, *Rbad = rmask + (size_t)(par * 2 + 1) * TH * NW;
            unsigned *Vany = vmask + (size_t)(par * 2 + 0) * OH * NW, *Vbad = vmask + (size_t)(par * 2 + 1) * OH * NW;
            if (ft == 0) s_any[par] = 0;
            if (ft < 16) s_rowbad[par * 16 + ft] = 0u;
            named_bar_sync(1, CONV_FMA_THREADS);                                   // every FMA warp is done with the previous comp tile
            // ---- convert: raw floats -> image widened to double, variance, flag bit rows.  Two copies of the loop: the zuds
            // call (rms map, no user mask on the convolved image) is a handful of instructions per pixel; everything else
            // (gain/read-noise variance model, -tmi / -imi mask of the convolved image) takes the general form
            int anyf = 0;
            const int x_in = nxi - gx0;                                            // tile columns inside the frame
            if (have_rms && !P.cmask_src) {
                for (int ty = fwarp; ty < TH; ty += NFW) {
                    const bool yin = gy0 + ty < nyi;
                    const float *__restrict__ ri = raw_img + ty * WR + sh, *__restrict__ rr = raw_rms + ty * WR + sh;
                    double *__restrict__ cd = compd + (size_t)ty * Wc;
                    float *__restrict__ cv = compv + (size_t)ty * Wv;
#ifdef CONV_CONVERT_SIMPLE
                    for (int txb = 0; txb < Wt; txb += 32) {
                        const int tx = txb + lane;
                        bool bad = false;
                        if (tx < Wt) {
                            const float pv = ri[tx], rv = rr[tx];
                            cd[tx] = (double)pv;
                            cv[tx] = __fmul_rn(rv, rv);
                            bad = yin && tx < x_in && (pv == fill || pv >= P.cU || pv <= P.cL);
                        }
                        const unsigned bb = __ballot_sync(0xffffffffu, bad);
                        if (lane == 0) { Rany[ty * NW + (txb >> 5)] = bb; Rbad[ty * NW + (txb >> 5)] = bb; if (bb) atomicOr(&s_rowbad[par * 16 + (ty >> 5)], 1u << (ty & 31)); }
                        anyf |= (bb != 0u);
                    }
#else
                    for (int txb = 0; txb < Wt; txb += 128) {                      // four 32-column segments in flight
                        float pv[4], rv[4];
#pragma unroll
                        for (int q = 0; q < 4; q++) {
                            const int tx = txb + 32 * q + lane;
                            pv[q] = tx < Wt ? ri[tx] : 0.f; rv[q] = tx < Wt ? rr[tx] : 0.f;
                        }
#pragma unroll
                        for (int q = 0; q < 4; q++) {
                            const int tx = txb + 32 * q + lane;
                            bool bad = false;                                      // every value flag carries INPUT_ISBAD (makeInputMask)
                            if (tx < Wt) {
                                cd[tx] = (double)pv[q];
                                cv[tx] = __fmul_rn(rv[q], rv[q]);
                                bad = yin && tx < x_in && (pv[q] == fill || pv[q] >= P.cU || pv[q] <= P.cL);
                            }
                            const unsigned bb = __ballot_sync(0xffffffffu, bad);
                            if (lane == 0 && txb + 32 * q < Wt) {
                                Rany[ty * NW + ((txb >> 5) + q)] = bb; Rbad[ty * NW + ((txb >> 5) + q)] = bb;
                                if (bb) atomicOr(&s_rowbad[par * 16 + (ty >> 5)], 1u << (ty & 31));
                            }
                            anyf |= (bb != 0u);
                        }
                    }
#endif
                }
            } else {
                for (int ty = fwarp; ty < TH; ty += NFW) {
                    const int gy = gy0 + ty;
                    for (int txb = 0; txb < Wt; txb += 32) {
                        const int tx = txb + lane;
                        int cm = 0;
                        if (tx < Wt) {
                            const float pv = raw_img[ty * WR + tx + sh];
                            float var;
                            if (have_rms) { const float rv = raw_rms[ty * WR + tx + sh]; var = __fmul_rn(rv, rv); }
                            else var = (float)__dadd_rn(__dmul_rn(fabs((double)pv), (double)P.c_invgain), c_qq);
                            compd[(size_t)ty * Wc + tx] = (double)pv;
                            compv[(size_t)ty * Wv + tx] = var;
                            if (tx < x_in && gy < nyi) {
                                cm = value_bits(pv, fill, P.cU, P.cL);
                                if (P.cmask_src) cm |= user_bits(P.cmask_src, P.cdt, (long long)gy * nxi + gx0 + tx);
                            }
                        }
                        const unsigned ba = __ballot_sync(0xffffffffu, cm != 0), bb = __ballot_sync(0xffffffffu, (cm & ZB_FLAG_INPUT_ISBAD) != 0);
                        if (lane == 0) { Rany[ty * NW + (txb >> 5)] = ba; Rbad[ty * NW + (txb >> 5)] = bb; if (bb) atomicOr(&s_rowbad[par * 16 + (ty >> 5)], 1u << (ty & 31)); }
                        anyf |= (ba != 0u);
                    }
                }
            }
            if (lane == 0 && anyf) s_any[par] = 1;
            __syncwarp();
            if (lane == 0) mbar_arrive(raw_empty);                                 // the producer may fetch the next tile's boxes
            // prefix sums of |K| along every kernel row of the tile's blocks, one thread per row (the uks/aks test below);
            // the kernels were requested when the previous tile's multiply ended and have normally landed by now
            mbar_wait(k_full, par);
            for (int k = CONV_FMA_THREADS - 1 - ft; k < NB * fw; k += CONV_FMA_THREADS) {
                const int b = k / fw, dy = k - b * fw;
                const KTap *kr = kbuf + (size_t)b * G.kstride + dy * fwp;
                double *pr = s_pre + (size_t)k * (fwp + 1);
                double acc = 0.0;
                pr[0] = 0.0;
                for (int dx = 0; dx < fw; dx++) { acc += fabs(kr[dx].k); pr[dx + 1] = acc; }
            }
            named_bar_sync(1, CONV_FMA_THREADS);
            if (s_any[par]) {                                                      // OR of the flag rows over each footprint's 2hw+1 rows
                for (int k = ft; k < 2 * OH * NW; k += CONV_FMA_THREADS) {
                    const int pl = k / (OH * NW), rem = k - pl * OH * NW, oy = rem / NW, w = rem - oy * NW;
                    const unsigned *src = (pl ? Rbad : Rany) + (size_t)oy * NW + w;
                    unsigned v = 0u;
                    for (int dy = 0; dy < fw; dy++) v |= src[(size_t)dy * NW];
                    (pl ? Vbad : Vany)[(size_t)oy * NW + w] = v;
                }
#ifdef CONV_V_BARSYNC
                named_bar_sync(1, CONV_FMA_THREADS);
#else
                __syncwarp();
                if (lane == 0) mbar_arrive(v_full);                                // read after the multiply: nobody waits here
#endif
            }
            // ---- multiply
            const bool active = lb < NB && T.bx0 + lbx < T.nbx && T.by0 + lby < T.nby;
            typedef typename std::conditional<VAR64, double, float>::type AV;
            double aq[TY][TX]; AV av[TY][TX];
#pragma unroll
            for (int oy = 0; oy < TY; oy++)
#pragma unroll
                for (int tt = 0; tt < TX; tt++) { aq[oy][tt] = 0.0; av[oy][tt] = (AV)0; }
            const bool any_bad = (s_rowbad[par * 16 + 0] | s_rowbad[par * 16 + 1] | s_rowbad[par * 16 + 2] | s_rowbad[par * 16 + 3] |
                                  s_rowbad[par * 16 + 4] | s_rowbad[par * 16 + 5] | s_rowbad[par * 16 + 6] | s_rowbad[par * 16 + 7] |
                                  s_rowbad[par * 16 + 8] | s_rowbad[par * 16 + 9] | s_rowbad[par * 16 + 10] | s_rowbad[par * 16 + 11] |
                                  s_rowbad[par * 16 + 12] | s_rowbad[par * 16 + 13]) != 0u;
            if (active) {
                const KTap *Kb = kbuf + (size_t)lb * G.kstride;
                const double *tbd = compd + (size_t)y0 * Wc + x0;
                const float *tbv = compv + (size_t)y0 * Wv + x0;
                for (int s = dy0; s < dy1 + TY - 1; s++) {
                    const double *trd = tbd + (size_t)s * Wc;
                    const float *trv = tbv + (size_t)s * Wv;
                    const bool steady = (s - (TY - 1) >= dy0) && (s < dy1);        // every output row of the thread sees this tile row
                    for (int ch = 0; ch < fwp; ch += CH) {
                        double wd[TX + CH - 1]; AV wv[TX + CH - 1];
                        conv_window<TX, CH, AV>(trd, trv, ch, wd, wv);
                        conv_chunk<TX, TY, CH, EXACT, VAR64, AV>(wd, wv, Kb, ch, s, fw, fwp, dy0, dy1, steady, aq, av);
                    }
                }
            }
            // ---- what the convolution contributes to the mask (alard.c:1304-1332), by the owners of each output: the
            // pixel's own bits with OUTPUT_ISBAD when it is bad; OK_CONV when a flagged pixel lies under the footprint, or
            // BAD_CONV | OUTPUT_ISBAD when too much of the kernel sits on bad pixels
            mbar_wait(res_empty, par ^ 1);
#ifndef CONV_V_BARSYNC
            if (s_any[par]) { mbar_wait(v_full, vit & 1); vit++; }                 // every warp's share of the V rows is written
#endif
#ifdef CONV_MASK_BAND0
            if (active && g == 0 && s_any[par]) {
#else
            if (active && s_any[par]) {                                            // the KS band owners of an output tile share its columns
#endif
                const float kfm = D->p.ker_frac_mask;
                const u64 fmask = (1ull << fw) - 1ull;
                const double aks_blk = any_bad ? kaks[(long long)T.r * per_region + (long long)(T.by0 + lby) * T.nbx + (T.bx0 + lbx)] : 1.0;
                const KTap *Kb = kbuf + (size_t)lb * G.kstride;
                const double *pre = s_pre + (size_t)lb * fw * (fwp + 1);
#pragma unroll 1
                for (int oy = 0; oy < TY; oy++) {
                    if (gyi * TY + oy >= B) break;
                    const int yy = y0 + oy;
                    const u64 wa = bits_at(Vany + (size_t)yy * NW, x0, TX + fw - 1), wb = bits_at(Vbad + (size_t)yy * NW, x0, TX + fw - 1);
                    const u64 wc = bits_at(Rany + (size_t)(yy + hw) * NW, x0 + hw, TX);
                    const u64 rows = any_bad ? bits_at(s_rowbad + par * 16, yy, fw) : 0ull;
#pragma unroll 1
#ifdef CONV_MASK_BAND0
                    for (int tt = 0; tt < TX; tt++) {
#else
                    for (int tt = g; tt < TX; tt += KS) {
#endif
                        if (gxi * TX + tt >= B) break;
                        int m = 0;
                        if ((wc >> tt) & 1ull) {
                            const float pv = (float)compd[(size_t)(yy + hw) * Wc + x0 + tt + hw];
                            int cm = value_bits(pv, fill, P.cU, P.cL);
                            if (P.cmask_src) cm |= user_bits(P.cmask_src, P.cdt, (long long)(gy0 + yy + hw) * nxi + gx0 + x0 + tt + hw);
                            m |= cm;
                            if (cm & ZB_FLAG_INPUT_ISBAD) m |= ZB_FLAG_OUTPUT_ISBAD;
                        }
                        if ((wa >> tt) & fmask) {
                            bool low = false;                                      // no bad tap: uks == aks exactly
                            if ((wb >> tt) & fmask) low = conv_kernel_fraction_low(pre, Kb, Rbad + (size_t)yy * NW, NW, x0 + tt, rows, fw, fwp, kfm, aks_blk);
                            m |= low ? (ZB_FLAG_OUTPUT_ISBAD | ZB_FLAG_BAD_CONV) : ZB_FLAG_OK_CONV;
                        }
                        resm[(size_t)yy * OW + x0 + tt] = (unsigned short)m;
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(k_empty);                                   // kernels (and, with the bar.sync above, comp) are free
            // ---- hand the sums to the epilogue warps: band 0 stores, the other bands of kernel rows add in band order
            for (int gb = 0; gb < KS; gb++) {
                if (active && g == gb) {
#pragma unroll
                    for (int oy = 0; oy < TY; oy++) {
                        if (gyi * TY + oy >= B) break;
#pragma unroll
                        for (int tt = 0; tt < TX; tt++) {
                            if (gxi * TX + tt >= B) break;
                            const size_t k = (size_t)(y0 + oy) * OW + x0 + tt;
                            if (gb == 0) { resq[k] = aq[oy][tt]; resv[k] = (float)av[oy][tt]; }
                            else { resq[k] += aq[oy][tt]; resv[k] = (float)((AV)resv[k] + av[oy][tt]); }
                        }
                    }
                }
                if (gb + 1 < KS) named_bar_sync(1, CONV_FMA_THREADS);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(res_full);
        }
    }
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

// ---- host side: tile geometry, tensor maps, launch -------------------------------------------------

typedef CUresult (*zb_encode_tiled_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                       const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                       CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static zb_encode_tiled_fn zb_encode_tiled() {
    static zb_encode_tiled_fn fn = nullptr;
    static int tried = 0;
    if (!tried) {
        void *f = nullptr; cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess) fn = (zb_encode_tiled_fn)f;
        tried = 1;
    }
    return fn;
}

// float image nx x ny, box WR x TH, out-of-bounds elements read as zero
static bool conv_tensor_map(CUtensorMap *tm, const float *img, int nx, int ny, int WR, int TH) {
    zb_encode_tiled_fn enc = zb_encode_tiled();
    if (!enc || !img || (nx & 3) || ((uintptr_t)img & 15) || WR > 256 || TH > 256) return false;
    const cuuint64_t gdim[2] = {(cuuint64_t)nx, (cuuint64_t)ny}, gstr[1] = {(cuuint64_t)nx * 4};
    const cuuint32_t box[2] = {(cuuint32_t)WR, (cuuint32_t)TH}, estr[2] = {1, 1};
    return enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void *)img, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
               CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

static inline int min_i(int a, int b) { return a < b ? a : b; }
static inline unsigned up128(size_t v) { return (unsigned)((v + 127) & ~(size_t)127); }

// Shared-memory plan of one tile shape; returns the total bytes.
static size_t conv_plan(ConvGeom &G, int B, int fw, int TX, int TY, int CH, bool have_rms) {
    const int nch = (fw + CH - 1) / CH;
    G.fwp = nch * CH;
    G.kstride = fw * G.fwp;
    G.Wt = (G.NBX - 1) * B + G.GX * TX + G.fwp - 1;
    G.TH = (G.NBY - 1) * B + G.GY * TY + fw - 1;
    G.WR = (G.Wt + 3 + 3) & ~3;      // + up to 3 columns: the box starts at a 16-byte aligned column (TMA traps otherwise)
    G.NW = (G.Wt + 31) / 32 + 2;
    // row pitches of the two planes: fewest shared-memory wavefronts of a window load, summed over every warp of a band
    // of kernel rows (double plane: 8-byte elements, a wavefront serves 16 lanes; float plane: 32 lanes).  Judged on the
    // first two warps only, the pitch left every window load with a 2-way conflict (ncu: 3.8 wavefronts per LDS.64).
    {
        const int tpb = G.GX * G.GY, Tg = CONV_FMA_THREADS / (G.KS > 0 ? G.KS : 1);
        for (int plane = 0; plane < 2; plane++) {
            const int lanes = plane == 0 ? 16 : 32;
            int best = -1, bestw = G.Wt;
            for (int wc = G.Wt; wc < G.Wt + 2 * lanes; wc++) {
                int total = 0;
                for (int u0 = 0; u0 < Tg; u0 += lanes) {
                    int cnt[32] = {0}, worst = 0;
                    long seen[32][8]; int nseen[32] = {0};
                    for (int l = 0; l < lanes; l++) {
                        const int u = u0 + l, lb = u / tpb, lt = u - lb * tpb;
                        if (u >= Tg || lb >= G.NBX * G.NBY) continue;
                        const int gx = lt / G.GY, gy = lt - gx * G.GY, lby = lb / G.NBX, lbx = lb - lby * G.NBX;
                        const long a = (long)(lby * B + gy * TY) * wc + lbx * B + gx * TX;
                        const int bank = (int)(a & (lanes - 1));
                        bool dup = false;                                  // lanes reading the same word share a wavefront
                        for (int k = 0; k < nseen[bank]; k++) dup |= seen[bank][k] == a;
                        if (!dup) { if (nseen[bank] < 8) seen[bank][nseen[bank]++] = a; cnt[bank]++; }
                    }
                    for (int k = 0; k < lanes; k++) worst = cnt[k] > worst ? cnt[k] : worst;
                    total += worst;
                }
                if (best < 0 || total < best) { best = total; bestw = wc; }
            }
            if (plane == 0) G.Wc = bestw; else G.Wv = bestw;
        }
    }
    const int NB = G.NBX * G.NBY, OW = G.NBX * B, OH = G.NBY * B;
    size_t o = 0;
    G.o_raw_img = up128(o); o = G.o_raw_img + (size_t)G.WR * G.TH * 4;
    G.o_raw_rms = up128(o); o = G.o_raw_rms + (have_rms ? (size_t)G.WR * G.TH * 4 : 0);
    G.o_compd = up128(o);   o = G.o_compd + (size_t)G.Wc * G.TH * 8;
    G.o_compv = up128(o);   o = G.o_compv + (size_t)G.Wv * G.TH * 4;
    G.o_k = up128(o);       o = G.o_k + (size_t)NB * G.kstride * sizeof(KTap);
    G.o_resq = up128(o);    o = G.o_resq + (size_t)OH * OW * 8;
    G.o_resv = up128(o);    o = G.o_resv + (size_t)OH * OW * 4;
    G.o_rmask = up128(o);   o = G.o_rmask + (size_t)4 * G.TH * G.NW * 4;
    G.o_vmask = up128(o);   o = G.o_vmask + (size_t)4 * OH * G.NW * 4;
    G.o_pre = up128(o);     o = G.o_pre + (size_t)NB * fw * (G.fwp + 1) * 8;
    G.o_resm = up128(o);    o = G.o_resm + (size_t)OH * OW * 2;
    G.o_misc = up128(o);    o = G.o_misc + 64 + 2 * 16 * 4 + 4 * 4;
    return o;
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
    if (fw > 55) { snprintf(ctx->err, sizeof ctx->err, "kernel width %d > 55 is not built", fw); return ZB_ERR_UNSUPPORTED; }

    // register tile per kernel-block size: (TX, TY, CH) with GX*TX >= B, GY*TY >= B
    int TXr, TYr, CHr;
    if (B <= 12)      { TXr = 4; TYr = 3; CHr = 4; }
    else if (B <= 24) { TXr = 7; TYr = 3; CHr = 7; }
    else              { TXr = 8; TYr = 3; CHr = 8; }
    const bool have_rms = P.conv_rms != nullptr;
    // The tile shape depends on the configuration only (kernel block size and width, rms map or not, -convvar):
    // searched once per context and configuration, not per pair (the search costs the host ~2 ms)
    ConvGeom G;
    const int gkey[6] = {B, fw, have_rms ? 1 : 0, p.convolve_variance ? 1 : 0, TXr, TYr};
    if (ctx->conv_geom_bytes == (int)sizeof(ConvGeom) && memcmp(ctx->conv_geom_key, gkey, sizeof gkey) == 0) {
        memcpy(&G, ctx->conv_geom, sizeof G);
    } else {
        memset(&G, 0, sizeof G);
        G.GX = (B + TXr - 1) / TXr; G.GY = (B + TYr - 1) / TYr;
        const int tpb = G.GX * G.GY;
        if (tpb > CONV_FMA_THREADS) { snprintf(ctx->err, sizeof ctx->err, "kc_step %d too large for the convolution kernel", B); return ZB_ERR_UNSUPPORTED; }
        // Tile shape: NBX x NBY blocks and KS bands of kernel rows.  256 FMA threads, one band per 256/KS of them.
        // Score = busy lanes x a mild preference for little halo; limited by shared memory and the TMA box (256).
        {
            static int env_nbx = -1, env_nby = 0, env_ks = 0;
            if (env_nbx < 0) { const char *e = getenv("ZB_CONV_TILE"); env_nbx = 0; if (e) sscanf(e, "%dx%dx%d", &env_nbx, &env_nby, &env_ks); }
            double best = -1; ConvGeom Gb = G;
            for (int ks = 1; ks <= CONV_MAX_KS; ks *= 2)
                for (int by = 1; by <= 8; by++)
                    for (int bx = 1; bx <= 8; bx++) {
                        if (env_nbx > 0 && !(bx == env_nbx && by == env_nby && ks == env_ks)) continue;
                        const int nb = bx * by;
                        if (nb * tpb > CONV_FMA_THREADS / ks || nb > CONV_MAX_NB || ks > fw) continue;
                        if (p.convolve_variance && ks > 1) continue;               // -convvar: the double variance sum is not split into bands
                        ConvGeom T = G; T.NBX = bx; T.NBY = by; T.KS = ks;
                        const size_t sm = conv_plan(T, B, fw, TXr, TYr, CHr, have_rms);
                        if (sm > (size_t)ZB_SMEM_CONV || T.WR > 256 || T.TH > 256 || T.TH > 16 * 32 - 64) continue;
                        const double util = (double)(nb * tpb * ks) / CONV_FMA_THREADS;
                        const double halo = (double)(nb * B * B) / ((double)T.Wt * T.TH);
                        const int lo = fw / ks, hi = (fw + ks - 1) / ks;               // band imbalance: the tile waits for the widest band
                        const double score = util * ((double)lo / hi * 0.5 + 0.5) * (0.85 + 0.15 * halo) * (1.0 - 0.02 * (ks - 1));
                        if (score > best) { best = score; Gb = T; }
                    }
            if (best < 0) { snprintf(ctx->err, sizeof ctx->err, "no convolution tile fits shared memory (kc_step %d, kernel width %d)", B, fw); return ZB_ERR_UNSUPPORTED; }
            G = Gb;
        }
        static_assert(sizeof(ConvGeom) <= sizeof(ctx->conv_geom), "conv_geom cache too small");
        memcpy(ctx->conv_geom, &G, sizeof G); memcpy(ctx->conv_geom_key, gkey, sizeof gkey); ctx->conv_geom_bytes = (int)sizeof(ConvGeom);
    }
    const size_t smem = conv_plan(G, B, fw, TXr, TYr, CHr, have_rms);
    G.nbx_max = nbx_max; G.nby_max = nby_max;
    if (blk_row1 > nby_max) blk_row1 = nby_max;
    if (blk_row0 < 0) blk_row0 = 0;
    G.blk_row0 = blk_row0; G.blk_row1 = blk_row1;
    G.have_rms = have_rms ? 1 : 0;
    int nt = 0;
    for (int r = 0; r < nR; r++) {
        const ZbRegion &R = H->reg[r];
        G.first[r] = nt;
        if (R.skip) continue;
        const int nbx = (R.rpx + B - 1) / B, nby = min_i((R.rpy + B - 1) / B, blk_row1);
        if (nby <= blk_row0) continue;
        nt += ((nbx + G.NBX - 1) / G.NBX) * ((nby - blk_row0 + G.NBY - 1) / G.NBY);
    }
    for (int r = nR; r <= ZB_MAX_REGIONS; r++) G.first[r] = nt;
    G.n_tiles = nt;

    dim3 gk((nbx_max * nby_max + 1 + MK_NB - 1) / MK_NB + 1, nR);
    k_make_kernels<<<gk, 256, 0, st>>>(ctx->d_dev, ctx->d_sol, ctx->d_kvec, G.fwp, ctx->d_kstat, ctx->d_kaks, nbx_max, nby_max,
                                       (KTap *)ctx->d_ktap, G.kstride); ZB_DBG("k_make_kernels");
    CUtensorMap tm_img, tm_rms;
    memset(&tm_img, 0, sizeof tm_img); memset(&tm_rms, 0, sizeof tm_rms);
    static int env_notma = -1;
    if (env_notma < 0) { const char *e = getenv("ZB_CONV_NOTMA"); env_notma = (e && e[0] == '1') ? 1 : 0; }
    G.use_tma = !env_notma && conv_tensor_map(&tm_img, P.conv, p.nx, p.ny, G.WR, G.TH) &&
                (!have_rms || conv_tensor_map(&tm_rms, P.conv_rms, p.nx, p.ny, G.WR, G.TH));
    int grid = ctx->n_sm > 0 ? ctx->n_sm : 148;
    if (grid > nt) grid = nt;
    zb_zero_small(ctx, ctx->d_tilectr, sizeof(int), st);
    cudaEventRecord(ctx->ev[10], st);          // k_convolve alone: the kernel bench.py's roofline is quoted on
    if (grid > 0) {
#define ZB_CONV_LAUNCH2(TX_, TY_, CH_, EX_, V64_) do { \
        cudaFuncSetAttribute(k_convolve<TX_, TY_, CH_, EX_, V64_>, cudaFuncAttributeMaxDynamicSharedMemorySize, ZB_SMEM_CONV); \
        k_convolve<TX_, TY_, CH_, EX_, V64_><<<grid, CONV_THREADS, smem, st>>>(tm_img, tm_rms, ctx->d_dev, P, G, ctx->d_sol, (const KTap *)ctx->d_ktap, \
            ctx->d_kstat, ctx->d_kaks, ctx->d_tilectr, ctx->d_rdiff, ctx->d_rnoise, ctx->d_omask, diff, noise, out_mask); } while (0)
#define ZB_CONV_LAUNCH(TX_, TY_, CH_) do { \
        if (p.convolve_variance) ZB_CONV_LAUNCH2(TX_, TY_, CH_, false, true); \
        else if ((fw % CH_) == 0) ZB_CONV_LAUNCH2(TX_, TY_, CH_, true, false); \
        else ZB_CONV_LAUNCH2(TX_, TY_, CH_, false, false); } while (0)
        if (B <= 12)      ZB_CONV_LAUNCH(4, 3, 4);
        else if (B <= 24) ZB_CONV_LAUNCH(7, 3, 7);
        else              ZB_CONV_LAUNCH(8, 3, 8);
#undef ZB_CONV_LAUNCH2
#undef ZB_CONV_LAUNCH
    }
    ZB_DBG("k_convolve");
    cudaEventRecord(ctx->ev[11], st);
    k_region_frame<<<dim3(64, nR), 256, 0, st>>>(ctx->d_dev, P, ctx->d_sol, ctx->d_rdiff, ctx->d_rnoise, ctx->d_omask, diff, noise, out_mask, row0, row1); ZB_DBG("k_region_frame");
    ctx->launches += 3;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return zb_cuda_fail(ctx, e, "convolution kernels", __FILE__, __LINE__);
    return ZB_OK;
}
