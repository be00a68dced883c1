// zb_stats.cu — per-region QA numbers that hotpants writes as FITS keywords.
//   k_final_sig / k_final_clip   alard.c:1027-1060 getFinalStampSig + main.c:1417-1435 (FSIGnn / FSCATnn)
//   k_qa_sample / k_qa_pass / k_qa_tail
//                                main.c:1479-1554: getStampStats3 (functions.c:763-1035) on the difference
//                                and noise images for GOOD and OK pixels, getNoiseStats3 (functions.c:719-761).
//                                Two passes over the frame, then the remaining clip iterations on short lists.
//   k_rescale_ok                 main.c:1557-1579 (-okn)
//
// The four getStampStats3 runs of a region read the same three arrays (difference, noise, mask),
// so they advance in lock step: one grid-wide pass reads every region pixel once and feeds all
// four; a tiny "decide" kernel between passes takes the reference's sequential decisions (clip
// loop termination, histogram re-binning).  HBM/L2-bound streaming: 12 bytes per pixel per pass.
// Block partial sums are combined in block order, so results are deterministic.
#include "zb_internal.h"
#include <stdlib.h>
#include "zb_hist.cuh"

// ------------------------------------------------------------------ FSIG / FSCAT

__global__ void __launch_bounds__(256)
k_final_sig(const ZbDev *__restrict__ D, ZbStamp *__restrict__ stamps,
            const float *__restrict__ rdiff, const float *__restrict__ rnoise, const int32_t *__restrict__ omask)
{
    extern __shared__ __align__(16) unsigned char dyn[];
    const int slot = blockIdx.x;
    ZbStamp &S = stamps[slot];
    if (S.nss <= 0 || S.sscnt >= S.nss) return;
    const ZbRegion R = D->reg[S.region];
    if (R.skip) return;
    const int hwss = D->p.hw_ksstamp, fwss = D->sz.fw_ksstamp, npix = fwss * fwss;
    double *term = (double *)dyn;
    signed char *ok = (signed char *)(term + npix);
    const int xR = S.xss[S.sscnt], yR = S.yss[S.sscnt];
    for (int k = threadIdx.x; k < npix; k += blockDim.x) {
        int j = k / fwss, i = k - j * fwss;
        long long idx = R.off + (long long)(xR - hwss + i) + (long long)R.rpx * (yR - hwss + j);
        float idat = rdiff[idx], indat = (float)(1. / rnoise[idx]);
        signed char f = 0;
        if (!(omask[idx] & ZB_FLAG_INPUT_ISBAD)) { f = 1; term[k] = (double)(idat * idat * indat * indat); }
        ok[k] = f;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double sig = 0; int nsig = 0;
        for (int k = 0; k < npix; k++) if (ok[k]) { nsig++; sig += term[k]; }
        S.sig1 = nsig > 0 ? sig / nsig : -1;
    }
}

// functions.c:1037-1107 sigma_clip on a short float array (one thread)
__device__ static void sigma_clip_small(const float *data, char *rej, int count, double *mean, double *stdev, int maxiter, float statSig)
{
    if (count == 0) { *mean = 0; *stdev = ZB_MAXVAL; return; }
    for (int i = 0; i < count; i++) rej[i] = 0;
    int cnt = 0, ncnt = count, iter = 0;
    while (ncnt != cnt && iter < maxiter) {
        cnt = ncnt;
        *mean = 0; *stdev = 0;
        for (int i = 0; i < count; i++) if (!rej[i]) { float d = data[i]; *mean += d; *stdev += (double)(d * d); }
        if (ncnt > 0) *mean /= ncnt; else { *mean = 0; *stdev = ZB_MAXVAL; return; }
        if (ncnt > 1) { *stdev = *stdev - ncnt * (*mean) * (*mean); *stdev = sqrt(*stdev / (double)(ncnt - 1)); }
        else { *stdev = ZB_MAXVAL; return; }
        ncnt = 0;
        double istd = 1. / (*stdev);
        for (int i = 0; i < count; i++)
            if (!rej[i]) { if ((fabs(data[i] - (*mean)) * istd) > statSig) rej[i] = 1; else ncnt++; }
        iter++;
    }
}

__global__ void k_final_clip(const ZbDev *__restrict__ D, const ZbStamp *__restrict__ stamps, const ZbFit *__restrict__ fit,
                             float *__restrict__ ss_scratch, char *__restrict__ rej_scratch, double *__restrict__ q)
{
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= D->n_regions || D->reg[r].skip) return;
    const int base = D->reg[r].stamp_base, nS = fit[r].n_stamps, cap = D->sz.n_stamps;
    float *ss = ss_scratch + (long long)r * cap;
    char *rej = rej_scratch + (long long)r * cap;
    int n = 0;
    for (int k = 0; k < nS; k++) {
        const ZbStamp &S = stamps[base + k];
        if (S.sscnt < S.nss) ss[n++] = (float)S.sig1;
    }
    double mean = 0, sd = 0;
    sigma_clip_small(ss, rej, n, &mean, &sd, 10, D->p.stat_sig);
    q[(long long)r * 64 + 0] = mean;
    q[(long long)r * 64 + 1] = sd;
}

int zb_launch_final_sig(zb_ctx *ctx, cudaStream_t st) {
    const zb_sizes &sz = ctx->sz;
    const int nR = sz.n_regions, npix = sz.fw_ksstamp * sz.fw_ksstamp;
    zb_zero_small(ctx, ctx->d_stats, (size_t)nR * 64 * sizeof(double), st);
    size_t sm = (size_t)npix * (sizeof(double) + 1) + 16;
    cudaFuncSetAttribute(k_final_sig, cudaFuncAttributeMaxDynamicSharedMemorySize, ZB_SMEM_OPTIN);
    k_final_sig<<<ctx->n_slots, 256, sm, st>>>(ctx->d_dev, ctx->d_stamps, ctx->d_rdiff, ctx->d_rnoise, ctx->d_omask); ZB_DBG("k_final_sig");
    // scratch for the clip lives behind the 64 result doubles of every region
    float *ss = (float *)(ctx->d_stats + (size_t)nR * 64);
    char *rej = (char *)(ss + (size_t)nR * sz.n_stamps);
    k_final_clip<<<(nR + 31) / 32, 32, 0, st>>>(ctx->d_dev, ctx->d_stamps, ctx->d_fit, ss, rej, ctx->d_stats); ZB_DBG("k_final_clip");
    ctx->launches += 2;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return zb_cuda_fail(ctx, e, "final sig kernels", __FILE__, __LINE__);
    return ZB_OK;
}

// ------------------------------------------------------------------ region statistics

// stat index: 0 diff/GOOD, 1 noise/GOOD, 2 diff/OK, 3 noise/OK (the order main.c:1479-1535 runs them)
__device__ static inline void qa_masks(int s, int &um, int &sm) {
    if (s < 2) { um = 0x0; sm = 0xffff; } else { um = 0xff; sm = ZB_FLAG_OUTPUT_ISBAD; }
}

__device__ static void qa_reset(ZbQaStat &Q, int s, int um, int sm) {
    Q.phase = 0; Q.status = 0; Q.maxiter = 5; Q.umask = um; Q.smask = sm; Q.use_noise = s & 1;
    Q.ncrit = 0; Q.cnt_prev = 0; Q.tries = 0; Q.imax = 0; Q.ticket = 0; Q.exact = 0;
    Q.clo = Q.llo = -3.402823466e38f; Q.chi = Q.lhi = 3.402823466e38f;
    Q.mean = Q.sd = Q.mode = Q.median = Q.fwhm = Q.lfwhm = Q.sum = 0;
    Q.base_s1 = Q.base_s2 = 0; Q.base_n = 0; Q.shell_bad = 0; Q.core_lo = Q.ncore_lo = 1.f; Q.core_hi = Q.ncore_hi = -1.f;
}

// The first 100 accepted random samples (functions.c:811-833), accepted in attempt order.
__global__ void __launch_bounds__(256)
k_qa_sample(const ZbDev *__restrict__ D, const float *__restrict__ rdiff, const float *__restrict__ rnoise,
            const int32_t *__restrict__ omask, ZbQaStat *__restrict__ qa, const double *__restrict__ ran, int n_ran, int hist)
{
    const int s = blockIdx.x, r = blockIdx.y, tid = threadIdx.x;
    const ZbRegion R = D->reg[r];
    ZbQaStat &Q = qa[r * 4 + s];
    __shared__ double work[100];
    __shared__ int wsum[8];
    __shared__ int s_acc;
    if (R.skip) { if (tid == 0) { Q.phase = 2; Q.status = -1; } return; }
    const int npts = R.rpx * R.rpy;
    const float *data = ((s & 1) ? rnoise : rdiff) + R.off;
    const int32_t *m = omask + R.off;
    int um, sm; qa_masks(s, um, sm);
    for (int k = tid; k < 256; k += 256) Q.bins[k] = 0;
    if (tid < 100) work[tid] = 0.0;
    if (tid == 0) s_acc = 0;
    __syncthreads();
    if (npts < 100) {
        if (tid == 0) { Q.phase = 2; Q.status = 4; Q.mean = Q.sd = Q.mode = Q.median = Q.fwhm = Q.lfwhm = Q.sum = 0; }
        return;
    }
    if (!hist) {
        // The reference's sample only sizes the histogram (functions.c:811-838).  Without the histogram a sample serves one
        // purpose: a robust guess of where the bulk of the values lies, the *core* whose complement pass 0 appends to the
        // shell lists (k_qa_pass<SHELL>).  It affects which values are listed, never a result.  GOOD-pixel statistics:
        // the first 256 draws of the same random sequence, median +- 2 x half the 16..84 % range.  OK-pixel statistics
        // (a percent of the pixels): no core at all, every value is listed.
        __shared__ float sv[256], so[256];
        float c_lo = 3.402823466e38f, c_hi = -3.402823466e38f;       // empty core
        if (s < 2) {
            int flag = 0; float rdat = 0.f;
            if (2 * tid + 1 < n_ran) {
                const int xr = (int)floor(ran[2 * tid] * R.rpx), yr = (int)floor(ran[2 * tid + 1] * R.rpy);
                rdat = data[xr + yr * R.rpx];
                flag = !zb_excluded(m[xr + yr * R.rpx], rdat, um, sm) && !(rdat * 0.0f != 0.0f);
            }
            const unsigned bal = __ballot_sync(0xffffffffu, flag);
            if ((tid & 31) == 0) wsum[tid >> 5] = __popc(bal);
            __syncthreads();
            int before = 0, n = 0;
            for (int w = 0; w < 8; w++) { if (w < (tid >> 5)) before += wsum[w]; n += wsum[w]; }
            const int pos = before + __popc(bal & ((1u << (tid & 31)) - 1u));
            if (flag) sv[pos] = rdat;
            __syncthreads();
            if (n >= 32) {
                if (tid < n) {                                          // rank sort: ties broken by position
                    const float v = sv[tid]; int rank = 0;
                    for (int j = 0; j < n; j++) { const float u = sv[j]; rank += (u < v) || (u == v && j < tid); }
                    so[rank] = v;
                }
                __syncthreads();
                if (tid == 0) {
                    const float med = so[n / 2], half = 0.5f * (so[(int)(0.84f * n)] - so[(int)(0.16f * n)]);
                    c_lo = med - 2.f * half; c_hi = med + 2.f * half;
                }
            }
        }
        if (tid == 0) {
            Q.npts_s = 0; Q.samp_med = Q.binsize = Q.bin1 = 0.;
            qa_reset(Q, s, um, sm);
            Q.ncore_lo = c_lo; Q.ncore_hi = c_hi;
        }
        return;
    }
    for (int base = 0; base < npts; base += 256) {
        const int a = base + tid;
        int flag = 0; double rdat = 0;
        if (a < npts && 2 * a + 1 < n_ran) {
            int xr = (int)floor(ran[2 * a] * R.rpx), yr = (int)floor(ran[2 * a + 1] * R.rpy);
            rdat = data[xr + yr * R.rpx];
            flag = !zb_excluded(m[xr + yr * R.rpx], rdat, um, sm);
        }
        unsigned bal = __ballot_sync(0xffffffffu, flag);
        if ((tid & 31) == 0) wsum[tid >> 5] = __popc(bal);
        __syncthreads();
        int before = s_acc;
        for (int w = 0; w < (tid >> 5); w++) before += wsum[w];
        int pos = before + __popc(bal & ((1u << (tid & 31)) - 1u));
        if (flag && pos < 100) work[pos] = rdat;
        __syncthreads();
        if (tid == 0) { int t = 0; for (int w = 0; w < 8; w++) t += wsum[w]; s_acc += t; }
        __syncthreads();
        if (s_acc >= 100) break;
    }
    if (tid == 0) {
        int n = s_acc < 100 ? s_acc : 100;
        for (int a = 1; a < n; a++) {
            double v = work[a]; int b = a - 1;
            while (b >= 0 && work[b] > v) { work[b + 1] = work[b]; b--; }
            work[b + 1] = v;
        }
        Q.npts_s = n;
        Q.samp_med = work[(int)(0.5f * n)];
        Q.binsize = (work[(int)(0.9f * n)] - work[(int)(0.5f * n)]) / (float)100;
        Q.bin1 = work[(int)(0.5f * n)] - 128. * Q.binsize;
        qa_reset(Q, s, um, sm);
    }
}

struct QaShared {
    int phase[4], ncrit[4], exact[4];
    float flo[4], fhi[4];
    double cm[4][8], ci[4][8], mean[4], isd[4], bin1[4], binsize[4], ibin[4];
    float core_lo[4], core_hi[4];
    int bins[4][256];
    double ra[8], rb[8]; int rc[8];
};

__device__ static void qa_block_sum(QaShared &S, double &a, double &b, int &c) {
    for (int o = 16; o > 0; o >>= 1) {
        a += __shfl_xor_sync(0xffffffffu, a, o);
        b += __shfl_xor_sync(0xffffffffu, b, o);
        c += __shfl_xor_sync(0xffffffffu, c, o);
    }
    const int w = threadIdx.x >> 5;
    __syncthreads();
    if ((threadIdx.x & 31) == 0) { S.ra[w] = a; S.rb[w] = b; S.rc[w] = c; }
    __syncthreads();
    double sa = 0, sb = 0; int sc = 0;
    for (int k = 0; k < 8; k++) { sa += S.ra[k]; sb += S.rb[k]; sc += S.rc[k]; }
    a = sa; b = sb; c = sc;
}

// ---- float intervals for the clip test --------------------------------------------------------
// The reference rejects a pixel when (fabs((double)x - mean) * istd) > statSig.  For finite mean and
// finite positive istd that predicate is monotone in x on either side of the mean, so the accepted
// floats form one interval [lo, hi]; its end points are found by bisection over the float ordering
// with the reference's own double expression, and the per-pixel test becomes two float compares
// with identical outcomes.
__device__ static inline int qa_f2o(float f) { int i = __float_as_int(f); return i >= 0 ? i : (int)0x80000000 - i; }
__device__ static inline float qa_o2f(int o) { return __int_as_float(o >= 0 ? o : (int)0x80000000 - o); }
__device__ static inline bool qa_keep(float x, double mean, double istd, float sig) { return !((fabs((double)x - mean) * istd) > sig); }

__device__ static bool qa_interval(double mean, double istd, float sig, float &lo, float &hi)
{
    if (!(istd > 0.) || istd * 0.0 != 0.0 || mean * 0.0 != 0.0 || !(fabs(mean) < 1e38)) return false;   // caller keeps the double test
    const float c = (float)mean;
    if (!qa_keep(c, mean, istd, sig)) { lo = 1.f; hi = -1.f; return true; }       // nothing passes
    int a = qa_f2o(c), b = qa_f2o(3.402823466e38f);
    if (qa_keep(qa_o2f(b), mean, istd, sig)) a = b;
    while (b - a > 1) { int m = a + (b - a) / 2; if (qa_keep(qa_o2f(m), mean, istd, sig)) a = m; else b = m; }
    hi = qa_o2f(a);
    a = qa_f2o(c); b = qa_f2o(-3.402823466e38f);
    if (qa_keep(qa_o2f(b), mean, istd, sig)) a = b;
    while (a - b > 1) { int m = b + (a - b) / 2; if (qa_keep(qa_o2f(m), mean, istd, sig)) a = m; else b = m; }
    lo = qa_o2f(a);
    return true;
}

// The sequential decisions of sigma_clip (functions.c:1037-1107) and of the re-binning loop
// (functions.c:879-1031) for one (region, stat); s == 4 finishes getNoiseStats3 on the first pass.
// Runs in the last CTA of a pass to finish (see k_qa_pass).
__device__ static void qa_decide(const ZbDev *D, ZbQaStat *qa, const ZbQaPart *P /* this statistic's nblk partial sums */, int nblk, double *q, int r, int s, int first, const int *bins,
                                 int pass, int hist)
{
    double s1 = 0, s2 = 0; long long n = 0;
    if (s == 4) {
        if (!first) return;
        for (int b = 0; b < nblk; b++) { s1 += P[b].s1; n += P[b].n; }
        q[(long long)r * 64 + 3] = (double)n;
        q[(long long)r * 64 + 2] = (n > 1) ? s1 / (float)n : ZB_MAXVAL;
        return;
    }
    ZbQaStat &Q = qa[r * 4 + s];
    if (Q.phase == 2) return;
    for (int b = 0; b < nblk; b++) { s1 += P[b].s1; s2 += P[b].s2; n += P[b].n; }
    const int ncnt = (int)n;
    const float sig = D->p.stat_sig;
    if (Q.phase == 0) {
        const int k = Q.ncrit;                          // passes done so far == iterations started
        const int after = hist ? 1 : 2;                 // the histogram outputs reach no product (see zb_launch_region_stats)
        if (pass >= 0) {                                // sums over the interval this pass kept, and the core its shell lists were cut at
            Q.base_s1 = s1; Q.base_s2 = s2; Q.base_n = ncnt;
            Q.core_lo = Q.ncore_lo; Q.core_hi = Q.ncore_hi;
        }
        if (k == 0 && ncnt == 0) { Q.mean = 0; Q.sd = ZB_MAXVAL; Q.status = 5; Q.phase = 2; return; }
        if (ncnt != Q.cnt_prev && k < Q.maxiter) {
            Q.cnt_prev = ncnt;
            if (ncnt > 0) Q.mean = s1 / ncnt;
            else { Q.mean = 0; Q.sd = ZB_MAXVAL; Q.status = 5; Q.phase = 2; return; }
            if (ncnt > 1) { double sd = s2 - ncnt * Q.mean * Q.mean; Q.sd = sqrt(sd / (double)(ncnt - 1)); }
            else { Q.sd = ZB_MAXVAL; Q.status = 5; Q.phase = 2; return; }
            Q.crit_mean[k] = Q.mean; Q.crit_istd[k] = 1. / Q.sd; Q.ncrit = k + 1;
            if (pass >= 0) {                            // a frame pass: values outside this core go to the shell list of the next one
                const double w = 0.75 * (double)sig * Q.sd;
                Q.ncore_lo = (float)(Q.mean - w); Q.ncore_hi = (float)(Q.mean + w);
            }
            float lo, hi;
            if (!Q.exact && qa_interval(Q.mean, 1. / Q.sd, sig, lo, hi)) {
                Q.llo = lo; Q.lhi = hi;
                Q.clo = fmaxf(Q.clo, lo); Q.chi = fminf(Q.chi, hi);
            } else Q.exact = 1;
            if (k + 1 >= Q.maxiter) Q.phase = after;    // iter == maxiter ends the loop whatever the next count is
        } else Q.phase = after;
        return;
    }
    // histogram phase: this pass filled the bins with the current bin1/binsize
    const int goodcnt = ncnt;
    if (goodcnt == 0) { Q.mode = Q.median = Q.samp_med; Q.fwhm = Q.lfwhm = 0.; Q.status = 2; Q.phase = 2; return; }
    if (Q.binsize == 0.) { Q.mode = Q.median = Q.samp_med; Q.fwhm = Q.lfwhm = 0.; Q.status = 3; Q.phase = 2; return; }
    HistOut o; int imax = Q.imax;
    hist_analyse(bins, goodcnt, Q.bin1, Q.binsize, imax, o);
    Q.imax = imax; Q.mode = o.mode;
    if (o.action == 0) { Q.sum = s1; Q.fwhm = o.fwhm; Q.lfwhm = o.lfwhm; Q.median = o.median; Q.status = 0; Q.phase = 2; return; }
    if (o.action == 1) { Q.bin1 = Q.bin1 - 128. * Q.binsize; Q.binsize = Q.binsize * 2.; }
    else { Q.binsize = Q.binsize / 3.; Q.bin1 = o.mode - 128. * Q.binsize; }
    Q.tries++;
    for (int b = 0; b < 256; b++) Q.bins[b] = 0;
    if (Q.tries >= 5) { Q.status = 1; Q.phase = 2; }
}

// SHELL (the pass after the first criterion): every kept value outside the criterion's core interval is also
// appended to the shell list of its (statistic, warp) — in pixel order, positions from warp ballots, so the
// lists and every sum formed from them are deterministic.  k_qa_tail runs the remaining clip iterations on them.
template <int MINB, int ROT, bool SHELL>
__global__ void __launch_bounds__(256, MINB)
k_qa_pass(const ZbDev *__restrict__ D, const float *__restrict__ rdiff, const float *__restrict__ rnoise,
          int32_t *__restrict__ omask, int32_t *__restrict__ out_mask, ZbQaStat *qa,
          ZbQaPart *part, double *__restrict__ q, int first, int use_div, int pass, int hist,
          float *__restrict__ shell, int *__restrict__ shell_n, int shell_cap)
{
    __shared__ QaShared S;
    __shared__ int s_last;
    const int r = blockIdx.y, tid = threadIdx.x;
    const ZbRegion R = D->reg[r];
    if (R.skip) return;
    if (tid < 4) {
        const ZbQaStat &Q = qa[r * 4 + tid];
        S.phase[tid] = Q.phase; S.ncrit[tid] = Q.ncrit; S.exact[tid] = Q.exact;
        for (int c = 0; c < 8; c++) { S.cm[tid][c] = Q.crit_mean[c]; S.ci[tid][c] = Q.crit_istd[c]; }
        S.mean[tid] = Q.mean; S.isd[tid] = 1. / Q.sd; S.bin1[tid] = Q.bin1; S.binsize[tid] = Q.binsize; S.ibin[tid] = use_div ? __longlong_as_double(0x7ff8000000000000ll) : 1. / Q.binsize;
        S.flo[tid] = Q.phase == 0 ? Q.clo : Q.llo; S.fhi[tid] = Q.phase == 0 ? Q.chi : Q.lhi;
        S.core_lo[tid] = Q.ncore_lo; S.core_hi[tid] = Q.ncore_hi;
    }
    for (int k = tid; k < 4 * 256; k += 256) (&S.bins[0][0])[k] = 0;
    __syncthreads();
    if (!first && S.phase[0] == 2 && S.phase[1] == 2 && S.phase[2] == 2 && S.phase[3] == 2) return;
    const float statSig = D->p.stat_sig;
    const int npix = R.rpx * R.rpy;
    const int nblk = gridDim.x;                     // CTAs per region: ZB_QA_NBLK_MIN .. ZB_QA_NBLK (zb_qa_nblk)
    double a1[4] = {0, 0, 0, 0}, a2[4] = {0, 0, 0, 0}; int an[4] = {0, 0, 0, 0};
    double nsum = 0; int nn = 0;
    // one statistic's share of a pixel; `s` is a compile-time constant at every call site
#define QA_ACC(s, xv, mk) do { \
        if (S.phase[s] != 2) { int um_, sm_; qa_masks(s, um_, sm_); \
            if (!zb_excluded(mk, xv, um_, sm_) && !((xv) * 0.0f != 0.0f)) { \
                bool keep_; \
                if (!S.exact[s]) keep_ = !((xv) < S.flo[s] || (xv) > S.fhi[s]); \
                else if (S.phase[s] == 0) { keep_ = true; for (int c_ = 0; c_ < S.ncrit[s]; c_++) keep_ &= !((fabs((xv) - S.cm[s][c_]) * S.ci[s][c_]) > statSig); } \
                else keep_ = !((fabs((xv) - S.mean[s]) * S.isd[s]) > statSig); \
                if (keep_) { \
                    if (S.phase[s] == 0) { a1[s] += (xv); a2[s] += (double)((xv) * (xv)); an[s]++; } \
                    else { \
                        /* floor((x - bin1) / binsize) as the reference forms it; the division is replaced by a */ \
                        /* multiplication whenever the quotient is provably not within rounding of an integer   */ \
                        const double t_ = (xv) - S.bin1[s]; \
                        const double q_ = t_ * S.ibin[s]; \
                        double fl_ = floor(q_); \
                        { const double fr_ = q_ - fl_, eps_ = fabs(q_) * 8.9e-16 + 1e-300; \
                          if (!(fr_ >= eps_ && fr_ <= 1.0 - eps_)) fl_ = floor(t_ / S.binsize[s]); } \
                        int index = (int)(fl_ + 1); \
                        index = (index < 0 ? 0 : index); index = (index > 255 ? 255 : index); \
                        const unsigned peers = __match_any_sync(__activemask(), index); \
                        if ((int)(__ffs(peers) - 1) == (threadIdx.x & 31)) atomicAdd(&S.bins[s][index], __popc(peers)); \
                        a1[s] += fabs((double)(xv)); an[s]++; } } } } } while (0)
    const int stride = nblk * 256;
    // every statistic either finished (empty interval: nothing is kept, its sums are not used) or in a clip pass
    bool fast = true;
    float f_lo[4], f_hi[4];
#pragma unroll
    for (int q = 0; q < 4; q++) {
        fast &= S.phase[q] == 2 || (S.phase[q] == 0 && !S.exact[q]);
        f_lo[q] = S.phase[q] == 2 ? __int_as_float(0x7f800000) : S.flo[q];
        f_hi[q] = S.phase[q] == 2 ? __int_as_float(0xff800000) : S.fhi[q];
    }
    const int lane = tid & 31;
    int scnt[4] = {0, 0, 0, 0};                     // SHELL: values appended so far to this warp's four segments
    float c_lo[4], c_hi[4];
#pragma unroll
    for (int q = 0; q < 4; q++) { c_lo[q] = S.core_lo[q]; c_hi[q] = S.core_hi[q]; }
    // warp-uniform trip count (lanes past the end carry excluded dummy pixels): the shell ballots need whole warps
    for (int kb0 = blockIdx.x * 256 + (tid & ~31); kb0 < npix; kb0 += 4 * stride) {
        const int kb = kb0 + lane;
        float dv[4], nv[4]; int mv[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {                   // 12 independent loads in flight per thread
            const int k = kb + u * stride;
            if (k < npix) { dv[u] = rdiff[R.off + k]; nv[u] = rnoise[R.off + k]; mv[u] = omask[R.off + k]; }
            else { dv[u] = 0.f; nv[u] = 0.f; mv[u] = 0xffff; }      // excluded from everything
        }
        // Common case — every statistic either finished or in a clip pass with a float interval, every
        // value of the batch finite: branch-light code with the interval ends in registers.  It adds the
        // same terms to the same per-thread sums in the same order as the generic code below (a rejected
        // pixel adds 0.0), so the two paths are interchangeable bit for bit.
        // (a NaN/Inf only matters under a mask value that lets the pixel into one of the four runs)
        bool nonfinite = false;
#pragma unroll
        for (int u = 0; u < 4; u++)
            nonfinite |= ((dv[u] * 0.0f != 0.0f) || (nv[u] * 0.0f != 0.0f))
                         && ((mv[u] & 0xffff) == 0 || ((mv[u] & 0xff) && !(mv[u] & ZB_FLAG_OUTPUT_ISBAD)));
        if (SHELL) {
            const bool plain = fast && !__any_sync(0xffffffffu, nonfinite);
            if (plain) {
                const unsigned lt = (1u << lane) - 1u;
                const size_t seg0 = ((size_t)(r * 4) * nblk + blockIdx.x) * 8 + (tid >> 5);     // statistic 0's segment
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    const float d = dv[u], nz = nv[u];
                    const int m0 = mv[u];
                    const bool good = (m0 & 0xffff) == 0;
                    const bool okp = !good && (m0 & 0xff) && !(m0 & ZB_FLAG_OUTPUT_ISBAD);
                    const double dd = (double)d, dn = (double)nz;
                    const double d2 = (double)(d * d), n2 = (double)(nz * nz);
                    const bool usable_d = !(fabs(dd) <= ZB_ZEROVAL), usable_n = !(fabs(dn) <= ZB_ZEROVAL);
                    if (first && good && !(fabsf(d) < 1e-10f)) {        // getNoiseStats3, functions.c:719-761
                        const float ndat = (float)(1. / nz);
                        nn++; nsum += (double)(d * d * ndat * ndat);
                    }
                    bool keep[4];
                    keep[0] = good && usable_d && !(d < f_lo[0] || d > f_hi[0]);
                    keep[1] = good && usable_n && !(nz < f_lo[1] || nz > f_hi[1]);
                    keep[2] = okp && usable_d && !(d < f_lo[2] || d > f_hi[2]);
                    keep[3] = okp && usable_n && !(nz < f_lo[3] || nz > f_hi[3]);
#pragma unroll
                    for (int s = 0; s < 4; s++) {
                        const float v = (s & 1) ? nz : d;
                        if ((s < 2) ? good : okp) {                     // the same terms, in the same order, as the plain form below
                            a1[s] += keep[s] ? ((s & 1) ? dn : dd) : 0.0; a2[s] += keep[s] ? ((s & 1) ? n2 : d2) : 0.0; an[s] += keep[s];
                        }
                        const bool sh = keep[s] && (v < c_lo[s] || v > c_hi[s]);
                        const unsigned b = __ballot_sync(0xffffffffu, sh);
                        if (b) {
                            const int pos = scnt[s] + __popc(b & lt);
                            if (sh && pos < shell_cap) shell[(seg0 + (size_t)s * nblk * 8) * shell_cap + pos] = v;
                            scnt[s] += __popc(b);
                        }
                    }
                }
                continue;
            }
            // a statistic clipping with the exact criteria, or a NaN/Inf in the batch: the lists are not complete
            if (lane == 0) for (int s = 0; s < 4; s++) atomicOr(&qa[r * 4 + s].shell_bad, 1);
        }
        if (!SHELL && fast && !nonfinite) {
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const float d = dv[u], nz = nv[u];
                const int m0 = mv[u];
                const bool good = (m0 & 0xffff) == 0;
                const bool okp = !good && (m0 & 0xff) && !(m0 & ZB_FLAG_OUTPUT_ISBAD);
                if (!(good || okp)) continue;
                const double dd = (double)d, dn = (double)nz;
                const double d2 = (double)(d * d), n2 = (double)(nz * nz);
                const bool usable_d = !(fabs(dd) <= ZB_ZEROVAL), usable_n = !(fabs(dn) <= ZB_ZEROVAL);
                if (good) {
                    if (first && !(fabsf(d) < 1e-10f)) {        // getNoiseStats3, functions.c:719-761
                        float ndat = (float)(1. / nz);
                        nn++; nsum += (double)(d * d * ndat * ndat);
                    }
                    const bool kd = usable_d && !(d < f_lo[0] || d > f_hi[0]), kn = usable_n && !(nz < f_lo[1] || nz > f_hi[1]);
                    a1[0] += kd ? dd : 0.0; a2[0] += kd ? d2 : 0.0; an[0] += kd;
                    a1[1] += kn ? dn : 0.0; a2[1] += kn ? n2 : 0.0; an[1] += kn;
                } else {
                    const bool kd = usable_d && !(d < f_lo[2] || d > f_hi[2]), kn = usable_n && !(nz < f_lo[3] || nz > f_hi[3]);
                    a1[2] += kd ? dd : 0.0; a2[2] += kd ? d2 : 0.0; an[2] += kd;
                    a1[3] += kn ? dn : 0.0; a2[3] += kn ? n2 : 0.0; an[3] += kn;
                }
            }
            continue;
        }
        // generic form (a statistic in its histogram phase or clipping with the exact criteria, or a NaN/Inf in
        // the batch): ONE copy of the code, the batch rotated through element 0 — unrolled four times this
        // kernel was 13.7 k instructions and stalled on instruction fetch
#pragma unroll (ROT ? 1 : 4)
        for (int u = 0; u < 4; u++) {
            const int k = kb + u * stride;
            if (k >= npix) break;
            const float d = dv[ROT ? 0 : u], nz = nv[ROT ? 0 : u];
            const int m0 = mv[ROT ? 0 : u];
            if (ROT) {
                dv[0] = dv[1]; dv[1] = dv[2]; dv[2] = dv[3];
                nv[0] = nv[1]; nv[1] = nv[2]; nv[2] = nv[3];
                mv[0] = mv[1]; mv[1] = mv[2]; mv[2] = mv[3];
            }
            const bool nfd = (d * 0.0f != 0.0f), nfn = (nz * 0.0f != 0.0f);
            if (!(nfd || nfn)) {
                // GOOD pixels (no mask bit at all) and OK pixels (a bit in 0xff, not OUTPUT_ISBAD) are disjoint
                if ((m0 & 0xffff) == 0) {
                    if (first && !(fabsf(d) < 1e-10f)) {        // getNoiseStats3, functions.c:719-761
                        float ndat = (float)(1. / nz);
                        nn++; nsum += (double)(d * d * ndat * ndat);
                    }
                    QA_ACC(0, d, m0); QA_ACC(1, nz, m0);
                } else if ((m0 & 0xff) && !(m0 & ZB_FLAG_OUTPUT_ISBAD)) {
                    QA_ACC(2, d, m0); QA_ACC(3, nz, m0);
                }
                continue;
            }
            // NaN/Inf side effect of getStampStats3 (functions.c:857-861), in the order the runs happen
            int ms[4] = {m0, m0, m0, m0}, mn = m0;
            const int mark = ZB_FLAG_INPUT_ISBAD | ZB_FLAG_ISNAN;
            int m = m0;
            if (nfd && !zb_excluded(m, d, 0x0, 0xffff)) m |= mark;
            ms[1] = m;
            if (nfn && !zb_excluded(m, nz, 0x0, 0xffff)) m |= mark;
            mn = m; ms[2] = m;
            if (nfd && !zb_excluded(m, d, 0xff, ZB_FLAG_OUTPUT_ISBAD)) m |= mark;
            ms[3] = m;
            if (nfn && !zb_excluded(m, nz, 0xff, ZB_FLAG_OUTPUT_ISBAD)) m |= mark;
            if (first && m != m0) {
                omask[R.off + k] = m;
                const int y = k / R.rpx, x = k - y * R.rpx;
                if (out_mask && x >= R.rx0 - R.bx0 && x <= R.own_x1 - R.bx0 && y >= R.ry0 - R.by0 && y <= R.own_y1 - R.by0)
                    out_mask[(long long)(R.by0 + y) * D->nx + (R.bx0 + x)] = m;
            }
            if (first && !zb_excluded(mn, d, 0x0, 0xffff)) {
                float ndat = (float)(1. / nz);
                nn++; nsum += (double)(d * d * ndat * ndat);
            }
            QA_ACC(0, d, ms[0]); QA_ACC(1, nz, ms[1]); QA_ACC(2, d, ms[2]); QA_ACC(3, nz, ms[3]);
        }
    }
#undef QA_ACC
    if (SHELL && lane == 0) {
#pragma unroll
        for (int s = 0; s < 4; s++) shell_n[((size_t)(r * 4 + s) * nblk + blockIdx.x) * 8 + (tid >> 5)] = scnt[s];
    }
#pragma unroll                                   // static indices keep a1/a2/an in registers
    for (int s = 0; s < 4; s++) {
        if (S.phase[s] == 2) continue;
        qa_block_sum(S, a1[s], a2[s], an[s]);
        if (tid == 0) { ZbQaPart &P = part[(r * 5 + s) * ZB_QA_NBLK + blockIdx.x]; P.s1 = a1[s]; P.s2 = a2[s]; P.n = an[s]; }
        if (S.phase[s] == 1) {
            __syncthreads();
            int v = S.bins[s][tid];
            if (v) atomicAdd(&qa[r * 4 + s].bins[tid], v);
        }
    }
    if (first) {
        double z = 0;
        qa_block_sum(S, nsum, z, nn);
        if (tid == 0) { ZbQaPart &P = part[(r * 5 + 4) * ZB_QA_NBLK + blockIdx.x]; P.s1 = nsum; P.s2 = 0; P.n = nn; }
    }
    // the last CTA of this region to get here takes the decisions for the next pass
    __threadfence();
    __syncthreads();
    if (tid == 0) s_last = (atomicAdd(&qa[r * 4].ticket, 1) == nblk - 1);
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    for (int k = tid; k < 4 * 256; k += 256) (&S.bins[0][0])[k] = __ldcg(&qa[r * 4 + (k >> 8)].bins[k & 255]);
    __shared__ ZbQaPart s_part[5 * ZB_QA_NBLK];
    for (int k = tid; k < 5 * ZB_QA_NBLK; k += 256) {
        const ZbQaPart *g = part + (size_t)r * 5 * ZB_QA_NBLK + k;
        s_part[k].s1 = __ldcg(&g->s1); s_part[k].s2 = __ldcg(&g->s2); s_part[k].n = __ldcg(&g->n);
    }
    __syncthreads();
    if (tid < 5) qa_decide(D, qa, s_part + tid * ZB_QA_NBLK, nblk, q, r, tid, first, tid < 4 ? S.bins[tid] : S.bins[0], pass, hist);
    if (tid == 0) qa[r * 4].ticket = 0;
}

// The clip iterations of one (region, statistic) that follow a shell pass, without another pass over the frame.  The
// pass left the sums over the values it kept (base) and, on the shell lists, every kept value outside its core.  As
// long as the accepted interval still contains the core, every value a later criterion rejects is on the lists, and the
// iteration's sums are the base sums minus the sums over the rejected list values.  The same decisions as a full pass
// (qa_decide); if the containment ever fails, or the lists are incomplete, the state is left where it is and the frame
// passes that follow carry on from it.
#define QA_TAIL_NT 1024
__global__ void __launch_bounds__(QA_TAIL_NT)
k_qa_tail(const ZbDev *__restrict__ D, ZbQaStat *qa, double *__restrict__ q,
          const float *__restrict__ shell, const int *__restrict__ shell_n, int shell_cap, int nblk, int pass)
{
    const int s = blockIdx.x, r = blockIdx.y, tid = threadIdx.x;
    if (D->reg[r].skip) return;
    ZbQaStat &Q = qa[r * 4 + s];
    __shared__ double w1[QA_TAIL_NT / 32], w2[QA_TAIL_NT / 32];
    __shared__ int wn[QA_TAIL_NT / 32];
    __shared__ int s_go, s_dummy_bins[1];
    __shared__ float s_lo, s_hi;
    __shared__ ZbQaPart s_one;
    const int nseg = nblk * 8;
    const size_t seg0 = (size_t)(r * 4 + s) * nseg;
    // uniform over the CTA (nobody has written yet); base sums, core and lists are those of the shell pass that has just run
    const int usable = Q.phase == 0 && !Q.shell_bad && !Q.exact;
    __syncthreads();
    if (tid == 0) Q.shell_bad = 0;                                               // for the next shell pass, if there is one
    if (!usable) return;
    // overflow of any segment: the list is incomplete
    int over = 0;
    for (int g = tid; g < nseg; g += QA_TAIL_NT) over |= shell_n[seg0 + g] > shell_cap;
    if (__syncthreads_or(over)) return;
    for (int it = 0; it < 8; it++) {
        if (tid == 0) {
            s_go = Q.phase == 0 && !Q.exact && Q.clo <= Q.core_lo && Q.chi >= Q.core_hi;
            s_lo = Q.clo; s_hi = Q.chi;
        }
        __syncthreads();
        if (!s_go) return;
        const float lo = s_lo, hi = s_hi;
        double r1 = 0, r2 = 0; int rn = 0;
        // fixed assignment: thread t takes segments t, t + QA_TAIL_NT, ... and walks each in order: the sums do not depend on timing
        for (int g = tid; g < nseg; g += QA_TAIL_NT) {
            const int n = shell_n[seg0 + g];
            const float *v = shell + (seg0 + g) * shell_cap;
            int k = 0;
            for (; k + 8 <= n; k += 8) {
                float x[8];
#pragma unroll
                for (int e = 0; e < 8; e++) x[e] = v[k + e];
#pragma unroll
                for (int e = 0; e < 8; e++)
                    if (x[e] < lo || x[e] > hi) { r1 += (double)x[e]; r2 += (double)(x[e] * x[e]); rn++; }
            }
            for (; k < n; k++) {
                const float x = v[k];
                if (x < lo || x > hi) { r1 += (double)x; r2 += (double)(x * x); rn++; }
            }
        }
        for (int o = 16; o > 0; o >>= 1) {
            r1 += __shfl_xor_sync(0xffffffffu, r1, o);
            r2 += __shfl_xor_sync(0xffffffffu, r2, o);
            rn += __shfl_xor_sync(0xffffffffu, rn, o);
        }
        if ((tid & 31) == 0) { w1[tid >> 5] = r1; w2[tid >> 5] = r2; wn[tid >> 5] = rn; }
        __syncthreads();
        if (tid == 0) {
            double t1 = 0, t2 = 0; int tn = 0;
            for (int w = 0; w < QA_TAIL_NT / 32; w++) { t1 += w1[w]; t2 += w2[w]; tn += wn[w]; }
            s_one.s1 = Q.base_s1 - t1; s_one.s2 = Q.base_s2 - t2; s_one.n = Q.base_n - tn;
            qa_decide(D, qa, &s_one, 1, q, r, s, 0, s_dummy_bins, -1, 0);        // not a frame pass: base sums and core stay
        }
        __syncthreads();
    }
}

__global__ void k_qa_finish(const ZbDev *__restrict__ D, const ZbQaStat *__restrict__ qa, double *__restrict__ q)
{
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= D->n_regions || D->reg[r].skip) return;
    const ZbQaStat *Q = qa + r * 4;
    double *o = q + (long long)r * 64;
    o[4] = Q[0].mean; o[5] = Q[0].sd; o[6] = Q[1].mean;
    o[7] = Q[2].mean; o[8] = Q[2].sd; o[9] = Q[3].mean;
    double nscalo = Q[0].sd / Q[1].mean;
    if (D->p.rescale_ok) nscalo = (Q[2].sd / Q[3].mean) / nscalo;       // main.c:1557-1561
    o[10] = nscalo;
}

// -okn: scale the noise of OK pixels (main.c:1563-1576); output pixels only
__global__ void k_rescale_ok(const ZbDev *__restrict__ D, const int32_t *__restrict__ omask, const double *__restrict__ q,
                             float *__restrict__ noise)
{
    const int r = blockIdx.y;
    const ZbRegion R = D->reg[r];
    if (R.skip) return;
    const double diffrat = q[(long long)r * 64 + 10];
    if (!(diffrat > 1)) return;
    const int npix = R.rpx * R.rpy;
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < npix; k += gridDim.x * blockDim.x) {
        const int m = omask[R.off + k];
        if (!((m & 0xff) && !(m & ZB_FLAG_OUTPUT_ISBAD))) continue;
        const int y = k / R.rpx, x = k - y * R.rpx;
        if (x >= R.rx0 - R.bx0 && x <= R.own_x1 - R.bx0 && y >= R.ry0 - R.by0 && y <= R.own_y1 - R.by0) {
            long long g = (long long)(R.by0 + y) * D->nx + (R.bx0 + x);
            noise[g] = (float)((double)noise[g] * diffrat);
        }
    }
}

int zb_launch_region_stats(zb_ctx *ctx, float *noise, int32_t *out_mask, cudaStream_t st)
{
    const int nR = ctx->sz.n_regions;
    // The keywords hotpants writes (DMEAN, DSIGE, DSIG and their OK-pixel twins, X2NRM, NX2NRM) come from the
    // sigma-clip phase of getStampStats3 and from getNoiseStats3.  Its histogram phase (functions.c:879-1031: mode,
    // median, FWHM) only feeds fprintf(stderr) lines of main.c:1479-1554 and has no side effect on any product, so
    // it — and the 100-value sample that sizes its bins — is left out unless ZB_QA_HIST=1 asks for the whole routine.
    static int minb = 0, use_div = 0, rot = -1;
    int hist, notail;               // read per call: the tests switch them (ZB_QA_NOTAIL=1: full passes only)
    { const char *e = getenv("ZB_QA_HIST"); hist = (e && e[0] == '1') ? 1 : 0; }
    { const char *e = getenv("ZB_QA_NOTAIL"); notail = (e && e[0] == '1') ? 1 : 0; }
    if (rot < 0) { const char *e = getenv("ZB_QA_ROT"); rot = e ? atoi(e) : 1; }
    { const char *e = getenv("ZB_QA_DIV"); use_div = (e && e[0] == '1') ? 1 : 0; }
    // 4 CTAs/SM: 9 regions x 64 CTAs = one resident wave (A/B on B200: 1.47 ms vs 1.65 at 3)
    if (!minb) { const char *e = getenv("ZB_QA_MINB"); minb = e ? atoi(e) : 4; }
    k_qa_sample<<<dim3(4, nR), 256, 0, st>>>(ctx->d_dev, ctx->d_rdiff, ctx->d_rnoise, ctx->d_omask, ctx->d_qa, ctx->d_ran, ctx->n_ran, hist); ZB_DBG("k_qa_sample");
    ctx->launches++;
    const int nblk = zb_qa_nblk(nR);
#define QA_ARGS(pass) ctx->d_dev, ctx->d_rdiff, ctx->d_rnoise, ctx->d_omask, out_mask, ctx->d_qa, ctx->d_qpart, ctx->d_stats, (pass) == 0, use_div, \
                      (pass), hist, ctx->d_qshell, ctx->d_qshell_n, ctx->qshell_cap
    // sigma clip: at most maxiter (5) iterations.  Pass 0 reads the frame: sums of everything, and every value outside the
    // sample's core onto the shell lists; k_qa_tail then runs the clip iterations on the lists.  If it has to hand a
    // statistic back (the accepted interval no longer contains the core, a list overflowed, NaNs in the region), pass 1
    // collects again around mean0 +- 0.75 statSig sd0 and the tail retries; passes 2-4 are plain passes that fall
    // through at once when nothing is left.  With the histogram: 5 + up to 5 plain passes, no shell lists.
    const int npass = hist ? 10 : 5;
    for (int pass = 0; pass < npass; pass++) {
        if (pass <= 1 && !hist) {
            k_qa_pass<3, 1, true><<<dim3(nblk, nR), 256, 0, st>>>(QA_ARGS(pass)); ZB_DBG("k_qa_pass shell");
            ctx->launches++;
            if (notail) continue;
            k_qa_tail<<<dim3(4, nR), QA_TAIL_NT, 0, st>>>(ctx->d_dev, ctx->d_qa, ctx->d_stats, ctx->d_qshell, ctx->d_qshell_n, ctx->qshell_cap, nblk, pass); ZB_DBG("k_qa_tail");
            ctx->launches++;
            continue;
        }
        if (!rot) k_qa_pass<4, 0, false><<<dim3(nblk, nR), 256, 0, st>>>(QA_ARGS(pass));       // ZB_QA_ROT=0: the generic form unrolled four times (A/B)
        else if (minb == 3) k_qa_pass<3, 1, false><<<dim3(nblk, nR), 256, 0, st>>>(QA_ARGS(pass));
        else if (minb == 4) k_qa_pass<4, 1, false><<<dim3(nblk, nR), 256, 0, st>>>(QA_ARGS(pass));
        else k_qa_pass<2, 1, false><<<dim3(nblk, nR), 256, 0, st>>>(QA_ARGS(pass));
        ZB_DBG("k_qa_pass");
        ctx->launches++;
    }
#undef QA_ARGS
    k_qa_finish<<<(nR + 31) / 32, 32, 0, st>>>(ctx->d_dev, ctx->d_qa, ctx->d_stats); ZB_DBG("k_qa_finish");
    ctx->launches++;
    if (ctx->p.rescale_ok) {
        k_rescale_ok<<<dim3(64, nR), 256, 0, st>>>(ctx->d_dev, ctx->d_omask, ctx->d_stats, noise); ZB_DBG("k_rescale_ok");
        ctx->launches++;
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return zb_cuda_fail(ctx, e, "region stats kernels", __FILE__, __LINE__);
    return ZB_OK;
}
