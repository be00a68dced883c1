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

#define NT 256

struct RedBuf { double a[NT / 32]; double b[NT / 32]; int c[NT / 32]; double ra, rb; int rc; };

// block-wide sums with a fixed reduction tree (deterministic)
__device__ static void block_sum(RedBuf &rb, double &a, double &b, int &c) {
    for (int o = 16; o > 0; o >>= 1) {
        a += __shfl_xor_sync(0xffffffffu, a, o);
        b += __shfl_xor_sync(0xffffffffu, b, o);
        c += __shfl_xor_sync(0xffffffffu, c, o);
    }
    int w = threadIdx.x >> 5;
    __syncthreads();
    if ((threadIdx.x & 31) == 0) { rb.a[w] = a; rb.b[w] = b; rb.c[w] = c; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double sa = 0, sb = 0; int sc = 0;
        for (int k = 0; k < NT / 32; k++) { sa += rb.a[k]; sb += rb.b[k]; sc += rb.c[k]; }
        rb.ra = sa; rb.rb = sb; rb.rc = sc;
    }
    __syncthreads();
    a = rb.ra; b = rb.rb; c = rb.rc;
}

__device__ static inline bool excluded(int mdat, double rdat, int umask, int smask) {
    return ((umask > 0) && (!(mdat & umask))) || ((smask > 0) && (mdat & smask)) || (fabs(rdat) <= ZB_ZEROVAL);
}

struct Shared {
    RedBuf rb;
    double work[100];
    int bins[256];
    double mean, sd, binsize, bin1, mode, median, fwhm, lfwhm, sum;
    int status, npts_s, cont, imax;
    // search state
    int cand, pcnt, brk;
    int xloc[2 * ZB_MAX_KSS], yloc[2 * ZB_MAX_KSS];
    double peaks[2 * ZB_MAX_KSS];
    unsigned long long best;
    int anybad;
};

// getStampStats3 on the shared-memory stamp.  Returns the reference's status code; fields of
// `S` that the reference would leave untouched on an early return keep their incoming values.
__device__ static int stats3_block(Shared &S, const float *pix, int *msk, unsigned char *state,
                                   int nPixX, int nPixY, int umask, int smask, int maxiter, float statSig,
                                   const double *__restrict__ ran, int n_ran,
                                   int32_t *gmask, int gx0, int gy0, int gpitch)
{
    const int tid = threadIdx.x;
    const int npix = nPixX * nPixY;
    if (npix < 100) return 4;
    if (tid == 0) {
        int i, goodcnt = 0, r = 0;
        for (i = 0; i < 100; i++) S.work[i] = 0.0;
        for (i = 0; (i < 100) && (goodcnt < npix) && (r + 1 < n_ran); i++, goodcnt++) {
            int xr = (int)floor(ran[r++] * nPixX);
            int yr = (int)floor(ran[r++] * nPixY);
            double rdat = pix[xr + yr * nPixX];
            int mdat = msk[xr + yr * nPixX];
            if (excluded(mdat, rdat, umask, smask)) i--;
            else S.work[i] = rdat;
        }
        int n = i;
        for (int a = 1; a < n; a++) {            // ascending sort of the n samples
            double v = S.work[a]; int b = a - 1;
            while (b >= 0 && S.work[b] > v) { S.work[b + 1] = S.work[b]; b--; }
            S.work[b + 1] = v;
        }
        S.npts_s = n;
        S.binsize = (S.work[(int)(0.9f * n)] - S.work[(int)(0.5f * n)]) / (float)100;
        S.bin1 = S.work[(int)(0.5f * n)] - 128. * S.binsize;
    }
    // valid pixels (NaN/Inf pixels get 0x88, functions.c:857-861)
    int cnt_local = 0;
    for (int k = tid; k < npix; k += NT) {
        double rdat = pix[k]; int mdat = msk[k];
        unsigned char st = 1;
        if (!excluded(mdat, rdat, umask, smask)) {
            if (rdat * 0.0 != 0.0) {
                msk[k] |= (ZB_FLAG_INPUT_ISBAD | ZB_FLAG_ISNAN);
                int y = k / nPixX, x = k - y * nPixX;
                atomicOr(&gmask[(long long)(gy0 + y) * gpitch + gx0 + x], ZB_FLAG_INPUT_ISBAD | ZB_FLAG_ISNAN);
            } else { st = 0; cnt_local++; }
        }
        state[k] = st;
    }
    double da = 0, db = 0;
    block_sum(S.rb, da, db, cnt_local);
    const int count = cnt_local;
    // sigma_clip
    if (count == 0) { if (tid == 0) { S.mean = 0; S.sd = ZB_MAXVAL; } __syncthreads(); return 5; }
    int cnt = 0, ncnt = count, iter = 0;
    while (ncnt != cnt && iter < maxiter) {
        cnt = ncnt;
        double s1 = 0, s2 = 0; int dummy = 0;
        for (int k = tid; k < npix; k += NT)
            if (!state[k]) { float d = pix[k]; s1 += d; s2 += (double)(d * d); }
        block_sum(S.rb, s1, s2, dummy);
        double mean, sd;
        if (ncnt > 0) mean = s1 / ncnt;
        else { if (tid == 0) { S.mean = 0; S.sd = ZB_MAXVAL; } __syncthreads(); return 5; }
        if (ncnt > 1) { sd = s2 - ncnt * mean * mean; sd = sqrt(sd / (double)(ncnt - 1)); }
        else { if (tid == 0) { S.mean = mean; S.sd = ZB_MAXVAL; } __syncthreads(); return 5; }
        if (tid == 0) { S.mean = mean; S.sd = sd; }
        double istd = 1. / sd;
        int keep = 0;
        for (int k = tid; k < npix; k += NT)
            if (!state[k]) {
                if ((fabs(pix[k] - mean) * istd) > statSig) state[k] = 2;
                else keep++;
            }
        double z0 = 0, z1 = 0;
        block_sum(S.rb, z0, z1, keep);
        ncnt = keep;
        iter++;
    }
    __syncthreads();
    const double mean = S.mean, isd = 1. / S.sd;
    // histogram loop
    int tries = 0;
    if (tid == 0) S.imax = 0;
    for (;;) {
        if (tries >= 5) return 1;
        for (int k = tid; k < 256; k += NT) S.bins[k] = 0;
        __syncthreads();
        const double bin1 = S.bin1, binsize = S.binsize;
        double ssum = 0, z = 0; int good = 0;
        for (int k = tid; k < npix; k += NT) {
            double rdat = pix[k]; int mdat = msk[k];
            if (excluded(mdat, rdat, umask, smask)) continue;
            if (rdat * 0.0 != 0.0) continue;
            if ((fabs(rdat - mean) * isd) > statSig) continue;
            int index = (int)(floor((rdat - bin1) / binsize) + 1);
            index = (index < 0 ? 0 : index);
            index = (index > 255 ? 255 : index);
            atomicAdd(&S.bins[index], 1);
            ssum += fabs(rdat);
            good++;
        }
        block_sum(S.rb, ssum, z, good);
        if (good == 0) { if (tid == 0) { S.mode = S.median = S.work[(int)(0.5f * S.npts_s)]; S.fwhm = S.lfwhm = 0.; } __syncthreads(); return 2; }
        if (binsize == 0.) { if (tid == 0) { S.mode = S.median = S.work[(int)(0.5f * S.npts_s)]; S.fwhm = S.lfwhm = 0.; } __syncthreads(); return 3; }
        if (tid == 0) {
            HistOut o; int imax = S.imax;
            hist_analyse(S.bins, good, bin1, binsize, imax, o);
            S.imax = imax;
            S.mode = o.mode;
            S.cont = o.action;
            if (o.action == 1) { S.bin1 = bin1 - 128. * binsize; S.binsize = binsize * 2.; }
            else if (o.action == 2) { S.binsize = binsize / 3.; S.bin1 = o.mode - 128. * S.binsize; }
            else { S.sum = ssum; S.fwhm = o.fwhm; S.lfwhm = o.lfwhm; S.median = o.median; }
        }
        __syncthreads();
        if (S.cont == 0) break;
        tries++;
    }
    return 0;
}

// functions.c:1676-1714: Hoare-partition index sort (explicit stack instead of recursion)
__device__ static void qsort_idx(const double *v, int *idx, int n) {
    int stack[64]; int sp = 0;
    stack[sp++] = 0; stack[sp++] = n - 1;
    while (sp > 0) {
        int hi = stack[--sp], lo = stack[--sp];
        double piv = v[idx[(lo + hi) / 2]];
        int i = lo - 1, j = hi + 1, t;
        for (;;) {
            while (v[idx[++i]] < piv) ;
            while (v[idx[--j]] > piv) ;
            if (i < j) { t = idx[j]; idx[j] = idx[i]; idx[i] = t; }
            else if (i == j) { ++i; break; }
            else break;
        }
        // the reference recurses (lo..j) first, then (i..hi); the two ranges are disjoint, so order is irrelevant
        if (i < hi) { stack[sp++] = i; stack[sp++] = hi; }
        if (lo < j) { stack[sp++] = lo; stack[sp++] = j; }
    }
}

__global__ void __launch_bounds__(NT)
k_stamp_search(const ZbDev *__restrict__ D, const float *__restrict__ img, int32_t *__restrict__ m_in,
               ZbCell *__restrict__ cells, int first_cell, const double *__restrict__ ran, int n_ran)
{
    extern __shared__ __align__(16) unsigned char dyn[];
    __shared__ Shared S;
    ZbCell &C = cells[first_cell + blockIdx.x];
    const ZbRegion R = D->reg[C.region];
    const int tid = threadIdx.x;
    const int xLen = C.spx, yLen = C.spy, npix = xLen * yLen;
    float *pix = (float *)dyn;
    int *msk = (int *)(pix + npix);
    unsigned char *state = (unsigned char *)(msk + npix);
    int32_t *gm = m_in + R.off;

    for (int k = tid; k < npix; k += NT) {
        int y = k / xLen, x = k - y * xLen;
        pix[k] = img[(long long)(R.by0 + C.sy0 + y) * D->nx + (R.bx0 + C.sx0 + x)];
        msk[k] = gm[(long long)(C.sy0 + y) * R.rpx + (C.sx0 + x)];
    }
    if (tid == 0) {
        S.mean = C.mean; S.sd = C.sd; S.mode = C.in_mode; S.fwhm = C.in_fwhm;
        S.median = C.median; S.lfwhm = C.lfwhm; S.sum = C.sum;
    }
    __syncthreads();
    int status = stats3_block(S, pix, msk, state, xLen, yLen, 0x0, 0xffff, 3, D->p.stat_sig, ran, n_ran,
                              gm, C.sx0, C.sy0, R.rpx);
    __syncthreads();

    // ---- getPsfCenters ----
    const int hwss = D->p.hw_ksstamp, nkss = D->p.n_ksstamps, fcnt = 2 * nkss;
    const float kft = C.kft;
    const double hiThresh = D->conv_template ? (double)D->tUK : (double)D->iUK;
    const int bbit2 = D->conv_template ? ZB_FLAG_T_SKIP : ZB_FLAG_I_SKIP;
    const int bbit = (D->conv_template ? (ZB_FLAG_T_BAD | ZB_FLAG_T_SKIP) : (ZB_FLAG_I_BAD | ZB_FLAG_I_SKIP)) | 0xbf;
    const double sky = S.mode, invdsky = 1. / S.fwhm;
    const float skyf = (float)sky, invdskyf = (float)invdsky;     // checkPsfCenter takes them as floats
    const double floor_ = sky + kft * S.fwhm;
    float dfrac = 0.9f;
    if (tid == 0) { S.pcnt = 0; S.brk = 0; }
    __syncthreads();
    for (;;) {
        if (S.pcnt >= fcnt) break;
        double loPsf = sky + (hiThresh - sky) * dfrac;
        loPsf = (loPsf > floor_) ? loPsf : floor_;
        int start = 0;
        for (;;) {
            // first raster index >= start that is a candidate
            if (tid == 0) S.cand = npix;
            __syncthreads();
            int mine = npix;
            for (int k = start + tid; k < npix; k += NT) {
                if (msk[k] & bbit) continue;
                double dpt = pix[k];
                if (dpt >= hiThresh) continue;
                if (((dpt - sky) * invdsky) < kft) continue;
                if (dpt > loPsf) { mine = k; break; }
            }
            if (mine < npix) atomicMin(&S.cand, mine);
            __syncthreads();
            const int cand = S.cand;
            if (cand >= npix) break;
            const int j = cand / xLen, i = cand - j * xLen;
            // centre on the peak flux in the +-hwss box (first pixel in raster order that holds the maximum)
            if (tid == 0) S.best = 0ull;
            __syncthreads();
            {
                const int bw = 2 * hwss + 1;
                unsigned long long best = 0ull;
                for (int q = tid; q < bw * bw; q += NT) {
                    int l = j - hwss + q / bw, k = i - hwss + q % bw;
                    if (l < 0 || l >= yLen || k < 0 || k >= xLen) continue;
                    int n2 = k + l * xLen;
                    if (msk[n2] & bbit) continue;
                    double dpt2 = pix[n2];
                    if (dpt2 >= hiThresh) continue;
                    if (((dpt2 - sky) * invdsky) < kft) continue;
                    if (dpt2 > (double)pix[cand]) {
                        // order-preserving key: larger value first, then smaller raster index
                        unsigned int fb = __float_as_uint(pix[n2]);
                        fb = (fb & 0x80000000u) ? ~fb : (fb | 0x80000000u);
                        unsigned long long key = ((unsigned long long)fb << 32) | (unsigned int)(0x7fffffff - n2);
                        if (key > best) best = key;
                    }
                }
                if (best) atomicMax(&S.best, best);
            }
            __syncthreads();
            int imax = i, jmax = j;
            if (S.best) { int n2 = 0x7fffffff - (int)(S.best & 0xffffffffu); jmax = n2 / xLen; imax = n2 - jmax * xLen; }
            // checkPsfCenter: any masked / saturated pixel in the box (inside the stamp) kills the candidate
            if (tid == 0) S.anybad = 0;
            __syncthreads();
            double dsum = 0, z = 0; int bad = 0;
            {
                const int bw = 2 * hwss + 1;
                for (int q = tid; q < bw * bw; q += NT) {
                    int l = jmax - hwss + q / bw, k = imax - hwss + q % bw;
                    if (l < 0 || l >= yLen || k < 0 || k >= xLen) continue;
                    int n2 = k + l * xLen;
                    double dpt2 = pix[n2];
                    if ((msk[n2] & bbit) || dpt2 >= hiThresh) { bad = 1; continue; }
                    if (((dpt2 - skyf) * invdskyf) > kft) dsum += dpt2;
                }
            }
            block_sum(S.rb, dsum, z, bad);
            if (!bad && dsum != 0.) {
                if (tid == 0) { S.xloc[S.pcnt] = imax; S.yloc[S.pcnt] = jmax; S.peaks[S.pcnt] = dsum; S.pcnt++; }
                const int bw = 2 * hwss + 1;
                for (int q = tid; q < bw * bw; q += NT) {
                    int l = jmax - hwss + q / bw, k = imax - hwss + q % bw;
                    if ((k > 0) && (k < xLen) && (l > 0) && (l < yLen)) msk[k + l * xLen] |= bbit2;
                }
            }
            __syncthreads();
            if (S.pcnt >= fcnt) break;
            start = cand + 1;
        }
        if (S.pcnt >= fcnt) break;
        if (loPsf == floor_) break;
        dfrac -= 0.2;
    }
    __syncthreads();
    if (tid == 0) {
        C.stat_status = status;
        C.mean = S.mean; C.sd = S.sd; C.mode = S.mode; C.fwhm = S.fwhm; C.median = S.median; C.lfwhm = S.lfwhm; C.sum = S.sum;
        int pcnt = S.pcnt, nss = 0;
        if (pcnt > 0) {
            int qs[2 * ZB_MAX_KSS];
            for (int q = 0; q < pcnt; q++) qs[q] = q;
            qsort_idx(S.peaks, qs, pcnt);
            for (int q = 0; q < pcnt && nss < nkss; q++, nss++) {
                C.xss[nss] = S.xloc[qs[pcnt - q - 1]] + C.sx0;
                C.yss[nss] = S.yloc[qs[pcnt - q - 1]] + C.sy0;
            }
        }
        C.nss = nss;
    }
}

int zb_launch_stamp_search(zb_ctx *ctx, const float *conv_img, int first_cell, int n_cells, cudaStream_t st)
{
    if (n_cells <= 0) return ZB_OK;
    size_t npx = (size_t)ctx->sz.fw_stamp * ctx->sz.fw_stamp;
    size_t smem = npx * (sizeof(float) + sizeof(int) + 1) + 16;
    if (smem > 200 * 1024) {
        snprintf(ctx->err, sizeof ctx->err, "stamp of %d x %d pixels does not fit shared memory", ctx->sz.fw_stamp, ctx->sz.fw_stamp);
        return ZB_ERR_UNSUPPORTED;
    }
    cudaFuncSetAttribute(k_stamp_search, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    k_stamp_search<<<n_cells, NT, smem, st>>>(ctx->d_dev, conv_img, ctx->d_min, ctx->d_cells, first_cell, ctx->d_ran, ctx->n_ran);
    ctx->launches++;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return zb_cuda_fail(ctx, e, "k_stamp_search", __FILE__, __LINE__);
    return ZB_OK;
}
