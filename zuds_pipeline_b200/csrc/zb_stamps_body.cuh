// zb_stamps_body.cuh — body of the stamp-search kernel, included by zb_stamps.cu once per CTA size
// (NT threads, NW = NT/32 candidates evaluated concurrently).  Not a stand-alone header.
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
    double work[100], work2[100];
    int bins[256];
    double mean, sd, binsize, bin1, mode, median, fwhm, lfwhm, sum;
    int status, npts_s, cont, imax;
    // search state
    int cand, pcnt, brk;
    int xloc[2 * ZB_MAX_KSS], yloc[2 * ZB_MAX_KSS];
    double peaks[2 * ZB_MAX_KSS];
    int cands[NW], ncand;
    int c_ok[NW], c_x[NW], c_y[NW], c_new[NW];
    int failed[ZB_NFAIL], nfail;
    double c_sum[NW];
};

// getStampStats3 on the shared-memory stamp.  Returns the reference's status code; fields of
// `S` that the reference would leave untouched on an early return keep their incoming values.
__device__ static int stats3_block(Shared &S, const float *pix, unsigned short *msk, unsigned char *state,
                                   int nPixX, int nPixY, int umask, int smask, int maxiter, float statSig,
                                   const double *__restrict__ ran, int n_ran,
                                   int32_t *gmask, int gx0, int gy0, int gpitch)
{
    const int tid = threadIdx.x;
    const int npix = nPixX * nPixY;
    if (npix < 100) return 4;
    // the first 100 accepted random samples, in attempt order (functions.c:811-833): attempt a uses
    // ran[2a], ran[2a+1]; acceptance does not depend on earlier attempts, so they are tested NT at a time
    if (tid < 100) { S.work[tid] = 0.0; S.work2[tid] = 0.0; }
    if (tid == 0) S.npts_s = 0;
    __syncthreads();
    for (int base = 0; base < npix; base += NT) {
        const int a = base + tid;
        int flag = 0; double rdat = 0;
        if (a < npix && 2 * a + 1 < n_ran) {
            int xr = (int)floor(ran[2 * a] * nPixX), yr = (int)floor(ran[2 * a + 1] * nPixY);
            rdat = pix[xr + yr * nPixX];
            flag = !excluded(msk[xr + yr * nPixX], rdat, umask, smask);
        }
        const unsigned bal = __ballot_sync(0xffffffffu, flag);
        if ((tid & 31) == 0) S.rb.c[tid >> 5] = __popc(bal);
        __syncthreads();
        int before = S.npts_s;
        for (int w = 0; w < (tid >> 5); w++) before += S.rb.c[w];
        const int pos = before + __popc(bal & ((1u << (tid & 31)) - 1u));
        if (flag && pos < 100) S.work2[pos] = rdat;
        __syncthreads();
        if (tid == 0) { int t = 0; for (int w = 0; w < NT / 32; w++) t += S.rb.c[w]; S.npts_s += t; }
        __syncthreads();
        if (S.npts_s >= 100) break;
    }
    {
        const int n = S.npts_s < 100 ? S.npts_s : 100;
        if (tid < n) {                               // rank sort (ascending; equal values keep any order)
            const double v = S.work2[tid]; int rank = 0;
            for (int q = 0; q < n; q++) { double u = S.work2[q]; rank += (u < v) || (u == v && q < tid); }
            S.work[rank] = v;
        }
        __syncthreads();
        if (tid == 0) {
            S.npts_s = n;
            S.binsize = (S.work[(int)(0.9f * n)] - S.work[(int)(0.5f * n)]) / (float)100;
            S.bin1 = S.work[(int)(0.5f * n)] - 128. * S.binsize;
        }
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
               ZbCell *__restrict__ cells, int first_cell, const double *__restrict__ ran, int n_ran,
               unsigned char *__restrict__ gscratch, size_t gscratch_per_cell, int dir_t)
{
    extern __shared__ __align__(16) unsigned char dyn_smem[];
    // stamps too large for shared memory (coadd-scale frames with a coarse stamp grid) work out of a
    // per-CTA slice of global scratch instead; same code, L1/L2-cached
    unsigned char *dyn = gscratch ? gscratch + (size_t)blockIdx.x * gscratch_per_cell : dyn_smem;
    __shared__ Shared S;
    ZbCell &C = cells[first_cell + blockIdx.x];
    const ZbRegion R = D->reg[C.region];
    const int tid = threadIdx.x;
    const int xLen = C.spx, yLen = C.spy, npix = xLen * yLen;
    float *pix = (float *)dyn;
    // the mask copy keeps the low 16 bits: every test here (smask 0xffff, the T/I bits, 0xbf) and every bit set here
    // lies below bit 16, and a user mask value with only higher bits set still carries INPUT_MASK (main.c:778-803).
    // 7 instead of 9 bytes per stamp pixel puts four CTAs of 81 x 81 stamps on an SM instead of three
    unsigned short *msk = (unsigned short *)(pix + npix);
    unsigned char *state = (unsigned char *)(msk + ((npix + 1) & ~1));
    unsigned *cbm = (unsigned *)(state + ((npix + 15) & ~15));      // candidate bitmap of the current pass
    int32_t *gm = m_in + R.off;

    for (int k = tid; k < npix; k += NT) {
        int y = k / xLen, x = k - y * xLen;
        pix[k] = img[(long long)(R.by0 + C.sy0 + y) * D->nx + (R.bx0 + C.sx0 + x)];
        msk[k] = (unsigned short)gm[(long long)(C.sy0 + y) * R.rpx + (C.sx0 + x)];
    }
    if (tid == 0) {
        S.mean = C.mean; S.sd = C.sd; S.mode = C.in_mode; S.fwhm = C.in_fwhm;
        S.median = C.median; S.lfwhm = C.lfwhm; S.sum = C.sum;
    }
    __syncthreads();
    const long long t_begin = clock64();
    int status = stats3_block(S, pix, msk, state, xLen, yLen, 0x0, 0xffff, 3, D->p.stat_sig, ran, n_ran,
                              gm, C.sx0, C.sy0, R.rpx);
    __syncthreads();

    // ---- getPsfCenters ----
    const int hwss = D->p.hw_ksstamp, nkss = D->p.n_ksstamps, fcnt = 2 * nkss;
    const float kft = C.kft;
    // dir_t: the image searched is the template (1) or the science image (0); with -c b both are searched, each
    // with its own pair of mask bits (functions.c:195,251-330), so the two searches do not see each other
    const double hiThresh = dir_t ? (double)D->tUK : (double)D->iUK;
    const int bbit2 = dir_t ? ZB_FLAG_T_SKIP : ZB_FLAG_I_SKIP;
    const int bbit = (dir_t ? (ZB_FLAG_T_BAD | ZB_FLAG_T_SKIP) : (ZB_FLAG_I_BAD | ZB_FLAG_I_SKIP)) | 0xbf;
    const double sky = S.mode, invdsky = 1. / S.fwhm;
    const float skyf = (float)sky, invdskyf = (float)invdsky;     // checkPsfCenter takes them as floats
    const double floor_ = sky + kft * S.fwhm;
    float dfrac = 0.9f;
    // Raster-greedy search, NW candidates at a time.  A candidate that fails changes nothing, so the
    // next NW candidates (in raster order, against the current mask) are evaluated concurrently, one
    // warp each; the first one that succeeds claims its box, the ones after it are re-evaluated.
    const int bw = 2 * hwss + 1, nbox = bw * bw;
    const int nwords = (npix + 31) >> 5;
    const int warp = tid >> 5, lane = tid & 31;
    // per-pixel tests that do not depend on the mask, evaluated once with the reference's expressions
    // (functions.c:574-581 and checkPsfCenter :476-490): bit0 usable for centring (below hiThresh and
    // >= kerFitThresh sigma in double), bit1 counted by checkPsfCenter (float sky / 1/fwhm), bit2 >= hiThresh
    unsigned char *fl = state;
    for (int k = tid; k < npix; k += NT) {
        const double dpt = pix[k];
        unsigned char f = 0;
        const bool hi = dpt >= hiThresh;
        if (!hi && !(((dpt - sky) * invdsky) < kft)) f |= 1;
        if (((dpt - skyf) * invdskyf) > kft) f |= 2;
        if (hi) f |= 4;
        fl[k] = f;
    }
    if (tid == 0) { S.pcnt = 0; S.nfail = 0; }
    const long long t_search0 = clock64();
    __syncthreads();
    for (;;) {
        if (S.pcnt >= fcnt) break;
        double loPsf = sky + (hiThresh - sky) * dfrac;
        loPsf = (loPsf > floor_) ? loPsf : floor_;
        int start = 0;
        bool rebuild = true;
        for (;;) {
            if (rebuild) {
                for (int base = 0; base < npix; base += NT) {           // one ballot = one 32-pixel word
                    const int k = base + tid;
                    int c = 0;
                    if (k < npix && !(msk[k] & bbit)) {
                        const double dpt = pix[k];
                        c = (dpt < hiThresh) && !(((dpt - sky) * invdsky) < kft) && (dpt > loPsf);
                    }
                    const unsigned bits = __ballot_sync(0xffffffffu, c);
                    if (lane == 0 && (k >> 5) < nwords) cbm[k >> 5] = bits;
                }
                rebuild = false;
                __syncthreads();
            }
            // warp 0 lists the next NW candidates at or after `start`
            if (warp == 0) {
                int found = 0;
                for (int w0 = start >> 5; w0 < nwords && found < NW; w0 += 32) {
                    const int w = w0 + lane;
                    unsigned bits = w < nwords ? cbm[w] : 0u;
                    if (w == (start >> 5)) bits &= ~0u << (start & 31);
                    int c = __popc(bits), inc = c;
                    for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
                    int ord = found + inc - c;                       // ordinal of this word's first candidate
                    while (bits && ord < NW) { int bpos = __ffs(bits) - 1; S.cands[ord++] = (w << 5) + bpos; bits &= bits - 1; }
                    found += __shfl_sync(0xffffffffu, inc, 31);
                }
                if (lane == 0) S.ncand = found < NW ? found : NW;
            }
            __syncthreads();
            const int ncand = S.ncand;
            if (ncand == 0) break;
            if (warp < ncand) {
                const int cand = S.cands[warp];
                const int j = cand / xLen, i = cand - j * xLen;
                const float cval = pix[cand];
                const int l0 = max(j - hwss, 0), l1 = min(j + hwss, yLen - 1);
                const int k0 = max(i - hwss, 0), k1 = min(i + hwss, xLen - 1);
                // centre on the peak flux in the +-hwss box: first pixel in raster order holding the maximum
                unsigned long long best = 0ull;
                for (int l = l0; l <= l1; l++)
                    for (int k = k0 + lane; k <= k1; k += 32) {
                        const int n2 = k + l * xLen;
                        if (!(fl[n2] & 1) || (msk[n2] & bbit)) continue;
                        const float v2 = pix[n2];
                        if (v2 > cval) {
                            unsigned int fb = __float_as_uint(v2);
                            fb = (fb & 0x80000000u) ? ~fb : (fb | 0x80000000u);
                            unsigned long long key = ((unsigned long long)fb << 32) | (unsigned int)(0x7fffffff - n2);
                            if (key > best) best = key;
                        }
                    }
                for (int o = 16; o > 0; o >>= 1) { unsigned long long t = __shfl_xor_sync(0xffffffffu, best, o); if (t > best) best = t; }
                int imax = i, jmax = j;
                if (best) { int n2 = 0x7fffffff - (int)(best & 0xffffffffu); jmax = n2 / xLen; imax = n2 - jmax * xLen; }
                // a centre that failed once fails again: pixel values are fixed and mask bits are only ever added
                const int cpos = imax + jmax * xLen;
                int known = 0;
                for (int q = lane; q < S.nfail; q += 32) known |= (S.failed[q] == cpos);
                known = __any_sync(0xffffffffu, known);
                double dsum = 0; int bad = known;
                if (!known) {
                    // checkPsfCenter: any masked / saturated pixel of the box (inside the stamp) kills the candidate
                    const int m0 = max(jmax - hwss, 0), m1 = min(jmax + hwss, yLen - 1);
                    const int q0 = max(imax - hwss, 0), q1 = min(imax + hwss, xLen - 1);
                    for (int l = m0; l <= m1 && !bad; l++)
                        for (int k = q0 + lane; k <= q1; k += 32) {
                            const int n2 = k + l * xLen;
                            const unsigned char f = fl[n2];
                            if ((msk[n2] & bbit) || (f & 4)) { bad = 1; break; }
                            if (f & 2) dsum += (double)pix[n2];
                        }
                    bad = __any_sync(0xffffffffu, bad);
                    for (int o = 16; o > 0; o >>= 1) dsum += __shfl_xor_sync(0xffffffffu, dsum, o);
                }
                if (lane == 0) { S.c_ok[warp] = (!bad && dsum != 0.); S.c_x[warp] = imax; S.c_y[warp] = jmax; S.c_sum[warp] = dsum;
                                 S.c_new[warp] = !known; }
            }
            __syncthreads();
            int hit = -1;
            for (int w = 0; w < ncand; w++) if (S.c_ok[w]) { hit = w; break; }
            if (tid == 0) {                                  // remember the centres that failed before the hit
                const int lim = hit < 0 ? ncand : hit;
                for (int w = 0; w < lim; w++)
                    if (S.c_new[w] && S.nfail < ZB_NFAIL) {
                        const int cpos = S.c_x[w] + S.c_y[w] * xLen; bool dup = false;
                        for (int q = 0; q < S.nfail; q++) dup |= (S.failed[q] == cpos);
                        if (!dup) S.failed[S.nfail++] = cpos;
                    }
            }
            if (hit < 0) { start = S.cands[ncand - 1] + 1; __syncthreads(); continue; }
            {
                const int imax = S.c_x[hit], jmax = S.c_y[hit];
                for (int q = tid; q < nbox; q += NT) {
                    const int l = jmax - hwss + q / bw, k = imax - hwss + q % bw;
                    if ((k > 0) && (k < xLen) && (l > 0) && (l < yLen)) msk[k + l * xLen] |= bbit2;
                }
                start = S.cands[hit] + 1;
                __syncthreads();
                if (tid == 0) { S.xloc[S.pcnt] = imax; S.yloc[S.pcnt] = jmax; S.peaks[S.pcnt] = S.c_sum[hit]; S.pcnt++; }
                rebuild = true;
                __syncthreads();
                if (S.pcnt >= fcnt) break;
            }
        }
        if (S.pcnt >= fcnt) break;
        if (loPsf == floor_) break;
        dfrac -= 0.2;
    }
    __syncthreads();
    if (tid == 0) {
        C.stat_status = status;
        C.clk_stats = (int)(t_search0 - t_begin); C.clk_search = (int)(clock64() - t_search0);
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

