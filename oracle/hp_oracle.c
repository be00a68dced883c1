/*
 * hp_oracle.c — CPU restatement of the reference hotpants 5.1.11 path as zuds
 * drives it (TEST INFRASTRUCTURE; see hp_oracle.h).  Plain C99, one region at a
 * time, no globals, in-memory arrays instead of FITS.  Every function cites the
 * reference lines it restates (paths relative to /root/reference/).
 *
 * Float/double mixing follows the reference exactly (it is compiled -ansi, so no
 * FMA contraction: build this file with -ffp-contract=off).
 *
 * Scope (same as the CUDA path): -c t | -c i, -fom v, -imi/-tmi masks, rms maps or
 * the gain/readnoise noise model, any -ng/-r/-rss/-ko/-bgo/-nrx/-nry/-nsx/-nsy.
 * Out of scope and rejected: -c b (check_stamps), -ssf/-cmp, -ki, -pca, -gd, -rf/-rkw.
 */
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <stdio.h>
#include "hp_oracle.h"

#define MAXVAL  1e10      /* hotpants/defaults.h:24 */
#define ZEROVAL 1e-10     /* hotpants/defaults.h:25 */

static int imin_(int a, int b) { return a < b ? a : b; }
static int imax_(int a, int b) { return a < b ? b : a; }

/* ------------------------------------------------------------------ params */

/* hotpants/vargs.c:55-128 with hotpants/defaults.h:27-80 */
void hp_oracle_params_default(zb_params *p) {
    static const double sig[3] = {0.7, 1.5, 3.0};
    static const int deg[3] = {6, 4, 2};
    int i;
    memset(p, 0, sizeof *p);
    p->ngauss = 3;
    for (i = 0; i < 3; i++) { p->deg[i] = deg[i]; p->beta[i] = (float)(1.0 / (2.0 * sig[i] * sig[i])); }
    p->t_uthresh = p->t_ukthresh = p->i_uthresh = p->i_ukthresh = 25000.f;
    p->t_lthresh = p->i_lthresh = 0.f;
    p->t_gain = p->i_gain = 1.f;
    p->t_rdnoise = p->i_rdnoise = 0.f;
    p->t_pedestal = p->i_pedestal = 0.f;
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

/* hotpants/main.c:382-428 (sizes), :460-463 (image limits), :535-555,:623-635 (regions) */
int hp_oracle_sizes(const zb_params *p, zb_sizes *s) {
    int i, j, nR = 0;
    memset(s, 0, sizeof *s);
    for (i = 0; i < p->ngauss; i++) s->n_comp_ker += ((p->deg[i] + 1) * (p->deg[i] + 2)) / 2;
    s->n_comp = ((p->ker_order + 1) * (p->ker_order + 2)) / 2;
    s->n_bg_vectors = ((p->bg_order + 1) * (p->bg_order + 2)) / 2;
    s->n_comp_total = s->n_comp_ker * s->n_comp + s->n_bg_vectors;
    s->mat_size = (s->n_comp_ker - 1) * s->n_comp + s->n_bg_vectors + 1;
    s->fw_kernel = 2 * p->hw_kernel + 1;
    s->fw_ksstamp = 2 * p->hw_ksstamp + 1;
    s->nstamp_x = p->nstamp_x; s->nstamp_y = p->nstamp_y;
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
    for (j = 0; j < p->nreg_y; j++)
        for (i = 0; i < p->nreg_x; i++, nR++) {
            int xMax = p->nx - 1, yMax = p->ny - 1;
            int rx0 = i * xMax / p->nreg_x, ry0 = j * yMax / p->nreg_y;
            int rx1 = imin_((i + 1) * xMax / p->nreg_x, xMax), ry1 = imin_((j + 1) * yMax / p->nreg_y, yMax);
            int buf = s->n_regions > 1 ? s->fw_stamp / 2 : p->hw_kernel;
            int bx0 = imax_(0, rx0 - buf), by0 = imax_(0, ry0 - buf);
            int bx1 = imin_(xMax, rx1 + buf), by1 = imin_(yMax, ry1 + buf);
            s->max_rpix_x = imax_(s->max_rpix_x, bx1 - bx0 + 1);
            s->max_rpix_y = imax_(s->max_rpix_y, by1 - by0 + 1);
        }
    return 0;
}

/* ------------------------------------------------------------ ran1 / sorting */

/* Numerical Recipes (1st ed.) ran1 as restated in hotpants/functions.c:1620-1662:
   three linear congruential generators and a 97-entry shuffle table. */
typedef struct { long ix1, ix2, ix3; double r[98]; int iff; } ran1_state;

static double ran1_next(ran1_state *s, int *idum) {
    const long M1 = 259200, IA1 = 7141, IC1 = 54773;
    const long M2 = 134456, IA2 = 8121, IC2 = 28411;
    const long M3 = 243000, IA3 = 4561, IC3 = 51349;
    const double RM1 = 1.0 / 259200, RM2 = 1.0 / 134456;
    double t; int j;
    if (*idum < 0 || s->iff == 0) {
        s->iff = 1;
        s->ix1 = (IC1 - (*idum)) % M1;
        s->ix1 = (IA1 * s->ix1 + IC1) % M1;
        s->ix2 = s->ix1 % M2;
        s->ix1 = (IA1 * s->ix1 + IC1) % M1;
        s->ix3 = s->ix1 % M3;
        for (j = 1; j <= 97; j++) {
            s->ix1 = (IA1 * s->ix1 + IC1) % M1;
            s->ix2 = (IA2 * s->ix2 + IC2) % M2;
            s->r[j] = (s->ix1 + s->ix2 * RM2) * RM1;
        }
        *idum = 1;
    }
    s->ix1 = (IA1 * s->ix1 + IC1) % M1;
    s->ix2 = (IA2 * s->ix2 + IC2) % M2;
    s->ix3 = (IA3 * s->ix3 + IC3) % M3;
    j = 1 + (int)((97 * s->ix3) / M3);
    t = s->r[j];
    s->r[j] = (s->ix1 + s->ix2 * RM2) * RM1;
    return t;
}

static ran1_state g_ran1;   /* the reference keeps this state in function statics */
double hp_oracle_ran1(int *idum) { return ran1_next(&g_ran1, idum); }

static int dcmp(const void *a, const void *b) {   /* functions.c:1717-1724 flcomp */
    double x = *(const double *)a, y = *(const double *)b;
    return x > y ? 1 : (x == y ? 0 : -1);
}

/* functions.c:1676-1714: Hoare-partition index sort, ascending */
static void qsort_idx(const double *v, int *idx, int lo, int hi) {
    double piv = v[idx[(lo + hi) / 2]];
    int i = lo - 1, j = hi + 1, t;
    for (;;) {
        while (v[idx[++i]] < piv) ;
        while (v[idx[--j]] > piv) ;
        if (i < j) { t = idx[j]; idx[j] = idx[i]; idx[i] = t; }
        else if (i == j) { ++i; break; }
        else break;
    }
    if (lo < j) qsort_idx(v, idx, lo, j);
    if (i < hi) qsort_idx(v, idx, i, hi);
}

/* ------------------------------------------------------------- sigma clip */

/* functions.c:1037-1107.  Sums are double over float data; d*d is a FLOAT product. */
int hp_oracle_sigma_clip(const float *data, int count, double *mean, double *stdev, int maxiter, float stat_sig) {
    int cnt = 0, ncnt = count, iter = 0, i;
    char *rej;
    if (count == 0) { *mean = 0; *stdev = MAXVAL; return 1; }
    rej = (char *)calloc((size_t)count, 1);
    while (ncnt != cnt && iter < maxiter) {
        double istd;
        cnt = ncnt;
        *mean = 0; *stdev = 0;
        for (i = 0; i < count; i++)
            if (!rej[i]) { float d = data[i]; *mean += d; *stdev += d * d; }
        if (ncnt > 0) *mean /= ncnt;
        else { *mean = 0; *stdev = MAXVAL; free(rej); return 2; }
        if (ncnt > 1) {
            *stdev = *stdev - ncnt * (*mean) * (*mean);
            *stdev = sqrt(*stdev / (double)(ncnt - 1));
        } else { *stdev = MAXVAL; free(rej); return 3; }
        ncnt = 0;
        istd = 1. / (*stdev);
        for (i = 0; i < count; i++)
            if (!rej[i]) {
                if ((fabs(data[i] - (*mean)) * istd) > stat_sig) rej[i] = 1;
                else ncnt++;
            }
        iter++;
    }
    free(rej);
    return 0;
}

/* ---------------------------------------------------------- region context */

typedef struct {
    int x0, y0, x, y;                 /* origin / centre, region coords */
    int nss, sscnt;
    int xss[ZB_MAX_KSS], yss[ZB_MAX_KSS];
    double *vec;                      /* [(nCompKer+nBG)][fwKSStamp^2] */
    double *kref;                     /* [fwKSStamp^2] */
    double *mat;                      /* [nC][nC], 1-indexed like the reference */
    double *scprod;                   /* [nC] */
    double sum, mean, median, mode, sd, fwhm, lfwhm, chi2;
} stamp_t;

typedef struct {
    const zb_params *p;
    zb_sizes sz;
    int rpx, rpy;                     /* rPixX, rPixY of the buffered region */
    int *m;                           /* mRData */
    float kft;                        /* kerFitThresh, reset per region (main.c:829) */
    double *kvec, *fx, *fy;           /* kernel_vec, filter_x, filter_y */
    double *kernel, *kcoef;           /* make_kernel scratch */
    float *temp;                      /* the reference's global float scratch */
    double **wxy;
} reg_t;

typedef struct { double sum, mean, median, mode, sd, fwhm, lfwhm; } stats_t;

/* functions.c:763-1035 getStampStats3.  `data` is an nPixX x nPixY array whose
   pixel (i,j) sits at region pixel (i+x0Reg, j+y0Reg) of the mask. */
static int stamp_stats3(reg_t *R, const float *data, int x0Reg, int y0Reg, int nPixX, int nPixY,
                        stats_t *st, int umask, int smask, int maxiter) {
    const int nstat = 100; const float ufstat = 0.9f, mfstat = 0.5f;
    const float statSig = R->p->stat_sig;
    double bin1, binsize, maxdens, sumx, sumxx, isd, lower, upper, mode_bin, rdat, ssum;
    int mdat, npts, bins[256], i, j, xr, yr, index, imax = 0, tries = 0, ilower, iupper, repeat, goodcnt, idum;
    float *sdat; double work[100];
    int rpx = R->rpx, *m = R->m;

    npts = nPixX * nPixY;
    if (npts < nstat) return 4;
    sdat = (float *)calloc((size_t)npts, sizeof(float));
    memset(work, 0, sizeof work);
    idum = -666;
    goodcnt = 0;
    for (i = 0; (i < nstat) && (goodcnt < npts); i++, goodcnt++) {
        xr = (int)floor(hp_oracle_ran1(&idum) * nPixX);
        yr = (int)floor(hp_oracle_ran1(&idum) * nPixY);
        rdat = data[xr + yr * nPixX];
        mdat = m[(xr + x0Reg) + (yr + y0Reg) * rpx];
        if (((umask > 0) && (!(mdat & umask))) || ((smask > 0) && (mdat & smask)) || (fabs(rdat) <= ZEROVAL)) i--;
        else work[i] = rdat;
    }
    qsort(work, (size_t)i, sizeof(double), dcmp);
    npts = i;
    binsize = (work[(int)(ufstat * npts)] - work[(int)(mfstat * npts)]) / (float)nstat;
    bin1 = work[(int)(mfstat * npts)] - 128. * binsize;

    goodcnt = 0;
    for (j = 0; j < nPixY; j++)
        for (i = 0; i < nPixX; i++) {
            rdat = data[i + j * nPixX];
            mdat = m[(i + x0Reg) + (j + y0Reg) * rpx];
            if (((umask > 0) && (!(mdat & umask))) || ((smask > 0) && (mdat & smask)) || (fabs(rdat) <= ZEROVAL)) continue;
            if (rdat * 0.0 != 0.0) { m[(i + x0Reg) + (j + y0Reg) * rpx] |= (ZB_FLAG_INPUT_ISBAD | ZB_FLAG_ISNAN); continue; }
            sdat[goodcnt++] = (float)rdat;
        }
    if (hp_oracle_sigma_clip(sdat, goodcnt, &st->mean, &st->sd, maxiter, statSig)) { free(sdat); return 5; }
    free(sdat);
    isd = 1. / st->sd;

    repeat = 1;
    while (repeat) {
        if (tries >= 5) return 1;
        for (i = 0; i < 256; i++) bins[i] = 0;
        ssum = sumx = sumxx = 0.;
        goodcnt = 0;
        for (j = 0; j < nPixY; j++)
            for (i = 0; i < nPixX; i++) {
                rdat = data[i + j * nPixX];
                mdat = m[(i + x0Reg) + (j + y0Reg) * rpx];
                if (((umask > 0) && (!(mdat & umask))) || ((smask > 0) && (mdat & smask)) || (fabs(rdat) <= ZEROVAL)) continue;
                if (rdat * 0.0 != 0.0) { m[(i + x0Reg) + (j + y0Reg) * rpx] |= (ZB_FLAG_INPUT_ISBAD | ZB_FLAG_ISNAN); continue; }
                if ((fabs(rdat - st->mean) * isd) > statSig) continue;
                index = (int)(floor((rdat - bin1) / binsize) + 1);
                index = (index < 0 ? 0 : index);
                index = (index > 255 ? 255 : index);
                bins[index]++;
                ssum += fabs(rdat);
                goodcnt += 1;
            }
        if (goodcnt == 0) { st->mode = st->median = work[(int)(mfstat * npts)]; st->fwhm = st->lfwhm = 0.; return 2; }
        if (binsize == 0.) { st->mode = st->median = work[(int)(mfstat * npts)]; st->fwhm = st->lfwhm = 0.; return 3; }

        /* mode: narrowest window holding ~10% of the points */
        sumx = maxdens = 0.;
        for (ilower = iupper = 1; iupper < 255; sumx -= bins[ilower++]) {
            while ((sumx < goodcnt / 10.) && (iupper < 255)) sumx += bins[iupper++];
            if (sumx / (iupper - ilower) > maxdens) { maxdens = sumx / (iupper - ilower); imax = ilower; }
        }
        if (imax < 0 || imax > 255) imax = 0;
        sumxx = sumx = 0.;
        for (i = imax; (sumx < goodcnt / 10.) && (i < 255); i++) { sumx += bins[i]; sumxx += i * bins[i]; }
        mode_bin = sumxx / sumx + 0.5;
        st->mode = bin1 + binsize * (mode_bin - 1.);

        imax = (int)floor(mode_bin);
        for (i = 0, sumx = 0.; i < imax; sumx += bins[i++]) ;
        sumx += bins[imax] * (mode_bin - imax);
        sumx /= goodcnt;

        lower = goodcnt * 0.25;
        upper = goodcnt * 0.75;
        sumx = 0.;
        for (i = 0; sumx < lower; sumx += bins[i++]) ;
        lower = i - (sumx - lower) / bins[i - 1];
        for (; sumx < upper; sumx += bins[i++]) ;
        upper = i - (sumx - upper) / bins[i - 1];

        if ((lower < 1.) || (upper > 255.)) { bin1 -= 128. * binsize; binsize *= 2.; tries++; repeat = 1; }
        else if ((upper - lower) < 40.) { binsize /= 3.; bin1 = st->mode - 128. * binsize; tries++; repeat = 1; }
        else repeat = 0;
    }
    st->sum = ssum;
    st->fwhm = binsize * (upper - lower) / 1.35;
    for (i = 0, sumx = 0; sumx < goodcnt / 2.; sumx += bins[i++]) ;
    st->median = i - (sumx - goodcnt / 2.) / bins[i - 1];
    st->lfwhm = binsize * (st->median - lower) * 2. / 1.35;
    st->median = bin1 + binsize * (st->median - 1.);
    return 0;
}

/* functions.c:719-761 getNoiseStats3.  The product is formed in FLOAT. */
static void noise_stats3(reg_t *R, const float *data, const float *noise, double *nnorm, int *nncount, int umask, int smask) {
    double nsum = 0; int i, n = 0, mdat; float ddat, ndat;
    for (i = R->rpx * R->rpy; i--;) {
        ddat = data[i]; mdat = R->m[i];
        if (((umask > 0) && (!(mdat & umask))) || ((smask > 0) && (mdat & smask)) || (fabs(ddat) <= ZEROVAL)) continue;
        ndat = (float)(1. / noise[i]);
        n += 1;
        nsum += ddat * ddat * ndat * ndat;
    }
    *nncount = n;
    *nnorm = (n > 1) ? nsum / (float)n : MAXVAL;
}

/* ------------------------------------------------------------ stamp search */

/* functions.c:445-502 checkPsfCenter: zero tolerance for masked / saturated pixels
   inside the sub-stamp box (only the part of the box that lies inside the stamp). */
static double check_psf_center(reg_t *R, const float *img, int imax, int jmax, int xLen, int yLen,
                               int sx0, int sy0, double hiThresh, float sky, float invdsky, int bbit, int bbit1) {
    int hwss = R->p->hw_ksstamp, l, k, brk = 0;
    double dmax2 = 0., dpt2;
    for (l = jmax - hwss; l <= jmax + hwss; l++) {
        if (l < 0 || l >= yLen) continue;
        for (k = imax - hwss; k <= imax + hwss; k++) {
            int nr2;
            if (k < 0 || k >= xLen) continue;
            nr2 = (k + sx0) + R->rpx * (l + sy0);
            if (R->m[nr2] & bbit) { brk = 1; dmax2 = 0; break; }
            dpt2 = img[nr2];
            if (dpt2 >= hiThresh) { R->m[nr2] |= bbit1; brk = 1; dmax2 = 0; break; }
            if (((dpt2 - sky) * invdsky) > R->kft) dmax2 += dpt2;
        }
        if (brk) break;
    }
    return dmax2;
}

/* functions.c:504-716 getPsfCenters: descending threshold sweep, raster scan, greedy
   claim of +-hwKSStamp boxes; keeps the nKSStamps best of up to 2*nKSStamps peaks. */
static int psf_centers(reg_t *R, stamp_t *s, const float *img, int xLen, int yLen, double hiThresh, int bbit1, int bbit2) {
    const zb_params *p = R->p;
    int hwss = p->hw_ksstamp, nkss = p->n_ksstamps;
    int i, j, k, l, nr, nr2, imax, jmax, sx0 = s->x0, sy0 = s->y0, pcnt = 0, brk = 0, fcnt = 2 * nkss;
    int bbit = bbit1 | bbit2 | 0xbf;
    double dmax, dmax2, dpt, dpt2, loPsf, floor_, sky = s->mode, invdsky = 1. / s->fwhm;
    float dfrac = 0.9f;
    int xloc[2 * ZB_MAX_KSS], yloc[2 * ZB_MAX_KSS], qs[2 * ZB_MAX_KSS];
    double peaks[2 * ZB_MAX_KSS];

    if (s->nss >= nkss) return 0;
    floor_ = sky + R->kft * s->fwhm;
    while (pcnt < fcnt) {
        loPsf = sky + (hiThresh - sky) * dfrac;
        loPsf = (loPsf > floor_) ? loPsf : floor_;
        for (j = 0; j < yLen; j++) {
            for (i = 0; i < xLen; i++) {
                nr = (i + sx0) + R->rpx * (j + sy0);
                if (R->m[nr] & bbit) continue;
                dpt = img[nr];
                if (dpt >= hiThresh) { R->m[nr] |= bbit1; continue; }
                if (((dpt - sky) * invdsky) < R->kft) continue;
                if (dpt > loPsf) {
                    dmax = dpt; imax = i; jmax = j;
                    for (l = j - hwss; l <= j + hwss; l++) {
                        if (l < 0 || l >= yLen) continue;
                        for (k = i - hwss; k <= i + hwss; k++) {
                            if (k < 0 || k >= xLen) continue;
                            nr2 = (k + sx0) + R->rpx * (l + sy0);
                            if (R->m[nr2] & bbit) continue;
                            dpt2 = img[nr2];
                            if (dpt2 >= hiThresh) { R->m[nr2] |= bbit1; continue; }
                            if (((dpt2 - sky) * invdsky) < R->kft) continue;
                            if (dpt2 > dmax) { dmax = dpt2; imax = k; jmax = l; }
                        }
                    }
                    dmax2 = check_psf_center(R, img, imax, jmax, xLen, yLen, sx0, sy0, hiThresh,
                                             (float)sky, (float)invdsky, bbit, bbit1);
                    if (dmax2 == 0.) continue;
                    xloc[pcnt] = imax; yloc[pcnt] = jmax; peaks[pcnt++] = dmax2;
                    for (l = jmax - hwss; l <= jmax + hwss; l++)
                        for (k = imax - hwss; k <= imax + hwss; k++)
                            if ((k > 0) && (k < xLen) && (l > 0) && (l < yLen))
                                R->m[(k + sx0) + R->rpx * (l + sy0)] |= bbit2;
                    if (pcnt >= fcnt) brk = 2;
                }
                if (brk == 2) break;
            }
            if (brk == 2) break;
        }
        if (loPsf == floor_) break;
        dfrac -= 0.2;
    }
    if (pcnt == 0) return 1;
    for (i = 0; i < pcnt; i++) qs[i] = i;
    qsort_idx(peaks, qs, 0, pcnt - 1);
    for (i = s->nss, j = 0; j < pcnt && i < nkss; i++, j++) {
        s->xss[i] = xloc[qs[pcnt - j - 1]] + sx0;
        s->yss[i] = yloc[qs[pcnt - j - 1]] + sy0;
        s->nss++;
    }
    return 0;
}

/* functions.c:187-361 buildStamps (getCenters = 1 branch) + cutStamp :363-388 */
/* phase: 1 = the statistics half of buildStamps (functions.c:207-249), 2 = the sub-stamp search half (:251-330),
   3 = both.  With -c b the reference runs both statistics of a grid cell before either search. */
static void build_stamp(reg_t *R, stamp_t *s, const float *img, int sXMin, int sXMax, int sYMin, int sYMax,
                        int rXBMin, int rYBMin, double hiThresh, int bbit1, int bbit2, int phase) {
    int sPixX = sXMax - sXMin + 1, sPixY = sYMax - sYMin + 1;
    int xMin = sXMin - rXBMin, yMin = sYMin - rYBMin, xMax = sXMax - rXBMin, yMax = sYMax - rYBMin, i, j;
    if ((phase & 1) && s->nss == 0) {
        float *area = (float *)calloc((size_t)sPixX * sPixY, sizeof(float));
        stats_t st;
        for (j = yMin; j <= yMax; j++)
            for (i = xMin; i <= xMax; i++) area[(i - xMin) + (j - yMin) * sPixX] = img[i + j * R->rpx];
        s->x0 = xMin; s->y0 = yMin;
        s->x = xMin + (xMax - xMin) / 2; s->y = yMin + (yMax - yMin) / 2;
        st.sum = s->sum; st.mean = s->mean; st.median = s->median; st.mode = s->mode;
        st.sd = s->sd; st.fwhm = s->fwhm; st.lfwhm = s->lfwhm;
        stamp_stats3(R, area, s->x0, s->y0, sPixX, sPixY, &st, 0x0, 0xffff, 3);
        s->sum = st.sum; s->mean = st.mean; s->median = st.median; s->mode = st.mode;
        s->sd = st.sd; s->fwhm = st.fwhm; s->lfwhm = st.lfwhm;
        free(area);
    }
    if (phase & 2) psf_centers(R, s, img, sPixX, sPixY, hiThresh, bbit1, bbit2);
}

/* ------------------------------------------------------------------- basis */

/* alard.c:31-48 getKernelVec + :134-197 kernel_vector.  Basis n = (gaussian g, x-degree,
   y-degree) in g -> px -> py order.  Even/even terms are normalised to unit sum and,
   for n > 0, have basis 0 subtracted. */
void hp_oracle_basis(const zb_params *p, double *kvec, double *fx, double *fy) {
    int hw = p->hw_kernel, fw = 2 * hw + 1, n = 0, g, px, py, ix, i, j;
    for (g = 0; g < p->ngauss; g++)
        for (px = 0; px <= p->deg[g]; px++)
            for (py = 0; py <= p->deg[g] - px; py++, n++) {
                double sx = 0.0, sy = 0.0, *v = kvec + (size_t)n * fw * fw;
                int dx = (px / 2) * 2 - px, dy = (py / 2) * 2 - py;
                for (ix = 0; ix < fw; ix++) {
                    double x = (double)(ix - hw);
                    double qe = exp(-x * x * p->beta[g]);
                    fx[ix + n * fw] = qe * pow(x, px);
                    fy[ix + n * fw] = qe * pow(x, py);
                    sx += fx[ix + n * fw];
                    sy += fy[ix + n * fw];
                }
                sx = 1. / sx; sy = 1. / sy;
                if (dx == 0 && dy == 0)
                    for (ix = 0; ix < fw; ix++) { fx[ix + n * fw] *= sx; fy[ix + n * fw] *= sy; }
                for (i = 0; i < fw; i++)
                    for (j = 0; j < fw; j++) v[i + fw * j] = fx[i + n * fw] * fy[j + n * fw];
                if (dx == 0 && dy == 0 && n > 0)
                    for (i = 0; i < fw * fw; i++) v[i] -= kvec[i];
            }
}

/* --------------------------------------------------------------- fillStamp */

/* alard.c:230-279 xy_conv_stamp.  The y pass accumulates into the FLOAT scratch
   `temp` (globals.h:68), i.e. rounds to float after every tap; the x pass is double. */
static void xy_conv_stamp(reg_t *R, stamp_t *s, const float *image, int n, int ren) {
    int hw = R->p->hw_kernel, hwss = R->p->hw_ksstamp, fw = R->sz.fw_kernel, fwss = R->sz.fw_ksstamp;
    int xi = s->xss[s->sscnt], yi = s->yss[s->sscnt], sub_width = fwss + fw - 1, i, j, xc, yc, xij;
    int npix = fwss * fwss;
    double *imc = s->vec + (size_t)n * npix, *v0 = s->vec;
    float *temp = R->temp;
    for (i = xi - hwss - hw; i <= xi + hwss + hw; i++)
        for (j = yi - hwss; j <= yi + hwss; j++) {
            xij = i - xi + sub_width / 2 + sub_width * (j - yi + hwss);
            temp[xij] = 0.0;
            for (yc = -hw; yc <= hw; yc++)
                temp[xij] += image[i + R->rpx * (j + yc)] * R->fy[hw - yc + n * fw];
        }
    for (j = -hwss; j <= hwss; j++)
        for (i = -hwss; i <= hwss; i++) {
            xij = i + hwss + fwss * (j + hwss);
            imc[xij] = 0.0;
            for (xc = -hw; xc <= hw; xc++)
                imc[xij] += temp[i + xc + sub_width / 2 + sub_width * (j + hwss)] * R->fx[hw - xc + n * fw];
        }
    if (ren) for (i = 0; i < npix; i++) imc[i] -= v0[i];
}

#define MAT(s, nC, i, j) ((s)->mat[(size_t)(i) * (nC) + (j)])

/* alard.c:50-132 fillStamp, with functions.c:390-443 cutSStamp, alard.c:384-434
   build_matrix0 and :436-481 build_scprod0 folded in. */
static int fill_stamp(reg_t *R, stamp_t *s, const float *imConv, const float *imRef) {
    const zb_params *p = R->p;
    int nK = R->sz.n_comp_ker, nC = nK + 2, hwss = p->hw_ksstamp, fwss = R->sz.fw_ksstamp, npix = fwss * fwss;
    int g, px, py, n = 0, i, j, k, xi, yi, di, dj, a, b, nv, xc, yc;
    float rPixX2 = 0.5f * R->rpx, rPixY2 = 0.5f * R->rpy;
    double sum = 0., q, p0;

    if (s->sscnt >= s->nss) return 1;
    for (g = 0; g < p->ngauss; g++)
        for (px = 0; px <= p->deg[g]; px++)
            for (py = 0; py <= p->deg[g] - px; py++, n++) {
                int dx = (px / 2) * 2 - px, dy = (py / 2) * 2 - py;
                xy_conv_stamp(R, s, imConv, n, (dx == 0 && dy == 0 && n > 0));
            }
    /* cutSStamp: reference-side pixels of the sub-stamp, and sum |pix| over non-bad pixels */
    xi = s->xss[s->sscnt]; yi = s->yss[s->sscnt];
    for (j = yi - hwss; j <= yi + hwss; j++)
        for (i = xi - hwss; i <= xi + hwss; i++) {
            float dpt = imRef[i + R->rpx * j];
            s->kref[(i - (xi - hwss)) + (j - (yi - hwss)) * fwss] = dpt;
            sum += (R->m[i + R->rpx * j] & ZB_FLAG_INPUT_ISBAD) ? 0 : fabs(dpt);
        }
    s->sum = sum;
    /* background monomials, alard.c:100-124 */
    di = xi - hwss; dj = yi - hwss;
    for (i = xi - hwss; i <= xi + hwss; i++) {
        double xf = (i - rPixX2) / rPixX2;
        for (j = yi - hwss; j <= yi + hwss; j++) {
            double yf = (j - rPixY2) / rPixY2, ax = 1.0, ay;
            nv = nK;
            for (a = 0; a <= p->bg_order; a++) {
                ay = 1.0;
                for (b = 0; b <= p->bg_order - a; b++) {
                    s->vec[(size_t)nv * npix + (i - di) + fwss * (j - dj)] = ax * ay;
                    ay *= yf; ++nv;
                }
                ax *= xf;
            }
        }
    }
    /* build_matrix0 */
    for (i = 0; i < nK; i++)
        for (j = 0; j <= i; j++) {
            const double *vi = s->vec + (size_t)i * npix, *vj = s->vec + (size_t)j * npix;
            for (q = 0.0, k = 0; k < npix; k++) q += vi[k] * vj[k];
            MAT(s, nC, i + 1, j + 1) = q;
        }
    for (i = 0; i < nK; i++) {
        const double *vi = s->vec + (size_t)i * npix, *vb = s->vec + (size_t)nK * npix;
        for (p0 = 0.0, k = 0; k < npix; k++) p0 += vi[k] * vb[k];
        MAT(s, nC, nK + 1, i + 1) = p0;
    }
    { const double *vb = s->vec + (size_t)nK * npix;
      for (q = 0.0, k = 0; k < npix; k++) q += vb[k] * vb[k];
      MAT(s, nC, nK + 1, nK + 1) = q; }
    /* build_scprod0: note x-outer / y-inner summation order */
    for (i = 0; i <= nK; i++) {
        const double *vi = s->vec + (size_t)i * npix;
        p0 = 0.0;
        for (xc = -hwss; xc <= hwss; xc++)
            for (yc = -hwss; yc <= hwss; yc++)
                p0 += vi[xc + hwss + fwss * (yc + hwss)] * imRef[xc + xi + R->rpx * (yc + yi)];
        s->scprod[i + 1] = p0;
    }
    return 0;
}

/* ---------------------------------------------------------- global system */

/* alard.c:843-965 build_matrix */
static void build_matrix(reg_t *R, stamp_t *st, int nS, double **M) {
    const zb_params *p = R->p;
    int nK = R->sz.n_comp_ker, nC = nK + 2, ncomp1 = nK - 1, ncomp2 = R->sz.n_comp, ncomp = ncomp1 * ncomp2;
    int nbg = R->sz.n_bg_vectors, mat_size = R->sz.mat_size, npix = R->sz.fw_ksstamp * R->sz.fw_ksstamp;
    int i, j, k, is, i1, i2, j1, j2, ibg, jbg, a, b;
    float rPixX2 = 0.5f * R->rpx, rPixY2 = 0.5f * R->rpy;
    for (i = 0; i <= mat_size; i++) for (j = 0; j <= mat_size; j++) M[i][j] = 0.0;
    for (i = 0; i < nS; i++) for (j = 0; j < ncomp2; j++) R->wxy[i][j] = 0.0;
    for (is = 0; is < nS; is++) {
        stamp_t *s = &st[is];
        double fx, fy, a1, a2, p0, q, *w = R->wxy[is];
        if (s->sscnt >= s->nss) continue;
        fx = (s->xss[s->sscnt] - rPixX2) / rPixX2;
        fy = (s->yss[s->sscnt] - rPixY2) / rPixY2;
        k = 0; a1 = 1.0;
        for (a = 0; a <= p->ker_order; a++) {
            a2 = 1.0;
            for (b = 0; b <= p->ker_order - a; b++) { w[k++] = a1 * a2; a2 *= fy; }
            a1 *= fx;
        }
        for (i = 0; i < ncomp; i++) {
            i1 = i / ncomp2; i2 = i - i1 * ncomp2;
            for (j = 0; j <= i; j++) {
                j1 = j / ncomp2; j2 = j - j1 * ncomp2;
                M[i + 2][j + 2] += w[i2] * w[j2] * MAT(s, nC, i1 + 2, j1 + 2);
            }
        }
        M[1][1] += MAT(s, nC, 1, 1);
        for (i = 0; i < ncomp; i++) {
            i1 = i / ncomp2; i2 = i - i1 * ncomp2;
            M[i + 2][1] += w[i2] * MAT(s, nC, i1 + 2, 1);
        }
        for (ibg = 0; ibg < nbg; ibg++) {
            const double *vb = s->vec + (size_t)(ncomp1 + ibg + 1) * npix;
            i = ncomp + ibg + 1;
            for (i1 = 1; i1 < ncomp1 + 1; i1++) {
                const double *vi = s->vec + (size_t)i1 * npix;
                for (p0 = 0.0, k = 0; k < npix; k++) p0 += vi[k] * vb[k];
                for (i2 = 0; i2 < ncomp2; i2++) M[i + 1][(i1 - 1) * ncomp2 + i2 + 2] += p0 * w[i2];
            }
            for (p0 = 0.0, k = 0; k < npix; k++) p0 += s->vec[k] * vb[k];
            M[i + 1][1] += p0;
            for (jbg = 0; jbg <= ibg; jbg++) {
                const double *vj = s->vec + (size_t)(ncomp1 + jbg + 1) * npix;
                for (q = 0.0, k = 0; k < npix; k++) q += vb[k] * vj[k];
                M[i + 1][ncomp + jbg + 2] += q;
            }
        }
    }
    for (i = 0; i < mat_size; i++) for (j = 0; j <= i; j++) M[j + 1][i + 1] = M[i + 1][j + 1];
}

/* alard.c:967-1025 build_scprod */
static void build_scprod(reg_t *R, stamp_t *st, int nS, const float *image, double *sol) {
    int nK = R->sz.n_comp_ker, ncomp1 = nK - 1, ncomp2 = R->sz.n_comp, ncomp = ncomp1 * ncomp2, nbg = R->sz.n_bg_vectors;
    int hwss = R->p->hw_ksstamp, fwss = R->sz.fw_ksstamp, npix = fwss * fwss, is, i, i1, i2, ibg, xc, yc;
    for (i = 0; i <= ncomp + nbg + 1; i++) sol[i] = 0.0;
    for (is = 0; is < nS; is++) {
        stamp_t *s = &st[is]; double p0, q, *w = R->wxy[is]; int xi, yi;
        if (s->sscnt >= s->nss) continue;
        xi = s->xss[s->sscnt]; yi = s->yss[s->sscnt];
        sol[1] += s->scprod[1];
        for (i1 = 1; i1 < ncomp1 + 1; i1++) {
            p0 = s->scprod[i1 + 1];
            for (i2 = 0; i2 < ncomp2; i2++) sol[(i1 - 1) * ncomp2 + i2 + 2] += p0 * w[i2];
        }
        for (ibg = 0; ibg < nbg; ibg++) {
            const double *vb = s->vec + (size_t)(ncomp1 + ibg + 1) * npix;
            q = 0.0;
            for (xc = -hwss; xc <= hwss; xc++)
                for (yc = -hwss; yc <= hwss; yc++)
                    q += vb[xc + hwss + fwss * (yc + hwss)] * image[xc + xi + R->rpx * (yc + yi)];
            sol[ncomp + ibg + 2] += q;
        }
    }
}

/* alard.c:1465-1531 ludcmp: Crout LU, implicit row scaling, partial pivoting (ties: last wins) */
static int lu_decompose(double **a, int n, int *indx) {
    int i, imax = 0, j, k; double big, dum, sum, t, *vv = (double *)malloc((size_t)(n + 1) * sizeof(double));
    for (i = 1; i <= n; i++) {
        big = 0.0;
        for (j = 1; j <= n; j++) if ((t = fabs(a[i][j])) > big) big = t;
        if (big == 0.) { free(vv); return 1; }
        vv[i] = 1.0 / big;
    }
    for (j = 1; j <= n; j++) {
        for (i = 1; i < j; i++) { sum = a[i][j]; for (k = 1; k < i; k++) sum -= a[i][k] * a[k][j]; a[i][j] = sum; }
        big = 0.0;
        for (i = j; i <= n; i++) {
            sum = a[i][j];
            for (k = 1; k < j; k++) sum -= a[i][k] * a[k][j];
            a[i][j] = sum;
            if ((dum = vv[i] * fabs(sum)) >= big) { big = dum; imax = i; }
        }
        if (j != imax) {
            for (k = 1; k <= n; k++) { dum = a[imax][k]; a[imax][k] = a[j][k]; a[j][k] = dum; }
            vv[imax] = vv[j];
        }
        indx[j] = imax;
        if (a[j][j] == 0.0) a[j][j] = 1.0e-20;
        if (j != n) { dum = 1.0 / a[j][j]; for (i = j + 1; i <= n; i++) a[i][j] *= dum; }
    }
    free(vv);
    return 0;
}

/* alard.c:1533-1552 lubksb */
static void lu_backsub(double **a, int n, const int *indx, double *b) {
    int i, ii = 0, ip, j; double sum;
    for (i = 1; i <= n; i++) {
        ip = indx[i]; sum = b[ip]; b[ip] = b[i];
        if (ii) for (j = ii; j <= i - 1; j++) sum -= a[i][j] * b[j];
        else if (sum) ii = i;
        b[i] = sum;
    }
    for (i = n; i >= 1; i--) {
        sum = b[i];
        for (j = i + 1; j <= n; j++) sum -= a[i][j] * b[j];
        b[i] = sum / a[i][i];
    }
}

/* ------------------------------------------------------ model / clip loop */

/* alard.c:1387-1415 get_background */
static double get_background(const reg_t *R, int xi, int yi, const double *sol) {
    int ncompBG = (R->sz.n_comp_ker - 1) * R->sz.n_comp + 1, i, j, k = 1;
    double bg = 0.0, ax = 1.0, ay;
    double xf = (xi - 0.5 * R->rpx) / (0.5 * R->rpx), yf = (yi - 0.5 * R->rpy) / (0.5 * R->rpy);
    for (i = 0; i <= R->p->bg_order; i++) {
        ay = 1.0;
        for (j = 0; j <= R->p->bg_order - i; j++) { bg += sol[ncompBG + k++] * ax * ay; ay *= yf; }
        ax *= xf;
    }
    return bg;
}

/* alard.c:1346-1385 make_kernel: kernel realised at region pixel (xi, yi); returns its sum */
static double make_kernel(reg_t *R, int xi, int yi, const double *sol) {
    int nK = R->sz.n_comp_ker, fw2 = R->sz.fw_kernel * R->sz.fw_kernel, i1, k = 2, ix, iy, i;
    double ax, ay, sum = 0.0;
    double xf = (xi - 0.5 * R->rpx) / (0.5 * R->rpx), yf = (yi - 0.5 * R->rpy) / (0.5 * R->rpy);
    for (i1 = 1; i1 < nK; i1++) {
        R->kcoef[i1] = 0.0; ax = 1.0;
        for (ix = 0; ix <= R->p->ker_order; ix++) {
            ay = 1.0;
            for (iy = 0; iy <= R->p->ker_order - ix; iy++) { R->kcoef[i1] += sol[k++] * ax * ay; ay *= yf; }
            ax *= xf;
        }
    }
    R->kcoef[0] = sol[1];
    for (i = 0; i < fw2; i++) R->kernel[i] = 0.0;
    for (i = 0; i < fw2; i++) {
        for (i1 = 0; i1 < nK; i1++) R->kernel[i] += R->kcoef[i1] * R->kvec[(size_t)i1 * fw2 + i];
        sum += R->kernel[i];
    }
    return sum;
}

/* alard.c:1417-1463 make_model: the model is accumulated in FLOAT (csModel = temp) */
static void make_model(reg_t *R, const stamp_t *s, const double *sol, float *model) {
    int nK = R->sz.n_comp_ker, npix = R->sz.fw_ksstamp * R->sz.fw_ksstamp, i1, k = 2, ix, iy, i;
    int xi = s->xss[s->sscnt], yi = s->yss[s->sscnt];
    double ax, ay, coeff;
    double xf = (xi - 0.5 * R->rpx) / (0.5 * R->rpx), yf = (yi - 0.5 * R->rpy) / (0.5 * R->rpy);
    for (i = 0; i < npix; i++) model[i] = 0.0;
    coeff = sol[1];
    for (i = 0; i < npix; i++) model[i] += coeff * s->vec[i];
    for (i1 = 1; i1 < nK; i1++) {
        const double *v = s->vec + (size_t)i1 * npix;
        coeff = 0.0; ax = 1.0;
        for (ix = 0; ix <= R->p->ker_order; ix++) {
            ay = 1.0;
            for (iy = 0; iy <= R->p->ker_order - ix; iy++) { coeff += sol[k++] * ax * ay; ay *= yf; }
            ax *= xf;
        }
        for (i = 0; i < npix; i++) model[i] += coeff * v[i];
    }
}

/* alard.c:1062-1140 getStampSig, figure of merit 'v' only */
static double stamp_sig(reg_t *R, stamp_t *s, const double *sol, const float *imNoise) {
    int hwss = R->p->hw_ksstamp, fwss = R->sz.fw_ksstamp, i, j, nsig = 0;
    int xR = s->xss[s->sscnt], yR = s->yss[s->sscnt];
    double bg = get_background(R, xR, yR, sol), sig1 = 0;
    float *temp = R->temp;
    make_model(R, s, sol, temp);
    for (j = 0; j < fwss; j++) {
        int y2 = yR - hwss + j;
        for (i = 0; i < fwss; i++) {
            int x2 = xR - hwss + i, idx = i + j * fwss;
            double tdat = temp[idx], idat = s->kref[idx], ndat = imNoise[x2 + R->rpx * y2];
            double diff = tdat - idat + bg;
            if ((R->m[x2 + R->rpx * y2] & ZB_FLAG_INPUT_ISBAD) || (fabs(idat) <= ZEROVAL)) continue;
            temp[idx] = (float)diff;
            if ((tdat * 0.0 != 0.0) || (idat * 0.0 != 0.0)) { R->m[x2 + R->rpx * y2] |= (ZB_FLAG_INPUT_ISBAD | ZB_FLAG_ISNAN); continue; }
            nsig++;
            sig1 += diff * diff / ndat;
        }
    }
    if (nsig > 0) { sig1 /= nsig; if (sig1 >= MAXVAL) sig1 = -1; }
    else sig1 = -1;
    return sig1;
}

/* alard.c:1143-1247 check_again */
static int check_again(reg_t *R, stamp_t *st, int nS, const double *sol, const float *imConv, const float *imRef,
                       const float *imNoise, double *meansig, double *scatter, int *nskipped) {
    float *ss = (float *)calloc((size_t)nS + 1, sizeof(float));
    int nss = 0, is, check = 0; double mean = 0, stdev = 0;
    *nskipped = 0;
    for (is = 0; is < nS; is++) {
        stamp_t *s = &st[is];
        if (s->sscnt < s->nss) {
            double sig1 = stamp_sig(R, s, sol, imNoise);
            if (sig1 == -1) { s->sscnt++; fill_stamp(R, s, imConv, imRef); check = 1; }
            else { s->chi2 = sig1; ss[nss++] = (float)sig1; }
        } else (*nskipped)++;
    }
    hp_oracle_sigma_clip(ss, nss, &mean, &stdev, 10, R->p->stat_sig);
    *meansig = mean; *scatter = stdev;
    for (is = 0; is < nS; is++) {
        stamp_t *s = &st[is];
        if (s->sscnt < s->nss) {
            /* a stamp bumped above keeps its previous chi2 here (alard.c:1180-1186,1224) */
            if ((s->chi2 - mean) > R->p->ker_sig_reject * stdev) { s->sscnt++; fill_stamp(R, s, imConv, imRef); check = 1; }
        }
    }
    free(ss);
    return check;
}

/* alard.c:316-381 fitKernel */
static int fit_kernel(reg_t *R, stamp_t *st, int nS, const float *imRef, const float *imConv, const float *imNoise,
                      double *sol, double *meansig, double *scatter, int *nskipped, int *iters, int *singular) {
    int n = R->sz.mat_size, i, check, *indx = (int *)calloc((size_t)n + 2, sizeof(int));
    double **M = (double **)malloc((size_t)(n + 1) * sizeof(double *));
    for (i = 0; i <= n; i++) M[i] = (double *)malloc((size_t)(n + 1) * sizeof(double));
    R->wxy = (double **)malloc((size_t)nS * sizeof(double *));
    for (i = 0; i < nS; i++) R->wxy[i] = (double *)malloc((size_t)R->sz.n_comp * sizeof(double));
    *iters = 0; *singular = 0;
    do {
        build_matrix(R, st, nS, M);
        build_scprod(R, st, nS, imRef, sol);
        /* a singular matrix makes ludcmp return early; lubksb still runs on it (alard.c:353-354) */
        if (lu_decompose(M, n, indx)) *singular = 1;
        lu_backsub(M, n, indx, sol);
        check = check_again(R, st, nS, sol, imConv, imRef, imNoise, meansig, scatter, nskipped);
        (*iters)++;
    } while (check);
    for (i = 0; i <= n; i++) free(M[i]);
    for (i = 0; i < nS; i++) free(R->wxy[i]);
    free(M); free(R->wxy); free(indx);
    return 0;
}

/* alard.c:483-724 check_stamps (-c b): every stamp is fitted alone with a constant kernel (the leading
   (nCompKer+1)^2 block of its own normal equations), stamps whose kernel sum lies more than kerSigReject sigma
   from the clipped mean are left out, the rest get one trial global fit, and the clipped mean of their
   getStampSig figures divided by the kernel sum at region pixel (0,0) is the merit of this direction.
   Figure of merit 'v' only (sig2 = sig3 = 0 then, alard.c:1082-1083,1127).  A direction without stamps returns
   666 (what the reference's fall-through yields for it, after reading unset memory on the way). */
static double check_stamps(reg_t *R, stamp_t *st, int nS, const float *imRef, const float *imNoise) {
    int nK = R->sz.n_comp_ker, nC = nK + 2, nComps = nK + 1, n = R->sz.mat_size;
    int i, im, jm, nks = 0, ntest = 0, mcnt1 = 0, mcnt2 = 0, *indx;
    double **cm, *cv, kmean = 0, kstdev = 0, merit1 = 0, merit2 = 0, sd1 = 0, sd2 = 0, **M, *tsol, *norm, *dif;
    float *ks, *m1, *m2; stamp_t *ts;
    if (nS <= 0) return 666;
    ks = (float *)calloc((size_t)nS, sizeof(float));
    norm = (double *)calloc((size_t)nS, sizeof(double)); dif = (double *)calloc((size_t)nS, sizeof(double));
    cm = (double **)malloc((size_t)nC * sizeof(double *));
    for (i = 0; i < nC; i++) cm[i] = (double *)calloc((size_t)nC, sizeof(double));
    cv = (double *)calloc((size_t)nC, sizeof(double));
    indx = (int *)calloc((size_t)(n > nC ? n : nC) + 2, sizeof(int));
    for (i = 0; i < nS; i++) {
        stamp_t *s = &st[i];
        for (im = 1; im <= nComps; im++) {
            cv[im] = s->scprod[im];
            for (jm = 1; jm <= im; jm++) { cm[im][jm] = MAT(s, nC, im, jm); cm[jm][im] = cm[im][jm]; }
        }
        lu_decompose(cm, nComps, indx);
        lu_backsub(cm, nComps, indx, cv);
        norm[i] = cv[1];
        ks[nks++] = (float)cv[1];
    }
    hp_oracle_sigma_clip(ks, nks, &kmean, &kstdev, 10, R->p->stat_sig);
    for (i = 0; i < nS; i++) { dif[i] = fabs((norm[i] - kmean) / kstdev); if (dif[i] < R->p->ker_sig_reject) ntest++; }
    ts = (stamp_t *)calloc((size_t)ntest + 1, sizeof(stamp_t));
    for (i = 0, ntest = 0; i < nS; i++) if (dif[i] < R->p->ker_sig_reject) ts[ntest++] = st[i];
    M = (double **)malloc((size_t)(n + 1) * sizeof(double *));
    for (i = 0; i <= n; i++) M[i] = (double *)calloc((size_t)n + 1, sizeof(double));
    R->wxy = (double **)malloc((size_t)nS * sizeof(double *));
    for (i = 0; i < nS; i++) R->wxy[i] = (double *)calloc((size_t)R->sz.n_comp, sizeof(double));
    tsol = (double *)calloc((size_t)R->sz.n_comp_total + 1, sizeof(double));
    build_matrix(R, ts, ntest, M);
    build_scprod(R, ts, ntest, imRef, tsol);
    lu_decompose(M, n, indx);
    lu_backsub(M, n, indx, tsol);
    kmean = make_kernel(R, 0, 0, tsol);
    m1 = (float *)calloc((size_t)ntest + 1, sizeof(float)); m2 = (float *)calloc((size_t)ntest + 1, sizeof(float));
    if (getenv("HP_ORACLE_DEBUG")) {
        fprintf(stderr, "CHK region nS %d kmean %.17g\n", nS, kmean);
        for (i = 0; i < nS; i++) fprintf(stderr, "CHK  stamp %d ksum %.17g excl %d\n", i, norm[i], !(dif[i] < R->p->ker_sig_reject));
    }
    for (i = 0; i < ntest; i++) {
        double sig1 = stamp_sig(R, &ts[i], tsol, imNoise);
        if (getenv("HP_ORACLE_DEBUG")) fprintf(stderr, "CHK  test %d sig1 %.17g\n", i, sig1);
        if (sig1 != -1 && sig1 <= MAXVAL) m1[mcnt1++] = (float)sig1;
        m2[mcnt2++] = 0.f;                                 /* sig2 and sig3 stay 0 with -fom v */
    }
    hp_oracle_sigma_clip(m1, mcnt1, &merit1, &sd1, 10, R->p->stat_sig);
    hp_oracle_sigma_clip(m2, mcnt2, &merit2, &sd2, 10, R->p->stat_sig);
    merit1 /= kmean; merit2 /= kmean;
    for (i = 0; i <= n; i++) free(M[i]);
    for (i = 0; i < nS; i++) free(R->wxy[i]);
    for (i = 0; i < nC; i++) free(cm[i]);
    free(M); free(R->wxy); R->wxy = NULL; free(cm); free(cv); free(indx); free(ks); free(norm); free(dif);
    free(ts); free(tsol); free(m1); free(m2);
    if (mcnt1 > 0) return merit1;
    if (mcnt2 > 0) return merit2;
    return 666;
}

/* ----------------------------------------------------------- convolution */

/* alard.c:1249-1344 spatial_convolve: one kernel per kcStep x kcStep block, realised at
   the block centre; image and variance convolved with double accumulation; the mask of
   the convolved image is OR-ed over the footprint. Writes R->m (output mask). */
static void spatial_convolve(reg_t *R, const float *image, float **variance, const double *sol, float *out, const int *cMask) {
    int xS = R->rpx, yS = R->rpy, hw = R->p->hw_kernel, fw = R->sz.fw_kernel, kc = R->sz.kc_step;
    int nsx = (int)ceil((double)xS / (double)kc), nsy = (int)ceil((double)yS / (double)kc);
    int i1, j1, i2, j2, i, j, i0, j0, ic, jc, ik, jk, nc, ni, mbit;
    float *vData = (float *)calloc((size_t)xS * yS, sizeof(float));
    const float *var = *variance;
    for (j1 = 0; j1 < nsy; j1++) {
        j0 = j1 * kc + hw;
        for (i1 = 0; i1 < nsx; i1++) {
            i0 = i1 * kc + hw;
            make_kernel(R, i0 + hw, j0 + hw, sol);
            for (j2 = 0; j2 < kc; j2++) {
                j = j0 + j2;
                if (j >= yS - hw) break;
                for (i2 = 0; i2 < kc; i2++) {
                    double q = 0.0, qv = 0.0, aks = 0.0, uks = 0.0, kk;
                    i = i0 + i2;
                    if (i >= xS - hw) break;
                    ni = i + xS * j; mbit = 0;
                    for (jc = j - hw; jc <= j + hw; jc++) {
                        jk = j - jc + hw;
                        for (ic = i - hw; ic <= i + hw; ic++) {
                            ik = i - ic + hw;
                            nc = ic + xS * jc;
                            kk = R->kernel[ik + jk * fw];
                            q += image[nc] * kk;
                            if (R->p->convolve_variance) qv += var[nc] * kk;
                            else qv += var[nc] * kk * kk;
                            mbit |= cMask[nc];
                            aks += fabs(kk);
                            if (!(cMask[nc] & ZB_FLAG_INPUT_ISBAD)) uks += fabs(kk);
                        }
                    }
                    out[ni] = (float)q;
                    vData[ni] = (float)qv;
                    R->m[ni] |= cMask[ni];
                    R->m[ni] |= ZB_FLAG_OUTPUT_ISBAD * ((cMask[ni] & ZB_FLAG_INPUT_ISBAD) > 0);
                    if (mbit) {
                        if ((uks / aks) < R->p->ker_frac_mask) R->m[ni] |= (ZB_FLAG_OUTPUT_ISBAD | ZB_FLAG_BAD_CONV);
                        else R->m[ni] |= ZB_FLAG_OK_CONV;
                    }
                }
            }
        }
    }
    free(*variance);
    *variance = vData;
}

/* alard.c:1027-1060 getFinalStampSig: float products, double sum */
static double final_stamp_sig(reg_t *R, const stamp_t *s, const float *imDiff, const float *imNoise) {
    int hwss = R->p->hw_ksstamp, fwss = R->sz.fw_ksstamp, i, j, nsig = 0;
    int xR = s->xss[s->sscnt], yR = s->yss[s->sscnt];
    double sig = 0;
    for (j = 0; j < fwss; j++)
        for (i = 0; i < fwss; i++) {
            int idx = (xR - hwss + i) + R->rpx * (yR - hwss + j);
            float idat = imDiff[idx], indat = (float)(1. / imNoise[idx]);
            if (R->m[idx] & ZB_FLAG_INPUT_ISBAD) continue;
            nsig++;
            sig += idat * idat * indat * indat;
        }
    return nsig > 0 ? sig / nsig : -1;
}

/* ------------------------------------------------------------ main driver */

/* functions.c:1172-1187 makeNoiseImage4 */
static void noise_model(const float *img, float invGain, float quad, float *out, int n) {
    double qquad = quad * quad; int i;
    for (i = n; i--;) out[i] = (float)(fabs(img[i]) * invGain + qquad);
}

static void crop_f(const float *src, int nx, int bx0, int by0, int rpx, int rpy, float *dst) {
    int y; for (y = 0; y < rpy; y++) memcpy(dst + (size_t)y * rpx, src + (size_t)(by0 + y) * nx + bx0, (size_t)rpx * sizeof(float));
}

static void value_bits(int *mask, const float *img, int n, float fill, float hi, float lo, int extra) {
    int l;
    for (l = n; l--;) {
        mask[l] |= (extra | ZB_FLAG_INPUT_ISBAD | ZB_FLAG_BAD_PIXVAL) * (img[l] == fill);
        mask[l] |= (extra | ZB_FLAG_INPUT_ISBAD | ZB_FLAG_SAT_PIXEL) * (img[l] >= hi);
        mask[l] |= (extra | ZB_FLAG_INPUT_ISBAD | ZB_FLAG_LOW_PIXEL) * (img[l] <= lo);
    }
}

static void frame_or(int *m, int rpx, int rpy, int w, int bits) {
    int l, k;
    for (l = 0; l < rpy; l++) {
        for (k = 0; k < w; k++) m[k + rpx * l] |= bits;
        for (k = rpx - w; k < rpx; k++) m[k + rpx * l] |= bits;
    }
    for (l = 0; l < w; l++) for (k = w; k < rpx - w; k++) m[k + rpx * l] |= bits;
    for (l = rpy - w; l < rpy; l++) for (k = w; k < rpx - w; k++) m[k + rpx * l] |= bits;
}

/* hotpants/main.c:606-1867 for one image pair (the per-region loop), on arrays. */
int hp_oracle_subtract(const zb_params *p,
                       const float *sci, const float *ref,
                       const float *sci_rms, const float *ref_rms,
                       const int32_t *sci_mask, const int32_t *ref_mask,
                       float *diff, float *noise, int32_t *out_mask,
                       zb_region_info *regions, double *kernel_sol, zb_stamp_info *stamps_out) {
    zb_sizes sz; reg_t R; int ir, jr, nR = 0, nK, nC, nvec, npix, fw, k, l;
    const int both = (p->force_convolve == 'b');
    int convTmpl = (p->force_convolve == 't');
    float tU = p->t_uthresh - p->t_pedestal, tUK = p->t_ukthresh - p->t_pedestal, tL = p->t_lthresh - p->t_pedestal;
    float iU = p->i_uthresh - p->i_pedestal, iUK = p->i_ukthresh - p->i_pedestal, iL = p->i_lthresh - p->i_pedestal;

    if (p->force_convolve != 't' && p->force_convolve != 'i' && !both) return ZB_ERR_UNSUPPORTED;
    if (p->fig_merit != 'v' || p->use_full_ss) return ZB_ERR_UNSUPPORTED;
    if (p->n_ksstamps > ZB_MAX_KSS || p->ngauss > ZB_MAX_GAUSS) return ZB_ERR_UNSUPPORTED;
    hp_oracle_sizes(p, &sz);
    memset(&R, 0, sizeof R);
    R.p = p; R.sz = sz;
    nK = sz.n_comp_ker; nC = nK + 2; nvec = nK + sz.n_bg_vectors; npix = sz.fw_ksstamp * sz.fw_ksstamp; fw = sz.fw_kernel;
    R.kvec = (double *)calloc((size_t)nK * fw * fw, sizeof(double));
    R.fx = (double *)calloc((size_t)nK * fw, sizeof(double));
    R.fy = (double *)calloc((size_t)nK * fw, sizeof(double));
    R.kernel = (double *)calloc((size_t)fw * fw, sizeof(double));
    R.kcoef = (double *)calloc((size_t)nK, sizeof(double));
    R.temp = (float *)calloc((size_t)(sz.fw_ksstamp + fw) * sz.fw_ksstamp, sizeof(float));
    hp_oracle_basis(p, R.kvec, R.fx, R.fy);

    memset(diff, 0, (size_t)p->nx * p->ny * sizeof(float));
    memset(noise, 0, (size_t)p->nx * p->ny * sizeof(float));
    if (out_mask) memset(out_mask, 0, (size_t)p->nx * p->ny * sizeof(int32_t));
    if (stamps_out) memset(stamps_out, 0, (size_t)sz.n_regions * sz.n_stamps * sizeof(zb_stamp_info));
    if (kernel_sol) memset(kernel_sol, 0, (size_t)sz.n_regions * (sz.n_comp_total + 1) * sizeof(double));

    for (jr = 0; jr < p->nreg_y; jr++)
    for (ir = 0; ir < p->nreg_x; ir++, nR++) {
        zb_region_info *ri = &regions[nR];
        int xMax = p->nx - 1, yMax = p->ny - 1;
        int rXMin = ir * xMax / p->nreg_x, rYMin = jr * yMax / p->nreg_y;
        int rXMax = imin_((ir + 1) * xMax / p->nreg_x, xMax), rYMax = imin_((jr + 1) * yMax / p->nreg_y, yMax);
        int buf = sz.n_regions > 1 ? sz.fw_stamp / 2 : p->hw_kernel;
        int bx0 = imax_(0, rXMin - buf), by0 = imax_(0, rYMin - buf), bx1 = imin_(xMax, rXMax + buf), by1 = imin_(yMax, rYMax + buf);
        int xBufLo = rXMin - bx0, yBufLo = rYMin - by0;
        int rpx = bx1 - bx0 + 1, rpy = by1 - by0 + 1, n = rpx * rpy, nS = 0, attempt, is, x, y;
        float *t, *i_, *o, *e, *nz; int *m, *mis, *mts, *cmask, *omask_other;
        stamp_t *st, *st2 = NULL; int nS2 = 0; double tMerit = 0, iMerit = 0;   /* -c b: st = template stamps, st2 = image stamps */
        double *sol, ksum, inv1, sum; stats_t sd_, sn_, sdm, snm;
        const float *imConv, *imRef; float *noiseOut; int iters = 0, singular = 0, nskip = 0;
        double meansig = 0, scatter = 0, meansigF = 0, scatterF = 0, x2norm; int nx2norm;

        memset(ri, 0, sizeof *ri);
        ri->rx_min = rXMin; ri->rx_max = rXMax; ri->ry_min = rYMin; ri->ry_max = rYMax;
        ri->bx_min = bx0; ri->bx_max = bx1; ri->by_min = by0; ri->by_max = by1;
        R.rpx = rpx; R.rpy = rpy;
        t = (float *)malloc((size_t)n * 4); i_ = (float *)malloc((size_t)n * 4);
        o = (float *)malloc((size_t)n * 4); e = (float *)malloc((size_t)n * 4); nz = (float *)malloc((size_t)n * 4);
        m = (int *)calloc((size_t)n, sizeof(int)); mis = (int *)calloc((size_t)n, sizeof(int)); mts = (int *)calloc((size_t)n, sizeof(int));
        R.m = m;
        crop_f(ref, p->nx, bx0, by0, rpx, rpy, t);
        crop_f(sci, p->nx, bx0, by0, rpx, rpy, i_);
        if (p->t_pedestal != 0.f || p->i_pedestal != 0.f)
            for (l = n; l--;) { t[l] -= p->t_pedestal; i_[l] -= p->i_pedestal; }
        /* approximate variance for the fit's figure of merit, main.c:739-770 */
        if (sci_rms) { crop_f(sci_rms, p->nx, bx0, by0, rpx, rpy, o); for (l = n; l--;) o[l] *= o[l]; }
        else noise_model(i_, (float)(1. / p->i_gain), p->i_rdnoise / p->i_gain, o, n);
        if (ref_rms) { crop_f(ref_rms, p->nx, bx0, by0, rpx, rpy, e); for (l = n; l--;) e[l] *= e[l]; }
        else noise_model(t, (float)(1. / p->t_gain), p->t_rdnoise / p->t_gain, e, n);
        for (l = n; l--;) o[l] += e[l];
        /* input masks, main.c:778-803 */
        if (sci_mask) for (y = 0; y < rpy; y++) for (x = 0; x < rpx; x++) {
            int v = sci_mask[(size_t)(by0 + y) * p->nx + bx0 + x]; l = x + rpx * y;
            mis[l] = v | (ZB_FLAG_INPUT_MASK * (v > 0)); m[l] |= mis[l];
        }
        if (ref_mask) for (y = 0; y < rpy; y++) for (x = 0; x < rpx; x++) {
            int v = ref_mask[(size_t)(by0 + y) * p->nx + bx0 + x]; l = x + rpx * y;
            mts[l] = v | (ZB_FLAG_INPUT_MASK * (v > 0)); m[l] |= mts[l];
        }
        /* makeInputMask functions.c:1388-1405 + spreadMask :1353-1386 */
        for (l = n; l--;) {
            m[l] |= (ZB_FLAG_INPUT_ISBAD | ZB_FLAG_BAD_PIXVAL) * (t[l] == p->fill_val || i_[l] == p->fill_val);
            m[l] |= (ZB_FLAG_INPUT_ISBAD | ZB_FLAG_SAT_PIXEL) * (t[l] >= tU || i_[l] >= iU);
            m[l] |= (ZB_FLAG_INPUT_ISBAD | ZB_FLAG_LOW_PIXEL) * (t[l] <= tL || i_[l] <= iL);
        }
        { int width = (int)(p->hw_kernel * p->kf_spread_mask1), w2 = width / 2, ii, jj;
          if (width > 0)
            for (y = 0; y < rpy; y++) for (x = 0; x < rpx; x++)
                if (m[x + rpx * y] & ZB_FLAG_INPUT_ISBAD)
                    for (k = -w2; k <= w2; k++) { ii = x + k; if (ii < 0 || ii >= rpx) continue;
                        for (l = -w2; l <= w2; l++) { jj = y + l; if (jj < 0 || jj >= rpy) continue;
                            m[ii + rpx * jj] |= ZB_FLAG_OK_CONV * (!(m[ii + rpx * jj] & ZB_FLAG_INPUT_ISBAD)); } } }
        frame_or(m, rpx, rpy, sz.s_border, ZB_FLAG_T_BAD | ZB_FLAG_I_BAD);   /* main.c:811-824 */

        /* stamp search with one retry at a lower threshold, main.c:829-998 */
        st = (stamp_t *)calloc((size_t)sz.n_stamps + 1, sizeof(stamp_t));
        if (both) st2 = (stamp_t *)calloc((size_t)sz.n_stamps + 1, sizeof(stamp_t));
        R.kft = p->ker_fit_thresh;
        for (attempt = 1; attempt <= 2; attempt++) {
            int flag = 0, gk, gl;
            nS = 0; nS2 = 0;
            for (gl = 0; gl < sz.nstamp_y; gl++)
                for (gk = 0; gk < sz.nstamp_x; gk++) {
                    int sXMin = bx0 + gk * rpx / sz.nstamp_x, sYMin = by0 + gl * rpy / sz.nstamp_y;
                    int sXMax = imin_(sXMin + sz.fw_stamp - 1, bx1), sYMax = imin_(sYMin + sz.fw_stamp - 1, by1);
                    stamp_t *s = &st[nS];
                    s->sscnt = s->nss = 0; s->chi2 = 0.0;       /* other stats persist in a reused slot */
                    if (both) {       /* functions.c:207-330: template statistics, image statistics, template search, image search */
                        stamp_t *s2 = &st2[nS2];
                        s2->sscnt = s2->nss = 0; s2->chi2 = 0.0;
                        build_stamp(&R, s, t, sXMin, sXMax, sYMin, sYMax, bx0, by0, tUK, ZB_FLAG_T_BAD, ZB_FLAG_T_SKIP, 1);
                        build_stamp(&R, s2, i_, sXMin, sXMax, sYMin, sYMax, bx0, by0, iUK, ZB_FLAG_I_BAD, ZB_FLAG_I_SKIP, 1);
                        build_stamp(&R, s, t, sXMin, sXMax, sYMin, sYMax, bx0, by0, tUK, ZB_FLAG_T_BAD, ZB_FLAG_T_SKIP, 2);
                        build_stamp(&R, s2, i_, sXMin, sXMax, sYMin, sYMax, bx0, by0, iUK, ZB_FLAG_I_BAD, ZB_FLAG_I_SKIP, 2);
                        if (s2->nss > 0) nS2++;
                    }
                    else if (convTmpl) build_stamp(&R, s, t, sXMin, sXMax, sYMin, sYMax, bx0, by0, tUK, ZB_FLAG_T_BAD, ZB_FLAG_T_SKIP, 3);
                    else               build_stamp(&R, s, i_, sXMin, sXMax, sYMin, sYMax, bx0, by0, iUK, ZB_FLAG_I_BAD, ZB_FLAG_I_SKIP, 3);
                    if (s->nss > 0) nS++;
                }
            if ((nS / (float)sz.n_stamps) < p->min_frac_good_stamps) flag = 1;
            if (both && (nS2 / (float)sz.n_stamps) < p->min_frac_good_stamps) flag = 1;     /* main.c:921-924 */
            if (!flag) break;
            if (attempt == 1 && p->scale_fit_thresh < 1.f) {   /* stamps are re-allocated (zeroed) only here */
                R.kft *= p->scale_fit_thresh;
                memset(st, 0, ((size_t)sz.n_stamps + 1) * sizeof(stamp_t));
                if (both) memset(st2, 0, ((size_t)sz.n_stamps + 1) * sizeof(stamp_t));
                for (l = n; l--;) m[l] &= ~0xa00;
                ri->flags |= ZB_REGION_RETRIED;
            }
        }
        if (both) {
            /* main.c:1026-1072: fill both sets, merit of each direction, the smaller one wins (:1076-1085) */
            stamp_t *sets[2]; int cnt[2], d;
            sets[0] = st; sets[1] = st2; cnt[0] = nS; cnt[1] = nS2;
            for (d = 0; d < 2 && (nS > 0 || nS2 > 0); d++) {
                for (is = 0; is < cnt[d]; is++) {
                    stamp_t *s = &sets[d][is];
                    s->vec = (double *)calloc((size_t)nvec * npix, sizeof(double));
                    s->kref = (double *)calloc((size_t)npix, sizeof(double));
                    s->mat = (double *)calloc((size_t)nC * nC, sizeof(double));
                    s->scprod = (double *)calloc((size_t)nC, sizeof(double));
                    s->sscnt = 0; fill_stamp(&R, s, d == 0 ? t : i_, d == 0 ? i_ : t);
                }
                if (d == 0) tMerit = check_stamps(&R, st, nS, i_, o);
                else        iMerit = check_stamps(&R, st2, nS2, t, o);
            }
            convTmpl = (tMerit < iMerit);
            ri->t_merit = tMerit; ri->i_merit = iMerit;
            if (!convTmpl) { stamp_t *q = st; st = st2; st2 = q; d = nS; nS = nS2; nS2 = d; }
            for (is = 0; is < nS2; is++) { free(st2[is].vec); free(st2[is].kref); free(st2[is].mat); free(st2[is].scprod); }
            free(st2); st2 = NULL;
        }
        ri->conv_template = convTmpl;
        ri->n_stamps = nS;
        if (nS == 0) {          /* main.c:1002-1009: region silently skipped (with -c b: the winning direction has no stamps) */
            ri->flags |= ZB_REGION_SKIPPED;
            free(t); free(i_); free(o); free(e); free(nz); free(m); free(mis); free(mts); free(st);
            continue;
        }
        imConv = convTmpl ? t : i_; imRef = convTmpl ? i_ : t;
        if (!both) for (is = 0; is < nS; is++) {
            st[is].vec = (double *)calloc((size_t)nvec * npix, sizeof(double));
            st[is].kref = (double *)calloc((size_t)npix, sizeof(double));
            st[is].mat = (double *)calloc((size_t)nC * nC, sizeof(double));
            st[is].scprod = (double *)calloc((size_t)nC, sizeof(double));
            st[is].sscnt = 0; fill_stamp(&R, &st[is], imConv, imRef);
        }
        sol = (double *)calloc((size_t)sz.n_comp_total + 1, sizeof(double));
        fit_kernel(&R, st, nS, imRef, imConv, o, sol, &meansig, &scatter, &nskip, &iters, &singular);
        if (singular) ri->flags |= ZB_REGION_SINGULAR;

        /* output stage, main.c:1088-1173 (conv template) / :1207-1300 (conv image) */
        for (l = n; l--;) o[l] = p->fill_val;
        cmask = convTmpl ? mts : mis; omask_other = convTmpl ? mis : mts;
        if (convTmpl) value_bits(cmask, t, n, p->fill_val, tU, tL, 0);
        else          value_bits(cmask, i_, n, p->fill_val, iU, iL, 0);
        memset(m, 0, (size_t)n * sizeof(int));
        if (convTmpl) { if (ref_rms) { crop_f(ref_rms, p->nx, bx0, by0, rpx, rpy, e); for (l = n; l--;) e[l] *= e[l]; }
                        else noise_model(t, (float)(1. / p->t_gain), p->t_rdnoise / p->t_gain, e, n); }
        else          { if (sci_rms) { crop_f(sci_rms, p->nx, bx0, by0, rpx, rpy, e); for (l = n; l--;) e[l] *= e[l]; }
                        else noise_model(i_, (float)(1. / p->i_gain), p->i_rdnoise / p->i_gain, e, n); }
        { float *ev = (float *)malloc((size_t)n * 4); memcpy(ev, e, (size_t)n * 4);
          spatial_convolve(&R, imConv, &ev, sol, o, cmask); memcpy(e, ev, (size_t)n * 4); free(ev); }
        for (l = p->hw_kernel; l < rpy - p->hw_kernel; l++)
            for (k = p->hw_kernel; k < rpx - p->hw_kernel; k++) o[k + rpx * l] += get_background(&R, k, l, sol);
        ksum = make_kernel(&R, rpx / 2, rpy / 2, sol);
        if (convTmpl) { if (sci_rms) { crop_f(sci_rms, p->nx, bx0, by0, rpx, rpy, nz); for (l = n; l--;) nz[l] *= nz[l]; }
                        else noise_model(i_, (float)(1. / p->i_gain), p->i_rdnoise / p->i_gain, nz, n); }
        else          { if (ref_rms) { crop_f(ref_rms, p->nx, bx0, by0, rpx, rpy, nz); for (l = n; l--;) nz[l] *= nz[l]; }
                        else noise_model(t, (float)(1. / p->t_gain), p->t_rdnoise / p->t_gain, nz, n); }
        for (l = n; l--;) nz[l] = (float)sqrt(nz[l] + e[l]);
        noiseOut = nz;
        if (convTmpl) value_bits(m, i_, n, p->fill_val, iU, iL, ZB_FLAG_OUTPUT_ISBAD);
        else          value_bits(m, t, n, p->fill_val, tU, tL, ZB_FLAG_OUTPUT_ISBAD);
        for (l = n; l--;) { m[l] |= omask_other[l]; m[l] |= ZB_FLAG_OUTPUT_ISBAD * ((omask_other[l] & ZB_FLAG_INPUT_ISBAD) > 0); }
        frame_or(m, rpx, rpy, p->hw_kernel, ZB_FLAG_OUTPUT_ISBAD);          /* main.c:1320-1331 */

        /* subtract / scale, main.c:1399-1415 and :1439-1454 */
        inv1 = 1. / ksum;
        for (l = p->hw_kernel; l < rpy - p->hw_kernel; l++)
            for (k = p->hw_kernel; k < rpx - p->hw_kernel; k++) {
                int q = k + rpx * l;
                if (convTmpl) {
                    o[q] -= i_[q];
                    if (p->phot_normalize == 't') { o[q] *= inv1; noiseOut[q] *= inv1; }
                    o[q] *= -1.;
                } else {
                    o[q] -= t[q];
                    if (p->phot_normalize == 'i') { o[q] *= inv1; noiseOut[q] *= inv1; }
                }
            }
        /* FSIG / FSCAT, main.c:1417-1435 */
        { float *t2 = (float *)calloc((size_t)nS + 1, sizeof(float)); int kk = 0;
          for (is = 0; is < nS; is++) if (st[is].sscnt < st[is].nss) { sum = final_stamp_sig(&R, &st[is], o, noiseOut); t2[kk++] = (float)sum; }
          hp_oracle_sigma_clip(t2, kk, &meansigF, &scatterF, 10, p->stat_sig); free(t2); }
        /* region QA statistics, main.c:1479-1554 */
        memset(&sd_, 0, sizeof sd_); memset(&sn_, 0, sizeof sn_); memset(&sdm, 0, sizeof sdm); memset(&snm, 0, sizeof snm);
        stamp_stats3(&R, o, 0, 0, rpx, rpy, &sd_, 0x0, 0xffff, 5);
        stamp_stats3(&R, noiseOut, 0, 0, rpx, rpy, &sn_, 0x0, 0xffff, 5);
        noise_stats3(&R, o, noiseOut, &x2norm, &nx2norm, 0x0, 0xffff);
        stamp_stats3(&R, o, 0, 0, rpx, rpy, &sdm, 0xff, ZB_FLAG_OUTPUT_ISBAD, 5);
        stamp_stats3(&R, noiseOut, 0, 0, rpx, rpy, &snm, 0xff, ZB_FLAG_OUTPUT_ISBAD, 5);
        ri->nscalo = sd_.sd / sn_.mean;
        if (p->rescale_ok) {                                                /* main.c:1557-1579 */
            double diffrat = (sdm.sd / snm.mean) / ri->nscalo;
            ri->nscalo = diffrat;
            if (diffrat > 1) for (l = n; l--;) if ((m[l] & 0xff) && (!(m[l] & ZB_FLAG_OUTPUT_ISBAD))) noiseOut[l] *= diffrat;
        }
        /* output masking, main.c:1584-1605 */
        if (p->kf_spread_mask2 >= 0)
            for (l = p->hw_kernel; l < rpy - p->hw_kernel; l++)
                for (k = p->hw_kernel; k < rpx - p->hw_kernel; k++)
                    if (m[k + rpx * l] & ZB_FLAG_OUTPUT_ISBAD) { o[k + rpx * l] = p->fill_val; noiseOut[k + rpx * l] = p->fill_val_noise; }
        /* write the un-buffered sub-rectangle, main.c:655-658 + functions.c:1515-1552 */
        for (y = 0; y <= rYMax - rYMin; y++)
            for (x = 0; x <= rXMax - rXMin; x++) {
                size_t src = (size_t)(xBufLo + x) + (size_t)rpx * (yBufLo + y), dst = (size_t)(rXMin + x) + (size_t)p->nx * (rYMin + y);
                diff[dst] = o[src]; noise[dst] = noiseOut[src];
                if (out_mask) out_mask[dst] = m[src];
            }
        ri->ksum = ksum; ri->sssig = meansig; ri->ssscat = scatter; ri->fsig = meansigF; ri->fscat = scatterF;
        ri->x2norm = x2norm; ri->nx2norm = nx2norm; ri->n_skipped = nskip; ri->fit_iters = iters;
        ri->dmean = sd_.mean; ri->dsige = sd_.sd; ri->dsig = sn_.mean;
        ri->dmeano = sdm.mean; ri->dsigeo = sdm.sd; ri->dsigo = snm.mean;
        if (kernel_sol) memcpy(kernel_sol + (size_t)nR * (sz.n_comp_total + 1), sol, ((size_t)sz.n_comp_total + 1) * sizeof(double));
        if (stamps_out)
            for (is = 0; is < nS; is++) {
                zb_stamp_info *so = &stamps_out[(size_t)nR * sz.n_stamps + is]; int q;
                so->x0 = st[is].x0; so->y0 = st[is].y0; so->nss = st[is].nss; so->sscnt = st[is].sscnt;
                for (q = 0; q < st[is].nss && q < ZB_MAX_KSS; q++) { so->xss[q] = st[is].xss[q]; so->yss[q] = st[is].yss[q]; }
                so->mode = st[is].mode; so->fwhm = st[is].fwhm; so->chi2 = st[is].chi2;
            }
        for (is = 0; is < nS; is++) { free(st[is].vec); free(st[is].kref); free(st[is].mat); free(st[is].scprod); }
        free(sol); free(t); free(i_); free(o); free(e); free(nz); free(m); free(mis); free(mts); free(st);
    }
    free(R.kvec); free(R.fx); free(R.fy); free(R.kernel); free(R.kcoef); free(R.temp);
    return 0;
}
