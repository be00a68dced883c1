// zb_hist.cuh — the histogram analysis of getStampStats3 (hotpants/functions.c:953-1031),
// shared by the stamp search (zb_stamps.cu) and the region QA statistics (zb_stats.cu).
#pragma once
#include "zb_internal.h"

__device__ static inline bool zb_excluded(int mdat, double rdat, int umask, int smask) {
    return ((umask > 0) && (!(mdat & umask))) || ((smask > 0) && (mdat & smask)) || (fabs(rdat) <= ZB_ZEROVAL);
}

struct HistOut { double mode, median, fwhm, lfwhm; int action; };   // action: 0 done, 1 widen, 2 narrow

// functions.c:953-1031: mode / quartiles / median from the 256-bin histogram (single thread)
__device__ static void hist_analyse(const int *bins, int goodcnt, double bin1, double binsize, int &imax, HistOut &o) {
    double sumx, sumxx, maxdens, mode_bin, lower, upper, med;
    int i, ilower, iupper;
    sumx = maxdens = 0.;
    for (ilower = iupper = 1; iupper < 255; sumx -= bins[ilower++]) {
        while ((sumx < goodcnt / 10.) && (iupper < 255)) sumx += bins[iupper++];
        if (sumx / (iupper - ilower) > maxdens) { maxdens = sumx / (iupper - ilower); imax = ilower; }
    }
    if (imax < 0 || imax > 255) imax = 0;
    sumxx = sumx = 0.;
    for (i = imax; (sumx < goodcnt / 10.) && (i < 255); i++) { sumx += bins[i]; sumxx += i * bins[i]; }
    mode_bin = sumxx / sumx + 0.5;
    o.mode = bin1 + binsize * (mode_bin - 1.);
    imax = (int)floor(mode_bin);
    lower = goodcnt * 0.25;
    upper = goodcnt * 0.75;
    sumx = 0.;
    for (i = 0; sumx < lower; sumx += bins[i++]) ;
    lower = i - (sumx - lower) / bins[i - 1];
    for (; sumx < upper; sumx += bins[i++]) ;
    upper = i - (sumx - upper) / bins[i - 1];
    if ((lower < 1.) || (upper > 255.)) { o.action = 1; return; }
    if ((upper - lower) < 40.) { o.action = 2; return; }
    o.action = 0;
    o.fwhm = binsize * (upper - lower) / 1.35;
    for (i = 0, sumx = 0; sumx < goodcnt / 2.; sumx += bins[i++]) ;
    med = i - (sumx - goodcnt / 2.) / bins[i - 1];
    o.lfwhm = binsize * (med - lower) * 2. / 1.35;
    o.median = bin1 + binsize * (med - 1.);
}

