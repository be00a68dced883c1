// zb_check.cu — `-c b`: which image of a region gets convolved (hotpants/alard.c:483-724 check_stamps,
// called from main.c:1026-1085 once per direction and region).
//
//   k_check_ksum    alard.c:521-548   every stamp fitted alone with a constant kernel: the leading
//                                     (nCompKer+1)^2 block of its own normal equations (build_matrix0 /
//                                     build_scprod0, here k_gram / k_rhs) through ludcmp + lubksb; the first
//                                     unknown is the stamp's kernel sum
//   (host)          alard.c:550-566   sigma_clip of the kernel sums, |ksum - mean| / sd per stamp: a few dozen
//                                     numbers per region, decided on the host like the stamp bookkeeping of main.c
//   trial fit       alard.c:574-616   build_matrix / build_scprod / ludcmp / lubksb over the stamps that passed:
//                                     the kernels of the real fit (zb_fit.cu), the others flagged ZbStamp::excl
//   k_check_kmean   alard.c:619       make_kernel(0, 0): the trial kernel's sum at region pixel (0,0)
//   k_stamp_sig     alard.c:637-650   getStampSig of every stamp of the trial fit (zb_fit.cu)
//   (host)          alard.c:651-721   sigma_clip of those figures, / kernel sum = the direction's merit
//
// Only -fom v is built (sig2 = sig3 = 0 then, alard.c:1082-1083).  Compiled with -fmad=false: the kernel sums
// feed a clip decision and the merits a comparison, both must match the reference's unfused doubles.
#include "zb_internal.h"
#include <math.h>
#include <stdlib.h>

// One CTA per stamp slot.  Right-looking LU with the reference's implicit-scaling pivot rule (ties: last row
// wins, alard.c:1504), every element receiving a_ij -= l_ik * u_kj for k ascending, product and difference rounded
// separately: the per-element operation sequence of the Crout loops, hence the same factors (as k_lu_* in zb_fit.cu).
// Forward/back substitution by one thread in the reference's order (n <= 161: a few thousand dependent operations).
__global__ void __launch_bounds__(256)
k_check_ksum(const ZbDev *__restrict__ D, const ZbStamp *__restrict__ stamps, const ZbFit *__restrict__ fit,
             const double *__restrict__ gram, const double *__restrict__ rhs, double *__restrict__ ksum)
{
    extern __shared__ __align__(16) unsigned char dyn[];
    const int slot = blockIdx.x;
    const ZbStamp &S = stamps[slot];
    if (S.nss <= 0 || !fit[S.region].active) return;
    const int nK = D->sz.n_comp_ker, nvec = nK + D->sz.n_bg_vectors, npair = nvec * (nvec + 1) / 2;
    const int n = nK + 1, ld = n + 1, tid = threadIdx.x, nt = blockDim.x;
    double *a = (double *)dyn;             // [n][ld]
    double *b = a + (size_t)n * ld;        // [n]
    double *vv = b + n;                    // [n]
    int *indx = (int *)(vv + n);           // [n]
    __shared__ int s_imax, s_sing;
    const double *g = gram + (long long)slot * npair;
    for (int e = tid; e < n * n; e += nt) {
        const int i = e / n, j = e - i * n, u = i > j ? i : j, v = i > j ? j : i;
        a[i * ld + j] = g[u * (u + 1) / 2 + v];
    }
    for (int i = tid; i < n; i += nt) b[i] = rhs[(long long)slot * nvec + i];
    if (tid == 0) s_sing = 0;
    __syncthreads();
    for (int i = tid; i < n; i += nt) {
        double big = 0.0;
        for (int j = 0; j < n; j++) { const double t = fabs(a[i * ld + j]); if (t > big) big = t; }
        if (big == 0.) s_sing = 1; else vv[i] = 1.0 / big;
    }
    __syncthreads();
    const bool singular = s_sing != 0;     // ludcmp returns before touching the matrix (alard.c:1480-1495); lubksb still runs
    for (int j = 0; j < n && !singular; j++) {
        if (tid == 0) {
            double big = 0.0; int imax = j;
            for (int i = j; i < n; i++) { const double dum = vv[i] * fabs(a[i * ld + j]); if (dum >= big) { big = dum; imax = i; } }
            s_imax = imax;
        }
        __syncthreads();
        const int imax = s_imax;
        if (imax != j) for (int k = tid; k < n; k += nt) { const double t = a[imax * ld + k]; a[imax * ld + k] = a[j * ld + k]; a[j * ld + k] = t; }
        __syncthreads();
        if (tid == 0) {
            if (imax != j) vv[imax] = vv[j];
            indx[j] = imax;
            if (a[j * ld + j] == 0.0) a[j * ld + j] = 1.0e-20;
        }
        __syncthreads();
        if (j != n - 1) {
            const double dum = 1.0 / a[j * ld + j];
            for (int i = j + 1 + tid; i < n; i += nt) a[i * ld + j] = __dmul_rn(a[i * ld + j], dum);
        }
        __syncthreads();
        const int m = n - j - 1;
        for (int e = tid; e < m * m; e += nt) {
            const int i = j + 1 + e / m, c = j + 1 + e % m;
            a[i * ld + c] = __dsub_rn(a[i * ld + c], __dmul_rn(a[i * ld + j], a[j * ld + c]));
        }
        __syncthreads();
    }
    if (tid == 0) {
        if (singular) for (int i = 0; i < n; i++) indx[i] = i;        // the reference reuses whatever its last ludcmp left there
        int ii = -1;
        for (int i = 0; i < n; i++) {
            const int ip = indx[i];
            double sum = b[ip]; b[ip] = b[i];
            if (ii >= 0) { for (int j = ii; j <= i - 1; j++) sum = __dsub_rn(sum, __dmul_rn(a[i * ld + j], b[j])); }
            else if (sum) ii = i;
            b[i] = sum;
        }
        for (int i = n - 1; i >= 0; i--) {
            double sum = b[i];
            for (int j = i + 1; j < n; j++) sum = __dsub_rn(sum, __dmul_rn(a[i * ld + j], b[j]));
            b[i] = sum / a[i * ld + i];
        }
        ksum[slot] = b[0];
    }
}

// make_kernel(0, 0, sol) (alard.c:1346-1385): coefficients of every basis at region pixel (0,0), the kernel
// pixel by pixel with its bases summed in order, then the pixels summed in order.  One CTA per region.
__global__ void __launch_bounds__(256)
k_check_kmean(const ZbDev *__restrict__ D, const ZbFit *__restrict__ fit, const double *__restrict__ sols,
              const double *__restrict__ kvec, double *__restrict__ kmean)
{
    extern __shared__ __align__(16) unsigned char dyn[];
    const int r = blockIdx.x;
    if (!fit[r].active) return;
    const ZbRegion R = D->reg[r];
    const int fw2 = D->sz.fw_kernel * D->sz.fw_kernel, nK = D->sz.n_comp_ker, nc2 = D->sz.n_comp, ko = D->p.ker_order;
    double *coef = (double *)dyn;          // [nK]
    double *ker = coef + nK;               // [fw2]
    const double *sol = sols + (long long)r * (D->sz.n_comp_total + 1);
    const double xf = (0 - 0.5 * R.rpx) / (0.5 * R.rpx), yf = (0 - 0.5 * R.rpy) / (0.5 * R.rpy);
    for (int i1 = threadIdx.x; i1 < nK; i1 += blockDim.x) {
        if (i1 == 0) { coef[0] = sol[1]; continue; }
        int k = 2 + (i1 - 1) * nc2;
        double c = 0.0, ax = 1.0;
        for (int ix = 0; ix <= ko; ix++) {
            double ay = 1.0;
            for (int iy = 0; iy <= ko - ix; iy++) { c = __dadd_rn(c, __dmul_rn(__dmul_rn(sol[k++], ax), ay)); ay = __dmul_rn(ay, yf); }
            ax = __dmul_rn(ax, xf);
        }
        coef[i1] = c;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < fw2; i += blockDim.x) {
        double acc = 0.0;
        for (int i1 = 0; i1 < nK; i1++) acc = __dadd_rn(acc, __dmul_rn(coef[i1], kvec[(long long)i1 * fw2 + i]));
        ker[i] = acc;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0;
        for (int i = 0; i < fw2; i++) s = __dadd_rn(s, ker[i]);
        kmean[r] = s;
    }
}

// functions.c:1037-1107 sigma_clip on a short host array
static void sigma_clip_host(const float *data, int count, double *mean, double *stdev, int maxiter, float statSig)
{
    if (count == 0) { *mean = 0; *stdev = ZB_MAXVAL; return; }
    char *rej = (char *)calloc((size_t)count, 1);
    int cnt = 0, ncnt = count, iter = 0;
    while (ncnt != cnt && iter < maxiter) {
        cnt = ncnt;
        *mean = 0; *stdev = 0;
        for (int i = 0; i < count; i++) if (!rej[i]) { float d = data[i]; *mean += d; *stdev += d * d; }
        if (ncnt > 0) *mean /= ncnt; else { *mean = 0; *stdev = ZB_MAXVAL; break; }
        if (ncnt > 1) { *stdev = *stdev - ncnt * (*mean) * (*mean); *stdev = sqrt(*stdev / (double)(ncnt - 1)); }
        else { *stdev = ZB_MAXVAL; break; }
        ncnt = 0;
        const double istd = 1. / (*stdev);
        for (int i = 0; i < count; i++)
            if (!rej[i]) { if ((fabs(data[i] - (*mean)) * istd) > statSig) rej[i] = 1; else ncnt++; }
        iter++;
    }
    free(rej);
}

#define CHECK_RC(x) do { int rc_ = (x); if (rc_ != ZB_OK) return rc_; } while (0)
#define SYNC(what) do { cudaError_t e_ = zb_stream_wait(ctx, st); if (e_ != cudaSuccess) return zb_cuda_fail(ctx, e_, what, __FILE__, __LINE__); } while (0)

// check_stamps for every region at once, one direction.  On entry the direction's stamps sit in their slots
// (h_stamps == d_stamps, sscnt = 0, filled by zb_launch_fill) and h_fit[r].n_stamps counts them.  merit[r] is
// left alone for regions without stamps in this direction (the caller presets 666, see zb_host.cu).
int zb_check_stamps(zb_ctx *ctx, const float *sci, const float *ref, const float *sci_rms, const float *ref_rms,
                    double *merit, cudaStream_t st)
{
    const zb_sizes &sz = ctx->sz;
    const ZbDev *H = ctx->h_dev;
    const zb_params *p = &ctx->p;
    const int nR = sz.n_regions, n = sz.n_comp_ker + 1;
    bool any = false;
    for (int r = 0; r < nR; r++) {
        ZbFit &F = ctx->h_fit[r];
        F.active = F.n_stamps > 0; F.singular = 0;
        any |= F.active != 0;
    }
    if (!any) return ZB_OK;
    zb_copy_small(ctx, ctx->d_fit, ctx->h_fit, (size_t)nR * sizeof(ZbFit), st);
    const size_t sm = ((size_t)n * (n + 1) + 2 * (size_t)n) * sizeof(double) + (size_t)n * sizeof(int);
    if (sm > ZB_SMEM_OPTIN) { snprintf(ctx->err, sizeof ctx->err, "-c b: %d kernel components too many for the per-stamp solve", n - 1); return ZB_ERR_UNSUPPORTED; }
    cudaFuncSetAttribute(k_check_ksum, cudaFuncAttributeMaxDynamicSharedMemorySize, ZB_SMEM_OPTIN);
    k_check_ksum<<<ctx->n_slots, 256, sm, st>>>(ctx->d_dev, ctx->d_stamps, ctx->d_fit, ctx->d_gram, ctx->d_rhs, ctx->d_check); ZB_DBG("k_check_ksum");
    ctx->launches++;
    zb_copy_small(ctx, ctx->h_check, ctx->d_check, (size_t)ctx->n_slots * sizeof(double), st);
    SYNC("check_stamps kernel sums");

    // kernel-sum clip (alard.c:550-566) and the stamps of the trial fit (:590-606)
    float *ks = (float *)malloc((size_t)(sz.n_stamps + 1) * sizeof(float));
    if (!ks) return ZB_ERR_NOMEM;
    for (int r = 0; r < nR; r++) {
        const ZbFit &F = ctx->h_fit[r];
        if (!F.active) continue;
        const int base = H->reg[r].stamp_base;
        for (int k = 0; k < F.n_stamps; k++) ks[k] = (float)ctx->h_check[base + k];
        double kmean = 0, kstdev = 0;
        sigma_clip_host(ks, F.n_stamps, &kmean, &kstdev, 10, p->stat_sig);
        for (int k = 0; k < F.n_stamps; k++) {
            const double dif = fabs((ctx->h_check[base + k] - kmean) / kstdev);
            ctx->h_stamps[base + k].excl = !(dif < p->ker_sig_reject);
        }
    }
    zb_copy_small(ctx, ctx->d_stamps, ctx->h_stamps, (size_t)ctx->n_slots * sizeof(ZbStamp), st);

    // trial global fit over the stamps that passed, their figures of merit, the trial kernel's sum at (0,0)
    CHECK_RC(zb_launch_fit_solve(ctx, st));
    CHECK_RC(zb_launch_stamp_sig(ctx, sci, ref, sci_rms, ref_rms, 0, ctx->n_slots, st));
    const size_t smk = ((size_t)sz.n_comp_ker + (size_t)sz.fw_kernel * sz.fw_kernel) * sizeof(double);
    cudaFuncSetAttribute(k_check_kmean, cudaFuncAttributeMaxDynamicSharedMemorySize, ZB_SMEM_OPTIN);
    k_check_kmean<<<nR, 256, smk, st>>>(ctx->d_dev, ctx->d_fit, ctx->d_sol, ctx->d_kvec, ctx->d_check + ctx->n_slots); ZB_DBG("k_check_kmean");
    ctx->launches++;
    zb_copy_small(ctx, ctx->h_stamps, ctx->d_stamps, (size_t)ctx->n_slots * sizeof(ZbStamp), st);
    zb_copy_small(ctx, ctx->h_check + ctx->n_slots, ctx->d_check + ctx->n_slots, (size_t)nR * sizeof(double), st);
    SYNC("check_stamps trial fit");

    static int dbg = -1;
    if (dbg < 0) { const char *e = getenv("ZB_CHECK_DEBUG"); dbg = (e && e[0] == '1') ? 1 : 0; }
    for (int r = 0; r < nR; r++) {
        const ZbFit &F = ctx->h_fit[r];
        if (!F.active) continue;
        const int base = H->reg[r].stamp_base;
        const double kmean = ctx->h_check[ctx->n_slots + r];
        int ntest = 0, mcnt1 = 0;
        if (dbg) {
            fprintf(stderr, "CHK region %d nS %d kmean %.17g\n", r, F.n_stamps, kmean);
            for (int k = 0; k < F.n_stamps; k++)
                fprintf(stderr, "CHK  stamp %d ksum %.17g excl %d sig1 %.17g\n", k, ctx->h_check[base + k], ctx->h_stamps[base + k].excl, ctx->h_stamps[base + k].sig1);
        }
        for (int k = 0; k < F.n_stamps; k++) {
            ZbStamp &S = ctx->h_stamps[base + k];
            if (!S.excl) {
                ntest++;
                if (S.sig1 != -1 && S.sig1 <= ZB_MAXVAL) ks[mcnt1++] = (float)S.sig1;
            }
            S.excl = 0; S.sig1 = 0;
        }
        double merit1 = 0, sd1 = 0;
        sigma_clip_host(ks, mcnt1, &merit1, &sd1, 10, p->stat_sig);
        // -fom v (alard.c:693-703): the variance merit if any stamp has one, else the (all-zero) sd merit, else 666
        if (mcnt1 > 0) merit[r] = merit1 / kmean;
        else if (ntest > 0) merit[r] = 0.0 / kmean;
        else merit[r] = 666;
    }
    free(ks);
    return ZB_OK;
}
