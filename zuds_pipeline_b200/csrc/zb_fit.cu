// zb_fit.cu — the least-squares side of the path, batched over all stamps / regions of a pair.
//
//   k_fill_conv     alard.c:230-279  xy_conv_stamp   separable basis convolution of a sub-stamp
//   k_fill_post     alard.c:81-85,273-276 (ren), functions.c:390-443 cutSStamp, alard.c:100-124 bg vectors
//   k_gram          alard.c:384-434  build_matrix0 (+ the bg dot products of build_matrix :925-953)
//   k_rhs           alard.c:436-481  build_scprod0 (+ bg right-hand sides of build_scprod :1008-1019)
//   k_wxy / k_build_matrix   alard.c:843-965 build_matrix, :967-1025 build_scprod
//   k_lu_solve      alard.c:1465-1552 ludcmp + lubksb
//   k_stamp_sig     alard.c:1062-1140 getStampSig (+ :1417-1463 make_model, :1387-1415 get_background)
//   k_clip          alard.c:1143-1247 check_again
//
// All of it is FP64 and compiled with -fmad=false.  Every reduction that feeds the solution is
// evaluated per element in the reference's own term order (one thread owns one output element
// and walks its terms sequentially; parallelism comes from the thousands of elements), so the
// fitted kernel coefficients come out bit-identical to the reference, not merely close.
#include "zb_internal.h"

// ------------------------------------------------------------------ fillStamp

#define FILL_ILP 4
__global__ void __launch_bounds__(256)
k_fill_conv(const ZbDev *__restrict__ D, const float *__restrict__ img, const ZbStamp *__restrict__ stamps,
            const int *__restrict__ jobs, const double *__restrict__ fxs, const double *__restrict__ fys,
            double *__restrict__ vec)
{
    extern __shared__ __align__(16) unsigned char dyn[];
    const int slot = jobs[blockIdx.x], n = blockIdx.y;
    const ZbStamp &S = stamps[slot];
    if (S.sscnt >= S.nss) return;
    const ZbRegion R = D->reg[S.region];
    const int hw = D->p.hw_kernel, hwss = D->p.hw_ksstamp, fw = D->sz.fw_kernel, fwss = D->sz.fw_ksstamp;
    const int sw = fwss + fw - 1;                     // sub_width
    const int xi = S.xss[S.sscnt], yi = S.yss[S.sscnt];
    // The float -> double conversion pipe (16 lanes/clk/SM) is what bounds this kernel, so every value is
    // widened ONCE when it is staged instead of once per tap: the patch as it is loaded, the y-pass result
    // when it is stored.  What remains per tap is the rounding of the running y-sum to float and back.
    double *fx = (double *)dyn, *fy = fx + fw;
    double *patch = fy + fw;                          // [sw][sw]   (double)pixel
    double *temp = patch + sw * sw;                   // [fwss][sw] (double)(float) y-pass sums
    const int tid = threadIdx.x, nt = blockDim.x;
    for (int k = tid; k < fw; k += nt) { fx[k] = fxs[k + n * fw]; fy[k] = fys[k + n * fw]; }
    const int px0 = xi - hwss - hw, py0 = yi - hwss - hw;
    for (int py = tid / sw, px = tid - (tid / sw) * sw, k = tid; k < sw * sw; k += nt) {
        patch[k] = (double)img[(long long)(R.by0 + py0 + py) * D->nx + (R.bx0 + px0 + px)];
        px += nt; while (px >= sw) { px -= sw; py++; }
    }
    __syncthreads();
    // y pass: float accumulator, rounded after every tap (the reference's scratch is float)
    // FILL_ILP outputs per thread advance together: the per-tap chain (DMUL, DADD, round to float, widen) is
    // ~100 cycles of dependent latency, and the staged doubles leave room for only a few CTAs per SM
    for (int k0 = tid; k0 < sw * fwss; k0 += FILL_ILP * nt) {
        double t[FILL_ILP]; int off[FILL_ILP];
#pragma unroll
        for (int u = 0; u < FILL_ILP; u++) {
            const int k = k0 + u * nt, r = k / sw, c = k - r * sw;          // r: row of the sub-stamp, c: column incl. halo
            t[u] = 0.0; off[u] = k < sw * fwss ? c + sw * r : 0;
        }
        for (int q = 0; q < fw; q++) {                                      // q = yc + hw
            const double f = fy[fw - 1 - q];
#pragma unroll
            for (int u = 0; u < FILL_ILP; u++) t[u] = (double)(float)(t[u] + patch[off[u] + sw * q] * f);
        }
#pragma unroll
        for (int u = 0; u < FILL_ILP; u++) if (k0 + u * nt < sw * fwss) temp[k0 + u * nt] = t[u];
    }
    __syncthreads();
    const int npix = fwss * fwss;
    double *out = vec + (long long)slot * (D->sz.n_comp_ker + D->sz.n_bg_vectors) * npix + (long long)n * npix;
    for (int k0 = tid; k0 < npix; k0 += FILL_ILP * nt) {
        double acc[FILL_ILP]; int off[FILL_ILP];
#pragma unroll
        for (int u = 0; u < FILL_ILP; u++) {
            const int k = k0 + u * nt, r = k / fwss, c = k - r * fwss;
            acc[u] = 0.0; off[u] = k < npix ? c + sw * r : 0;
        }
        for (int q = 0; q < fw; q++) {                                      // q = xc + hw
            const double f = fx[fw - 1 - q];
#pragma unroll
            for (int u = 0; u < FILL_ILP; u++) acc[u] += temp[off[u] + q] * f;
        }
#pragma unroll
        for (int u = 0; u < FILL_ILP; u++) if (k0 + u * nt < npix) out[k0 + u * nt] = acc[u];
    }
}

__global__ void __launch_bounds__(256)
k_fill_post(const ZbDev *__restrict__ D, const float *__restrict__ refimg, const ZbStamp *__restrict__ stamps,
            const int *__restrict__ jobs, const int *__restrict__ ren, double *__restrict__ vec, double *__restrict__ kref)
{
    const int slot = jobs[blockIdx.x];
    const ZbStamp &S = stamps[slot];
    if (S.sscnt >= S.nss) return;
    const ZbRegion R = D->reg[S.region];
    const int hwss = D->p.hw_ksstamp, fwss = D->sz.fw_ksstamp, npix = fwss * fwss;
    const int nK = D->sz.n_comp_ker, nvec = nK + D->sz.n_bg_vectors, bgo = D->p.bg_order;
    const int xi = S.xss[S.sscnt], yi = S.yss[S.sscnt];
    double *v = vec + (long long)slot * nvec * npix;
    double *kr = kref + (long long)slot * npix;
    const float rPixX2 = 0.5f * R.rpx, rPixY2 = 0.5f * R.rpy;
    for (int k = threadIdx.x; k < npix; k += blockDim.x) {
        int r = k / fwss, c = k - r * fwss;
        int i = xi - hwss + c, j = yi - hwss + r;
        double v0 = v[k];
        for (int n = 1; n < nK; n++)
            if (ren[n]) v[(long long)n * npix + k] -= v0;
        kr[k] = (double)refimg[(long long)(R.by0 + j) * D->nx + (R.bx0 + i)];
        // background monomials; the normalised coordinate is formed in FLOAT (alard.c:60-63,105,109)
        double xf = (double)(((float)i - rPixX2) / rPixX2);
        double yf = (double)(((float)j - rPixY2) / rPixY2);
        double ax = 1.0; int nv = nK;
        for (int a = 0; a <= bgo; a++) {
            double ay = 1.0;
            for (int b = 0; b <= bgo - a; b++) { v[(long long)nv * npix + k] = ax * ay; ay *= yf; ++nv; }
            ax *= xf;
        }
    }
}

// Gram matrix of the stamp's vectors, lower triangle: G[i(i+1)/2+j] = sum_k v_i[k] v_j[k], k ascending
// (alard.c:405-431 sums the pixels in storage order).  One CTA per stamp; a thread owns a 4x4 tile of
// (i, j) pairs and keeps its 16 running sums in registers; products and sums are separate roundings
// (-fmad=false) so each G element sees exactly the reference's operation sequence.  Chunks of GRAM_KC
// pixels of all vectors are staged in shared memory, the next chunk is prefetched into registers
// while the current one is consumed.
#define GRAM_KC 32
#define GRAM_T 4
#define GRAM_NT 128
#define GRAM_PF 16            // prefetch registers per thread: GRAM_NT*GRAM_PF >= GRAM_KC*nvec_pad (nvec <= 64)
__global__ void __launch_bounds__(GRAM_NT)
k_gram(const ZbDev *__restrict__ D, const ZbStamp *__restrict__ stamps, const int *__restrict__ jobs,
       const double *__restrict__ vec, double *__restrict__ gram)
{
    extern __shared__ __align__(16) unsigned char dyn[];
    const int slot = jobs[blockIdx.x];
    const ZbStamp &S = stamps[slot];
    if (S.sscnt >= S.nss) return;
    const int fwss = D->sz.fw_ksstamp, npix = fwss * fwss;
    const int nvec = D->sz.n_comp_ker + D->sz.n_bg_vectors;
    const int nt4 = (nvec + GRAM_T - 1) / GRAM_T, ld = nt4 * GRAM_T;       // padded vector count
    const int ntile = nt4 * (nt4 + 1) / 2;
    const double *v = vec + (long long)slot * nvec * npix;
    double *g = gram + (long long)slot * (nvec * (nvec + 1) / 2);
    double *s = (double *)dyn;                         // [GRAM_KC][ld]
    const int tid = threadIdx.x;
    for (int tbase = 0; tbase < ntile; tbase += GRAM_NT) {
        const int tile = tbase + tid;
        int ti = -1, tj = 0;
        if (tile < ntile) {
            ti = (int)((sqrt(8.0 * tile + 1.0) - 1.0) * 0.5);
            while (ti * (ti + 1) / 2 > tile) ti--;
            while ((ti + 1) * (ti + 2) / 2 <= tile) ti++;
            tj = tile - ti * (ti + 1) / 2;
        }
        double acc[GRAM_T][GRAM_T];
#pragma unroll
        for (int a = 0; a < GRAM_T; a++)
#pragma unroll
            for (int c = 0; c < GRAM_T; c++) acc[a][c] = 0.0;
        double pf[GRAM_PF];
        const int nelem = GRAM_KC * ld;
        // prefetch chunk 0
#pragma unroll
        for (int q = 0; q < GRAM_PF; q++) {
            int e = q * GRAM_NT + tid; pf[q] = 0.0;
            if (e < nelem) { int n = e / GRAM_KC, kk = e - n * GRAM_KC; if (n < nvec && kk < npix) pf[q] = v[(long long)n * npix + kk]; }
        }
        for (int k0 = 0; k0 < npix; k0 += GRAM_KC) {
            __syncthreads();
#pragma unroll
            for (int q = 0; q < GRAM_PF; q++) {
                int e = q * GRAM_NT + tid;
                if (e < nelem) { int n = e / GRAM_KC, kk = e - n * GRAM_KC; s[kk * ld + n] = pf[q]; }
            }
            __syncthreads();
            const int k1 = k0 + GRAM_KC;
            if (k1 < npix) {
#pragma unroll
                for (int q = 0; q < GRAM_PF; q++) {
                    int e = q * GRAM_NT + tid; pf[q] = 0.0;
                    if (e < nelem) { int n = e / GRAM_KC, kk = e - n * GRAM_KC; if (n < nvec && k1 + kk < npix) pf[q] = v[(long long)n * npix + k1 + kk]; }
                }
            }
            if (ti >= 0) {
                const int kc = min(GRAM_KC, npix - k0);
                for (int kk = 0; kk < kc; kk++) {
                    const double2 a01 = *(const double2 *)&s[kk * ld + ti * GRAM_T], a23 = *(const double2 *)&s[kk * ld + ti * GRAM_T + 2];
                    const double2 b01 = *(const double2 *)&s[kk * ld + tj * GRAM_T], b23 = *(const double2 *)&s[kk * ld + tj * GRAM_T + 2];
                    const double av[4] = {a01.x, a01.y, a23.x, a23.y}, bv[4] = {b01.x, b01.y, b23.x, b23.y};
#pragma unroll
                    for (int a = 0; a < GRAM_T; a++)
#pragma unroll
                        for (int c = 0; c < GRAM_T; c++) acc[a][c] += av[a] * bv[c];
                }
            }
        }
        if (ti >= 0) {
#pragma unroll
            for (int a = 0; a < GRAM_T; a++)
#pragma unroll
                for (int c = 0; c < GRAM_T; c++) {
                    const int i = ti * GRAM_T + a, j = tj * GRAM_T + c;
                    if (i < nvec && j <= i) g[i * (i + 1) / 2 + j] = acc[a][c];
                }
        }
    }
}

// rhs[i] = sum over the sub-stamp of v_i * imRef with x outer / y inner (alard.c:459-466): one chain of
// fwss^2 dependent additions per vector.  A warp owns one vector: all lanes fetch 4 columns at a time
// (32-byte segments of every row), form the products and park them in shared memory in chain order;
// lane 0 then adds them in that order while the warp already fetches the next columns.
#define RHS_WARPS 8
#define RHS_COLS 4
__global__ void __launch_bounds__(RHS_WARPS * 32)
k_rhs(const ZbDev *__restrict__ D, const ZbStamp *__restrict__ stamps, const int *__restrict__ jobs,
      const double *__restrict__ vec, const double *__restrict__ kref, double *__restrict__ rhs)
{
    extern __shared__ __align__(16) unsigned char dyn[];
    const int slot = jobs[blockIdx.x];
    const ZbStamp &S = stamps[slot];
    if (S.sscnt >= S.nss) return;
    const int fwss = D->sz.fw_ksstamp, npix = fwss * fwss;
    const int nvec = D->sz.n_comp_ker + D->sz.n_bg_vectors;
    const int chunk = RHS_COLS * fwss;
    double *kr = (double *)dyn;                          // [npix], row-major like kref
    double *bufs = kr + npix;                            // [RHS_WARPS][2][chunk]
    const double *krg = kref + (long long)slot * npix;
    for (int k = threadIdx.x; k < npix; k += blockDim.x) kr[k] = krg[k];
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int i = blockIdx.y * RHS_WARPS + warp;
    if (i >= nvec) return;
    const double *v = vec + ((long long)slot * nvec + i) * npix;
    double *buf = bufs + (size_t)warp * 2 * chunk;
    const int per = (chunk + 31) / 32;                   // elements per lane per chunk (<= 8 for fwss <= 63)
    double pr[8];
    // element e of a chunk starting at column x0: row yc = e / RHS_COLS, column x0 + e % RHS_COLS
#define RHS_FETCH(x0) \
    _Pragma("unroll") for (int q = 0; q < 8; q++) { \
        int e = q * 32 + lane; pr[q] = 0.0; \
        if (q < per && e < chunk) { int yc = e / RHS_COLS, xc = (x0) + (e - yc * RHS_COLS); \
            if (xc < fwss) pr[q] = v[xc + fwss * yc] * kr[xc + fwss * yc]; } }
    RHS_FETCH(0)
    double p0 = 0.0;
    int par = 0;
    for (int x0 = 0; x0 < fwss; x0 += RHS_COLS, par ^= 1) {
        double *b = buf + par * chunk;
#pragma unroll
        for (int q = 0; q < 8; q++) {
            int e = q * 32 + lane;
            if (q < per && e < chunk) { int yc = e / RHS_COLS, dx = e - yc * RHS_COLS; b[dx * fwss + yc] = pr[q]; }   // chain order
        }
        __syncwarp();
        if (x0 + RHS_COLS < fwss) { RHS_FETCH(x0 + RHS_COLS) }
        if (lane == 0) {
            const int cnt = min(RHS_COLS, fwss - x0) * fwss;
            for (int t = 0; t < cnt; t++) p0 += b[t];
        }
    }
#undef RHS_FETCH
    if (lane == 0) rhs[(long long)slot * nvec + i] = p0;
}

int zb_launch_fill(zb_ctx *ctx, const float *conv_img, const float *ref_img, const int *d_jobs, int n_jobs, cudaStream_t st)
{
    if (n_jobs <= 0) return ZB_OK;
    const zb_sizes &sz = ctx->sz;
    const int fw = sz.fw_kernel, fwss = sz.fw_ksstamp, sw = fwss + fw - 1, npix = fwss * fwss;
    const int nvec = sz.n_comp_ker + sz.n_bg_vectors;
    size_t sm1 = (2 * fw + (size_t)sw * sw + (size_t)sw * fwss) * sizeof(double);
    const int ldv = ((nvec + GRAM_T - 1) / GRAM_T) * GRAM_T;
    size_t sm3 = (size_t)GRAM_KC * ldv * sizeof(double);
    size_t sm4 = ((size_t)npix + (size_t)RHS_WARPS * 2 * RHS_COLS * fwss) * sizeof(double);
    if (GRAM_KC * ldv > GRAM_NT * GRAM_PF || RHS_COLS * fwss > 256) {
        snprintf(ctx->err, sizeof ctx->err, "%d basis vectors / sub-stamp width %d exceed the fit kernels' limits", nvec, fwss);
        return ZB_ERR_UNSUPPORTED;
    }
    if (sm1 > ZB_SMEM_OPTIN || sm3 > ZB_SMEM_OPTIN || sm4 > ZB_SMEM_OPTIN) {
        snprintf(ctx->err, sizeof ctx->err, "sub-stamp of %d pixels too large for shared memory", fwss);
        return ZB_ERR_UNSUPPORTED;
    }
    cudaFuncSetAttribute(k_fill_conv, cudaFuncAttributeMaxDynamicSharedMemorySize, ZB_SMEM_OPTIN);
    cudaFuncSetAttribute(k_gram, cudaFuncAttributeMaxDynamicSharedMemorySize, ZB_SMEM_OPTIN);
    cudaFuncSetAttribute(k_rhs, cudaFuncAttributeMaxDynamicSharedMemorySize, ZB_SMEM_OPTIN);
    k_fill_conv<<<dim3(n_jobs, sz.n_comp_ker), 256, sm1, st>>>(ctx->d_dev, conv_img, ctx->d_stamps, d_jobs, ctx->d_fx, ctx->d_fy, ctx->d_vec); ZB_DBG("k_fill_conv");
    k_fill_post<<<n_jobs, 256, 0, st>>>(ctx->d_dev, ref_img, ctx->d_stamps, d_jobs, ctx->d_ren, ctx->d_vec, ctx->d_kref); ZB_DBG("k_fill_post");
    k_gram<<<n_jobs, GRAM_NT, sm3, st>>>(ctx->d_dev, ctx->d_stamps, d_jobs, ctx->d_vec, ctx->d_gram); ZB_DBG("k_gram");
    k_rhs<<<dim3(n_jobs, (nvec + RHS_WARPS - 1) / RHS_WARPS), RHS_WARPS * 32, sm4, st>>>(ctx->d_dev, ctx->d_stamps, d_jobs, ctx->d_vec, ctx->d_kref, ctx->d_rhs); ZB_DBG("k_rhs");
    ctx->launches += 4;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return zb_cuda_fail(ctx, e, "fill kernels", __FILE__, __LINE__);
    return ZB_OK;
}

// --------------------------------------------------------------- global system

// spatial weights of each stamp at its current sub-stamp centre (alard.c:886-903)
__global__ void k_wxy(const ZbDev *__restrict__ D, const ZbStamp *__restrict__ stamps, int n_slots, double *__restrict__ wxy)
{
    int slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= n_slots) return;
    const ZbStamp &S = stamps[slot];
    if (S.nss <= 0 || S.sscnt >= S.nss) return;
    const ZbRegion R = D->reg[S.region];
    const float rPixX2 = 0.5f * R.rpx, rPixY2 = 0.5f * R.rpy;
    double fx = (double)(((float)S.xss[S.sscnt] - rPixX2) / rPixX2);
    double fy = (double)(((float)S.yss[S.sscnt] - rPixY2) / rPixY2);
    double *w = wxy + (long long)slot * D->sz.n_comp;
    int k = 0; double a1 = 1.0;
    for (int a = 0; a <= D->p.ker_order; a++) {
        double a2 = 1.0;
        for (int b = 0; b <= D->p.ker_order - a; b++) { w[k++] = a1 * a2; a2 *= fy; }
        a1 *= fx;
    }
}

// per-stamp normal matrices are packed lower triangles: gram(u, v), u >= v, sits at u (u + 1) / 2 + v

// The bulk of the normal matrix: elements (I, J) = (1 + i1*nc2 + wi, 1 + j1*nc2 + wj) with j1 <= i1, i.e.
// sum over stamps of w[wi] * w[wj] * gram(i1+1, j1+1) (alard.c:905-913).  One thread per element, every kind of element through one
// switch, was ~20 instructions per term (three loads, the stamp's state, a switch); at ko = 4 that made this step the largest
// consumer of SM time of a pair (n = 722: 2.8 of 7.3 ms, profiles/r02_launches_c1b_summary.txt, r02_fit_ab_s25.txt).  Here a thread owns one (wi, wj) and BM_R pairs
// (i1, j1): per stamp it forms w[wi] * w[wj] once and adds its product with BM_R gram entries, still one element per
// accumulator, stamps ascending, (w * w) * g then the add, each rounded: the same sums bit for bit.
#define BM_R 8
#define BM_CS 16                                     // stamps staged per round
#define BM_GROUPS(PP) ((PP) >= 256 ? 1 : (256 / (PP) > 32 ? 32 : 256 / (PP)))   // at most 32: the staged record stays small at ko <= 1
__global__ void __launch_bounds__(256)
k_build_matrix_main(const ZbDev *__restrict__ D, const ZbStamp *__restrict__ stamps, const ZbFit *__restrict__ fit,
                    const double *__restrict__ gram, const double *__restrict__ wxy, double *__restrict__ mat)
{
    extern __shared__ __align__(16) unsigned char bm_dyn[];
    __shared__ int s_nact;
    const int r = blockIdx.y;
    if (!fit[r].active) return;
    const int n = D->sz.mat_size, nK = D->sz.n_comp_ker, nc2 = D->sz.n_comp, PP = nc2 * nc2;
    const int nvec = nK + D->sz.n_bg_vectors, npair = nvec * (nvec + 1) / 2;
    const int nb1 = nK - 1, nuv = nb1 * (nb1 + 1) / 2;
    const int base = D->reg[r].stamp_base, nS = fit[r].n_stamps, tid = threadIdx.x;
    const int G = BM_GROUPS(PP);                     // (wi, wj) groups side by side in a CTA
    const int NE = G * BM_R, L = NE + nc2;           // staged per stamp: the CTA's gram entries, then the stamp's weights
    double *buf = (double *)bm_dyn;                  // [BM_CS][L]
    int *s_act = (int *)(buf + (size_t)BM_CS * L);   // slots of the stamps that enter the sums, ascending
    int *s_goff = s_act + D->sz.n_stamps;            // [NE] packed-gram offset of every (i1, j1) pair of the CTA, -1 beyond the last
    if (tid == 0) {
        int c = 0;
        for (int k = 0; k < nS; k++) { const ZbStamp &S = stamps[base + k]; if (!(S.sscnt >= S.nss || S.excl)) s_act[c++] = base + k; }
        s_nact = c;
    }
    const int qc = blockIdx.x * NE;                  // first (i1, j1) pair of the CTA, triangular order
    for (int e = tid; e < NE; e += 256) {
        const int q = qc + e;
        int off = -1;
        if (q < nuv) {
            int i1 = (int)((sqrt(8.0 * q + 1.0) - 1.0) * 0.5);
            while ((i1 + 1) * (i1 + 2) / 2 <= q) i1++;
            while (i1 * (i1 + 1) / 2 > q) i1--;
            const int j1 = q - i1 * (i1 + 1) / 2, u = i1 + 1, v = j1 + 1;
            off = u * (u + 1) / 2 + v;
        }
        s_goff[e] = off;
    }
    __syncthreads();
    const int nA = s_nact;
    const int grp = PP >= 256 ? 0 : tid / PP;
    const bool worker = grp < G && qc + grp * BM_R < nuv;
    double *M = mat + (long long)r * n * n;
    for (int p0 = 0; p0 < PP; p0 += 256) {           // one round unless ko >= 5 (more than 256 (wi, wj) pairs)
        const int p = PP >= 256 ? p0 + tid : tid - grp * PP;
        const bool mine = worker && p < PP;
        const int wi = mine ? p / nc2 : 0, wj = mine ? p - wi * nc2 : 0;
        double acc[BM_R];
#pragma unroll
        for (int e = 0; e < BM_R; e++) acc[e] = 0.0;
        for (int c0 = 0; c0 < nA; c0 += BM_CS) {
            const int ns = min(BM_CS, nA - c0);
            __syncthreads();                         // the previous round's values have been consumed
            for (int idx = tid; idx < ns * L; idx += 256) {
                const int sidx = idx / L, o = idx - sidx * L, slot = s_act[c0 + sidx];
                double v = 0.0;
                if (o < NE) { const int off = s_goff[o]; if (off >= 0) v = gram[(long long)slot * npair + off]; }
                else v = wxy[(long long)slot * nc2 + (o - NE)];
                buf[idx] = v;
            }
            __syncthreads();
            if (mine)
                for (int k = 0; k < ns; k++) {
                    const double *b = buf + (size_t)k * L;
                    const double ww = __dmul_rn(b[NE + wi], b[NE + wj]);
#pragma unroll
                    for (int e = 0; e < BM_R; e++) acc[e] = __dadd_rn(acc[e], __dmul_rn(ww, b[grp * BM_R + e]));
                }
        }
        if (mine) {
#pragma unroll
            for (int e = 0; e < BM_R; e++) {
                const int off = s_goff[grp * BM_R + e];
                if (off < 0) continue;
                int u = (int)((sqrt(8.0 * off + 1.0) - 1.0) * 0.5);
                while ((u + 1) * (u + 2) / 2 <= off) u++;
                while (u * (u + 1) / 2 > off) u--;
                const int v = off - u * (u + 1) / 2;            // off = u(u+1)/2 + v, 1 <= v <= u
                const int I = 1 + (u - 1) * nc2 + wi, J = 1 + (v - 1) * nc2 + wj;
                if (J <= I) { M[(long long)I * n + J] = acc[e]; M[(long long)J * n + I] = acc[e]; }
            }
        }
    }
}

// Everything else: column 0, the background rows and the right-hand side (alard.c:915-953, 967-1025).
// A term is value or w * value with value one of gram(u, 0), rhs[u], gram(nK + ibg, v): those are staged per round
// of BM_CS stamps together with the stamp's weights, so an element's sum (stamps ascending, one add per stamp) reads
// shared memory only; with one global load chain per stamp and element this step took 94 us at 100 stamps per region
// and 0.9 ms at 1024 (profiles/r02_launches_c1b_after_summary.txt).
__global__ void __launch_bounds__(256)
k_build_matrix_edges(const ZbDev *__restrict__ D, const ZbStamp *__restrict__ stamps, const ZbFit *__restrict__ fit,
                     const double *__restrict__ gram, const double *__restrict__ rhs, const double *__restrict__ wxy,
                     double *__restrict__ mat, double *__restrict__ sol, int cs)
{
    extern __shared__ __align__(16) unsigned char bm_dyn[];
    __shared__ int s_nact;
    const int r = blockIdx.y;
    if (!fit[r].active) return;
    const int n = D->sz.mat_size, nK = D->sz.n_comp_ker, nc2 = D->sz.n_comp, ncomp = (nK - 1) * nc2;
    const int nbg = D->sz.n_bg_vectors, nvec = nK + nbg, npair = nvec * (nvec + 1) / 2;
    const int base = D->reg[r].stamp_base, nS = fit[r].n_stamps, tid = threadIdx.x;
    const int OW = (2 + nbg) * nvec, L = OW + nc2;   // per stamp: gram(u,0) | rhs[u] | gram(nK+ibg, v) rows | weights
    double *buf = (double *)bm_dyn;                  // [cs][L], cs stamps staged per round
    int *s_act = (int *)(buf + (size_t)cs * L);
    if (tid == 0) {
        int c = 0;
        for (int k = 0; k < nS; k++) { const ZbStamp &S = stamps[base + k]; if (!(S.sscnt >= S.nss || S.excl)) s_act[c++] = base + k; }
        s_nact = c;
    }
    // this thread's element: where its value sits in a staged record (ia), which weight multiplies it (iw, -1: none)
    const int idx = blockIdx.x * 256 + tid;
    int I = -1, J = -1, ia = 0, iw = -1;
    if (idx < n) {                                   // column 0
        I = idx; J = 0;
        if (I == 0) ia = 0;
        else if (I <= ncomp) { const int i1 = (I - 1) / nc2; iw = (I - 1) - i1 * nc2; ia = i1 + 1; }
        else ia = nK + (I - ncomp - 1);
    } else if (idx < 2 * n) {                        // right-hand side
        I = idx - n; J = n;
        if (I == 0) ia = nvec;
        else if (I <= ncomp) { const int i1 = (I - 1) / nc2; iw = (I - 1) - i1 * nc2; ia = nvec + i1 + 1; }
        else ia = nvec + nK + (I - ncomp - 1);
    } else {                                         // background rows, J = 1 .. I
        const int e = idx - 2 * n, ibg = e / (n - 1);
        if (ibg < nbg) {
            const int Ib = ncomp + 1 + ibg, Jb = e - ibg * (n - 1) + 1;
            if (Jb <= Ib) {
                I = Ib; J = Jb;
                if (J <= ncomp) { const int j1 = (J - 1) / nc2; iw = (J - 1) - j1 * nc2; ia = (2 + ibg) * nvec + j1 + 1; }
                else ia = (2 + ibg) * nvec + nK + (J - ncomp - 1);
            }
        }
    }
    __syncthreads();
    const int nA = s_nact;
    double acc = 0.0;
    for (int c0 = 0; c0 < nA; c0 += cs) {
        const int ns = min(cs, nA - c0);
        __syncthreads();
        for (int q = tid; q < ns * L; q += 256) {
            const int sidx = q / L, o = q - sidx * L, slot = s_act[c0 + sidx];
            const double *g = gram + (long long)slot * npair;
            double v = 0.0;
            if (o < nvec) v = g[o * (o + 1) / 2];
            else if (o < 2 * nvec) v = rhs[(long long)slot * nvec + (o - nvec)];
            else if (o < OW) { const int ibg = (o - 2 * nvec) / nvec, vv = (o - 2 * nvec) - ibg * nvec, u = nK + ibg; if (vv <= u) v = g[u * (u + 1) / 2 + vv]; }
            else v = wxy[(long long)slot * nc2 + (o - OW)];
            buf[q] = v;
        }
        __syncthreads();
        if (I >= 0)
            for (int k = 0; k < ns; k++) {
                const double *b = buf + (size_t)k * L;
                acc = __dadd_rn(acc, iw >= 0 ? __dmul_rn(b[OW + iw], b[ia]) : b[ia]);
            }
    }
    if (I < 0) return;
    if (J == n) sol[(long long)r * (D->sz.n_comp_total + 1) + I + 1] = acc;     // 1-indexed like kernelSol
    else {
        double *M = mat + (long long)r * n * n;
        M[(long long)I * n + J] = acc;
        M[(long long)J * n + I] = acc;
    }
}

// ---------------------------------------------------------------------------- LU
// alard.c:1465-1552 ludcmp + lubksb, blocked so that the O(n^3) part runs on every SM while each
// matrix element still receives exactly the reference's operations in the reference's order:
// a_ij -= l_ik * u_kj for k ascending (product rounded, then the subtraction rounded), pivot = last
// row among the maxima of vv[i]*|a_ik|, l_ik = a_ik * (1/a_kk).  A right-looking blocked LU applies
// the k's panel by panel in ascending order, so the per-element sequence is unchanged and the
// factors (and the solution) are bit-identical to the Crout loops of the reference.
//   k_lu_init    row scale factors vv, singular test
//   k_lu_panel   unblocked factorisation of one LU_NB-column panel in shared memory (one CTA per region)
//   k_lu_swap_u  the panel's row swaps on the columns outside it + the U12 triangular solve
//   k_lu_trail   A22 -= L21 * U12 on 32x32 tiles
//   k_lu_subst   permutation, forward and back substitution
#define LU_NB 32
#define LU_PT 512

__global__ void __launch_bounds__(128)
k_lu_init(const ZbDev *__restrict__ D, ZbFit *__restrict__ fit, const double *__restrict__ mat, double *__restrict__ vvs)
{
    const int r = blockIdx.y;
    if (!fit[r].active) return;
    const int n = D->sz.mat_size, i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (i >= n) return;
    const double *row = mat + (long long)r * n * n + (long long)i * n;
    double big = 0.0;
    for (int j = lane; j < n; j += 32) { double t = fabs(row[j]); if (t > big) big = t; }
    for (int o = 16; o > 0; o >>= 1) { double t = __shfl_xor_sync(0xffffffffu, big, o); if (t > big) big = t; }
    if (lane) return;
    if (big == 0.) fit[r].singular = 1; else vvs[(long long)r * n + i] = 1.0 / big;
}

__global__ void __launch_bounds__(LU_PT)
k_lu_panel(const ZbDev *__restrict__ D, const ZbFit *__restrict__ fit, double *__restrict__ mat,
           double *__restrict__ vvs, int *__restrict__ indxs, int k0)
{
    extern __shared__ __align__(16) unsigned char dyn[];
    const int r = blockIdx.x;
    if (!fit[r].active || fit[r].singular) return;
    const int n = D->sz.mat_size, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nb = min(LU_NB, n - k0), m = n - k0, ld = LU_NB + 1;
    double *a = mat + (long long)r * n * n;
    double *P = (double *)dyn;                    // [m][ld]: rows k0.., columns k0..k0+nb-1
    double *vv = P + (size_t)m * ld;              // [m]
    __shared__ double s_v[2][LU_PT / 32]; __shared__ int s_i[2][LU_PT / 32];
    for (int e = tid; e < m * nb; e += LU_PT) { int i = e / nb, c = e - i * nb; P[i * ld + c] = a[(long long)(k0 + i) * n + k0 + c]; }
    for (int i = tid; i < m; i += LU_PT) vv[i] = vvs[(long long)r * n + k0 + i];
    __syncthreads();
    for (int kk = 0; kk < nb; kk++) {
        // pivot: products are >= 0, ties -> the larger row index (the reference scans upward with >=)
        double bestv = -1.0; int besti = -1;
        for (int i = kk + tid; i < m; i += LU_PT) {
            double dum = vv[i] * fabs(P[i * ld + kk]);
            if (dum >= bestv) { bestv = dum; besti = i; }
        }
        for (int o = 16; o > 0; o >>= 1) {
            double ov = __shfl_xor_sync(0xffffffffu, bestv, o);
            int oi = __shfl_xor_sync(0xffffffffu, besti, o);
            if (ov > bestv || (ov == bestv && oi > besti)) { bestv = ov; besti = oi; }
        }
        const int par = kk & 1;                   // double-buffered partials: one barrier per reduction
        if (lane == 0) { s_v[par][warp] = bestv; s_i[par][warp] = besti; }
        __syncthreads();
        // every warp finishes the reduction for itself
        double bv = lane < LU_PT / 32 ? s_v[par][lane] : -1.0; int bi = lane < LU_PT / 32 ? s_i[par][lane] : -1;
        for (int o = 16; o > 0; o >>= 1) {
            double ov = __shfl_xor_sync(0xffffffffu, bv, o);
            int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ov > bv || (ov == bv && oi > bi)) { bv = ov; bi = oi; }
        }
        const int imax = bi;
        if (imax != kk && tid < nb) { double t = P[imax * ld + tid]; P[imax * ld + tid] = P[kk * ld + tid]; P[kk * ld + tid] = t; }
        if (tid == LU_PT - 1) {
            if (imax != kk) vv[imax] = vv[kk];
            indxs[(long long)r * n + k0 + kk] = k0 + imax;
        }
        __syncthreads();
        double piv = P[kk * ld + kk];
        if (piv == 0.0) { piv = 1.0e-20; if (tid == 0) P[kk * ld + kk] = piv; }   // row kk is not read again as a row operand
        if (k0 + kk != n - 1) {
            const double dum = 1.0 / piv;
            // a thread owns whole rows: scale l_ik, then update the rest of its panel row
            for (int i = kk + 1 + tid; i < m; i += LU_PT) {
                double *row = P + i * ld;
                const double l = row[kk] * dum;
                row[kk] = l;
                // eight columns at a time, every load ahead of the first store: a store to `row` might alias the pivot
                // row for all the compiler knows, and an in-order warp then waits out one load-multiply-subtract
                // chain per column
                const double *prow = P + kk * ld;
                int c = kk + 1;
                for (; c + 8 <= nb; c += 8) {
                    double pv[8], rv[8];
#pragma unroll
                    for (int q = 0; q < 8; q++) { pv[q] = prow[c + q]; rv[q] = row[c + q]; }
#pragma unroll
                    for (int q = 0; q < 8; q++) rv[q] = __dsub_rn(rv[q], __dmul_rn(l, pv[q]));
#pragma unroll
                    for (int q = 0; q < 8; q++) row[c + q] = rv[q];
                }
                for (; c < nb; c++) row[c] = __dsub_rn(row[c], __dmul_rn(l, prow[c]));
            }
        }
        __syncthreads();
    }
    for (int e = tid; e < m * nb; e += LU_PT) { int i = e / nb, c = e - i * nb; a[(long long)(k0 + i) * n + k0 + c] = P[i * ld + c]; }
    for (int i = tid; i < m; i += LU_PT) vvs[(long long)r * n + k0 + i] = vv[i];
}

// The panel's row interchanges on the columns outside it, then (right of the panel) the U12 solve
// u_kj = a_kj - sum_{k'<k in panel} l_kk' u_k'j.  The LU_NB interchanges are composed into one
// permutation of the <= 2*LU_NB rows involved, so a column's loads are all independent.
__global__ void __launch_bounds__(128)
k_lu_swap_u(const ZbDev *__restrict__ D, const ZbFit *__restrict__ fit, double *__restrict__ mat,
            const int *__restrict__ indxs, int k0)
{
    __shared__ double L[LU_NB][LU_NB + 1];
    __shared__ int rows[2 * LU_NB], src[2 * LU_NB], s_nrow;
    const int r = blockIdx.y;
    if (!fit[r].active || fit[r].singular) return;
    const int n = D->sz.mat_size, nb = min(LU_NB, n - k0);
    double *a = mat + (long long)r * n * n;
    for (int e = threadIdx.x; e < nb * nb; e += blockDim.x) { int i = e / nb, c = e - i * nb; L[i][c] = a[(long long)(k0 + i) * n + k0 + c]; }
    if (threadIdx.x < 32) {
        // positions 0..nb-1 are the panel rows; pivots outside the panel are appended once each.  Warp 0 composes the
        // permutation: the panel's pivots are fetched once (one per lane), the look-up among the appended rows is a
        // ballot.  (One thread reading a pivot from global memory per step and scanning the list was two thirds of
        // this kernel's 32 us.)
        const int lane = threadIdx.x;
        const int my_ip = lane < nb ? indxs[(long long)r * n + k0 + lane] : 0;
        if (lane < nb) { rows[lane] = k0 + lane; src[lane] = lane; }
        __syncwarp();
        int nrow = nb;
        for (int kk = 0; kk < nb; kk++) {
            const int ip = __shfl_sync(0xffffffffu, my_ip, kk);
            int pos;
            if (ip < k0 + nb) pos = ip - k0;
            else {
                const unsigned hit = __ballot_sync(0xffffffffu, nb + lane < nrow && rows[nb + lane] == ip);   // at most nb <= 32 appended rows
                if (hit) pos = nb + __ffs(hit) - 1;
                else { pos = nrow; if (lane == 0) { rows[nrow] = ip; src[nrow] = nrow; } nrow++; }
            }
            __syncwarp();
            if (lane == 0) { const int t = src[pos]; src[pos] = src[kk]; src[kk] = t; }      // after all swaps position p holds original row rows[src[p]]
            __syncwarp();
        }
        if (lane == 0) s_nrow = nrow;
    }
    __syncthreads();
    int j = blockIdx.x * blockDim.x + threadIdx.x;           // index among the n - nb outside columns
    if (j >= n - nb) return;
    if (j >= k0) j += nb;
    const int nrow = s_nrow;
    double val[2 * LU_NB];
#pragma unroll
    for (int q = 0; q < 2 * LU_NB; q++) val[q] = q < nrow ? a[(long long)rows[q] * n + j] : 0.0;
    // scatter through shared-free register shuffle: write position p <- val[src[p]] (dynamic index => local memory is fine here)
    double u[LU_NB];
#pragma unroll
    for (int q = 0; q < 2 * LU_NB; q++) {
        if (q < nrow) {
            const double x = val[src[q]];
            if (q >= nb) a[(long long)rows[q] * n + j] = x;
            else u[q] = x;
        }
    }
    if (j < k0) {
#pragma unroll
        for (int kk = 0; kk < LU_NB; kk++) if (kk < nb) a[(long long)(k0 + kk) * n + j] = u[kk];
        return;
    }
#pragma unroll
    for (int kk = 0; kk < LU_NB; kk++) {
        if (kk < nb) {
            double sum = u[kk];
#pragma unroll
            for (int q = 0; q < LU_NB; q++) if (q < kk) sum -= L[kk][q] * u[q];
            u[kk] = sum;
            a[(long long)(k0 + kk) * n + j] = sum;
        }
    }
}

__global__ void __launch_bounds__(256)
k_lu_trail(const ZbDev *__restrict__ D, const ZbFit *__restrict__ fit, double *__restrict__ mat, int k0)
{
    __shared__ double Ls[32][LU_NB + 1];         // rows of the tile x panel columns
    __shared__ double Us[LU_NB][32 + 1];         // panel rows x columns of the tile
    const int r = blockIdx.z;
    if (!fit[r].active || fit[r].singular) return;
    const int n = D->sz.mat_size, nb = min(LU_NB, n - k0), t0 = k0 + nb;
    const int i0 = t0 + blockIdx.y * 32, j0 = t0 + blockIdx.x * 32;
    if (i0 >= n || j0 >= n) return;
    double *a = mat + (long long)r * n * n;
    const int tid = threadIdx.x;
    for (int e = tid; e < 32 * LU_NB; e += 256) {
        int i = e / LU_NB, c = e - i * LU_NB;
        Ls[i][c] = (i0 + i < n && c < nb) ? a[(long long)(i0 + i) * n + k0 + c] : 0.0;
        int kk = e / 32, jj = e - kk * 32;
        Us[kk][jj] = (j0 + jj < n && kk < nb) ? a[(long long)(k0 + kk) * n + j0 + jj] : 0.0;
    }
    __syncthreads();
    const int jj = tid & 31, ib = tid >> 5;       // 8 row groups x 32 columns, 4 rows each
    if (j0 + jj >= n) return;
    // the four rows of a thread advance together: each element's chain (product rounded, then the difference, kk
    // ascending) is unchanged, but four independent chains are in flight instead of one
    double v[4];
#pragma unroll
    for (int q = 0; q < 4; q++) v[q] = (i0 + ib * 4 + q < n) ? a[(long long)(i0 + ib * 4 + q) * n + j0 + jj] : 0.0;
    for (int kk = 0; kk < nb; kk++) {
        const double uu = Us[kk][jj];
#pragma unroll
        for (int q = 0; q < 4; q++) v[q] = __dsub_rn(v[q], __dmul_rn(Ls[ib * 4 + q][kk], uu));
    }
#pragma unroll
    for (int q = 0; q < 4; q++) if (i0 + ib * 4 + q < n) a[(long long)(i0 + ib * 4 + q) * n + j0 + jj] = v[q];
}

// lubksb.  Forward: column sweep in blocks of 32 (per row the terms are subtracted for j ascending).
// Back: the reference subtracts a_ij*x_j for j ascending from i+1, so row i cannot start before
// x_{i+1} exists: a chain of n^2/2 dependent subtractions.  One thread runs the chain on products
// that the rest of the CTA prepares one row ahead (every x_j with j > i+1 is already known).
#define LU_ST 256
__global__ void __launch_bounds__(LU_ST)
k_lu_subst(const ZbDev *__restrict__ D, const ZbFit *__restrict__ fit, const double *__restrict__ mat,
           const int *__restrict__ indxs, double *__restrict__ sol)
{
    extern __shared__ __align__(16) unsigned char dyn[];
    const int r = blockIdx.x;
    if (!fit[r].active || fit[r].singular) return;
    const int n = D->sz.mat_size, tid = threadIdx.x;
    const double *a = mat + (long long)r * n * n;
    double *bg = sol + (long long)r * (D->sz.n_comp_total + 1) + 1;
    const int *indx = indxs + (long long)r * n;
    double *b = (double *)dyn;                    // [n]
    double *prod = b + n;                         // [2][n]
    double *dg = prod + 2 * n, *sd1 = dg + n;     // diagonal and first super-diagonal of U
    int *sidx = (int *)(sd1 + n);                 // the pivots, fetched in parallel: the permutation below walks them one by one
    __shared__ double Lb[32][33];
    for (int i = tid; i < n; i += LU_ST) {
        b[i] = bg[i];
        dg[i] = a[(long long)i * n + i];
        sd1[i] = i + 1 < n ? a[(long long)i * n + i + 1] : 0.0;
        sidx[i] = indx[i];
    }
    __syncthreads();
    if (tid == 0)
        for (int i = 0; i < n; i++) { int ip = sidx[i]; double t = b[ip]; b[ip] = b[i]; b[i] = t; }
    __syncthreads();
    for (int j0 = 0; j0 < n; j0 += 32) {
        const int w = min(32, n - j0);
        for (int e = tid; e < w * w; e += LU_ST) { int i = e / w, c = e - i * w; Lb[i][c] = a[(long long)(j0 + i) * n + j0 + c]; }
        __syncthreads();
        if (tid < 32)                              // the diagonal block, one warp in lock step
            for (int c = 0; c < w - 1; c++) {
                const double bj = b[j0 + c];
                if (tid > c && tid < w) b[j0 + tid] -= Lb[tid][c] * bj;
                __syncwarp();
            }
        __syncthreads();
        for (int i = j0 + w + tid; i < n; i += LU_ST) {
            const double *row = a + (long long)i * n + j0;
            double v = b[i];
            for (int c = 0; c < w; c++) v -= row[c] * b[j0 + c];
            b[i] = v;
        }
        __syncthreads();
    }
    // back substitution
    for (int i = n - 1; i >= 0; i--) {
        double *p = prod + (size_t)(i & 1) * n;    // products of row i for j >= i+2, prepared during row i+1
        if (i == n - 1) { /* nothing to subtract */ }
        if (tid == 0) {
            double sum = b[i];
            if (i + 1 < n) sum -= sd1[i] * b[i + 1];
            // the chain of dependent subtractions is the critical path of the whole solve: the next eight products are
            // in registers before the current eight are subtracted, so only the adder's latency is paid per term
            int j = i + 2;
            if (j + 8 <= n) {
                double q0 = p[j], q1 = p[j + 1], q2 = p[j + 2], q3 = p[j + 3], q4 = p[j + 4], q5 = p[j + 5], q6 = p[j + 6], q7 = p[j + 7];
                for (j += 8; j + 8 <= n; j += 8) {
                    const double r0 = p[j], r1 = p[j + 1], r2 = p[j + 2], r3 = p[j + 3], r4 = p[j + 4], r5 = p[j + 5], r6 = p[j + 6], r7 = p[j + 7];
                    sum -= q0; sum -= q1; sum -= q2; sum -= q3; sum -= q4; sum -= q5; sum -= q6; sum -= q7;
                    q0 = r0; q1 = r1; q2 = r2; q3 = r3; q4 = r4; q5 = r5; q6 = r6; q7 = r7;
                }
                sum -= q0; sum -= q1; sum -= q2; sum -= q3; sum -= q4; sum -= q5; sum -= q6; sum -= q7;
            }
            for (; j < n; j++) sum -= p[j];
            b[i] = sum / dg[i];
        } else if (tid >= 32 && i >= 1) {
            // prepare row i-1: products with x_j, j >= i+1 (all known while thread 0 works on row i).  Warps 1-7 only:
            // with lanes 1-31 of warp 0 taking part, thread 0's chain ran behind its own warp's global loads (the two
            // sides of a divergent warp take turns): 2-3 k cycles per matrix row, as much as the chain itself
            // (ncu source view, profiles/r02_ncu_k_lu_subst_c1b.txt)
            const double *rown = a + (long long)(i - 1) * n;
            double *pn = prod + (size_t)((i - 1) & 1) * n;
            for (int j = i + 1 + (tid - 32); j < n; j += LU_ST - 32) pn[j] = rown[j] * b[j];
        }
        __syncthreads();
    }
    for (int i = tid; i < n; i += LU_ST) bg[i] = b[i];
}

// ------------------------------------------------------------ model and clipping

// noise^2 used as the fit's figure-of-merit weight (hotpants/main.c:739-770): squares and the sum of the two images'
// terms in float; the gain model (makeNoiseImage4, functions.c:1172-1187) multiplies and adds in DOUBLE (C's fabs
// returns double) and rounds once -- fabs(float) here would be C++'s float overload and round the product first
__device__ static inline float approx_var(const ZbDev *D, const float *sci, const float *ref,
                                          const float *sci_rms, const float *ref_rms, long long g)
{
    float o, e;
    if (sci_rms) { float v = sci_rms[g]; o = v * v; }
    else { float inv = (float)(1. / D->p.i_gain), quad = D->p.i_rdnoise / D->p.i_gain; double qq = quad * quad; o = (float)(fabs((double)sci[g]) * (double)inv + qq); }
    if (ref_rms) { float v = ref_rms[g]; e = v * v; }
    else { float inv = (float)(1. / D->p.t_gain), quad = D->p.t_rdnoise / D->p.t_gain; double qq = quad * quad; e = (float)(fabs((double)ref[g]) * (double)inv + qq); }
    return o + e;
}

#define SIG_PX 10
__global__ void __launch_bounds__(256)
k_stamp_sig(const ZbDev *__restrict__ D, ZbStamp *__restrict__ stamps, const ZbFit *__restrict__ fit,
            const double *__restrict__ vec, const double *__restrict__ kref, const double *__restrict__ sols,
            const float *__restrict__ sci, const float *__restrict__ ref,
            const float *__restrict__ sci_rms, const float *__restrict__ ref_rms, int32_t *__restrict__ m_in, int slot0)
{
    extern __shared__ __align__(16) unsigned char dyn[];
    const int slot = slot0 + blockIdx.x;
    ZbStamp &S = stamps[slot];
    if (S.nss <= 0 || S.sscnt >= S.nss || S.excl) return;
    if (!fit[S.region].active) return;
    const ZbRegion R = D->reg[S.region];
    const int hwss = D->p.hw_ksstamp, fwss = D->sz.fw_ksstamp, npix = fwss * fwss;
    const int nK = D->sz.n_comp_ker, nvec = nK + D->sz.n_bg_vectors, ko = D->p.ker_order, bgo = D->p.bg_order;
    const double *sol = sols + (long long)S.region * (D->sz.n_comp_total + 1);
    double *term = (double *)dyn;                     // [npix]
    double *coef = term + npix;                       // [nK]
    signed char *ok = (signed char *)(coef + nK);     // [npix]
    __shared__ double s_bg;
    const int xR = S.xss[S.sscnt], yR = S.yss[S.sscnt];
    const double xf = (xR - 0.5 * R.rpx) / (0.5 * R.rpx), yf = (yR - 0.5 * R.rpy) / (0.5 * R.rpy);
    const int nc2 = D->sz.n_comp;
    for (int i1 = threadIdx.x; i1 < nK; i1 += blockDim.x) {
        if (i1 == 0) { coef[0] = sol[1]; continue; }
        int k = 2 + (i1 - 1) * nc2;
        double c = 0.0, ax = 1.0;
        for (int ix = 0; ix <= ko; ix++) {
            double ay = 1.0;
            for (int iy = 0; iy <= ko - ix; iy++) { c += sol[k++] * ax * ay; ay *= yf; }
            ax *= xf;
        }
        coef[i1] = c;
    }
    if (threadIdx.x == 0) {
        int ncompBG = (nK - 1) * nc2 + 1, k = 1;
        double bg = 0.0, ax = 1.0;
        for (int i = 0; i <= bgo; i++) {
            double ay = 1.0;
            for (int j = 0; j <= bgo - i; j++) { bg += sol[ncompBG + k++] * ax * ay; ay *= yf; }
            ax *= xf;
        }
        s_bg = bg;
    }
    __syncthreads();
    const double bg = s_bg;
    const double *v = vec + (long long)slot * nvec * npix;
    const double *kr = kref + (long long)slot * npix;
    int32_t *gm = m_in + R.off;
    // make_model keeps its running model in the reference's FLOAT scratch (alard.c:1441-1459): the sum over
    // the bases is rounded to float after every term.  SIG_PX pixels per thread advance together so the
    // 49 dependent rounding steps of one pixel overlap with those of the others.
    for (int kbase = 0; kbase < npix; kbase += 256 * SIG_PX) {
        float m[SIG_PX];
#pragma unroll
        for (int q = 0; q < SIG_PX; q++) m[q] = 0.0f;
        for (int n = 0; n < nK; n++) {
            const double c = coef[n];
            const double *vn = v + (long long)n * npix + kbase + threadIdx.x;
#pragma unroll
            for (int q = 0; q < SIG_PX; q++)
                if (kbase + threadIdx.x + 256 * q < npix) m[q] = (float)((double)m[q] + c * vn[256 * q]);
        }
#pragma unroll
        for (int q = 0; q < SIG_PX; q++) {
            const int k = kbase + threadIdx.x + 256 * q;
            if (k >= npix) continue;
            int j = k / fwss, i = k - j * fwss;
            int x2 = xR - hwss + i, y2 = yR - hwss + j;
            double tdat = m[q], idat = kr[k];
            double ndat = approx_var(D, sci, ref, sci_rms, ref_rms, (long long)(R.by0 + y2) * D->nx + (R.bx0 + x2));
            double diff = tdat - idat + bg;
            signed char f = 0;
            if (!((gm[(long long)y2 * R.rpx + x2] & ZB_FLAG_INPUT_ISBAD) || (fabs(idat) <= ZB_ZEROVAL))) {
                if ((tdat * 0.0 != 0.0) || (idat * 0.0 != 0.0))
                    atomicOr(&gm[(long long)y2 * R.rpx + x2], ZB_FLAG_INPUT_ISBAD | ZB_FLAG_ISNAN);
                else { f = 1; term[k] = diff * diff / ndat; }
            }
            ok[k] = f;
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double sig1 = 0; int nsig = 0;
        for (int k = 0; k < npix; k++) if (ok[k]) { nsig++; sig1 += term[k]; }
        if (nsig > 0) { sig1 /= nsig; if (sig1 >= ZB_MAXVAL) sig1 = -1; }
        else sig1 = -1;
        S.sig1 = sig1;
    }
}

// functions.c:1037-1107 sigma_clip, single thread, on a small float array
__device__ static void sigma_clip_seq(const float *data, char *rej, int count, double *mean, double *stdev, int maxiter, float statSig)
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

// check_again (alard.c:1143-1247), one CTA per region.  The per-stamp steps are independent and run one stamp per
// thread; the sigma clip in between sums in stamp order (sequential by definition) and is done by one thread on
// shared memory.  As one thread per region walking global memory this took 137 us at 100 stamps per region and
// ~1.4 ms at 1024, once per clip iteration (profiles/r02_launches_c1b_after_summary.txt).
__global__ void __launch_bounds__(256)
k_clip(const ZbDev *__restrict__ D, ZbStamp *__restrict__ stamps, ZbFit *__restrict__ fit, int *__restrict__ jobs, int *__restrict__ njobs)
{
    extern __shared__ __align__(16) unsigned char clip_dyn[];
    const int r = blockIdx.x, tid = threadIdx.x;
    ZbFit &F = fit[r];
    if (!F.active) return;
    const int base = D->reg[r].stamp_base, nS = F.n_stamps, cap = D->sz.n_stamps;
    double *chi2 = (double *)clip_dyn;               // [cap] figure the second test uses (a bumped stamp keeps its previous one)
    float *val = (float *)(chi2 + cap);              // [cap] (float)sig1 of the stamps that enter the clip
    float *ss = val + cap;                           // [cap] the same, compacted in stamp order
    int *sc = (int *)(ss + cap);                     // [cap] sscnt
    int *ns_ = sc + cap;                             // [cap] nss
    char *live = (char *)(ns_ + cap);                // [cap] enters the clip
    char *rej = live + cap;                          // [cap]
    char *refill = rej + cap;                        // [cap]
    __shared__ double s_mean, s_sd;
    __shared__ int s_check, s_nskip;
    if (tid == 0) { s_check = 0; s_nskip = 0; }
    __syncthreads();
    for (int k = tid; k < nS; k += 256) {
        ZbStamp &S = stamps[base + k];
        int sscnt = S.sscnt; const int nss = S.nss;
        char lv = 0, rf = 0;
        double c2 = S.chi2;
        if (sscnt < nss) {
            const double sig1 = S.sig1;
            if (sig1 == -1) { sscnt++; rf = 1; s_check = 1; }
            else { c2 = sig1; lv = 1; val[k] = (float)sig1; }
        } else atomicAdd(&s_nskip, 1);
        chi2[k] = c2; sc[k] = sscnt; ns_[k] = nss; live[k] = lv; refill[k] = rf;
    }
    __syncthreads();
    if (tid == 0) {
        int nss = 0;
        for (int k = 0; k < nS; k++) if (live[k]) ss[nss++] = val[k];
        double mean = 0, stdev = 0;
        sigma_clip_seq(ss, rej, nss, &mean, &stdev, 10, D->p.stat_sig);
        s_mean = mean; s_sd = stdev;
    }
    __syncthreads();
    const double mean = s_mean, stdev = s_sd;
    for (int k = tid; k < nS; k += 256) {
        if (sc[k] < ns_[k] && (chi2[k] - mean) > D->p.ker_sig_reject * stdev) { sc[k]++; refill[k] = 1; s_check = 1; }
        ZbStamp &S = stamps[base + k];
        S.sscnt = sc[k]; S.refill = refill[k]; S.chi2 = chi2[k];
    }
    __syncthreads();
    if (tid == 0) {
        int cnt = 0;
        for (int k = 0; k < nS; k++) cnt += (refill[k] && sc[k] < ns_[k]);
        int q = cnt ? atomicAdd(njobs, cnt) : 0;
        for (int k = 0; k < nS; k++) if (refill[k] && sc[k] < ns_[k]) jobs[q++] = base + k;
        F.meansig = mean; F.scatter = stdev; F.n_skipped = s_nskip; F.iters++;
        F.active = s_check;
    }
}

// wxy + global normal equations + LU solve (alard.c:343-360) for every active region
int zb_launch_fit_solve(zb_ctx *ctx, cudaStream_t st)
{
    const zb_sizes &sz = ctx->sz;
    const int n = sz.mat_size, nR = sz.n_regions;
    k_wxy<<<(ctx->n_slots + 127) / 128, 128, 0, st>>>(ctx->d_dev, ctx->d_stamps, ctx->n_slots, ctx->d_wxy); ZB_DBG("k_wxy");
    {
        const int PP = sz.n_comp * sz.n_comp, G = BM_GROUPS(PP);
        const int nb1 = sz.n_comp_ker - 1, nuv = nb1 * (nb1 + 1) / 2;
        const size_t sm_main = (size_t)BM_CS * (G * BM_R + sz.n_comp) * sizeof(double) + (size_t)(sz.n_stamps + G * BM_R + 1) * sizeof(int);
        // the edge records grow with the background order ((2 + nbg) * nvec values per stamp): fewer stamps per round then
        const size_t rec_edge = ((size_t)(2 + sz.n_bg_vectors) * (sz.n_comp_ker + sz.n_bg_vectors) + sz.n_comp) * sizeof(double);
        const size_t act_bytes = (size_t)(sz.n_stamps + 1) * sizeof(int);
        int cs_edge = BM_CS;
        while (cs_edge > 1 && cs_edge * rec_edge + act_bytes > (size_t)ZB_SMEM_OPTIN / 2) cs_edge /= 2;
        const size_t sm_edge = cs_edge * rec_edge + act_bytes;
        if (sm_main > ZB_SMEM_OPTIN || sm_edge > ZB_SMEM_OPTIN) { snprintf(ctx->err, sizeof ctx->err, "%d stamps per region too many for the normal-matrix kernels", sz.n_stamps); return ZB_ERR_UNSUPPORTED; }
        cudaFuncSetAttribute(k_build_matrix_main, cudaFuncAttributeMaxDynamicSharedMemorySize, ZB_SMEM_OPTIN);
        cudaFuncSetAttribute(k_build_matrix_edges, cudaFuncAttributeMaxDynamicSharedMemorySize, ZB_SMEM_OPTIN);
        if (nuv > 0)
            k_build_matrix_main<<<dim3((nuv + G * BM_R - 1) / (G * BM_R), nR), 256, sm_main, st>>>(
                ctx->d_dev, ctx->d_stamps, ctx->d_fit, ctx->d_gram, ctx->d_wxy, ctx->d_mat); ZB_DBG("k_build_matrix_main");
        const int nedge = 2 * n + sz.n_bg_vectors * (n - 1);
        k_build_matrix_edges<<<dim3((nedge + 255) / 256, nR), 256, sm_edge, st>>>(ctx->d_dev, ctx->d_stamps, ctx->d_fit, ctx->d_gram, ctx->d_rhs, ctx->d_wxy, ctx->d_mat, ctx->d_sol, cs_edge); ZB_DBG("k_build_matrix_edges");
        ctx->launches++;
    }
    k_lu_init<<<dim3((n + 3) / 4, nR), 128, 0, st>>>(ctx->d_dev, ctx->d_fit, ctx->d_mat, ctx->d_vv); ZB_DBG("k_lu_init");
    const size_t smp = ((size_t)n * (LU_NB + 1) + n) * sizeof(double);
    if (smp > ZB_SMEM_OPTIN) { snprintf(ctx->err, sizeof ctx->err, "normal matrix of %d rows too large for the LU panel kernel", n); return ZB_ERR_UNSUPPORTED; }
    cudaFuncSetAttribute(k_lu_panel, cudaFuncAttributeMaxDynamicSharedMemorySize, ZB_SMEM_OPTIN);
    cudaFuncSetAttribute(k_lu_subst, cudaFuncAttributeMaxDynamicSharedMemorySize, ZB_SMEM_OPTIN);
    // The factorisation is 3 launches per 32 columns (70 per solve at n = 735, one solve per clip iteration, 38 iterations
    // at configs[3]): with 32 contexts the host's launch rate, not the GPU, set the pace.  The sequence depends only on
    // (n, regions) and on workspace pointers that live as long as the configuration, so it is captured once into a
    // graph and replayed with one launch.  ZB_LU_GRAPH=0: plain launches.
    static int use_graph = -1;
    if (use_graph < 0) { const char *e = getenv("ZB_LU_GRAPH"); use_graph = (e && e[0] == '0') ? 0 : 1; }
    int nlaunch = 0;
    auto enqueue = [&](cudaStream_t s) {
        for (int k0 = 0; k0 < n; k0 += LU_NB) {
            const int nb = n - k0 < LU_NB ? n - k0 : LU_NB, rest = n - k0 - nb;
            k_lu_panel<<<nR, LU_PT, ((size_t)(n - k0) * (LU_NB + 1) + (n - k0)) * sizeof(double), s>>>(ctx->d_dev, ctx->d_fit, ctx->d_mat, ctx->d_vv, ctx->d_indx, k0);
            nlaunch++;
            if (n - nb > 0) { k_lu_swap_u<<<dim3((n - nb + 127) / 128, nR), 128, 0, s>>>(ctx->d_dev, ctx->d_fit, ctx->d_mat, ctx->d_indx, k0); nlaunch++; }
            if (rest > 0) { k_lu_trail<<<dim3((rest + 31) / 32, (rest + 31) / 32, nR), 256, 0, s>>>(ctx->d_dev, ctx->d_fit, ctx->d_mat, k0); nlaunch++; }
        }
        k_lu_subst<<<nR, LU_ST, (size_t)5 * n * sizeof(double) + (size_t)n * sizeof(int), s>>>(ctx->d_dev, ctx->d_fit, ctx->d_mat, ctx->d_indx, ctx->d_sol);
        nlaunch++;
    };
    if (use_graph && !(ctx->lu_exec && ctx->lu_key[0] == n && ctx->lu_key[1] == nR)) {
        if (ctx->lu_exec) { cudaGraphExecDestroy(ctx->lu_exec); ctx->lu_exec = NULL; }
        // captured on the context's own stream (a legacy default stream cannot be captured); nothing runs during capture
        cudaGraph_t g = NULL;
        if (cudaStreamBeginCapture(ctx->stream, cudaStreamCaptureModeThreadLocal) == cudaSuccess) {
            enqueue(ctx->stream);
            if (cudaStreamEndCapture(ctx->stream, &g) == cudaSuccess && g) {
                if (cudaGraphInstantiate(&ctx->lu_exec, g, 0) != cudaSuccess) ctx->lu_exec = NULL;
                cudaGraphDestroy(g);
            }
        }
        (void)cudaGetLastError();
        ctx->lu_key[0] = n; ctx->lu_key[1] = nR;
        nlaunch = 0;
    }
    if (use_graph && ctx->lu_exec) {
        if (cudaGraphLaunch(ctx->lu_exec, st) != cudaSuccess) return zb_cuda_fail(ctx, cudaGetLastError(), "LU graph launch", __FILE__, __LINE__);
        for (int k0 = 0; k0 < n; k0 += LU_NB) { const int nb = n - k0 < LU_NB ? n - k0 : LU_NB; nlaunch += 1 + (n - nb > 0) + (n - k0 - nb > 0); }
        nlaunch++;                                   // kernels launched (by one graph launch)
    } else enqueue(st);
    ZB_DBG("LU");
    ctx->launches += nlaunch + 3;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return zb_cuda_fail(ctx, e, "fit solve", __FILE__, __LINE__);
    return ZB_OK;
}

// getStampSig (alard.c:1062-1140) for the stamp slots [slot0, slot1)
int zb_launch_stamp_sig(zb_ctx *ctx, const float *sci, const float *ref, const float *sci_rms, const float *ref_rms,
                        int slot0, int slot1, cudaStream_t st)
{
    if (slot1 <= slot0) return ZB_OK;
    const zb_sizes &sz = ctx->sz;
    const int npix = sz.fw_ksstamp * sz.fw_ksstamp;
    size_t sm = (size_t)npix * (sizeof(double) + 1) + (size_t)sz.n_comp_ker * sizeof(double) + 16;
    cudaFuncSetAttribute(k_stamp_sig, cudaFuncAttributeMaxDynamicSharedMemorySize, ZB_SMEM_OPTIN);
    k_stamp_sig<<<slot1 - slot0, 256, sm, st>>>(ctx->d_dev, ctx->d_stamps, ctx->d_fit, ctx->d_vec, ctx->d_kref, ctx->d_sol,
                                                 sci, ref, sci_rms, ref_rms, ctx->d_min, slot0); ZB_DBG("k_stamp_sig");
    ctx->launches++;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return zb_cuda_fail(ctx, e, "stamp sig", __FILE__, __LINE__);
    return ZB_OK;
}

// check_again's decisions (alard.c:1143-1247); leaves the slots to refill in d_jobs / d_njobs
int zb_launch_clip(zb_ctx *ctx, cudaStream_t st)
{
    const zb_sizes &sz = ctx->sz;
    const int nR = sz.n_regions;
    zb_zero_small(ctx, ctx->d_njobs, sizeof(int), st);
    const size_t sm = (size_t)sz.n_stamps * (sizeof(double) + 2 * sizeof(float) + 2 * sizeof(int) + 3) + 16;
    if (sm > ZB_SMEM_OPTIN) { snprintf(ctx->err, sizeof ctx->err, "%d stamps per region too many for the clip kernel", sz.n_stamps); return ZB_ERR_UNSUPPORTED; }
    cudaFuncSetAttribute(k_clip, cudaFuncAttributeMaxDynamicSharedMemorySize, ZB_SMEM_OPTIN);
    k_clip<<<nR, 256, sm, st>>>(ctx->d_dev, ctx->d_stamps, ctx->d_fit, ctx->d_jobs, ctx->d_njobs); ZB_DBG("k_clip");
    ctx->launches++;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return zb_cuda_fail(ctx, e, "clip", __FILE__, __LINE__);
    return ZB_OK;
}

int zb_launch_fit_iteration(zb_ctx *ctx, const float *conv_img, const float *ref_img,
                            const float *sci, const float *ref, const float *sci_rms, const float *ref_rms, cudaStream_t st)
{
    (void)conv_img; (void)ref_img;
    int rc = zb_launch_fit_solve(ctx, st);
    if (rc != ZB_OK) return rc;
    rc = zb_launch_stamp_sig(ctx, sci, ref, sci_rms, ref_rms, 0, ctx->n_slots, st);
    if (rc != ZB_OK) return rc;
    return zb_launch_clip(ctx, st);
}
