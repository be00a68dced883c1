// zb_bkg.cu — zuds.utils.quick_background_estimate on the GPU (zuds/utils.py:32-53):
//     bkgpix = data[mask == 0];  bkg = median(bkgpix);  bkgstd = 1.4826 * median(|bkgpix - bkg|)
// evaluated as numpy evaluates it on the float32 pixels zuds holds: the median of float32 values is a float32
// (the two central values of an even count are added and halved in float32), the deviations |x - bkg| are
// float32 subtractions, and only the factor 1.4826 promotes to double.  Both medians are EXACT order
// statistics found by radix selection on order-preserving integer keys, 11 bits per pass:
//     x      -> 32-bit key of the float    (3 passes: 11 + 11 + 10 bits)
//     |x-b|  -> 32-bit key of the float    (3 passes), recomputed from x on the fly, never stored
// A pass streams the pixels once (HBM-bound: 4 B + mask bytes per pixel), builds the digit histogram
// of the keys that still match the prefix of either wanted rank in shared memory, and the last CTA
// to finish narrows the two prefixes.  Integer work throughout, so the result is bit-exact.
#include "zb_internal.h"
#include <math.h>

#define BK_BINS 2048
#define BK_NT 256

struct BkState {
    unsigned long long prefix[2];   // key bits fixed so far (high bits), one per wanted rank
    long long k[2];                 // rank still to find inside the prefix
    long long n;                    // unmasked pixels
    long long n_nan;                // unmasked NaNs (numpy's median is NaN then)
    int ticket, pad;
    double bkg, mad;
    double val[2];                  // the two central order statistics of the current selection
    unsigned int hist[2][BK_BINS];
};

__device__ static inline unsigned int key32(float f) {
    unsigned int u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ static inline float unkey32(unsigned int k) {
    return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

// sky pixels share their leading key bits, so whole warps hit one bin: aggregate before the shared atomic
__device__ static inline void bk_hist_add(unsigned int *h, unsigned int dg) {
    const unsigned peers = __match_any_sync(__activemask(), dg);
    if ((int)(__ffs(peers) - 1) == (int)(threadIdx.x & 31)) atomicAdd(&h[dg], (unsigned int)__popc(peers));
}

// One radix pass.  mode 0: keys of x, mode 1: keys of |x - bkg| (32 bits, held in the high half of a 64-bit key).  `shift` = position of this pass's digit, `bits` its width; keys whose bits above
// the digit equal prefix[r] are counted for rank r.  first: nothing fixed yet, every key matches, and the
// unmasked / NaN counts are taken too.
__global__ void __launch_bounds__(BK_NT)
k_bkg_pass(const float *__restrict__ data, const void *__restrict__ mask, int mdt, long long n,
           BkState *__restrict__ st, int mode, int shift, int bits, int first, int last)
{
    __shared__ unsigned int h[2][BK_BINS];
    __shared__ int s_last;
    __shared__ long long s_cnt[2];
    const int tid = threadIdx.x;
    for (int i = tid; i < 2 * BK_BINS; i += BK_NT) (&h[0][0])[i] = 0;
    if (tid < 2) s_cnt[tid] = 0;
    __syncthreads();
    const unsigned long long p0 = st->prefix[0], p1 = st->prefix[1];
    const bool same = first || p0 == p1;
    const double bkg = st->bkg;
    const unsigned int dmask = (1u << bits) - 1u;
    const int hi = shift + bits;                       // bits above the digit
    long long cnt = 0, nan = 0;
    for (long long i = (long long)blockIdx.x * BK_NT + tid; i < n; i += (long long)gridDim.x * BK_NT) {
        if (zb_read_mask(mask, mdt, i) != 0) continue;
        const float x = data[i];
        if (x != x) { nan++; cnt++; continue; }
        cnt++;
        unsigned long long key;
        if (mode == 0) key = (unsigned long long)key32(x) << 32;
        else key = (unsigned long long)key32(fabsf(__fsub_rn(x, (float)bkg))) << 32;
        const unsigned long long top = hi >= 64 ? 0ull : (key >> hi);
        const unsigned int dg = (unsigned int)(key >> shift) & dmask;
        if (first || top == p0) bk_hist_add(h[0], dg);
        if (!same && top == p1) bk_hist_add(h[1], dg);
    }
    if (first) {
        for (int o = 16; o > 0; o >>= 1) { cnt += __shfl_xor_sync(0xffffffffu, cnt, o); nan += __shfl_xor_sync(0xffffffffu, nan, o); }
        if ((tid & 31) == 0) { atomicAdd((unsigned long long *)&s_cnt[0], (unsigned long long)cnt); atomicAdd((unsigned long long *)&s_cnt[1], (unsigned long long)nan); }
    }
    __syncthreads();
    for (int i = tid; i < BK_BINS; i += BK_NT) {
        if (h[0][i]) atomicAdd(&st->hist[0][i], h[0][i]);
        if (!same && h[1][i]) atomicAdd(&st->hist[1][i], h[1][i]);
    }
    if (first && tid == 0 && mode == 0) {
        atomicAdd((unsigned long long *)&st->n, (unsigned long long)s_cnt[0]);
        atomicAdd((unsigned long long *)&st->n_nan, (unsigned long long)s_cnt[1]);
    }
    __threadfence();
    __syncthreads();
    if (tid == 0) s_last = (atomicAdd(&st->ticket, 1) == (int)gridDim.x - 1);
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    // ---- the last CTA narrows the prefixes: warp r handles rank r
    const int warp = tid >> 5, lane = tid & 31;
    if (warp < 2) {
        const int r = warp;
        if (first && lane == 0) {                      // ranks of the two central values (np.median)
            const long long m = __ldcg(&st->n);
            st->k[r] = r == 0 ? (m - 1) / 2 : m / 2;
        }
        __syncwarp();
        const unsigned int *hh = st->hist[(same) ? 0 : r];
        long long k = __ldcg(&st->k[r]);
        // 2048 bins, 64 per lane: lane totals -> exclusive scan -> the lane holding rank k -> its bin
        long long tot = 0;
        for (int b = 0; b < BK_BINS / 32; b++) tot += __ldcg(&hh[lane * (BK_BINS / 32) + b]);
        long long inc = tot;
        for (int o = 1; o < 32; o <<= 1) { long long t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
        const long long before = inc - tot;
        const bool mine = k >= before && k < inc;
        const unsigned ball = __ballot_sync(0xffffffffu, mine);
        if (ball && lane == __ffs(ball) - 1) {
            long long c = before; int dg = 0;
            for (int b = 0; b < BK_BINS / 32; b++) {
                const long long v = __ldcg(&hh[lane * (BK_BINS / 32) + b]);
                if (k < c + v) { dg = lane * (BK_BINS / 32) + b; break; }
                c += v;
            }
            const unsigned long long pre = (first ? 0ull : st->prefix[r]);
            const unsigned long long np = (pre << bits) | (unsigned long long)dg;
            st->prefix[r] = np; st->k[r] = k - c;
            if (last) st->val[r] = (double)unkey32((unsigned int)(np << shift >> 32));
        }
    }
    __syncthreads();
    for (int i = tid; i < 2 * BK_BINS; i += BK_NT) (&st->hist[0][0])[i] = 0;
    if (tid == 0) st->ticket = 0;
    if (last && tid == 0) {
        __threadfence();
        const long long m = st->n;
        double v;
        if (m <= 0 || st->n_nan > 0) v = __longlong_as_double(0x7ff8000000000000ll);     // numpy: median of nothing / of NaNs is NaN
        else v = (double)__fmul_rn(__fadd_rn((float)st->val[0], (float)st->val[1]), 0.5f);      // np.mean of two float32
        if (mode == 0) st->bkg = v; else st->mad = v;
    }
}

static int bkg_run(zb_ctx *ctx, const float *data, const void *mask, int mdt, long long n, double *bkg, double *bkgstd, cudaStream_t st)
{
    if (!ctx->d_bkg) {
        if (cudaMalloc((void **)&ctx->d_bkg, sizeof(BkState)) != cudaSuccess) { snprintf(ctx->err, sizeof ctx->err, "cudaMalloc of the background-estimate state failed"); return ZB_ERR_NOMEM; }
        if (cudaMallocHost((void **)&ctx->h_bkg, 2 * sizeof(double)) != cudaSuccess) { snprintf(ctx->err, sizeof ctx->err, "cudaMallocHost failed"); return ZB_ERR_NOMEM; }
    }
    BkState *S = (BkState *)ctx->d_bkg;
    zb_zero_small(ctx, S, sizeof(BkState), st);
    long long want = (n + BK_NT * 16 - 1) / (BK_NT * 16);
    const int grid = (int)(want < 1 ? 1 : (want > 148 * 8 ? 148 * 8 : want));
    // median of x: 32-bit keys in the high half of the 64-bit key
    k_bkg_pass<<<grid, BK_NT, 0, st>>>(data, mask, mdt, n, S, 0, 53, 11, 1, 0);
    k_bkg_pass<<<grid, BK_NT, 0, st>>>(data, mask, mdt, n, S, 0, 42, 11, 0, 0);
    k_bkg_pass<<<grid, BK_NT, 0, st>>>(data, mask, mdt, n, S, 0, 32, 10, 0, 1);
    // median of |x - bkg| (float32 deviations): the same three passes
    k_bkg_pass<<<grid, BK_NT, 0, st>>>(data, mask, mdt, n, S, 1, 53, 11, 1, 0);
    k_bkg_pass<<<grid, BK_NT, 0, st>>>(data, mask, mdt, n, S, 1, 42, 11, 0, 0);
    k_bkg_pass<<<grid, BK_NT, 0, st>>>(data, mask, mdt, n, S, 1, 32, 10, 0, 1);
    ctx->launches += 6;
    zb_copy_small(ctx, &ctx->h_bkg[0], &S->bkg, sizeof(double), st);
    zb_copy_small(ctx, &ctx->h_bkg[1], &S->mad, sizeof(double), st);
    cudaError_t e = zb_stream_wait(ctx, st);
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e != cudaSuccess) return zb_cuda_fail(ctx, e, "background estimate", __FILE__, __LINE__);
    *bkg = ctx->h_bkg[0];
    *bkgstd = 1.4826 * ctx->h_bkg[1];
    return ZB_OK;
}

static size_t bk_mask_elem(int dt) { return dt == ZB_MASK_U8 ? 1 : (dt == ZB_MASK_I16 ? 2 : 4); }

extern "C" int zb_background_estimate(zb_ctx *ctx, const float *data, const void *mask, int mask_dtype, long long n,
                                      double *bkg, double *bkgstd, void *stream)
{
    if (!ctx || !data || !bkg || !bkgstd || n < 0) { if (ctx) snprintf(ctx->err, sizeof ctx->err, "null argument"); return ZB_ERR_PARAM; }
    if (mask && !(mask_dtype == ZB_MASK_U8 || mask_dtype == ZB_MASK_I16 || mask_dtype == ZB_MASK_I32)) { snprintf(ctx->err, sizeof ctx->err, "bad mask dtype %d", mask_dtype); return ZB_ERR_PARAM; }
    if (cudaSetDevice(ctx->device) != cudaSuccess) return zb_cuda_fail(ctx, cudaGetLastError(), "cudaSetDevice", __FILE__, __LINE__);
    return bkg_run(ctx, data, mask, mask_dtype, n, bkg, bkgstd, zb_pick_stream(ctx, stream));
}

// host arrays in (what zuds holds: FITS data and mask), staged through the context's buffers
extern "C" int zb_background_estimate_host(zb_ctx *ctx, const float *data, const void *mask, int mask_dtype, long long n,
                                           double *bkg, double *bkgstd)
{
    if (!ctx || !data || !bkg || !bkgstd || n < 0) { if (ctx) snprintf(ctx->err, sizeof ctx->err, "null argument"); return ZB_ERR_PARAM; }
    if (mask && !(mask_dtype == ZB_MASK_U8 || mask_dtype == ZB_MASK_I16 || mask_dtype == ZB_MASK_I32)) { snprintf(ctx->err, sizeof ctx->err, "bad mask dtype %d", mask_dtype); return ZB_ERR_PARAM; }
    if (cudaSetDevice(ctx->device) != cudaSuccess) return zb_cuda_fail(ctx, cudaGetLastError(), "cudaSetDevice", __FILE__, __LINE__);
    const size_t nb = (size_t)n * 4, mb = mask ? (size_t)n * bk_mask_elem(mask_dtype) : 0;
    int rc = zb_stage_buf(ctx, 0, nb ? nb : 16);
    if (rc != ZB_OK) return rc;
    if (mask) { rc = zb_stage_buf(ctx, 4, mb ? mb : 16); if (rc != ZB_OK) return rc; }
    cudaMemcpyAsync(ctx->d_stage[0], data, nb, cudaMemcpyHostToDevice, ctx->stream);
    if (mask) cudaMemcpyAsync(ctx->d_stage[4], mask, mb, cudaMemcpyHostToDevice, ctx->stream);
    return bkg_run(ctx, (const float *)ctx->d_stage[0], mask ? ctx->d_stage[4] : NULL, mask_dtype, n, bkg, bkgstd, ctx->stream);
}

// ------------------------------------------------------------------------------------------------------------------
// zb_background_estimate_static — the same estimate for callers that may not allocate (code under CUDA graph capture,
// tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The radix-selection state that bkg_run keeps in a lazily cudaMalloc'ed BkState lives here in a
// file-scope device array that every GPU loading the library holds until exit; no context, no pinned host buffer, no
// synchronisation: the two results are written to a device buffer of the caller, ordered on the caller's stream.
// In host code the array's NAME is only a host-side stand-in for the symbol, not a device address (a kernel handed it
// faults, or silently uses the wrong memory where the GPU can reach host memory): the device address is obtained once
// per device with cudaGetSymbolAddress, with that device current, and that pointer is what the kernels get.
// Every call on a device shares the array and finds the previous call's contents in it: the whole state is cleared on
// the stream at the start of each call, and two calls that overlap on the GPU (different streams without an ordering
// between them) race.
#include <mutex>

static __device__ __align__(256) unsigned char g_bkg_static[ZB_BKG_STATIC_SCRATCH_BYTES];
static_assert(sizeof(BkState) <= ZB_BKG_STATIC_SCRATCH_BYTES, "the static scratch array is smaller than the selection state");

#define BK_MAX_DEVICES 64
static std::mutex g_bkg_static_mu;
static BkState *g_bkg_static_ptr[BK_MAX_DEVICES];      // device address of g_bkg_static on each device, NULL until resolved

__global__ void k_bkg_emit(const BkState *__restrict__ st, double *__restrict__ out)
{
    out[0] = st->bkg;
    out[1] = 1.4826 * st->mad;          // the one double operation of zuds/utils.py:52 (this file is built with -fmad=false)
}

static int bkg_static_fail(char *msg, int msglen, int rc, const char *what, cudaError_t e)
{
    if (msg && msglen > 0) {
        if (e != cudaSuccess) snprintf(msg, (size_t)msglen, "%s: %s", what, cudaGetErrorString(e));
        else snprintf(msg, (size_t)msglen, "%s", what);
    }
    return rc;
}

extern "C" int zb_background_estimate_static(const float *data, const void *mask, int mask_dtype, long long n,
                                             double *out, void *stream, char *msg, int msglen)
{
    if (msg && msglen > 0) msg[0] = 0;
    if (!data || !out || n < 0) return bkg_static_fail(msg, msglen, ZB_ERR_PARAM, "zb_background_estimate_static: null data/out pointer or negative n", cudaSuccess);
    if (n > ZB_BKG_STATIC_MAX_PIXELS) {
        char buf[192];
        snprintf(buf, sizeof buf, "zb_background_estimate_static: %lld pixels exceed the maximum of %lld the static scratch's 32-bit counters are sized for",
                 n, (long long)ZB_BKG_STATIC_MAX_PIXELS);
        return bkg_static_fail(msg, msglen, ZB_ERR_PARAM, buf, cudaSuccess);
    }
    if (mask && !(mask_dtype == ZB_MASK_U8 || mask_dtype == ZB_MASK_I16 || mask_dtype == ZB_MASK_I32))
        return bkg_static_fail(msg, msglen, ZB_ERR_PARAM, "zb_background_estimate_static: bad mask dtype", cudaSuccess);
    if (stream == ZB_STREAM_OWN) return bkg_static_fail(msg, msglen, ZB_ERR_PARAM, "zb_background_estimate_static: ZB_STREAM_OWN needs a context; pass a stream (NULL = legacy default stream)", cudaSuccess);
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return bkg_static_fail(msg, msglen, ZB_ERR_CUDA, "cudaGetDevice", e);
    if (dev < 0 || dev >= BK_MAX_DEVICES) return bkg_static_fail(msg, msglen, ZB_ERR_UNSUPPORTED, "zb_background_estimate_static: device ordinal out of range", cudaSuccess);
    BkState *S;
    {
        // set-up, once per device, with that device current (it is: the calling thread's current device)
        std::lock_guard<std::mutex> lk(g_bkg_static_mu);
        if (!g_bkg_static_ptr[dev]) {
            void *p = NULL;
            e = cudaGetSymbolAddress(&p, g_bkg_static);
            if (e != cudaSuccess || !p) return bkg_static_fail(msg, msglen, ZB_ERR_CUDA, "cudaGetSymbolAddress(g_bkg_static)", e);
            g_bkg_static_ptr[dev] = (BkState *)p;
        }
        S = g_bkg_static_ptr[dev];
    }
    cudaStream_t st = stream ? (cudaStream_t)stream : cudaStreamLegacy;
    e = cudaMemsetAsync(S, 0, sizeof(BkState), st);            // prefixes, ranks, counts, ticket and histograms start at zero
    if (e != cudaSuccess) return bkg_static_fail(msg, msglen, ZB_ERR_CUDA, "cudaMemsetAsync of the static scratch", e);
    long long want = (n + BK_NT * 16 - 1) / (BK_NT * 16);
    const int grid = (int)(want < 1 ? 1 : (want > 148 * 8 ? 148 * 8 : want));
    k_bkg_pass<<<grid, BK_NT, 0, st>>>(data, mask, mask_dtype, n, S, 0, 53, 11, 1, 0);
    k_bkg_pass<<<grid, BK_NT, 0, st>>>(data, mask, mask_dtype, n, S, 0, 42, 11, 0, 0);
    k_bkg_pass<<<grid, BK_NT, 0, st>>>(data, mask, mask_dtype, n, S, 0, 32, 10, 0, 1);
    k_bkg_pass<<<grid, BK_NT, 0, st>>>(data, mask, mask_dtype, n, S, 1, 53, 11, 1, 0);
    k_bkg_pass<<<grid, BK_NT, 0, st>>>(data, mask, mask_dtype, n, S, 1, 42, 11, 0, 0);
    k_bkg_pass<<<grid, BK_NT, 0, st>>>(data, mask, mask_dtype, n, S, 1, 32, 10, 0, 1);
    k_bkg_emit<<<1, 1, 0, st>>>(S, out);
    e = cudaGetLastError();
    if (e != cudaSuccess) return bkg_static_fail(msg, msglen, ZB_ERR_CUDA, "zb_background_estimate_static: kernel launch", e);
    return ZB_OK;
}
