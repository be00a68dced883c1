// zb_align.cu — alignment of the reference image (and its mask) onto the science image's pixel grid, the step
// immediately before the subtraction path.  The reference shells out to SWarp once per image
// and once per mask (zuds/swarp.py:107-204, zuds/fitsfile.py:290-314) with astromatic/makecoadd/default.swarp:
// RESAMPLING_TYPE LANCZOS3, OVERSAMPLING 0, FSCALASTRO_TYPE FIXED, WEIGHT_TYPE NONE, SUBTRACT_BACK N, one input
// image.  SWarp's source is not part of the reference tree; this is a restatement of its documented resampling
// for that configuration (oracle/align_oracle.py restates it on the CPU and the tests compare the two):
//   * every output pixel centre is carried to the sky through the target's WCS and back to a (fractional) pixel
//     position of the source image through the source's WCS: gnomonic (TAN) projection with a CD matrix (positions
//     within 1e-6 pixel of a pixel centre are snapped to it);
//   * the value is the 6 x 6 Lanczos-3 interpolation around that position, L(d) = sinc(d) sinc(d/3) for |d| < 3,
//     the six weights of each axis normalised to unit sum, times the fixed flux scale |det CD_dst| / |det CD_src|;
//   * an output pixel whose 6 x 6 footprint is not entirely inside the source frame gets value 0 and weight 0
//     (weight 1 otherwise): the weight map zuds turns into mask bit 16 (zuds/mask.py:26-33);
//   * masks are carried by nearest neighbour (flag bits cannot be interpolated) and receive `nodata_bit` where the
//     weight is 0.
// A gather stencil: HBM-bound for the output (4 B + 4 B per pixel written) with the 36 source reads served by
// L1/L2; the transcendental part (2 x sincos, atan2, sqrt per pixel in double) is what costs, ~1 ms per quadrant.
#include "zb_internal.h"
#include <math.h>

#define ZB_D2R 0.017453292519943295769
#define ZB_PI  3.14159265358979323846

struct AlignMap {
    zb_wcs src, dst;
    double isrc[4];      // inverse CD matrix of the source (pixels per degree)
    double flux_scale;
};

// output pixel (0-based i, j) -> source position (0-based, fractional)
__device__ static inline void align_map(const AlignMap &M, int i, int j, double *sx, double *sy)
{
    const double px = (double)(i + 1) - M.dst.crpix1, py = (double)(j + 1) - M.dst.crpix2;
    const double xi = (M.dst.cd11 * px + M.dst.cd12 * py) * ZB_D2R, eta = (M.dst.cd21 * px + M.dst.cd22 * py) * ZB_D2R;
    double s0, c0;
    sincos(M.dst.crval2 * ZB_D2R, &s0, &c0);
    const double den = c0 - eta * s0;
    const double ra = M.dst.crval1 * ZB_D2R + atan2(xi, den);
    const double dec = atan2(s0 + eta * c0, sqrt(xi * xi + den * den));
    double s1, c1, sd, cd, sa, ca;
    sincos(M.src.crval2 * ZB_D2R, &s1, &c1);
    sincos(dec, &sd, &cd);
    sincos(ra - M.src.crval1 * ZB_D2R, &sa, &ca);
    const double cosc = s1 * sd + c1 * cd * ca;
    const double xs = cd * sa / cosc / ZB_D2R, es = (c1 * sd - s1 * cd * ca) / cosc / ZB_D2R;      // degrees on the source tangent plane
    double x = M.isrc[0] * xs + M.isrc[1] * es + M.src.crpix1 - 1.0, y = M.isrc[2] * xs + M.isrc[3] * es + M.src.crpix2 - 1.0;
    // the round trip through the sky leaves ~1e-10 pixel of rounding: a position that close to a pixel centre IS the
    // pixel centre (grids offset by whole pixels then give an exact copy, and the footprint test cannot flip on noise)
    const double rx = rint(x), ry = rint(y);
    if (fabs(x - rx) < 1e-6) x = rx;
    if (fabs(y - ry) < 1e-6) y = ry;
    *sx = x; *sy = y;
}

__device__ static inline double lanczos3(double d)
{
    if (d == 0.0) return 1.0;
    if (fabs(d) >= 3.0) return 0.0;
    const double a = ZB_PI * d, b = a / 3.0;
    return (sin(a) / a) * (sin(b) / b);
}

__global__ void __launch_bounds__(256)
k_align_lanczos3(const float *__restrict__ src, int snx, int sny, AlignMap M, int dnx, int dny,
                 float *__restrict__ out, float *__restrict__ wout)
{
    const int i = blockIdx.x * 32 + (threadIdx.x & 31), j = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (i >= dnx || j >= dny) return;
    double sx, sy;
    align_map(M, i, j, &sx, &sy);
    const double fx = floor(sx), fy = floor(sy);
    const long long gi = (long long)j * dnx + i;
    // the 6 x 6 footprint ix-2 .. ix+3, iy-2 .. iy+3 must lie inside the source (NaN positions fail the test too)
    if (!(fx - 2.0 >= 0.0 && fx + 3.0 <= (double)(snx - 1) && fy - 2.0 >= 0.0 && fy + 3.0 <= (double)(sny - 1))) {
        out[gi] = 0.f;
        if (wout) wout[gi] = 0.f;
        return;
    }
    const int ix = (int)fx, iy = (int)fy;
    const double dx = sx - fx, dy = sy - fy;
    double wx[6], wy[6], sumx = 0.0, sumy = 0.0;
#pragma unroll
    for (int k = 0; k < 6; k++) {
        wx[k] = lanczos3(dx - (double)(k - 2)); wy[k] = lanczos3(dy - (double)(k - 2));
        sumx += wx[k]; sumy += wy[k];
    }
    double acc = 0.0;
#pragma unroll
    for (int b = 0; b < 6; b++) {
        const float *row = src + (long long)(iy - 2 + b) * snx + (ix - 2);
        double r = 0.0;
#pragma unroll
        for (int a = 0; a < 6; a++) r += wx[a] * (double)row[a];
        acc += wy[b] * r;
    }
    out[gi] = (float)(acc / (sumx * sumy) * M.flux_scale);
    if (wout) wout[gi] = 1.f;
}

__global__ void __launch_bounds__(256)
k_align_mask(const void *__restrict__ src, int dt, int snx, int sny, AlignMap M, int dnx, int dny,
             int32_t *__restrict__ out, int nodata_bit)
{
    const int i = blockIdx.x * 32 + (threadIdx.x & 31), j = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (i >= dnx || j >= dny) return;
    double sx, sy;
    align_map(M, i, j, &sx, &sy);
    const double fx = floor(sx), fy = floor(sy);
    int v = 0;
    const double nx_ = floor(sx + 0.5), ny_ = floor(sy + 0.5);
    if (nx_ >= 0.0 && nx_ <= (double)(snx - 1) && ny_ >= 0.0 && ny_ <= (double)(sny - 1))
        v = zb_read_mask(src, dt, (long long)ny_ * snx + (long long)nx_);
    // the same footprint test as the image: weight 0 there -> the no-data bit (zuds/mask.py:26-33)
    if (!(fx - 2.0 >= 0.0 && fx + 3.0 <= (double)(snx - 1) && fy - 2.0 >= 0.0 && fy + 3.0 <= (double)(sny - 1))) v |= nodata_bit;
    out[(long long)j * dnx + i] = v;
}

static int make_map(zb_ctx *ctx, const zb_wcs *s, const zb_wcs *d, AlignMap *M)
{
    const double det = s->cd11 * s->cd22 - s->cd12 * s->cd21, detd = d->cd11 * d->cd22 - d->cd12 * d->cd21;
    if (!(fabs(det) > 0.0) || !(fabs(detd) > 0.0)) { snprintf(ctx->err, sizeof ctx->err, "singular CD matrix"); return ZB_ERR_PARAM; }
    M->src = *s; M->dst = *d;
    M->isrc[0] = s->cd22 / det; M->isrc[1] = -s->cd12 / det; M->isrc[2] = -s->cd21 / det; M->isrc[3] = s->cd11 / det;
    M->flux_scale = fabs(detd) / fabs(det);
    return ZB_OK;
}

extern "C" int zb_align_lanczos3(zb_ctx *ctx, const float *src, int snx, int sny, const zb_wcs *src_wcs,
                                 int dnx, int dny, const zb_wcs *dst_wcs, float *out, float *out_weight, void *stream)
{
    if (!ctx || !src || !src_wcs || !dst_wcs || !out || snx < 6 || sny < 6 || dnx < 1 || dny < 1) { if (ctx) snprintf(ctx->err, sizeof ctx->err, "bad argument"); return ZB_ERR_PARAM; }
    if (cudaSetDevice(ctx->device) != cudaSuccess) return zb_cuda_fail(ctx, cudaGetLastError(), "cudaSetDevice", __FILE__, __LINE__);
    AlignMap M;
    int rc = make_map(ctx, src_wcs, dst_wcs, &M);
    if (rc != ZB_OK) return rc;
    cudaStream_t st = zb_pick_stream(ctx, stream);
    k_align_lanczos3<<<dim3((dnx + 31) / 32, (dny + 7) / 8), 256, 0, st>>>(src, snx, sny, M, dnx, dny, out, out_weight);
    ctx->launches++;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return zb_cuda_fail(ctx, e, "k_align_lanczos3", __FILE__, __LINE__);
    return ZB_OK;
}

extern "C" int zb_align_mask(zb_ctx *ctx, const void *src_mask, int dtype, int snx, int sny, const zb_wcs *src_wcs,
                             int dnx, int dny, const zb_wcs *dst_wcs, int32_t *out, int nodata_bit, void *stream)
{
    if (!ctx || !src_mask || !src_wcs || !dst_wcs || !out || snx < 6 || sny < 6 || dnx < 1 || dny < 1) { if (ctx) snprintf(ctx->err, sizeof ctx->err, "bad argument"); return ZB_ERR_PARAM; }
    if (!(dtype == ZB_MASK_U8 || dtype == ZB_MASK_I16 || dtype == ZB_MASK_I32)) { snprintf(ctx->err, sizeof ctx->err, "bad mask dtype %d", dtype); return ZB_ERR_PARAM; }
    if (cudaSetDevice(ctx->device) != cudaSuccess) return zb_cuda_fail(ctx, cudaGetLastError(), "cudaSetDevice", __FILE__, __LINE__);
    AlignMap M;
    int rc = make_map(ctx, src_wcs, dst_wcs, &M);
    if (rc != ZB_OK) return rc;
    cudaStream_t st = zb_pick_stream(ctx, stream);
    k_align_mask<<<dim3((dnx + 31) / 32, (dny + 7) / 8), 256, 0, st>>>(src_mask, dtype, snx, sny, M, dnx, dny, out, nodata_bit);
    ctx->launches++;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return zb_cuda_fail(ctx, e, "k_align_mask", __FILE__, __LINE__);
    return ZB_OK;
}
