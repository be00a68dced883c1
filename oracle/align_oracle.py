"""CPU restatement of the alignment step (TEST INFRASTRUCTURE; only tests/ may import it).

What zuds runs SWarp for before every subtraction (zuds/swarp.py:107-204, zuds/fitsfile.py:290-314) with the shipped
configuration astromatic/makecoadd/default.swarp: RESAMPLING_TYPE LANCZOS3, OVERSAMPLING 0, FSCALASTRO_TYPE FIXED,
WEIGHT_TYPE NONE, SUBTRACT_BACK N, one input image per call.  SWarp (astromatic.net, unpinned by the reference: its
Dockerfile installs the distribution's package) is NOT in /root/reference, and there is no golden vector of its
output there either, so parity of zb_align_* is pinned to THIS restatement of the published algorithm only:

* target pixel centre -> sky through the target WCS -> fractional source pixel through the source WCS (gnomonic
  projection, CD matrix); a position within 1e-6 pixel of a pixel centre is snapped to it;
* 6 x 6 Lanczos-3 interpolation, L(d) = sinc(d) sinc(d/3), |d| < 3, each axis' six weights normalised to unit sum,
  times the fixed flux scale |det CD_dst| / |det CD_src|;
* value 0 / weight 0 where the footprint leaves the source frame; masks by nearest neighbour, no-data bit where the
  weight is 0 (zuds/mask.py:26-33).
"""
import numpy as np

D2R = np.pi / 180.0


def source_positions(src_wcs, dst_wcs, dny, dnx):
    """0-based fractional source coordinates (sx, sy) of every target pixel; wcs = dict(crpix1, crpix2, crval1, crval2, cd11..cd22)."""
    j, i = np.mgrid[0:dny, 0:dnx].astype(np.float64)
    px, py = i + 1 - dst_wcs["crpix1"], j + 1 - dst_wcs["crpix2"]
    xi = (dst_wcs["cd11"] * px + dst_wcs["cd12"] * py) * D2R
    eta = (dst_wcs["cd21"] * px + dst_wcs["cd22"] * py) * D2R
    s0, c0 = np.sin(dst_wcs["crval2"] * D2R), np.cos(dst_wcs["crval2"] * D2R)
    den = c0 - eta * s0
    ra = dst_wcs["crval1"] * D2R + np.arctan2(xi, den)
    dec = np.arctan2(s0 + eta * c0, np.sqrt(xi * xi + den * den))
    s1, c1 = np.sin(src_wcs["crval2"] * D2R), np.cos(src_wcs["crval2"] * D2R)
    sd, cd = np.sin(dec), np.cos(dec)
    sa, ca = np.sin(ra - src_wcs["crval1"] * D2R), np.cos(ra - src_wcs["crval1"] * D2R)
    cosc = s1 * sd + c1 * cd * ca
    xs = cd * sa / cosc / D2R
    es = (c1 * sd - s1 * cd * ca) / cosc / D2R
    det = src_wcs["cd11"] * src_wcs["cd22"] - src_wcs["cd12"] * src_wcs["cd21"]
    inv = (src_wcs["cd22"] / det, -src_wcs["cd12"] / det, -src_wcs["cd21"] / det, src_wcs["cd11"] / det)
    sx = inv[0] * xs + inv[1] * es + src_wcs["crpix1"] - 1.0
    sy = inv[2] * xs + inv[3] * es + src_wcs["crpix2"] - 1.0
    # positions within 1e-6 pixel of a pixel centre are that pixel centre (rounding of the sky round trip)
    sx = np.where(np.abs(sx - np.rint(sx)) < 1e-6, np.rint(sx), sx)
    sy = np.where(np.abs(sy - np.rint(sy)) < 1e-6, np.rint(sy), sy)
    return sx, sy


def lanczos3(d):
    d = np.asarray(d, np.float64)
    a = np.pi * d
    with np.errstate(divide="ignore", invalid="ignore"):
        v = (np.sin(a) / a) * (np.sin(a / 3.0) / (a / 3.0))
    v = np.where(d == 0.0, 1.0, v)
    return np.where(np.abs(d) >= 3.0, 0.0, v)


def align_lanczos3(src, src_wcs, dst_shape, dst_wcs):
    """(aligned float32 image, float32 weight map)."""
    sny, snx = src.shape
    dny, dnx = dst_shape
    sx, sy = source_positions(src_wcs, dst_wcs, dny, dnx)
    fx, fy = np.floor(sx), np.floor(sy)
    ok = (fx - 2 >= 0) & (fx + 3 <= snx - 1) & (fy - 2 >= 0) & (fy + 3 <= sny - 1)
    ix, iy = np.where(ok, fx, 2).astype(np.int64), np.where(ok, fy, 2).astype(np.int64)
    dx, dy = sx - fx, sy - fy
    wx = np.stack([lanczos3(dx - (k - 2)) for k in range(6)])
    wy = np.stack([lanczos3(dy - (k - 2)) for k in range(6)])
    acc = np.zeros((dny, dnx), np.float64)
    s64 = src.astype(np.float64)
    for b in range(6):
        r = np.zeros((dny, dnx), np.float64)
        for a in range(6):
            r += wx[a] * s64[iy - 2 + b, ix - 2 + a]
        acc += wy[b] * r
    det_s = src_wcs["cd11"] * src_wcs["cd22"] - src_wcs["cd12"] * src_wcs["cd21"]
    det_d = dst_wcs["cd11"] * dst_wcs["cd22"] - dst_wcs["cd12"] * dst_wcs["cd21"]
    out = acc / (wx.sum(0) * wy.sum(0)) * (abs(det_d) / abs(det_s))
    return np.where(ok, out, 0.0).astype(np.float32), ok.astype(np.float32)


def align_mask(mask, src_wcs, dst_shape, dst_wcs, nodata_bit=2 ** 16):
    sny, snx = mask.shape
    dny, dnx = dst_shape
    sx, sy = source_positions(src_wcs, dst_wcs, dny, dnx)
    fx, fy = np.floor(sx), np.floor(sy)
    ok = (fx - 2 >= 0) & (fx + 3 <= snx - 1) & (fy - 2 >= 0) & (fy + 3 <= sny - 1)
    nx_, ny_ = np.floor(sx + 0.5), np.floor(sy + 0.5)
    inside = (nx_ >= 0) & (nx_ <= snx - 1) & (ny_ >= 0) & (ny_ <= sny - 1)
    v = np.where(inside, mask.astype(np.int64)[np.where(inside, ny_, 0).astype(np.int64), np.where(inside, nx_, 0).astype(np.int64)], 0)
    return (v | np.where(ok, 0, nodata_bit)).astype(np.int32)
