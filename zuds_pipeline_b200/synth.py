"""Seeded synthetic ZTF-like quadrant pairs.

No pixel data ship with the reference (zuds/tests/conftest.py:14-21 downloads
its cutouts), so every parity test, golden vector and benchmark in this repo
uses pairs made here.  All arrays are float32 C-order ``(ny, nx)``; a full ZTF
quadrant is ``(3080, 3072)`` (zuds/tests/fixtures.py:82-83).

The generator reproduces what reaches hotpants on the zuds path:
  * science image: background-subtracted + ``BKG_VAL`` = 150 (zuds/hotpants.py:27-30)
  * reference image already aligned to the science grid (zuds/subtraction.py:109)
  * per-pixel rms maps, ``BIG_RMS`` where masked (zuds/image.py:179-190)
  * boolean bad-pixel mask ``(scimask | refmask) & BAD_SUM > 0``
    (zuds/subtraction.py:135-142, zuds/mask.py:65)
"""
from __future__ import annotations

import dataclasses
import numpy as np

BKG_VAL = 150.0                       # zuds/constants.py:23
BIG_RMS = float(np.sqrt(50000.0))     # zuds/constants.py:3
BAD_BITS = (0, 2, 3, 4, 5, 7, 8, 9, 10, 16, 17)   # zuds/constants.py:45
BAD_SUM = int(sum(2 ** b for b in BAD_BITS))      # 202749

FWHM2SIG = 1.0 / 2.3548200450309493


@dataclasses.dataclass
class Pair:
    sci: np.ndarray        # float32 (ny, nx)  science, sky = 150
    ref: np.ndarray        # float32           aligned reference, sky = 150
    sci_rms: np.ndarray    # float32
    ref_rms: np.ndarray    # float32
    sci_mask: np.ndarray   # int16   ZTF-style bits
    ref_mask: np.ndarray   # int32   bit 16 = no data after alignment
    bpm: np.ndarray        # uint8   (sci_mask | ref_mask) & BAD_SUM > 0
    seeing: float          # FWHM in pixels (header SEEING)
    seed: int


def _add_stars(img, xs, ys, peaks, sig_of):
    ny, nx = img.shape
    for x, y, p in zip(xs, ys, peaks):
        s = sig_of(x, y)
        r = int(np.ceil(5.0 * s)) + 1
        x0, x1 = max(0, int(x) - r), min(nx, int(x) + r + 1)
        y0, y1 = max(0, int(y) - r), min(ny, int(y) + r + 1)
        if x0 >= x1 or y0 >= y1:
            continue
        gx = np.exp(-0.5 * ((np.arange(x0, x1) - x) / s) ** 2)
        gy = np.exp(-0.5 * ((np.arange(y0, y1) - y) / s) ** 2)
        img[y0:y1, x0:x1] += p * gy[:, None] * gx[None, :]


def make_pair(seed=1000, shape=(3080, 3072), seeing=4.0, nstars=None, ntransients=20,
              gain=6.0, ref_noise=2.5, flux_ratio=1.05, bad_frac=1e-3, ncols_bad=6, border=12,
              ref_sharpness=0.8, seeing_gradient=0.04):
    """One synthetic (science, reference) pair.  Deterministic in ``seed``.

    ``ref_sharpness``: reference PSF width as a fraction of the science seeing; ``seeing_gradient``: relative change
    of the science PSF width from the frame centre to its left/right edge (a large value with ``ref_sharpness`` = 1
    makes the science image the sharper one on the left and the broader one on the right: the -c b cases)."""
    rng = np.random.default_rng(seed)
    ny, nx = shape
    if nstars is None:
        nstars = max(8, int(round(4000.0 * (nx * ny) / (3072.0 * 3080.0))))

    xs = rng.uniform(0, nx - 1, nstars)
    ys = rng.uniform(0, ny - 1, nstars)
    peaks = np.exp(rng.uniform(np.log(60.0), np.log(6000.0), nstars))

    sig_ref = ref_sharpness * seeing * FWHM2SIG

    def sig_sci(x, y):
        u = 2.0 * x / max(nx - 1, 1) - 1.0
        v = 2.0 * y / max(ny - 1, 1) - 1.0
        return seeing * (1.0 + seeing_gradient * u + 0.03 * v * v) * FWHM2SIG

    ref_sig = np.zeros(shape, np.float64)
    _add_stars(ref_sig, xs, ys, peaks, lambda x, y: sig_ref)
    sci_sig = np.zeros(shape, np.float64)
    sci_peaks = [p * flux_ratio * (sig_ref / sig_sci(x, y)) ** 2 for x, y, p in zip(xs, ys, peaks)]
    _add_stars(sci_sig, xs, ys, sci_peaks, sig_sci)

    # transients: in the science image only
    tx = rng.uniform(30, nx - 31, ntransients)
    ty = rng.uniform(30, ny - 31, ntransients)
    tp = np.exp(rng.uniform(np.log(40.0), np.log(800.0), ntransients))
    _add_stars(sci_sig, tx, ty, tp, sig_sci)

    ref_rms = np.full(shape, ref_noise, np.float64)
    sci_rms = np.sqrt((BKG_VAL + sci_sig) / gain)
    ref = BKG_VAL + ref_sig + ref_noise * rng.standard_normal(shape)
    sci = BKG_VAL + sci_sig + sci_rms * rng.standard_normal(shape)

    # masks (ZTF bit meanings: zuds/constants.py:48-66)
    sci_mask = np.zeros(shape, np.int16)
    bad = rng.random(shape) < bad_frac
    which = rng.random(shape) < 0.5
    sci_mask[bad & which] |= 1 << 2          # low responsivity
    sci_mask[bad & ~which] |= 1 << 9         # dead
    for c in rng.integers(border + 2, max(border + 3, nx - border - 2), ncols_bad):
        sci_mask[:, int(c)] |= 1 << 4        # noisy column
    sci_mask[sci_sig > 30.0] |= 1 << 1       # SExtractor detection: NOT a bad bit
    sci_mask[(sci_sig + BKG_VAL) >= 5000.0] |= 1 << 8   # saturated cores
    dead = (sci_mask & (1 << 9)) != 0
    sci[dead] = BKG_VAL * 0.02               # dead pixels read low

    ref_mask = np.zeros(shape, np.int32)
    edge = np.zeros(shape, bool)
    wob = (border + 3.0 * np.sin(np.arange(max(nx, ny)) / 37.0)).astype(int)
    for x in range(nx):
        edge[:wob[x], x] = True
        edge[ny - wob[x]:, x] = True
    for y in range(ny):
        edge[y, :wob[y]] = True
        edge[y, nx - wob[y]:] = True
    ref_mask[edge] |= 1 << 16                # SWarp no-data section
    ref[edge] = 0.0

    bpm = (((sci_mask.astype(np.int32) | ref_mask) & BAD_SUM) > 0)
    sci_rms = np.where(((sci_mask.astype(np.int32) & BAD_SUM) > 0), BIG_RMS, sci_rms)
    ref_rms = np.where(((ref_mask & BAD_SUM) > 0), BIG_RMS, ref_rms)

    return Pair(sci=sci.astype(np.float32), ref=ref.astype(np.float32),
                sci_rms=sci_rms.astype(np.float32), ref_rms=ref_rms.astype(np.float32),
                sci_mask=sci_mask, ref_mask=ref_mask, bpm=bpm.astype(np.uint8),
                seeing=float(seeing), seed=int(seed))
