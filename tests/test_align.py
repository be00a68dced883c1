"""Alignment of the reference onto the science grid: the CPU restatement of SWarp's LANCZOS3
resampling (oracle/align_oracle.py: parity is pinned to that restatement only, SWarp itself is not in the reference
tree) against known answers, and — under -m gpu — zb_align_lanczos3 / zb_align_mask against the restatement."""
import importlib.util
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_spec = importlib.util.spec_from_file_location("align_oracle", os.path.join(ROOT, "oracle", "align_oracle.py"))
AO = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(AO)

SCALE = 1.01 / 3600.0          # ZTF: 1.01 arcsec per pixel


def wcs(crpix=(100.0, 120.0), crval=(150.0, 30.0), rot=0.0, scale=SCALE, flip=True):
    c, s = np.cos(np.radians(rot)), np.sin(np.radians(rot))
    sx = -scale if flip else scale
    return dict(crpix1=crpix[0], crpix2=crpix[1], crval1=crval[0], crval2=crval[1],
                cd11=sx * c, cd12=-scale * s, cd21=sx * s, cd22=scale * c)


def image(ny=240, nx=200, seed=1):
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:ny, 0:nx]
    img = 150 + 3 * rng.normal(size=(ny, nx))
    for _ in range(30):
        cx, cy, a = rng.uniform(10, nx - 10), rng.uniform(10, ny - 10), rng.uniform(50, 3000)
        img += a * np.exp(-0.5 * ((x - cx) ** 2 + (y - cy) ** 2) / 1.7 ** 2)
    return img.astype(np.float32)


def test_oracle_integer_shift_is_a_copy():
    img = image()
    w0 = wcs()
    w1 = wcs(crpix=(100.0 - 7, 120.0 + 4))            # the same sky on a grid shifted by (+7, -4) pixels
    out, wgt = AO.align_lanczos3(img, w0, img.shape, w1)
    ny, nx = img.shape
    ok = wgt > 0
    want = np.zeros_like(img)
    want[4:ny, 0:nx - 7] = img[0:ny - 4, 7:nx]         # target (i, j) looks at source (i + 7, j - 4)
    assert ok.sum() > 0.8 * img.size
    # positions snap to pixel centres, where the Lanczos weights are (0,0,1,0,0,0) up to sin(k pi) ~ 1e-16
    assert np.abs(out[ok] - want[ok]).max() <= 1e-6 * np.abs(want).max()
    assert not out[~ok].any()
    m = (np.arange(img.size).reshape(img.shape) % 7 == 0).astype(np.int16) * 5
    mo = AO.align_mask(m, w0, img.shape, w1)
    assert np.array_equal(mo[ok] & 0xffff, want_mask(m, 7, 4)[ok]) and ((mo[~ok] & 2 ** 16) != 0).all()


def want_mask(m, dx, dy):
    ny, nx = m.shape
    w = np.zeros((ny, nx), np.int32)
    w[dy:ny, 0:nx - dx] = m[0:ny - dy, dx:nx]
    return w


def test_oracle_half_pixel_shift_uses_symmetric_weights():
    img = image()
    out, wgt = AO.align_lanczos3(img, wcs(), img.shape, wcs(crpix=(100.0 - 0.5, 120.0)))
    w = AO.lanczos3(0.5 - (np.arange(6) - 2.0))
    w /= w.sum()
    assert np.allclose(w, w[::-1])                      # symmetric about the half pixel
    j, i = 100, 50
    want = float((w * img[j, i - 2:i + 4].astype(np.float64)).sum())
    assert wgt[j, i] == 1 and abs(out[j, i] - want) <= 1e-5 * abs(want)


def test_oracle_conserves_flux_under_rescaling():
    img = image() - 150.0
    out, wgt = AO.align_lanczos3(img, wcs(), (120, 100), wcs(crpix=(50.5, 60.5), scale=2 * SCALE))
    inner = wgt > 0
    # target pixels are 4x larger: the fixed flux scale keeps surface brightness x area
    assert abs(out[inner].sum() / img.sum() - 1.0) < 0.05


@pytest.mark.gpu
@pytest.mark.parametrize("kind", ["shift", "half", "rotate", "rescale", "offfield"])
def test_gpu_alignment_matches_the_restatement(gpu_ctx, kind):
    import torch
    from zuds_pipeline_b200 import params as P
    img = image(300, 280, seed=5)
    w0 = wcs(crpix=(140.0, 150.0))
    w1, shape = {"shift": (wcs(crpix=(140.0 - 7, 150.0 + 4)), img.shape),
                 "half": (wcs(crpix=(139.5, 150.25)), img.shape),
                 "rotate": (wcs(crpix=(128.0, 131.0), rot=3.0, crval=(150.002, 30.001)), (260, 256)),
                 "rescale": (wcs(crpix=(70.0, 75.0), scale=2 * SCALE), (150, 140)),
                 "offfield": (wcs(crpix=(140.0, 150.0), crval=(151.0, 30.0)), img.shape)}[kind]
    want, wwant = AO.align_lanczos3(img, w0, shape, w1)

    def zw(d):
        z = P.ZbWcs()
        for k, v in d.items():
            setattr(z, k, v)
        return z
    out, wgt = gpu_ctx.align_lanczos3(torch.from_numpy(img).cuda(), zw(w0), shape, zw(w1))
    out, wgt = out.cpu().numpy(), wgt.cpu().numpy()
    assert np.array_equal(wgt, wwant)
    if kind == "offfield":
        assert not wgt.any() and not out.any()
    assert np.abs(out - want).max() <= 2e-6 * max(1.0, np.abs(want).max())
    m = ((np.arange(img.size).reshape(img.shape) * 2654435761 % 97) < 3).astype(np.int16) * 4 + 1
    mo = gpu_ctx.align_mask(torch.from_numpy(m).cuda(), zw(w0), shape, zw(w1)).cpu().numpy()
    mw = AO.align_mask(m, w0, shape, w1)
    # nearest-neighbour ties (a target centre within rounding of a source pixel edge) may fall either way
    assert (mo != mw).mean() < 1e-4 and np.array_equal(mo & 2 ** 16, mw & 2 ** 16)


@pytest.mark.gpu
def test_aligned_to_resamples_a_shifted_reference(tmp_path):
    from zuds_pipeline_b200 import fitsio, image as IM
    img = image(200, 180, seed=9)
    hs, hr = fitsio.Header(), fitsio.Header()
    for h, w in ((hs, wcs(crpix=(90.0, 100.0))), (hr, wcs(crpix=(93.0, 98.0)))):
        for k, v in (("CTYPE1", "RA---TAN"), ("CTYPE2", "DEC--TAN"), ("CRPIX1", w["crpix1"]), ("CRPIX2", w["crpix2"]),
                     ("CRVAL1", w["crval1"]), ("CRVAL2", w["crval2"]), ("CD1_1", w["cd11"]), ("CD1_2", w["cd12"]),
                     ("CD2_1", w["cd21"]), ("CD2_2", w["cd22"])):
            h[k] = v
    sci = IM.CalibratableImageBase(img.copy(), hs, "sci.fits")
    ref = IM.CalibratableImageBase(img.copy(), hr, "ref.fits")
    ref.mask_image = IM.MaskImageBase(np.zeros(img.shape, np.int16), hr.copy(), "ref.mask.fits")
    al = ref.aligned_to(sci)
    ok = al.weight_map > 0
    # the reference grid is the science grid shifted by (+3, -2) pixels: sci pixel (i, j) sees ref pixel (i + 3, j - 2)
    want = np.zeros_like(img)
    want[2:, :-3] = img[:-2, 3:]
    assert ok.sum() > 0.8 * img.size and np.abs(al.data[ok] - want[ok]).max() <= 1e-6 * np.abs(want).max()
    assert al.header["CRPIX1"] == hs["CRPIX1"] and ((al.mask_image.data[~ok] & 2 ** 16) != 0).all()
    assert ref.aligned_to(ref) is ref
