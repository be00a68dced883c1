"""File-backed image objects: the part of zuds' FITS/DB object model that the subtraction path touches.

The reference's ``FITSImage`` / ``CalibratableImageBase`` / ``MaskImageBase`` (zuds/fitsfile.py:69-206,
zuds/image.py:136-208, zuds/mask.py:26-72) are SQLAlchemy-mapped records; the ORM, archive and WCS
machinery is out of scope here.  These classes keep the attribute names the path
uses (``data``, ``header``, ``header_comments``, ``local_path``, ``basename``, ``ismapped``,
``map_to_local_file``, ``save``, ``from_file``, ``mask_image``, ``rms_image``, ``weight_image``,
``boolean``, ``parent_image``, ``aligned_to``) so that ``prepare_hotpants`` and ``from_images`` read
like the reference's.
"""
from __future__ import annotations

import os

import numpy as np

from . import fitsio
from .synth import BAD_SUM, BIG_RMS

__all__ = ["FITSImage", "MaskImageBase", "CalibratableImageBase", "quick_background_estimate"]


def quick_background_estimate(image, nsamp=None, mask_image=None, ctx=None):
    """zuds/utils.py:32-53.  Median and 1.4826 x MAD of the pixels whose mask value is zero, evaluated in the
    pixels' dtype as the reference's numpy does (float32 medians and deviations; the 1.4826 factor in double,
    numpy 1.x scalar promotion).

    ``ctx`` (a ``_lib.Context``): run it on the GPU (``zb_background_estimate_host``: exact radix
    selection, same two numbers bit for bit); without one this is the reference's own numpy code."""
    if mask_image is None:
        mask_image = image.mask_image
    if ctx is not None and nsamp is None:
        return ctx.background_estimate(np.asarray(image.data), np.asarray(mask_image.data))
    bkgpix = np.asarray(image.data)[np.asarray(mask_image.data) == 0]
    if nsamp is not None:
        bkgpix = np.random.choice(bkgpix, size=nsamp)
    bkg = np.median(bkgpix)
    mad = np.median(np.abs(bkgpix - bkg))
    return float(bkg), 1.4826 * float(mad)


class FITSImage:
    """An array + header, optionally mapped to a FITS file on disk (zuds/fitsfile.py:69-206)."""

    def __init__(self, data=None, header=None, basename=None):
        self._data = data
        self.header = header if header is not None else fitsio.Header()
        self.header_comments = {}
        self.basename = basename
        self._path = None
        self.parent_image = None

    # -- disk mapping ---------------------------------------------------------
    @property
    def local_path(self):
        if self._path is None:
            raise FileNotFoundError("image %r is not mapped to a file" % self.basename)
        return self._path

    @property
    def ismapped(self):
        return self._path is not None

    def map_to_local_file(self, path):
        self._path = str(path)
        if self.basename is None:
            self.basename = os.path.basename(self._path)

    @classmethod
    def from_file(cls, path, **kw):
        obj = cls()
        obj.map_to_local_file(path)
        obj.basename = os.path.basename(str(path))
        obj.load()
        return obj

    def load(self):
        data, hdr = fitsio.read_image(self.local_path)
        self._data, self.header = data, hdr
        for key in ("NAXIS1", "NAXIS2"):
            if key not in self.header:
                self.header[key] = data.shape[1] if key == "NAXIS1" else data.shape[0]

    def save(self):
        data = np.asarray(self.data)
        if data.dtype == np.bool_:
            data = data.astype(np.uint8)              # boolean masks go to disk as 8-bit (zuds/fitsfile.py)
        elif data.dtype.kind == "f":
            data = data.astype(np.float32)
        elif data.dtype.itemsize == 8:
            data = data.astype(np.int32)
        hdr = self.header.copy()
        for k, c in self.header_comments.items():
            if k in hdr:
                hdr[k] = (hdr[k], c)
        fitsio.write_image(self.local_path, data, hdr)

    # -- data -----------------------------------------------------------------
    @property
    def data(self):
        if self._data is None and self.ismapped:
            self.load()
        return self._data

    @data.setter
    def data(self, value):
        self._data = value
        ny, nx = np.shape(value)
        self.header["NAXIS1"], self.header["NAXIS2"] = int(nx), int(ny)

    def aligned_to(self, other, tmpdir='/tmp', device=None):
        """This image on the pixel grid of ``other`` (zuds/fitsfile.py:290-314, where SWarp does it: LANCZOS3
        resampling of the image, zuds/swarp.py:107-204; the mask follows with bit 16 where nothing was covered,
        zuds/mask.py:26-33).  Runs on the GPU (``zb_align_lanczos3`` / ``zb_align_mask``) from the two headers' world
        coordinates (TAN projection).  Images without world coordinates that already share the grid pass through."""
        from . import params as P
        have_wcs = all(k in self.header and k in other.header for k in ("CRPIX1", "CRPIX2", "CRVAL1", "CRVAL2"))
        if not have_wcs:
            if np.shape(self.data) != np.shape(other.data):
                raise ValueError("images %s and %s have different shapes and no world coordinates to align them with"
                                 % (self.basename, other.basename))
            return self
        src_wcs, dst_wcs = P.ZbWcs.from_header(self.header), P.ZbWcs.from_header(other.header)
        same = np.shape(self.data) == np.shape(other.data) and all(
            getattr(src_wcs, f) == getattr(dst_wcs, f) for f, _ in P.ZbWcs._fields_)
        if same:
            return self                                       # already on the target grid: SWarp would copy it
        import torch
        from . import hotpants
        ctx = hotpants._context(hotpants.DEVICE if device is None else device)
        dev = torch.device("cuda", ctx.device)
        dst_shape = tuple(np.shape(other.data))
        out, wgt = ctx.align_lanczos3(torch.from_numpy(np.ascontiguousarray(self.data, np.float32)).to(dev), src_wcs, dst_shape, dst_wcs)
        new = type(self)() if not isinstance(self, MaskImageBase) else FITSImage()
        new.header = self.header.copy()
        for k in ("CRPIX1", "CRPIX2", "CRVAL1", "CRVAL2", "CD1_1", "CD1_2", "CD2_1", "CD2_2", "CDELT1", "CDELT2",
                  "PC1_1", "PC1_2", "PC2_1", "PC2_2", "CTYPE1", "CTYPE2"):
            new.header.remove(k)
            if k in other.header:
                new.header[k] = other.header[k]
        new.header_comments = dict(self.header_comments)
        new.data = out.cpu().numpy()
        new.basename = self.basename.replace(".fits", "_aligned_to_%s.remap.fits" % other.basename[:-5])
        new.parent_image = self
        new.weight_map = wgt.cpu().numpy()
        mask = getattr(self, "mask_image", None)
        if mask is not None:
            m = np.ascontiguousarray(mask.data)
            if m.dtype.name not in ("uint8", "int16", "int32"):
                m = m.astype(np.int32)
            md = ctx.align_mask(torch.from_numpy(m).to(dev), src_wcs, dst_shape, dst_wcs, nodata_bit=2 ** 16)
            nm = MaskImageBase(md.cpu().numpy(), mask.header.copy(), mask.basename.replace(".fits", "_aligned_to_%s.remap.fits" % other.basename[:-5]))
            nm.header_comments = dict(mask.header_comments)
            nm.parent_image = mask
            new.mask_image = nm
        return new


class MaskImageBase(FITSImage):
    """Integer bit mask (zuds/mask.py:26-72)."""

    @property
    def boolean(self):
        """True where a pixel is unusable: ``(data & BAD_SUM) > 0`` (zuds/mask.py:41-72)."""
        try:
            return self._boolean
        except AttributeError:
            b = FITSImage()
            b.data = (np.asarray(self.data).astype(np.int64) & BAD_SUM) > 0
            hdr = self.header.copy()
            b.header.cards = [c for c in hdr.cards if c[0] not in ("NAXIS1", "NAXIS2")] + \
                             [c for c in b.header.cards]
            b.header_comments = dict(self.header_comments)
            b.basename = self.basename.replace('.fits', '.bpm.fits')
            self._boolean = b
        return self._boolean

    def update_from_weight_map(self, weight_image):
        """Bit 16 where the (SWarp) weight is zero (zuds/mask.py:26-33)."""
        mskarr = np.asarray(self.data).copy()
        mskarr[np.asarray(weight_image.data) == 0] += 2 ** 16
        self.data = mskarr
        self.__dict__.pop("_boolean", None)


class CalibratableImageBase(FITSImage):
    """Science / reference image with a mask and an rms or weight map (zuds/image.py:136-208)."""

    mask_image = None

    def _sibling(self, suffix, data):
        im = FITSImage()
        im.basename = self.basename.replace('.fits', suffix)
        im.data = data
        im.header.cards = [c for c in self.header.cards if c[0] not in ("NAXIS1", "NAXIS2")] + list(im.header.cards)
        im.header_comments = dict(self.header_comments)
        if self.ismapped:
            im.map_to_local_file(os.path.join(os.path.dirname(self.local_path), im.basename))
            im.save()
        return im

    @property
    def weight_image(self):
        try:
            return self._weightimg
        except AttributeError:
            ind = np.asarray(self.mask_image.boolean.data, bool)
            wgt = np.empty(ind.shape, dtype='<f4')
            wgt[~ind] = 1 / np.asarray(self.rms_image.data)[~ind] ** 2
            wgt[ind] = 0.
            if 'SATURATE' in self.header:
                wgt[np.asarray(self.data) >= 0.9 * self.header['SATURATE']] = 0.
            self._weightimg = self._sibling('.weight.fits', wgt)
        return self._weightimg

    @weight_image.setter
    def weight_image(self, im):
        self._weightimg = im

    @property
    def rms_image(self):
        try:
            return self._rmsimg
        except AttributeError:
            if not hasattr(self, '_weightimg'):
                raise ValueError("image %s has neither an rms map nor a weight map (the reference would run "
                                 "SExtractor here, zuds/image.py:203-205)" % self.basename)
            ind = np.asarray(self.mask_image.boolean.data, bool)
            rms = np.empty(ind.shape, dtype='<f4')
            with np.errstate(divide="ignore"):
                rms[~ind] = 1 / np.sqrt(np.asarray(self.weight_image.data)[~ind])
            rms[ind] = BIG_RMS
            if 'SATURATE' in self.header:
                rms[np.asarray(self.data) >= 0.9 * self.header['SATURATE']] = BIG_RMS
            self._rmsimg = self._sibling('.rms.fits', rms)
        return self._rmsimg

    @rms_image.setter
    def rms_image(self, im):
        self._rmsimg = im

    # header-backed identifiers the reference copies into the subtraction (zuds/subtraction.py:193-201)
    def _hdr(self, key):
        return self.header.get(key)

    field = property(lambda self: self._hdr("FIELD"))
    ccdid = property(lambda self: self._hdr("CCDID"))
    qid = property(lambda self: self._hdr("QID"))
    fid = property(lambda self: self._hdr("FID"))

    @classmethod
    def from_file(cls, path, load_others=True, **kw):
        """Like the reference's ``from_file``: sibling ``.rms.fits`` / ``.weight.fits`` / ``.mask.fits``
        files in the same directory are attached when present (zuds/image.py:100-134)."""
        obj = super().from_file(path)
        if load_others:
            base = str(path)
            for suffix, attr, klass in (('.rms.fits', '_rmsimg', FITSImage), ('.weight.fits', '_weightimg', FITSImage),
                                        ('.mask.fits', 'mask_image', MaskImageBase)):
                sib = base.replace('.fits', suffix)
                if sib != base and os.path.exists(sib):
                    setattr(obj, attr, klass.from_file(sib))
        return obj
