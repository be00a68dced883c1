"""Drop-in for ``zuds.subtraction``: ``sub_name`` and ``Subtraction.from_images``.

Same call signature, keyword arguments (``subtract_back``, ``nreg_side``, ``hotpants_kws``), file products
(``sub.<sci>_<ref>.fits``, ``.rms.fits``, ``.mask.fits``), header keywords and mask-bit semantics as
zuds/subtraction.py:25-226.  What differs is everything in between: the reference writes its inputs to a scratch
directory, fork/execs ``hotpants`` on them (zuds/subtraction.py:162) and reads the products back; here the images
are uploaded ONCE, every step runs on tensors resident in HBM (``subtract_on_device``: mask OR / boolean, rms from
weight, background estimates, the subtraction itself, the bit-17 flag), and the products come back ONCE to be
written as FITS.  No Postgres records, no archive copy (out of scope).
"""
from __future__ import annotations

import os
import shutil
import uuid
from pathlib import Path

import numpy as np

from . import fitsio, params as P
from .image import CalibratableImageBase, FITSImage, MaskImageBase
from .synth import BAD_SUM, BIG_RMS, BKG_VAL

__all__ = ['sub_name', 'subtract_on_device', 'Subtraction', 'SingleEpochSubtraction', 'MultiEpochSubtraction']

APER_KEY = 'APCOR4'           # zuds/constants.py
HOTPANTS_BIT = 17             # zuds/subtraction.py:170-176
SENTINEL = np.float32(1e-30)  # what hotpants writes into masked difference pixels (-fi default)


def sub_name(frame, template):
    """zuds/subtraction.py:25-37."""
    frame = f'{frame}'
    template = f'{template}'
    refp = os.path.basename(template)[:-5]
    newp = os.path.basename(frame)[:-5]
    outdir = os.path.dirname(frame)
    subp = '_'.join([newp, refp])
    return os.path.join(outdir, 'sub.%s.fits' % subp)


# ------------------------------------------------------------------------------------------ device path

class DeviceProducts:
    """What ``subtract_on_device`` leaves in HBM: ``diff``, ``noise`` (float32), ``mask`` (int32: refmask | scimask |
    bit 17 where hotpants masked the difference), ``boolean`` (uint8, the -imi image); plus ``command`` (the hotpants
    command line zuds would have run), ``args`` (it parsed) and ``result`` (per-region QA, kernel table, stamps)."""


def subtract_on_device(ctx, sci, ref, sci_rms, ref_rms, sci_mask, ref_mask, seeing, nreg_side=3, hotpants_kws=None,
                       names=None, stream=None):
    """The arithmetic of ``Subtraction.from_images`` on torch CUDA tensors, nothing returning to the host in between.

    zuds/subtraction.py:135 (mask OR), zuds/mask.py:42-72 (boolean), zuds/hotpants.py:65-73 (background levels ->
    -il/-tl), :75-93 (command line), zuds/subtraction.py:162 (hotpants), :167-177 (bit 17).  ``names``: the file
    names that go into the command line (and so into the DIFFCMD keyword); placeholders when None."""
    hotpants_kws = hotpants_kws or {}
    ny, nx = sci.shape
    badpix, boolean = ctx.mask_combine(sci_mask, ref_mask, BAD_SUM, stream=stream)
    scibkg, scistd = ctx.background_estimate(sci, sci_mask)
    refbkg, refstd = ctx.background_estimate(ref, ref_mask)
    il, tl = scibkg - 10 * scistd, refbkg - 10 * refstd
    n = names or {}
    command = P.hotpants_command(n.get('sci', 'sci.fits'), n.get('ref', 'ref.fits'), n.get('out', 'sub.fits'),
                                 n.get('sci_rms', 'sci.rms.fits'), n.get('ref_rms', 'ref.rms.fits'), n.get('bpm', 'sub.mask.bpm.fits'),
                                 seeing, nx, ny, il, tl, nreg_side=nreg_side, hotpants_kws=hotpants_kws)
    args = P.parse_hotpants_args(command.split()[1:])
    args.params.nx, args.params.ny = nx, ny
    res = ctx.subtract(args.params, sci, ref, sci_rms, ref_rms, boolean, None, want_mask=False, stream=stream)
    ctx.mask_flag_sentinel(badpix, res.diff, SENTINEL, 2 ** HOTPANTS_BIT, stream=stream)
    out = DeviceProducts()
    out.diff, out.noise, out.mask, out.boolean = res.diff, res.noise, badpix, boolean
    out.command, out.args, out.result = command, args, res
    return out


class _Scratch:
    """Per-call private directory, removed on exit whatever happens (zuds/subtraction.py:68,224)."""

    def __init__(self, tmpdir):
        self.path = Path(tmpdir) / uuid.uuid4().hex

    def __enter__(self):
        self.path.mkdir(exist_ok=True, parents=True)
        return self

    def __exit__(self, *exc):
        shutil.rmtree(self.path, ignore_errors=True)
        return False

    def adopt(self, image, need_noise_map):
        """A private copy of ``image`` and its companions, reloaded from the copies: nothing another process does to
        the originals while this subtraction runs can reach it (zuds/subtraction.py:68-99)."""
        files = [image.local_path, image.mask_image.local_path]
        maps = [getattr(image, a) for a in ('_rmsimg', '_weightimg') if hasattr(image, a)]
        if need_noise_map and not maps:
            raise ValueError('Science image must have a weight map or rms map defined prior to subtraction.')
        files += [m.local_path for m in maps]
        for f in files:
            shutil.copy(f, self.path)

        def reopened(src, klass):
            """``src`` as an object of ``klass`` mapped to the private copy.  Pixels and header already in memory are
            taken over as they are (what was loaded from the original a moment ago is what the copy holds);
            an image that was never loaded is read from the copy."""
            path = os.path.join(self.path, src.basename)
            if src._data is None:
                return klass.from_file(path, load_others=False) if klass is CalibratableImageBase else klass.from_file(path)
            obj = klass()
            obj.map_to_local_file(path)
            obj.basename = src.basename
            obj._data, obj.header = src._data, src.header.copy()
            obj.header_comments = dict(src.header_comments)
            return obj
        mine = reopened(image, CalibratableImageBase)
        mine.mask_image = reopened(image.mask_image, MaskImageBase)
        for attr in ('_rmsimg', '_weightimg'):
            if hasattr(image, attr):
                setattr(mine, attr, reopened(getattr(image, attr), FITSImage))
        return mine


def _noise_map_on_device(ctx, image, data_dev, up):
    """``image.rms_image`` as a device tensor + the file name zuds would have given it: the rms map itself when the
    image has one, else 1/sqrt(weight) with BIG_RMS on bad / near-saturated pixels (zuds/image.py:179-190)."""
    name = image.local_path.replace('.fits', '.rms.fits')
    if hasattr(image, '_rmsimg'):
        return up(np.asarray(image._rmsimg.data, np.float32)), image._rmsimg.local_path
    if not hasattr(image, '_weightimg'):
        raise ValueError("image %s has neither an rms map nor a weight map" % image.basename)
    _, own_bad = ctx.mask_combine(up(np.asarray(image.mask_image.data)), None, BAD_SUM, want_or=False)
    rms = ctx.rms_from_weight(up(np.asarray(image._weightimg.data, np.float32)), bad=own_bad, data=data_dev,
                              saturate=image.header.get('SATURATE'), big_rms=BIG_RMS)
    return rms, name


class Subtraction(CalibratableImageBase):
    """A difference image with its rms map, mask, target and reference."""

    reference_image = None
    target_image = None
    device = 0                  # CUDA device used by from_images
    last_result = None          # _lib.Result of the most recent from_images (per-region QA, kernel table, stamps)

    @property
    def mjd(self):
        return self.target_image.header.get('OBSMJD') if self.target_image is not None else None

    @classmethod
    def from_images(cls, sci, ref, data_product=False, tmpdir='/tmp', **kwargs):
        import torch
        from . import hotpants

        subtract_back = kwargs.get('subtract_back', True)
        nreg_side = kwargs.get('nreg_side', 3)
        hotpants_kws = kwargs.get('hotpants_kws', {})
        if data_product:
            raise NotImplementedError("archive() (HTTP/tape copy, zuds/archive.py) is outside this path")
        if 'SEEING' not in sci.header:
            raise KeyError("SEEING (FWHM in pixels) must be in the science header; zuds.seeing.estimate_seeing "
                           "queries external catalogues and is outside this path")

        ctx = hotpants._context(cls.device)
        dev = torch.device('cuda', cls.device)

        def up(a):
            a = np.ascontiguousarray(a)
            if a.dtype == np.bool_:
                a = a.view(np.uint8)
            elif a.dtype == np.uint16:
                a = a.astype(np.int32)
            elif a.dtype.kind == 'f' and a.dtype != np.float32:
                a = a.astype(np.float32)
            elif a.dtype.kind in 'iu' and a.dtype.name not in ('uint8', 'int16', 'int32'):
                a = a.astype(np.int32)
            return torch.from_numpy(a).to(dev)

        with _Scratch(tmpdir) as ws:
            my_sci = ws.adopt(sci, need_noise_map=True)
            my_ref = ws.adopt(ref, need_noise_map=False)
            aligned = my_ref.aligned_to(my_sci, tmpdir=tmpdir)          # SWarp in the reference (:109)
            aligned.parent_image = my_ref
            outname = sub_name(my_sci.local_path, my_ref.local_path)

            # SExtractor background subtraction of the science image (zuds/hotpants.py:31-40), a hook upstream of the path
            target = my_sci
            if subtract_back:
                if hotpants.background_subtractor is None:
                    raise NotImplementedError(
                        "subtract_back=True needs SExtractor's background map, which is upstream of the subtraction "
                        "path: set zuds_pipeline_b200.hotpants.background_subtractor or pass subtract_back=False")
                bkg = hotpants.background_subtractor(my_sci)
                target = FITSImage(np.asarray(my_sci.data, np.float32) - np.asarray(bkg, np.float32) + np.float32(BKG_VAL),
                                   my_sci.header.copy(), my_sci.basename.replace('.fits', '.bkgsub.fits'))
                target.map_to_local_file(os.path.join(ws.path, target.basename))

            # one upload, every step on the device, one download
            d_sci, d_ref = up(np.asarray(target.data, np.float32)), up(np.asarray(aligned.data, np.float32))
            d_scimask, d_refmask = up(np.asarray(sci.mask_image.data)), up(np.asarray(aligned.mask_image.data))
            d_scirms, scirms_name = _noise_map_on_device(ctx, my_sci, d_sci, up)
            d_refrms, refrms_name = _noise_map_on_device(ctx, getattr(aligned, 'parent_image', None) or aligned, d_ref, up)
            bpm_name = outname.replace('.fits', '.mask.bpm.fits')
            names = {'sci': target.local_path, 'ref': aligned.local_path, 'out': outname, 'sci_rms': scirms_name,
                     'ref_rms': refrms_name, 'bpm': os.path.join(ws.path, os.path.basename(bpm_name))}
            prod = subtract_on_device(ctx, d_sci, d_ref, d_scirms, d_refrms, d_scimask, d_refmask, sci.header['SEEING'],
                                      nreg_side=nreg_side, hotpants_kws=hotpants_kws, names=names)
            torch.cuda.synchronize(dev)
            result = prod.result
            result.diff, result.noise, result.mask = prod.diff.cpu().numpy(), prod.noise.cpu().numpy(), None
            submask_data = prod.mask.cpu().numpy()

            # the three products, written where hotpants would have written them, then published next to the science image
            table_hdus = hotpants.write_outputs(prod.args, my_sci.header, result, argv0='hotpants')
            submask = MaskImageBase(submask_data, sci.mask_image.header.copy(), os.path.basename(outname.replace('.fits', '.mask.fits')))
            submask.header_comments = dict(sci.mask_image.header_comments)
            submask.header['BIT%d' % HOTPANTS_BIT] = HOTPANTS_BIT
            submask.header_comments['BIT%d' % HOTPANTS_BIT] = 'MASKED BY HOTPANTS (1e-30) / DG'
            submask.map_to_local_file(outname.replace('.fits', '.mask.fits'))
            submask.save()
            final_out = os.path.join(os.path.dirname(sci.local_path), os.path.basename(outname))
            for suffix in ('.fits', '.rms.fits', '.mask.fits'):
                shutil.copy(outname.replace('.fits', suffix), final_out.replace('.fits', suffix))

        def product(klass, path, pixels):
            """The object ``klass.from_file(path)`` would give: the header as the file holds it, the pixels that were
            just written to it (37.8 MB per product are not read back and byte-swapped again)."""
            obj = klass()
            obj.map_to_local_file(path)
            obj.basename = os.path.basename(path)
            obj.header, obj._data = fitsio.read_header(path), pixels
            return obj
        sub = product(cls, final_out, result.diff)
        sub._table_hdus = table_hdus
        sub.mask_image = product(MaskImageBase, final_out.replace('.fits', '.mask.fits'), submask_data)
        sub.mask_image.parent_image = sub
        sub._rmsimg = product(FITSImage, final_out.replace('.fits', '.rms.fits'), result.noise)
        for key in ('FIELD', 'CCDID', 'QID', 'FID'):                    # zuds/subtraction.py:193-201
            v = sci.header.get(key)
            if v is not None:
                sub.header[key] = v
                sub.mask_image.header[key] = v
        sub.reference_image, sub.target_image = ref, sci
        sub.header['SEEING'] = sci.header['SEEING']
        if 'SEEING' in sci.header_comments:
            sub.header_comments['SEEING'] = sci.header_comments['SEEING']
        if issubclass(cls, SingleEpochSubtraction) and 'MAGZP' in sci.header:
            sub.header['MAGZP'] = sci.header['MAGZP']
            if APER_KEY in sci.header:
                sub.header[APER_KEY] = sci.header[APER_KEY]
        sub.last_result = result
        sub.last_command = prod.command
        sub.save()
        sub.mask_image.save()
        return sub

    def save(self):
        """Re-write the primary HDU keeping the -hki kernel table HDU the subtraction wrote."""
        extra = getattr(self, '_table_hdus', None)
        if extra is None:
            extra = ()
            if self.ismapped and os.path.exists(self.local_path):
                hdus = fitsio.read_hdus(self.local_path)
                if len(hdus) > 1:
                    structural = {"XTENSION", "BITPIX", "NAXIS", "NAXIS1", "NAXIS2", "PCOUNT", "GCOUNT", "TFIELDS"}
                    extra = []
                    for h, d in hdus[1:]:
                        h.cards = [c for c in h.cards if c[0] not in structural]      # write_image supplies them
                        extra.append(("bintable", h, d))
                    extra = tuple(extra)
        data = np.asarray(self.data, np.float32)
        hdr = self.header.copy()
        for k, c in self.header_comments.items():
            if k in hdr:
                hdr[k] = (hdr[k], c)
        fitsio.write_image(self.local_path, data, hdr, extra_hdus=extra)


class SingleEpochSubtraction(Subtraction):
    """zuds/subtraction.py:229-258 (calibrated: carries MAGZP / APCOR4)."""


class MultiEpochSubtraction(Subtraction):
    """zuds/subtraction.py:261-319.  The reference builds these by SWarp-coadding single-epoch subtractions
    (:283-319); that coadd is not built here, and a pairwise ``from_images`` under this name
    would pose as it, so the class refuses."""

    @classmethod
    def from_images(cls, *a, **kw):
        raise NotImplementedError("MultiEpochSubtraction is a SWarp coadd of single-epoch subtractions in the reference "
                                  "(zuds/subtraction.py:283-319); only SingleEpochSubtraction.from_images is built")
