"""Drop-in for ``zuds.subtraction``: ``sub_name`` and ``Subtraction.from_images``.

Same call signature, keyword arguments (``subtract_back``, ``nreg_side``, ``hotpants_kws``), file
products (``sub.<sci>_<ref>.fits``, ``.rms.fits``, ``.mask.fits``), header keywords and mask-bit
semantics as zuds/subtraction.py:25-226 — with the fork/exec of the ``hotpants`` binary
(zuds/subtraction.py:162) replaced by the in-process GPU path (``hotpants.run_hotpants``), and without
the Postgres records and the archive copy (out of scope).
"""
from __future__ import annotations

import os
import shutil
import uuid
from pathlib import Path

import numpy as np

from . import fitsio
from .image import CalibratableImageBase, FITSImage, MaskImageBase

__all__ = ['sub_name', 'Subtraction', 'SingleEpochSubtraction', 'MultiEpochSubtraction']

APER_KEY = 'APCOR4'           # zuds/constants.py


def sub_name(frame, template):
    """zuds/subtraction.py:25-37."""
    frame = f'{frame}'
    template = f'{template}'
    refp = os.path.basename(template)[:-5]
    newp = os.path.basename(frame)[:-5]
    outdir = os.path.dirname(frame)
    subp = '_'.join([newp, refp])
    return os.path.join(outdir, 'sub.%s.fits' % subp)


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
        from .hotpants import prepare_hotpants, run_hotpants

        subtract_new_back = kwargs.get('subtract_back', True)
        nreg_side = kwargs.get('nreg_side', 3)
        hotpants_kws = kwargs.get('hotpants_kws', {})

        directory = Path(tmpdir) / uuid.uuid4().hex
        directory.mkdir(exist_ok=True, parents=True)
        try:
            # copy the inputs into the directory and reload them: transactional isolation (:68-99)
            shutil.copy(sci.local_path, directory)
            shutil.copy(sci.mask_image.local_path, directory)
            if hasattr(sci, '_rmsimg'):
                shutil.copy(sci.rms_image.local_path, directory)
            elif hasattr(sci, '_weightimg'):
                shutil.copy(sci.weight_image.local_path, directory)
            else:
                raise ValueError('Science image must have a weight map or '
                                 'rms map defined prior to subtraction.')
            shutil.copy(ref.local_path, directory)
            shutil.copy(ref.mask_image.local_path, directory)
            shutil.copy(ref.weight_image.local_path, directory)

            sciname = os.path.join(directory, sci.basename)
            scimaskn = os.path.join(directory, sci.mask_image.basename)
            refname = os.path.join(directory, ref.basename)
            refmaskn = os.path.join(directory, ref.mask_image.basename)

            transact_sci = CalibratableImageBase.from_file(sciname)
            transact_scimask = MaskImageBase.from_file(scimaskn)
            transact_ref = CalibratableImageBase.from_file(refname)
            transact_refmask = MaskImageBase.from_file(refmaskn)
            transact_sci.mask_image = transact_scimask
            transact_ref.mask_image = transact_refmask

            outname = sub_name(transact_sci.local_path, transact_ref.local_path)
            outmask = outname.replace('.fits', '.mask.fits')

            # the reference resamples with SWarp here (:109); aligned inputs pass through
            remapped_ref = transact_ref.aligned_to(transact_sci, tmpdir=tmpdir)
            remapped_refmask = remapped_ref.mask_image
            remapped_ref.parent_image = transact_ref

            submask = MaskImageBase()
            submask.basename = os.path.basename(outmask)
            submask.map_to_local_file(outmask)
            badpix = np.asarray(remapped_refmask.data).astype(np.int32) | np.asarray(sci.mask_image.data).astype(np.int32)
            submask.header = sci.mask_image.header.copy()
            submask.header_comments = dict(sci.mask_image.header_comments)
            submask.data = badpix
            submask.save()
            submask.boolean.map_to_local_file(directory / submask.boolean.basename)
            submask.boolean.save()

            command = prepare_hotpants(transact_sci, remapped_ref, outname, submask.boolean, directory,
                                       tmpdir=tmpdir, nreg_side=nreg_side,
                                       subtract_new_back=subtract_new_back, hotpants_kws=hotpants_kws)

            final_dir = os.path.dirname(sci.local_path)
            final_out = os.path.join(final_dir, os.path.basename(outname))
            product_map = {
                outname: final_out,
                outname.replace('.fits', '.rms.fits'): final_out.replace('.fits', '.rms.fits'),
                outname.replace('.fits', '.mask.fits'): final_out.replace('.fits', '.mask.fits')
            }

            # run HOTPANTS (in process, on the GPU) -- zuds/subtraction.py:162
            result = run_hotpants(command, device=cls.device)

            # bit 17 where hotpants masked the difference (:167-177)
            sd = result.diff
            hotbad = np.zeros_like(submask.data, dtype=int)
            hotbad[sd == np.float32(1e-30)] = 2 ** 17
            submask.data = np.asarray(submask.data) | hotbad
            submask.header['BIT17'] = 17
            submask.header_comments['BIT17'] = 'MASKED BY HOTPANTS (1e-30) / DG'
            submask.save()

            for f in product_map:
                shutil.copy(f, product_map[f])

            sub = cls.from_file(final_out, load_others=False)
            finalsubmask = MaskImageBase.from_file(final_out.replace('.fits', '.mask.fits'))
            sub._rmsimg = FITSImage.from_file(final_out.replace('.fits', '.rms.fits'))

            for key in ('FIELD', 'CCDID', 'QID', 'FID'):
                v = sci.header.get(key)
                if v is not None:
                    sub.header[key] = v
                    finalsubmask.header[key] = v
            sub.mask_image = finalsubmask
            finalsubmask.parent_image = sub
            sub.reference_image = ref
            sub.target_image = sci
            sub.header['SEEING'] = sci.header['SEEING']
            if 'SEEING' in sci.header_comments:
                sub.header_comments['SEEING'] = sci.header_comments['SEEING']
            if issubclass(cls, SingleEpochSubtraction) and 'MAGZP' in sci.header:
                sub.header['MAGZP'] = sci.header['MAGZP']
                if APER_KEY in sci.header:
                    sub.header[APER_KEY] = sci.header[APER_KEY]
            sub.last_result = result
            sub.save()
            sub.mask_image.save()
            if data_product:
                raise NotImplementedError("archive() (HTTP/tape copy, zuds/archive.py) is outside this path")
        finally:
            shutil.rmtree(directory, ignore_errors=True)
        return sub

    def save(self):
        """Re-write the primary HDU keeping the -hki kernel table HDU the subtraction wrote."""
        extra = ()
        if self.ismapped and os.path.exists(self.local_path):
            hdus = fitsio.read_hdus(self.local_path)
            if len(hdus) > 1:
                extra = tuple(("bintable", h, d) for h, d in hdus[1:])
        data = np.asarray(self.data, np.float32)
        hdr = self.header.copy()
        for k, c in self.header_comments.items():
            if k in hdr:
                hdr[k] = (hdr[k], c)
        fitsio.write_image(self.local_path, data, hdr, extra_hdus=extra)


class SingleEpochSubtraction(Subtraction):
    """zuds/subtraction.py:229-258 (calibrated: carries MAGZP / APCOR4)."""


class MultiEpochSubtraction(Subtraction):
    """zuds/subtraction.py:261-319.  The reference builds these by SWarp-coadding single-epoch
    subtractions; only the pairwise ``from_images`` inherited above is on this path."""
