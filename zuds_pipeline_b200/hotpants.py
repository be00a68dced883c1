"""Drop-in for ``zuds.hotpants`` + the ``hotpants`` executable it drives.

``prepare_hotpants`` keeps the reference signature and returns the same command string
(zuds/hotpants.py:15-95).  ``run_hotpants(command)`` replaces
``subprocess.check_call(command.split())`` (zuds/subtraction.py:162): it parses the flags the way
hotpants/vargs.c does, reads the FITS inputs, runs the subtraction on the GPU through the C ABI
(libzb200.so, include/zb200.h) and writes the difference image (+ ``-hki`` kernel table), the
``-oni`` noise image and optionally the ``-omi`` mask with the keywords hotpants/main.c writes.
``main(argv)`` is the same thing as a command-line program (bin/hotpants).

There is no CPU compute path: without libzb200.so and a CUDA device ``run_hotpants`` raises.
"""
from __future__ import annotations

import math
import os
import socket
import subprocess
import sys
import time

import numpy as np

from . import fitsio, params as P
from .synth import BIG_RMS, BKG_VAL

__all__ = ["prepare_hotpants", "run_hotpants", "build_headers", "write_outputs", "main", "VERSION"]

VERSION = "5.1.11"          # hotpants/main.c:109; zuds parses it from line 2 of the usage text (zuds/constants.py:91-95)

# zuds/hotpants.py:27: the science image is background-subtracted with SExtractor before the call.
# SExtractor is an external binary upstream of this path; a caller that has one plugs it in here:
# background_subtractor(sci) -> image object whose pixels are background-subtracted (BKG_VAL is added here).
background_subtractor = None


def prepare_hotpants(sci, ref, outname, submask, directory, tmpdir='/tmp',
                     nreg_side=3, subtract_new_back=True, hotpants_kws=None):
    """Same call and same result as zuds.hotpants.prepare_hotpants: the hotpants command line.

    ``sci`` / ``ref`` / ``submask`` are image objects of ``zuds_pipeline_b200.image`` (or anything with
    the attributes used below: ``local_path``, ``header``, ``data``, ``rms_image``, ``mask_image``).
    """
    from .image import quick_background_estimate
    os.makedirs(str(directory), exist_ok=True)            # zuds.utils.initialize_directory
    if hotpants_kws is None:
        hotpants_kws = {}

    if subtract_new_back:
        if background_subtractor is None:
            raise NotImplementedError(
                "subtract_new_back=True needs SExtractor's background map, which is upstream of the "
                "subtraction path: set zuds_pipeline_b200.hotpants.background_subtractor or pass "
                "subtract_back=False with a background-subtracted science image")
        scimbkg = background_subtractor(sci)
        scimbkg.data = scimbkg.data + np.float32(BKG_VAL)
        scimbkg.save()
    else:
        scimbkg = sci

    scipath = scimbkg.local_path
    if 'SEEING' not in sci.header:
        raise KeyError("SEEING (FWHM in pixels) must be in the science header; zuds.seeing.estimate_seeing "
                       "queries external catalogues and is outside this path")
    seepix = sci.header['SEEING']
    r = 2.5 * seepix
    rss = 6. * seepix
    nsx = sci.header['NAXIS1'] / 100.
    nsy = sci.header['NAXIS2'] / 100.

    scirms = sci.rms_image
    parent = getattr(ref, 'parent_image', None) or ref
    refrms = parent.rms_image.aligned_to(scirms, tmpdir=tmpdir)
    for rms in (scirms, refrms):
        if not rms.ismapped:
            rms.map_to_local_file(os.path.join(str(directory), rms.basename))
            rms.save()

    bctx = _prestep_context()
    scibkg, scibkgstd = quick_background_estimate(scimbkg, mask_image=sci.mask_image, ctx=bctx)
    refbkg, refbkgstd = quick_background_estimate(ref, ctx=bctx)
    subrms = outname.replace('.fits', '.rms.fits')
    il = scibkg - 10 * scibkgstd
    tl = refbkg - 10 * refbkgstd
    satlev = 5e3

    syscall = f'hotpants -inim {scipath} -hki -n i -c t ' \
              f'-tmplim {ref.local_path} -outim {outname} ' \
              f'-tu {satlev} -iu {satlev}  -tl {tl} -il {il} -r {r} ' \
              f'-rss {rss} -tni {refrms.local_path} ' \
              f'-ini {scirms.local_path} ' \
              f'-imi {submask.local_path}  -v 0 -oni {subrms} ' \
              f'-fin {BIG_RMS} -nsx {nsx / nreg_side} -nsy {nsy / nreg_side} ' \
              f'-nrx {nreg_side} -nry {nreg_side} '
    for key in hotpants_kws:
        syscall += f' -{key} {hotpants_kws[key]}'
    if 'bgo' not in hotpants_kws:
        syscall += ' -bgo 0'
    if 'ko' not in hotpants_kws:
        syscall += ' -ko 4'
    return syscall


# ----------------------------------------------------------------------------- header keywords

def _flt_g(v, digits):
    """cfitsio fits_write_key_flt / fits_update_key_flt with NEGATIVE decimals: %.*G, '.' appended if bare."""
    s = "%.*G" % (digits, float(np.float32(v)))
    if "." not in s and "E" not in s and "N" not in s:
        s += "."
    return fitsio.Raw(s)


def _flt_e(v, decimals):
    """cfitsio fits_write_key_flt with POSITIVE decimals: %.*E."""
    return fitsio.Raw("%.*E" % (decimals, float(np.float32(v))))


def copy_header(src):
    """hotpants/functions.c:1407-1430 hp_fits_copy_header: everything after the mandatory
    SIMPLE/BITPIX/NAXIS* cards except EXTEND and the two standard FITS COMMENT cards."""
    out = fitsio.Header()
    for k, v, c in src.cards:
        if k in ("SIMPLE", "BITPIX", "NAXIS", "NAXIS1", "NAXIS2", "EXTEND", "END"):
            continue
        if k == "COMMENT" and (str(v).lstrip().startswith("FITS (Flexible Image Transport System) format is")
                               or str(v).lstrip().startswith("and Astrophysics', volume 376, page 3")):
            continue
        out.append(k, v, c)
    return out


REGION_COMMENTS = (("KSUM", "Kernel Sum"), ("SSSIG", "Average Figure of Merit across Stamps"),
                   ("SSSCAT", "Stdev in Figure of Merit"), ("FSIG", "Final SSSIG"), ("FSCAT", "Final SSSCAT"),
                   ("X2NRM", "1/N * SUM (diff/noise)^2"), ("NX2NRM", "Number of pixels in X2NRM"),
                   ("DMEAN", "Mean of diff image; good pixels"), ("DSIGE", "Stdev of diff image; good pixels"),
                   ("DSIG", "Mean of noise image; good pixels"), ("DMEANO", "Mean of diff image; OK pixels"),
                   ("DSIGEO", "Stdev of diff image; OK pixels"), ("DSIGO", "Mean of noise image; OK pixels"))


def build_headers(args, sci_header, result, argv0="hotpants", now=None):
    """(diff header, noise header, kernel-table header) as hotpants/main.c:243-276,346-351,561-579,
    1719-1819,1874-1953 write them.  ``result`` carries ``sizes``, ``regions`` (zb_region_info[]) ."""
    p, sz = args.params, result.sizes
    d = copy_header(sci_header)
    d.append("COMMENT", "")
    d.append("SOFTNAME", "HOTPanTS", "The software that differenced this image")
    d.append("SOFTVERS", VERSION, "Version")
    d.append("SOFTAUTH", "Andrew Becker <becker@astro.washington.edu>", "Maintainer")
    d.append("SOFTINST", "University of Washington", "Institute")
    d.append("COMMENT", "")
    user = None if os.environ.get("USERNAME") else os.environ.get("LOGNAME")      # main.c:251-255 (sic)
    d.append("AUTHOR", user or "unknown", "Who ran the software")
    d.append("ORIGIN", socket.gethostname()[:79], "Where it was done")
    tm = time.gmtime(now)
    d.append("DATE", "%04d-%02d-%02dT%02d:%02d:%02d" % tm[:6], "When it was started (GMT)")
    d.append("COMMENT", "")
    d["BZERO"] = _flt_g(0.0, 5)
    d["BSCALE"] = _flt_g(1.0, 5)
    n = copy_header(sci_header)
    n["OBJECT"] = "HOTPanTS Noise Image"
    n["GAIN"] = (_flt_g(1.0, 1), "No gain in noise image")
    n["RDNOISE"] = (_flt_g(0.0, 1), "No rdnoise in noise image")
    n.append("MASKVAL", _flt_g(p.fill_val_noise, 6), "Value of Masked Pixels")
    for i in range(sz.n_regions):
        r = result.regions[i]
        if r.flags & 0x1:                      # region skipped before any keyword is written (main.c:1002-1009)
            continue
        d.append("REGION%02d" % i, "[%d:%d,%d:%d]" % (r.rx_min + 1, r.rx_max + 1, r.ry_min + 1, r.ry_max + 1))
        d.append("CONVOL%02d" % i, "TEMPLATE" if r.conv_template else "IMAGE")
        vals = {"KSUM": r.ksum, "SSSIG": r.sssig, "SSSCAT": r.ssscat, "FSIG": r.fsig, "FSCAT": r.fscat, "X2NRM": r.x2norm,
                "DMEAN": r.dmean, "DSIGE": r.dsige, "DSIG": r.dsig, "DMEANO": r.dmeano, "DSIGEO": r.dsigeo, "DSIGO": r.dsigo}
        for key, comment in REGION_COMMENTS:
            if key in ("FSIG", "FSCAT") and chr(p.fig_merit) != "v":
                continue
            d.append("%s%02d" % (key, i), ("%d" % r.nx2norm) if key == "NX2NRM" else ("%.4f" % vals[key]), comment)
        if p.rescale_ok:
            d.append("NSCALO%02d" % i, "%.4f" % r.nscalo, "Rescaling factor of OK noise pixels")
        n.append("NSCALO%02d" % i, "%.4f" % r.nscalo, "Rescaling factor of OK noise pixels")
    d.append("DIFFCMD", " ".join([argv0] + list(args.argv)))
    d.append("TARGET", args.files["image"], "HOTPanTS : Input Image")
    d.append("TEMPLATE", args.files["template"], "HOTPanTS : Reference Image")
    d.append("DIFFIM", args.files["outim"], "HOTPanTS : Difference Image")
    d.append("PHOTNORM", chr(p.phot_normalize), "Direction of photometric normalization")
    d.append("NREGION", int(p.nreg_x * p.nreg_y), "Number of independent regions")
    d.append("MASKVAL", _flt_g(p.fill_val, 6), "Value of Masked Pixels")
    t = None
    if args.hki:
        d.append("KERINFO", True)
        t = fitsio.Header()
        for k in range(sz.n_regions):
            t.append("TTYPE%d" % (k + 1), "Region%d" % k, "label for field %3d" % (k + 1))
            t.append("TFORM%d" % (k + 1), "D10.8", "data format of field: 8-byte DOUBLE")
        t.append("EXTNAME", "Convolution Kernel Information", "name of this binary table extension")
        t.append("OBJECT", "Convolution Kernel Information")
        t.append("NGAUSS", int(p.ngauss), "Number of Gaussian Basis Functions")
        for k in range(p.ngauss):
            t.append("DGAUSS%d" % (k + 1), int(p.deg[k]), "Polynomial Degree")
            t.append("SGAUSS%d" % (k + 1), _flt_e(1. / math.sqrt(2. * p.beta[k]), 8), "Gaussian Sigma")
        t.append("FWKERN", int(sz.fw_kernel), "Kernel Width in Pixels")
        t.append("CKORDER", int(p.ker_order), "Spatial Convolution Order")
        t.append("BGORDER", int(p.bg_order), "Background Order")
        t.append("KERINFO", False)            # main.c:1951-1953: lands in the table HDU (file pointer still there)
    else:
        d.append("KERINFO", False)
        d.append("KERINFO", False)            # written twice without -hki/-oki (main.c:1909-1912,1951-1953)
    return d, n, t


# ----------------------------------------------------------------------------- execution

class HotpantsError(subprocess.CalledProcessError):
    """What ``subprocess.check_call`` raises when the reference executable exits non-zero
    (scripts/dosub.py:204-209 catches it per image and moves on)."""


_ctx_cache = {}

# Where the wrapper's own pre-step (quick_background_estimate: two medians and two MADs of 9.5 Mpx per
# pair, ~0.4 s of numpy in the reference) runs: "gpu" (zb_background_estimate; the default, and it fails
# loudly without a CUDA device like the rest of the path) or "host" (the reference's numpy code, only when
# asked for: ZB_PRESTEP=host or hotpants.prestep = "host"; announced on stderr).  Both give the same
# numbers bit for bit, so the hotpants command line does not depend on the choice.
prestep = os.environ.get("ZB_PRESTEP", "gpu")
DEVICE = int(os.environ.get("ZB_DEVICE", os.environ.get("LOCAL_RANK", "0")))
_host_prestep_announced = False


def _prestep_context():
    global _host_prestep_announced
    if prestep == "host":
        if not _host_prestep_announced:
            sys.stderr.write("zuds_pipeline_b200.hotpants: quick_background_estimate on the host (numpy) as requested\n")
            _host_prestep_announced = True
        return None
    if prestep != "gpu":
        raise ValueError("hotpants.prestep must be 'gpu' or 'host', not %r" % (prestep,))
    return _context(DEVICE)


def _context(device):
    """One zb_ctx per (device, host thread): a ctx serves one thread at a time (include/zb200.h)."""
    import threading
    from . import _lib
    key = (device, threading.get_ident())
    if key not in _ctx_cache:
        _ctx_cache[key] = _lib.Context(device)
    return _ctx_cache[key]


def _read_float(path):
    data, hdr = fitsio.read_image(path)
    return np.ascontiguousarray(data, dtype=np.float32), hdr


def write_outputs(args, sci_header, result, argv0="hotpants"):
    """Write -outim (+ -hki table), -oni and -omi exactly where the command line says.  Returns the extension HDUs
    of -outim as ``fitsio.write_image`` takes them (the kernel table, or nothing)."""
    d, n, t = build_headers(args, sci_header, result, argv0=argv0)
    sz = result.sizes
    extra = ()
    if t is not None:
        extra = (("bintable", t, np.ascontiguousarray(result.sol.T)),)       # nCompTotal+1 rows x nR columns
    fitsio.write_image(args.files["outim"], np.asarray(result.diff, np.float32), d, extra_hdus=extra)
    if "noise_image" in args.files:
        fitsio.write_image(args.files["noise_image"], np.asarray(result.noise, np.float32), n)
    if "out_mask" in args.files and result.mask is not None:
        m = np.asarray(result.mask).copy()
        for i in range(sz.n_regions):
            r = result.regions[i]
            if r.flags & 0x1:        # never written by the reference: BZERO'd zeros read back as 32768
                m[r.ry_min:r.ry_max + 1, r.rx_min:r.rx_max + 1] = 0x8000
        mh = copy_header(sci_header)
        mh["OBJECT"] = "Hotpants output Mask Image"
        fitsio.write_image(args.files["out_mask"], (m & 0xffff).astype(np.int32), mh)
    if "savexy" in args.files and getattr(result, "stamps", None) is not None:
        write_savexy(args.files["savexy"], result)
    return extra


def write_savexy(name, result):
    """-savexy: the ds9 region files of hotpants/functions.c:75-127 (``name``: the sub-stamp in use per stamp,
    ``name.all``: every sub-stamp found, ``name.skipped``: the ones given up), one block per region that was fitted.
    The file is started by region 0 and appended to by the others (functions.c:84-92); a region convolving the image
    is written one pixel further up and right than one convolving the template (main.c:1176 vs :1297)."""
    sz = result.sizes
    fw = sz.fw_ksstamp
    for i in range(sz.n_regions):
        r = result.regions[i]
        if r.flags & 0x1:
            continue
        off = 1 if r.conv_template else 2
        mode = "w" if i == 0 else "a"
        with open(name, mode) as used, open(name + ".all", mode) as alls, open(name + ".skipped", mode) as skipped:
            for f in (alls, used, skipped):
                f.write("#%4s %4s\n" % ("X", "Y"))
            for k in range(r.n_stamps):
                s = result.stamps[i * sz.n_stamps + k]
                for q in range(s.nss):
                    x, y = s.xss[q] + r.bx_min + off, s.yss[q] + r.by_min + off
                    alls.write("image;box(%4d,%4d,%d,%d) # color=yellow\n" % (x, y, fw, fw))
                    if q == s.sscnt:
                        used.write("image;box(%4d,%4d,%d,%d) # color=green\n" % (x, y, fw, fw))
                    if q < s.sscnt:
                        skipped.write("image;box(%4d,%4d,%d,%d) # color=red\n" % (x, y, fw, fw))


def run_hotpants(command, device=None, ctx=None):
    """Drop-in for ``subprocess.check_call(command.split())`` with ``command`` from prepare_hotpants.

    Returns the ``_lib.Result`` (arrays + per-region QA) after writing the FITS products."""
    from . import _lib
    argv = command.split() if isinstance(command, str) else list(command)
    try:
        args = P.parse_hotpants_args(argv[1:])
    except (ValueError, NotImplementedError) as e:
        sys.stderr.write("%s\n" % e)
        raise HotpantsError(1, argv) from e
    if args.no_clobber and os.path.exists(args.files["outim"]):
        raise HotpantsError(105, argv)                    # cfitsio FILE_NOT_CREATED
    try:
        sci, sci_hdr = _read_float(args.files["image"])
        ref, _ = _read_float(args.files["template"])
        if sci.shape != ref.shape:
            sys.stderr.write("Image Sizes Different, Aborting...\n")      # main.c:150-160 region
            raise HotpantsError(1, argv)
        sci_rms = _read_float(args.files["i_noise"])[0] if "i_noise" in args.files else None
        ref_rms = _read_float(args.files["t_noise"])[0] if "t_noise" in args.files else None

        def rd_mask(key):
            if key not in args.files:
                return None
            m = fitsio.read_image(args.files[key])[0]
            if m.dtype.kind == "f":
                m = np.rint(m).astype(np.int32)       # cfitsio's fits_read_subset_int rounds to nearest (main.c:781,792)
            return np.ascontiguousarray(m.astype(np.int32) if m.dtype.name not in _lib.MASK_DTYPES else m)
        sci_mask, ref_mask = rd_mask("i_mask"), rd_mask("t_mask")
    except (IOError, OSError, KeyError) as e:
        sys.stderr.write("%s\n" % e)
        raise HotpantsError(104, argv) from e             # cfitsio FILE_NOT_OPENED
    p = args.params
    p.ny, p.nx = sci.shape
    c = ctx or _context(DEVICE if device is None else device)
    try:
        res = c.subtract_host(p, sci, ref, sci_rms, ref_rms, sci_mask, ref_mask, want_mask=True)
    except _lib.ZbError as e:
        sys.stderr.write("%s\n" % e)
        raise HotpantsError(1, argv) from e
    write_outputs(args, sci_hdr, res, argv0=argv[0])
    if args.verbose >= 0:
        sys.stderr.write("SUCCESS\n")                     # main.c:1997
    return res


USAGE = """Usage : hotpants [options]
Version %s
Required options:
   [-inim fitsfile]  : comparison image to be differenced
   [-tmplim fitsfile]: template image
   [-outim fitsfile] : output difference image

Options built into the B200 path (see hotpants/vargs.c:130-260 for their meaning):
   -tu -tuk -tl -tg -tr -tp -tni -tmi   -iu -iuk -il -ig -ir -ip -ini -imi
   -ki(n/a) -r -kcs -ft -sft -nft -mins -mous -omi -gd(n/a) -nrx -nry -nsx -nsy -nss -rss -savexy
   -c [t|i] -n [t|i|u] -fom [v] -sconv(n/a) -ko -bgo -ssig -ks -kfm -okn -fi -fin -convvar
   -oni -hki -nc -ng -v
""" % VERSION


def main(argv=None):
    """The ``hotpants`` program: exit status 0 + ``SUCCESS`` on stderr, non-zero otherwise."""
    argv = list(sys.argv if argv is None else argv)
    if len(argv) < 2:
        sys.stderr.write(USAGE)       # zuds/constants.py:91-95 reads 'Version x.y.z' from line 2 of this text
        return 1
    try:
        run_hotpants(argv)
    except HotpantsError as e:
        return int(e.returncode) or 1
    return 0


if __name__ == "__main__":
    sys.exit(main())
