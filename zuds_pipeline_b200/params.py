"""hotpants flags <-> ``zb_params`` (ctypes mirror of include/zb200.h).

``parse_hotpants_args`` restates hotpants/vargs.c:262-473 for the flags zuds passes
(zuds/hotpants.py:77-93), including the two parsing quirks that change results:
``sscanf("%d")`` truncates float-valued integer flags (``-r 6.25`` -> 6,
``-nsx 10.24`` -> 10) and ``sscanf("%f")`` rounds to float32.
"""
from __future__ import annotations

import ctypes as C
import re
import numpy as np

ZB_MAX_GAUSS = 8
ZB_MAX_KSS = 8


class ZbParams(C.Structure):
    _fields_ = [
        ("nx", C.c_int32), ("ny", C.c_int32),
        ("hw_kernel", C.c_int32), ("hw_ksstamp", C.c_int32), ("n_ksstamps", C.c_int32),
        ("ker_order", C.c_int32), ("bg_order", C.c_int32),
        ("nreg_x", C.c_int32), ("nreg_y", C.c_int32), ("nstamp_x", C.c_int32), ("nstamp_y", C.c_int32),
        ("kc_step", C.c_int32),
        ("ngauss", C.c_int32),
        ("deg", C.c_int32 * ZB_MAX_GAUSS),
        ("beta", C.c_float * ZB_MAX_GAUSS),
        ("t_uthresh", C.c_float), ("t_ukthresh", C.c_float), ("t_lthresh", C.c_float),
        ("i_uthresh", C.c_float), ("i_ukthresh", C.c_float), ("i_lthresh", C.c_float),
        ("t_gain", C.c_float), ("t_rdnoise", C.c_float), ("i_gain", C.c_float), ("i_rdnoise", C.c_float),
        ("t_pedestal", C.c_float), ("i_pedestal", C.c_float),
        ("ker_fit_thresh", C.c_float), ("scale_fit_thresh", C.c_float), ("min_frac_good_stamps", C.c_float),
        ("stat_sig", C.c_float), ("ker_sig_reject", C.c_float), ("ker_frac_mask", C.c_float),
        ("kf_spread_mask1", C.c_float), ("kf_spread_mask2", C.c_float),
        ("fill_val", C.c_float), ("fill_val_noise", C.c_float),
        ("force_convolve", C.c_int32), ("phot_normalize", C.c_int32), ("fig_merit", C.c_int32),
        ("convolve_variance", C.c_int32), ("rescale_ok", C.c_int32), ("use_full_ss", C.c_int32),
    ]


class ZbSizes(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "n_comp_ker", "n_comp", "n_bg_vectors", "n_comp_total", "mat_size",
        "fw_kernel", "fw_ksstamp", "fw_stamp", "kc_step",
        "nstamp_x", "nstamp_y", "n_stamps", "n_regions", "s_border", "max_rpix_x", "max_rpix_y")]


class ZbRegionInfo(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "rx_min", "rx_max", "ry_min", "ry_max", "bx_min", "bx_max", "by_min", "by_max",
        "conv_template", "flags", "n_stamps", "n_skipped", "fit_iters", "nx2norm")] + \
        [(n, C.c_double) for n in (
            "ksum", "sssig", "ssscat", "fsig", "fscat", "x2norm",
            "dmean", "dsige", "dsig", "dmeano", "dsigeo", "dsigo", "nscalo", "t_merit", "i_merit")]


class ZbStampInfo(C.Structure):
    _fields_ = [("x0", C.c_int32), ("y0", C.c_int32), ("nss", C.c_int32), ("sscnt", C.c_int32),
                ("xss", C.c_int32 * ZB_MAX_KSS), ("yss", C.c_int32 * ZB_MAX_KSS),
                ("mode", C.c_double), ("fwhm", C.c_double), ("chi2", C.c_double)]


class ZbWcs(C.Structure):
    """zb_wcs (include/zb200.h): TAN projection, FITS conventions."""
    _fields_ = [(n, C.c_double) for n in ("crpix1", "crpix2", "crval1", "crval2", "cd11", "cd12", "cd21", "cd22")]

    @classmethod
    def from_header(cls, hdr):
        """CRPIXi / CRVALi and the CD matrix (CDi_j, or CDELTi with an optional PCi_j) of a FITS header; raises KeyError
        without them and NotImplementedError for anything but a plain gnomonic projection."""
        for ax in ("1", "2"):
            ctype = str(hdr.get("CTYPE" + ax, ""))
            if ctype and not (ctype.endswith("TAN") or ctype.endswith("TPV")):
                raise NotImplementedError("projection %r: only gnomonic (TAN) world coordinates are built" % ctype)
        if any(k.startswith("PV1_") or k.startswith("PV2_") for k in hdr.keys()):
            raise NotImplementedError("TPV distortion polynomials (PVi_j) are not built")
        w = cls()
        w.crpix1, w.crpix2 = float(hdr["CRPIX1"]), float(hdr["CRPIX2"])
        w.crval1, w.crval2 = float(hdr["CRVAL1"]), float(hdr["CRVAL2"])
        if "CD1_1" in hdr:
            w.cd11, w.cd12 = float(hdr["CD1_1"]), float(hdr.get("CD1_2", 0.0))
            w.cd21, w.cd22 = float(hdr.get("CD2_1", 0.0)), float(hdr["CD2_2"])
        else:
            d1, d2 = float(hdr["CDELT1"]), float(hdr["CDELT2"])
            w.cd11, w.cd12 = d1 * float(hdr.get("PC1_1", 1.0)), d1 * float(hdr.get("PC1_2", 0.0))
            w.cd21, w.cd22 = d2 * float(hdr.get("PC2_1", 0.0)), d2 * float(hdr.get("PC2_2", 1.0))
        return w


class ZbHostPair(C.Structure):
    """zb_host_pair (include/zb200.h): the host pointers of one pair for zb_subtract_batch."""
    _fields_ = [("sci", C.c_void_p), ("ref", C.c_void_p), ("sci_rms", C.c_void_p), ("ref_rms", C.c_void_p),
                ("sci_mask", C.c_void_p), ("ref_mask", C.c_void_p),
                ("sci_mask_dtype", C.c_int32), ("ref_mask_dtype", C.c_int32),
                ("diff", C.c_void_p), ("noise", C.c_void_p), ("out_mask", C.c_void_p)]


def default_params(nx=0, ny=0):
    """Defaults of every flag, hotpants/vargs.c:55-128 + defaults.h:27-80."""
    p = ZbParams()
    p.nx, p.ny = nx, ny
    p.ngauss = 3
    for i, (d, s) in enumerate(((6, 0.7), (4, 1.5), (2, 3.0))):
        p.deg[i] = d
        p.beta[i] = np.float32(1.0 / (2.0 * s * s))
    p.t_uthresh = p.t_ukthresh = p.i_uthresh = p.i_ukthresh = 25000.0
    p.t_lthresh = p.i_lthresh = 0.0
    p.t_gain = p.i_gain = 1.0
    p.hw_kernel = 10
    p.ker_fit_thresh, p.min_frac_good_stamps, p.scale_fit_thresh = 20.0, 0.1, 0.5
    p.kf_spread_mask1 = p.kf_spread_mask2 = 1.0
    p.nreg_x = p.nreg_y = 1
    p.nstamp_x = p.nstamp_y = 10
    p.n_ksstamps, p.hw_ksstamp = 3, 15
    p.ker_order, p.bg_order = 2, 1
    p.stat_sig, p.ker_sig_reject, p.ker_frac_mask = 3.0, 2.0, 0.99
    p.fill_val, p.fill_val_noise = 1e-30, 0.0
    p.force_convolve, p.phot_normalize, p.fig_merit = ord("b"), ord("t"), ord("v")
    return p


_INT_RE = re.compile(r"\s*[+-]?\d+")


def _scan_d(s, old):
    """sscanf("%d"): leading integer part, value unchanged when nothing matches."""
    m = _INT_RE.match(s)
    return int(m.group(0)) if m else old


def _scan_f(s, old):
    """sscanf("%f"): nearest float32."""
    try:
        return float(np.float32(float(s)))
    except ValueError:
        return old


_FILE_FLAGS = {"inim": "image", "tmplim": "template", "outim": "outim", "tni": "t_noise", "ini": "i_noise",
               "tmi": "t_mask", "imi": "i_mask", "oni": "noise_image", "omi": "out_mask", "savexy": "savexy",
               "oki": "kernel_out"}
_FLOAT_FLAGS = {"tu": "t_uthresh", "tuk": "t_ukthresh", "tl": "t_lthresh", "tg": "t_gain", "tr": "t_rdnoise",
                "tp": "t_pedestal", "iu": "i_uthresh", "iuk": "i_ukthresh", "il": "i_lthresh", "ig": "i_gain",
                "ir": "i_rdnoise", "ip": "i_pedestal", "ft": "ker_fit_thresh", "sft": "scale_fit_thresh",
                "nft": "min_frac_good_stamps", "mins": "kf_spread_mask1", "mous": "kf_spread_mask2",
                "ssig": "stat_sig", "ks": "ker_sig_reject", "kfm": "ker_frac_mask", "fi": "fill_val",
                "fin": "fill_val_noise"}
_INT_FLAGS = {"r": "hw_kernel", "kcs": "kc_step", "nrx": "nreg_x", "nry": "nreg_y", "nsx": "nstamp_x",
              "nsy": "nstamp_y", "nss": "n_ksstamps", "rss": "hw_ksstamp", "ko": "ker_order", "bgo": "bg_order"}
_CHAR_FLAGS = {"c": "force_convolve", "n": "phot_normalize", "fom": "fig_merit"}
_BOOL_FLAGS = {"okn": "rescale_ok", "convvar": "convolve_variance"}
_UNSUPPORTED = {"ki", "gd", "rf", "rkw", "ssf", "cmp", "afssc", "sconv", "ef", "sum", "ond", "nim", "ndm", "oci",
                "cim", "sht", "obs", "obz", "nsht", "nbs", "nbz", "allm", "pca"}


class HotpantsArgs:
    """Result of parsing a hotpants command line: ``params`` + file names + switches."""

    def __init__(self):
        self.params = default_params()
        self.files = {}
        self.hki = False
        self.no_clobber = False
        self.verbose = 1
        self.argv = []


def parse_hotpants_args(argv):
    """Restates hotpants/vargs.c:262-473.  ``argv`` excludes the program name."""
    a = HotpantsArgs()
    a.argv = list(argv)
    p = a.params
    tuk_set = iuk_set = False
    it = iter(argv)
    for tok in it:
        if not tok.startswith("-"):
            raise ValueError("Unexpected string encountered on command line : %s" % tok)
        f = tok[1:].lower()
        if f in _FILE_FLAGS:
            a.files[_FILE_FLAGS[f]] = next(it)
        elif f in _FLOAT_FLAGS:
            name = _FLOAT_FLAGS[f]
            setattr(p, name, _scan_f(next(it), getattr(p, name)))
            tuk_set |= f == "tuk"
            iuk_set |= f == "iuk"
        elif f in _INT_FLAGS:
            name = _INT_FLAGS[f]
            setattr(p, name, _scan_d(next(it), getattr(p, name)))
        elif f in _CHAR_FLAGS:
            setattr(p, _CHAR_FLAGS[f], ord(next(it)[0]))
        elif f in _BOOL_FLAGS:
            setattr(p, _BOOL_FLAGS[f], 1)
        elif f == "hki":
            a.hki = True
        elif f == "nc":
            a.no_clobber = True
        elif f == "v":
            a.verbose = _scan_d(next(it), a.verbose)
        elif f == "ng":
            p.ngauss = _scan_d(next(it), p.ngauss)
            if p.ngauss > ZB_MAX_GAUSS:
                raise NotImplementedError("-ng %d > %d" % (p.ngauss, ZB_MAX_GAUSS))
            for i in range(p.ngauss):
                p.deg[i] = _scan_d(next(it), 0)
                s = np.float32(float(next(it)))
                p.beta[i] = np.float32(1.0 / (2.0 * np.float64(s) * np.float64(s)))
        elif f in _UNSUPPORTED:
            raise NotImplementedError("hotpants option -%s is outside the B200 path's scope" % f)
        else:
            raise ValueError("Unknown option : %s" % tok)
    # NB: -tu/-iu do not move tUKThresh/iUKThresh (vargs.c:68-69,74-75): they keep 25000 unless -tuk/-iuk
    for need in ("image", "template", "outim"):
        if need not in a.files:
            raise ValueError("FATAL ERROR %s : required command line option" % need)
    return a


def hotpants_command(scipath, refpath, outname, scirms, refrms, submask, seeing, nx, ny, il, tl,
                     nreg_side=3, hotpants_kws=None, exe="hotpants"):
    """The command string of zuds/hotpants.py:77-93, character for character."""
    from .synth import BIG_RMS
    hotpants_kws = hotpants_kws or {}
    r = 2.5 * seeing
    rss = 6.0 * seeing
    nsx = nx / 100.0
    nsy = ny / 100.0
    satlev = 5e3
    subrms = outname.replace(".fits", ".rms.fits")
    syscall = f'{exe} -inim {scipath} -hki -n i -c t ' \
              f'-tmplim {refpath} -outim {outname} ' \
              f'-tu {satlev} -iu {satlev}  -tl {tl} -il {il} -r {r} ' \
              f'-rss {rss} -tni {refrms} ' \
              f'-ini {scirms} ' \
              f'-imi {submask}  -v 0 -oni {subrms} ' \
              f'-fin {np.sqrt(50000.)} -nsx {nsx / nreg_side} -nsy {nsy / nreg_side} ' \
              f'-nrx {nreg_side} -nry {nreg_side} '
    for key in hotpants_kws:
        syscall += f' -{key} {hotpants_kws[key]}'
    if 'bgo' not in hotpants_kws:
        syscall += ' -bgo 0'
    if 'ko' not in hotpants_kws:
        syscall += ' -ko 4'
    assert abs(float(np.sqrt(50000.)) - BIG_RMS) == 0
    return syscall


def quick_background_estimate(data, mask):
    """zuds/utils.py:32-53: median and 1.4826*MAD of the unmasked pixels, in the pixels' own dtype as numpy
    evaluates it (float32 data: float32 median and deviations; only the 1.4826 factor is double, which is what
    numpy 1.x's scalar promotion gives the reference)."""
    bkgpix = np.asarray(data)[np.asarray(mask) == 0]
    bkg = np.median(bkgpix)
    mad = np.median(np.abs(bkgpix - bkg))
    return float(bkg), 1.4826 * float(mad)


def params_for_pair(pair, nreg_side=3, hotpants_kws=None):
    """zb_params for a synthetic pair exactly as the zuds command line would set them
    (zuds/hotpants.py:38-93 -> hotpants/vargs.c)."""
    ny, nx = pair.sci.shape
    scibkg, scistd = quick_background_estimate(pair.sci, pair.sci_mask)
    refbkg, refstd = quick_background_estimate(pair.ref, pair.ref_mask)
    il, tl = scibkg - 10 * scistd, refbkg - 10 * refstd
    cmd = hotpants_command("sci.fits", "ref.fits", "sub.fits", "sci.rms.fits", "ref.rms.fits", "bpm.fits",
                           pair.seeing, nx, ny, il, tl, nreg_side=nreg_side, hotpants_kws=hotpants_kws)
    a = parse_hotpants_args(cmd.split()[1:])
    a.params.nx, a.params.ny = nx, ny
    return a.params
