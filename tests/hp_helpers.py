"""Test helpers: drive the CPU oracle (oracle/libhp_oracle.so) and, where it exists,
the reference binary oracle/_ref/hotpants on synthetic pairs.  TEST CODE ONLY."""
from __future__ import annotations

import ctypes as C
import os
import re
import subprocess
import tempfile

import numpy as np

from zuds_pipeline_b200 import fitsio, params as P, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "libhp_oracle.so")
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "hotpants")
GOLDEN = os.path.join(ROOT, "tests", "golden")

_fp = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")


def load_oracle():
    if not os.path.exists(ORACLE_SO):
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "port"])
    lib = C.CDLL(ORACLE_SO)
    lib.hp_oracle_subtract.restype = C.c_int
    lib.hp_oracle_subtract.argtypes = [C.POINTER(P.ZbParams)] + [C.c_void_p] * 9 + \
        [C.POINTER(P.ZbRegionInfo), C.c_void_p, C.POINTER(P.ZbStampInfo)]
    lib.hp_oracle_sizes.argtypes = [C.POINTER(P.ZbParams), C.POINTER(P.ZbSizes)]
    lib.hp_oracle_ran1.restype = C.c_double
    lib.hp_oracle_ran1.argtypes = [C.POINTER(C.c_int)]
    return lib


class Result:
    pass


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def run_oracle(lib, p, sci, ref, sci_rms, ref_rms, sci_mask=None, ref_mask=None):
    """Call hp_oracle_subtract; returns Result(diff, noise, mask, regions, sol, stamps, sizes)."""
    sz = P.ZbSizes()
    lib.hp_oracle_sizes(C.byref(p), C.byref(sz))
    ny, nx = sci.shape
    r = Result()
    r.sizes = sz
    r.diff = np.zeros((ny, nx), np.float32)
    r.noise = np.zeros((ny, nx), np.float32)
    r.mask = np.zeros((ny, nx), np.int32)
    r.regions = (P.ZbRegionInfo * sz.n_regions)()
    r.sol = np.zeros((sz.n_regions, sz.n_comp_total + 1), np.float64)
    r.stamps = (P.ZbStampInfo * (sz.n_regions * sz.n_stamps))()
    m1 = None if sci_mask is None else np.ascontiguousarray(sci_mask, np.int32)
    m2 = None if ref_mask is None else np.ascontiguousarray(ref_mask, np.int32)
    arrs = [np.ascontiguousarray(a, np.float32) if a is not None else None for a in (sci, ref, sci_rms, ref_rms)]
    rc = lib.hp_oracle_subtract(C.byref(p), _ptr(arrs[0]), _ptr(arrs[1]), _ptr(arrs[2]), _ptr(arrs[3]),
                                _ptr(m1), _ptr(m2), _ptr(r.diff), _ptr(r.noise), _ptr(r.mask),
                                r.regions, _ptr(r.sol), r.stamps)
    r.rc = rc
    return r


def stamp_table(res):
    """[(region, index, x0, y0, nss, sscnt, (xss...), (yss...))] for the kept stamps."""
    out = []
    ns = res.sizes.n_stamps
    for ir in range(res.sizes.n_regions):
        for k in range(res.regions[ir].n_stamps):
            s = res.stamps[ir * ns + k]
            out.append((ir, k, s.x0, s.y0, s.nss, s.sscnt, tuple(s.xss[:s.nss]), tuple(s.yss[:s.nss])))
    return out


def zuds_flags(pair, nreg_side=3, hotpants_kws=None):
    """il/tl thresholds exactly as zuds/hotpants.py:65-73 computes them."""
    scibkg, scistd = P.quick_background_estimate(pair.sci, pair.sci_mask)
    refbkg, refstd = P.quick_background_estimate(pair.ref, pair.ref_mask)
    return scibkg - 10 * scistd, refbkg - 10 * refstd


def write_pair(pair, d):
    hdr = fitsio.Header()
    hdr["SEEING"] = pair.seeing
    hdr["FIELD"] = 600
    hdr["CCDID"] = 1
    hdr["QID"] = 1
    hdr["FID"] = 2
    names = {}
    for n, a in (("sci", pair.sci), ("ref", pair.ref), ("sci.rms", pair.sci_rms), ("ref.rms", pair.ref_rms),
                 ("bpm", pair.bpm)):
        names[n] = os.path.join(d, n + ".fits")
        fitsio.write_image(names[n], a, hdr)
    return names


def zuds_command(pair, d, nreg_side=3, hotpants_kws=None, exe="hotpants", extra=""):
    names = write_pair(pair, d)
    il, tl = zuds_flags(pair)
    ny, nx = pair.sci.shape
    cmd = P.hotpants_command(names["sci"], names["ref"], os.path.join(d, "sub.fits"), names["sci.rms"],
                             names["ref.rms"], names["bpm"], pair.seeing, nx, ny, il, tl,
                             nreg_side=nreg_side, hotpants_kws=hotpants_kws, exe=exe)
    return cmd + extra


_BOX = re.compile(r"image;box\(\s*(-?\d+),\s*(-?\d+),")


def run_reference(pair, nreg_side=3, hotpants_kws=None, workdir=None, case=None):
    """Run oracle/_ref/hotpants (the unmodified reference) on ``pair`` through FITS files.
    Adds -omi (internal int mask) and -savexy (stamp centres) to the zuds command.
    With ``case`` (tests/cases.py) the command comes from cases.command()."""
    assert os.path.exists(REF_BIN), "build it with: make -C oracle ref"
    tmp = workdir or tempfile.mkdtemp(prefix="hpref_")
    extra = " -omi %s -savexy %s" % (os.path.join(tmp, "sub.mask.fits"), os.path.join(tmp, "xy"))
    if case is not None:
        import cases as CS
        names = write_pair(pair, tmp)
        names["out"] = os.path.join(tmp, "sub.fits")
        cmd = CS.command(case, pair, names, exe=REF_BIN) + extra
    else:
        cmd = zuds_command(pair, tmp, nreg_side, hotpants_kws, exe=REF_BIN, extra=extra)
    pr = subprocess.run(cmd.split(), capture_output=True, text=True)
    if pr.returncode != 0:
        raise RuntimeError("reference hotpants failed: %s" % pr.stderr[-2000:])
    r = Result()
    r.cmd = cmd
    r.stderr = pr.stderr
    hdus = fitsio.read_hdus(os.path.join(tmp, "sub.fits"))
    r.diff, r.header = hdus[0][1], hdus[0][0]
    r.sol = hdus[1][1].T.copy() if len(hdus) > 1 else None          # [region][nCompTotal+1]
    r.table_header = hdus[1][0] if len(hdus) > 1 else None
    rms = os.path.join(tmp, "sub.rms.fits")
    r.noise, r.noise_header = (fitsio.read_image(rms)[0], fitsio.read_image(rms)[1]) if os.path.exists(rms) else (None, None)
    r.mask = fitsio.read_image(os.path.join(tmp, "sub.mask.fits"))[0].astype(np.int32)

    def boxes(path):
        if not os.path.exists(path):
            return []
        return [tuple(map(int, m.groups())) for m in _BOX.finditer(open(path).read())]
    r.xy_used = boxes(os.path.join(tmp, "xy"))
    r.xy_all = boxes(os.path.join(tmp, "xy.all"))
    r.workdir = tmp
    return r


def params_from_command(cmd, nx, ny):
    a = P.parse_hotpants_args(cmd.split()[1:])
    a.params.nx, a.params.ny = nx, ny
    return a


def oracle_xy(res):
    """Stamp centres in the -savexy convention (1-indexed image coords): (all, used)."""
    allc, used = [], []
    ns = res.sizes.n_stamps
    for ir in range(res.sizes.n_regions):
        ri = res.regions[ir]
        for k in range(ri.n_stamps):
            s = res.stamps[ir * ns + k]
            for q in range(s.nss):
                # savexy is called with rXBMin for -c t but with pixMin (= rXBMin+1) for -c i
                # (hotpants/main.c:1176 vs :1292): reproduce that off-by-one here
                off = 1 if ri.conv_template else 2
                c = (s.xss[q] + ri.bx_min + off, s.yss[q] + ri.by_min + off)
                allc.append(c)
                if q == s.sscnt:
                    used.append(c)
    return allc, used
