"""Parity cases shared by the CPU (oracle vs golden) and GPU (CUDA vs oracle) tests.

Each case is a seeded synthetic pair + the hotpants command line zuds would build for it
(zuds/hotpants.py:77-93) + optional extra flags.  Small enough for the oracle to finish in
seconds; the full-size BASELINE.json configs are exercised by the GPU tests through the same table.
"""
from __future__ import annotations

import dataclasses

from zuds_pipeline_b200 import params as P, synth


@dataclasses.dataclass
class Case:
    name: str
    shape: tuple
    seeing: float
    nreg_side: int = 1
    hotpants_kws: dict = dataclasses.field(default_factory=dict)
    synth_kws: dict = dataclasses.field(default_factory=dict)
    seed: int = 1000
    use_rms: bool = True          # False: drop -tni/-ini -> gain/rdnoise noise model (functions.c:1172)
    use_mask: bool = True         # False: drop -imi
    replace: dict = dataclasses.field(default_factory=dict)   # flag -> new value in the zuds command
    flags: tuple = ()             # bare switches appended to the command (-okn, -convvar)
    swap: bool = False            # exchange the roles of the two images (the "template" is then the broader one)


SMALL = [
    # BASELINE configs[0] shape at reduced size: hw 10, rss 24, ko 2, bgo 0, convolve template
    Case("c1_small", (512, 512), 4.0, 1, {"ko": 2}, dict(nstars=300, bad_frac=2e-4)),
    # zuds-actual: seeing 2 -> hw 5, rss 12, default ko 4, 2x2 regions
    Case("zuds_actual_2x2", (448, 480), 2.0, 2, {}, dict(nstars=500, bad_frac=2e-4)),
    # convolve the science image, photometric normalisation to the template, bg order 1
    Case("conv_image", (400, 400), 2.4, 1, {"ko": 1, "bgo": 1}, dict(nstars=260, bad_frac=2e-4),
         replace={"-c": "i", "-n": "t"}),
    # no rms maps: variance from gain / read noise
    Case("gain_model", (384, 384), 2.0, 1, {"ko": 1, "tg": 6.0, "ig": 6.0, "tr": 8.0, "ir": 9.0},
         dict(nstars=240, bad_frac=2e-4), use_rms=False),
    # no usable stamps at all: region skipped, outputs stay zero (main.c:1002-1009)
    Case("no_stamps", (320, 320), 4.0, 1, {"ko": 2}, dict(nstars=8, bad_frac=2e-3)),
    # few stamps: the one-shot threshold retry (main.c:948-995)
    Case("retry", (512, 512), 3.0, 1, {"ko": 0, "ft": 1900.0}, dict(nstars=300, bad_frac=1e-4)),
    # different basis: 2 Gaussians, 3x3 regions, non-default clipping
    Case("two_gauss_3x3", (600, 640), 2.0, 3, {"ko": 1, "ng": "2 4 0.8 2 2.0", "ks": 1.5},
         dict(nstars=900, bad_frac=1e-4)),
    # cubic kernel variation, quadratic background, 2x1 regions on a non-square frame
    Case("ko3_bgo2_2x1", (420, 560), 2.0, 1, {"ko": 3, "bgo": 2}, dict(nstars=520, bad_frac=2e-4), replace={"-nrx": 2}),
    # no photometric scaling of the difference (-n u), science image convolved, no mask file
    Case("norm_u_nomask", (384, 400), 2.4, 1, {"ko": 1}, dict(nstars=260, bad_frac=2e-4), use_mask=False,
         replace={"-c": "i", "-n": "u"}),
]

# Options of the path that zuds does not pass by default.  Pinned against the reference on CPU (oracle vs golden);
# the GPU parity tests take them over from here once a round has GPU time to run them.
EXTRA = [
    # -okn: noise of OK pixels rescaled by the region's diffrat (main.c:1557-1576); -convvar: variance convolved
    # with the kernel instead of its square (alard.c:1300-1306)
    Case("okn_convvar", (400, 416), 2.2, 1, {"ko": 1}, dict(nstars=260, bad_frac=3e-4), flags=("-okn", "-convvar")),
    # the numeric knobs away from their defaults: 2 sub-stamps per stamp, stricter bad-tap fraction, wider input-mask
    # dilation, no output masking (-mous < 0, main.c:1584), 2.5-sigma statistics, other fill values
    Case("knobs", (440, 420), 2.0, 1, {"ko": 1, "nss": 2, "kfm": 0.9, "mins": 2.0, "mous": -1.0, "ssig": 2.5,
                                       "fi": -7.0, "fin": 99.0}, dict(nstars=420, bad_frac=3e-4)),
    # kernel re-evaluated every 7 pixels instead of every 2 hw + 1 (-kcs, alard.c:1272-1278)
    Case("kcs7", (384, 400), 2.0, 1, {"ko": 2, "kcs": 7}, dict(nstars=300, bad_frac=2e-4)),
    # ADU pedestals taken off both images before anything else (-tp / -ip, main.c:212-218,728-733), gain/read-noise model
    Case("pedestals", (384, 400), 2.2, 1, {"ko": 1, "tp": 20.0, "ip": 35.0, "tg": 6.0, "ig": 6.0, "tr": 8.0, "ir": 9.0},
         dict(nstars=260, bad_frac=2e-4), use_rms=False),
]

# -c b: hotpants' own default, the direction chosen per region by check_stamps (alard.c:483-724).  The synthetic
# reference is the sharper image, so the template wins unless the images are swapped.
BOTH = [
    Case("dir_b_2x2", (448, 480), 2.0, 2, {}, dict(nstars=500, bad_frac=2e-4), replace={"-c": "b"}),
    Case("dir_b_swapped", (400, 416), 2.4, 1, {"ko": 1, "bgo": 1}, dict(nstars=300, bad_frac=2e-4),
         replace={"-c": "b"}, swap=True),
    # no rms maps, 2 sub-stamps per stamp, stricter kernel-sum clip: stamps are dropped from the trial fit
    Case("dir_b_gain_3x1", (420, 600), 2.0, 1, {"ko": 2, "tg": 6.0, "ig": 6.0, "tr": 8.0, "ir": 9.0, "ks": 1.2, "nss": 2},
         dict(nstars=700, bad_frac=2e-4), use_rms=False, replace={"-c": "b", "-nry": 3}),
    # science PSF narrower than the reference's on the left, broader on the right: the two regions go opposite ways
    Case("dir_b_mixed_2x1", (400, 640), 2.4, 1, {"ko": 1}, dict(nstars=600, bad_frac=2e-4, ref_sharpness=1.0, seeing_gradient=0.35),
         replace={"-c": "b", "-nrx": 2}),
    # too few stamps at the first threshold in both directions: the one-shot retry re-searches both sets (main.c:921-995)
    Case("dir_b_retry", (512, 512), 3.0, 1, {"ko": 0, "ft": 1900.0}, dict(nstars=300, bad_frac=1e-4), replace={"-c": "b"}),
]

TINY_FULL = Case("tiny_full", (200, 208), 1.3, 1, {"ko": 1}, dict(nstars=120, bad_frac=2e-4))

# BASELINE.json configs at full size (GPU tests / bench only)
FULL = {
    "c1": Case("c1", (3080, 3072), 4.0, 3, {"ko": 2}),
    "c1b": Case("c1b", (3080, 3072), 2.0, 3, {}),
    "c4": Case("c4", (3080, 3072), 6.0, 1, {"ko": 4, "nsx": 32, "nsy": 32, "r": 15, "rss": 15},
               replace={"-c": "i"}),
    # coadd-scale frame, one global fit: the row-tile configuration (BASELINE.json configs[4])
    "c5": Case("c5", (6160, 6144), 4.0, 1, {"ko": 2, "nsx": 20, "nsy": 20}),
    # configs[1] with hotpants' own -c b on a pair whose science PSF is the narrower one on the left and the broader
    # one on the right: the 3x3 regions do not all go the same way
    "c1_dir_b": Case("c1_dir_b", (3080, 3072), 4.0, 3, {"ko": 2}, dict(ref_sharpness=1.0, seeing_gradient=0.35),
                     replace={"-c": "b"}),
}


def make_pair(case):
    pair = synth.make_pair(case.seed, case.shape, case.seeing, **case.synth_kws)
    if case.swap:
        pair = dataclasses.replace(pair, sci=pair.ref, ref=pair.sci, sci_rms=pair.ref_rms, ref_rms=pair.sci_rms)
    return pair


def command(case, pair, names=None, exe="hotpants"):
    """The zuds command line for this case (file names are placeholders unless given)."""
    names = names or {k: k + ".fits" for k in ("sci", "ref", "sci.rms", "ref.rms", "bpm")}
    ny, nx = pair.sci.shape
    scibkg, scistd = P.quick_background_estimate(pair.sci, pair.sci_mask)
    refbkg, refstd = P.quick_background_estimate(pair.ref, pair.ref_mask)
    il, tl = scibkg - 10 * scistd, refbkg - 10 * refstd
    out = names.get("out", "sub.fits")
    cmd = P.hotpants_command(names["sci"], names["ref"], out, names["sci.rms"], names["ref.rms"], names["bpm"],
                             pair.seeing, nx, ny, il, tl, nreg_side=case.nreg_side,
                             hotpants_kws=case.hotpants_kws, exe=exe)
    toks = cmd.split()
    if not case.use_rms:
        for f in ("-tni", "-ini"):
            i = toks.index(f)
            del toks[i:i + 2]
    if not case.use_mask:
        i = toks.index("-imi")
        del toks[i:i + 2]
    for f, v in case.replace.items():
        i = toks.index(f)
        toks[i + 1] = str(v)
    toks += list(case.flags)
    return " ".join(toks)


def params_of(case, pair):
    ny, nx = pair.sci.shape
    a = P.parse_hotpants_args(command(case, pair).split()[1:])
    a.params.nx, a.params.ny = nx, ny
    return a.params


def inputs_of(case, pair):
    """(sci, ref, sci_rms, ref_rms, sci_mask, ref_mask) as the command line selects them."""
    return (pair.sci, pair.ref, pair.sci_rms if case.use_rms else None, pair.ref_rms if case.use_rms else None,
            pair.bpm if case.use_mask else None, None)
