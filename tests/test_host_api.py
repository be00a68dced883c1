"""Host side of the drop-in boundary: flag parsing, command building, FITS products and the
zuds-level API mirrors.  CPU tests feed the header/FITS writer with ORACLE results (the writer does
not care who computed them); GPU tests go through from_images / the hotpants CLI end to end."""
import os
import subprocess
import sys

import numpy as np
import pytest

import cases as CS
import hp_helpers as H
from zuds_pipeline_b200 import fitsio, hotpants, image, params as P, subtraction, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_flag_parsing_quirks():
    a = P.parse_hotpants_args("-inim a -tmplim b -outim c -r 6.25 -rss 15.0 -nsx 10.24 -nsy 10.266666 -tu 5000.0 -iu 5000.0 "
                              "-tl 124.2502586669922 -fin 223.60679774997897 -ko 4 -bgo 0 -c t -n i -hki -v 0".split())
    p = a.params
    assert (p.hw_kernel, p.hw_ksstamp, p.nstamp_x, p.nstamp_y) == (6, 15, 10, 10)      # sscanf("%d") truncation
    assert p.t_uthresh == 5000.0 and p.t_ukthresh == 25000.0 and p.i_ukthresh == 25000.0  # -tu does not move -tuk
    assert p.t_lthresh == np.float32(124.2502586669922) and p.fill_val_noise == np.float32(223.60679774997897)
    assert (chr(p.force_convolve), chr(p.phot_normalize), a.hki, a.verbose) == ("t", "i", True, 0)
    with pytest.raises(ValueError):
        P.parse_hotpants_args("-inim a -tmplim b".split())                             # -outim is required
    with pytest.raises(ValueError):
        P.parse_hotpants_args("-inim a -tmplim b -outim c -bogus 1".split())
    with pytest.raises(NotImplementedError):
        P.parse_hotpants_args("-inim a -tmplim b -outim c -pca 1".split())


def _objects(pair, d):
    """Image objects on disk the way zuds holds them before from_images."""
    hdr = fitsio.Header()
    for k, v in (("SEEING", pair.seeing), ("FIELD", 600), ("CCDID", 1), ("QID", 1), ("FID", 2), ("MAGZP", 26.3)):
        hdr[k] = v
    paths = {}
    for n, a in (("sci", pair.sci), ("sci.mask", pair.sci_mask), ("sci.rms", pair.sci_rms),
                 ("ref", pair.ref), ("ref.mask", pair.ref_mask), ("ref.rms", pair.ref_rms)):
        paths[n] = os.path.join(d, n + ".fits")
        fitsio.write_image(paths[n], a, hdr)
    sci = image.CalibratableImageBase.from_file(paths["sci"])
    ref = image.CalibratableImageBase.from_file(paths["ref"])
    assert sci.mask_image is not None and hasattr(sci, "_rmsimg") and hasattr(ref, "_rmsimg")
    ref.weight_image                  # the reference copies ref.weight_image (zuds/subtraction.py:86)
    return sci, ref


def test_prepare_hotpants_builds_the_zuds_command(tmp_path):
    case = CS.SMALL[0]
    pair = CS.make_pair(case)
    sci, ref = _objects(pair, str(tmp_path))
    submask = image.MaskImageBase()
    submask.basename = "sub.mask.fits"
    submask.data = pair.sci_mask.astype(np.int32) | pair.ref_mask
    submask.boolean.map_to_local_file(str(tmp_path / "sub.mask.bpm.fits"))
    out = str(tmp_path / "sub.fits")
    cmd = hotpants.prepare_hotpants(sci, ref, out, submask.boolean, tmp_path / "w", nreg_side=1,
                                    subtract_new_back=False, hotpants_kws={"ko": 2})
    ny, nx = pair.sci.shape
    il, tl = H.zuds_flags(pair)
    want = P.hotpants_command(sci.local_path, ref.local_path, out, sci.rms_image.local_path, ref.rms_image.local_path,
                              submask.boolean.local_path, pair.seeing, nx, ny, il, tl, nreg_side=1, hotpants_kws={"ko": 2})
    assert cmd == want
    assert np.array_equal(submask.boolean.data, pair.bpm.astype(bool))
    with pytest.raises(NotImplementedError):
        hotpants.prepare_hotpants(sci, ref, out, submask.boolean, tmp_path / "w")      # SExtractor is upstream


def test_rms_weight_roundtrip(tmp_path):
    case = CS.TINY_FULL
    pair = CS.make_pair(case)
    sci, _ = _objects(pair, str(tmp_path))
    w = sci.weight_image.data
    bad = (pair.sci_mask.astype(np.int32) & synth.BAD_SUM) > 0
    assert (w[bad] == 0).all() and np.allclose(w[~bad], 1 / pair.sci_rms[~bad] ** 2, rtol=1e-6)
    other = image.CalibratableImageBase.from_file(sci.local_path, load_others=False)
    other.mask_image = sci.mask_image
    other.weight_image = sci.weight_image
    r = other.rms_image.data
    assert (r[bad] == np.float32(synth.BIG_RMS)).all() and np.allclose(r[~bad], pair.sci_rms[~bad], rtol=1e-6)


def test_sub_name():
    assert subtraction.sub_name("/d/ztf_1_sciimg.fits", "/r/ref_2.fits") == "/d/sub.ztf_1_sciimg_ref_2.fits"


@pytest.mark.parametrize("case", [CS.SMALL[1], CS.SMALL[4], CS.BOTH[3]], ids=lambda c: c.name)
def test_fits_products_match_reference_binary(case, oracle, tmp_path):
    """Keyword set / order / values and pixel data of the written FITS files against the files the
    unmodified reference binary wrote for the same command (volatile AUTHOR/ORIGIN/DATE values aside)."""
    if not os.path.exists(H.REF_BIN):
        pytest.skip("oracle/_ref/hotpants not built on this host")
    pair = CS.make_pair(case)
    ref = H.run_reference(pair, case=case)
    args = P.parse_hotpants_args(ref.cmd.split()[1:])
    ny, nx = pair.sci.shape
    args.params.nx, args.params.ny = nx, ny
    res = H.run_oracle(oracle, args.params, *CS.inputs_of(case, pair))
    out = tmp_path / "mine"
    out.mkdir()
    for k in ("outim", "noise_image", "out_mask", "savexy"):
        args.files[k] = str(out / os.path.basename(args.files[k]))
    sci_hdr = fitsio.read_image(args.files["image"])[1]
    hotpants.write_outputs(args, sci_hdr, res, argv0=ref.cmd.split()[0])
    mine = fitsio.read_hdus(args.files["outim"])
    theirs = fitsio.read_hdus(os.path.join(ref.workdir, "sub.fits"))
    assert len(mine) == len(theirs) == 2
    volatile = {"AUTHOR", "ORIGIN", "DATE", "DIFFIM", "DIFFCMD"}
    for (hm, dm), (ht, dt) in zip(mine, theirs):
        km = [k for k in hm.keys() if k not in ("COMMENT", "")]
        kt = [k for k in ht.keys() if k not in ("COMMENT", "")]
        skip = {"XTENSION", "PCOUNT", "GCOUNT", "TFIELDS", "BITPIX", "NAXIS", "NAXIS1", "NAXIS2", "SIMPLE", "EXTEND"}
        assert [k for k in km if k not in skip and not k.startswith(("TUNIT",))] == \
               [k for k in kt if k not in skip and not k.startswith(("TUNIT",))]
        for k in km:
            if k in volatile or k in skip:
                continue
            vm, vt = hm[k], ht[k]
            if isinstance(vm, float) or isinstance(vt, float):
                assert float(vm) == pytest.approx(float(vt), rel=1e-6), k
            else:
                assert vm == vt, k
        assert np.array_equal(np.asarray(dm), np.asarray(dt))
    assert np.array_equal(fitsio.read_image(args.files["noise_image"])[0], ref.noise)
    hn, hr = fitsio.read_image(args.files["noise_image"])[1], ref.noise_header
    for k in ("OBJECT", "GAIN", "RDNOISE", "MASKVAL"):
        assert hn[k] == pytest.approx(hr[k]) if isinstance(hn[k], float) else hn[k] == hr[k]
    assert [k for k in hn.keys() if k.startswith("NSCALO")] == [k for k in hr.keys() if k.startswith("NSCALO")]
    assert np.array_equal(fitsio.read_image(args.files["out_mask"])[0].astype(np.int32) & 0xffff, ref.mask & 0xffff)
    # -savexy: the three ds9 region files, byte for byte (functions.c:75-127); a run whose only region is skipped writes none
    for suffix in ("", ".all", ".skipped"):
        theirs_xy = os.path.join(ref.workdir, "xy" + suffix)
        if os.path.exists(theirs_xy):
            assert open(args.files["savexy"] + suffix).read() == open(theirs_xy).read(), suffix
        else:
            assert not os.path.exists(args.files["savexy"] + suffix)


def test_usage_text_has_version_on_line_2(capsys):
    assert hotpants.main(["hotpants"]) == 1
    err = capsys.readouterr().err.splitlines()
    assert err[1].split() == ["Version", "5.1.11"]           # zuds/constants.py:91-95 parses exactly this


# ------------------------------------------------------------------------------ GPU

@pytest.mark.gpu
def test_from_images_end_to_end(oracle, tmp_path):
    case = CS.SMALL[0]
    pair = CS.make_pair(case)
    d = tmp_path / "data"
    d.mkdir()
    sci, ref = _objects(pair, str(d))
    sub = subtraction.SingleEpochSubtraction.from_images(sci, ref, tmpdir=str(tmp_path / "tmp"), subtract_back=False,
                                                         nreg_side=1, hotpants_kws={"ko": 2})
    final = subtraction.sub_name(sci.local_path, ref.local_path)
    assert sub.local_path == final
    for suffix in (".fits", ".rms.fits", ".mask.fits"):
        assert os.path.exists(final.replace(".fits", suffix))
    assert not os.listdir(tmp_path / "tmp")                   # transactional temp dir removed
    o = H.run_oracle(oracle, CS.params_of(case, pair), *CS.inputs_of(case, pair))
    sig = np.maximum(o.noise, 1e-3)
    assert (np.abs(sub.data - o.diff) <= 1e-4 * sig).all()
    assert (np.abs(sub.rms_image.data - o.noise) <= 1e-4 * sig).all()
    want_mask = (pair.sci_mask.astype(np.int64) | pair.ref_mask) | ((o.diff == np.float32(1e-30)) * 2 ** 17)
    assert np.array_equal(sub.mask_image.data.astype(np.int64), want_mask)
    assert sub.mask_image.header["BIT17"] == 17
    for k, v in (("FIELD", 600), ("CCDID", 1), ("QID", 1), ("FID", 2), ("SEEING", pair.seeing), ("MAGZP", 26.3)):
        assert sub.header[k] == v
    assert sub.header["SOFTNAME"] == "HOTPanTS" and sub.header["KSUM00"] == "%.4f" % o.regions[0].ksum
    assert len(fitsio.read_hdus(final)) == 2                  # -hki table survived the final save
    assert sub.reference_image is ref and sub.target_image is sci


@pytest.mark.gpu
def test_hotpants_cli_shim(oracle, tmp_path):
    case = CS.SMALL[2]                                        # -c i -n t -bgo 1
    pair = CS.make_pair(case)
    names = H.write_pair(pair, str(tmp_path))
    names["out"] = str(tmp_path / "sub.fits")
    cmd = CS.command(case, pair, names, exe=os.path.join(ROOT, "bin", "hotpants")) + " -omi %s" % (tmp_path / "m.fits")
    pr = subprocess.run([sys.executable] + cmd.split(), capture_output=True, text=True)
    assert pr.returncode == 0 and "SUCCESS" in pr.stderr        # main.c:1997 prints it at every verbosity
    o = H.run_oracle(oracle, CS.params_of(case, pair), *CS.inputs_of(case, pair))
    diff = fitsio.read_image(names["out"])[0]
    sig = np.maximum(o.noise, 1e-3)
    assert (np.abs(diff - o.diff) <= 1e-4 * sig).all()
    assert np.array_equal(fitsio.read_image(str(tmp_path / "m.fits"))[0].astype(np.int32) & 0xffff, o.mask & 0xffff)
    bad = subprocess.run([sys.executable, os.path.join(ROOT, "bin", "hotpants"), "-inim", "nope.fits", "-tmplim", "x", "-outim", "y"],
                         capture_output=True, text=True)
    assert bad.returncode != 0


@pytest.mark.gpu
def test_hotpants_cli_shim_default_direction(oracle, tmp_path):
    """No -c on the command line: hotpants' own default, -c b (vargs.c:55-128); the two regions of this frame end up
    convolved in opposite directions and CONVOLnn says so."""
    case = [c for c in CS.BOTH if c.name == "dir_b_mixed_2x1"][0]
    pair = CS.make_pair(case)
    names = H.write_pair(pair, str(tmp_path))
    names["out"] = str(tmp_path / "sub.fits")
    toks = CS.command(case, pair, names, exe=os.path.join(ROOT, "bin", "hotpants")).split()
    i = toks.index("-c")
    del toks[i:i + 2]
    pr = subprocess.run([sys.executable] + toks, capture_output=True, text=True)
    assert pr.returncode == 0 and "SUCCESS" in pr.stderr, pr.stderr[-2000:]
    o = H.run_oracle(oracle, CS.params_of(case, pair), *CS.inputs_of(case, pair))
    diff, hdr = fitsio.read_image(names["out"])
    assert (hdr["CONVOL00"], hdr["CONVOL01"]) == ("IMAGE", "TEMPLATE")
    assert (np.abs(diff - o.diff) <= 1e-4 * np.maximum(o.noise, 1e-3)).all()


def test_import_sets_hardware_queue_default():
    """the binding asks for 32 CUDA work queues before the first CUDA call unless the caller chose a value"""
    import os
    from zuds_pipeline_b200 import _lib  # noqa: F401
    assert os.environ.get("CUDA_DEVICE_MAX_CONNECTIONS") is not None


def test_fits_scaling_cards_do_not_travel(tmp_path):
    """Round-1 advisor finding: BSCALE/BZERO were applied on read but left in the header, so a saved copy was
    scaled twice; unsigned 16-bit masks could not be written at all."""
    from zuds_pipeline_b200 import fitsio
    raw = (np.arange(12, dtype=np.int16).reshape(3, 4))
    h = fitsio.Header()
    h["BSCALE"] = 2.0
    h["BZERO"] = 100.0
    # write the raw integers with scaling cards by hand (write_image itself refuses to emit non-neutral cards)
    path = str(tmp_path / "scaled.fits")
    fitsio.write_image(path, raw)
    buf = bytearray(open(path, "rb").read())
    cards = ("BSCALE  = %20s" % "2.0").ljust(80) + ("BZERO   = %20s" % "100.0").ljust(80)
    end = buf.index(b"END     ")
    buf[end:end + 240] = (cards + "END".ljust(80)).encode()
    open(path, "wb").write(buf)
    data, hdr = fitsio.read_image(path)
    assert np.array_equal(data, raw * 2.0 + 100.0)
    assert "BSCALE" not in hdr and "BZERO" not in hdr
    fitsio.write_image(str(tmp_path / "copy.fits"), data, hdr)
    again, _ = fitsio.read_image(str(tmp_path / "copy.fits"))
    assert np.array_equal(again, data)
    # unsigned 16-bit: int16 + BZERO 32768 on disk, uint16 back, also through a second save
    m = np.array([[0, 1, 32768, 65535]], np.uint16)
    fitsio.write_image(str(tmp_path / "m.fits"), m)
    back, mh = fitsio.read_image(str(tmp_path / "m.fits"))
    assert back.dtype == np.uint16 and np.array_equal(back, m) and "BZERO" not in mh
    fitsio.write_image(str(tmp_path / "m2.fits"), back, mh)
    assert np.array_equal(fitsio.read_image(str(tmp_path / "m2.fits"))[0], m)


def test_read_header_alone_equals_the_header_of_a_full_read(tmp_path):
    """from_images re-opens the products it has just written by their headers only (the pixels are in memory)."""
    from zuds_pipeline_b200 import fitsio
    h = fitsio.Header()
    for i in range(300):                                      # more than the first chunk read_header tries
        h["KEY%04d" % i if i < 100 else "K%d" % i] = i * 0.5
    h["OBJECT"] = "a string value"
    for arr in (np.arange(12, dtype=np.float32).reshape(3, 4), np.arange(12, dtype=np.uint16).reshape(3, 4)):
        path = str(tmp_path / ("x_%s.fits" % arr.dtype.name))
        fitsio.write_image(path, arr, h)
        full = fitsio.read_image(path)[1]
        assert fitsio.read_header(path).cards == full.cards


def test_background_estimate_follows_float32_numpy():
    """zuds/utils.py:32-53 on float32 pixels: np.median stays float32 (the two central values of an even count are
    averaged in float32) and so do the deviations; only the 1.4826 factor is double."""
    from zuds_pipeline_b200 import params as P
    rng = np.random.default_rng(5)
    x = (150 + 3 * rng.normal(size=10000)).astype(np.float32)
    m = np.zeros(x.shape, np.uint8)
    bkg, std = P.quick_background_estimate(x, m)
    b32 = np.median(x)
    assert b32.dtype == np.float32 and bkg == float(b32)
    assert std == 1.4826 * float(np.median(np.abs(x - b32)))
