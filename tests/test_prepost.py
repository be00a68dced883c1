"""The wrapper's elementwise pre/post steps on the GPU against the numpy statements they replace
(zuds/mask.py:42-72, zuds/subtraction.py:135,167-177, zuds/image.py:137-208), bit for bit, and the device-resident
route through the zuds-level API (subtraction.subtract_on_device) against prepare_hotpants + the oracle."""
import numpy as np
import pytest

import cases as CS
import hp_helpers as H
from zuds_pipeline_b200 import params as P, subtraction, synth

pytestmark = pytest.mark.gpu


def _t(a):
    import torch
    a = np.ascontiguousarray(a)
    return torch.from_numpy(a.view(np.uint8) if a.dtype == np.bool_ else a).cuda()


@pytest.mark.parametrize("dts", [(np.int16, np.int32), (np.int32, np.int32), (np.uint8, np.int16)])
def test_mask_combine_and_boolean(gpu_ctx, dts):
    rng = np.random.default_rng(3)
    hi = 200 if dts[0] == np.uint8 else 2 ** 15 - 1
    a = rng.integers(0, hi, size=(257, 131)).astype(dts[0])
    b = (rng.integers(0, 2 ** 18, size=(257, 131)) * (rng.random((257, 131)) < 0.3)).astype(dts[1])
    o, boo = gpu_ctx.mask_combine(_t(a), _t(b), synth.BAD_SUM)
    want = a.astype(np.int32) | b.astype(np.int32)
    assert np.array_equal(o.cpu().numpy(), want)
    assert np.array_equal(boo.cpu().numpy(), ((want & synth.BAD_SUM) > 0).astype(np.uint8))
    o2, b2 = gpu_ctx.mask_combine(_t(a), None, synth.BAD_SUM)              # a single mask: its own boolean (zuds/mask.py)
    assert np.array_equal(o2.cpu().numpy(), a.astype(np.int32))
    assert np.array_equal(b2.cpu().numpy(), ((a.astype(np.int32) & synth.BAD_SUM) > 0).astype(np.uint8))


def test_flag_sentinel_is_exact_equality(gpu_ctx):
    diff = np.array([1e-30, 1.0000001e-30, 0.0, -1e-30, np.nan, 3.5], np.float32)
    diff[1] = np.nextafter(np.float32(1e-30), np.float32(1))
    mask = np.arange(6, dtype=np.int32)
    got = gpu_ctx.mask_flag_sentinel(_t(mask), _t(diff)).cpu().numpy()
    want = mask.copy()
    want[diff == np.float32(1e-30)] |= 2 ** 17
    assert np.array_equal(got, want) and got[0] == 2 ** 17 and got[1] == 1


def test_rms_weight_round_trip_matches_numpy(gpu_ctx):
    rng = np.random.default_rng(7)
    n = 50000
    w = rng.uniform(0, 0.2, n).astype(np.float32)
    w[::97] = 0.0                                              # 1/sqrt(0) = inf, as numpy gives
    bad = (rng.random(n) < 0.01).astype(np.uint8)
    data = rng.uniform(0, 60000, n).astype(np.float32)
    sat = 50000.0
    with np.errstate(divide="ignore"):
        want = (1 / np.sqrt(w)).astype(np.float32)
    want[bad > 0] = synth.BIG_RMS
    want[data >= 0.9 * sat] = synth.BIG_RMS
    got = gpu_ctx.rms_from_weight(_t(w), _t(bad), _t(data), sat).cpu().numpy()
    assert np.array_equal(got, want)
    rms = rng.uniform(0.5, 300, n).astype(np.float32)
    want = (1 / rms ** 2).astype(np.float32)
    want[bad > 0] = 0.0
    want[data >= 0.9 * sat] = 0.0
    got = gpu_ctx.weight_from_rms(_t(rms), _t(bad), _t(data), sat).cpu().numpy()
    assert np.array_equal(got, want)
    assert np.array_equal(gpu_ctx.weight_from_rms(_t(rms)).cpu().numpy(), (1 / rms ** 2).astype(np.float32))


def test_subtract_on_device_matches_the_wrapper_and_the_oracle(oracle, gpu_ctx):
    """Device tensors in, device tensors out: same command line as prepare_hotpants builds from the same numbers,
    pixels / masks as the oracle run with that command."""
    import torch
    case = CS.SMALL[0]
    pair = CS.make_pair(case)
    prod = subtraction.subtract_on_device(gpu_ctx, _t(pair.sci), _t(pair.ref), _t(pair.sci_rms), _t(pair.ref_rms),
                                          _t(pair.sci_mask), _t(pair.ref_mask), pair.seeing, nreg_side=case.nreg_side,
                                          hotpants_kws=case.hotpants_kws)
    torch.cuda.synchronize()
    # the command: thresholds from the two masked background estimates, exactly as zuds/hotpants.py:65-93
    scibkg, scistd = P.quick_background_estimate(pair.sci, pair.sci_mask)
    refbkg, refstd = P.quick_background_estimate(pair.ref, pair.ref_mask)
    ny, nx = pair.sci.shape
    want_cmd = P.hotpants_command("sci.fits", "ref.fits", "sub.fits", "sci.rms.fits", "ref.rms.fits", "sub.mask.bpm.fits", pair.seeing,
                                  nx, ny, scibkg - 10 * scistd, refbkg - 10 * refstd, nreg_side=case.nreg_side, hotpants_kws=case.hotpants_kws)
    assert prod.command == want_cmd
    o = H.run_oracle(oracle, CS.params_of(case, pair), *CS.inputs_of(case, pair))
    assert np.array_equal(prod.boolean.cpu().numpy(), pair.bpm)
    diff, noise = prod.diff.cpu().numpy(), prod.noise.cpu().numpy()
    sig = np.maximum(o.noise.astype(np.float64), 1e-3)
    assert (np.abs(diff.astype(np.float64) - o.diff) <= 1e-4 * sig).all() and (np.abs(noise.astype(np.float64) - o.noise) <= 1e-4 * sig).all()
    want_mask = (pair.sci_mask.astype(np.int64) | pair.ref_mask) | ((o.diff == np.float32(1e-30)) * 2 ** 17)
    assert np.array_equal(prod.mask.cpu().numpy().astype(np.int64), want_mask)
    assert H.stamp_table(o) == H.stamp_table(prod.result)


def test_from_images_accepts_a_weight_map_for_the_reference(tmp_path):
    """A reference with only a weight map (what zuds' coadds carry): its rms map is made on the device
    (zuds/image.py:179-190) and the result equals the run with the rms map given."""
    import os
    from test_host_api import _objects
    from zuds_pipeline_b200 import fitsio
    case = CS.SMALL[0]
    pair = CS.make_pair(case)
    out = []
    for mode in ("rms", "weight"):
        d = tmp_path / mode
        d.mkdir()
        sci, ref = _objects(pair, str(d))
        if mode == "weight":
            os.remove(ref.rms_image.local_path)
            del ref._rmsimg
        sub = subtraction.SingleEpochSubtraction.from_images(sci, ref, tmpdir=str(tmp_path / "tmp"), subtract_back=False,
                                                             nreg_side=1, hotpants_kws={"ko": 2})
        out.append((np.asarray(sub.data), np.asarray(sub.rms_image.data), np.asarray(sub.mask_image.data)))
    # 1/sqrt(1/rms^2) is rms to within an ulp, so the two runs agree to the tolerance, masks exactly
    assert np.array_equal(out[0][2], out[1][2])
    sig = np.maximum(out[0][1], 1e-3)
    assert (np.abs(out[0][0] - out[1][0]) <= 1e-3 * sig).all()
