"""Wrapper pre-step on the GPU: zb_background_estimate against the reference's numpy code
(zuds/utils.py:32-53 as restated in params.quick_background_estimate), bit for bit."""
import numpy as np
import pytest

import cases as CS
from zuds_pipeline_b200 import params as P

pytestmark = pytest.mark.gpu


def _same(a, b):
    return all((x == y) or (x != x and y != y) for x, y in zip(a, b))


def _host(data, mask):
    if mask is None:
        mask = np.zeros(data.shape, np.uint8)
    with np.errstate(all="ignore"):
        import warnings
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            return P.quick_background_estimate(data, mask)


def test_synthetic_pair_all_mask_dtypes(gpu_ctx):
    pair = CS.make_pair(CS.SMALL[0])
    for data, mask in ((pair.sci, pair.sci_mask), (pair.ref, pair.ref_mask), (pair.sci, pair.bpm), (pair.ref, None),
                       (pair.sci, pair.sci_mask.astype(np.int32)), (pair.sci, pair.bpm.astype(bool))):
        assert _same(gpu_ctx.background_estimate(data, mask), _host(data, mask))


@pytest.mark.parametrize("n", [1, 2, 3, 4, 7, 100, 101, 4096, 65537])
def test_small_counts_even_and_odd(gpu_ctx, n):
    rng = np.random.default_rng(n)
    data = (rng.normal(0, 1, n) * np.exp(rng.uniform(-20, 20, n))).astype(np.float32)      # both signs, 17 decades
    mask = (rng.uniform(size=n) < 0.3).astype(np.int16) * 5
    assert _same(gpu_ctx.background_estimate(data, mask), _host(data, mask))
    assert _same(gpu_ctx.background_estimate(data, None), _host(data, None))


def test_degenerate_inputs(gpu_ctx):
    data = np.full(1000, 150.0, np.float32)
    assert _same(gpu_ctx.background_estimate(data, None), (150.0, 0.0))
    assert _same(gpu_ctx.background_estimate(data, np.ones(1000, np.uint8)), (float("nan"), float("nan")))     # nothing unmasked
    data[17] = np.nan
    assert _same(gpu_ctx.background_estimate(data, None), (float("nan"), float("nan")))                        # numpy: NaN poisons the median
    m = np.zeros(1000, np.uint8)
    m[17] = 1
    assert _same(gpu_ctx.background_estimate(data, m), (150.0, 0.0))                                           # ... unless it is masked
    data = np.array([-0.0, 0.0, -1e-30, 1e-30, np.inf, -np.inf, 3.0, -3.0], np.float32)
    assert _same(gpu_ctx.background_estimate(data, None), _host(data, None))
    ties = np.repeat(np.array([1.0, 2.0, 2.0, 3.0], np.float32), 250)
    assert _same(gpu_ctx.background_estimate(ties, None), _host(ties, None))


def test_full_quadrant_device_and_host_calls(gpu_ctx):
    import time
    import torch
    pair = CS.make_pair(CS.FULL["c1"])
    t0 = time.perf_counter()
    want = _host(pair.sci, pair.sci_mask)
    t_host = time.perf_counter() - t0
    assert _same(gpu_ctx.background_estimate(pair.sci, pair.sci_mask), want)
    d, m = torch.from_numpy(pair.sci).cuda(), torch.from_numpy(pair.sci_mask).cuda()
    gpu_ctx.background_estimate(d, m)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    got = gpu_ctx.background_estimate(d, m)
    t_dev = time.perf_counter() - t0
    assert _same(got, want)
    print("quick_background_estimate 3080x3072: numpy %.1f ms, GPU (device-resident) %.3f ms" % (1e3 * t_host, 1e3 * t_dev))


def test_prepare_hotpants_command_is_unchanged_by_the_gpu_prestep(tmp_path):
    from test_host_api import _objects
    from zuds_pipeline_b200 import hotpants, image
    pair = CS.make_pair(CS.SMALL[0])
    sci, ref = _objects(pair, str(tmp_path))
    submask = image.MaskImageBase()
    submask.basename = "sub.mask.fits"
    submask.data = pair.sci_mask.astype(np.int32) | pair.ref_mask
    submask.boolean.map_to_local_file(str(tmp_path / "sub.mask.bpm.fits"))
    cmds = []
    for mode in ("host", "gpu"):
        hotpants.prestep = mode
        try:
            cmds.append(hotpants.prepare_hotpants(sci, ref, str(tmp_path / "sub.fits"), submask.boolean, tmp_path / "w", nreg_side=1,
                                                  subtract_new_back=False, hotpants_kws={"ko": 2}))
        finally:
            hotpants.prestep = "gpu"
    assert cmds[0] == cmds[1]
