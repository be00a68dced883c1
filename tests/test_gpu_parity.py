"""Parity of the CUDA path (through the C ABI) against the CPU oracle and the reference's golden
vectors.  The bars (BASELINE.json north_star): masks and selected stamps bit-exact; difference and
noise within 1e-4 x local sigma; kernel coefficients within 1e-5 relative; sentinel pixels exact."""
import numpy as np
import pytest

import cases as CS
import hp_helpers as H
from test_oracle_golden import check_against_golden, input_hash, keywords_mismatch, load_golden, region_keywords

pytestmark = pytest.mark.gpu

TOL_SIGMA = 1e-4      # |delta diff|, |delta noise| <= TOL_SIGMA * noise_ref
TOL_COEF = 1e-5       # |delta coef| <= TOL_COEF * max|coef| of the region


def compare(o, g):
    """o: oracle Result, g: CUDA Result (numpy arrays)."""
    assert H.stamp_table(o) == H.stamp_table(g), "selected stamps / sscnt differ"
    for r in range(o.sizes.n_regions):
        ro, rg = o.regions[r], g.regions[r]
        assert (ro.flags, ro.n_stamps, ro.fit_iters, ro.n_skipped) == (rg.flags, rg.n_stamps, rg.fit_iters, rg.n_skipped)
        assert ro.conv_template == rg.conv_template, "region %d convolved the other way" % r
        for a, b in ((ro.t_merit, rg.t_merit), (ro.i_merit, rg.i_merit)):      # -c b: check_stamps' merits (0 otherwise)
            assert abs(a - b) <= 1e-12 * abs(a) or (a != a and b != b), "merit %r vs %r in region %d" % (a, b, r)
    # the last figure of merit of every stamp (getStampSig): it feeds the clip decisions, so it should agree far
    # below float precision (a float-rounded intermediate would show up at 6e-8)
    ns = o.sizes.n_stamps
    for r in range(o.sizes.n_regions):
        for k in range(o.regions[r].n_stamps):
            a, b = o.stamps[r * ns + k].chi2, g.stamps[r * ns + k].chi2
            assert abs(a - b) <= 1e-9 * abs(a) or (a != a and b != b), "chi2 %r vs %r, region %d stamp %d" % (a, b, r, k)
    assert np.array_equal(o.mask, g.mask), "hotpants int mask differs in %d pixels" % int((o.mask != g.mask).sum())
    scale = np.abs(o.sol).max(axis=1, keepdims=True) + 1e-300
    assert (np.abs(o.sol - g.sol) / scale).max() <= TOL_COEF
    sent = o.diff == np.float32(1e-30)
    assert np.array_equal(sent, g.diff == np.float32(1e-30)), "1e-30 sentinel pixels differ"
    # -convvar can drive the convolved variance negative next to BIG_RMS pixels: the reference then writes NaN noise
    # (sqrt of a negative sum, main.c:1320); NaNs must sit on the same pixels, everything else within tolerance
    nan_n, nan_d = np.isnan(o.noise), np.isnan(o.diff)
    assert np.array_equal(nan_n, np.isnan(g.noise)), "NaN noise pixels differ"
    assert np.array_equal(nan_d, np.isnan(g.diff)), "NaN difference pixels differ"
    sig = np.maximum(np.where(nan_n, 1.0, o.noise.astype(np.float64)), 1e-3)
    dd = np.where(nan_d, 0.0, np.abs(o.diff.astype(np.float64) - g.diff.astype(np.float64))) / sig
    dn = np.where(nan_n, 0.0, np.abs(o.noise.astype(np.float64) - g.noise.astype(np.float64))) / sig
    assert dd.max() <= TOL_SIGMA, "diff off by %.3g sigma" % dd.max()
    assert dn.max() <= TOL_SIGMA, "noise off by %.3g sigma" % dn.max()
    assert np.array_equal(o.noise[sent], g.noise[sent], equal_nan=True)
    bad = keywords_mismatch(region_keywords(g), region_keywords(o), exact=False)
    assert not bad, bad
    for r in range(o.sizes.n_regions):
        a, b = o.regions[r].nscalo, g.regions[r].nscalo
        assert abs(a - b) <= 2e-4 + 1e-4 * abs(a) or (a != a and b != b)


@pytest.mark.parametrize("case", CS.SMALL + CS.EXTRA + CS.BOTH + [CS.TINY_FULL], ids=lambda c: c.name)
def test_small_cases_host_call(case, oracle, gpu_ctx):
    pair = CS.make_pair(case)
    p = CS.params_of(case, pair)
    ins = CS.inputs_of(case, pair)
    o = H.run_oracle(oracle, p, *ins)
    g = gpu_ctx.subtract_host(p, *ins)
    assert gpu_ctx.launch_count() > 0
    compare(o, g)
    gold = load_golden(case.name)
    if input_hash(case, pair) != gold["input_sha256"]:
        pytest.fail("synthetic generator drifted on this host: golden vectors of %s not comparable" % case.name)
    check_against_golden(case, g, gold, exact=False)


@pytest.mark.parametrize("case", CS.SMALL[:3], ids=lambda c: c.name)
def test_small_cases_device_call(case, oracle, gpu_ctx):
    """torch CUDA tensors in and out: only data_ptr() crosses the C ABI."""
    import torch
    pair = CS.make_pair(case)
    p = CS.params_of(case, pair)
    ins = CS.inputs_of(case, pair)
    o = H.run_oracle(oracle, p, *ins)
    dev = [None if a is None else torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in ins]
    g = gpu_ctx.subtract(p, *dev)
    torch.cuda.synchronize()
    g.diff, g.noise, g.mask = g.diff.cpu().numpy(), g.noise.cpu().numpy(), g.mask.cpu().numpy()
    compare(o, g)


def test_device_call_orders_after_async_uploads(oracle, gpu_ctx):
    """Inputs still in flight on the caller's (default) stream when zb_subtract is called: stream handle 0 must mean
    the legacy default stream, not the context's private stream (round-1 advisor finding)."""
    import torch
    case = CS.SMALL[0]
    pair = CS.make_pair(case)
    p = CS.params_of(case, pair)
    ins = CS.inputs_of(case, pair)
    o = H.run_oracle(oracle, p, *ins)
    pinned = [None if a is None else torch.from_numpy(np.ascontiguousarray(a)).pin_memory() for a in ins]
    torch.cuda.synchronize()
    junk = torch.empty(64 << 20, dtype=torch.float32, device="cuda")
    for _ in range(8):
        junk.normal_()                              # keep the default stream busy so the uploads really are pending
    dev = [None if t is None else t.cuda(non_blocking=True) for t in pinned]
    g = gpu_ctx.subtract(p, *dev)                    # no synchronize in between
    torch.cuda.synchronize()
    g.diff, g.noise, g.mask = g.diff.cpu().numpy(), g.noise.cpu().numpy(), g.mask.cpu().numpy()
    compare(o, g)


def test_mask_dtypes_agree(gpu_ctx):
    case = CS.SMALL[0]
    pair = CS.make_pair(case)
    p = CS.params_of(case, pair)
    ins = list(CS.inputs_of(case, pair))
    base = gpu_ctx.subtract_host(p, *ins)
    for dt in (np.int16, np.int32, np.bool_):
        ins[4] = pair.bpm.astype(dt)
        g = gpu_ctx.subtract_host(p, *ins)
        assert np.array_equal(g.mask, base.mask) and np.array_equal(g.diff, base.diff)


def test_repeatable_and_context_reuse(gpu_ctx):
    """Same ctx, alternating configurations: the workspace is rebuilt and results are bit-identical."""
    a, b = CS.SMALL[0], CS.SMALL[1]
    pa, pb = CS.make_pair(a), CS.make_pair(b)
    r1 = gpu_ctx.subtract_host(CS.params_of(a, pa), *CS.inputs_of(a, pa))
    gpu_ctx.subtract_host(CS.params_of(b, pb), *CS.inputs_of(b, pb))
    r2 = gpu_ctx.subtract_host(CS.params_of(a, pa), *CS.inputs_of(a, pa))
    assert np.array_equal(r1.diff, r2.diff) and np.array_equal(r1.noise, r2.noise) and np.array_equal(r1.mask, r2.mask)
    assert np.array_equal(r1.sol, r2.sol)


def test_batch_call_equals_single_calls(gpu_ctx):
    """zb_subtract_batch (copies of neighbouring pairs overlap the computation) gives what zb_subtract_host gives pair by pair."""
    base = CS.SMALL[0]
    cases = [CS.Case(base.name, base.shape, base.seeing, base.nreg_side, base.hotpants_kws, base.synth_kws, seed=3000 + k) for k in range(5)]
    pairs = [CS.make_pair(c) for c in cases]
    params = [CS.params_of(c, p) for c, p in zip(cases, pairs)]
    ins = [CS.inputs_of(c, p) for c, p in zip(cases, pairs)]
    single = [gpu_ctx.subtract_host(p, *i) for p, i in zip(params, ins)]
    for want_mask in (True, False):
        got = gpu_ctx.subtract_batch(params, ins, want_mask=want_mask)
        assert len(got) == 5
        for a, b in zip(single, got):
            assert np.array_equal(a.diff, b.diff) and np.array_equal(a.noise, b.noise) and np.array_equal(a.sol, b.sol)
            assert (b.mask is None) == (not want_mask)
            if want_mask:
                assert np.array_equal(a.mask, b.mask)
            assert H.stamp_table(a) == H.stamp_table(b)


def test_bad_params_are_refused(gpu_ctx):
    from zuds_pipeline_b200 import _lib
    case = CS.SMALL[0]
    pair = CS.make_pair(case)
    p = CS.params_of(case, pair)
    p.hw_ksstamp = 3                      # rss < r : vargs.c FATAL ERROR
    with pytest.raises(_lib.ZbError) as e:
        gpu_ctx.subtract_host(p, *CS.inputs_of(case, pair))
    assert e.value.code == -1


def test_nan_and_constant_inputs(oracle, gpu_ctx):
    """Edge inputs the reference tolerates: NaN pixels, fill-value pixels, a dead (constant) quadrant."""
    case = CS.SMALL[0]
    pair = CS.make_pair(case)
    pair.sci[100:104, 200:203] = np.nan
    pair.ref[300, 310] = np.inf
    pair.sci[50, 60] = np.float32(1e-30)
    pair.ref[400:, 400:] = 150.0
    p = CS.params_of(case, pair)
    ins = CS.inputs_of(case, pair)
    o = H.run_oracle(oracle, p, *ins)
    g = gpu_ctx.subtract_host(p, *ins)
    assert H.stamp_table(o) == H.stamp_table(g)
    assert np.array_equal(o.mask, g.mask)
    fin = np.isfinite(o.diff) & np.isfinite(o.noise)
    assert np.array_equal(fin, np.isfinite(g.diff) & np.isfinite(g.noise))
    sig = np.maximum(o.noise.astype(np.float64), 1e-3)
    assert (np.abs(o.diff.astype(np.float64) - g.diff)[fin] / sig[fin]).max() <= TOL_SIGMA


QA_FIELDS = ("dmean", "dsige", "dsig", "dmeano", "dsigeo", "dsigo", "x2norm", "nx2norm", "nscalo")


def _qa_numbers(g):
    return [[float(getattr(g.regions[i], f)) for f in QA_FIELDS] for i in range(g.sizes.n_regions)]


@pytest.mark.parametrize("nan_pixels", [False, True])
def test_region_statistics_routes_agree(nan_pixels, oracle, gpu_ctx, monkeypatch):
    """The region statistics by their three routes: clip iterations >= 2 on the shell lists (default), full passes
    only (ZB_QA_NOTAIL=1), and getStampStats3 with its histogram phase (ZB_QA_HIST=1).  The two full-pass routes add
    the same terms in the same order: identical numbers; the shell route forms each iteration's sums as a
    difference, so it may differ in the last bits.  All of them are the oracle's keywords."""
    case = CS.SMALL[0]
    pair = CS.make_pair(case)
    if nan_pixels:                     # a NaN under a clean mask enters the OK-pixel runs: the shell lists must be given up
        pair.sci[100:104, 200:203] = np.nan
        pair.ref[300, 310] = np.inf
    p = CS.params_of(case, pair)
    ins = CS.inputs_of(case, pair)
    o = H.run_oracle(oracle, p, *ins)
    base = gpu_ctx.subtract_host(p, *ins)
    monkeypatch.setenv("ZB_QA_NOTAIL", "1")
    full = gpu_ctx.subtract_host(p, *ins)
    monkeypatch.delenv("ZB_QA_NOTAIL")
    monkeypatch.setenv("ZB_QA_HIST", "1")
    hist = gpu_ctx.subtract_host(p, *ins)
    monkeypatch.delenv("ZB_QA_HIST")
    assert _qa_numbers(full) == _qa_numbers(hist)
    for a, b in zip(_qa_numbers(base), _qa_numbers(full)):
        for x, y in zip(a, b):
            assert x == y or abs(x - y) <= 1e-10 * max(abs(x), abs(y)), (a, b)
    for g in (base, full, hist):
        assert np.array_equal(g.mask, o.mask)
        bad = keywords_mismatch(region_keywords(g), region_keywords(o), exact=False)
        assert not bad, bad


@pytest.mark.parametrize("name", ["c1", "c1b", "c4"])
def test_full_size_quadrant(name, oracle, gpu_ctx):
    """BASELINE.json configs[0]/[1] (and the zuds-actual variant) and configs[3] (hw 15, order 4,
    convolve-science, 32x32 stamp grid) at 3080x3072 against the oracle."""
    case = CS.FULL[name]
    pair = CS.make_pair(case)
    p = CS.params_of(case, pair)
    ins = CS.inputs_of(case, pair)
    o = H.run_oracle(oracle, p, *ins)
    g = gpu_ctx.subtract_host(p, *ins)
    compare(o, g)
    # zuds-level product: bit 17 where the difference carries the sentinel (zuds/subtraction.py:167-174)
    zo = (pair.sci_mask.astype(np.int64) | pair.ref_mask) | ((o.diff == np.float32(1e-30)) * 2 ** 17)
    zg = (pair.sci_mask.astype(np.int64) | pair.ref_mask) | ((g.diff == np.float32(1e-30)) * 2 ** 17)
    assert np.array_equal(zo, zg)


def test_full_size_properties(gpu_ctx):
    """Size-independent properties at full size: idempotent re-run, the photometric scaling only
    rescales (-n t vs -n i), and swapping the mask dtype changes nothing."""
    case = CS.FULL["c1"]
    pair = CS.make_pair(case)
    p = CS.params_of(case, pair)
    ins = CS.inputs_of(case, pair)
    a = gpu_ctx.subtract_host(p, *ins)
    b = gpu_ctx.subtract_host(p, *ins)
    assert np.array_equal(a.diff, b.diff) and np.array_equal(a.mask, b.mask) and np.array_equal(a.noise, b.noise)
    p2 = CS.params_of(case, pair)
    p2.phot_normalize = ord("t")
    c = gpu_ctx.subtract_host(p2, *ins)
    assert np.array_equal(a.mask, c.mask) and np.array_equal(a.sol, c.sol)
    good = (a.mask & 0x8000) == 0
    # every pixel scaled by 1/ksum of ITS region; check on the first region's interior
    r = a.regions[0]
    sl = (slice(r.ry_min + 30, r.ry_max - 30), slice(r.rx_min + 30, r.rx_max - 30))
    ratio = c.diff[sl][good[sl]].astype(np.float64) * r.ksum
    assert np.allclose(ratio, a.diff[sl][good[sl]], rtol=2e-6, atol=1e-5)
