"""zb_background_estimate_static (include/zb200.h): the background estimate for callers that cannot allocate.  It has
no counterpart in the reference; its oracle is the existing entry point,
zb_background_estimate, which tests/test_bkg.py holds bit for bit against the reference's numpy statement.

Collected last on purpose: a failure here hides nothing else.  The test stands on its
own in a fresh process (it creates what it needs) and after the other tests (the static scratch then holds whatever
earlier calls left in it; every call clears it)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _same(a, b):
    return all((x == y) or (x != x and y != y) for x, y in zip(a, b))


def _inputs(seed, n, mask_dtype):
    rng = np.random.default_rng(seed)
    data = (150.0 + 12.0 * rng.standard_normal(n)).astype(np.float32)
    data[rng.integers(0, n, n // 50)] += np.float32(4000.0)                       # stars
    mask = (rng.uniform(size=n) < 0.2).astype(mask_dtype) * 3
    return data, mask


def test_static_scratch_variant_equals_existing_entry_point(gpu_ctx):
    import torch
    from zuds_pipeline_b200 import _lib
    dev = torch.device("cuda", gpu_ctx.device)
    torch.cuda.set_device(dev)                                                    # the entry point runs on the current device
    out = torch.full((2,), -1.0, dtype=torch.float64, device=dev)
    # two calls in a row on different inputs (sizes, mask types, one without a mask): the second finds the first's state
    calls = [_inputs(11, 300 * 1000 + 7, np.uint8), _inputs(12, 64 * 1024, np.int32), (_inputs(13, 5000, np.int16)[0], None)]
    tensors = [(torch.from_numpy(d).to(dev), None if m is None else torch.from_numpy(m).to(dev)) for d, m in calls]
    want = [gpu_ctx.background_estimate(d, m) for d, m in tensors]                # existing entry point, same device tensors
    torch.cuda.synchronize(dev)
    got, free = [], []
    for d, m in tensors:
        free.append(torch.cuda.mem_get_info(dev)[0])
        _lib.background_estimate_static(d, m, out)
        torch.cuda.synchronize(dev)
        free.append(torch.cuda.mem_get_info(dev)[0])
        got.append(tuple(out.cpu().tolist()))
    for g, w in zip(got, want):
        assert _same(g, w), (g, w)                                                # exact order statistics: identical, as test_bkg.py asks
    assert np.isfinite(got[0][0]) and got[0][1] > 0
    # after the first call (which resolves the symbol and loads the kernels) a call leaves the device's free memory alone
    assert free[2] == free[3] and free[4] == free[5], free
    # and once more on the first input: nothing of the calls in between survives in the scratch
    _lib.background_estimate_static(tensors[0][0], tensors[0][1], out)
    torch.cuda.synchronize(dev)
    assert _same(tuple(out.cpu().tolist()), want[0])


def test_static_scratch_variant_refuses_what_it_cannot_do(gpu_ctx):
    import torch
    from zuds_pipeline_b200 import _lib
    dev = torch.device("cuda", gpu_ctx.device)
    torch.cuda.set_device(dev)
    d = torch.zeros(1024, dtype=torch.float32, device=dev)
    out = torch.full((2,), -1.0, dtype=torch.float64, device=dev)
    with pytest.raises(_lib.ZbError) as e:                                        # refused before anything is launched
        _lib.background_estimate_static(d, None, out, n=_lib.BKG_STATIC_MAX_PIXELS + 1)
    assert e.value.code == -1 and "exceed the maximum" in str(e.value)
    with pytest.raises(_lib.ZbError) as e:
        _lib.background_estimate_static(d, None, out, stream=_lib.OwnStream())    # no context, so no context stream
    assert e.value.code == -1
    torch.cuda.synchronize(dev)
    assert out.cpu().tolist() == [-1.0, -1.0]                                     # nothing ran
    _lib.background_estimate_static(d, None, out)                                 # the entry point still works afterwards
    torch.cuda.synchronize(dev)
    assert out.cpu().tolist() == [0.0, 0.0]
