"""Options of the path outside the zuds default command line (tests/cases.py:EXTRA) through the CUDA path, against the
oracle and the reference's golden vectors.  The oracle side of these cases is pinned on CPU (test_oracle_golden.py).

History: three cases were added when round 1 had seconds of GPU time left.  In the one B200 run they got, `ko3_bgo2_2x1`
and `norm_u_nomask` passed and moved into tests/cases.py:SMALL; `okn_convvar` did not: with -convvar the reference itself
writes NaN noise where the convolved variance goes negative next to BIG_RMS pixels (2122 pixels in this case), and
test_gpu_parity.compare() was not NaN-aware then.  compare() now requires NaNs on the same pixels; this case has not run
on a GPU since.  `knobs` and `kcs7` were added after the round's GPU time was spent and have never run on one.  All three
stay non-strict xfails (the file sorts last so that a fault here cannot disturb other tests).
Whoever sees one pass on a B200 moves it into SMALL; if it fails, the message names the parity difference to fix.
"""
import pytest

import cases as CS
import hp_helpers as H
from test_gpu_parity import compare
from test_oracle_golden import check_against_golden, input_hash, load_golden

pytestmark = pytest.mark.gpu


@pytest.mark.xfail(strict=False, reason="not yet confirmed on a GPU (see the module docstring)")
@pytest.mark.parametrize("case", CS.EXTRA, ids=lambda c: c.name)
def test_extra_cases_host_call(case, oracle, gpu_ctx):
    pair = CS.make_pair(case)
    p = CS.params_of(case, pair)
    ins = CS.inputs_of(case, pair)
    o = H.run_oracle(oracle, p, *ins)
    g = gpu_ctx.subtract_host(p, *ins)
    assert gpu_ctx.launch_count() > 0
    compare(o, g)
    gold = load_golden(case.name)
    if input_hash(case, pair) == gold["input_sha256"]:
        check_against_golden(case, g, gold, exact=False)
