"""Options of the path outside the zuds default command line (tests/cases.py:EXTRA: -okn + -convvar, ko 3 / bgo 2 on
2x1 regions, -n u without a mask file) through the CUDA path, against the oracle and the reference's golden vectors.

The oracle side of these cases is pinned on CPU (test_oracle_golden.py).  The cases were added when round 1 had no GPU
time left, so the CUDA side has not run yet: non-strict xfail keeps an unexpected difference from masking the rest of
the suite, and the file sorts last so that a fault here cannot disturb other tests.  Whoever sees them pass on a B200
moves them into test_gpu_parity.py's table; whoever sees one fail has a parity bug to fix.
"""
import pytest

import cases as CS
import hp_helpers as H
from test_gpu_parity import compare
from test_oracle_golden import check_against_golden, input_hash, load_golden

pytestmark = pytest.mark.gpu


@pytest.mark.xfail(strict=False, reason="first GPU run of these cases decides (added without GPU time left in round 1)")
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
