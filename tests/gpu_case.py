"""GPU-box helper: run named parity cases (tests/cases.py) through the C ABI and report differences
against the oracle.  Usage: python tests/gpu_case.py c1_small [zuds_actual_2x2 ...]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import cases as CS          # noqa: E402
import hp_helpers as H      # noqa: E402
from zuds_pipeline_b200 import _lib  # noqa: E402


def main():
    names = sys.argv[1:] or [c.name for c in CS.SMALL]
    allc = {c.name: c for c in CS.SMALL + [CS.TINY_FULL] + list(CS.FULL.values())}
    ctx = _lib.Context(0)
    lib = H.load_oracle()
    for nm in names:
        case = allc[nm]
        pair = CS.make_pair(case)
        p = CS.params_of(case, pair)
        ins = CS.inputs_of(case, pair)
        g = ctx.subtract_host(p, *ins)
        print(nm, "gpu ok", {k: round(v, 3) for k, v in ctx.timings().items()}, flush=True)
        if os.environ.get("NO_ORACLE"):
            continue
        o = H.run_oracle(lib, p, *ins)
        print("   stamps equal", H.stamp_table(o) == H.stamp_table(g), "sol bit-equal", np.array_equal(o.sol, g.sol),
              "mask equal", np.array_equal(o.mask, g.mask), "diff bit-equal", np.array_equal(o.diff, g.diff, equal_nan=True),
              "noise bit-equal", np.array_equal(o.noise, g.noise, equal_nan=True), flush=True)
        for r in range(o.sizes.n_regions):
            ro, rg = o.regions[r], g.regions[r]
            for f in ("ksum", "sssig", "fsig", "fscat", "x2norm", "nx2norm", "dmean", "dsige", "dsig", "dmeano", "dsigeo", "dsigo", "nscalo"):
                if "%.4f" % getattr(ro, f) != "%.4f" % getattr(rg, f):
                    print("   region", r, f, getattr(ro, f), getattr(rg, f))


if __name__ == "__main__":
    main()
