"""One arm of the k_convolve arithmetic comparison (profiles/r02_conv_arith.md): for each full-size workload, the
kernel's CUDA-event time (mean of 4 serial subtractions) and the worst pixel against the oracle.
    ZB_LIB=build/<variant>.so python tools/conv_ab.py <label> c1 c1b c4"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import cases as CS          # noqa: E402
import hp_helpers as H      # noqa: E402
from zuds_pipeline_b200 import _lib  # noqa: E402


def main():
    label, names = sys.argv[1], sys.argv[2:]
    ctx = _lib.Context(0)
    lib = H.load_oracle()
    for nm in names:
        case = CS.FULL[nm]
        pair = CS.make_pair(case)
        p = CS.params_of(case, pair)
        ins = CS.inputs_of(case, pair)
        ms = []
        for i in range(6):
            g = ctx.subtract_host(p, *ins)
            if i >= 2:
                ms.append(ctx.timings()["k_convolve"])
        rec = {"arm": label, "workload": nm, "k_convolve_ms": round(float(np.mean(ms)), 4), "ms_all": [round(m, 4) for m in ms]}
        if not os.environ.get("NO_ORACLE"):
            o = H.run_oracle(lib, p, *ins)
            sig = np.maximum(o.noise.astype(np.float64), 1e-3)
            ok = np.isfinite(o.diff) & np.isfinite(o.noise)
            dd = np.abs(o.diff.astype(np.float64) - g.diff.astype(np.float64))[ok] / sig[ok]
            dn = np.abs(o.noise.astype(np.float64) - g.noise.astype(np.float64))[ok] / sig[ok]
            rec.update(worst_diff_sigma=float(dd.max()), worst_noise_sigma=float(dn.max()),
                       diff_pixels_not_identical=int((o.diff != g.diff).sum()), noise_pixels_not_identical=int((o.noise != g.noise).sum()),
                       pixels=int(o.diff.size), mask_equal=bool(np.array_equal(o.mask, g.mask)),
                       stamps_equal=H.stamp_table(o) == H.stamp_table(g))
        print(json.dumps(rec), flush=True)


if __name__ == "__main__":
    main()
