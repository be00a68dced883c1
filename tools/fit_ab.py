"""Serial stage times of full-size workloads for one build of the library (A/B of fit-chain changes):
mean of 4 subtractions after 2 warm-ups, from the context's CUDA events.
    ZB_LIB=build/<variant>.so python tools/fit_ab.py <label> c1 c1b"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import cases as CS          # noqa: E402
from zuds_pipeline_b200 import _lib  # noqa: E402


def main():
    label, names = sys.argv[1], sys.argv[2:]
    ctx = _lib.Context(0)
    for nm in names:
        case = CS.FULL[nm]
        pair = CS.make_pair(case)
        p = CS.params_of(case, pair)
        ins = CS.inputs_of(case, pair)
        acc = {}
        for i in range(6):
            g = ctx.subtract_host(p, *ins)
            if i >= 2:
                for k, v in ctx.timings().items():
                    acc.setdefault(k, []).append(v)
        rec = {"arm": label, "workload": nm, "fit_iters": [g.regions[r].fit_iters for r in range(g.sizes.n_regions)],
               "stage_ms": {k: round(float(np.mean(v)), 4) for k, v in acc.items() if not k.endswith("kclk")},
               "sol_sha": __import__("hashlib").sha256(np.ascontiguousarray(g.sol).tobytes()).hexdigest()[:16]}
        print(json.dumps(rec), flush=True)


if __name__ == "__main__":
    main()
