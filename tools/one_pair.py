"""One full-size subtraction of a workload (for captures under ncu):   python tools/one_pair.py c1b"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import cases as CS          # noqa: E402
from zuds_pipeline_b200 import _lib  # noqa: E402

case = CS.FULL[sys.argv[1] if len(sys.argv) > 1 else "c1"]
pair = CS.make_pair(case)
p = CS.params_of(case, pair)
ctx = _lib.Context(0)
g = ctx.subtract_host(p, *CS.inputs_of(case, pair))
print("done", [g.regions[r].fit_iters for r in range(g.sizes.n_regions)])
