#!/bin/bash
# diagnostic: intermediates of check_stamps, GPU vs oracle, on the case whose merit differed at 2e-9
mkdir -p gpurun_out
ZB_CHECK_DEBUG=1 HP_ORACLE_DEBUG=1 timeout 300 python - > gpurun_out/s23_diag.txt 2>&1 <<'PY'
import sys
sys.path.insert(0, "tests")
import cases as CS, hp_helpers as H
from zuds_pipeline_b200 import _lib
lib = H.load_oracle()
ctx = _lib.Context(0)
for c in CS.BOTH[2:3]:
    pair = CS.make_pair(c); p = CS.params_of(c, pair); ins = CS.inputs_of(c, pair)
    sys.stderr.write("ORACLE %s\n" % c.name); sys.stderr.flush()
    o = H.run_oracle(lib, p, *ins)
    sys.stderr.write("GPU %s\n" % c.name); sys.stderr.flush()
    g = ctx.subtract_host(p, *ins)
    for r in range(o.sizes.n_regions):
        print(c.name, r, repr(o.regions[r].t_merit), repr(g.regions[r].t_merit), repr(o.regions[r].i_merit), repr(g.regions[r].i_merit))
PY
echo rc $?; wc -l gpurun_out/s23_diag.txt
