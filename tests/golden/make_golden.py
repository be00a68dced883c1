"""Generate the golden vectors under tests/golden/ by running the UNMODIFIED reference
(oracle/_ref/hotpants, built from /root/reference/hotpants by oracle/Makefile) on the seeded
synthetic cases of tests/cases.py.  Run in the build container:

    make -C oracle ref && python tests/golden/make_golden.py

What is stored per case (golden_<name>.json): the exact command line, a sha256 of the inputs (so a
host whose numpy produces different synthetic pixels is detected, not mis-compared), sha256 of the
reference's diff / noise / 16-bit mask images, the -hki kernel table, the -savexy stamp centres,
every per-region header keyword, for -c b cases the two merits per region that the reference prints
("Result : merit = %.3f", main.c:1037,1053), and an 8x8 centre patch of diff and noise (the same kind of
known-answer check the reference's own test uses, zuds/tests/suite/test_sub.py:8-36).
tiny_full.npz additionally carries complete inputs and outputs of one very small case.
"""
import hashlib
import json
import os
import re
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import cases as CS          # noqa: E402
import hp_helpers as H      # noqa: E402


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def input_hash(case, pair):
    h = hashlib.sha256()
    for a in CS.inputs_of(case, pair):
        if a is not None:
            h.update(np.ascontiguousarray(a).tobytes())
    return h.hexdigest()


REGION_KEYS = ("REGION", "CONVOL", "KSUM", "SSSIG", "SSSCAT", "FSIG", "FSCAT", "X2NRM", "NX2NRM",
               "DMEAN", "DSIGE", "DSIG", "DMEANO", "DSIGEO", "DSIGO")


def golden_of(case):
    pair = CS.make_pair(case)
    ref = H.run_reference(pair, case=case)
    ny, nx = pair.sci.shape
    cy, cx = ny // 2, nx // 2
    g = {
        "name": case.name,
        "command": " ".join(os.path.basename(t) if t.endswith(".fits") or t.endswith("hotpants") else t
                            for t in ref.cmd.split()[:-4]),
        "input_sha256": input_hash(case, pair),
        "diff_sha256": sha(ref.diff.astype(np.float32)),
        "noise_sha256": sha(ref.noise.astype(np.float32)),
        "mask_sha256": sha((ref.mask & 0xffff).astype(np.int32)),
        "sol": None if ref.sol is None else [[repr(float(v)) for v in row] for row in ref.sol],
        "xy_all": ref.xy_all, "xy_used": ref.xy_used,
        "diff_patch": [[repr(float(v)) for v in row] for row in ref.diff[cy - 4:cy + 4, cx - 4:cx + 4]],
        "noise_patch": [[repr(float(v)) for v in row] for row in ref.noise[cy - 4:cy + 4, cx - 4:cx + 4]],
        "n_masked_1e30": int((ref.diff == np.float32(1e-30)).sum()),
        "region_keys": {k: ref.header[k] for k in ref.header.keys()
                        if any(k.startswith(p) and k[len(p):].isdigit() for p in REGION_KEYS)},
        "global_keys": {k: ref.header[k] for k in ("PHOTNORM", "NREGION", "MASKVAL", "KERINFO", "SOFTNAME", "SOFTVERS")
                        if k in ref.header},
        "noise_keys": {k: ref.noise_header[k] for k in ref.noise_header.keys()
                       if k in ("OBJECT", "GAIN", "RDNOISE", "MASKVAL") or k.startswith("NSCALO")},
        "table_keys": {} if ref.table_header is None else
                      {k: ref.table_header[k] for k in ref.table_header.keys()
                       if k in ("NGAUSS", "FWKERN", "CKORDER", "BGORDER", "KERINFO", "EXTNAME")
                       or k.startswith("DGAUSS") or k.startswith("SGAUSS")},
    }
    if "-c b" in ref.cmd:
        g["merits"] = re.findall(r"Result : merit = (\S+)", ref.stderr)      # template, image per region that has stamps
    return g, pair, ref


def main():
    only = set(sys.argv[1:])
    for case in CS.SMALL + CS.EXTRA + CS.BOTH + [CS.TINY_FULL]:
        if only and case.name not in only:
            continue
        g, pair, ref = golden_of(case)
        with open(os.path.join(HERE, "golden_%s.json" % case.name), "w") as f:
            json.dump(g, f, indent=1, sort_keys=True)
        print(case.name, "stamps used", len(ref.xy_used), "masked", g["n_masked_1e30"])
        if case is CS.TINY_FULL:
            np.savez_compressed(os.path.join(HERE, "tiny_full.npz"),
                                sci=pair.sci, ref=pair.ref, sci_rms=pair.sci_rms, ref_rms=pair.ref_rms, bpm=pair.bpm,
                                diff=ref.diff.astype(np.float32), noise=ref.noise.astype(np.float32),
                                mask=(ref.mask & 0xffff).astype(np.uint16),
                                sol=ref.sol, xy_all=np.array(ref.xy_all), xy_used=np.array(ref.xy_used))


if __name__ == "__main__":
    main()
