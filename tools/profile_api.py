"""Where the time of one zuds-level call goes: dosub.do_one -> SingleEpochSubtraction.from_images on FITS files
(single thread, cProfile, top of the cumulative list).   python tools/profile_api.py [npairs]"""
import cProfile
import os
import pstats
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import cases as CS          # noqa: E402
from zuds_pipeline_b200 import dosub, fitsio  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 3
    case = CS.FULL["c1"]
    pair = CS.make_pair(case)
    with tempfile.TemporaryDirectory(prefix="zb_api_") as d:
        hdr = fitsio.Header()
        for k, v in (("SEEING", pair.seeing), ("FIELD", 600), ("CCDID", 1), ("QID", 1), ("FID", 2), ("MAGZP", 26.3)):
            hdr[k] = v
        lines = []
        for i in range(n + 1):
            sd = os.path.join(d, "p%03d" % i)
            os.mkdir(sd)
            for nm, a in (("sci", pair.sci), ("sci.mask", pair.sci_mask), ("sci.rms", pair.sci_rms),
                          ("ref", pair.ref), ("ref.mask", pair.ref_mask), ("ref.rms", pair.ref_rms),
                          ("ref.weight", (1.0 / pair.ref_rms.astype(np.float64) ** 2).astype(np.float32))):
                fitsio.write_image(os.path.join(sd, nm + ".fits"), a, hdr)
            lines.append("%s %s" % (os.path.join(sd, "sci.fits"), os.path.join(sd, "ref.fits")))
        kws = dict(case.hotpants_kws)

        def work(line):
            return dosub.do_one(line, tmpdir=d, quiet=True, subtract_back=False, nreg_side=case.nreg_side, hotpants_kws=kws)
        work(lines[0])                                       # warm-up: context, workspace
        pr = cProfile.Profile()
        t0 = time.perf_counter()
        pr.enable()
        for ln in lines[1:]:
            work(ln)
        pr.disable()
        dt = time.perf_counter() - t0
    print("%.3f s per pair (single thread)" % (dt / n))
    pstats.Stats(pr).sort_stats("cumulative").print_stats(32)


if __name__ == "__main__":
    main()
