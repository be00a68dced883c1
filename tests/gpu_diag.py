"""GPU-box diagnostic: run the CUDA path and the CPU oracle on the same synthetic pair and report,
stage by stage, where they first differ.  Usage: python tests/gpu_diag.py ny nx seeing nreg "{kws}" """
import sys, time, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import hp_helpers as H
from zuds_pipeline_b200 import synth, params as P, _lib


def main():
    ny, nx, seeing, nreg = int(sys.argv[1]), int(sys.argv[2]), float(sys.argv[3]), int(sys.argv[4])
    kws = eval(sys.argv[5]) if len(sys.argv) > 5 else {}
    pair = synth.make_pair(1000, (ny, nx), seeing)
    il, tl = H.zuds_flags(pair)
    cmd = P.hotpants_command("sci", "ref", "sub.fits", "srms", "rrms", "bpm", seeing, nx, ny, il, tl, nreg, kws)
    a = H.params_from_command(cmd, nx, ny)
    p = a.params
    lib = H.load_oracle()
    t = time.time(); o = H.run_oracle(lib, p, pair.sci, pair.ref, pair.sci_rms, pair.ref_rms, pair.bpm, None)
    print("oracle %.2fs rc=%d" % (time.time() - t, o.rc))
    ctx = _lib.Context(0)
    for it in range(2):
        t = time.time(); g = ctx.subtract_host(p, pair.sci, pair.ref, pair.sci_rms, pair.ref_rms, pair.bpm, None)
        print("gpu host-call %.3fs" % (time.time() - t), {k: round(v, 3) for k, v in ctx.timings().items()})
    so, sg = H.stamp_table(o), H.stamp_table(g)
    print("stamps equal:", so == sg, len(so), len(sg))
    if so != sg:
        for x, y in zip(so, sg):
            if x != y:
                print("  first stamp diff:", x, y); break
    for r in range(o.sizes.n_regions):
        ro, rg = o.regions[r], g.regions[r]
        print("region", r, "nst", ro.n_stamps, rg.n_stamps, "iters", ro.fit_iters, rg.fit_iters, "flags", ro.flags, rg.flags,
              "ksum %.10f %.10f" % (ro.ksum, rg.ksum), "sssig %.6f %.6f" % (ro.sssig, rg.sssig), "skip", ro.n_skipped, rg.n_skipped)
    ds = np.abs(o.sol - g.sol)
    print("sol: bit-equal", np.array_equal(o.sol, g.sol), "max abs", ds.max(), "max rel-to-maxcoef", (ds.max(axis=1) / np.abs(o.sol).max(axis=1)).max())
    print("mask: equal", np.array_equal(o.mask, g.mask), "ndiff", int((o.mask != g.mask).sum()))
    if not np.array_equal(o.mask, g.mask):
        yy, xx = np.nonzero(o.mask != g.mask)
        for k in range(min(5, len(yy))):
            print("   mask diff at", yy[k], xx[k], hex(o.mask[yy[k], xx[k]]), hex(g.mask[yy[k], xx[k]]))
    for name, A, B in (("diff", o.diff, g.diff), ("noise", o.noise, g.noise)):
        d = np.abs(A.astype(np.float64) - B.astype(np.float64))
        sig = np.maximum(o.noise.astype(np.float64), 1e-3)
        print(name, ": bit-equal", np.array_equal(A, B), "ndiff", int((A != B).sum()), "max abs", d.max(), "max |d|/sigma", (d / sig).max())
        if d.max() > 0:
            k = np.argmax(d); print("    worst at", np.unravel_index(k, d.shape), A.flat[k], B.flat[k])
    for r in range(o.sizes.n_regions):
        ro, rg = o.regions[r], g.regions[r]
        for f in ("fsig", "fscat", "x2norm", "nx2norm", "dmean", "dsige", "dsig", "dmeano", "dsigeo", "dsigo", "nscalo"):
            vo, vg = getattr(ro, f), getattr(rg, f)
            if "%.4f" % vo != "%.4f" % vg:
                print("  region", r, f, vo, vg)
    print("launches", ctx.launch_count())


if __name__ == "__main__":
    main()
