"""Experiment: does bulk host<->device DMA traffic slow the subtraction kernels down?  Runs the device-resident
batch (as bench.py's `value`) with and without a background thread that keeps both copy directions busy with
dummy pinned buffers.  Usage: python tools/exp_dma.py [workers]"""
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402
import cases as CS  # noqa: E402
from zuds_pipeline_b200.batch import GpuPool  # noqa: E402


class NullStream:
    cuda_stream = -1


def main():
    W = int(sys.argv[1]) if len(sys.argv) > 1 else 16
    case = CS.FULL["c1"]
    pair = CS.make_pair(case)
    p = CS.params_of(case, pair)
    ins = [None if a is None else torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in CS.inputs_of(case, pair)]
    ny, nx = case.shape
    pool = GpuPool(0, W)
    outs = [(torch.empty((ny, nx), dtype=torch.float32, device="cuda"), torch.empty((ny, nx), dtype=torch.float32, device="cuda"),
             torch.empty((ny, nx), dtype=torch.int32, device="cuda")) for _ in range(W)]
    slot = {id(c): i for i, c in enumerate(pool.ctxs)}

    def job(ctx):
        return ctx.subtract(p, *ins, out=outs[slot[id(ctx)]], stream=NullStream())

    last = {}

    def rate(n):
        t0 = time.perf_counter()
        _, tim = pool.run([job] * n)
        dt = time.perf_counter() - t0
        for k in ("k_convolve", "fit_loop", "stamp_search", "fill_stamps", "region_stats", "input_mask"):
            last[k] = round(float(np.mean([t[k] for t in tim])), 3)
        return n / dt

    rate(2 * W)
    print("quiet   : %.1f pairs/s" % rate(6 * W), last, flush=True)
    stop = threading.Event()
    moved = [0, 0]

    def noise(direction):
        hb = torch.empty(128 << 20, dtype=torch.uint8).pin_memory()
        db = torch.empty(128 << 20, dtype=torch.uint8, device="cuda")
        st = torch.cuda.Stream()
        with torch.cuda.stream(st):
            while not stop.is_set():
                if direction == 0:
                    db.copy_(hb, non_blocking=True)
                else:
                    hb.copy_(db, non_blocking=True)
                st.synchronize()
                moved[direction] += hb.numel()
    for which, label in (((0,), "h2d busy"), ((1,), "d2h busy"), ((0, 1), "both busy")):
        stop.clear()
        moved[0] = moved[1] = 0
        ths = [threading.Thread(target=noise, args=(d,)) for d in which]
        for t in ths:
            t.start()
        time.sleep(0.3)
        t0 = time.perf_counter()
        m0 = list(moved)
        r = rate(6 * W)
        dt = time.perf_counter() - t0
        stop.set()
        for t in ths:
            t.join()
        print("%-8s: %.1f pairs/s  (background copies: h2d %.1f GB/s, d2h %.1f GB/s)" % (label, r, (moved[0] - m0[0]) / dt / 1e9, (moved[1] - m0[1]) / dt / 1e9), last, flush=True)
    pool.close()


if __name__ == "__main__":
    main()
