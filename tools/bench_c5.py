"""BASELINE.json configs[4]: one coadd-scale 6160x6144 pair, ONE global fit (-nrx 1 -nry 1 -nsx 20 -nsy 20),
split into row tiles over the GPUs of one box (zuds_pipeline_b200/tiles.py: P2P halo exchange,
all-reduced per-stamp normal equations).  Strong scaling: the work is one frame whatever N is.

    python tools/bench_c5.py                         # N = 1: zb_subtract on the whole frame
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/bench_c5.py

Every rank holds only its own rows on the device when the timed region starts; a step = halo exchange
+ search + merged fit + convolution of the own block rows + gather on rank 0 + QA statistics.  Timed
with CUDA events on the session stream, max over ranks; rank 0 prints one JSON line and (with
--check) compares the tiled products with the single-GPU zb_subtract of the same frame bit for bit.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--check", action="store_true")
    ap.add_argument("--profile", action="store_true", help="one extra step with a stream sync at every phase boundary")
    ap.add_argument("--shape", type=int, nargs=2, default=None)
    a = ap.parse_args()
    import torch
    import torch.distributed as dist
    import cases as CS
    from zuds_pipeline_b200 import _lib, tiles

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    case = CS.FULL["c5"]
    if a.shape:
        case = CS.Case("c5_custom", tuple(a.shape), 4.0, 1, {"ko": 2, "nsx": 20, "nsy": 20})
    t0 = time.time()
    pair = CS.make_pair(case)
    p = CS.params_of(case, pair)
    ins = CS.inputs_of(case, pair)
    t_synth = time.time() - t0
    ny, nx = pair.sci.shape
    names = ("sci", "ref", "sci_rms", "ref_rms", "sci_mask", "ref_mask")
    ctx = _lib.Context(local)
    st = torch.cuda.Stream(device=dev)
    infos = tiles.plan_all(p, world)
    me = infos[rank]

    def own_rows():
        """full-frame device tensors holding ONLY this rank's rows (the rest poisoned)"""
        out = []
        for arr in ins:
            if arr is None:
                out.append(None)
                continue
            arr = np.ascontiguousarray(arr.view(np.uint8) if arr.dtype == np.bool_ else arr)
            t = torch.empty((ny, nx), dtype=torch.from_numpy(arr[:1]).dtype, device=dev)
            t.fill_(float("nan") if t.dtype == torch.float32 else 77)
            t[me.row0:me.row1] = torch.from_numpy(arr[me.row0:me.row1]).to(dev)
            out.append(t)
        return out

    with torch.cuda.stream(st):
        mine = own_rows()
        pristine = [None if t is None else t.clone() for t in mine]
        comm = tiles.DistComm() if world > 1 else None
        times, stats, res = [], {}, None
        for it in range(a.warmup + a.steps):
            for t, q in zip(mine, pristine):          # halo rows back to "not received yet"
                if t is not None:
                    t.copy_(q)
            st.synchronize()
            if world > 1:
                dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st)
            if world > 1:
                res = tiles.subtract_tiled(ctx, comm, p, *mine, stats=stats)
            else:
                res = ctx.subtract(p, *mine, stream=st)
            e1.record(st)
            st.synchronize()
            ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
            if world > 1:
                dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            if it >= a.warmup:
                times.append(float(ms.item()))
        prof = None
        if a.profile:
            if world > 1:
                for t, q in zip(mine, pristine):
                    if t is not None:
                        t.copy_(q)
                st.synchronize()
                dist.barrier()
                ps = {"profile": True}
                tiles.subtract_tiled(ctx, comm, p, *mine, stats=ps)
                allp = [None] * world
                dist.all_gather_object(allp, ps.get("phases_ms"))
                prof = allp
            else:
                prof = [{k: round(v, 3) for k, v in ctx.timings().items()}]
        ok = None
        if a.check and rank == 0:
            full = [None if arr is None else torch.from_numpy(np.ascontiguousarray(arr.view(np.uint8) if arr.dtype == np.bool_ else arr)).to(dev) for arr in ins]
            ctx2 = _lib.Context(local)
            g = ctx2.subtract(p, *full, stream=st)
            st.synchronize()
            ok = bool(torch.equal(g.mask, res.mask) and torch.equal(g.diff.view(torch.int32), res.diff.view(torch.int32))
                      and torch.equal(g.noise.view(torch.int32), res.noise.view(torch.int32)) and np.array_equal(g.sol, res.sol))
            ctx2.close()
    if rank == 0:
        ms = float(np.median(times))
        line = {"metric": "coadd-scale frame subtractions/sec (one global fit, row tiles)", "value": 1000.0 / ms, "unit": "frames/s",
                "n_gpus": world, "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms, "ms_all": [round(t, 3) for t in times],
                "higher_is_better": True, "scaling": "strong", "dtype": "f64", "data": "synthetic",
                "config": {"workload": "c5: %dx%d pair, -c t -n i -r 10 -rss 24 -ko 2 -bgo 0 -nrx 1 -nry 1 -nsx 20 -nsy 20 (BASELINE configs[4])" % (ny, nx),
                           "tiles": [i.as_dict() for i in infos] if world > 1 else None,
                           "collectives": "P2P halo rows (inputs, then search-claimed mask rows), all-reduce(SUM) of search records / per-stamp Gram+rhs / per-stamp sigma, gather of output rows on rank 0"},
                "n_stamps": int(res.regions[0].n_stamps), "fit_iters": int(res.regions[0].fit_iters),
                "tile_stats_rank0": {k: v for k, v in stats.items() if k != "info"}, "bit_equal_to_single_gpu": ok, "phases_ms_per_rank": prof,
                "synth_s": round(t_synth, 1)}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
