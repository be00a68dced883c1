#!/usr/bin/env python
"""bench.py — ZTF quadrant subtractions per second on N B200s (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W [--workload c1|c1b|c1_dir_b|c4|c5|batch256]   # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K ... [--workload ...]           # the reference hotpants on the host cores

Workloads ("Notation and sizes"; the same string goes into `config.workload` of both arms):
  c1        BASELINE configs[1] (default): 3080x3072 pair, -c t -n i -r 10 -rss 24 -ko 2 -bgo 0, 3x3 regions
  c1b       the same frame as zuds really calls it at seeing 2 px: -r 5 -rss 12 -ko 4
  c4        BASELINE configs[3]: -c i -r 15 -rss 15 -ko 4, one region, 32x32 stamps
  batch256  BASELINE configs[2]: 256 DISTINCT seeded pairs (seeds 1000..1255) sharded over the ranks by image
  c5        BASELINE configs[4]: one 6160x6144 frame, one global fit, row tiles over the ranks (strong scaling)
A "step" is one batch of pair subtractions per GPU (c5: one frame).  Pairs shard by image across ranks with no
collective: scaling is "weak"; c5 exchanges halo rows and all-reduces per-stamp normal equations.

  value     pairs/s with the inputs already resident in HBM (zb_subtract on device pointers)
  e2e       pairs/s through the host-buffer C-ABI call (zb_subtract_batch) as zuds calls hotpants (no -omi): every
            subtraction copies its five input arrays from pinned host memory and its two output images back
  e2e_api   pairs/s through the zuds-level API mirror (dosub.do_one -> Subtraction.from_images) from FITS files on
            local disk to FITS products on local disk (N = 1, bounded sample)
  roofline  the convolution/output kernel (k_convolve), ALGORITHMIC bytes 29 B/pixel over its
            CUDA-event duration, against MEASURED_PEAKS.json's HBM copy bandwidth; the FP32 pipe that really
            bounds it at hw >= 5 is reported beside it
  cpu_baseline  the unmodified reference binary oracle/_ref/hotpants (or the oracle port) on one full pair on one core
"""
from __future__ import annotations

import argparse
import csv
import json
import os
import statistics
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

# hardware work queues for the concurrent contexts (see zuds_pipeline_b200/_lib.py); must precede CUDA initialisation
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "ZTF quadrant subtractions/sec"
UNIT = "quadrant-pairs/s"
ALG_BYTES_PER_PIXEL = 29          # 4 float inputs + 1 mask byte in, 2 floats + int32 mask out
FP32_PEAK_TFMA = 36.06            # measured on this pool's B200 with tools/fp32_peak.cu (profiles/r02_fp32_peak_b200.txt): 124 FFMA/clk/SM

WORKLOADS = {
    "c1": ("c1", "c1: 3080x3072 quadrant pair, -c t -n i -r 10 -rss 24 -ko 2 -bgo 0, 3x3 regions x 10x10 stamps (BASELINE configs[1])"),
    "c1b": ("c1b", "c1b: 3080x3072 quadrant pair as zuds calls it at seeing 2 px, -c t -n i -r 5 -rss 12 -ko 4 -bgo 0, 3x3 regions x 10x10 stamps"),
    "c4": ("c4", "c4: 3080x3072 quadrant pair, -c i -n i -r 15 -rss 15 -ko 4 -bgo 0, 1 region x 32x32 stamps (BASELINE configs[3])"),
    "batch256": ("c1", "batch256: 256 distinct 3080x3072 pairs (seeds 1000..1255) as c1, sharded by image over the ranks (BASELINE configs[2])"),
    "c1_dir_b": ("c1_dir_b", "c1_dir_b: c1 with hotpants' own default -c b (direction chosen per region by check_stamps) on a pair whose science PSF is narrower than the reference's on the left and broader on the right"),
    "c5": ("c5", "c5: one 6160x6144 pair, -c t -n i -r 10 -rss 24 -ko 2 -bgo 0 -nrx 1 -nry 1 -nsx 20 -nsy 20, row tiles over the ranks (BASELINE configs[4])"),
}


def workload_case(name="c1"):
    import cases as CS
    return CS.FULL[WORKLOADS[name][0]]


def ncu_traffic(workload):
    """dram__bytes_read.sum + dram__bytes_write.sum of one k_convolve launch from the committed raw ncu export
    (profiles/r02_ncu_k_convolve_<workload>.csv = `ncu -i ... --page raw --csv`), or (None, None)."""
    path = os.path.join(ROOT, "profiles", "r02_ncu_k_convolve_%s.csv" % workload)
    if not os.path.exists(path):
        return None, None
    try:
        with open(path, newline="") as f:
            rows = list(csv.reader(f))
        hdr = rows[0]
        units = rows[1] if len(rows) > 1 and not rows[1][0].strip().isdigit() else None
        kcol = hdr.index("Kernel Name")
        data = [r for r in rows[1:] if r and r[0].strip().isdigit() and "k_convolve" in r[kcol]]
        if not data:
            return None, None
        scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        tot = []
        for r in data:
            t = 0.0
            for name in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                c = hdr.index(name)
                t += float(r[c].replace(",", "")) * (scale.get(units[c], 1.0) if units else 1.0)
            tot.append(t)
        return statistics.mean(tot), os.path.relpath(path, ROOT)
    except (ValueError, IndexError, OSError):
        return None, None


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


# --------------------------------------------------------------------------- clocks

class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            out = self.proc.communicate(timeout=5)[0]
        except subprocess.TimeoutExpired:
            self.proc.kill()
            out = self.proc.communicate()[0]
        sm, mx, reasons = [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for line in out.splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# --------------------------------------------------------------------------- CPU side

def write_pair_fits(pair, d):
    import hp_helpers as H
    return H.write_pair(pair, d)


def reference_command(case, pair, names, outdir, tag):
    import cases as CS
    import hp_helpers as H
    n = dict(names)
    n["out"] = os.path.join(outdir, "sub_%s.fits" % tag)
    exe = H.REF_BIN if os.path.exists(H.REF_BIN) else None
    return exe, (CS.command(case, pair, n, exe=exe or "hotpants"))


def run_cpu_once(case, pair, names, outdir, nproc):
    """`nproc` concurrent single-threaded hotpants processes (how zuds runs it: srun -n 64 -c1,
    nersc/controller.py:101), each on one full pair.  Returns wall seconds and the kind used."""
    import hp_helpers as H
    if os.path.exists(H.REF_BIN):
        procs = []
        t0 = time.perf_counter()
        for k in range(nproc):
            _, cmd = reference_command(case, pair, names, outdir, "p%d" % k)
            procs.append(subprocess.Popen(cmd.split(), stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL))
        rcs = [p.wait() for p in procs]
        dt = time.perf_counter() - t0
        if any(rcs):
            raise RuntimeError("reference hotpants exited with %s" % rcs)
        return dt, "reference"
    # oracle port (single process, `nproc` threads: ctypes releases the GIL)
    import cases as CS
    lib = H.load_oracle()
    p = CS.params_of(case, pair)
    ins = CS.inputs_of(case, pair)
    ths = [threading.Thread(target=H.run_oracle, args=(lib, p) + tuple(ins)) for _ in range(nproc)]
    t0 = time.perf_counter()
    for t in ths:
        t.start()
    for t in ths:
        t.join()
    return time.perf_counter() - t0, "port"


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def bench_reference(args):
    """--impl reference: the reference's own CPU implementation of the workload on all host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import cases as CS
    case = workload_case(args.workload)
    pair = CS.make_pair(case)
    import hp_helpers as H
    nproc = args.cpu_procs or host_cores()
    if args.workload == "c5":
        nproc = min(nproc, 8)          # 16 concurrent coadd-scale processes leave no room for their page cache
    if not os.path.exists(H.REF_BIN):
        nproc = 1                      # the oracle port keeps hotpants' global ran1 state: one thread only
    with tempfile.TemporaryDirectory(prefix="zb_ref_") as d:
        names = write_pair_fits(pair, d)
        for _ in range(min(args.warmup, 1)):      # host code has no clocks/caches to warm beyond one pass
            run_cpu_once(case, pair, names, d, nproc)
        times = []
        kind = "reference"
        for _ in range(args.steps):
            dt, kind = run_cpu_once(case, pair, names, d, nproc)
            times.append(dt)
    total = sum(times)
    value = nproc * len(times) / total
    c5 = args.workload == "c5"
    unit = "frames/s" if c5 else UNIT
    line = {
        "impl": "reference", "metric": "coadd-scale frame subtractions/sec (one global fit)" if c5 else METRIC,
        "value": value, "unit": unit, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times),
        "higher_is_better": True, "scaling": "strong" if c5 else "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOADS[args.workload][1]},
        "run": {"pairs_per_step": nproc},
        "cpu_baseline": {"value": value, "unit": unit, "cores": nproc, "kind": kind,
                         "sample": "%d concurrent single-threaded hotpants processes per step, one full pair each (FITS in/out on local disk)" % nproc},
        "e2e": {"value": value, "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


# --------------------------------------------------------------------------- GPU side: pairs

def _case_with_seed(case, seed):
    import cases as CS
    return CS.Case(case.name, case.shape, case.seeing, case.nreg_side, case.hotpants_kws, case.synth_kws, seed=seed,
                   use_rms=case.use_rms, use_mask=case.use_mask, replace=case.replace, flags=case.flags)


def _make_one(arg):
    """(worker process) one seeded pair of the workload + its zb_params bytes."""
    import cases as CS
    name, seed = arg
    c = _case_with_seed(workload_case(name), seed)
    pair = CS.make_pair(c)
    p = CS.params_of(c, pair)
    return seed, CS.inputs_of(c, pair), bytes(p)


def make_pairs(workload, seeds, nproc):
    """Distinct seeded pairs, generated in worker processes BEFORE this process touches CUDA."""
    from zuds_pipeline_b200 import params as P
    out = {}
    if nproc > 1 and len(seeds) > 2:
        import multiprocessing as mp
        with mp.get_context("fork").Pool(min(nproc, len(seeds))) as pool:
            for seed, ins, pb in pool.imap_unordered(_make_one, [(workload, s) for s in seeds]):
                out[seed] = (ins, P.ZbParams.from_buffer_copy(pb))
    else:
        for sd in seeds:
            seed, ins, pb = _make_one((workload, sd))
            out[seed] = (ins, P.ZbParams.from_buffer_copy(pb))
    return [out[sd] for sd in seeds]


def api_leg(case, pair, device, npairs, threads):
    """pairs/s through dosub.do_one -> SingleEpochSubtraction.from_images: FITS on local disk in, FITS products out."""
    from concurrent.futures import ThreadPoolExecutor
    from zuds_pipeline_b200 import dosub, fitsio
    from zuds_pipeline_b200.subtraction import SingleEpochSubtraction
    SingleEpochSubtraction.device = device
    with tempfile.TemporaryDirectory(prefix="zb_api_") as d:
        hdr = fitsio.Header()
        for k, v in (("SEEING", pair.seeing), ("FIELD", 600), ("CCDID", 1), ("QID", 1), ("FID", 2), ("MAGZP", 26.3)):
            hdr[k] = v
        lines = []
        for i in range(npairs + threads):
            sd = os.path.join(d, "p%03d" % i)
            os.mkdir(sd)
            for n, a in (("sci", pair.sci), ("sci.mask", pair.sci_mask), ("sci.rms", pair.sci_rms),
                         ("ref", pair.ref), ("ref.mask", pair.ref_mask), ("ref.rms", pair.ref_rms),
                         ("ref.weight", (1.0 / pair.ref_rms.astype(np.float64) ** 2).astype(np.float32))):
                fitsio.write_image(os.path.join(sd, n + ".fits"), a, hdr)
            lines.append("%s %s" % (os.path.join(sd, "sci.fits"), os.path.join(sd, "ref.fits")))
        kws = dict(case.hotpants_kws)

        def work(line):
            return dosub.do_one(line, tmpdir=d, quiet=True, subtract_back=False, nreg_side=case.nreg_side, hotpants_kws=kws)
        with ThreadPoolExecutor(max_workers=threads) as ex:
            list(ex.map(work, lines[:threads]))                    # warm-up: contexts, workspaces
            t0 = time.perf_counter()
            subs = list(ex.map(work, lines[threads:]))
            dt = time.perf_counter() - t0
        assert all(sb is not None for sb in subs)
    return {"value": npairs / dt, "unit": UNIT, "pairs": npairs, "threads": threads,
            "call": "zuds_pipeline_b200.dosub.do_one -> SingleEpochSubtraction.from_images (FITS files on local disk in and out, wrapper steps included)"}


def bench_pairs(args):
    import cases as CS
    from zuds_pipeline_b200 import batch

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    pinned_cores, box_cores = None, host_cores()
    if world > 1 and not args.no_pin:
        # one rank per GPU, each on its own share of the box's cores (what a production launcher does):
        # the worker threads of a rank yield to each other, not to another rank's
        try:
            avail = sorted(os.sched_getaffinity(0))
            share = len(avail) // world
            if share >= 1:
                pinned_cores = avail[local * share:(local + 1) * share]
                os.sched_setaffinity(0, pinned_cores)
        except (AttributeError, OSError):
            pinned_cores = None
    case = workload_case(args.workload)
    ny, nx = case.shape
    if args.workload == "batch256":
        seeds = batch.shard(list(range(1000, 1000 + args.npairs)), rank, world)
    else:
        seeds = [1000 + rank * 2, 1000 + rank * 2 + 1]
    t_gen = time.perf_counter()
    made = make_pairs(args.workload, seeds, len(pinned_cores) if pinned_cores else box_cores)      # before CUDA starts (fork)
    t_gen = time.perf_counter() - t_gen

    import torch
    import torch.distributed as dist
    from zuds_pipeline_b200 import _lib
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the subtraction path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    npairs_distinct = len(seeds)
    params, dev_in, pin_in = [], [], []
    for ins, p in made:
        params.append(p)
        pinned = [None if a is None else torch.from_numpy(np.ascontiguousarray(a.view(np.uint8) if a.dtype == np.bool_ else a)).pin_memory()
                  for a in ins]
        pin_in.append(pinned)
        dev_in.append([None if t is None else t.cuda(non_blocking=True) for t in pinned])
    torch.cuda.synchronize()
    del made

    # contexts per GPU: each is a host thread that polls its stream and yields the core between polls
    # (zb_stream_wait, ZB_SYNC=yield), so the count does not depend on the cores: 24-32 is where the rate stops growing
    W = args.workers or 32
    B = npairs_distinct if args.workload == "batch256" else (args.batch or 4 * W)
    from zuds_pipeline_b200.batch import GpuPool
    pool = GpuPool(local, W)
    dev_out = [(torch.empty((ny, nx), dtype=torch.float32, device="cuda"), torch.empty((ny, nx), dtype=torch.float32, device="cuda"),
                torch.empty((ny, nx), dtype=torch.int32, device="cuda")) for _ in range(W)]
    # two pinned output slots per context: zb_subtract_batch sends pair n's products home while pair n+1 is computed
    pin_out = [[(torch.empty((ny, nx), dtype=torch.float32).pin_memory(), torch.empty((ny, nx), dtype=torch.float32).pin_memory())
                for _ in range(2)] for _ in range(W)]
    slot_of = {id(c): i for i, c in enumerate(pool.ctxs)}

    def dev_job(k):
        def fn(ctx):
            return ctx.subtract(params[k], *dev_in[k], out=dev_out[slot_of[id(ctx)]], stream=_lib.OwnStream())
        return fn

    def host_batch_job(ks):
        def fn(ctx):
            o = pin_out[slot_of[id(ctx)]]
            ins = [[None if t is None else t.numpy() for t in pin_in[k]] for k in ks]
            # zuds' command line has no -omi (zuds/hotpants.py:77-93): the int mask stays on the device
            outs = [(o[j & 1][0].numpy(), o[j & 1][1].numpy(), None) for j in range(len(ks))]
            return ctx.subtract_batch([params[k] for k in ks], ins, outs=outs)
        return fn

    def step(make_job):
        res, tim = pool.run([make_job(i % npairs_distinct) for i in range(B)])
        return res, [t.get("k_convolve", 0.0) for t in tim]

    def host_step():
        # one zb_subtract_batch per context and step: its share of the step's pairs, pipelined inside the context
        jobs = [list(range(B))[w::W] for w in range(W)]
        pool.run([host_batch_job([i % npairs_distinct for i in js]) for js in jobs if js])

    # ---- device-resident throughput ("value")
    for _ in range(args.warmup):
        step(dev_job)
    barrier()
    clocks = ClockSampler(local)
    if not args.no_clocks:
        clocks.start()
    l0 = pool.launches()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    conv_ms = []
    t0 = time.perf_counter()
    step_ms = []
    for _ in range(args.steps):
        ts = time.perf_counter()
        _, ms = step(dev_job)
        step_ms.append(1e3 * (time.perf_counter() - ts))
        conv_ms += ms
    e1.record()
    if args.verbose and rank == 0:
        sys.stderr.write("device-resident step ms: %s\n" % " ".join("%.1f" % t for t in step_ms))
    barrier()
    wall = time.perf_counter() - t0
    dev_s = e0.elapsed_time(e1) / 1e3
    launches = pool.launches() - l0
    clk = clocks.stop()
    # ---- end to end through the host-buffer call ("e2e")
    for _ in range(max(1, args.warmup // 2)):
        host_step()
    barrier()
    h0, h1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    h0.record()
    for _ in range(args.steps):
        host_step()
    h1.record()
    barrier()
    host_s = h0.elapsed_time(h1) / 1e3
    # ---- the host link, measured alone: what bounds e2e when the copies cannot hide behind compute
    link = {}
    try:
        hb = torch.empty(256 << 20, dtype=torch.uint8).pin_memory()
        db = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
        for name, src, dst in (("h2d_gbs", hb, db), ("d2h_gbs", db, hb)):
            best = 0.0
            for _ in range(4):
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                dst.copy_(src, non_blocking=True)
                b.record()
                b.synchronize()
                best = max(best, hb.numel() / (a.elapsed_time(b) * 1e-3) / 1e9)
            link[name] = round(best, 2)
        # every rank at once, both directions at once, in the proportions of the e2e leg (one pair's inputs up, one
        # pair's products down, nothing computed): what the host side (DRAM, PCIe roots) gives N GPUs together
        barrier()
        in_b = min(hb.numel(), sum(t.numel() * t.element_size() for t in pin_in[0] if t is not None))
        out_b = sum(t.numel() * t.element_size() for t in pin_out[0][0])
        sa, sb = torch.cuda.Stream(), torch.cuda.Stream()
        hb2 = torch.empty(out_b, dtype=torch.uint8).pin_memory()
        db2 = torch.empty(out_b, dtype=torch.uint8, device="cuda")
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record()
        sa.wait_event(a)
        sb.wait_event(a)
        nrep = 8
        for _ in range(nrep):
            with torch.cuda.stream(sa):
                db[:in_b].copy_(hb[:in_b], non_blocking=True)
            with torch.cuda.stream(sb):
                hb2.copy_(db2, non_blocking=True)
        torch.cuda.current_stream().wait_stream(sa)
        torch.cuda.current_stream().wait_stream(sb)
        b.record()
        b.synchronize()
        link["copy_only_replay_pairs_per_s_per_gpu"] = round(nrep / (a.elapsed_time(b) * 1e-3), 1)
        link["copy_only_replay_gbs_both_ways_per_gpu"] = round(nrep * (in_b + out_b) / (a.elapsed_time(b) * 1e-3) / 1e9, 2)
        del hb, db, hb2, db2
    except RuntimeError as e:          # pinned allocation refused: report, do not fail the bench
        link = {"error": str(e)[:120]}
    # ---- the convolution kernel alone, one context, nothing else on the GPU (roofline)
    serial_ms = []
    for i in range(6):
        pool.ctxs[0].subtract(params[i % npairs_distinct], *dev_in[i % npairs_distinct], out=dev_out[0], stream=_lib.OwnStream())
        if i >= 2:
            serial_ms.append(pool.ctxs[0].timings())
    stage_ms = {k: statistics.mean(t[k] for t in serial_ms) for k in serial_ms[0]}

    tmax = torch.tensor([dev_s, host_s], dtype=torch.float64, device="cuda")
    nper = torch.tensor([float(B)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(nper, op=dist.ReduceOp.SUM)
    dev_s, host_s = float(tmax[0]), float(tmax[1])
    pairs_total = args.steps * int(round(float(nper[0])))
    in_bytes = sum(t.numel() * t.element_size() for t in pin_in[0] if t is not None)
    out_bytes = sum(t.numel() * t.element_size() for t in pin_out[0][0])
    hbm_peak, peak_src = peaks()
    alg_bytes = ALG_BYTES_PER_PIXEL * nx * ny
    conv_s = stage_ms["k_convolve"] / 1e3
    achieved = alg_bytes / conv_s / 1e9
    fw = 2 * params[0].hw_kernel + 1
    fma = 2 * fw * fw * nx * ny                       # one multiply-add per tap for the image and one for the variance
    traffic, traffic_src = ncu_traffic(args.workload)
    line = {
        "metric": METRIC, "value": pairs_total / dev_s, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * dev_s / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64 (image convolution and fit, float x double products summed in double as the reference does); the variance sum of k_convolve in f32", "data": "synthetic",
        "config": {"workload": WORKLOADS[args.workload][1]},
        "run": {"pairs_per_step_per_gpu": B, "contexts_per_gpu": W, "host_cores": box_cores,
                "rank_cores": None if pinned_cores is None else len(pinned_cores), "distinct_pairs_per_gpu": npairs_distinct,
                "l2": "inputs of one pair are %.0f MB (> 126 MB L2) and consecutive jobs take different pairs" % (in_bytes / 1e6),
                "sharding": "by image pair, no collective", "synth_s": round(t_gen, 1)},
        "e2e": {"value": pairs_total / host_s, "unit": UNIT, "h2d_bytes_per_step": in_bytes * B, "d2h_bytes_per_step": out_bytes * B,
                "call": "zb_subtract_batch (one call per context and step: H2D of pair n+1 and D2H of pair n-1 overlap pair n), out_mask NULL as in the zuds command line (no -omi)",
                "host_link": link,
                "link_bound_pairs_per_s": (round(world * min(link["h2d_gbs"] * 1e9 / in_bytes, link["d2h_gbs"] * 1e9 / out_bytes), 1)
                                           if "h2d_gbs" in link else None),
                "copy_only_replay_pairs_per_s": (round(world * link["copy_only_replay_pairs_per_s_per_gpu"], 1)
                                                 if "copy_only_replay_pairs_per_s_per_gpu" in link else None)},
        "gpu_launches": launches,
        "clocks": clk,
        "roofline": {"kernel": "k_convolve", "bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                     "frac": achieved / hbm_peak, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                     "alg_bytes_per_launch": alg_bytes, "kernel_ms": stage_ms["k_convolve"],
                     "kernel_ms_how": "CUDA events around the launch on the context's stream, mean of 4 isolated serial subtractions after the timed region",
                     "kernel_ms_in_timed_region": statistics.mean(conv_ms) if conv_ms else None,
                     # what really limits the kernel at hw >= 5: the FP32 pipe (2 (2hw+1)^2 FMA per pixel, packed as
                     # fma.rn.f32x2); peak measured on this pool's B200 with tools/fp32_peak.cu
                     "limited_by": "fp32 pipe" if fw >= 11 else "hbm",
                     "fp32": {"fma_per_launch": fma, "achieved": fma / conv_s / 1e12, "peak": FP32_PEAK_TFMA, "unit": "T FMA/s",
                              "frac": fma / conv_s / 1e12 / FP32_PEAK_TFMA,
                              "peak_source": "measured (tools/fp32_peak.cu, profiles/r02_fp32_peak_b200.txt)"},
                     "alg_flops_conv": 2 * fma, "frac_of_fp32_peak_74.4T": 2 * fma / conv_s / 74.4e12},
        "stage_ms_serial": {k: round(v, 4) for k, v in stage_ms.items()},
        "wall_s": wall,
    }
    if rank == 0 and world == 1 and not args.no_cpu:
        c = _case_with_seed(case, seeds[0])
        pair0 = CS.make_pair(c)
        with tempfile.TemporaryDirectory(prefix="zb_cpu_") as d:
            names = write_pair_fits(pair0, d)
            dt, kind = run_cpu_once(c, pair0, names, d, 1)
        line["cpu_baseline"] = {"value": 1.0 / dt, "unit": UNIT, "cores": 1, "kind": kind,
                                "sample": "1 full %dx%d pair, one single-threaded hotpants process incl. FITS I/O (%.1f s)" % (ny, nx, dt)}
        if not args.no_api:
            try:
                line["e2e_api"] = api_leg(c, pair0, local, args.api_pairs, args.api_threads)
            except Exception as e:                     # the API leg is informative: never lose the line over it
                line["e2e_api"] = {"error": "%s: %s" % (type(e).__name__, str(e)[:300])}
    pool.close()
    if rank == 0:
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


# --------------------------------------------------------------------------- GPU side: row tiles (configs[4])

def bench_c5(args):
    """One coadd-scale frame, one global fit, split into row tiles over the ranks (zuds_pipeline_b200/tiles.py).
    Strong scaling: the work is one frame whatever N is.  Every rank holds only its own rows; `value` starts with
    them resident in HBM, `e2e` uploads them from pinned host memory and brings the products back to rank 0's host."""
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
    case = workload_case("c5")
    pair = CS.make_pair(case)
    p = CS.params_of(case, pair)
    ins = [None if a is None else np.ascontiguousarray(a.view(np.uint8) if a.dtype == np.bool_ else a) for a in CS.inputs_of(case, pair)]
    ny, nx = pair.sci.shape
    ctx = _lib.Context(local)
    st = torch.cuda.Stream(device=dev)
    infos = tiles.plan_all(p, world)
    me = infos[rank]
    pin_rows = [None if a is None else torch.from_numpy(a[me.row0:me.row1]).pin_memory() for a in ins]
    pin_out = None
    if rank == 0:
        pin_out = (torch.empty((ny, nx), dtype=torch.float32).pin_memory(), torch.empty((ny, nx), dtype=torch.float32).pin_memory())

    def fresh(upload):
        """full-frame device tensors holding ONLY this rank's rows (the rest poisoned = not received yet)"""
        out = []
        for a, pr in zip(ins, pin_rows):
            if a is None:
                out.append(None)
                continue
            t = torch.empty((ny, nx), dtype=pr.dtype, device=dev)
            t.fill_(float("nan") if t.dtype == torch.float32 else 77)
            if upload:
                t[me.row0:me.row1].copy_(pr, non_blocking=True)
            out.append(t)
        return out

    clocks = ClockSampler(local)
    clk = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
    with torch.cuda.stream(st):
        comm = tiles.DistComm() if world > 1 else None
        stats, res = {}, None
        times = {"value": [], "e2e": []}
        launches = 0
        for mode in ("value", "e2e"):
            for it in range(args.warmup + args.steps):
                mine = fresh(mode == "value")              # value: resident before the clock starts
                st.synchronize()
                if world > 1:
                    dist.barrier()
                if mode == "value" and it == args.warmup and not args.no_clocks:
                    clocks.start()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                l0 = ctx.launch_count()
                e0.record(st)
                if mode == "e2e":
                    for t, pr in zip(mine, pin_rows):
                        if t is not None:
                            t[me.row0:me.row1].copy_(pr, non_blocking=True)
                if world > 1:
                    res = tiles.subtract_tiled(ctx, comm, p, *mine, stats=stats)
                else:
                    res = ctx.subtract(p, *mine, stream=st)
                if mode == "e2e" and rank == 0:
                    pin_out[0].copy_(res.diff, non_blocking=True)
                    pin_out[1].copy_(res.noise, non_blocking=True)
                e1.record(st)
                st.synchronize()
                ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
                if world > 1:
                    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
                if it >= args.warmup:
                    times[mode].append(float(ms.item()))
                    if mode == "value":
                        launches += ctx.launch_count() - l0
            if mode == "value" and not args.no_clocks:
                clk = clocks.stop()
        ok = None
        if args.check and rank == 0:
            full = [None if a is None else torch.from_numpy(a).to(dev) for a in ins]
            ctx2 = _lib.Context(local)
            g = ctx2.subtract(p, *full, stream=st)
            st.synchronize()
            ok = bool(torch.equal(g.mask, res.mask) and torch.equal(g.diff.view(torch.int32), res.diff.view(torch.int32))
                      and torch.equal(g.noise.view(torch.int32), res.noise.view(torch.int32)) and np.array_equal(g.sol, res.sol))
            ctx2.close()
    if rank == 0:
        tv, te = sum(times["value"]) / 1e3, sum(times["e2e"]) / 1e3
        in_bytes = sum(a.nbytes for a in ins if a is not None)
        line = {"metric": "coadd-scale frame subtractions/sec (one global fit)", "value": args.steps / tv, "unit": "frames/s",
                "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tv / args.steps,
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "f64 (image convolution and fit, float x double products summed in double as the reference does); the variance sum of k_convolve in f32",
                "data": "synthetic", "config": {"workload": WORKLOADS["c5"][1]},
                "run": {"tiles": [i.as_dict() for i in infos] if world > 1 else None,
                        "collectives": "P2P halo rows (inputs, then search-claimed mask rows), all-reduce(SUM) of search records / per-stamp Gram+rhs / per-stamp sigma, gather of output rows on rank 0",
                        "n_stamps": int(res.regions[0].n_stamps), "fit_iters": int(res.regions[0].fit_iters),
                        "tile_stats_rank0": {k: v for k, v in stats.items() if k != "info"},
                        "l2": "one frame's inputs are %.0f MB (> 126 MB L2)" % (in_bytes / 1e6)},
                "e2e": {"value": args.steps / te, "unit": "frames/s", "h2d_bytes_per_step": in_bytes, "d2h_bytes_per_step": 2 * nx * ny * 4,
                        "call": "own rows from pinned host memory on every rank, products to rank 0's host"},
                "gpu_launches": launches, "clocks": clk, "bit_equal_to_single_gpu": ok,
                "ms_all": [round(t, 3) for t in times["value"]]}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c1", choices=sorted(WORKLOADS))
    ap.add_argument("--workers", type=int, default=0, help="concurrent contexts (streams) per GPU (0: 32)")
    ap.add_argument("--batch", type=int, default=0, help="pair subtractions per step per GPU (0: 4 per context)")
    ap.add_argument("--npairs", type=int, default=256, help="batch256: distinct pairs over all ranks")
    ap.add_argument("--cpu-procs", type=int, default=0, help="reference arm: concurrent hotpants processes (0 = all cores)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline and e2e_api legs")
    ap.add_argument("--no-api", action="store_true", help="skip the e2e_api leg")
    ap.add_argument("--api-pairs", type=int, default=16)
    ap.add_argument("--api-threads", type=int, default=4)
    ap.add_argument("--check", action="store_true", help="c5: compare the tiled products with one GPU bit for bit")
    ap.add_argument("--no-pin", action="store_true", help="(N > 1) do not bind each rank to its share of the cores")
    ap.add_argument("--no-clocks", action="store_true", help="(diagnostics) do not sample clocks during the timed region")
    ap.add_argument("--verbose", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        bench_reference(args)
    elif args.workload == "c5":
        bench_c5(args)
    else:
        bench_pairs(args)


if __name__ == "__main__":
    main()
