#!/usr/bin/env python
"""bench.py — ZTF quadrant subtractions per second on N B200s (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K ...  # the reference hotpants on the host cores

Workload (config.workload "c1"): BASELINE.json configs[1] — a 3080x3072 float32 synthetic quadrant
pair with rms maps and the boolean bad-pixel mask, hotpants flags as zuds builds them
(-c t -n i, 3 Gaussians, -r 10 -rss 24 -ko 2 -bgo 0, 3x3 regions of 10x10 stamps).
A "step" is one batch of `--batch` pair subtractions per GPU.  Pairs shard by image across ranks
with no collective: scaling is "weak".

  value     pairs/s with the inputs already resident in HBM (zb_subtract on device pointers)
  e2e       pairs/s through the host-buffer C-ABI call (zb_subtract_host): every subtraction copies
            its five input arrays from pinned host memory and its three outputs back
  roofline  the convolution/output kernel (k_convolve), ALGORITHMIC bytes 29 B/pixel
            over its CUDA-event duration, against MEASURED_PEAKS.json's HBM copy bandwidth
  cpu_baseline  the unmodified reference binary oracle/_ref/hotpants (or the oracle port) on one
            full pair on one host core
"""
from __future__ import annotations

import argparse
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
SHAPE = (3080, 3072)
ALG_BYTES_PER_PIXEL = 29          # 4 float inputs + 1 mask byte in, 2 floats + int32 mask out
# dram__bytes_read.sum + dram__bytes_write.sum of one k_convolve launch at C1, from the ncu --set full
# capture summarised in profiles/r01_ncu_k_convolve_s17.txt (251.88 MB + 188.61 MB)
FP64_PEAK_TDFMA = 16.77
NCU_CONV_TRAFFIC_BYTES = 440489472


def workload_case():
    import cases as CS
    return CS.FULL["c1"]


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
    """--impl reference: the reference's own CPU implementation on all host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import cases as CS
    case = workload_case()
    pair = CS.make_pair(case)
    import hp_helpers as H
    nproc = args.cpu_procs or host_cores()
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
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "c1: 3080x3072 quadrant pair, -c t -n i -r 10 -rss 24 -ko 2 -bgo 0, 3x3 regions x 10x10 stamps",
                   "pairs_per_step": nproc},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": nproc, "kind": kind,
                         "sample": "%d concurrent single-threaded hotpants processes per step, one full pair each (FITS in/out on local disk)" % nproc},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


# --------------------------------------------------------------------------- GPU side

def bench_b200(args):
    import torch
    import torch.distributed as dist
    import cases as CS
    from zuds_pipeline_b200 import _lib

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the subtraction path has no CPU fallback")
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
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    case = workload_case()
    ny, nx = case.shape
    npairs_distinct = 2
    pairs, params, dev_in, pin_in = [], [], [], []
    for k in range(npairs_distinct):
        c = CS.Case(case.name, case.shape, case.seeing, case.nreg_side, case.hotpants_kws, case.synth_kws,
                    seed=1000 + rank * npairs_distinct + k)
        pair = CS.make_pair(c)
        pairs.append(pair)
        params.append(CS.params_of(c, pair))
        ins = CS.inputs_of(c, pair)
        pinned = [None if a is None else torch.from_numpy(np.ascontiguousarray(a)).pin_memory() for a in ins]
        pin_in.append(pinned)
        dev_in.append([None if t is None else t.cuda(non_blocking=True) for t in pinned])
    torch.cuda.synchronize()

    # contexts per GPU: each is a host thread that polls its stream and yields the core between polls
    # (zb_stream_wait, ZB_SYNC=yield), so the count does not depend on the cores: 24-32 is where the rate stops
    # growing, with 16 cores or with 4 (taskset -c 0-3: 225.9 / 193.7 pairs/s; busy-waiting needed a core per
    # thread and gave 196 / 151 with 4 contexts on 4 cores)
    cores = box_cores
    W = args.workers or 32
    B = args.batch or 4 * W
    from zuds_pipeline_b200.batch import GpuPool
    pool = GpuPool(local, W)
    dev_out = [(torch.empty((ny, nx), dtype=torch.float32, device="cuda"), torch.empty((ny, nx), dtype=torch.float32, device="cuda"),
                torch.empty((ny, nx), dtype=torch.int32, device="cuda")) for _ in range(W)]
    pin_out = [(torch.empty((ny, nx), dtype=torch.float32).pin_memory(), torch.empty((ny, nx), dtype=torch.float32).pin_memory(),
                torch.empty((ny, nx), dtype=torch.int32).pin_memory()) for _ in range(W)]
    slot_of = {id(c): i for i, c in enumerate(pool.ctxs)}

    def dev_job(k):
        def fn(ctx):
            return ctx.subtract(params[k], *dev_in[k], out=dev_out[slot_of[id(ctx)]], stream=_NullStream())
        return fn

    def host_job(k):
        def fn(ctx):
            o = pin_out[slot_of[id(ctx)]]
            ins = [None if t is None else t.numpy() for t in pin_in[k]]
            return ctx.subtract_host(params[k], *ins, out=(o[0].numpy(), o[1].numpy(), o[2].numpy()))
        return fn

    def step(make_job):
        res, tim = pool.run([make_job(i % npairs_distinct) for i in range(B)])
        return res, [t.get("k_convolve", 0.0) for t in tim]

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
        step(host_job)
    barrier()
    h0, h1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    h0.record()
    for _ in range(args.steps):
        step(host_job)
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
        del hb, db
    except RuntimeError as e:          # pinned allocation refused: report, do not fail the bench
        link = {"error": str(e)[:120]}
    # ---- the convolution kernel alone, one context, nothing else on the GPU (roofline)
    serial_ms = []
    for i in range(6):
        pool.ctxs[0].subtract(params[i % npairs_distinct], *dev_in[i % npairs_distinct], out=dev_out[0], stream=_NullStream())
        if i >= 2:
            serial_ms.append(pool.ctxs[0].timings())
    stage_ms = {k: statistics.mean(t[k] for t in serial_ms) for k in serial_ms[0]}

    tmax = torch.tensor([dev_s, host_s], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    dev_s, host_s = float(tmax[0]), float(tmax[1])
    pairs_total = world * args.steps * B
    in_bytes = sum(t.numel() * t.element_size() for t in pin_in[0] if t is not None)
    out_bytes = sum(t.numel() * t.element_size() for t in pin_out[0])
    hbm_peak, peak_src = peaks()
    alg_bytes = ALG_BYTES_PER_PIXEL * nx * ny
    conv_s = stage_ms["k_convolve"] / 1e3
    achieved = alg_bytes / conv_s / 1e9
    line = {
        "metric": METRIC, "value": pairs_total / dev_s, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * dev_s / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "c1: 3080x3072 quadrant pair, -c t -n i -r 10 -rss 24 -ko 2 -bgo 0, 3x3 regions x 10x10 stamps (BASELINE configs[1])",
                   "pairs_per_step_per_gpu": B, "contexts_per_gpu": W, "host_cores": cores, "rank_cores": None if pinned_cores is None else len(pinned_cores), "distinct_pairs_per_gpu": npairs_distinct,
                   "l2": "inputs of one pair are 161 MB (> 126 MB L2) and consecutive jobs alternate pairs",
                   "sharding": "by image pair, no collective"},
        "e2e": {"value": pairs_total / host_s, "unit": UNIT, "h2d_bytes_per_step": in_bytes * B, "d2h_bytes_per_step": out_bytes * B,
                "host_link": link,
                "link_bound_pairs_per_s": (round(min(link["h2d_gbs"] * 1e9 / in_bytes, link["d2h_gbs"] * 1e9 / out_bytes), 1)
                                           if "h2d_gbs" in link else None)},
        "gpu_launches": launches,
        "clocks": clk,
        "roofline": {"kernel": "k_convolve", "bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                     "frac": achieved / hbm_peak, "traffic": NCU_CONV_TRAFFIC_BYTES, "peak_source": peak_src,
                     "alg_bytes_per_launch": alg_bytes, "kernel_ms": stage_ms["k_convolve"],
                     "kernel_ms_in_timed_region": statistics.mean(conv_ms) if conv_ms else None,
                     "fp64_dfma_per_launch": 2 * (2 * params[0].hw_kernel + 1) ** 2 * nx * ny,
                     # the pipe that actually bounds this kernel: FP64 FMA, peak measured on this pool's B200 with
                     # tools/fp64_peak.cu (profiles/r01_fp64_peak_b200.txt: 57.7 DFMA/clk/SM = 16.77 T DFMA/s at 1965 MHz)
                     "fp64": {"achieved": 2 * (2 * params[0].hw_kernel + 1) ** 2 * nx * ny / (stage_ms["k_convolve"] * 1e-3) / 1e12,
                              "peak": FP64_PEAK_TDFMA, "unit": "T DFMA/s",
                              "frac": 2 * (2 * params[0].hw_kernel + 1) ** 2 * nx * ny / (stage_ms["k_convolve"] * 1e-3) / 1e12 / FP64_PEAK_TDFMA,
                              "peak_source": "measured (tools/fp64_peak.cu)"},
                     # SURVEY.md 8(d)'s figure: ALG_FLOPS_CONV = 4 nx ny (2hw+1)^2 against the 74.4 TFLOP/s FP32 CUDA-core
                     # peak; the products are float x double accumulated in double (alard.c:1289-1306), so the FP64 pipe above is the bound
                     "alg_flops_conv": 4 * (2 * params[0].hw_kernel + 1) ** 2 * nx * ny,
                     "frac_of_fp32_peak_74.4T": 4 * (2 * params[0].hw_kernel + 1) ** 2 * nx * ny / (stage_ms["k_convolve"] * 1e-3) / 74.4e12,
                     "note": "compute-bound at hw=10 (2x441 FP64 FMA per pixel, reference semantics); HBM fraction reported as the north star asks"},
        "stage_ms_serial": {k: round(v, 4) for k, v in stage_ms.items()},
        "wall_s": wall,
    }
    if rank == 0 and world == 1 and not args.no_cpu:
        with tempfile.TemporaryDirectory(prefix="zb_cpu_") as d:
            import hp_helpers as H
            names = write_pair_fits(pairs[0], d)
            dt, kind = run_cpu_once(case, pairs[0], names, d, 1)
        line["cpu_baseline"] = {"value": 1.0 / dt, "unit": UNIT, "cores": 1, "kind": kind,
                                "sample": "1 full 3080x3072 pair, one single-threaded hotpants process incl. FITS I/O (%.1f s)" % dt}
    pool.close()
    if rank == 0:
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


class _NullStream:
    """ZB_STREAM_OWN (include/zb200.h): the ctx runs on its own non-blocking stream (= _lib.OwnStream)."""
    cuda_stream = -1


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workers", type=int, default=0, help="concurrent contexts (streams) per GPU (0: 32)")
    ap.add_argument("--batch", type=int, default=0, help="pair subtractions per step per GPU (0: 4 per context)")
    ap.add_argument("--cpu-procs", type=int, default=0, help="reference arm: concurrent hotpants processes (0 = all cores)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-pin", action="store_true", help="(N > 1) do not bind each rank to its share of the cores")
    ap.add_argument("--no-clocks", action="store_true", help="(diagnostics) do not sample clocks during the timed region")
    ap.add_argument("--verbose", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        bench_reference(args)
    else:
        bench_b200(args)


if __name__ == "__main__":
    main()
