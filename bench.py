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
import queue
import statistics
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "ZTF quadrant subtractions/sec"
UNIT = "quadrant-pairs/s"
SHAPE = (3080, 3072)
ALG_BYTES_PER_PIXEL = 29          # 4 float inputs + 1 mask byte in, 2 floats + int32 mask out


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


# --------------------------------------------------------------------------- worker pool

class Pool:
    """`n` host threads, each owning one zb_ctx (one stream + workspace) on the same GPU.  ctypes
    releases the GIL inside the C call, so the contexts overlap on the device."""

    def __init__(self, device, n):
        from zuds_pipeline_b200 import _lib
        self.ctxs = [_lib.Context(device) for _ in range(n)]
        self.q = queue.Queue()
        self.done = queue.Queue()
        self.threads = [threading.Thread(target=self._run, args=(c,), daemon=True) for c in self.ctxs]
        for t in self.threads:
            t.start()

    def _run(self, ctx):
        while True:
            job = self.q.get()
            if job is None:
                return
            try:
                self.done.put((job[0], job[1](ctx), ctx.timings().get("k_convolve", 0.0)))
            except Exception as e:       # surface worker failures in the main thread
                self.done.put((job[0], e, 0.0))

    def run_batch(self, fns):
        for i, fn in enumerate(fns):
            self.q.put((i, fn))
        out, conv_ms = [None] * len(fns), []
        for _ in fns:
            i, r, ms = self.done.get()
            if isinstance(r, Exception):
                raise r
            out[i] = r
            conv_ms.append(ms)
        return out, conv_ms

    def launches(self):
        return sum(c.launch_count() for c in self.ctxs)

    def close(self):
        for _ in self.threads:
            self.q.put(None)
        for t in self.threads:
            t.join()
        for c in self.ctxs:
            c.close()
