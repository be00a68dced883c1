"""Batch execution on one GPU: several zb_ctx (stream + workspace) driven by host threads.

The reference gets its throughput from N independent single-threaded `hotpants` processes, one per
image pair (`srun -n 64 -c1`, nersc/controller.py:101; zuds/mpi.py:36-64).  Here one process owns
one GPU (one rank per GPU, pairs sharded by image, no collective) and keeps it busy by running
`contexts` subtractions concurrently: the fit stage of a pair is a chain of small kernels with
host-side decisions between them, and it overlaps with the bandwidth/compute-heavy stages of the
other pairs.  ctypes releases the GIL inside the C-ABI call, so plain threads are enough.
"""
from __future__ import annotations

import queue
import threading

from . import _lib

__all__ = ["GpuPool", "shard"]


def shard(items, rank, world):
    """The slice of ``items`` rank ``rank`` of ``world`` owns: contiguous, sizes differing by at most
    one, exactly ``numpy.array_split`` as zuds/mpi.py:60 uses it."""
    n = len(items)
    base, extra = divmod(n, world)
    start = rank * base + min(rank, extra)
    return items[start:start + base + (1 if rank < extra else 0)]


class GpuPool:
    """``contexts`` worker threads on CUDA device ``device``; jobs are callables ``fn(ctx) -> result``."""

    def __init__(self, device=0, contexts=4):
        self.device = device
        self.ctxs = [_lib.Context(device) for _ in range(contexts)]
        self._q = queue.Queue()
        self._done = queue.Queue()
        self._threads = [threading.Thread(target=self._run, args=(c,), daemon=True) for c in self.ctxs]
        for t in self._threads:
            t.start()

    def slot(self, ctx):
        return self.ctxs.index(ctx)

    def _run(self, ctx):
        while True:
            job = self._q.get()
            if job is None:
                return
            i, fn = job
            try:
                self._done.put((i, fn(ctx), ctx.timings()))
            except BaseException as e:            # handed to the caller of run()
                self._done.put((i, e, {}))

    def run(self, fns, return_exceptions=False):
        """Run the jobs; returns (results in job order, per-job stage timings)."""
        fns = list(fns)
        for i, fn in enumerate(fns):
            self._q.put((i, fn))
        out, tim = [None] * len(fns), [None] * len(fns)
        first_error = None
        for _ in fns:
            i, r, t = self._done.get()
            if isinstance(r, BaseException) and not return_exceptions and first_error is None:
                first_error = r
            out[i], tim[i] = r, t
        if first_error is not None:
            raise first_error
        return out, tim

    def subtract_host_many(self, jobs, return_exceptions=False):
        """``jobs``: iterable of (params, sci, ref, sci_rms, ref_rms, sci_mask, ref_mask) host arrays."""
        def mk(j):
            return lambda ctx: ctx.subtract_host(*j)
        return self.run([mk(j) for j in jobs], return_exceptions=return_exceptions)[0]

    def launches(self):
        return sum(c.launch_count() for c in self.ctxs)

    def close(self):
        for _ in self._threads:
            self._q.put(None)
        for t in self._threads:
            t.join()
        for c in self.ctxs:
            c.close()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
