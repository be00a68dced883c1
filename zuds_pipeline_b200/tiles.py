"""Row-tile subtraction of ONE oversized frame over several GPUs (BASELINE.json configs[4]).

The reference has no counterpart: hotpants is one single-threaded process per frame and zuds only
shards whole images over MPI ranks (zuds/mpi.py:36-64).  A coadd-scale frame fitted by a single
region (``-nrx 1 -nry 1``) cannot be sharded that way, so here the frame is cut into row tiles, one
rank (process + GPU) per tile, over the ``zb_tile_*`` entry points of include/zb200.h:

* halo rows of the inputs travel peer to peer (``Comm.sendrecv``: NCCL send/recv over NVLink);
* stamp-search records, per-stamp normal equations (Gram matrix + right-hand side) and per-stamp
  figures of merit are merged with one ``all_reduce(SUM)`` each — every entry is non-zero on exactly
  one rank, so the sum is exact and the global fit, built on every rank in the reference's stamp
  order (alard.c:843-1025), is bit-identical to the single-GPU ``zb_subtract``;
* mask rows changed by the stamp search travel with a second halo exchange;
* output rows are gathered on rank 0, which runs the frame-wide QA statistics (main.c:1417-1554).

``DistComm`` is the production communicator (``torch.distributed``, NCCL); ``LocalComm`` runs the
ranks as threads of one process on one GPU and exists for the single-GPU parity tests.
"""
from __future__ import annotations

import ctypes as C
import threading

import numpy as np

from . import _lib
from . import params as P

__all__ = ["ZbTileInfo", "plan", "plan_all", "row_transfers", "exchange_rows", "DistComm", "LocalComm",
           "subtract_tiled", "run_local"]

TILE_MASK_IN, TILE_RDIFF, TILE_RNOISE, TILE_OMASK = 0, 1, 2, 3


class ZbTileInfo(C.Structure):
    """include/zb200.h: zb_tile_info"""
    _fields_ = [(n, C.c_int32) for n in (
        "rank", "world", "cell_row0", "cell_row1", "row0", "row1", "in_row0", "in_row1",
        "blk_row0", "blk_row1", "out_row0", "out_row1", "halo", "n_cells", "cell_words",
        "n_slots", "gram_len", "rhs_len")]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


def _bind(lib):
    if getattr(lib, "_tiles_bound", False):
        return lib
    vp, ip = C.c_void_p, C.POINTER(C.c_int)
    lib.zb_tile_plan.argtypes = [C.POINTER(P.ZbParams), C.c_int, C.c_int, C.POINTER(ZbTileInfo)]
    lib.zb_tile_begin.argtypes = [vp, C.POINTER(P.ZbParams), C.c_int, C.c_int, vp, vp, vp, vp, vp, C.c_int, vp, C.c_int,
                                  vp, vp, vp, vp]
    lib.zb_tile_begin_retry.argtypes = [vp]
    lib.zb_tile_cells_export.argtypes = [vp, vp]
    lib.zb_tile_cells_import.argtypes = [vp, vp]
    lib.zb_tile_resolve.argtypes = [vp, ip, ip]
    lib.zb_tile_research.argtypes = [vp, C.c_int]
    lib.zb_tile_rows.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, vp]
    lib.zb_tile_fill.argtypes = [vp]
    lib.zb_tile_normal_export.argtypes = [vp, vp, vp]
    lib.zb_tile_normal_import.argtypes = [vp, vp, vp]
    lib.zb_tile_solve.argtypes = [vp]
    lib.zb_tile_sig_export.argtypes = [vp, vp]
    lib.zb_tile_sig_import.argtypes = [vp, vp]
    lib.zb_tile_clip.argtypes = [vp, ip]
    lib.zb_tile_convolve.argtypes = [vp]
    lib.zb_tile_finish.argtypes = [vp, C.POINTER(P.ZbRegionInfo), vp, C.POINTER(P.ZbStampInfo)]
    lib._tiles_bound = True
    return lib


def plan(p, world, rank):
    """Who owns what (pure host arithmetic, zb_tile_plan; needs no GPU)."""
    info = ZbTileInfo()
    rc = _bind(_lib.load()).zb_tile_plan(C.byref(p), int(world), int(rank), C.byref(info))
    if rc != 0:
        raise _lib.ZbError(rc, "zb_tile_plan(world=%d, rank=%d): row tiles need nreg_x = nreg_y = 1" % (world, rank))
    return info


def plan_all(p, world):
    return [plan(p, world, r) for r in range(world)]


def row_transfers(infos, rank, have=("row0", "row1"), need=("in_row0", "in_row1")):
    """Rows to move so that every rank holds its ``need`` range, given that rank q holds its ``have``
    range.  Returns (sends, recvs): lists of (peer, r0, r1), in a fixed global order on both sides."""
    def cut(a0, a1, b0, b1):
        lo, hi = max(a0, b0), min(a1, b1)
        return (lo, hi) if hi > lo else None
    me = infos[rank]
    sends, recvs = [], []
    for q, other in enumerate(infos):
        if q == rank:
            continue
        c = cut(getattr(me, have[0]), getattr(me, have[1]), getattr(other, need[0]), getattr(other, need[1]))
        if c:
            sends.append((q, c[0], c[1]))
        c = cut(getattr(other, have[0]), getattr(other, have[1]), getattr(me, need[0]), getattr(me, need[1]))
        if c:
            recvs.append((q, c[0], c[1]))
    return sends, recvs


def _bytes_view(t):
    import torch
    return t.view(torch.uint8) if t.dtype != torch.uint8 else t


def exchange_rows(tensors, infos, rank, comm, have=("row0", "row1"), need=("in_row0", "in_row1")):
    """Halo exchange of full-frame [ny, nx] tensors (any dtype; rows are contiguous byte ranges)."""
    sends, recvs = row_transfers(infos, rank, have, need)
    s, r = [], []
    for t in tensors:
        if t is None:
            continue
        for q, a, b in sends:
            s.append((q, _bytes_view(t[a:b])))
        for q, a, b in recvs:
            r.append((q, _bytes_view(t[a:b])))
    comm.sendrecv(s, r)
    return sum(x.numel() for _, x in s), sum(x.numel() for _, x in r)


class DistComm:
    """torch.distributed communicator: NCCL for CUDA tensors (P2P send/recv and all-reduce ride
    NVLink/NVSwitch), gloo for the CPU tests."""

    def __init__(self, group=None):
        import torch.distributed as dist
        self.dist, self.group = dist, group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)

    def all_reduce_sum(self, t):
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM, group=self.group)
        return t

    def sendrecv(self, sends, recvs):
        ops = [self.dist.P2POp(self.dist.isend, t, q, self.group) for q, t in sends]
        ops += [self.dist.P2POp(self.dist.irecv, t, q, self.group) for q, t in recvs]
        if ops:
            for req in self.dist.batch_isend_irecv(ops):
                req.wait()

    def barrier(self):
        if self.world > 1:
            self.dist.barrier(group=self.group)


class LocalComm:
    """``world`` ranks as threads of ONE process (one GPU, one stream per rank).  Collectives are
    barriers plus device copies; used by the single-GPU parity tests of the tile protocol."""

    class _Shared:
        def __init__(self, world):
            self.world = world
            self.bar = threading.Barrier(world)
            self.slots = [None] * world
            self.mail = {}
            self.lock = threading.Lock()

    def __init__(self, shared, rank):
        self.sh, self.rank, self.world = shared, rank, shared.world

    @staticmethod
    def _sync(t):
        if t.is_cuda:
            import torch
            torch.cuda.current_stream(t.device).synchronize()

    def all_reduce_sum(self, t):
        sh = self.sh
        self._sync(t)
        sh.slots[self.rank] = t
        sh.bar.wait()
        tot = sh.slots[0].clone()
        for q in range(1, self.world):
            tot += sh.slots[q]
        self._sync(tot)
        sh.bar.wait()
        t.copy_(tot)
        self._sync(t)
        sh.bar.wait()
        return t

    def sendrecv(self, sends, recvs):
        sh = self.sh
        for q, t in sends:
            self._sync(t)
            with sh.lock:
                sh.mail.setdefault((self.rank, q), []).append(t)
        sh.bar.wait()
        taken = {}
        for q, t in recvs:
            i = taken.get(q, 0)
            t.copy_(sh.mail[(q, self.rank)][i])
            taken[q] = i + 1
            self._sync(t)
        sh.bar.wait()
        if self.rank == 0:
            sh.mail.clear()
        sh.bar.wait()

    def barrier(self):
        self.sh.bar.wait()


def _chk(ctx, rc):
    if rc != 0:
        ctx._err(rc)


def subtract_tiled(ctx, comm, p, sci, ref, sci_rms=None, ref_rms=None, sci_mask=None, ref_mask=None,
                   exchange_inputs=True, stats=None):
    """One frame, ``comm.world`` ranks.  Every argument is a full-frame [ny, nx] CUDA tensor on this
    rank's device; only rows [row0, row1) of ``plan(p, world, rank)`` need to hold data when
    ``exchange_inputs`` (the halo rows are fetched from their owners), else rows [in_row0, in_row1).
    Runs on torch's current stream.  Returns a ``_lib.Result`` (diff, noise, mask, regions, sol,
    stamps) on rank 0 and None elsewhere."""
    import torch
    lib = _bind(ctx.lib)
    rank, world = comm.rank, comm.world
    ny, nx = sci.shape
    assert (p.nx, p.ny) == (nx, ny)
    dev = sci.device
    for t in (sci, ref, sci_rms, ref_rms, sci_mask, ref_mask):
        if t is not None:
            assert t.is_cuda and t.is_contiguous() and t.device.index == ctx.device and tuple(t.shape) == (ny, nx)
    infos = plan_all(p, world)
    me = infos[rank]
    stream = torch.cuda.current_stream(dev)
    if stream.cuda_stream == 0:
        raise ValueError("subtract_tiled must run under a non-default torch.cuda.stream (the one its collectives use)")
    moved = [0, 0]
    profile = bool(stats is not None and stats.get("profile"))
    phases, t_last = {}, [0.0]
    if profile:
        import time
        stream.synchronize()
        t_last[0] = time.perf_counter()

    def mark(name):
        """profile mode only: wall-clock per phase with a stream sync at each boundary (perturbs the run)"""
        if profile:
            stream.synchronize()
            t = time.perf_counter()
            phases[name] = phases.get(name, 0.0) + (t - t_last[0]) * 1e3
            t_last[0] = t

    def count(sr):
        moved[0] += sr[0]
        moved[1] += sr[1]

    def ptr(t):
        return None if t is None else C.c_void_p(t.data_ptr())

    def md(t):
        return 0 if t is None else _lib.MASK_DTYPES[str(t.dtype).replace("torch.", "")]

    if exchange_inputs and world > 1:
        count(exchange_rows([sci, ref, sci_rms, ref_rms, sci_mask, ref_mask], infos, rank, comm))
    mark("halo_inputs")
    diff = torch.empty((ny, nx), dtype=torch.float32, device=dev)
    noise = torch.empty((ny, nx), dtype=torch.float32, device=dev)
    mask = torch.empty((ny, nx), dtype=torch.int32, device=dev)
    h = ctx.h
    _chk(ctx, lib.zb_tile_begin(h, C.byref(p), world, rank, ptr(sci), ptr(ref), ptr(sci_rms), ptr(ref_rms),
                               ptr(sci_mask), md(sci_mask), ptr(ref_mask), md(ref_mask),
                               ptr(diff), ptr(noise), ptr(mask), C.c_void_p(stream.cuda_stream)))
    mark("mask+search")
    # ---- stamp search records: merge, replicated stale-slot walk / retry (main.c:909-998)
    words = np.zeros(me.n_cells * me.cell_words, np.int32)
    action, cell = C.c_int(0), C.c_int(-1)
    rounds = 0
    while True:
        _chk(ctx, lib.zb_tile_cells_export(h, words.ctypes.data_as(C.c_void_p)))
        if world > 1:
            t = torch.from_numpy(words).to(dev)
            comm.all_reduce_sum(t)
            words = np.ascontiguousarray(t.cpu().numpy())
        _chk(ctx, lib.zb_tile_cells_import(h, words.ctypes.data_as(C.c_void_p)))
        _chk(ctx, lib.zb_tile_resolve(h, C.byref(action), C.byref(cell)))
        rounds += 1
        if action.value == 0:
            break
        if action.value == 1:
            _chk(ctx, lib.zb_tile_research(h, cell.value))
        else:
            _chk(ctx, lib.zb_tile_begin_retry(h))
        if rounds > 4 * me.n_cells + 8:
            raise _lib.ZbError(-5, "stamp resolve did not converge")
    mark("merge_records+resolve")
    # ---- the search wrote claim bits into the mask rows of its stamps: refresh the halo rows
    scratch_i = torch.empty((ny, nx), dtype=torch.int32, device=dev)
    if world > 1:
        _chk(ctx, lib.zb_tile_rows(h, TILE_MASK_IN, 0, me.row0, me.row1, ptr(scratch_i[me.row0:me.row1])))
        count(exchange_rows([scratch_i], infos, rank, comm))
        for a, b in ((me.in_row0, me.row0), (me.row1, me.in_row1)):
            if b > a:
                _chk(ctx, lib.zb_tile_rows(h, TILE_MASK_IN, 1, a, b, ptr(scratch_i[a:b])))
    mark("halo_mask")
    # ---- fit loop (alard.c:316-381): per-stamp terms merged, global system replicated
    _chk(ctx, lib.zb_tile_fill(h))
    mark("fill")
    gram = torch.empty(me.n_slots * me.gram_len, dtype=torch.float64, device=dev)
    rhs = torch.empty(me.n_slots * me.rhs_len, dtype=torch.float64, device=dev)
    sig = torch.empty(me.n_slots, dtype=torch.float64, device=dev)
    active = C.c_int(1)
    iters = 0
    while active.value:
        if world > 1:
            _chk(ctx, lib.zb_tile_normal_export(h, ptr(gram), ptr(rhs)))
            comm.all_reduce_sum(gram)
            comm.all_reduce_sum(rhs)
            _chk(ctx, lib.zb_tile_normal_import(h, ptr(gram), ptr(rhs)))
        mark("allreduce_normal")
        _chk(ctx, lib.zb_tile_solve(h))
        mark("solve+sig")
        if world > 1:
            _chk(ctx, lib.zb_tile_sig_export(h, ptr(sig)))
            comm.all_reduce_sum(sig)
            _chk(ctx, lib.zb_tile_sig_import(h, ptr(sig)))
        mark("allreduce_sig")
        _chk(ctx, lib.zb_tile_clip(h, C.byref(active)))
        mark("clip+refill")
        iters += 1
        if iters > 4 * me.n_slots * 8:
            raise _lib.ZbError(-5, "fit loop did not converge")
    # ---- fused convolution / output pass on the rank's block rows
    _chk(ctx, lib.zb_tile_convolve(h))
    mark("convolve")
    # ---- gather rows on rank 0: the three products and the three pre-masking arrays the QA statistics read
    if world > 1:
        scratch_f = [torch.empty((ny, nx), dtype=torch.float32, device=dev) for _ in range(2)]
        work = [(TILE_RDIFF, scratch_f[0]), (TILE_RNOISE, scratch_f[1]), (TILE_OMASK, scratch_i)]
        if me.out_row1 > me.out_row0 and rank != 0:
            for which, buf in work:
                _chk(ctx, lib.zb_tile_rows(h, which, 0, me.out_row0, me.out_row1, ptr(buf[me.out_row0:me.out_row1])))
        allt = [diff, noise, mask] + [b for _, b in work]
        sends, recvs = [], []
        if rank != 0:
            if me.out_row1 > me.out_row0:
                sends = [(0, _bytes_view(t[me.out_row0:me.out_row1])) for t in allt]
        else:
            for q in range(1, world):
                o = infos[q]
                if o.out_row1 > o.out_row0:
                    recvs += [(q, _bytes_view(t[o.out_row0:o.out_row1])) for t in allt]
        comm.sendrecv(sends, recvs)
        moved[0] += sum(x.numel() for _, x in sends)
        moved[1] += sum(x.numel() for _, x in recvs)
        if rank == 0:
            lo = infos[0].out_row1
            for which, buf in work:
                _chk(ctx, lib.zb_tile_rows(h, which, 1, lo, ny, ptr(buf[lo:ny])))
    mark("gather")
    if stats is not None:
        stats.update(fit_iters=iters, resolve_rounds=rounds, sent_bytes=moved[0], recv_bytes=moved[1], info=me.as_dict())
    if rank != 0:
        stream.synchronize()
        if profile:
            stats["phases_ms"] = {k: round(v, 3) for k, v in phases.items()}
        return None
    r = ctx._outputs(p)
    _chk(ctx, lib.zb_tile_finish(h, r.regions, r.sol.ctypes.data_as(C.c_void_p), r.stamps))
    mark("qa_stats")
    if profile:
        stats["phases_ms"] = {k: round(v, 3) for k, v in phases.items()}
    r.diff, r.noise, r.mask = diff, noise, mask
    return r


def run_local(world, p, arrays, device=0):
    """Run ``subtract_tiled`` with ``world`` ranks as threads on ONE GPU (LocalComm).  ``arrays``:
    dict of full-frame torch CUDA tensors (sci, ref, sci_rms, ref_rms, sci_mask, ref_mask; None ok).
    Each rank gets its OWN copy holding only its rows (the rest poisoned), so the test exercises the
    halo exchanges.  Returns (result of rank 0, per-rank stats)."""
    import torch
    shared = LocalComm._Shared(world)
    out, errs, stats = [None] * world, [None] * world, [dict() for _ in range(world)]
    infos = plan_all(p, world)

    def body(rank):
        ctx = None
        try:
            torch.cuda.set_device(device)
            ctx = _lib.Context(device)
            comm = LocalComm(shared, rank)
            me = infos[rank]
            st = torch.cuda.Stream(device=device)
            with torch.cuda.stream(st):
                mine = {}
                for k, v in arrays.items():
                    if v is None:
                        mine[k] = None
                        continue
                    t = torch.full_like(v, 77) if v.dtype != torch.float32 else torch.full_like(v, float("nan"))
                    t[me.row0:me.row1] = v[me.row0:me.row1]
                    mine[k] = t
                st.synchronize()
                r = subtract_tiled(ctx, comm, p, mine["sci"], mine["ref"], mine.get("sci_rms"), mine.get("ref_rms"),
                                   mine.get("sci_mask"), mine.get("ref_mask"), stats=stats[rank])
                st.synchronize()
                out[rank] = r
        except BaseException as e:      # surfaced to the caller below
            errs[rank] = e
            shared.bar.abort()
        finally:
            if ctx is not None and rank != 0:
                ctx.close()
    ths = [threading.Thread(target=body, args=(r,)) for r in range(world)]
    for t in ths:
        t.start()
    for t in ths:
        t.join()
    for e in errs:
        if e is not None and not isinstance(e, threading.BrokenBarrierError):
            raise e
    for e in errs:
        if e is not None:
            raise e
    return out[0], stats
