"""Batch driver: the GPU counterpart of scripts/dosub.py + zuds/mpi.py.

    torchrun --nproc-per-node 8 -m zuds_pipeline_b200.dosub pairs.txt [--contexts 4] [--nreg-side 3] [--ko 2]

``pairs.txt`` lists ``<science.fits> <reference.fits>`` per line (mask / rms / weight siblings next to
them, as `CalibratableImageBase.from_file` expects).  The list is split over ranks exactly as
``zuds.get_my_share_of_work`` does (SLURM job-array split first, then ``np.array_split`` over ranks,
zuds/mpi.py:47-60) — with RANK/WORLD_SIZE from the launcher instead of mpi4py, one rank per GPU, no
collective on the data path.  Per pair the semantics of scripts/dosub.py:190-211 are kept: a failing
pair is reported and skipped, an existing product is not redone (PredecessorError, :83-94).
"""
from __future__ import annotations

import argparse
import os
import sys
import time
import traceback

from . import batch
from .image import CalibratableImageBase
from .subtraction import SingleEpochSubtraction, sub_name

__all__ = ["get_my_share_of_work", "do_one", "main", "PredecessorError"]


class PredecessorError(Exception):
    """The product already exists (scripts/dosub.py:93)."""


def default_reader(fname):
    with open(fname) as f:
        return [ln.strip() for ln in f if ln.strip() and not ln.startswith("#")]


def get_my_share_of_work(fname, reader=default_reader, rank=None, size=None):
    """zuds/mpi.py:36-64 without MPI: rank/size come from the launcher's environment."""
    rank = int(os.environ.get("RANK", "0")) if rank is None else rank
    size = int(os.environ.get("WORLD_SIZE", "1")) if size is None else size
    files = reader(fname)
    if os.getenv('SLURM_ARRAY_JOB_ID') is not None:
        idx = int(os.getenv('SLURM_ARRAY_TASK_ID'))
        ntasks = int(os.getenv('SLURM_ARRAY_TASK_MAX')) + 1
        files = batch.shard(files, idx, ntasks)
    return batch.shard(files, rank, size)


def do_one(line, subclass=SingleEpochSubtraction, tmpdir='/tmp', quiet=False, **kwargs):
    """One pair: load, refuse to redo, subtract (scripts/dosub.py:26-187 minus catalog/DB steps)."""
    tstart = time.time()
    scipath, refpath = line.split()[:2]
    sci = CalibratableImageBase.from_file(scipath)
    ref = CalibratableImageBase.from_file(refpath)
    if not quiet:
        print(f'load: {time.time() - tstart:.2f} sec to load input images', flush=True)
    final = sub_name(sci.local_path, ref.local_path)
    if os.path.exists(final):
        raise PredecessorError(f'{os.path.basename(final)} already has a predecessor')
    substart = time.time()
    sub = subclass.from_images(sci, ref, data_product=False, tmpdir=tmpdir, **kwargs)
    if not quiet:
        print(f'sub: {time.time() - substart:.2f} sec to make {sub.basename}', flush=True)
    return sub


def main(argv=None):
    ap = argparse.ArgumentParser(prog="dosub")
    ap.add_argument("infile")
    ap.add_argument("--tmpdir", default="/tmp")
    ap.add_argument("--nreg-side", type=int, default=3)
    ap.add_argument("--ko", type=int, default=None)
    ap.add_argument("--bgo", type=int, default=None)
    ap.add_argument("--contexts", type=int, default=4, help="concurrent subtractions per GPU")
    args = ap.parse_args(argv)
    local = int(os.environ.get("LOCAL_RANK", "0"))
    SingleEpochSubtraction.device = local
    kws = {}
    if args.ko is not None:
        kws["ko"] = args.ko
    if args.bgo is not None:
        kws["bgo"] = args.bgo
    imgs = get_my_share_of_work(args.infile)
    from concurrent.futures import ThreadPoolExecutor

    def work(line):
        # run_hotpants keeps one zb_ctx per (device, worker thread)
        try:
            return do_one(line, tmpdir=args.tmpdir, subtract_back=False, nreg_side=args.nreg_side, hotpants_kws=kws)
        except Exception:
            traceback.print_exception(*sys.exc_info())
            return None
    t0 = time.time()
    with ThreadPoolExecutor(max_workers=args.contexts) as ex:
        done = [s for s in ex.map(work, imgs) if s is not None]
    print(f'rank {os.environ.get("RANK", "0")}: {len(done)}/{len(imgs)} subtractions in {time.time() - t0:.2f} sec', flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
