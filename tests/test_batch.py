"""Sharding of the pair list over ranks (one rank per GPU, no data-path collective): host logic on CPU,
including a world_size-2 gloo run; the GPU pool itself under -m gpu."""
import os

import numpy as np
import pytest

from zuds_pipeline_b200 import batch, dosub


def test_shard_is_array_split():
    items = ["p%03d" % i for i in range(23)]
    for world in (1, 2, 3, 8, 30):
        parts = [batch.shard(items, r, world) for r in range(world)]
        want = [list(a) for a in np.array_split(items, world)]
        assert parts == want
        assert sum(parts, []) == items
    assert batch.shard([], 0, 4) == []


def test_share_of_work_env_and_job_array(tmp_path, monkeypatch):
    f = tmp_path / "pairs.txt"
    f.write_text("# comment\n" + "".join("sci%d.fits ref.fits\n" % i for i in range(10)))
    monkeypatch.setenv("RANK", "1")
    monkeypatch.setenv("WORLD_SIZE", "4")
    mine = dosub.get_my_share_of_work(str(f))
    assert mine == ["sci3.fits ref.fits", "sci4.fits ref.fits", "sci5.fits ref.fits"]
    monkeypatch.setenv("SLURM_ARRAY_JOB_ID", "77")
    monkeypatch.setenv("SLURM_ARRAY_TASK_ID", "1")
    monkeypatch.setenv("SLURM_ARRAY_TASK_MAX", "1")
    mine = dosub.get_my_share_of_work(str(f), rank=0, size=2)         # second half of the list, first rank
    assert mine == ["sci5.fits ref.fits", "sci6.fits ref.fits", "sci7.fits ref.fits"]


def _gloo_worker(rank, world, port, listfile, out):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mine = dosub.get_my_share_of_work(listfile)
    gathered = [None] * world
    dist.all_gather_object(gathered, mine)                 # test-only check; the data path has no collective
    dist.barrier()
    if rank == 0:
        with open(out, "w") as f:
            f.write(repr(gathered))
    dist.destroy_process_group()


def test_two_ranks_gloo_cover_the_list_once(tmp_path):
    import torch.multiprocessing as mp
    f = tmp_path / "pairs.txt"
    lines = ["sci%d.fits ref%d.fits" % (i, i % 3) for i in range(9)]
    f.write_text("\n".join(lines) + "\n")
    out = tmp_path / "gathered.txt"
    port = 29500 + (os.getpid() % 1000)
    mp.spawn(_gloo_worker, args=(2, port, str(f), str(out)), nprocs=2, join=True)
    gathered = eval(out.read_text())
    assert gathered[0] + gathered[1] == lines and len(gathered[0]) == 5 and len(gathered[1]) == 4


@pytest.mark.gpu
def test_gpu_pool_matches_single_context(gpu_ctx):
    import cases as CS
    jobs = []
    for case in CS.SMALL[:4]:
        pair = CS.make_pair(case)
        jobs.append((CS.params_of(case, pair),) + tuple(CS.inputs_of(case, pair)))
    single = [gpu_ctx.subtract_host(*j) for j in jobs]
    with batch.GpuPool(0, contexts=3) as pool:
        many = pool.subtract_host_many(jobs * 2)
        assert pool.launches() > 0
    for k, r in enumerate(many):
        s = single[k % len(jobs)]
        assert np.array_equal(r.diff, s.diff) and np.array_equal(r.noise, s.noise) and np.array_equal(r.mask, s.mask)


@pytest.mark.gpu
def test_dosub_driver_end_to_end(tmp_path):
    import cases as CS
    from test_host_api import _objects
    lines = []
    for k, case in enumerate(CS.SMALL[:2]):
        d = tmp_path / ("p%d" % k)
        d.mkdir()
        sci, ref = _objects(CS.make_pair(case), str(d))
        lines.append("%s %s" % (sci.local_path, ref.local_path))
    lines.append("%s %s" % (tmp_path / "missing.fits", tmp_path / "missing_ref.fits"))     # fails, is skipped
    listfile = tmp_path / "pairs.txt"
    listfile.write_text("\n".join(lines) + "\n")
    assert dosub.main([str(listfile), "--nreg-side", "1", "--ko", "1", "--tmpdir", str(tmp_path / "t"), "--contexts", "2"]) == 0
    for ln in lines[:2]:
        s, r = ln.split()
        from zuds_pipeline_b200.subtraction import sub_name
        assert os.path.exists(sub_name(s, r))
    with pytest.raises(dosub.PredecessorError):
        dosub.do_one(lines[0], tmpdir=str(tmp_path / "t"), subtract_back=False, nreg_side=1)
