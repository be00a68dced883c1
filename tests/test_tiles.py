"""Row-tile sessions (BASELINE.json configs[4]): ownership arithmetic and the halo-exchange schedule on
CPU (including a world_size-2 gloo run of the communicator), and — under -m gpu — the tile protocol
against the single-GPU path and the oracle, bit for bit."""
import os

import numpy as np
import pytest

import cases as CS
from zuds_pipeline_b200 import _lib, params as P, tiles


def _params(shape, hw=10, rss=24, nsx=10, nsy=10, ko=2):
    p = P.default_params(shape[1], shape[0])
    p.hw_kernel, p.hw_ksstamp, p.ker_order, p.bg_order = hw, rss, ko, 0
    p.nreg_x = p.nreg_y = 1
    p.nstamp_x, p.nstamp_y = nsx, nsy
    p.force_convolve, p.phot_normalize = ord("t"), ord("i")
    return p


C5 = dict(shape=(6160, 6144), hw=10, rss=24, nsx=20, nsy=20)


@pytest.mark.parametrize("world", [1, 2, 3, 4, 8])
@pytest.mark.parametrize("geom", [C5, dict(shape=(700, 640), hw=5, rss=12, nsx=6, nsy=7), dict(shape=(3080, 3072), hw=15, rss=15, nsx=32, nsy=32)],
                         ids=["c5", "small", "c4"])
def test_plan_partitions_rows_cells_blocks(world, geom):
    p = _params(**geom)
    sz = _lib.sizes_of(p)
    infos = tiles.plan_all(p, world)
    ny, hw, B = p.ny, p.hw_kernel, sz.kc_step
    nby = (ny + B - 1) // B
    for key, end in (("cell_row", sz.nstamp_y), ("row", ny), ("blk_row", nby), ("out_row", ny)):
        assert getattr(infos[0], key + "0") == 0 and getattr(infos[-1], key + "1") == end
        for a, b in zip(infos, infos[1:]):
            assert getattr(a, key + "1") == getattr(b, key + "0")          # contiguous, disjoint, complete
        for i in infos:
            assert getattr(i, key + "0") <= getattr(i, key + "1")
    width = int(p.hw_kernel * p.kf_spread_mask1)
    w2 = width // 2 if width > 0 else 0
    for i in infos:
        assert 0 <= i.in_row0 <= i.row0 <= i.row1 <= i.in_row1 <= ny
        # every stamp rectangle of the rank's grid-cell rows lies in its own rows (main.c:849-852) ...
        for gl in range(i.cell_row0, i.cell_row1):
            sy0 = gl * ny // sz.nstamp_y
            sy1 = min(sy0 + sz.fw_stamp - 1, ny - 1)
            assert i.row0 <= sy0 and sy1 < i.row1
            # ... and its sub-stamps (+ convolution halo, + mask dilation) in the rows it needs
            assert i.in_row0 <= max(0, sy0 - sz.s_border - w2) and min(ny, sy1 + 1 + sz.s_border + w2) <= i.in_row1
        # convolution blocks of the rank: output rows, and input rows +- hw (alard.c:1286-1306)
        if i.blk_row1 > i.blk_row0:
            first, last = hw + i.blk_row0 * B, min(hw + i.blk_row1 * B, ny - hw)
            if last > first:
                assert i.out_row0 <= first and last <= max(i.out_row1, ny - hw)
                assert i.in_row0 <= max(0, first - hw - w2) and min(ny, last + hw + w2) <= i.in_row1
        assert i.n_cells == sz.n_stamps and i.cell_words * 4 >= 100
        nvec = sz.n_comp_ker + sz.n_bg_vectors
        assert (i.gram_len, i.rhs_len) == (nvec * (nvec + 1) // 2, nvec)


def test_plan_rejects_multi_region_and_bad_rank():
    p = _params((700, 640), hw=5, rss=12)
    p.nreg_x = 2
    with pytest.raises(_lib.ZbError):
        tiles.plan(p, 2, 0)
    p.nreg_x = 1
    with pytest.raises(_lib.ZbError):
        tiles.plan(p, 2, 2)


@pytest.mark.parametrize("world", [2, 3, 8])
def test_row_transfers_are_matched_and_sufficient(world):
    p = _params(**C5)
    infos = tiles.plan_all(p, world)
    plans = [tiles.row_transfers(infos, r) for r in range(world)]
    for r, (sends, recvs) in enumerate(plans):
        got = np.zeros(p.ny, bool)
        got[infos[r].row0:infos[r].row1] = True
        for q, a, b in recvs:
            assert (r, a, b) in plans[q][0]                 # the owner sends exactly this range to me
            assert infos[q].row0 <= a and b <= infos[q].row1
            got[a:b] = True
        assert got[infos[r].in_row0:infos[r].in_row1].all()
        for q, a, b in sends:
            assert (r, a, b) in plans[q][1]


def _gloo_worker(rank, world, port, out):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    comm = tiles.DistComm()
    p = _params((700, 640), hw=5, rss=12, nsx=6, nsy=7)
    infos = tiles.plan_all(p, world)
    me = infos[rank]
    full = [torch.arange(p.ny * p.nx, dtype=torch.float32).reshape(p.ny, p.nx),
            (torch.arange(p.ny * p.nx, dtype=torch.int32).reshape(p.ny, p.nx) % 31000).to(torch.int16),
            (torch.arange(p.ny * p.nx, dtype=torch.int64).reshape(p.ny, p.nx) % 251).to(torch.uint8)]
    mine = []
    for f in full:
        t = torch.full_like(f, 7)
        t[me.row0:me.row1] = f[me.row0:me.row1]
        mine.append(t)
    tiles.exchange_rows(mine + [None], infos, rank, comm)
    ok = all(torch.equal(t[me.in_row0:me.in_row1], f[me.in_row0:me.in_row1]) for t, f in zip(mine, full))
    # the "non-zero on exactly one rank" merge used for search records / normal equations / figures of merit
    words = torch.zeros(40, dtype=torch.int32)
    words[rank * 20:(rank + 1) * 20] = torch.arange(20, dtype=torch.int32) - 7 + 1000 * rank
    comm.all_reduce_sum(words)
    want = torch.cat([torch.arange(20, dtype=torch.int32) - 7, torch.arange(20, dtype=torch.int32) - 7 + 1000])
    ok = ok and torch.equal(words, want)
    g = torch.zeros(6, dtype=torch.float64)
    g[rank * 3:(rank + 1) * 3] = torch.tensor([0.1, -1e-300, 3e200], dtype=torch.float64) * (rank + 1)
    comm.all_reduce_sum(g)
    ok = ok and torch.equal(g, torch.tensor([0.1, -1e-300, 3e200, 0.2, -2e-300, 6e200], dtype=torch.float64))
    flags = [None] * world
    dist.all_gather_object(flags, bool(ok))
    if rank == 0:
        with open(out, "w") as f:
            f.write(repr(flags))
    dist.destroy_process_group()


def test_two_ranks_gloo_halo_exchange_and_merge(tmp_path):
    import torch.multiprocessing as mp
    out = tmp_path / "ok.txt"
    port = 29600 + (os.getpid() % 1000)
    mp.spawn(_gloo_worker, args=(2, port, str(out)), nprocs=2, join=True)
    assert eval(out.read_text()) == [True, True]


# ------------------------------------------------------------------------------------------ GPU

TILE_CASES = [
    # one region, enough grid-cell rows to split 2 and 3 ways
    CS.Case("tile_c1_small", (700, 640), 2.0, 1, {"ko": 2, "nsx": 6, "nsy": 7}, dict(nstars=900, bad_frac=2e-4)),
    # convolve the science image, bg order 1, clipping active
    CS.Case("tile_conv_image", (640, 600), 2.4, 1, {"ko": 1, "bgo": 1, "nsx": 5, "nsy": 6, "ks": 1.2}, dict(nstars=800, bad_frac=2e-4),
            replace={"-c": "i", "-n": "t"}),
    # few stamps: the one-shot threshold retry crosses ranks
    CS.Case("tile_retry", (512, 512), 3.0, 1, {"ko": 0, "ft": 1900.0}, dict(nstars=300, bad_frac=1e-4)),
    # no stamps anywhere
    CS.Case("tile_no_stamps", (320, 320), 4.0, 1, {"ko": 2}, dict(nstars=8, bad_frac=2e-3)),
]


def _run_tiled(case, world):
    import torch
    pair = CS.make_pair(case)
    p = CS.params_of(case, pair)
    ins = CS.inputs_of(case, pair)
    names = ("sci", "ref", "sci_rms", "ref_rms", "sci_mask", "ref_mask")
    arrays = {k: (None if a is None else torch.from_numpy(np.ascontiguousarray(a.view(np.uint8) if a.dtype == np.bool_ else a)).cuda())
              for k, a in zip(names, ins)}
    r, stats = tiles.run_local(world, p, arrays)
    torch.cuda.synchronize()
    r.diff, r.noise, r.mask = r.diff.cpu().numpy(), r.noise.cpu().numpy(), r.mask.cpu().numpy()
    return p, ins, r, stats


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 3])
@pytest.mark.parametrize("case", TILE_CASES, ids=lambda c: c.name)
def test_tiles_equal_single_gpu_and_oracle(case, world, oracle, gpu_ctx):
    import hp_helpers as H
    from test_gpu_parity import compare
    p, ins, t, stats = _run_tiled(case, world)
    g = gpu_ctx.subtract_host(p, *ins)
    assert H.stamp_table(g) == H.stamp_table(t)
    assert np.array_equal(g.sol, t.sol), "fitted coefficients must not depend on the tiling"
    assert np.array_equal(g.mask, t.mask)
    assert np.array_equal(g.diff, t.diff, equal_nan=True) and np.array_equal(g.noise, t.noise, equal_nan=True)
    for f in ("flags", "n_stamps", "fit_iters", "n_skipped", "ksum", "sssig", "ssscat", "fsig", "fscat", "x2norm", "nx2norm",
              "dmean", "dsige", "dsig", "dmeano", "dsigeo", "dsigo", "nscalo"):
        a, b = getattr(g.regions[0], f), getattr(t.regions[0], f)
        assert a == b or (a != a and b != b), f
    if world > 1 and t.regions[0].n_stamps > 0:
        assert all(s["sent_bytes"] > 0 or s["recv_bytes"] > 0 for s in stats)
    o = H.run_oracle(oracle, p, *ins)
    compare(o, t)


@pytest.mark.gpu
def test_tiles_c5_full_size_against_single_gpu(gpu_ctx):
    """BASELINE.json configs[4] at full size, 4 row tiles (threads on one GPU) vs the whole frame."""
    import hp_helpers as H
    case = CS.FULL["c5"]
    p, ins, t, stats = _run_tiled(case, 4)
    g = gpu_ctx.subtract_host(p, *ins)
    assert t.regions[0].n_stamps > 100
    assert H.stamp_table(g) == H.stamp_table(t)
    assert np.array_equal(g.sol, t.sol) and np.array_equal(g.mask, t.mask)
    assert np.array_equal(g.diff, t.diff, equal_nan=True) and np.array_equal(g.noise, t.noise, equal_nan=True)
    assert g.regions[0].fsig == t.regions[0].fsig and g.regions[0].dsig == t.regions[0].dsig
