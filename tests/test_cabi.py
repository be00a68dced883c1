"""The C-ABI library loads and exports every symbol include/zb200.h declares; host-side logic
(parameter defaults, checks, derived sizes) agrees with the oracle.  No GPU needed, no compute calls."""
import ctypes as C
import os
import re

import pytest

from zuds_pipeline_b200 import _lib, params as P

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    src = open(os.path.join(ROOT, "include", "zb200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(zb_[a-z0-9_]+)\s*\(", src)))


def test_every_declared_symbol_is_exported():
    lib = _lib.load()
    names = declared_functions()
    assert len(names) >= 12
    for n in names:
        assert hasattr(lib, n), "libzb200.so does not export %s" % n
    assert set(_lib.EXPORTS) == set(names)


def test_struct_layouts_match_header():
    lib = _lib.load()
    lib.zb_abi_sizeof.restype = C.c_int
    lib.zb_abi_sizeof.argtypes = [C.c_int]
    assert lib.zb_abi_sizeof(0) == C.sizeof(P.ZbParams)
    assert lib.zb_abi_sizeof(1) == C.sizeof(P.ZbSizes)
    assert lib.zb_abi_sizeof(2) == C.sizeof(P.ZbRegionInfo)
    assert lib.zb_abi_sizeof(3) == C.sizeof(P.ZbStampInfo)
    from zuds_pipeline_b200 import tiles
    assert lib.zb_abi_sizeof(4) == C.sizeof(tiles.ZbTileInfo)
    assert lib.zb_abi_sizeof(5) == C.sizeof(P.ZbHostPair)


def test_defaults_match_python_and_oracle(oracle):
    lib = _lib.load()
    a, b, c = P.ZbParams(), P.default_params(), P.ZbParams()
    lib.zb_params_default(C.byref(a))
    oracle.hp_oracle_params_default(C.byref(c))
    for name, _ in P.ZbParams._fields_:
        va, vb, vc = getattr(a, name), getattr(b, name), getattr(c, name)
        if hasattr(va, "__len__"):
            assert list(va) == list(vb) == list(vc), name
        else:
            assert va == vb == vc, name


@pytest.mark.parametrize("shape,kw", [((3080, 3072), {}), ((3080, 3072), {"hw_kernel": 15, "hw_ksstamp": 15, "nstamp_x": 32, "nstamp_y": 32}),
                                      ((6160, 6144), {"nstamp_x": 20, "nstamp_y": 20}), ((200, 208), {"hw_kernel": 3, "hw_ksstamp": 7}),
                                      ((3080, 3072), {"nreg_x": 3, "nreg_y": 3, "hw_ksstamp": 24})])
def test_sizes_match_oracle(shape, kw, oracle):
    p = P.default_params(shape[1], shape[0])
    for k, v in kw.items():
        setattr(p, k, v)
    s1, s2 = _lib.sizes_of(p), P.ZbSizes()
    oracle.hp_oracle_sizes(C.byref(p), C.byref(s2))
    for name, _ in P.ZbSizes._fields_:
        assert getattr(s1, name) == getattr(s2, name), name


def test_param_checks_follow_vargs():
    p = P.default_params(512, 512)
    p.force_convolve = ord("t")
    assert _lib.check_params(p)[0] == 0
    bad = [("hw_kernel", 0), ("ker_order", -1), ("bg_order", -1), ("scale_fit_thresh", 1.5), ("min_frac_good_stamps", -0.1),
           ("t_gain", -1.0), ("force_convolve", ord("x")), ("phot_normalize", ord("q"))]
    for name, v in bad:
        q = P.default_params(512, 512)
        q.force_convolve = ord("t")
        setattr(q, name, v)
        rc, msg = _lib.check_params(q)
        assert rc == -1 and msg, name
    q = P.default_params(512, 512)
    q.force_convolve = ord("t")
    q.hw_ksstamp = 5
    rc, msg = _lib.check_params(q)
    assert rc == -1 and "conflicting" in msg                    # vargs.c: rss < r
    q = P.default_params(512, 512)                              # hotpants' own defaults: -c b, -fom v
    assert _lib.check_params(q)[0] == 0
    q.fig_merit = ord("s")                                      # a valid hotpants option outside this build
    assert _lib.check_params(q)[0] == -2


def test_create_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(_lib.ZbError):
        _lib.Context(0)
