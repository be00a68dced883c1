import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(autouse=True)
def _host_prestep_without_gpu():
    """The product's wrapper pre-step runs on the GPU and raises without one.  The CPU-only tests of the host
    logic (command strings, FITS products of the wrapper classes) ask for the numpy restatement explicitly."""
    from zuds_pipeline_b200 import hotpants
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except ImportError:
        has_gpu = False
    old = hotpants.prestep
    if not has_gpu:
        hotpants.prestep = "host"
    yield
    hotpants.prestep = old


@pytest.fixture(scope="session")
def oracle():
    import hp_helpers as H
    return H.load_oracle()


@pytest.fixture(scope="session")
def gpu_ctx():
    from zuds_pipeline_b200 import _lib
    ctx = _lib.Context(0)
    yield ctx
    ctx.close()
