import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")
    config.addinivalue_line("markers", "multigpu: needs >= 2 GPUs on one box")


def _gpu_count() -> int:
    try:
        import ctypes

        cuda = ctypes.CDLL("libcuda.so.1")
        if cuda.cuInit(0) != 0:
            return 0
        n = ctypes.c_int(0)
        cuda.cuDeviceGetCount(ctypes.byref(n))
        return n.value
    except OSError:
        return 0


GPU_COUNT = _gpu_count()


@pytest.fixture(scope="session")
def dl():
    """The C-ABI library initialised as a 1-rank world on cuda:0.

    No GPU -> skip. GPU but no library -> error (never a silent fallback)."""
    if GPU_COUNT == 0:
        pytest.skip("no CUDA device visible")
    from dlnetbench_b200 import capi

    capi.lib()  # raises DlnbError if libdlnb.so is missing
    capi.init(rank=0, world=1, device=0, session="pytest-%d" % os.getpid())
    yield capi
    capi.finalize()


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle_capi

    oracle_capi.lib()
    return oracle_capi
