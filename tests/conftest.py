import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session", params=[(0, 1), (0, 0), (1, 1)], ids=["general", "hostloop", "smallpath"])
def ctx(request):
    """One GPU context per test session (GPU tests only).  Every test runs three times: with the general
    level-by-level path forced (small_path=0) and its level loop on the device (CUDA graph WHILE/SWITCH
    nodes); the same with the host-driven loop (loop=0); and with the default dispatch, where graphs that
    live in L2 take the whole-batch cooperative kernel (msbfs_small.cuh)."""
    import pgq_b200
    c = pgq_b200.Context(0)
    c.set_option("small_path", request.param[0])
    c.set_option("loop", request.param[1])
    c.set_option("loop_min_batches", 0)  # the device loop for every call, however small (default: only when it pays)
    yield c
    c.close()


def gpu_count():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0
