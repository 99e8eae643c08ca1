import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "my-nope-nerf_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def golden():
    import torch

    def load(name):
        return torch.load(os.path.join(GOLDEN, name + ".pt"), weights_only=False)
    return load


@pytest.fixture(scope="session", autouse=True)
def _sanitizer_cuda_core_chamfer():
    """tools/sanitizer_round.sh racecheck: NNB_SANITIZER_CUDA_CORES=1 runs every Chamfer search on the CUDA-core sweeps (racecheck
    does not model the mbarrier / async-proxy synchronisation of the tcgen05 kernels)."""
    if os.environ.get("NNB_SANITIZER_CUDA_CORES") == "1":
        from model import _cabi
        old = _cabi.lib().nnb_debug_chamfer_pass1(1)
        yield
        _cabi.lib().nnb_debug_chamfer_pass1(old)
    else:
        yield
