import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _cuda_backend():
    import torch
    from cyten_b200.backend import B200AbelianBackend
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return B200AbelianBackend("cuda:0")


def _oracle_backend():
    from cyten_b200.backend import B200AbelianBackend
    from oracle import cabi_numpy
    return B200AbelianBackend("cpu", _library=cabi_numpy)


@pytest.fixture(params=[pytest.param("oracle"), pytest.param("cuda", marks=pytest.mark.gpu)])
def backend(request):
    """The host logic of cyten_b200 over (a) the numpy oracle of the C-ABI on CPU -- checks host
    logic + oracle against the golden vectors -- and (b) the real CUDA library on the GPU."""
    return _oracle_backend() if request.param == "oracle" else _cuda_backend()


@pytest.fixture
def cuda_backend():
    return _cuda_backend()


@pytest.fixture
def oracle_backend():
    return _oracle_backend()
