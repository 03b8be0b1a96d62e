import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def cases():
    from tests import util
    return util


@pytest.fixture(autouse=True)
def _emulated_library(monkeypatch):
    """FB200_EMULATE=1 python -m pytest tests -m gpu: run the GPU parity tests on the CPU against the CUDA-on-CPU emulation of
    tests/emu (test infrastructure; add FB200_EMU_ASAN=1 and LD_PRELOAD=$(g++ -print-file-name=libasan.so) for heap checks).
    Without the variable nothing is patched: the tests load the product library and need a CUDA device."""
    if os.environ.get("FB200_EMULATE"):
        from fesom2_b200 import api
        from tests import emu
        monkeypatch.setattr(api, "_lib", emu.load())
    yield
