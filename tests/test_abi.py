"""CPU test: the C-ABI library loads and exports every symbol include/fesom_b200.h declares; without a GPU
the compute entry points fail loudly (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

from fesom2_b200 import api, build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "fesom_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(psolver_init|psolve|psolver_final|fb200_\w+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    so = build.build_cuda()
    L = C.CDLL(so)
    names = declared_symbols()
    assert len(names) >= 20
    for nm in names:
        assert hasattr(L, nm), nm
    assert set(names) == set(api.EXPORTS)


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from tests import util
    hm = util.build_model("band")
    with pytest.raises(api.FesomB200Error, match="no CUDA device"):
        api.TracerEngine(hm)
    with pytest.raises(api.FesomB200Error, match="no CUDA device"):
        api.SSHSolver(hm)
