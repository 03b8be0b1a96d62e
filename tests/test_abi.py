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


def test_product_never_refers_to_the_test_emulation_or_the_oracle():
    """tests/emu (CUDA-on-CPU emulation) and oracle/ are test infrastructure: no product source may import, load or name them
    as something to run (bench.py's cpu_baseline / --impl reference legs and __graft_entry__.smoke() are the stated exceptions
    for the oracle)."""
    pkg = os.path.join(ROOT, "fesom2_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if not f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp")):
                continue
            src = open(os.path.join(dp, f)).read()
            assert "FB200_EMULATE" not in src and "libfesom_b200_emu" not in src and "tests.emu" not in src and "from tests" not in src, f
            assert "import oracle" not in src and "from oracle" not in src and "liboracle" not in src, f
            assert "import harness" not in src and "from harness" not in src, f          # host mirror, partitioner, synthetic meshes: harness/
    for f in ("bench.py", "__graft_entry__.py"):
        src = open(os.path.join(ROOT, f)).read()
        assert "FB200_EMULATE" not in src and "tests.emu" not in src and "libfesom_b200_emu" not in src, f


def test_psolver_init_refuses_a_multi_process_launch_without_comm_init(monkeypatch):
    """the reference's psolver_init gets the MPI communicator; here fb200_comm_init introduces the ranks.  A process started by a
    launcher as one of several that never called it must not silently solve the first partition on every rank (error 47,
    raised before any CUDA call)."""
    from tests import util
    hm = util.build_model("pi")
    hm.compute_ssh_rhs()
    monkeypatch.setenv("WORLD_SIZE", "4")
    monkeypatch.delenv("FB200_SINGLE_RANK", raising=False)
    with pytest.raises(api.FesomB200Error, match="fb200_comm_init was not called"):
        api.SSHSolver(hm)
    api.lib().fb200_clear_status()


def test_fortran_interface_file_covers_the_host_facing_abi():
    """include/fesom_b200_interface.F90 (= the module text of INTEGRATION.md) declares every entry point a Fortran host calls; what
    it leaves out are the test / diagnostic entry points listed here, and psolver_init / psolve, which FESOM declares itself"""
    f90 = open(os.path.join(ROOT, "include", "fesom_b200_interface.F90")).read()
    have = set(re.findall(r"name='(\w+)'", f90))
    diagnostic = {"fb200_amg_host_build", "fb200_amg_host_dims", "fb200_amg_host_fetch", "fb200_amg_host_free", "fb200_amg_host_levels",
                  "fb200_debug_fetch", "fb200_kernel_launches", "fb200_last_ms", "fb200_profile", "fb200_profile_report",
                  "fb200_psolve_amg_folded", "fb200_psolve_amg_info", "fb200_psolver_use_graph", "fb200_selftest_div", "fb200_status_message"}
    names = set(declared_symbols())
    assert have <= names
    assert names - have - diagnostic == {"psolver_init", "psolve", "psolver_final"}
    doc = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    for nm in have:
        assert nm in doc, nm
