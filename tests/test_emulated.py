"""CPU execution of the CUDA kernels' SOURCE through the CUDA-on-CPU emulation of tests/emu (test infrastructure, see
tests/emu/cuda_emu.h): fesom2_b200/csrc/*.cu are rewritten into plain C++ (launches -> fiber
scheduler) and compiled with g++ -ffp-contract=off, then the SAME parity tests the GPU suite runs are called against that
library.  It checks the kernels' logic and arithmetic order before GPU time is spent; it says nothing about races,
occupancy or speed, and it is not a fallback: the product library never loads it (tests/test_abi.py checks that the
product fails without a device)."""
import pytest


@pytest.fixture()
def emulated(monkeypatch):
    from fesom2_b200 import api
    from tests import emu
    monkeypatch.setattr(api, "_lib", emu.load())
    yield
    

def test_emulated_tracer_step_is_bit_identical_to_oracle(emulated):
    import numpy as np
    from fesom2_b200 import api
    from oracle import binding as ob
    from tests import util
    A, B = util.build_model("pi"), util.build_model("pi")
    o = ob.TracerOracle(A)
    eng = api.TracerEngine(B)
    eng.push_step_state()
    g_tr, g_tro = B["tr_arr"].copy(), B["tr_arr_old"].copy()
    for s in range(2):
        o.step()
        eng.solve_tracers_ale(g_tr, g_tro)
        assert np.array_equal(g_tr, A["tr_arr"]) and np.array_equal(g_tro, A["tr_arr_old"]), s
    eng.close()


@pytest.mark.parametrize("resident", [False, True])
def test_emulated_relax_zonal_temp(emulated, resident):
    from tests import test_gpu_zz_zonal as t
    t.test_relax_zonal_temp_matches_oracle(resident)


@pytest.mark.parametrize("accuracy", [4])
def test_emulated_output_means(emulated, accuracy):
    from tests import test_gpu_zz_means as t
    t.test_device_means_match_stream_restatement(accuracy)


@pytest.mark.parametrize("name,kw", [("band", {"which_ale": "zlevel", "water_flux": 2e-7}), ("soufflet", {})])
def test_emulated_ssh_assembly(emulated, name, kw):
    from tests import test_gpu_zz_ssh_assembly as t
    t.test_device_ssh_assembly_matches_host_mirror(name, kw)


@pytest.mark.parametrize("name,kw", [("pi", {"which_ale": "zstar", "w_split": True, "w_max_cfl": 0.2, "umax": 0.5})])
def test_emulated_ale_producers(emulated, name, kw):
    from tests import test_gpu_tracer as t
    t.test_device_ale_producers_match_host_mirror(name, kw)


@pytest.mark.parametrize("name", ["pi"])
def test_emulated_psolve(emulated, name):
    """the SSH solver's kernels (persistent reducing SpMV, block-Jacobi and multigrid V-cycle) under the emulation, CUDA
    graphs off: the same checks against the oracle and the reference's golden solution as on the GPU"""
    from tests import test_gpu_ssh as t
    t.test_psolve_matches_oracle(name, 4)
    t.test_psolve_matches_reference_golden(name)
    t.test_psolve_multigrid_matches_reference_and_oracle(name)


def test_emulated_psolve_value_refresh(emulated):
    from tests import test_gpu_ssh as t
    t.test_value_refresh_and_warm_start()
    t.test_psolve_multigrid_single_level_is_jacobi()


def test_emulated_tile_edge_order_is_bit_identical(emulated, monkeypatch):
    """FB200_EDGE_ORDER=tile (experimental storage order of the edge tables and edge fields, context.cu) must not change a
    bit: tracers over two steps, the antidiffusive edge fluxes returned in reference order, the ALE producers and the SSH
    assembly, which index edge records through the node slots"""
    import numpy as np
    from fesom2_b200 import api
    from oracle import binding as ob
    from tests import util
    adf = {}
    for order in ("ref", "tile"):
        monkeypatch.setenv("FB200_EDGE_ORDER", order)
        A, B = util.build_model("pi"), util.build_model("pi")
        o = ob.TracerOracle(A)
        eng = api.TracerEngine(B)
        perm = eng.debug("edge_perm")
        assert np.array_equal(np.sort(perm), np.arange(perm.size))
        assert np.array_equal(perm, np.arange(perm.size)) == (order == "ref")          # the permutation is in force
        eng.push_step_state()
        g_tr, g_tro = B["tr_arr"].copy(), B["tr_arr_old"].copy()
        for s in range(2):
            o.step()
            eng.solve_tracers_ale(g_tr, g_tro)
            assert np.array_equal(g_tr, A["tr_arr"]) and np.array_equal(g_tro, A["tr_arr_old"]), (order, s)
        adf[order] = eng.debug("adf_h").copy()
        eng.close()
    assert np.abs(adf["ref"]).max() > 0 and np.array_equal(adf["ref"], adf["tile"])
    from tests import test_gpu_tracer as t1, test_gpu_zz_ssh_assembly as t2
    t1.test_device_ale_producers_match_host_mirror("band", {"which_ale": "zstar", "water_flux": 2e-7})
    t2.test_device_ssh_assembly_matches_host_mirror("pi", {"which_ale": "zstar"})


@pytest.mark.parametrize("name,opts", [("pi", {"Redi": True, "use_sw_pene": True}),
                                       ("band", {"smooth_bh_tra": True, "gamma_tra": (0.0002, 0.05, 3.0), "Redi": True})])
def test_emulated_redi_and_biharmonic_filter(emulated, name, opts):
    """k_redi_expl (diff_ver_part_redi_expl, the term round 1 lacked) and k_bh (diff_part_bh) against the oracle, bit for bit"""
    from tests import test_gpu_tracer as t
    t.test_redi_and_shortwave_match_oracle(name, opts)
