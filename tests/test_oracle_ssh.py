"""CPU tests of the SSH-solve oracle: the restated psolve.c row scaling + bicgstab_ras.c Krylov loop (oracle/ssh.c)
against (a) the reference's OWN psolve.c + pARMS (oracle/_ref, built from /root/reference in the authoring
container) and (b) the committed golden solution vectors that build produced (tests/golden/ssh_*.npz, made by
tests/golden/make_ssh_golden.py), which is what pins the oracle on the GPU box where /root/reference is absent.
Tolerance: both solvers stop at ||D^-1(b - A x)||_2 < 1e-10 (bicgstab_ras.c:78,146,220), so solutions agree to
cond(D^-1 A) * 1e-10 / ||x||; measured <= 3e-8 relative on these matrices (soufflet is the worst), asserted at
1e-6."""
import os

import numpy as np
import pytest

from oracle import binding as ob
from tests import util

GOLD = util.GOLD


def system(name):
    hm = util.build_model(name)
    hm.compute_ssh_rhs()
    rp = hm["ssh_rowptr"]
    n = hm.ipar("myDim_nod2D")
    return hm, (rp - rp[0]).astype(np.int32), (hm["ssh_colind"] - 1).astype(np.int32), hm["ssh_values"].copy(), hm["ssh_rhs"][:n].copy()


def scaled_residual(rp, ci, vals, x, b):
    import scipy.sparse as sp
    n = rp.size - 1
    M = sp.csr_matrix((vals.copy(), ci.copy(), rp.copy()), shape=(n, n))      # scipy may sort indices in place
    sc = 1.0 / np.asarray(abs(M).sum(axis=1)).ravel()
    return np.linalg.norm(sc * (b - M @ x))


@pytest.mark.parametrize("name", ["pi", "soufflet", "band"])
@pytest.mark.parametrize("block", [1, 4])
def test_oracle_bicgstab_converges(name, block):
    hm, rp, ci, vals, b = system(name)
    x, its, res, rc = ob.ssh_solve(rp, ci, vals, b, np.zeros_like(b), pc_kind=1, block=block)
    assert rc == 0 and res < 1e-10 and its < 500
    assert scaled_residual(rp, ci, vals, x, b) < 2e-10


@pytest.mark.parametrize("name", ["pi", "soufflet"])
def test_oracle_matches_golden_reference_solution(name):
    """golden x = output of the reference's psolve (pARMS BiCGStab + RAS + ILU(2)) on this system"""
    g = np.load(os.path.join(GOLD, "ssh_%s.npz" % name))
    hm, rp, ci, vals, b = system(name)
    assert np.array_equal(rp, g["rowptr0"]) and np.array_equal(ci, g["colind0"])
    assert util.rel_err(vals, g["values"]) < 1e-14 and util.rel_err(b, g["rhs"]) < 1e-14
    x, its, res, rc = ob.ssh_solve(rp, ci, vals, b, np.zeros_like(b), pc_kind=1, block=1)
    assert util.rel_err(x, g["x_ref"]) < 1e-6
    # warm start from the reference's solution: already converged, zero iterations (bicgstab_ras.c:146)
    x2, its2, _, _ = ob.ssh_solve(rp, ci, vals, b, g["x_ref"], pc_kind=1, block=1)
    assert its2 <= 1 and util.rel_err(x2, g["x_ref"]) < 1e-12


@pytest.mark.parametrize("name", ["pi", "soufflet", "band"])
def test_oracle_matches_live_reference_build(name):
    if ob.ref_lib() is None:
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    hm, rp, ci, vals, b = system(name)
    ref = ob.RefSolver(rp, ci, vals)
    x_ref = ref.solve(b, vals, np.zeros_like(b), new_values=0)
    assert scaled_residual(rp, ci, vals, x_ref, b) < 1e-10
    x, its, res, rc = ob.ssh_solve(rp, ci, vals, b, np.zeros_like(b), pc_kind=1, block=1)
    assert util.rel_err(x, x_ref) < 1e-6
    # value refresh with a kept preconditioner (psolve.c:177-200, new=1): perturb the matrix as update_stiff_mat_ale does
    hm["dhe"][:] = 0.05 * np.sin(np.arange(hm["dhe"].size))
    hm.stiff_mat_update()
    vals2 = hm["ssh_values"].copy()
    x_ref2 = ref.solve(b, vals2, x_ref, new_values=1)
    x2, _, _, _ = ob.ssh_solve(rp, ci, vals2, b, x, pc_kind=1, block=1)
    assert scaled_residual(rp, ci, vals2, x_ref2, b) < 1e-10
    assert util.rel_err(x2, x_ref2) < 1e-6
