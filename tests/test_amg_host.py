"""CPU tests of the multigrid SETUP of the SSH preconditioner (fesom2_b200/csrc/ssh_amg.cu, host part — no CUDA call):
the hierarchy the library builds is read back through fb200_amg_host_* and used here in a numpy restatement of what the
device does with it (bicgstab_ras.c's right-preconditioned BiCGStab on the row-scaled system of psolve.c:58-65, one
V(1,1) damped-Jacobi cycle as the preconditioner).  Checks: Galerkin identity A_{l+1} = R A P, R = P^T, partition of
unity of the prolongator on interior rows, convergence to the reference's tolerance in far fewer iterations than
block-Jacobi, and the solution against the golden vectors of the reference's own psolve.c + pARMS."""
import os

import numpy as np
import pytest
import scipy.sparse as sp

from fesom2_b200 import api
from tests import util
from tests.test_oracle_ssh import scaled_residual, system


def vcycle(H, lev, b, omega=0.8):
    A = H[lev]["A"]
    dinv = 1.0 / A.diagonal()
    x = omega * dinv * b
    if "P" not in H[lev]:
        return x
    r = b - A @ x
    x = x + H[lev]["P"] @ vcycle(H, lev + 1, H[lev]["R"] @ r, omega)
    return x + omega * dinv * (b - A @ x)


def bicgstab(As, y, M, tol=1e-10, maxit=500):
    """bicgstab_ras.c:116-230 with a zero initial guess"""
    x = np.zeros_like(y)
    r = y.copy()
    r0 = y.copy()
    p = r.copy()
    omega, rho = r @ r, r0 @ r
    its = 0
    if omega < tol * tol:
        return x, 0, np.sqrt(omega)
    pt = M(p)
    while its < maxit:
        v = As @ pt
        alpha = rho / (r0 @ v)
        r = r - alpha * v
        st = M(r)
        t = As @ st
        sigma = (t @ r) / (t @ t)
        r = r - sigma * t
        x = x + alpha * pt + sigma * st
        omega = r @ r
        its += 1
        if omega < tol * tol:
            break
        rho_n = r0 @ r
        beta = (rho_n * alpha) / (rho * sigma)
        rho = rho_n
        p = r + beta * (p - sigma * v)
        pt = M(p)
    return x, its, np.sqrt(omega)


def solve_with_hierarchy(rp, ci, vals, b, **kw):
    n = rp.size - 1
    A = sp.csr_matrix((vals.copy(), ci.copy(), rp.copy()), shape=(n, n))
    D = np.asarray(abs(A).sum(axis=1)).ravel()
    H = api.amg_host_hierarchy(rp, ci, vals, **kw)
    As = sp.diags(1.0 / D) @ A
    x, its, res = bicgstab(As, b / D, lambda v: vcycle(H, 0, D * v))
    return H, A, x, its, res


@pytest.mark.parametrize("name", ["pi", "soufflet", "band"])
def test_hierarchy_is_galerkin(name):
    hm, rp, ci, vals, b = system(name)
    H = api.amg_host_hierarchy(rp, ci, vals, min_rows=50)
    n = rp.size - 1
    A = sp.csr_matrix((vals.copy(), ci.copy(), rp.copy()), shape=(n, n))
    assert abs(H[0]["A"] - A).max() == 0.0
    assert len(H) >= 2
    for lev in range(len(H) - 1):
        P, R, Af, Ac = H[lev]["P"], H[lev]["R"], H[lev]["A"], H[lev + 1]["A"]
        assert abs(R - P.T).max() == 0.0
        G = (R @ Af @ P).tocsr()
        assert abs(G - Ac).max() <= 1e-12 * abs(Ac).max()
        assert Ac.shape[0] < 0.8 * Af.shape[0]
        assert abs(Ac - Ac.T).max() <= 1e-10 * abs(Ac).max()           # symmetric like the fine matrix
        assert (Ac.diagonal() > 0).all()
    assert "P" not in H[-1]


@pytest.mark.parametrize("name", ["pi", "soufflet"])
def test_multigrid_bicgstab_matches_reference_golden(name):
    g = np.load(os.path.join(util.GOLD, "ssh_%s.npz" % name))
    hm, rp, ci, vals, b = system(name)
    H, A, x, its, res = solve_with_hierarchy(rp, ci, vals, b, min_rows=50)
    assert res < 1e-10 and scaled_residual(rp, ci, vals, x, b) < 2e-10
    assert util.rel_err(x, g["x_ref"]) < 1e-6
    from oracle import binding as ob
    _, its_bj, _, _ = ob.ssh_solve(rp, ci, vals, b, np.zeros_like(b), pc_kind=1, block=1)
    assert its < its_bj, (its, its_bj)


def test_multigrid_iterations_are_mesh_independent():
    """the point of the multigrid: the iteration count stays flat while block-Jacobi's grows with the mesh"""
    counts = []
    for nlon in (96, 192, 384):
        hm = util.build_model("band", nlon=nlon, nlat=nlon // 2, nl=8, dt=600.0)
        hm.compute_ssh_rhs()
        rp = hm["ssh_rowptr"]
        n = hm.ipar("myDim_nod2D")
        rp0, ci0 = (rp - rp[0]).astype(np.int32), (hm["ssh_colind"] - 1).astype(np.int32)
        H, A, x, its, res = solve_with_hierarchy(rp0, ci0, hm["ssh_values"].copy(), hm["ssh_rhs"][:n].copy())
        assert res < 1e-10
        counts.append(its)
    assert max(counts) <= 20 and counts[-1] <= counts[0] + 6, counts


def folded_cycle(H, lev, b):
    """the cycle as the device applies it by default (ssh_amg.cuh amg_fold_cycle): b_c = M b down, x = T b + Q x_c up; the coarsest
    level's damped-Jacobi step sits inside the Q of the level above; a small last level is collapsed into x = C b (C kept as T)"""
    if "M" not in H[lev]:
        return H[lev]["T"] @ b
    bc = H[lev]["M"] @ b
    xc = bc if "T" not in H[lev + 1] else folded_cycle(H, lev + 1, bc)
    return H[lev]["T"] @ b + H[lev]["Q"] @ xc


@pytest.mark.parametrize("name", ["pi", "band", "farclike"])
def test_folded_cycle_is_the_same_operator(name):
    """M = R (I - A W), T = (I + S) W, Q = S P (W = omega D^-1, S = I - W A) multiplied out on the host reproduce the
    step-by-step V(1,1) cycle to rounding, with fewer stored entries than 2 A + P + R on the fine level"""
    hm, rp, ci, vals, b = system(name)
    H = api.amg_host_hierarchy(rp, ci, vals, min_rows=50)
    assert len(H) >= 2 and "T" in H[0] and "T" not in H[-1]
    assert "M" not in H[-2] and "Q" not in H[-2]                   # the last level above the coarsest is collapsed (all test meshes are small)
    rng = np.random.default_rng(5)
    for _ in range(3):
        v = rng.standard_normal(b.size)
        ref = vcycle(H, 0, v)
        got = folded_cycle(H, 0, v)
        assert np.abs(got - ref).max() <= 1e-12 * np.abs(ref).max()
    for lev in range(len(H) - 1):
        h = H[lev]
        for nm in "MTQ":
            if nm in h:
                assert (h[nm].data != 0.0).all()                   # exact zeros are not stored
        assert h["T"].shape == h["A"].shape
        if "M" in h:
            assert h["M"].shape == h["R"].shape and h["Q"].shape == h["P"].shape
    if "M" in H[0]:
        old = 2 * H[0]["A"].nnz + H[0]["P"].nnz + H[0]["R"].nnz
        new = H[0]["M"].nnz + H[0]["T"].nnz + H[0]["Q"].nnz
        assert new < 1.05 * old, (new, old)
