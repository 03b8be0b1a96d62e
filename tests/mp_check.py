"""Multi-rank parity check, launched by tests/test_gpu_multi.py (and by hand) as
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 --master-port P tests/mp_check.py [mesh] [--ale]
One rank per GPU.  Every rank builds the same global model, partitions it N ways (src/gen_comm.F90 semantics), keeps its
own partition, runs the tracer step and the SSH solve through the C ABI with NCCL halo exchange / all-reduce, and
compares its own nodes with the oracle on the UNDIVIDED mesh: tracers bit for bit, SSH to the solver tolerance.
--ale adds the device-side ALE producers (fb200_ale_step_begin / _end with their halo exchanges) against the undivided host mirror."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def check_device_ale(name, rank, world, part_in):
    """the device-side ALE producers (fb200_ale_step_begin / _end) on this rank's partition, with the reference's four
    exchanges over NCCL, against the host-side mirror on the UNDIVIDED mesh: three coupled steps, own and halo entries bit
    for bit; the tracers stepped on the device-produced state against the oracle on the undivided mesh."""
    from fesom2_b200 import api
    from harness import partition as P
    from oracle import binding as ob
    from tests import util
    kw = dict(umax=0.05, water_flux=0.0 if name == "soufflet" else 2e-7)
    wsplit = "--wsplit" in sys.argv                           # Courant numbers above w_max_cfl: Wvel_i != 0, Wvel_e / Wvel_i exchanged too
    if wsplit:
        kw.update(w_split=True, w_max_cfl=0.2, umax=0.5)
    A = util.build_model(name, **kw)                          # undivided: host mirror + oracle
    G = util.build_model(name, **kw)                          # the same state, partitioned
    part = part_in if part_in is not None else P.contiguous_partition(G, world)
    r = P.build_partition(G, part, world)[rank]
    lm = P.local_model(G, r, part)
    nod, el = r.nod - 1, r.elem[:r.myDim_elem2D] - 1
    myN = r.myDim_nod2D
    o = ob.TracerOracle(A)
    eng = api.TracerEngine(lm)
    eng.push_step_state()
    eng.push_tracers()
    eng.ale_init()
    out = np.zeros_like(lm["tr_arr"])
    linfs = A.ipar("which_ale") == 0
    bad = []
    eng.ssh_assembly_init()
    rng = np.random.default_rng(11)
    lon = A["geo_coord_nod2D"][:, 0]
    elem_all = r.elem - 1
    rpA, rpL = A["ssh_rowptr"] - 1, lm["ssh_rowptr"] - 1
    for s in range(3):
        if s > 0:
            # the SSH assembly on the device (update_stiff_mat_ale, compute_ssh_rhs_ale, update_vel) with its exchanges of
            # ssh_rhs and UV, against the undivided host mirror; UV_rhs and d_eta stand in for the momentum step and the solve
            if not linfs:
                A.stiff_mat_update()
                eng.update_stiff_mat()
                for i in range(myN):                       # row i of this partition = row nod[i] of the undivided matrix
                    if not np.array_equal(lm["ssh_values"][rpL[i]:rpL[i + 1]], A["ssh_values"][rpA[nod[i]]:rpA[nod[i] + 1]]):
                        bad.append((s, "ssh_values")); break
            uvr = 1e-3 * rng.standard_normal(A["UV_rhs"].shape)
            A["UV_rhs"][:] = uvr
            lm["UV_rhs"][:] = uvr[elem_all]
            A.compute_ssh_rhs()
            eng.compute_ssh_rhs()
            if not np.array_equal(lm["ssh_rhs"], A["ssh_rhs"][nod]): bad.append((s, "ssh_rhs"))
            d = 1e-2 * np.sin(3 * lon + s)
            A["d_eta"][:] = d
            lm["d_eta"][:] = d[nod]
            util.update_vel_host(A)
            eng.update_vel()
            if not np.array_equal(lm["UV"], A["UV"][elem_all]):
                w = np.flatnonzero((lm["UV"] != A["UV"][elem_all]).any(axis=(1, 2)))
                bad.append((s, "UV", int(w.min()), int(w.max()), int(w.size), r.myDim_elem2D, r.eDim_elem2D, r.eXDim_elem2D))
            A.compute_hbar()
            A.vert_vel()
            eng.ale_step_begin(uv_resident=True)
            if not np.array_equal(lm["hbar"], A["hbar"][nod]): bad.append((s, "hbar"))
            sro_n = slice(0, myN) if linfs else slice(None)   # linfs: the reference leaves partial sums in the halo entries
            if not np.array_equal(lm["ssh_rhs_old"][sro_n], A["ssh_rhs_old"][nod][sro_n]): bad.append((s, "ssh_rhs_old"))
            if not np.array_equal(lm["dhe"][:r.myDim_elem2D], A["dhe"][el]): bad.append((s, "dhe"))
            for nm in ("Wvel", "hnode_new") + (("Wvel_e", "Wvel_i") if wsplit else ()):
                if not np.array_equal(eng.debug(nm), A[nm][nod]): bad.append((s, nm))
            if wsplit and not np.abs(A["Wvel_i"]).max() > 0: bad.append((s, "Wvel_i is zero: the split is not exercised"))
        o.step()
        eng.solve_tracers_ale_resident(out)
        if not np.array_equal(out, A["tr_arr"][:, nod]): bad.append((s, "tr_arr"))
        A.update_thickness()
        eng.ale_step_end()
        for nm in ("hnode", "Z_3d_n", "zbar_3d_n"):
            if not np.array_equal(eng.debug(nm), A[nm][nod]): bad.append((s, nm))
        if not np.array_equal(eng.debug("helem")[:r.myDim_elem2D], A["helem"][el]): bad.append((s, "helem"))
    eng.close()
    return "ok" if not bad else repr(bad)


def run_check(name, dist, rank, world, ale=False, emulate=False, steps=2):
    """one mesh: this rank's partition through the C ABI (NCCL halo exchange / all-reduce) against the oracle on the undivided
    mesh.  Returns this rank's verdict dict; the caller gathers them.  `dist` is an initialised torch.distributed (any backend):
    it only carries the NCCL id."""
    from fesom2_b200 import api
    from harness import partition as P
    from oracle import binding as ob
    from tests import util
    hm = util.build_model(name)
    hm.compute_ssh_rhs()
    if name == "pi" and world == 2:
        part = np.load(os.path.join(util.GOLD, "pi_dist2.npz"))["part"]      # the reference's own METIS partition
        part_fixed = part
    elif P.metis_lib() is not None and "--strips" not in sys.argv:
        part = P.metis_partition(hm, world)                                  # the reference's partitioner on this mesh
        part_fixed = part
    else:
        part = P.contiguous_partition(hm, world)
        part_fixed = None
    ranks = P.build_partition(hm, part, world)
    r = ranks[rank]
    lm = P.local_model(hm, r, part)
    api.comm_init(dist)
    eng = api.TracerEngine(lm)
    eng.push_step_state()
    tr, tro = lm["tr_arr"].copy(), lm["tr_arr_old"].copy()
    eng.push_tracers(tr, tro)
    for _ in range(steps):
        eng.solve()
    eng.pull_tracers(tr, tro)
    eng.close()
    ale_v = check_device_ale(name, rank, world, part_fixed) if ale else "ok"
    # oracle on the undivided mesh
    o = ob.TracerOracle(hm)
    for _ in range(steps):
        o.step()
    own = r.nod[:r.myDim_nod2D] - 1
    alln = r.nod - 1
    # SSH solve on the partitioned system
    sol = np.zeros_like(lm["d_eta"])
    s = api.SSHSolver(lm, pc="bj")
    s.solve(new_values=1, sol=sol)
    its, res = s.info()
    # the same solve with the multigrid preconditioner
    sol_mg = np.zeros_like(lm["d_eta"])
    s2 = api.SSHSolver(lm, ident=2, pc="amg", amg=(0.08, 0.8, 4, 50))
    s2.solve(new_values=1, sol=sol_mg)
    its_mg, res_mg = s2.info()
    rp = hm["ssh_rowptr"]
    n = hm.ipar("myDim_nod2D")
    x_ref, its_ref, res_ref, _ = ob.ssh_solve(rp - rp[0], hm["ssh_colind"] - 1, hm["ssh_values"], hm["ssh_rhs"][:n], np.zeros(n), 1, 1)
    ok_own = bool(np.array_equal(tr[:, :r.myDim_nod2D], hm["tr_arr"][:, own]) and np.array_equal(tro[:, :r.myDim_nod2D], hm["tr_arr_old"][:, own]))
    ok_halo = bool(np.array_equal(tr, hm["tr_arr"][:, alln]))
    err_t = float(np.abs(tr[:, :r.myDim_nod2D] - hm["tr_arr"][:, own]).max())
    err_s = float(util.rel_err(sol[:r.myDim_nod2D], x_ref[own]))
    err_mg = float(util.rel_err(sol_mg[:r.myDim_nod2D], x_ref[own]))
    verdict = dict(rank=rank, mesh=name, own_bitexact=ok_own, halo_bitexact=ok_halo, tracer_maxabs=err_t, ssh_rel=err_s, its=its, its_ref=its_ref,
                   res=res, ssh_rel_mg=err_mg, its_mg=its_mg, res_mg=res_mg, ale=ale_v, launches=int(api.lib().fb200_kernel_launches()))
    api.lib().psolver_final()
    api.comm_finalize()
    return verdict


def verdict_ok(v):
    return bool(v["own_bitexact"] and v["halo_bitexact"] and v["ssh_rel"] < 1e-6 and v["res"] < 1e-10
                and v["ssh_rel_mg"] < 1e-6 and v["res_mg"] < 1e-10 and v["ale"] == "ok")


def main():
    import torch
    import torch.distributed as dist
    from fesom2_b200 import api
    names = [a for a in sys.argv[1:] if not a.startswith("--")]
    name = names[0] if names else "pi"
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    emulate = "--emulate" in sys.argv                         # CPU: the kernels' source through tests/emu, NCCL over files
    if emulate:
        from tests import emu
        api._lib = emu.load()
    else:
        torch.cuda.set_device(local)
    dist.init_process_group("gloo")                           # only used to broadcast the NCCL id and gather verdicts
    verdict = run_check(name, dist, rank, world, ale="--ale" in sys.argv, emulate=emulate)
    allv = [None] * world
    dist.all_gather_object(allv, verdict)
    if rank == 0:
        for v in allv:
            print("MPCHECK", v)
        print("MPCHECK_RESULT", "PASS" if all(verdict_ok(v) for v in allv) else "FAIL")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
