#!/usr/bin/env python
"""bench.py — FESOM2 tracer FCT step + SSH solve on B200 (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--nodes M] [--nl L]

One "step" = one pass of the hot path over the synthetic ocean: psolve (SSH elliptic solve, zero initial guess,
matrix values refreshed) followed by solve_tracers_ale for all tracers (AB2, gradients, FCT advection, horizontal
diffusion, implicit vertical diffusion, salinity clamp).  Prints ONE JSON line.

  value      tracer node-level updates/s = wet (node, layer) pairs x num_tracers x K / device time of the K tracer
             steps, state resident in HBM (CUDA events on the launching stream, max over ranks)
  ssh        SSH solve ms/step (device time of psolve incl. row scaling; matrix values, right-hand side and solution in HBM),
             iterations, final scaled residual
  e2e        the same two calls through the C ABI with HOST (pinned) buffers: fb200_push_state_early/_late +
             psolve + fb200_solve_tracers_ale_resident, host<->device copies inside the timed region
  roofline   dominant kernel: algorithmic bytes per launch / measured launch duration vs MEASURED_PEAKS.json
  cpu_baseline / --impl reference : the oracle (C restatement, tracer step) and the reference's own
             psolve.c + pARMS (oracle/_ref, SSH solve) on the host cores, on a bounded sample of the workload
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "tracer node-level updates/s (FCT adv+diff)"
UNIT = "node-level updates/s"

# Algorithmic HBM bytes per wet (node, layer) of ONE launch of each kernel: every array the
# kernel must read or write counted once (perfect reuse inside the kernel), 2 elements and 3 edges per node.
ALGO_BYTES = {
    "k_ab2": 24, "k_grad_elem": 40, "k_lo": 118, "k_ho_edge": 144, "k_ho_edge_fb": 3, "k_fct_bounds": 72, "k_update": 120, "k_diff_edge": 88,
    "copy_told": 16, "k_impl_diff": 48, "k_clamp_salt": 16, "k_vert_impl": 40, "k_grad_z": 24,
}
# launches of the column kernels that carry two tracers are profiled under "<name>_x2": twice the per-tracer fields, the
# tracer-independent ones (volume flux inputs UV and helem, W, area, level depths, thicknesses, areasvol) once
ALGO_BYTES.update({"k_ho_edge_x2": 2 * 144 - 48, "k_diff_edge_x2": 2 * 88 - 32, "k_lo_x2": 2 * 118 - 56, "k_grad_elem_x2": 80,
                   "k_fct_bounds_x2": 2 * 72 - 8, "k_ab2_x2": 48, "copy_told_x2": 32, "k_update_x2": 2 * 120 - 24, "k_redi_expl": 64, "k_bh": 48,
                   "k_impl_diff_x2": 2 * 48 - 16})          # Kv and the level depths once for both tracers
# SSH kernels: bytes per matrix row per launch (7-wide SELL rows: 7 x (8 B value + 4 B column) + vectors); the
# block-Jacobi kernels also read their bs x bs inverse once per apply = bs x 8 B per row (added in algo_bytes_row)
ALGO_BYTES_ROW = {
    "k_spmv_alpha": 7 * 12 + 8 + 8 + 8, "k_spmv_sigma": 7 * 12 + 8 + 8 + 8, "k_r_update_pc": 8 * 4, "k_x_update": 8 * 8,
    "k_p_update_pc": 8 * 5, "k_scale_rows": 7 * 8 * 2 + 8 + 8, "k_init_res": 7 * 12 + 8 * 6, "k_pc": 16,
}
PC_KERNELS = ("k_r_update_pc", "k_p_update_pc", "k_pc")


def ncu_traffic(name, units):
    """DRAM bytes per launch of `name` at this run's size, from the committed ncu capture (profiles/r02_ncu_traffic.json:
    dram__bytes_read.sum + dram__bytes_write.sum per unit at the captured size x this run's units); None if absent"""
    name = {"k_ho_edge_x2": "k_edge_adv2", "k_diff_edge_x2": "k_edge_diff2", "k_lo_x2": "k_lo2", "k_update": "k_update2", "k_fct_bounds": "k_fct_bounds2",
            "k_ab2": "k_ab2_g", "k_ab2_x2": "k_ab2_g", "k_impl_diff_x2": "k_impl_diff", "k_ho_edge": "k_edge_adv2", "k_diff_edge": "k_edge_diff2", "k_lo": "k_lo2",
            "k_update_x2": "k_update2", "k_fct_bounds_x2": "k_fct_bounds2"}.get(name, name)
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "r02_ncu_traffic.json")))
        return t[name]["dram_bytes_per_unit"] * units
    except (OSError, ValueError, KeyError):
        return None


def algo_bytes_row(name, block):
    b = ALGO_BYTES_ROW.get(name, 0)
    return b + 8 * block if name in PC_KERNELS else b


def amg_bytes(levels, folded=False):
    """Algorithmic bytes per launch of the multigrid kernels from the hierarchy's sizes [(rows, A, P, R stored SELL entries)]:
    every stored matrix entry = 8 B value + 4 B column, every vector a kernel reads or writes counted once.  The kernels
    of the coarse levels share one profile name (k_amg_coarse): its figure is the mean over the coarse launches of a cycle."""
    if not levels:
        return {}
    out = {"k_amg_enter": (52 if folded else 60) * levels[0][0]}
    coarse, ncoarse = 0, 0
    for l in range(len(levels) - 1):
        n, a, p, r = levels[l]
        nc = levels[l + 1][0]
        if folded:      # (a, p, r) = stored entries of T, Q, M: down b_c = M b, up x = T b + Q x_c
            per = {"down": 12 * r + 8 * n + 8 * nc, "up": 12 * (a + p) + 16 * n + 8 * nc}
        else:
            per = {"resid": 12 * a + 24 * n, "restrict": 12 * r + 8 * n + 24 * nc, "prolong": 12 * p + 16 * n + 8 * nc, "post": 12 * a + 32 * n}
        for k, v in per.items():
            if l == 0:
                out["k_amg_%s0" % k] = v
            else:
                coarse += v
                ncoarse += 1
    if ncoarse:
        out["k_amg_coarse"] = coarse / ncoarse
    return out


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--nodes", type=int, default=0, help="target surface nodes of the global mesh (default: the --config's size)")
    ap.add_argument("--nl", type=int, default=0, help="number of levels (layers = nl-1; default: the --config's)")
    ap.add_argument("--cpu-nodes", type=int, default=120_000, help="surface nodes of the CPU-baseline sample")
    ap.add_argument("--cpu-cores", type=int, default=0, help="processes of the CPU baseline (0 = all host cores)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--block", type=int, default=16, help="block-Jacobi block size of the SSH preconditioner (--pc bj)")
    ap.add_argument("--e2e-host-state", action="store_true", help="also time the end-to-end step in which the HOST produces Wvel, hnode_new, "
                    "helem and the level depths every step (fb200_push_state_late) instead of the device-side ALE producers")
    ap.add_argument("--partition", default="metis", choices=["metis", "strips"], help="N > 1: the reference's own partitioner (src/fort_part.c + "
                    "METIS 5.1.0, oracle/_ref/libmetis_part.so) or contiguous strips of the node numbering")
    ap.add_argument("--no-parity", action="store_true", help="N > 1: skip the partitioned-vs-undivided parity check before the timed region")
    ap.add_argument("--cpu-steps", type=int, default=5, help="timed steps of the cpu_baseline leg")
    ap.add_argument("--pc", default="amg", choices=["amg", "bj"], help="SSH preconditioner: aggregation multigrid or block-Jacobi")
    ap.add_argument("--config", default="sphere5m", choices=["sphere5m", "core2like", "farclike"],
                    help="BASELINE.json configs: 5 (default, ~5 M-node band sphere x 48 layers), 3 (CORE2's size: ~127 k nodes x 47 layers, one GPU) "
                         "or 4 (farc's size and character: ~640 k nodes x 47 layers, rows 3x denser north of 65 N; METIS partitions for N > 1); "
                         "--nodes / --nl override the sizes")
    ap.add_argument("--shuffle", action="store_true", help="random node and element numbering (no locality at all: the worst case for the gathers; "
                                                           "the default numbering is row-major)")
    ap.add_argument("--rcm", action="store_true", help="the randomly numbered mesh after an offline reverse Cuthill-McKee renumbering of its node graph "
                                                       "(elements sorted by their smallest node): what INTEGRATION.md recommends for badly numbered meshes")
    a = ap.parse_args()
    dn, dl = {"sphere5m": (5_000_000, 49), "core2like": (127_000, 48), "farclike": (640_000, 48)}[a.config]
    a.nodes = a.nodes or dn
    a.nl = a.nl or dl
    return a


def mem_available_gb():
    try:
        for line in open("/proc/meminfo"):
            if line.startswith("MemAvailable"):
                return int(line.split()[1]) / 1e6
    except OSError:
        pass
    return 0.0


def band_dims(target_nodes):
    """nlon x nlat (= nlon/2) lon/lat band whose wet node count is ~target (about 8.5 % becomes land)"""
    nlon = int(round((2.0 * target_nodes / 0.915) ** 0.5))
    nlon += nlon % 2
    return nlon, nlon // 2


MESH_KW = {}          # set by main() from --config / --shuffle: extra arguments of the synthetic mesh (harness/synth.py lonlat_band_mesh)


def build_workload(target_nodes, nl, seed=12345):
    from tests import util
    nlon, nlat = band_dims(target_nodes)
    if MESH_KW.get("refine_north"):                               # rows get denser towards the north: 120 x 90 rows give 9750 nodes
        f = (target_nodes / 9750.0) ** 0.5
        nlon, nlat = int(round(120 * f)) // 2 * 2, int(round(90 * f))
    t0 = time.time()
    hm = util.build_model("band", nlon=nlon, nlat=nlat, nl=nl, seed=seed, dt=600.0, **MESH_KW)
    hm.compute_ssh_rhs()
    info = dict(nlon=nlon, nlat=nlat, nl=nl, nod2D=hm.ipar("myDim_nod2D"), elem2D=hm.ipar("myDim_elem2D"),
                edge2D=hm.ipar("myDim_edge2D"), setup_s=round(time.time() - t0, 1))
    info["wet_node_layers"] = int(util.wet_mask(hm).sum())
    return hm, info


class ClockSampler(threading.Thread):
    """nvidia-smi clocks + throttle reasons while the timed region runs (B200_PROFILING.md clocks line)"""

    def __init__(self, gpu_index):
        super().__init__(daemon=True)
        self.gpu = gpu_index
        self.stop_flag = False
        self.sm, self.reasons, self.max_mhz = [], set(), None

    def query(self):
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
            "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        try:
            out = subprocess.run(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + q, "--format=csv,noheader,nounits"],
                                 stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True, timeout=5).stdout.strip()
            f = [x.strip() for x in out.split(",")]
            self.sm.append(float(f[0]))
            self.max_mhz = float(f[1])
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[2:6]):
                if v.lower() == "active":
                    self.reasons.add(name)
        except Exception:
            pass

    def run(self):
        while not self.stop_flag:
            self.query()
            time.sleep(0.05)

    def result(self):
        if not self.sm:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}
        return {"sm_mhz": float(np.median(self.sm)), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.sm)}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (copy bandwidth)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md: 6.65 TB/s)"


def pin(arrays):
    """page-lock the host arrays the e2e leg copies from/to (cudaHostRegister); returns the registered list"""
    import torch
    rt = torch.cuda.cudart()
    done = []
    for a in arrays:
        if a.nbytes == 0:
            continue
        try:
            r = rt.cudaHostRegister(a.ctypes.data, a.nbytes, 0)
            if int(r) == 0:
                done.append(a)
        except Exception:
            pass
    return done


def unpin(arrays):
    import torch
    rt = torch.cuda.cudart()
    for a in arrays:
        try:
            rt.cudaHostUnregister(a.ctypes.data)
        except Exception:
            pass


STATE = ["UV", "Wvel", "hnode", "hnode_new", "helem", "Z_3d_n", "zbar_3d_n", "Kv", "Ki",
         "heat_flux", "water_flux", "virtual_salt", "relax_salt", "real_salt_flux"]


def run_b200(args):
    import torch
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from fesom2_b200 import api

    # ---- N > 1: before anything is timed, this run's partitioned device path (NCCL halo exchanges, distributed psolve, the
    # device-side ALE producers and SSH assembly with their exchanges) is compared with the UNDIVIDED oracle / host mirror on
    # two small meshes (what tests/mp_check.py does; the driver's GPU test box has a single GPU and skips those tests)
    parity = None
    if world > 1 and not args.no_parity:
        from tests import mp_check
        parity = {"meshes": [], "ok": True}
        for name in ("band", "pi"):
            v = mp_check.run_check(name, dist, rank, world, ale=True)
            allv = [None] * world
            dist.all_gather_object(allv, v)
            ok = all(mp_check.verdict_ok(x) for x in allv)
            parity["ok"] = parity["ok"] and ok
            parity["meshes"].append({"mesh": name, "ok": ok, "tracer_bitexact": all(x["own_bitexact"] and x["halo_bitexact"] for x in allv),
                                     "ssh_rel": max(max(x["ssh_rel"], x["ssh_rel_mg"]) for x in allv), "ale": all(x["ale"] == "ok" for x in allv),
                                     "iterations_bj": allv[0]["its"], "iterations_mg": allv[0]["its_mg"],
                                     "detail": [x["ale"] for x in allv if x["ale"] != "ok"][:2]})

    # ---- workload (host side = what the Fortran driver owns).  N > 1: the SAME global mesh split into N partitions
    # (strong scaling, BASELINE.json config 5): rank 0 builds + partitions it and ships every rank its partition.
    target = args.nodes
    need_gb = 19e-6 * target * (args.nl / 49.0)                    # measured ~18 kB of host state per surface node
    avail = mem_available_gb()
    reduced = False
    if avail and need_gb * 1.25 > avail:
        target = int(target * avail / (need_gb * 1.25))
        reduced = True
    from tests import util
    if world == 1:
        hm, info = build_workload(target, args.nl, seed=12345)
    else:
        from harness import partition as P
        gloo = dist.new_group(backend="gloo")
        t0 = time.time()
        ghm, info = (build_workload(target, args.nl, seed=12345) if rank == 0 else (None, None))
        box = [info]
        dist.broadcast_object_list(box, src=0, group=gloo)
        info = dict(box[0])
        part = None
        part_kind = [args.partition]
        if rank == 0:
            if args.partition == "metis" and P.metis_lib() is not None:
                tp = time.time()
                part = P.metis_partition(ghm, world)
                part_kind = ["METIS 5.1.0 through the reference's fort_part.c (weighted by 2D + 3D node counts, check_partitioning applied), %.0f s" % (time.time() - tp)]
            else:
                part = P.contiguous_partition(ghm, world)
                part_kind = ["contiguous strips of the node numbering" + ("" if args.partition == "strips" else " (oracle/_ref/libmetis_part.so not built)")]
        dist.broadcast_object_list(part_kind, src=0, group=gloo)
        info["partition"] = part_kind[0]
        hm = P.scatter_local_models(ghm, part, dist, group=gloo)
        del ghm
        own = hm.ipar("myDim_nod2D")
        info["wet_node_layers"] = int(util.wet_mask(hm)[:own].sum())
        info["nod2D_global"] = info["nod2D"]
        info["nod2D"] = own
        info["halo_nodes"] = hm.ipar("eDim_nod2D")
        info["setup_s"] = round(time.time() - t0, 1)
        api.comm_init(dist)
    ntr = hm.ipar("num_tracers")
    n_own = hm.ipar("myDim_nod2D")
    units_per_step = info["wet_node_layers"] * ntr

    stream = torch.cuda.current_stream()
    api.set_stream(stream.cuda_stream)
    eng = api.TracerEngine(hm)
    solver = api.SSHSolver(hm, block=args.block, pc=args.pc)
    amg_levels = solver.amg_levels()
    amg_folded = bool(api.lib().fb200_psolve_amg_folded(api._i(solver.ident))) if amg_levels else False
    eng.push_step_state()
    eng.push_tracers()
    L = api.lib()
    sol = np.zeros_like(hm["d_eta"])

    # the kernel-only legs take the SSH system from HBM like every other input of the timed region (psolve accepts device pointers);
    # the e2e leg below passes the host's arrays
    d_vals = torch.from_numpy(np.ascontiguousarray(hm.arr("ssh_values"))).cuda()
    d_rhs = torch.from_numpy(np.ascontiguousarray(hm.arr("ssh_rhs"))).cuda()
    d_sol = torch.zeros(sol.size, dtype=torch.float64, device="cuda")

    def one_step(ev=None):
        d_sol.zero_()
        if ev:
            ev[0].record(stream)
        solver.solve_ptr(d_rhs.data_ptr(), d_vals.data_ptr(), d_sol.data_ptr(), new_values=1)
        if ev:
            ev[1].record(stream)
        eng.solve()
        if ev:
            ev[2].record(stream)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    # psolve takes HOST arrays by definition (psolve.c:152): page-lock them like the Fortran host would (cudaHostRegister)
    ssh_pinned = pin([hm.arr("ssh_values"), hm.arr("ssh_rhs"), sol])
    for _ in range(args.warmup):
        one_step()
    launches0 = L.fb200_kernel_launches()
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(args.steps)]
    sampler = ClockSampler(local)
    barrier()
    sampler.start()
    t_wall = time.time()
    for k in range(args.steps):
        one_step(evs[k])
    barrier()
    t_wall = time.time() - t_wall
    sampler.stop_flag = True
    sampler.join()
    launches = L.fb200_kernel_launches() - launches0
    ssh_ms = sum(e[0].elapsed_time(e[1]) for e in evs)
    tra_ms = sum(e[1].elapsed_time(e[2]) for e in evs)
    tot_ms = evs[0][0].elapsed_time(evs[-1][2])
    its, res = solver.info()

    # ---- per-kernel timing pass (untimed for the headline; CUDA events around every launch)
    api.profile(2)
    for _ in range(2):
        one_step()
    torch.cuda.synchronize()
    api.profile(0)
    prof = api.profile_report()

    # ---- e2e: host buffers through the C ABI, copies inside the timed region.  The headline path keeps the ALE producers on the
    # device (fb200_ale_step_begin/_end: compute_hbar_ale, vert_vel_ale, update_thickness_ale): per step the host sends UV
    # (after update_vel), Kv, the surface fluxes and the SSH system, and receives tr_arr, hbar, ssh_rhs_old, dhe and d_eta.
    e2e = None
    if not args.no_e2e:
        host_arrays = [hm.arr(nm) for nm in STATE] + [hm.arr("tr_arr"), hm.arr("tr_arr_old"), hm.arr("hbar"), hm.arr("ssh_rhs_old"), hm.arr("dhe")]
        pinned = pin(host_arrays) + ssh_pinned
        tr_h = hm.arr("tr_arr").copy()
        pinned += pin([tr_h])
        ksteps = max(1, min(args.steps, 5))
        small = sum(hm.arr(nm).nbytes for nm in ("heat_flux", "water_flux", "virtual_salt", "relax_salt", "real_salt_flux"))

        def timed(step_fn):
            step_fn()
            barrier()
            a_, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a_.record(stream)
            for _ in range(ksteps):
                step_fn()
            b_.record(stream)
            barrier()
            ms = a_.elapsed_time(b_) / ksteps
            if dist is not None:                                    # max over ranks
                t_ = torch.tensor([ms], dtype=torch.float64, device="cuda")
                dist.all_reduce(t_, op=dist.ReduceOp.MAX)
                ms = float(t_[0])
            return ms

        host_state = None
        if args.e2e_host_state:                                     # first: the device-side producers are not initialised yet
            # the same step with the host producing Wvel, hnode_new, helem and the level depths: all of STATE but the old
            # thickness (the resident hnode_new of the previous step) and Ki (static without Fer_GM / Redi) crosses PCIe
            eng.push_step_state()
            eng.push_tracers()

            def host_step():
                sol[:] = 0.0
                eng.push_state_early(hnode_resident=True, ki_resident=True)
                solver.solve(new_values=1, sol=sol)
                eng.push_state_late()                               # H2D UV, Wvel, hnode_new, helem
                eng.solve_tracers_ale_resident(tr_h)

            hs_ms = timed(host_step)
            host_state = {"ms_per_step": hs_ms, "steps": ksteps,
                          "h2d": sum(hm.arr(nm).nbytes for nm in STATE if nm not in ("hnode", "Ki")) + hm.arr("ssh_values").nbytes + 2 * 8 * n_own,
                          "d2h": hm.arr("tr_arr").nbytes + 8 * n_own}
        eng.push_step_state()
        eng.push_tracers()                                          # once: initial condition
        eng.ale_init()

        def e2e_step():
            sol[:] = 0.0
            eng.push_state_early(hnode_resident=True, ki_resident=True, depths_resident=True)    # H2D Kv + surface fluxes: enqueued only, overlaps psolve
            solver.solve(new_values=1, sol=sol)                    # H2D values, rhs, guess; D2H solution
            eng.ale_step_begin()                                    # H2D UV, water_flux; hbar / Wvel / hnode_new on the device; D2H hbar, ssh_rhs_old, dhe
            eng.solve_tracers_ale_resident(tr_h)                    # step on the resident tracers, D2H tr_arr
            eng.ale_step_end()                                      # hnode, zbar_3d_n, Z_3d_n, helem for the next step

        e2e_ms = timed(e2e_step)
        h2d = hm.arr("UV").nbytes + hm.arr("water_flux").nbytes + hm.arr("Kv").nbytes + small + hm.arr("ssh_values").nbytes + 2 * 8 * n_own
        d2h = hm.arr("tr_arr").nbytes + hm.arr("hbar").nbytes + hm.arr("ssh_rhs_old").nbytes + hm.arr("dhe").nbytes + 8 * n_own
        unpin(pinned)
        ssh_pinned = []
        e2e = {"ms_per_step": e2e_ms, "steps": ksteps, "h2d": h2d, "d2h": d2h, "pinned_arrays": len(pinned), "host_state": host_state}
    unpin(ssh_pinned)

    # ---- max over ranks
    vals = torch.tensor([tra_ms, ssh_ms, tot_ms, e2e["ms_per_step"] if e2e else 0.0, float(units_per_step)],
                        dtype=torch.float64, device="cuda")
    if dist is not None:
        mx = vals.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = vals.clone()
        dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        tra_ms, ssh_ms, tot_ms, e2e_ms = [float(x) for x in mx[:4]]
        units_all = float(sm[4])
    else:
        e2e_ms = e2e["ms_per_step"] if e2e else 0.0
        units_all = float(units_per_step)
    return dict(rank=rank, world=world, hm=hm, info=info, reduced=reduced, units_all=units_all, tra_ms=tra_ms, ssh_ms=ssh_ms,
                tot_ms=tot_ms, t_wall=t_wall, its=its, res=res, prof=prof, e2e=e2e, e2e_ms=e2e_ms, launches=launches,
                clocks=sampler.result(), ntr=ntr, dist=dist, amg_levels=amg_levels, amg_folded=amg_folded, parity=parity)


def _cpu_worker(idx, cpu_nodes, nl, steps, warmup, barrier, out, mesh_kw=None):
    """One host core = one subdomain of the reference's MPI decomposition: build a sample of the workload, then time
    the oracle's tracer step and the reference's own psolve.c + pARMS (oracle/_ref) on it, all workers stepping at the
    same time so they share the memory system like the ranks of one node would."""
    try:
        from oracle import binding as ob
        ob.use_timing_build(True)                  # copies compiled with the reference's own optimisation flags (harness/build.py)
        MESH_KW.update(mesh_kw or {})              # spawned: the parent's globals do not travel
        hm, info = build_workload(cpu_nodes, nl)
        ntr = hm.ipar("num_tracers")
        o = ob.TracerOracle(hm)
        rp = hm["ssh_rowptr"]
        n = hm.ipar("myDim_nod2D")
        rp0, ci0 = (rp - rp[0]).astype(np.int32), (hm["ssh_colind"] - 1).astype(np.int32)
        vals, rhs = hm["ssh_values"], hm["ssh_rhs"][:n].copy()
        ref = ob.RefSolver(rp0, ci0, vals) if ob.ref_lib() is not None else None
        t_tr, t_ssh = 0.0, 0.0
        barrier.wait()
        for k in range(warmup + steps):
            t0 = time.perf_counter()
            if ref is not None:
                ref.solve(rhs, vals, np.zeros(n), new_values=1)
            else:
                ob.ssh_solve(rp0, ci0, vals, rhs, np.zeros(n), pc_kind=1, block=1)
            t1 = time.perf_counter()
            o.step()
            t2 = time.perf_counter()
            if k >= warmup:
                t_ssh += t1 - t0
                t_tr += t2 - t1
        out.put(dict(idx=idx, t_tr=t_tr, t_ssh=t_ssh, units=info["wet_node_layers"] * ntr, rows=n, ntr=ntr, info=info,
                     ssh_ref=ref is not None))
    except BaseException as e:          # never leave the parent waiting
        try:
            barrier.abort()
        except Exception:
            pass
        out.put(dict(idx=idx, error=repr(e)))


def host_cores():
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def cpu_reference(args, steps, warmup, nodes_per_core=None):
    """The reference's CPU path on ALL host cores, on a bounded sample of the workload.  FESOM2 parallelises over MPI
    ranks only (one subdomain per core); neither MPI nor a Fortran compiler exists here, so every core runs one
    subdomain-sized sample (--cpu-nodes surface nodes) in its own process at the same time, with no halo exchange
    (which favours the CPU side).  Tracer step = the oracle's C restatement of solve_tracers_ale ("port"); SSH solve =
    the reference's OWN psolve.c + pARMS when oracle/_ref was built, else the oracle's BiCGStab.  value = sum of the
    per-core throughputs."""
    import multiprocessing as mp
    cores = host_cores() if args.cpu_cores <= 0 else args.cpu_cores
    cpu_nodes = args.cpu_nodes if nodes_per_core is None else nodes_per_core
    gb_per_node = 3.5 / 120_000 * args.nl / 49                                   # host memory of one sample, per surface node
    # the samples of all cores together cover as much of the GPU arm's mesh as host memory and a few minutes of run time allow
    cpu_nodes = int(max(20_000, min(cpu_nodes, mem_available_gb() * 0.6 / (cores * gb_per_node), 150.0 / ((steps + warmup) * 2.2e-5))))
    from harness import build as _b0
    _b0.build_oracle_timing()                                                    # once, before the workers load it
    _b0.build_host()
    ctx = mp.get_context("spawn")                                                # the parent may hold a CUDA context
    barrier, out = ctx.Barrier(cores), ctx.Queue()
    procs = [ctx.Process(target=_cpu_worker, args=(i, cpu_nodes, args.nl, steps, warmup, barrier, out, dict(MESH_KW))) for i in range(cores)]
    for p_ in procs:
        p_.start()
    res = [out.get() for _ in procs]
    for p_ in procs:
        p_.join()
    bad = [r for r in res if "error" in r]
    if bad:
        raise RuntimeError("cpu baseline worker failed: %s" % bad[0]["error"])
    info, ntr = res[0]["info"], res[0]["ntr"]
    value = sum(r["units"] * steps / r["t_tr"] for r in res)
    t_tr = max(r["t_tr"] for r in res)
    t_ssh = max(r["t_ssh"] for r in res)
    n = res[0]["rows"]
    ref = res[0]["ssh_ref"]
    sample = "%d processes (one per host core), each: band mesh %dx%d nl=%d, %d surface nodes, %d wet node-layers x %d tracers, %d steps; " \
             "tracer step = C restatement (oracle), SSH = %s; no halo exchange between the samples" % (
                 cores, info["nlon"], info["nlat"], args.nl, info["nod2D"], info["wet_node_layers"], ntr, steps,
                 "reference psolve.c+pARMS (oracle/_ref), 1 subdomain per process" if ref else "oracle BiCGStab")
    from harness import build as _b
    flags = {"tracer_step": "gcc " + " ".join(_b.ORACLE_TIMING_FLAGS) + " (the reference's own: src/CMakeLists.txt:73; built on this host)",
             "ssh": "gcc -O3 -finline-functions -march=x86-64-v3 (oracle/Makefile.ref, timing copy; built where /root/reference exists)"}
    return dict(value=value, unit=UNIT, cores=cores, kind="port", ssh_kind="reference" if ref else "port", sample=sample, flags=flags,
                sample_surface_nodes_total=info["nod2D"] * cores, steps_timed=steps,
                tracer_ms_per_step=1e3 * t_tr / steps, ssh_ms_per_step=1e3 * t_ssh / steps, ssh_rows=n * cores,
                ssh_us_per_row=1e6 * t_ssh / steps / (n * cores), per_core_value=value / cores), info


def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.config == "farclike":
        MESH_KW["refine_north"] = 3.0
    if args.shuffle:
        MESH_KW["shuffle"] = True
    if args.rcm:
        MESH_KW["rcm"] = True
    workload = "%s: synthetic lon/lat-band sphere (|lat|<=80, ~8%% land, ragged bathymetry%s), ~%d surface nodes in total x %d layers, " \
               "2 tracers, MFCT+QR4C+FCT, zstar, %s numbering" % (
                   {"sphere5m": "BASELINE config 5", "core2like": "BASELINE config 3 stand-in (CORE2's size)", "farclike": "BASELINE config 4 stand-in (farc's size)"}[args.config],
                   ", rows 3x denser north of 65 N" if args.config == "farclike" else "", args.nodes, args.nl - 1,
                   "random, then reverse Cuthill-McKee" if args.rcm else "random (shuffled)" if args.shuffle else "row-major")
    if args.impl == "reference":
        if rank != 0:
            return
        cb, info = cpu_reference(args, max(1, args.steps), max(0, min(args.warmup, 1)), nodes_per_core=max(20_000, args.nodes // max(1, host_cores() if args.cpu_cores <= 0 else args.cpu_cores)))
        line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": cb["tracer_ms_per_step"] + cb["ssh_ms_per_step"], "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": workload, "sample": cb["sample"]},
                "ssh": {"ms_per_step": cb["ssh_ms_per_step"], "rows": cb["ssh_rows"], "us_per_row": cb["ssh_us_per_row"]},
                "cpu_baseline": cb,
                "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        print(json.dumps(line))
        return

    r = run_b200(args)
    if r["dist"] is not None:
        r["dist"].barrier()
    if rank != 0:
        if r["dist"] is not None:
            r["dist"].destroy_process_group()
        return
    info = r["info"]
    K = args.steps
    value = r["units_all"] * K / (r["tra_ms"] * 1e-3)
    peak, peak_src = peaks()
    # dominant kernel of the step
    prof = r["prof"]
    wet = info["wet_node_layers"]
    rows = info["nod2D"]
    dom = max(prof.items(), key=lambda kv: kv[1][1]) if prof else None
    roof = None
    amg_b = amg_bytes(r["amg_levels"], r["amg_folded"])

    def launch_bytes(name):
        if name in ALGO_BYTES:
            return ALGO_BYTES[name] * wet
        if name in amg_b:
            return amg_b[name]
        return algo_bytes_row(name, args.block) * rows

    if dom:
        name, (cnt, ms) = dom
        per = ALGO_BYTES.get(name)
        bytes_launch = launch_bytes(name)
        avg_ms = ms / cnt
        ach = bytes_launch / (avg_ms * 1e-3) / 1e9
        roof = {"bound": "hbm", "kernel": name, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                "traffic": ncu_traffic(name, wet if per is not None else rows),
                "traffic_source": "profiles/r02_ncu_traffic.json (ncu --set full at 1 M nodes, DRAM bytes per unit x units of this run)",
                "peak_source": peak_src, "algorithmic_bytes_per_launch": bytes_launch,
                "avg_launch_ms": avg_ms, "share_of_step": ms / sum(v[1] for v in prof.values())}
    kern = {k: {"launches": v[0], "ms": round(v[1], 4),
                "GBps": round(launch_bytes(k) * v[0] / (v[1] * 1e-3) / 1e9, 1) if v[1] > 0 else None}
            for k, v in sorted(prof.items(), key=lambda kv: -kv[1][1])}
    step_bytes = 344.0 * wet * r["ntr"]                     # 344 B per (node, layer, tracer)
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": r["world"], "steps": K, "warmup": args.warmup,
            "ms_per_step": r["tot_ms"] / K, "higher_is_better": True, "scaling": "strong" if r["world"] > 1 else "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": workload, "mesh": {k: info[k] for k in ("nlon", "nlat", "nl", "nod2D", "elem2D", "edge2D", "wet_node_layers", "nod2D_global", "halo_nodes") if k in info},
                       "mesh_is": "global mesh" if r["world"] == 1 else "partition of rank 0", "reduced_for_host_memory": r["reduced"], "l2": "inputs larger than L2 (per-field >> 126 MB)",
                       "parallelism": "1 domain per GPU" if r["world"] == 1 else "the same global mesh in %d node partitions (%s; gen_comm.F90 halo rules), NCCL halo exchange, peer-memory or NCCL reductions" % (r["world"], info.get("partition", "?")),
                       "ssh_initial_guess": "zero every step", "setup_s": info["setup_s"],
                       "ssh_preconditioner": ("aggregation multigrid V(1,1), %s = %s" % ("folded cycle, levels (rows, T, Q, M entries)" if r["amg_folded"] else "levels (rows, A, P, R entries)", r["amg_levels"])) if args.pc == "amg"
                       else "block-Jacobi, %d rows per block" % args.block},
            "tracer_ms_per_step": r["tra_ms"] / K,
            "ssh": {"ms_per_step": r["ssh_ms"] / K, "iterations": r["its"], "scaled_residual": r["res"], "rows": rows,
                    "ms_per_iteration": r["ssh_ms"] / K / max(1, r["its"])},
            "step_roofline": {"algorithmic_GB_per_tracer_step": step_bytes / 1e9, "achieved_GBps": step_bytes / (r["tra_ms"] / K * 1e-3) / 1e9,
                              "frac": step_bytes / (r["tra_ms"] / K * 1e-3) / 1e9 / peak},
            "whole_step": {"value": r["units_all"] * K / (r["tot_ms"] * 1e-3), "unit": UNIT, "what": "the same units over psolve + tracer step: divide by the N = 1 figure x N for the parallel efficiency of the whole step"},
            "roofline": roof, "kernels": kern, "gpu_launches": int(r["launches"]), "clocks": r["clocks"],
            "wall_s_timed_region": r["t_wall"]}
    if r["e2e"]:
        line["e2e"] = {"value": r["units_all"] / (r["e2e_ms"] * 1e-3), "unit": UNIT, "ms_per_step": r["e2e_ms"],
                       "h2d_bytes_per_step": int(r["e2e"]["h2d"]), "d2h_bytes_per_step": int(r["e2e"]["d2h"]),
                       "steps": r["e2e"]["steps"], "pinned_arrays": r["e2e"]["pinned_arrays"]}
        line["e2e"]["what"] = "psolve + fb200_ale_step_begin + fb200_solve_tracers_ale_resident + fb200_ale_step_end with pinned HOST arrays: UV, Kv, surface fluxes and the SSH system in; tr_arr, hbar, ssh_rhs_old, dhe, d_eta out (compute_hbar_ale, vert_vel_ale, update_thickness_ale on the device)"
        if r["e2e"].get("host_state"):
            ea = r["e2e"]["host_state"]
            line["e2e_host_state"] = {"value": r["units_all"] / (ea["ms_per_step"] * 1e-3), "unit": UNIT, "ms_per_step": ea["ms_per_step"],
                                      "h2d_bytes_per_step": int(ea["h2d"]), "d2h_bytes_per_step": int(ea["d2h"]), "steps": ea["steps"],
                                      "what": "the host produces Wvel, hnode_new, helem and the level depths every step (fb200_push_state_late)"}
    if r["world"] == 1 and not args.no_cpu_baseline:
        cb, _ = cpu_reference(args, max(1, args.cpu_steps), 1)
        line["cpu_baseline"] = cb
    if r.get("parity") is not None:
        line["parity"] = r["parity"]
    print(json.dumps(line))
    if r["dist"] is not None:
        r["dist"].destroy_process_group()
    if r.get("parity") is not None and not r["parity"]["ok"]:
        sys.stderr.write("bench.py: the partitioned run does NOT reproduce the undivided oracle: %s\n" % (r["parity"],))
        sys.exit(3)


if __name__ == "__main__":
    main()
