"""Caller side of the C ABI (include/fesom_b200.h) for Python: what the reference's Fortran routines
`solve_tracers_ale` (src/oce_ale_tracer.F90:101) and `solve_ssh_ale` (src/oce_ale.F90:2210) do with the library,
written with ctypes.  Every call goes through the extern "C" entry points with plain pointers — the same
calls the Fortran shim in INTEGRATION.md makes.  There is no CPU fallback: loading fails loudly if the CUDA
library has not been built, and every compute call raises if the library reports an error.
"""
import ctypes as C
import os

import numpy as np

from . import build as _build

_lib = None

HOR = {"MUSCL": 0, "MFCT": 1, "UPW1": 2}
VER = {"QR4C": 0, "CDIFF": 1, "PPM": 2, "UPW1": 3}
LIM = {"FCT": 0, "NONE": 1}

EXPORTS = ["psolver_init", "psolve", "psolver_final", "fb200_psolve_info", "fb200_psolver_set_block",
           "fb200_init_dims", "fb200_init_connectivity", "fb200_init_levels", "fb200_init_geometry",
           "fb200_set_options", "fb200_set_bh_filter", "fb200_init_commit", "fb200_push_step_state", "fb200_push_state_early", "fb200_push_state_late",
           "fb200_ale_init", "fb200_ale_step_begin", "fb200_ale_step_end", "fb200_ssh_assembly_init", "fb200_update_stiff_mat", "fb200_compute_ssh_rhs", "fb200_update_vel", "fb200_push_tracers", "fb200_set_clim", "fb200_set_zonal_relax", "fb200_push_kpp_nonlocal", "fb200_tracer_extrema", "fb200_tracer_extrema_noalloc", "fb200_mean_accumulate", "fb200_mean_fetch",
           "fb200_solve_tracers", "fb200_pull_tracers", "fb200_solve_tracers_ale", "fb200_solve_tracers_ale_resident", "fb200_debug_fetch",
           "fb200_finalize", "fb200_status", "fb200_status_message", "fb200_clear_status",
           "fb200_kernel_launches", "fb200_last_ms", "fb200_set_stream", "fb200_profile", "fb200_profile_report",
           "fb200_selftest_div", "fb200_init_halo", "fb200_comm_unique_id", "fb200_comm_init", "fb200_comm_finalize", "fb200_psolver_use_graph",
           "fb200_psolver_set_pc", "fb200_psolver_set_amg", "fb200_psolve_amg_info", "fb200_psolve_amg_folded", "fb200_amg_host_build", "fb200_amg_host_levels",
           "fb200_amg_host_dims", "fb200_amg_host_fetch", "fb200_amg_host_free"]


class FesomB200Error(RuntimeError):
    pass


def lib(rebuild=False):
    global _lib
    if _lib is None or rebuild:
        so = _build.CUDA_SO
        if os.environ.get("FB200_SO"):                      # a variant build for A/B measurements (build.build_cuda(out=...))
            so = os.environ["FB200_SO"]
        elif rebuild or not os.path.exists(so) or _build.nvcc_path() is not None:
            so = _build.build_cuda()
        if not os.path.exists(so):
            raise FesomB200Error("libfesom_b200.so is missing and nvcc is unavailable: build it with __graft_entry__.build()")
        L = C.CDLL(so)
        L.fb200_status.restype = C.c_int
        L.fb200_status_message.restype = C.c_char_p
        L.fb200_kernel_launches.restype = C.c_longlong
        L.fb200_last_ms.restype = C.c_double
        L.fb200_last_ms.argtypes = [C.c_char_p]
        L.fb200_debug_fetch.argtypes = [C.c_char_p, C.c_void_p]
        L.fb200_set_stream.argtypes = [C.c_void_p]
        L.fb200_profile_report.argtypes = [C.c_char_p, C.c_void_p]
        L.fb200_profile_report.restype = C.c_int
        _lib = L
    return _lib


def _check(L):
    if L.fb200_status() != 0:
        msg = L.fb200_status_message().decode()
        code = L.fb200_status()
        L.fb200_clear_status()
        raise FesomB200Error("fesom_b200 error %d: %s" % (code, msg))


def _i(v):
    return C.byref(C.c_int(int(v)))


def _d(v):
    return C.byref(C.c_double(float(v)))


def _p(a):
    return C.c_void_p(a.ctypes.data)


def set_stream(cuda_stream_ptr):
    """run all later library work on the caller's CUDA stream (an integer cudaStream_t, e.g. torch's .cuda_stream)"""
    lib().fb200_set_stream(C.c_void_p(int(cuda_stream_ptr) if cuda_stream_ptr else None))


def selftest_div(n=1 << 24):
    """number of operand pairs on which the kernels' reciprocal-based division differs from IEEE division"""
    bad = C.c_longlong(-1)
    lib().fb200_selftest_div(C.byref(C.c_longlong(int(n))), C.byref(bad))
    _check(lib())
    return bad.value


def profile(mode):
    """0 off, 1 on, 2 on + reset"""
    lib().fb200_profile(_i(mode))


def profile_report():
    """{kernel: (launches, total_ms)} since the last reset"""
    buf = C.create_string_buffer(1 << 16)
    n = lib().fb200_profile_report(buf, C.byref(C.c_int(len(buf))))
    out = {}
    if n > 0:
        for line in buf.value.decode().splitlines():
            nm, cnt, ms = line.split()
            out[nm] = (int(cnt), float(ms))
    return out


def comm_init(dist=None, rank=None, world=None):
    """collective: NCCL communicator of the library (fb200_comm_unique_id on rank 0, broadcast, fb200_comm_init).
    `dist` is an initialised torch.distributed module (any backend) used only to broadcast the 128-byte id — the
    Fortran host does the same with MPI_Bcast."""
    L = lib()
    rank = dist.get_rank() if rank is None else rank
    world = dist.get_world_size() if world is None else world
    buf = C.create_string_buffer(128)
    if rank == 0:
        L.fb200_comm_unique_id(buf)
    obj = [bytes(buf.raw)]
    if world > 1:
        dist.broadcast_object_list(obj, src=0)
    idb = C.create_string_buffer(obj[0], 128)
    L.fb200_comm_init(_i(world), _i(rank), idb)
    _check(L)


def comm_finalize():
    lib().fb200_comm_finalize()


class TracerEngine:
    """Device-side `solve_tracers_ale`.  `hm` is a HostModel (the Fortran driver's state)."""

    def __init__(self, hm, tra_adv_hor="MFCT", tra_adv_ver="QR4C", tra_adv_lim="FCT", tra_adv_ph=1.0, tra_adv_pv=1.0,
                 w_split=None, i_vert_diff=True, Redi=False, use_sw_pene=False, do_hor_diff=True, clamp_salt=True,
                 tracer_ids=(0, 1), epsilon=0.1, smooth_bh_tra=False, gamma_tra=(0.0005, 0.0125, 0.0)):
        self.L = lib()
        self.hm = hm
        L = self.L
        L.fb200_clear_status()
        ntr = hm.ipar("num_tracers")
        L.fb200_init_dims(_i(hm.ipar("nl")), _i(hm.ipar("myDim_nod2D")), _i(hm.ipar("eDim_nod2D")), _i(hm.ipar("myDim_elem2D")),
                          _i(hm.ipar("eDim_elem2D")), _i(hm.ipar("eXDim_elem2D")), _i(hm.ipar("myDim_edge2D")), _i(ntr), _i(0))
        a = hm.arr
        L.fb200_init_connectivity(_p(a("edges")), _p(a("edge_tri")), _p(a("elem2D_nodes")), _p(a("nod_in_elem2D_num")),
                                  _p(a("nod_in_elem2D")), _i(hm.ipar("nod_in_elem2D_max")), _p(a("edge_up_dn_tri")))
        L.fb200_init_levels(_p(a("nlevels")), _p(a("ulevels")), _p(a("nlevels_nod2D")), _p(a("ulevels_nod2D")),
                            _p(a("nlevels_nod2D_min")), _p(a("ulevels_nod2D_max")), _p(a("nboundary_lay")))
        L.fb200_init_geometry(_p(a("edge_cross_dxdy")), _p(a("edge_dxdy")), _p(a("elem_cos")), _p(a("elem_area")),
                              _p(a("gradient_sca")), _p(a("area")), _p(a("areasvol")), _p(a("zbar_n_bot")))
        ids = (C.c_int * 8)(*([int(t) for t in tracer_ids] + [301] * (8 - len(tracer_ids))))
        ws = hm.ipar("w_split") if w_split is None else int(w_split)
        L.fb200_set_options(_i(HOR[tra_adv_hor]), _i(VER[tra_adv_ver]), _i(LIM[tra_adv_lim]), _d(tra_adv_ph), _d(tra_adv_pv),
                            _i(ws), _i(i_vert_diff), _i(Redi), _i(use_sw_pene), _i(do_hor_diff), _i(clamp_salt),
                            _d(hm.par("is_nonlinfs")), _d(hm.par("dt")), _d(epsilon), ids)
        L.fb200_set_bh_filter(_i(int(smooth_bh_tra)), _d(gamma_tra[0]), _d(gamma_tra[1]), _d(gamma_tra[2]))
        _check(L)
        try:
            npes = hm.ipar("npes")
        except RuntimeError:
            npes = 1
        if npes > 1:                                   # com_nod2D / com_elem2D_full of this partition
            cn = [a("com_nod2D_" + k) for k in ("rPE", "rptr", "rlist", "sPE", "sptr", "slist")]
            ce = [a("com_elem2D_full_" + k) for k in ("rPE", "rptr", "rlist", "sPE", "sptr", "slist")]
            L.fb200_init_halo(_i(hm.ipar("com_nod2D_rPEnum")), _p(cn[0]), _p(cn[1]), _p(cn[2]),
                              _i(hm.ipar("com_nod2D_sPEnum")), _p(cn[3]), _p(cn[4]), _p(cn[5]),
                              _i(hm.ipar("com_elem2D_full_rPEnum")), _p(ce[0]), _p(ce[1]), _p(ce[2]),
                              _i(hm.ipar("com_elem2D_full_sPEnum")), _p(ce[3]), _p(ce[4]), _p(ce[5]))
            _check(L)
        L.fb200_init_commit()
        _check(L)

    def push_step_state(self):
        a = self.hm.arr
        self.L.fb200_push_step_state(_p(a("UV")), _p(a("Wvel")), _p(a("Wvel_e")), _p(a("Wvel_i")), _p(a("hnode")), _p(a("hnode_new")),
                                     _p(a("helem")), _p(a("Z_3d_n")), _p(a("zbar_3d_n")), _p(a("Kv")), _p(a("Ki")),
                                     _p(a("slope_tapered")), _p(a("heat_flux")), _p(a("water_flux")), _p(a("virtual_salt")),
                                     _p(a("relax_salt")), _p(a("real_salt_flux")), _p(a("sw_3d")))
        _check(self.L)

    def push_state_early(self, hnode_resident=False, ki_resident=False, depths_resident=False):
        """the half of the step state that exists before solve_ssh_ale; returns at once, the copies overlap psolve.
        hnode_resident: pass NULL for hnode (the library's hnode_new of the previous step is the old thickness);
        ki_resident: pass NULL for Ki (unchanged since the last push)"""
        a = self.hm.arr
        null = C.c_void_p(None)
        self.L.fb200_push_state_early(null if hnode_resident else _p(a("hnode")), null if depths_resident else _p(a("Z_3d_n")),
                                      null if depths_resident else _p(a("zbar_3d_n")), _p(a("Kv")),
                                      null if ki_resident else _p(a("Ki")),
                                      _p(a("slope_tapered")), _p(a("heat_flux")), _p(a("water_flux")), _p(a("virtual_salt")),
                                      _p(a("relax_salt")), _p(a("real_salt_flux")), _p(a("sw_3d")))
        _check(self.L)

    def push_state_late(self):
        """the half the SSH solution produces (UV, Wvel*, hnode_new, helem); returns when all of the state is resident"""
        a = self.hm.arr
        self.L.fb200_push_state_late(_p(a("UV")), _p(a("Wvel")), _p(a("Wvel_e")), _p(a("Wvel_i")), _p(a("hnode_new")), _p(a("helem")))
        _check(self.L)

    def set_clim(self, clim_relax, relax2clim, Tclim, Sclim):
        """relax_to_clim inside the step: relax2clim (N), Tclim/Sclim (N, nl-1) numpy arrays, pushed once"""
        self._clim = (np.ascontiguousarray(relax2clim), np.ascontiguousarray(Tclim), np.ascontiguousarray(Sclim))
        self.L.fb200_set_clim(_d(clim_relax), _p(self._clim[0]), _p(self._clim[1]), _p(self._clim[2]))
        _check(self.L)

    def push_kpp_nonlocal(self, gT, gS, b2, b3, ref_sss=34.7, ref_sss_local=False, enable=True):
        """KPP non-local transport arrays of this step, (N, nl) each (see fb200_push_kpp_nonlocal)"""
        self.L.fb200_push_kpp_nonlocal(_i(int(enable)), _p(gT), _p(gS), _p(b2), _p(b3), _d(ref_sss), _i(int(ref_sss_local)))
        _check(self.L)

    def set_zonal_relax(self, ztem, Tclim, lat0, dy, tau_inv, enable=True):
        """relax_zonal_temp of the soufflet channel instead of relax_to_clim: ztem (nybins, nl-1), Tclim (N, nl-1)"""
        self._zonal = (np.ascontiguousarray(ztem, dtype=np.float64), np.ascontiguousarray(Tclim, dtype=np.float64),
                       np.ascontiguousarray(self.hm.arr("coord_nod2D"), dtype=np.float64))
        self.L.fb200_set_zonal_relax(_i(int(enable)), _i(self._zonal[0].shape[0]), _d(lat0), _d(dy), _d(tau_inv), _p(self._zonal[2]),
                                     _p(self._zonal[0]), _p(self._zonal[1]))
        _check(self.L)

    def ale_init(self):
        """device-side compute_hbar_ale / vert_vel_ale / update_thickness_ale: hand over the current surface elevation"""
        hm = self.hm
        self.L.fb200_ale_init(_i(hm.ipar("which_ale")), _d(hm.par("w_max_cfl")), _p(hm.arr("hbar")))
        _check(self.L)

    def ale_step_begin(self, uv_resident=False):
        """H2D UV (unless fb200_update_vel has just produced it on the device) + water_flux; hbar, ssh_rhs_old, dhe land in
        the host model's arrays"""
        a = self.hm.arr
        self.L.fb200_ale_step_begin(None if uv_resident else _p(a("UV")), _p(a("water_flux")), _p(a("hbar")), _p(a("ssh_rhs_old")), _p(a("dhe")))
        _check(self.L)

    def ssh_assembly_init(self, g=9.81):
        """device-side compute_ssh_rhs_ale / update_stiff_mat_ale / update_vel: hand over ssh_stiff and g, alpha, theta"""
        hm = self.hm
        self.L.fb200_ssh_assembly_init(_p(hm.arr("ssh_rowptr")), _p(hm.arr("ssh_colind_loc")), _p(hm.arr("ssh_values")), _p(hm.arr("dhe")),
                                       _p(hm.arr("ssh_rhs_old")), _d(g), _d(hm.par("alpha")), _d(hm.par("theta")))
        _check(self.L)

    def update_stiff_mat(self):
        """update_stiff_mat_ale from the resident dhe; the host model's ssh_values receive the result"""
        self.L.fb200_update_stiff_mat(_p(self.hm.arr("ssh_values")))
        _check(self.L)

    def compute_ssh_rhs(self):
        """H2D UV_rhs; compute_ssh_rhs_ale on the device; the host model's ssh_rhs receives the result"""
        self.L.fb200_compute_ssh_rhs(_p(self.hm.arr("UV_rhs")), _p(self.hm.arr("ssh_rhs")))
        _check(self.L)

    def update_vel(self):
        """update_vel with the host model's d_eta; the host model's UV receives the new velocity"""
        self.L.fb200_update_vel(_p(self.hm.arr("d_eta")), _p(self.hm.arr("UV")))
        _check(self.L)

    def ale_step_end(self):
        self.L.fb200_ale_step_end()
        _check(self.L)

    def extrema(self, tr_num):
        """(min, max) of the resident tracer tr_num (1-based) over own wet cells"""
        lo, hi = C.c_double(), C.c_double()
        self.L.fb200_tracer_extrema(_i(tr_num), C.byref(lo), C.byref(hi))
        _check(self.L)
        return lo.value, hi.value

    def extrema_noalloc(self, tr_num, nblocks=0):
        """the same scan through fb200_tracer_extrema_noalloc (static device scratch, no allocation); nblocks <= 0: default grid"""
        lo, hi = C.c_double(), C.c_double()
        self.L.fb200_tracer_extrema_noalloc(_i(tr_num), _i(nblocks), C.byref(lo), C.byref(hi))
        _check(self.L)
        return lo.value, hi.value

    def mean_accumulate(self, tr_num, accuracy=8):
        """update_means for the output stream of tracer tr_num (1-based), accuracy 8 (real64) or 4 (real32)"""
        self.L.fb200_mean_accumulate(_i(tr_num), _i(accuracy))
        _check(self.L)

    def mean_fetch(self, tr_num, accuracy=8):
        """(myDim_nod2D, nl-1) mean over the accumulated steps in the stream's precision; clears the accumulator"""
        hm = self.hm
        out = np.zeros((hm.ipar("myDim_nod2D"), hm.ipar("nl") - 1), dtype=np.float64 if accuracy == 8 else np.float32)
        self.L.fb200_mean_fetch(_i(tr_num), _p(out))
        _check(self.L)
        return out

    def push_tracers(self, tr=None, tro=None):
        tr = self.hm.arr("tr_arr") if tr is None else tr
        tro = self.hm.arr("tr_arr_old") if tro is None else tro
        self.L.fb200_push_tracers(_p(tr), _p(tro))
        _check(self.L)

    def solve(self):
        self.L.fb200_solve_tracers()
        _check(self.L)

    def pull_tracers(self, tr=None, tro=None):
        tr = self.hm.arr("tr_arr") if tr is None else tr
        tro = self.hm.arr("tr_arr_old") if tro is None else tro
        self.L.fb200_pull_tracers(_p(tr), _p(tro))
        _check(self.L)

    def solve_tracers_ale(self, tr=None, tro=None):
        """push + solve + pull: the drop-in for `call solve_tracers_ale(mesh)` with host-resident tracers"""
        tr = self.hm.arr("tr_arr") if tr is None else tr
        tro = self.hm.arr("tr_arr_old") if tro is None else tro
        self.L.fb200_solve_tracers_ale(_p(tr), _p(tro))
        _check(self.L)

    def solve_tracers_ale_resident(self, tr_out=None):
        """one step on the device-resident tracers; only tr_arr comes back to the host"""
        tr_out = self.hm.arr("tr_arr") if tr_out is None else tr_out
        self.L.fb200_solve_tracers_ale_resident(_p(tr_out))
        _check(self.L)

    def debug(self, name):
        hm = self.hm
        nl = hm.ipar("nl")
        N = hm.ipar("myDim_nod2D") + hm.ipar("eDim_nod2D")
        if name == "edge_perm":
            shape = (hm.ipar("myDim_edge2D"),)
        elif name == "adf_h":
            shape = (hm.ipar("myDim_edge2D"), nl - 1)
        elif name in ("tr_xy_x", "tr_xy_y"):
            shape = (hm.ipar("myDim_elem2D") + hm.ipar("eDim_elem2D") + hm.ipar("eXDim_elem2D"), nl - 1)
        elif name == "helem":
            shape = (hm.ipar("myDim_elem2D") + hm.ipar("eDim_elem2D") + hm.ipar("eXDim_elem2D"), nl - 1)
        elif name in ("adf_v", "Wvel", "Wvel_e", "Wvel_i", "zbar_3d_n"):
            shape = (N, nl)
        else:
            shape = (N, nl - 1)
        out = np.zeros(shape)
        self.L.fb200_debug_fetch(name.encode(), _p(out))
        _check(self.L)
        return out

    def last_ms(self):
        return self.L.fb200_last_ms(b"tracers")

    def close(self):
        self.L.fb200_finalize()


def amg_host_hierarchy(rp, ci, vals, theta=0.08, max_levels=4, min_rows=200):
    """The multigrid setup of the SSH preconditioner run on the host (no CUDA call) and read back as scipy CSR matrices:
    [{"A": .., "P": .., "R": ..}, ...] (P, R absent on the coarsest level).  For the CPU tests of the setup logic."""
    import scipy.sparse as sp
    L = lib()
    rp = np.ascontiguousarray(rp, dtype=np.int32)
    ci = np.ascontiguousarray(ci, dtype=np.int32)
    vals = np.ascontiguousarray(vals, dtype=np.float64)
    h = L.fb200_amg_host_build(_i(len(rp) - 1), _p(rp), _p(ci), _p(vals), _d(theta), _i(max_levels), _i(min_rows))
    out = []
    for lev in range(L.fb200_amg_host_levels(_i(h))):
        d = {}
        for which, nm in enumerate("APRMTQ"):
            nr, nc, nnz = C.c_int(), C.c_int(), C.c_int()
            L.fb200_amg_host_dims(_i(h), _i(lev), _i(which), C.byref(nr), C.byref(nc), C.byref(nnz))
            if nr.value == 0:
                continue
            r_, c_, v_ = np.zeros(nr.value + 1, np.int32), np.zeros(nnz.value, np.int32), np.zeros(nnz.value)
            L.fb200_amg_host_fetch(_i(h), _i(lev), _i(which), _p(r_), _p(c_), _p(v_))
            d[nm] = sp.csr_matrix((v_, c_, r_), shape=(nr.value, nc.value))
        out.append(d)
    L.fb200_amg_host_free(_i(h))
    return out


class SSHSolver:
    """`solve_ssh_ale` (src/oce_ale.F90:2295-2336): psolver_init on first use, psolve per step."""
    SOLBICGS_RAS, PCRAS, PCILUK = 6, 2, 1        # lib/parms/include/fparms.h:25-32

    PC = {"bj": 0, "amg": 1}

    def __init__(self, hm, ident=1, block=1, soltol=1e-10, maxiter=2000, reuse=1, library=None, pc=None, amg=None):
        """pc: "bj" (block-Jacobi, `block` rows per block) | "amg" (aggregation multigrid) | None = library default /
        FB200_SSH_PC; amg = (theta, omega, max_levels, min_rows) overrides the multigrid defaults"""
        self.L = lib() if library is None else library
        self.hm = hm
        self.ident = ident
        L = self.L
        if library is None:
            L.fb200_clear_status()
            L.fb200_psolver_set_block(_i(block))
            L.fb200_psolver_set_pc(_i(-1 if pc is None else self.PC[pc]))       # the setters are sticky: always state both
            amg = (0.08, 0.8, 4, 200) if amg is None else amg
            L.fb200_psolver_set_amg(_d(amg[0]), _d(amg[1]), _i(amg[2]), _i(amg[3]))
        rp = hm.arr("ssh_rowptr")
        self._part = (hm.arr("part") - 1).astype(np.int32)                 # part-1
        self._rp = (rp - rp[0]).astype(np.int32)                            # rowptr(:)-rowptr(1)
        self._ci = (hm.arr("ssh_colind") - 1).astype(np.int32)             # colind-1
        vals = hm.arr("ssh_values")
        fcomm = C.c_int(0)
        L.psolver_init(_i(ident), _i(self.SOLBICGS_RAS), _i(self.PCRAS), _i(self.PCILUK), _i(2), _i(3), _d(1e-8),
                       _i(maxiter), _i(15), _d(soltol), _p(self._part), _p(self._rp), _p(self._ci), _p(vals), _i(reuse),
                       C.byref(fcomm))
        if library is None:
            _check(L)

    def solve(self, new_values=1, rhs=None, sol=None):
        hm = self.hm
        rhs = hm.arr("ssh_rhs") if rhs is None else rhs
        sol = hm.arr("d_eta") if sol is None else sol
        self.L.psolve(_i(self.ident), _p(rhs), _p(hm.arr("ssh_values")), _p(sol), _i(new_values))
        if hasattr(self.L, "fb200_status"):
            _check(self.L)
        return sol

    def solve_ptr(self, rhs_ptr, vals_ptr, sol_ptr, new_values=1):
        """psolve with raw addresses: the library takes host OR device pointers for rhs, vals and sol (unified addressing),
        so a host that assembles the system on the device (fb200_update_stiff_mat / fb200_compute_ssh_rhs) solves without
        crossing PCIe.  sol: in = initial guess, out = solution."""
        self.L.psolve(_i(self.ident), C.c_void_p(rhs_ptr), C.c_void_p(vals_ptr), C.c_void_p(sol_ptr), _i(new_values))
        if hasattr(self.L, "fb200_status"):
            _check(self.L)

    def info(self):
        its = C.c_int()
        res = C.c_double()
        self.L.fb200_psolve_info(_i(self.ident), C.byref(its), C.byref(res))
        return its.value, res.value

    def amg_levels(self):
        """[(rows, A entries, P entries, R entries)] of the multigrid hierarchy ([] with block-Jacobi)"""
        out = []
        for lev in range(8):
            rows, a, p, r = C.c_int(), C.c_longlong(), C.c_longlong(), C.c_longlong()
            self.L.fb200_psolve_amg_info(_i(self.ident), _i(lev), C.byref(rows), C.byref(a), C.byref(p), C.byref(r))
            if rows.value == 0:
                break
            out.append((rows.value, a.value, p.value, r.value))
        return out

    def last_ms(self):
        return self.L.fb200_last_ms(b"psolve")
