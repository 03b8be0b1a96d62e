"""ORACLE — TEST INFRASTRUCTURE ONLY (see oracle/oracle.h).

ctypes binding of liboracle.so (the C restatement of the reference's tracer step and SSH solve) and of
oracle/_ref/libpsolve_ref.so (the reference's own psolve.c + pARMS, compiled serially from
/root/reference).  Imported by tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs only.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_lib = None

HOR = {"MUSCL": 0, "MFCT": 1, "UPW1": 2}
VER = {"QR4C": 0, "CDIFF": 1, "PPM": 2, "UPW1": 3}
LIM = {"FCT": 0, "NONE": 1}

MESH_INT = ["edges", "edge_tri", "elem2D_nodes", "nlevels", "ulevels", "nlevels_nod2D", "ulevels_nod2D",
            "nlevels_nod2D_min", "ulevels_nod2D_max", "nod_in_elem2D", "nod_in_elem2D_num", "edge_up_dn_tri",
            "nboundary_lay"]
MESH_DBL = ["edge_cross_dxdy", "edge_dxdy", "elem_cos", "elem_area", "gradient_sca", "area", "areasvol", "zbar_n_bot"]
STATE_DBL = ["UV", "Wvel", "Wvel_e", "Wvel_i", "hnode", "hnode_new", "helem", "Z_3d_n", "zbar_3d_n", "Kv", "Ki",
             "slope_tapered", "heat_flux", "water_flux", "virtual_salt", "relax_salt", "real_salt_flux", "sw_3d"]
DIMS = ["nl", "myDim_nod2D", "eDim_nod2D", "myDim_elem2D", "eDim_elem2D", "eXDim_elem2D", "myDim_edge2D"]


_timing = False


def use_timing_build(on=True):
    """bench.py's CPU arm only: load the copies compiled with the reference's own flags (liboracle_timing.so built on this
    machine with -O3 -march=native -finline-functions; oracle/_ref/libpsolve_ref_timing.so) instead of the parity
    checkers (-O2 -ffp-contract=off).  Call before the first lib() / ref_lib()."""
    global _timing
    assert _lib is None and _ref is None
    _timing = bool(on)


def lib():
    global _lib
    if _lib is None:
        from harness import build
        L = C.CDLL(build.build_oracle_timing() if _timing else build.build_oracle())
        L.oracle_create.restype = C.c_void_p
        L.oracle_free.argtypes = [C.c_void_p]
        L.oracle_set_int.argtypes = [C.c_void_p, C.c_char_p, C.c_int]
        L.oracle_set_dbl.argtypes = [C.c_void_p, C.c_char_p, C.c_double]
        L.oracle_set_ptr.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p]
        L.oracle_get_ptr.argtypes = [C.c_void_p, C.c_char_p]
        L.oracle_get_ptr.restype = C.c_void_p
        L.oracle_alloc_work.argtypes = [C.c_void_p]
        for nm in ("o_phase1", "o_phase2", "o_phase3", "o_phase4", "o_phase5"):
            getattr(L, nm).argtypes = [C.c_void_p, C.c_int]
            getattr(L, nm).restype = None
        L.o_phase_final.argtypes = [C.c_void_p]
        L.o_phase_final.restype = None
        L.o_solve_tracers_ale.argtypes = [C.c_void_p]
        L.o_solve_tracers_ale.restype = None
        L.o_ssh_solve.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                  C.c_double, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_double)]
        L.o_spmv.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.o_spmv.restype = None
        _lib = L
    return _lib


class TracerOracle:
    """Binds a HostModel's arrays (no copies) into an oracle context.  tr_arr / tr_arr_old are updated in place."""

    def __init__(self, hm, tra_adv_hor="MFCT", tra_adv_ver="QR4C", tra_adv_lim="FCT", tra_adv_ph=1.0, tra_adv_pv=1.0,
                 w_split=None, i_vert_diff=True, Redi=False, use_sw_pene=False, do_hor_diff=True, clamp_salt=True,
                 tracer_ids=(0, 1), fct_cavity_neutral=False, smooth_bh_tra=False, gamma_tra=(0.0005, 0.0125, 0.0), bh_split=False):
        self.L = lib()
        self.hm = hm
        self.o = C.c_void_p(self.L.oracle_create())
        self._keep = []
        si, sd, sp = self.L.oracle_set_int, self.L.oracle_set_dbl, self.L.oracle_set_ptr
        for d in DIMS:
            assert si(self.o, d.encode(), hm.ipar(d)) == 0
        si(self.o, b"nod_in_elem2D_max", hm.ipar("nod_in_elem2D_max"))
        si(self.o, b"num_tracers", hm.ipar("num_tracers"))
        for nm in MESH_INT + MESH_DBL + STATE_DBL + ["tr_arr", "tr_arr_old"]:
            a = hm.arr(nm)
            self._keep.append(a)
            assert sp(self.o, nm.encode(), a.ctypes.data) == 0, nm
        sd(self.o, b"dt", hm.par("dt"))
        sd(self.o, b"is_nonlinfs", hm.par("is_nonlinfs"))
        sd(self.o, b"tra_adv_ph", tra_adv_ph)
        sd(self.o, b"tra_adv_pv", tra_adv_pv)
        si(self.o, b"tra_adv_hor", HOR[tra_adv_hor])
        si(self.o, b"tra_adv_ver", VER[tra_adv_ver])
        si(self.o, b"tra_adv_lim", LIM[tra_adv_lim])
        si(self.o, b"w_split", hm.ipar("w_split") if w_split is None else int(w_split))
        si(self.o, b"i_vert_diff", int(i_vert_diff))
        si(self.o, b"Redi", int(Redi))
        si(self.o, b"use_sw_pene", int(use_sw_pene))
        si(self.o, b"do_hor_diff", int(do_hor_diff))
        si(self.o, b"clamp_salt", int(clamp_salt))
        si(self.o, b"fct_cavity_neutral", int(fct_cavity_neutral))
        si(self.o, b"smooth_bh_tra", int(smooth_bh_tra))
        si(self.o, b"bh_split", int(bh_split))
        for k in range(3):
            sd(self.o, ("gamma%d_tra" % k).encode(), float(gamma_tra[k]))
        for k, tid in enumerate(tracer_ids):
            si(self.o, ("tracer_id%d" % k).encode(), int(tid))
        assert self.L.oracle_alloc_work(self.o) == 0

    def __del__(self):
        try:
            self.L.oracle_free(self.o)
        except Exception:
            pass

    def set_clim(self, clim_relax, relax2clim, Tclim, Sclim):
        """climatology relaxation (relax_to_clim, oce_tracer_mod.F90:87-119); arrays are bound, not copied"""
        self._keep += [relax2clim, Tclim, Sclim]
        self.L.oracle_set_dbl(self.o, b"clim_relax", float(clim_relax))
        for nm, a in (("relax2clim", relax2clim), ("Tclim", Tclim), ("Sclim", Sclim)):
            assert self.L.oracle_set_ptr(self.o, nm.encode(), a.ctypes.data) == 0

    def set_zonal_relax(self, ztem, coord_nod2D, Tclim, lat0, dy, tau_inv):
        """relax_zonal_temp of the soufflet channel (toy_channel_soufflet.F90:78-103) instead of relax_to_clim:
        ztem (nybins, nl-1), coord_nod2D (N, 2) [rad], Tclim (N, nl-1); arrays are bound, not copied"""
        self._keep += [ztem, coord_nod2D, Tclim]
        self.L.oracle_set_int(self.o, b"relax_zonal", 1)
        for nm, v in (("zonal_lat0", lat0), ("zonal_dy", dy), ("zonal_tau_inv", tau_inv)):
            assert self.L.oracle_set_dbl(self.o, nm.encode(), float(v)) == 0
        for nm, a in (("zonal_ztem", ztem), ("coord_nod2D", coord_nod2D), ("Tclim", Tclim)):
            assert a.flags.c_contiguous and a.dtype == np.float64
            assert self.L.oracle_set_ptr(self.o, nm.encode(), a.ctypes.data) == 0

    def set_kpp_nonlocal(self, gT, gS, b2, b3, ref_sss=34.7, ref_sss_local=False):
        """KPP non-local transport in diff_ver_part_impl_ale (oce_ale_tracer.F90:688-781); (N, nl) arrays, bound not copied"""
        self._keep += [gT, gS, b2, b3]
        self.L.oracle_set_int(self.o, b"use_kpp_nonlclflx", 1)
        self.L.oracle_set_int(self.o, b"ref_sss_local", int(ref_sss_local))
        self.L.oracle_set_dbl(self.o, b"ref_sss", float(ref_sss))
        for nm, a in (("kpp_gT", gT), ("kpp_gS", gS), ("kpp_b2", b2), ("kpp_b3", b3)):
            assert self.L.oracle_set_ptr(self.o, nm.encode(), a.ctypes.data) == 0

    def work(self, name, fshape):
        """numpy view (reversed Fortran shape) of an oracle-owned work array"""
        p = self.L.oracle_get_ptr(self.o, name.encode())
        n = int(np.prod(fshape))
        buf = (C.c_double * n).from_address(p)
        return np.frombuffer(buf, dtype=np.float64).reshape(tuple(fshape)[::-1])

    def step(self):
        self.L.o_solve_tracers_ale(self.o)

    def phase(self, k, tr_num):
        getattr(self.L, "o_phase%d" % k)(self.o, int(tr_num))

    def final(self):
        self.L.o_phase_final(self.o)


class MeanStream:
    """one output stream of io_meandata.F90 (update_means :768-794, compute_means/clean_meanarrays :894-900) over a
    (n, nl-1) field in real64 (accuracy 8) or real32 (accuracy 4)"""

    def __init__(self, shape, accuracy=8):
        self.accuracy = int(accuracy)
        self.acc = np.zeros(shape, dtype=np.float64 if accuracy == 8 else np.float32)
        self.count = 0

    def update(self, field):
        f = np.ascontiguousarray(field, dtype=np.float64)
        assert f.shape == self.acc.shape
        lib().o_update_means(self.accuracy, C.c_void_p(self.acc.ctypes.data), C.c_void_p(f.ctypes.data), C.c_longlong(f.size))
        self.count += 1

    def output(self):
        out = np.zeros_like(self.acc)
        lib().o_compute_means(self.accuracy, C.c_void_p(self.acc.ctypes.data), C.c_void_p(out.ctypes.data), C.c_longlong(out.size), self.count)
        self.count = 0
        return out


def ssh_solve(rowptr0, colind0, values, rhs, x0, pc_kind=1, block=1, soltol=1e-10, maxits=2000):
    L = lib()
    rp = np.ascontiguousarray(rowptr0, dtype=np.int32)
    ci = np.ascontiguousarray(colind0, dtype=np.int32)
    v = np.ascontiguousarray(values, dtype=np.float64)
    b = np.ascontiguousarray(rhs, dtype=np.float64)
    x = np.array(x0, dtype=np.float64, copy=True)
    its = C.c_int()
    res = C.c_double()
    rc = L.o_ssh_solve(rp.size - 1, rp.ctypes.data, ci.ctypes.data, v.ctypes.data, b.ctypes.data, x.ctypes.data,
                       int(pc_kind), int(block), float(soltol), int(maxits), C.byref(its), C.byref(res))
    return x, its.value, res.value, rc


# ---------------------------------------------------------------------------------------------------
_ref = None


def ref_lib():
    """the reference's own psolver_init/psolve (src/psolve.c + lib/parms), or None if it was never built"""
    global _ref
    if _ref is None:
        from harness import build
        so = build.build_ref()
        if _timing and so is not None and os.path.exists(build.REF_TIMING_SO):
            so = build.REF_TIMING_SO
        if so is None or not os.path.exists(so):
            return None
        R = C.CDLL(so)
        P = C.POINTER
        R.psolver_init.argtypes = [P(C.c_int)] * 6 + [P(C.c_double)] + [P(C.c_int)] * 2 + [P(C.c_double)] + \
                                  [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, P(C.c_int), P(C.c_int)]
        R.psolver_init.restype = None
        R.psolve.argtypes = [P(C.c_int), C.c_void_p, C.c_void_p, C.c_void_p, P(C.c_int)]
        R.psolve.restype = None
        _ref = R
    return _ref


class RefSolver:
    """The reference's own psolver_init/psolve (oracle/_ref) on one subdomain.  The MPI shim reports two ranks,
    the second one owning no rows (oracle/mpi_stub/mpi_stub.c), so `part` has three entries."""
    _next_id = 1

    def __init__(self, rowptr0, colind0, values, soltol=1e-10, maxits=2000, ilulevel=2, fillin=3, droptol=1e-8):
        R = ref_lib()
        if R is None:
            raise RuntimeError("oracle/_ref/libpsolve_ref.so has not been built (needs /root/reference)")
        self.R = R
        self.rp = np.ascontiguousarray(rowptr0, dtype=np.int32)
        self.ci = np.ascontiguousarray(colind0, dtype=np.int32)
        self.n = self.rp.size - 1
        self.part = np.array([0, self.n, self.n], dtype=np.int32)
        self.id = RefSolver._next_id
        RefSolver._next_id += 1
        assert self.id < 10, "psolve.c keeps at most NSOL=10 solvers"
        v = np.ascontiguousarray(values, dtype=np.float64)
        i = lambda x: C.byref(C.c_int(int(x)))          # noqa: E731
        d = lambda x: C.byref(C.c_double(float(x)))     # noqa: E731
        # SOLBICGS_RAS=6, PCRAS=2, PCILUK=1 (lib/parms/include/fparms.h:25-32; oce_ale.F90:2330)
        R.psolver_init(i(self.id), i(6), i(2), i(1), i(ilulevel), i(fillin), d(droptol), i(maxits), i(15), d(soltol),
                       self.part.ctypes.data, self.rp.ctypes.data, self.ci.ctypes.data, v.ctypes.data, i(1), i(0))

    def solve(self, rhs, values, x0, new_values=1):
        x = np.array(x0, dtype=np.float64, copy=True)
        b = np.ascontiguousarray(rhs, dtype=np.float64)
        v = np.ascontiguousarray(values, dtype=np.float64)
        self.R.psolve(C.byref(C.c_int(self.id)), b.ctypes.data, v.ctypes.data, x.ctypes.data, C.byref(C.c_int(int(new_values))))
        return x
