"""ctypes front of the host-side mirror (harness/host/*.cpp).

`HostModel` plays the part of the reference's Fortran driver state: it owns the mesh (t_mesh), the
partition-independent derived geometry, the ALE arrays and the model state in the reference's own memory
layout (column-major, level fastest, 1-based index values).  Arrays are exposed as numpy views whose
*numpy* shape is the reversed Fortran shape (so `tr_arr[tr, node, level]`), i.e. no copies are made and
`arr.ctypes.data` is exactly the pointer the Fortran caller would pass over the C ABI.

Reference: src/oce_mesh.F90 (mesh_setup), src/oce_setup_step.F90 (ocean_setup), src/oce_ale.F90.
"""
import ctypes as C
import os

import numpy as np

from . import build as _build

_lib = None


def lib():
    global _lib
    if _lib is None:
        so = _build.build_host()
        L = C.CDLL(so)
        L.fh_last_error.restype = C.c_char_p
        L.fh_create.restype = C.c_void_p
        L.fh_free.argtypes = [C.c_void_p]
        L.fh_set_param.argtypes = [C.c_void_p, C.c_char_p, C.c_double]
        L.fh_get_param.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_double)]
        L.fh_get_array.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_void_p), C.POINTER(C.c_int),
                                   C.POINTER(C.c_int), C.POINTER(C.c_int64)]
        L.fh_new_array.argtypes = [C.c_void_p, C.c_char_p, C.c_int, C.c_int, C.POINTER(C.c_int64)]
        L.fh_list_arrays.argtypes = [C.c_void_p, C.c_char_p, C.c_int]
        L.fh_mesh_set_global.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int,
                                         C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                         C.c_void_p, C.c_void_p, C.c_int, C.c_double, C.c_double, C.c_double, C.c_double]
        for nm in ("fh_mesh_find_edges", "fh_mesh_find_levels", "fh_mesh_derive", "fh_ale_init_thickness",
                   "fh_ale_vert_vel", "fh_ale_update_thickness", "fh_ale_compute_hbar", "fh_muscl_adv_init",
                   "fh_stiff_mat_init", "fh_stiff_mat_update", "fh_compute_ssh_rhs"):
            getattr(L, nm).argtypes = [C.c_void_p]
        L.fh_ale_init.argtypes = [C.c_void_p, C.c_char_p, C.c_int, C.c_double]
        L.fh_state_alloc.argtypes = [C.c_void_p, C.c_int]
        _lib = L
    return _lib


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class HostModel:
    def __init__(self):
        self._L = lib()
        self._h = C.c_void_p(self._L.fh_create())

    def __del__(self):
        try:
            self._L.fh_free(self._h)
        except Exception:
            pass

    def _ck(self, rc):
        if rc != 0:
            raise RuntimeError(self._L.fh_last_error().decode())

    # ---- parameters and arrays
    def set(self, **kw):
        for k, v in kw.items():
            self._ck(self._L.fh_set_param(self._h, k.encode(), float(v)))

    def par(self, name):
        v = C.c_double()
        self._ck(self._L.fh_get_param(self._h, name.encode(), C.byref(v)))
        return v.value

    def ipar(self, name):
        return int(self.par(name))

    def arr(self, name):
        """numpy view, numpy shape = reversed Fortran shape"""
        p = C.c_void_p()
        dt = C.c_int()
        nd = C.c_int()
        shp = (C.c_int64 * 4)()
        self._ck(self._L.fh_get_array(self._h, name.encode(), C.byref(p), C.byref(dt), C.byref(nd), shp))
        shape = tuple(int(shp[k]) for k in range(nd.value))[::-1]
        n = int(np.prod(shape)) if shape else 0
        ctype = C.c_int32 if dt.value == 0 else C.c_double
        if n == 0:
            return np.zeros(shape, dtype=np.int32 if dt.value == 0 else np.float64)
        buf = (ctype * n).from_address(p.value)
        return np.frombuffer(buf, dtype=np.int32 if dt.value == 0 else np.float64).reshape(shape)

    def new(self, name, fshape, dtype=np.float64):
        """allocate a zeroed array with FORTRAN shape `fshape`; returns the numpy view"""
        shp = (C.c_int64 * len(fshape))(*fshape)
        self._ck(self._L.fh_new_array(self._h, name.encode(), 0 if dtype == np.int32 else 1, len(fshape), shp))
        return self.arr(name)

    def names(self):
        buf = C.create_string_buffer(1 << 16)
        self._ck(self._L.fh_list_arrays(self._h, buf, len(buf)))
        return [s for s in buf.value.decode().split("\n") if s]

    def has(self, name):
        return name in self.names()

    def __getitem__(self, name):
        return self.arr(name)

    # ---- mesh
    def set_global_mesh(self, lon_deg, lat_deg, tri, zbar, depth, nlevels_nod=None, nlevels_elem=None,
                        edges=None, edge_tri=None, edge2D_in=0, force_rotation=False, cyclic_deg=360.0,
                        euler=(50.0, 15.0, -90.0)):
        lon = np.ascontiguousarray(lon_deg, dtype=np.float64)
        lat = np.ascontiguousarray(lat_deg, dtype=np.float64)
        tri = np.ascontiguousarray(tri, dtype=np.int32)          # (elem2D, 3), 1-based
        zbar = np.ascontiguousarray(zbar, dtype=np.float64)
        depth = np.ascontiguousarray(depth, dtype=np.float64)
        nn = None if nlevels_nod is None else np.ascontiguousarray(nlevels_nod, dtype=np.int32)
        ne = None if nlevels_elem is None else np.ascontiguousarray(nlevels_elem, dtype=np.int32)
        ed = None if edges is None else np.ascontiguousarray(edges, dtype=np.int32)
        et = None if edge_tri is None else np.ascontiguousarray(edge_tri, dtype=np.int32)
        self._ck(self._L.fh_mesh_set_global(
            self._h, lon.size, _ptr(lon), _ptr(lat), tri.shape[0], _ptr(tri), zbar.size, _ptr(zbar), _ptr(depth),
            _ptr(nn), _ptr(ne), 0 if ed is None else ed.shape[0], int(edge2D_in), _ptr(ed), _ptr(et),
            int(bool(force_rotation)), float(cyclic_deg), float(euler[0]), float(euler[1]), float(euler[2])))
        if ed is None:
            self._ck(self._L.fh_mesh_find_edges(self._h))
        if nn is None:
            self._ck(self._L.fh_mesh_find_levels(self._h))

    def derive(self):
        self._ck(self._L.fh_mesh_derive(self._h))

    def setup_ocean(self, dt, which_ale="zstar", use_partial_cell=True, partial_cell_thresh=0.0, alpha=1.0, theta=1.0,
                    num_tracers=2, w_split=False, w_max_cfl=1.0, K_hor=10.0):
        """ocean_setup: array_setup, init_ale, init_stiff_mat_ale, muscl_adv_init, init_thickness_ale
        (src/oce_setup_step.F90:20-170)"""
        self.set(dt=dt, alpha=alpha, theta=theta, w_split=int(w_split), w_max_cfl=w_max_cfl, K_hor=K_hor)
        self._ck(self._L.fh_state_alloc(self._h, int(num_tracers)))
        self._ck(self._L.fh_ale_init(self._h, which_ale.encode(), int(use_partial_cell), float(partial_cell_thresh)))
        self._ck(self._L.fh_stiff_mat_init(self._h))
        self._ck(self._L.fh_muscl_adv_init(self._h))
        self._ck(self._L.fh_ale_init_thickness(self._h))

    def init_thickness(self):
        self._ck(self._L.fh_ale_init_thickness(self._h))

    def vert_vel(self):
        self._ck(self._L.fh_ale_vert_vel(self._h))

    def update_thickness(self):
        self._ck(self._L.fh_ale_update_thickness(self._h))

    def compute_hbar(self):
        self._ck(self._L.fh_ale_compute_hbar(self._h))

    def stiff_mat_update(self):
        self._ck(self._L.fh_stiff_mat_update(self._h))

    def compute_ssh_rhs(self):
        self._ck(self._L.fh_compute_ssh_rhs(self._h))


# ---------------------------------------------------------------------------------------------------
# mesh files of the reference (test/meshes/<name>): nod2d.out, elem2d.out, aux3d.out, nlvls.out,
# elvls.out, edges.out, edge_tri.out, edgenum.out  (formats: oce_mesh.F90:266-307, 412-436, 482-533,
# 1456-1499; find_levels :699-893)
def read_mesh_files(path):
    def rd(name):
        with open(os.path.join(path, name)) as f:
            return f.read().split()
    t = rd("nod2d.out")
    n = int(t[0])
    a = np.array(t[1:1 + 4 * n], dtype=np.float64).reshape(n, 4)
    lon, lat = a[:, 1].copy(), a[:, 2].copy()
    t = rd("elem2d.out")
    ne = int(t[0])
    tri = np.array(t[1:1 + 3 * ne], dtype=np.int32).reshape(ne, 3)
    t = rd("aux3d.out")
    nl = int(t[0])
    zbar = np.array(t[1:1 + nl], dtype=np.float64)
    depth = np.array(t[1 + nl:1 + nl + n], dtype=np.float64)
    nlv = np.array(rd("nlvls.out"), dtype=np.int32)
    elv = np.array(rd("elvls.out"), dtype=np.int32)
    t = rd("edgenum.out")
    edge2D, edge2D_in = int(t[0]), int(t[1])
    edges = np.array(rd("edges.out"), dtype=np.int32).reshape(edge2D, 2)
    edge_tri = np.array(rd("edge_tri.out"), dtype=np.int32).reshape(edge2D, 2)
    return dict(lon=lon, lat=lat, tri=tri, zbar=zbar, depth=depth, nlevels_nod=nlv, nlevels_elem=elv,
                edges=edges, edge_tri=edge_tri, edge2D_in=edge2D_in)


def model_from_files(path, force_rotation=True, cyclic_deg=360.0, cartesian=False):
    m = read_mesh_files(path)
    h = HostModel()
    if cartesian:
        h.set(cartesian=1)
    h.set_global_mesh(m["lon"], m["lat"], m["tri"], m["zbar"], m["depth"], m["nlevels_nod"], m["nlevels_elem"],
                      m["edges"], m["edge_tri"], m["edge2D_in"], force_rotation=force_rotation, cyclic_deg=cyclic_deg)
    h.derive()
    return h
