// Plain C entry points of the host-side mirror, bound from Python with ctypes (fesom2_b200/host.py).
#include "host.hpp"

using namespace fh;
static thread_local std::string g_err;

#define GUARD(body) try { body; return 0; } catch (const std::exception& e) { g_err = e.what(); return 1; }

extern "C" {

const char* fh_last_error() { return g_err.c_str(); }
void* fh_create() { return new Host(); }
void fh_free(void* h) { delete static_cast<Host*>(h); }

int fh_set_param(void* h, const char* name, double v) { GUARD(static_cast<Host*>(h)->p[name] = v) }
int fh_get_param(void* h, const char* name, double* v) { GUARD(*v = static_cast<Host*>(h)->pd(name)) }

// dtype: 0 = int32, 1 = float64; shape has room for 4 entries
int fh_get_array(void* hh, const char* name, void** ptr, int* dtype, int* ndim, int64_t* shape) {
    GUARD({
        Host& h = *static_cast<Host*>(hh);
        auto it = h.a.find(name);
        if (it == h.a.end()) throw std::runtime_error(std::string("host: no array '") + name + "'");
        *ptr = it->second.buf.data(); *dtype = it->second.dtype; *ndim = int(it->second.shape.size());
        for (size_t k = 0; k < it->second.shape.size(); k++) shape[k] = it->second.shape[k];
    })
}
int fh_new_array(void* hh, const char* name, int dtype, int ndim, const int64_t* shape) {
    GUARD({
        Host& h = *static_cast<Host*>(hh);
        std::vector<int64_t> s(shape, shape + ndim);
        if (dtype == 0) h.I(name, s); else h.D(name, s);
    })
}
int fh_list_arrays(void* hh, char* out, int cap) {
    GUARD({
        Host& h = *static_cast<Host*>(hh);
        std::string s;
        for (auto& kv : h.a) { s += kv.first; s += "\n"; }
        if (int(s.size()) + 1 > cap) throw std::runtime_error("list buffer too small");
        memcpy(out, s.c_str(), s.size() + 1);
    })
}

int fh_mesh_set_global(void* h, int nod2D, const double* lon, const double* lat, int elem2D, const int* tri, int nl,
                       const double* zbar, const double* depth, const int* nlevels_nod, const int* nlevels_elem,
                       int edge2D, int edge2D_in, const int* edges, const int* edge_tri,
                       int force_rotation, double cyclic_deg, double alphaE, double betaE, double gammaE) {
    GUARD(mesh_set_global(*static_cast<Host*>(h), nod2D, lon, lat, elem2D, tri, nl, zbar, depth, nlevels_nod, nlevels_elem,
                          edge2D, edge2D_in, edges, edge_tri, force_rotation, cyclic_deg, alphaE, betaE, gammaE))
}
int fh_mesh_find_edges(void* h) { GUARD(mesh_find_edges(*static_cast<Host*>(h))) }
int fh_mesh_find_levels(void* h) { GUARD(mesh_find_levels(*static_cast<Host*>(h))) }
int fh_mesh_derive(void* h) { GUARD(mesh_derive(*static_cast<Host*>(h))) }
int fh_ale_init(void* h, const char* which_ale, int use_partial_cell, double thresh) { GUARD(ale_init(*static_cast<Host*>(h), which_ale, use_partial_cell, thresh)) }
int fh_ale_init_thickness(void* h) { GUARD(ale_init_thickness(*static_cast<Host*>(h))) }
int fh_ale_vert_vel(void* h) { GUARD(ale_vert_vel(*static_cast<Host*>(h))) }
int fh_ale_update_thickness(void* h) { GUARD(ale_update_thickness(*static_cast<Host*>(h))) }
int fh_ale_compute_hbar(void* h) { GUARD(ale_compute_hbar(*static_cast<Host*>(h))) }
int fh_state_alloc(void* h, int num_tracers) { GUARD(state_alloc(*static_cast<Host*>(h), num_tracers)) }
int fh_muscl_adv_init(void* h) { GUARD(muscl_adv_init(*static_cast<Host*>(h))) }
int fh_stiff_mat_init(void* h) { GUARD(stiff_mat_init(*static_cast<Host*>(h))) }
int fh_stiff_mat_update(void* h) { GUARD(stiff_mat_update(*static_cast<Host*>(h))) }
int fh_compute_ssh_rhs(void* h) { GUARD(compute_ssh_rhs(*static_cast<Host*>(h))) }

}  // extern "C"
