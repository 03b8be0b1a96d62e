// Host-side mirror of the reference's Fortran setup code (mesh, ALE geometry, MUSCL tables,
// SSH stiffness matrix, per-step producers).  It owns arrays in EXACTLY the layout the reference's
// Fortran driver would pass across the C ABI: column-major, (level, entity) with level fastest,
// index arrays holding 1-based values.  Nothing in here runs on the GPU and nothing in here is on the
// timed path; it is the caller-side of include/fesom_b200.h when no Fortran compiler is available.
//
// Reference: src/oce_mesh.F90, src/oce_ale.F90, src/oce_muscl_adv.F90, src/fvom_init.F90,
//            src/gen_modules_partitioning.F90 (cited per function in the .cpp files).
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <map>
#include <string>
#include <vector>
#include <stdexcept>
#include <algorithm>

namespace fh {

struct Arr {
    int dtype = 1;                 // 0 = int32, 1 = float64
    std::vector<int64_t> shape;    // Fortran shape (first index fastest)
    std::vector<char> buf;
    int64_t count() const { int64_t c = 1; for (auto s : shape) c *= s; return c; }
    double* d() { return reinterpret_cast<double*>(buf.data()); }
    int* i() { return reinterpret_cast<int*>(buf.data()); }
};

// Constants of the reference (src/oce_modules.F90:11-19).  pi is the truncated literal on purpose.
constexpr double PI = 3.14159265358979;
constexpr double RAD = PI / 180.0;
constexpr double R_EARTH = 6367500.0;
constexpr double G_ACC = 9.81;
constexpr double VCPW = 4.2e6;

struct Host {
    std::map<std::string, Arr> a;
    std::map<std::string, double> p;   // scalar parameters (ints stored as doubles)

    double* D(const std::string& name, std::vector<int64_t> shape) {
        Arr& r = a[name];
        r.dtype = 1; r.shape = shape;
        r.buf.assign(size_t(r.count()) * 8, 0);
        return r.d();
    }
    int* I(const std::string& name, std::vector<int64_t> shape) {
        Arr& r = a[name];
        r.dtype = 0; r.shape = shape;
        r.buf.assign(size_t(r.count()) * 4, 0);
        return r.i();
    }
    bool has(const std::string& name) const { return a.count(name) != 0; }
    double* d(const std::string& name) {
        auto it = a.find(name);
        if (it == a.end() || it->second.dtype != 1) throw std::runtime_error("host: no f64 array '" + name + "'");
        return it->second.d();
    }
    int* i(const std::string& name) {
        auto it = a.find(name);
        if (it == a.end() || it->second.dtype != 0) throw std::runtime_error("host: no i32 array '" + name + "'");
        return it->second.i();
    }
    int pi(const std::string& name) const {
        auto it = p.find(name);
        if (it == p.end()) throw std::runtime_error("host: no parameter '" + name + "'");
        return int(it->second);
    }
    double pd(const std::string& name) const {
        auto it = p.find(name);
        if (it == p.end()) throw std::runtime_error("host: no parameter '" + name + "'");
        return it->second;
    }
};

inline void trim_cyclic(double& b, double cyclic_length) {   // gen_modules_rotate_grid.F90:206-213
    if (b > cyclic_length / 2.0) b -= cyclic_length;
    if (b < -cyclic_length / 2.0) b += cyclic_length;
}

// mesh.cpp
void mesh_set_global(Host& h, int nod2D, const double* lon_deg, const double* lat_deg, int elem2D, const int* tri,
                     int nl, const double* zbar, const double* depth, const int* nlevels_nod, const int* nlevels_elem,
                     int edge2D, int edge2D_in, const int* edges, const int* edge_tri,
                     int force_rotation, double cyclic_deg, double alphaE, double betaE, double gammaE);
void mesh_find_edges(Host& h);
void mesh_find_levels(Host& h);
void mesh_derive(Host& h);
void mesh_restrict(Host& g, Host& loc, int npes, int mype, const int* mylist_nod, int myDim_nod, int eDim_nod,
                   const int* mylist_elem, int myDim_elem, int eDim_elem, int eXDim_elem,
                   const int* mylist_edge, int myDim_edge, int eDim_edge);
// ale.cpp
void ale_init(Host& h, const char* which_ale, int use_partial_cell, double partial_cell_thresh);
void ale_init_thickness(Host& h);
void ale_vert_vel(Host& h);
void ale_update_thickness(Host& h);
void ale_compute_hbar(Host& h);
// setup.cpp
void muscl_adv_init(Host& h);
void stiff_mat_init(Host& h);
void stiff_mat_update(Host& h);
void compute_ssh_rhs(Host& h);
void state_alloc(Host& h, int num_tracers);

}  // namespace fh
