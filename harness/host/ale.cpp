// ALE vertical geometry and the per-step producers of the tracer path's inputs
// (Wvel/Wvel_e/Wvel_i, hnode_new, helem, zbar_3d_n, Z_3d_n).  Caller-side mirror of src/oce_ale.F90.
#include "host.hpp"

namespace fh {
#define A2(p, ld, i, j) (p)[(int64_t)((j) - 1) * (ld) + ((i) - 1)]
#define A3(p, l1, l2, i, j, k) (p)[((int64_t)((k) - 1) * (l2) + ((j) - 1)) * (l1) + ((i) - 1)]

// init_ale + init_bottom_elem_thickness + init_bottom_node_thickness + init_surface_* (oce_ale.F90:82-544)
void ale_init(Host& h, const char* which_ale, int use_partial_cell, double partial_cell_thresh) {
    const int N = h.pi("myDim_nod2D") + h.pi("eDim_nod2D");
    const int myE = h.pi("myDim_elem2D"), E = myE + h.pi("eDim_elem2D");
    const int nl = h.pi("nl");
    int ale = 0;
    if (!strcmp(which_ale, "linfs")) ale = 0; else if (!strcmp(which_ale, "zlevel")) ale = 1;
    else if (!strcmp(which_ale, "zstar")) ale = 2; else throw std::runtime_error("unsupported which_ale");
    h.p["which_ale"] = ale; h.p["is_nonlinfs"] = ale == 0 ? 0.0 : 1.0;
    h.p["use_partial_cell"] = use_partial_cell;
    const int* en = h.i("elem2D_nodes"); const int* nle = h.i("nlevels"); const int* nln = h.i("nlevels_nod2D");
    const int* uln = h.i("ulevels_nod2D");
    const int* nien = h.i("nod_in_elem2D_num"); const int* nie = h.i("nod_in_elem2D"); const int maxn = h.pi("nod_in_elem2D_max");
    const double* zbar = h.d("zbar"); const double* Z = h.d("Z"); const double* depth = h.d("depth");
    h.D("hnode", {nl - 1, N}); h.D("hnode_new", {nl - 1, N}); h.D("ssh_rhs_old", {N});
    h.D("hbar", {N}); h.D("hbar_old", {N}); h.D("helem", {nl - 1, myE}); h.D("dhe", {myE});
    double* zbar_3d_n = h.D("zbar_3d_n", {nl, N}); double* Z_3d_n = h.D("Z_3d_n", {nl - 1, N});
    double* bet = h.D("bottom_elem_thickness", {myE});
    double* zeb = h.D("zbar_e_bot", {E}); double* zes = h.D("zbar_e_srf", {E});
    double* bnt = h.D("bottom_node_thickness", {N});
    double* znb = h.D("zbar_n_bot", {N}); double* zns = h.D("zbar_n_srf", {N});
    // init_bottom_elem_thickness :198-312  (all local elements so that halo values need no exchange)
    for (int e = 1; e <= E; e++) {
        int nlev = nle[e - 1];
        double zb;
        if (use_partial_cell) {
            double dd = (depth[A2(en, 3, 1, e) - 1] + depth[A2(en, 3, 2, e) - 1] + depth[A2(en, 3, 3, e) - 1]) / 3.0;
            if (zbar[nlev - 2] - zbar[nlev - 1] <= partial_cell_thresh) zb = zbar[nlev - 1];
            else if (dd < zbar[nlev - 1]) {
                if (nlev == nl) zb = std::max(dd, zbar[nlev - 1] + (zbar[nlev - 1] - Z[nlev - 2]));
                else zb = std::max(Z[nlev - 1], dd);
            } else zb = std::min(Z[nlev - 2], dd);
        } else zb = zbar[nlev - 1];
        zeb[e - 1] = zb;
        if (e <= myE) bet[e - 1] = zbar[nlev - 2] - zb;
    }
    // init_bottom_node_thickness :316-418
    for (int n = 1; n <= N; n++) {
        int nlev = nln[n - 1];
        if (use_partial_cell) {
            double mn = 1e300;
            for (int k = 1; k <= nien[n - 1]; k++) { int e = A2(nie, maxn, k, n); if (e <= E) mn = std::min(mn, zeb[e - 1]); }
            znb[n - 1] = mn;
        } else znb[n - 1] = zbar[nlev - 1];
        bnt[n - 1] = zbar[nlev - 2] - znb[n - 1];
    }
    for (int n = 0; n < N; n++) zns[n] = zbar[0];
    for (int e = 0; e < E; e++) zes[e] = zbar[0];
    // (cavity surface depths: use_cavity_partial_cell = .false. path, zbar_*_srf = zbar(ulevel))
    {
        const int* ule = h.i("ulevels");
        for (int e = 0; e < E; e++) if (ule[e] > 1) zes[e] = zbar[ule[e] - 1];
        for (int n = 0; n < N; n++) if (uln[n] > 1) zns[n] = zbar[uln[n] - 1];
    }
    // :150-190
    for (int n = 1; n <= N; n++) {
        int nzmin = uln[n - 1], nzmax = nln[n - 1];
        for (int nz = 1; nz <= nzmin - 1; nz++) A2(zbar_3d_n, nl, nz, n) = zbar[nz - 1];
        A2(zbar_3d_n, nl, nzmin, n) = zns[n - 1];
        for (int nz = nzmin + 1; nz <= nzmax - 1; nz++) A2(zbar_3d_n, nl, nz, n) = zbar[nz - 1];
        A2(zbar_3d_n, nl, nzmax, n) = znb[n - 1];
        for (int nz = 1; nz <= nzmin - 1; nz++) A2(Z_3d_n, nl - 1, nz, n) = Z[nz - 1];
        A2(Z_3d_n, nl - 1, nzmin, n) = A2(zbar_3d_n, nl, nzmin, n) + (A2(zbar_3d_n, nl, nzmin + 1, n) - zns[n - 1]) / 2;
        for (int nz = nzmin + 1; nz <= nzmax - 2; nz++) A2(Z_3d_n, nl - 1, nz, n) = Z[nz - 1];
        A2(Z_3d_n, nl - 1, nzmax - 1, n) = A2(zbar_3d_n, nl, nzmax - 1, n) + (znb[n - 1] - A2(zbar_3d_n, nl, nzmax - 1, n)) / 2;
    }
}

// init_thickness_ale (oce_ale.F90:549-795); hbar / hbar_old may have been set by the caller beforehand
void ale_init_thickness(Host& h) {
    const int N = h.pi("myDim_nod2D") + h.pi("eDim_nod2D"), myE = h.pi("myDim_elem2D"), nl = h.pi("nl");
    const int ale = h.pi("which_ale");
    const double dt = h.pd("dt"), alpha = h.pd("alpha");
    const int* en = h.i("elem2D_nodes"); const int* nle = h.i("nlevels"); const int* ule = h.i("ulevels");
    const int* nln = h.i("nlevels_nod2D"); const int* uln = h.i("ulevels_nod2D"); const int* nlmin = h.i("nlevels_nod2D_min");
    const double* zbar = h.d("zbar"); const double* areasvol = h.d("areasvol");
    double* hbar = h.d("hbar"); double* hbar_old = h.d("hbar_old"); double* ssh_rhs_old = h.d("ssh_rhs_old");
    double* hnode = h.d("hnode"); double* hnode_new = h.d("hnode_new"); double* helem = h.d("helem"); double* dhe = h.d("dhe");
    const double* z3 = h.d("zbar_3d_n"); const double* bnt = h.d("bottom_node_thickness"); const double* bet = h.d("bottom_elem_thickness");
    const double* zes = h.d("zbar_e_srf");
    double* eta_n = h.has("eta_n") ? h.d("eta_n") : h.D("eta_n", {N});
    for (int n = 1; n <= N; n++) ssh_rhs_old[n - 1] = (hbar[n - 1] - hbar_old[n - 1]) * A2(areasvol, nl, uln[n - 1], n) / dt;
    for (int n = 0; n < N; n++) eta_n[n] = alpha * hbar_old[n] + (1.0 - alpha) * hbar[n];
    const int L = nl - 1;
    if (ale == 0 || ale == 1) {   // linfs :590-633, zlevel :637-690
        for (int n = 1; n <= N; n++) {
            int nzmin = uln[n - 1], nzmax = nln[n - 1] - 1;
            for (int nz = nzmin; nz <= nzmax - 1; nz++) A2(hnode, L, nz, n) = A2(z3, nl, nz, n) - A2(z3, nl, nz + 1, n);
            if (ale == 1 && nzmin == 1) A2(hnode, L, nzmin, n) = hbar[n - 1] + (A2(z3, nl, nzmin, n) - A2(z3, nl, nzmin + 1, n));
            A2(hnode, L, nzmax, n) = bnt[n - 1];
        }
        for (int e = 1; e <= myE; e++) {
            int nzmin = ule[e - 1], nzmax = nle[e - 1] - 1;
            if (ale == 1) dhe[e - 1] = (hbar[A2(en, 3, 1, e) - 1] + hbar[A2(en, 3, 2, e) - 1] + hbar[A2(en, 3, 3, e) - 1]) / 3.0;
            A2(helem, L, nzmin, e) = zes[e - 1] - zbar[nzmin];
            if (ale == 1 && nzmin == 1) A2(helem, L, nzmin, e) = dhe[e - 1] + (zes[e - 1] - zbar[nzmin]);
            for (int nz = nzmin + 1; nz <= nzmax - 1; nz++) A2(helem, L, nz, e) = zbar[nz - 1] - zbar[nz];
            A2(helem, L, nzmax, e) = bet[e - 1];
        }
    } else {                      // zstar :694-783
        for (int n = 1; n <= N; n++) {
            int nzmin = uln[n - 1], nzmax = nln[n - 1] - 1;
            if (nzmin == 1) {
                double dd = zbar[nzmin - 1] - zbar[nlmin[n - 1] - 2];
                for (int nz = nzmin; nz <= nlmin[n - 1] - 2; nz++) A2(hnode, L, nz, n) = (zbar[nz - 1] - zbar[nz]) * (1.0 + hbar[n - 1] / dd);
                for (int nz = nlmin[n - 1] - 1; nz <= nzmax - 1; nz++) A2(hnode, L, nz, n) = zbar[nz - 1] - zbar[nz];
            } else
                for (int nz = nzmin; nz <= nzmax - 1; nz++) A2(hnode, L, nz, n) = A2(z3, nl, nz, n) - A2(z3, nl, nz + 1, n);
            A2(hnode, L, nzmax, n) = bnt[n - 1];
        }
        for (int e = 1; e <= myE; e++) {
            int nzmin = ule[e - 1], nzmax = nle[e - 1] - 1;
            int n1 = A2(en, 3, 1, e), n2 = A2(en, 3, 2, e), n3 = A2(en, 3, 3, e);
            if (nzmin == 1) {
                dhe[e - 1] = (hbar[n1 - 1] + hbar[n2 - 1] + hbar[n3 - 1]) / 3.0;
                A2(helem, L, nzmin, e) = (A2(hnode, L, nzmin, n1) + A2(hnode, L, nzmin, n2) + A2(hnode, L, nzmin, n3)) / 3.0;
            } else A2(helem, L, nzmin, e) = zes[e - 1] - zbar[nzmin];
            for (int nz = nzmin + 1; nz <= nzmax - 1; nz++)
                A2(helem, L, nz, e) = (A2(hnode, L, nz, n1) + A2(hnode, L, nz, n2) + A2(hnode, L, nz, n3)) / 3.0;
            A2(helem, L, nzmax, e) = bet[e - 1];
        }
    }
    memcpy(hnode_new, hnode, sizeof(double) * size_t(L) * N);
}

// compute_hbar_ale (oce_ale.F90:1585-1676)
void ale_compute_hbar(Host& h) {
    const int myN = h.pi("myDim_nod2D"), N = myN + h.pi("eDim_nod2D"), myE = h.pi("myDim_elem2D"), myEd = h.pi("myDim_edge2D"), nl = h.pi("nl");
    const int L = nl - 1, ale = h.pi("which_ale");
    const double dt = h.pd("dt");
    const int* edges = h.i("edges"); const int* edge_tri = h.i("edge_tri"); const int* en = h.i("elem2D_nodes");
    const int* nle = h.i("nlevels"); const int* ule = h.i("ulevels"); const int* uln = h.i("ulevels_nod2D");
    const double* ecd = h.d("edge_cross_dxdy"); const double* UV = h.d("UV"); const double* helem = h.d("helem");
    const double* areasvol = h.d("areasvol"); const double* water_flux = h.d("water_flux");
    double* sro = h.d("ssh_rhs_old"); double* hbar = h.d("hbar"); double* hbar_old = h.d("hbar_old"); double* dhe = h.d("dhe");
    for (int n = 0; n < N; n++) sro[n] = 0.0;
    for (int ed = 1; ed <= myEd; ed++) {
        int e1 = A2(edge_tri, 2, 1, ed), e2 = A2(edge_tri, 2, 2, ed);
        double c1 = 0.0, c2 = 0.0;
        double dX1 = A2(ecd, 4, 1, ed), dY1 = A2(ecd, 4, 2, ed);
        for (int nz = ule[e1 - 1]; nz <= nle[e1 - 1] - 1; nz++)
            c1 = c1 + (A3(UV, 2, L, 2, nz, e1) * dX1 - A3(UV, 2, L, 1, nz, e1) * dY1) * A2(helem, L, nz, e1);
        if (e2 > 0) {
            double dX2 = A2(ecd, 4, 3, ed), dY2 = A2(ecd, 4, 4, ed);
            for (int nz = ule[e2 - 1]; nz <= nle[e2 - 1] - 1; nz++)
                c2 = c2 - (A3(UV, 2, L, 2, nz, e2) * dX2 - A3(UV, 2, L, 1, nz, e2) * dY2) * A2(helem, L, nz, e2);
        }
        sro[A2(edges, 2, 1, ed) - 1] = sro[A2(edges, 2, 1, ed) - 1] + (c1 + c2);
        sro[A2(edges, 2, 2, ed) - 1] = sro[A2(edges, 2, 2, ed) - 1] - (c1 + c2);
    }
    if (ale != 0) for (int n = 1; n <= myN; n++) sro[n - 1] = sro[n - 1] - water_flux[n - 1] * A2(areasvol, nl, uln[n - 1], n);
    memcpy(hbar_old, hbar, sizeof(double) * N);
    for (int n = 1; n <= myN; n++) hbar[n - 1] = hbar_old[n - 1] + sro[n - 1] * dt / A2(areasvol, nl, uln[n - 1], n);
    for (int e = 1; e <= myE; e++) {
        if (ule[e - 1] > 1) dhe[e - 1] = 0.0;
        else {
            int n1 = A2(en, 3, 1, e) - 1, n2 = A2(en, 3, 2, e) - 1, n3 = A2(en, 3, 3, e) - 1;
            dhe[e - 1] = ((hbar[n1] - hbar_old[n1]) + (hbar[n2] - hbar_old[n2]) + (hbar[n3] - hbar_old[n3])) / 3.0;
        }
    }
}

// vert_vel_ale (oce_ale.F90:1692-2204): linfs, plain zlevel (no local-zstar fallback) and zstar
void ale_vert_vel(Host& h) {
    const int myN = h.pi("myDim_nod2D"), N = myN + h.pi("eDim_nod2D"), myEd = h.pi("myDim_edge2D"), nl = h.pi("nl");
    const int L = nl - 1, ale = h.pi("which_ale");
    const double dt = h.pd("dt");
    const bool w_split = h.pi("w_split"); const double w_max_cfl = h.pd("w_max_cfl");
    const int* edges = h.i("edges"); const int* edge_tri = h.i("edge_tri");
    const int* nle = h.i("nlevels"); const int* ule = h.i("ulevels");
    const int* nln = h.i("nlevels_nod2D"); const int* uln = h.i("ulevels_nod2D"); const int* nlmin = h.i("nlevels_nod2D_min");
    const double* ecd = h.d("edge_cross_dxdy"); const double* UV = h.d("UV"); const double* helem = h.d("helem");
    const double* area = h.d("area"); const double* water_flux = h.d("water_flux");
    const double* hbar = h.d("hbar"); const double* hbar_old = h.d("hbar_old"); const double* z3 = h.d("zbar_3d_n");
    const double* hnode = h.d("hnode"); double* hnode_new = h.d("hnode_new");
    double* W = h.d("Wvel"); double* We = h.d("Wvel_e"); double* Wi = h.d("Wvel_i"); double* CFL = h.d("CFL_z");
    for (int64_t q = 0; q < int64_t(nl) * N; q++) W[q] = 0.0;
    for (int ed = 1; ed <= myEd; ed++) {
        int n1 = A2(edges, 2, 1, ed), n2 = A2(edges, 2, 2, ed);
        int e1 = A2(edge_tri, 2, 1, ed), e2 = A2(edge_tri, 2, 2, ed);
        double dX1 = A2(ecd, 4, 1, ed), dY1 = A2(ecd, 4, 2, ed);
        for (int nz = nle[e1 - 1] - 1; nz >= ule[e1 - 1]; nz--) {
            double c1 = (A3(UV, 2, L, 2, nz, e1) * dX1 - A3(UV, 2, L, 1, nz, e1) * dY1) * A2(helem, L, nz, e1);
            A2(W, nl, nz, n1) = A2(W, nl, nz, n1) + c1;
            A2(W, nl, nz, n2) = A2(W, nl, nz, n2) - c1;
        }
        if (e2 > 0) {
            double dX2 = A2(ecd, 4, 3, ed), dY2 = A2(ecd, 4, 4, ed);
            for (int nz = nle[e2 - 1] - 1; nz >= ule[e2 - 1]; nz--) {
                double c2 = -(A3(UV, 2, L, 2, nz, e2) * dX2 - A3(UV, 2, L, 1, nz, e2) * dY2) * A2(helem, L, nz, e2);
                A2(W, nl, nz, n1) = A2(W, nl, nz, n1) + c2;
                A2(W, nl, nz, n2) = A2(W, nl, nz, n2) - c2;
            }
        }
    }
    for (int n = 1; n <= myN; n++)
        for (int nz = nln[n - 1] - 1; nz >= uln[n - 1]; nz--) A2(W, nl, nz, n) = A2(W, nl, nz, n) + A2(W, nl, nz + 1, n);
    for (int n = 1; n <= myN; n++)
        for (int nz = uln[n - 1]; nz <= nln[n - 1] - 1; nz++) A2(W, nl, nz, n) = A2(W, nl, nz, n) / A2(area, nl, nz, n);
    if (ale == 1) {          // :2005-2020 normal zlevel case
        for (int n = 1; n <= myN; n++) {
            int nzmin = uln[n - 1];
            if (nzmin == 1) {
                double dh = hbar[n - 1] - hbar_old[n - 1];
                A2(W, nl, nzmin, n) = A2(W, nl, nzmin, n) - dh / dt;
                A2(hnode_new, L, nzmin, n) = A2(hnode, L, nzmin, n) + dh;
            }
            A2(W, nl, nzmin, n) = A2(W, nl, nzmin, n) - water_flux[n - 1];
        }
    } else if (ale == 2) {   // :2028-2086
        for (int n = 1; n <= myN; n++) {
            int nzmin = uln[n - 1], nzmax = nlmin[n - 1] - 1;
            if (nzmin == 1) {
                double dd1 = A2(z3, nl, nzmax, n);
                double dd = A2(z3, nl, nzmin, n) - dd1;
                dd = (hbar[n - 1] - hbar_old[n - 1]) / dd;
                double dddt = dd / dt;
                for (int nz = nzmin; nz <= nzmax - 1; nz++) {
                    A2(W, nl, nz, n) = A2(W, nl, nz, n) - (A2(z3, nl, nz, n) - dd1) * dddt;
                    A2(hnode_new, L, nz, n) = A2(hnode, L, nz, n) + (A2(z3, nl, nz, n) - A2(z3, nl, nz + 1, n)) * dd;
                }
            }
            A2(W, nl, nzmin, n) = A2(W, nl, nzmin, n) - water_flux[n - 1];
        }
    }
    // (halo exchange of Wvel / hnode_new is the caller's business in the multi-partition case)
    for (int n = 1; n <= N; n++) A2(CFL, nl, 1, n) = 0.0;
    for (int n = 1; n <= N; n++)
        for (int nz = uln[n - 1]; nz <= nln[n - 1] - 1; nz++) {
            double c1 = fabs(A2(W, nl, nz, n) * dt / A2(hnode_new, L, nz, n));
            double c2 = fabs(A2(W, nl, nz + 1, n) * dt / A2(hnode_new, L, nz, n));
            A2(CFL, nl, nz, n) = A2(CFL, nl, nz, n) + c1;
            A2(CFL, nl, nz + 1, n) = c2;
        }
    for (int n = 1; n <= N; n++)
        for (int nz = uln[n - 1]; nz <= nln[n - 1]; nz++) {
            double c1 = 1.0, c2 = 0.0;
            if (w_split && A2(CFL, nl, nz, n) > w_max_cfl) {
                double dd = std::max(A2(CFL, nl, nz, n) - w_max_cfl, 0.0) / std::max(w_max_cfl, 1.e-12);
                c1 = 1.0 / (1.0 + dd); c2 = dd / (1.0 + dd);
            }
            A2(We, nl, nz, n) = c1 * A2(W, nl, nz, n);
            A2(Wi, nl, nz, n) = c2 * A2(W, nl, nz, n);
        }
}

// update_thickness_ale (oce_ale.F90:800-993): zstar and plain zlevel
void ale_update_thickness(Host& h) {
    const int N = h.pi("myDim_nod2D") + h.pi("eDim_nod2D"), myE = h.pi("myDim_elem2D"), nl = h.pi("nl");
    const int L = nl - 1, ale = h.pi("which_ale");
    const int* en = h.i("elem2D_nodes"); const int* nle = h.i("nlevels"); const int* ule = h.i("ulevels");
    const int* uln = h.i("ulevels_nod2D"); const int* nlmin = h.i("nlevels_nod2D_min"); const int* ulmax = h.i("ulevels_nod2D_max");
    double* hnode = h.d("hnode"); const double* hnode_new = h.d("hnode_new"); double* helem = h.d("helem");
    double* z3 = h.d("zbar_3d_n"); double* Z3 = h.d("Z_3d_n");
    if (ale == 1) {
        for (int e = 1; e <= myE; e++) {
            int nzmin = ule[e - 1]; if (nzmin > 1) continue;
            int n1 = A2(en, 3, 1, e), n2 = A2(en, 3, 2, e), n3 = A2(en, 3, 3, e);
            A2(helem, L, nzmin, e) = (A2(hnode_new, L, nzmin, n1) + A2(hnode_new, L, nzmin, n2) + A2(hnode_new, L, nzmin, n3)) / 3.0;
        }
        for (int n = 1; n <= N; n++) {
            int nzmin = ulmax[n - 1]; if (nzmin > 1) continue;
            A2(hnode, L, nzmin, n) = A2(hnode_new, L, nzmin, n);
            A2(z3, nl, nzmin, n) = A2(z3, nl, nzmin + 1, n) + A2(hnode_new, L, nzmin, n);
            A2(Z3, L, nzmin, n) = A2(z3, nl, nzmin + 1, n) + A2(hnode_new, L, nzmin, n) / 2.0;
        }
    } else if (ale == 2) {
        for (int n = 1; n <= N; n++) {
            int nzmin = uln[n - 1], nzmax = nlmin[n - 1] - 2;
            if (nzmin > 1) continue;
            for (int nz = nzmax; nz >= nzmin; nz--) {
                A2(hnode, L, nz, n) = A2(hnode_new, L, nz, n);
                A2(z3, nl, nz, n) = A2(z3, nl, nz + 1, n) + A2(hnode_new, L, nz, n);
                A2(Z3, L, nz, n) = A2(z3, nl, nz + 1, n) + A2(hnode_new, L, nz, n) / 2.0;
            }
        }
        for (int e = 1; e <= myE; e++) {
            int nzmin = ule[e - 1], nzmax = nle[e - 1] - 1;
            if (nzmin > 1) continue;
            int n1 = A2(en, 3, 1, e), n2 = A2(en, 3, 2, e), n3 = A2(en, 3, 3, e);
            for (int nz = nzmin; nz <= nzmax - 1; nz++)
                A2(helem, L, nz, e) = (A2(hnode, L, nz, n1) + A2(hnode, L, nz, n2) + A2(hnode, L, nz, n3)) / 3.0;
        }
    }
}

}  // namespace fh
