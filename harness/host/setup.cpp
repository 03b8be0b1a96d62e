// Setup-time tables of the hot path that the Fortran host owns and passes over the C ABI:
// MUSCL up/downwind triangles + nboundary_lay (src/oce_muscl_adv.F90), the SSH stiffness matrix in CSR
// (src/oce_ale.F90 init_stiff_mat_ale / update_stiff_mat_ale), the SSH right-hand side
// (compute_ssh_rhs_ale) and the state arrays of array_setup (src/oce_setup_step.F90:179-340).
#include "host.hpp"

namespace fh {
#define A2(p, ld, i, j) (p)[(int64_t)((j) - 1) * (ld) + ((i) - 1)]
#define A3(p, l1, l2, i, j, k) (p)[((int64_t)((k) - 1) * (l2) + ((j) - 1)) * (l1) + ((i) - 1)]

void state_alloc(Host& h, int num_tracers) {          // oce_setup_step.F90:205-323
    const int N = h.pi("myDim_nod2D") + h.pi("eDim_nod2D");
    const int E = h.pi("myDim_elem2D") + h.pi("eDim_elem2D");
    const int Eall = E + h.pi("eXDim_elem2D");
    const int nl = h.pi("nl");
    h.p["num_tracers"] = num_tracers;
    h.D("UV", {2, nl - 1, E}); h.D("UV_rhs", {2, nl - 1, E});
    h.D("eta_n", {N}); h.D("d_eta", {N}); h.D("ssh_rhs", {N});
    h.D("Wvel", {nl, N}); h.D("Wvel_e", {nl, N}); h.D("Wvel_i", {nl, N}); h.D("CFL_z", {nl, N});
    h.D("tr_arr", {nl - 1, N, num_tracers}); h.D("tr_arr_old", {nl - 1, N, num_tracers});
    h.D("del_ttf", {nl - 1, N});
    h.D("heat_flux", {N}); h.D("water_flux", {N}); h.D("relax_salt", {N}); h.D("virtual_salt", {N}); h.D("real_salt_flux", {N});
    h.D("Kv", {nl, N});
    h.D("tr_xy", {2, nl - 1, Eall}); h.D("tr_z", {nl, N});
    h.D("slope_tapered", {3, nl - 1, N});
    h.D("sw_3d", {nl, N});
    double* Ki = h.D("Ki", {nl - 1, N});
    const double K_hor = h.p.count("K_hor") ? h.pd("K_hor") : 10.0;
    const double* res = h.d("mesh_resolution");
    for (int n = 1; n <= N; n++) {                  // :329-331  Ki = K_hor*(res/100km)**2
        double r = res[n - 1] / 100000.0;
        double v = K_hor * (r * r);
        for (int nz = 1; nz <= nl - 1; nz++) A2(Ki, nl - 1, nz, n) = v;
    }
}

// init_stiff_mat_ale (oce_ale.F90:1088-1354).  colind keeps LOCAL 1-based node numbers in "colind_loc";
// "colind" holds the global PE-contiguous numbering (for one partition both coincide).
void stiff_mat_init(Host& h) {
    const int myN = h.pi("myDim_nod2D"), N = myN + h.pi("eDim_nod2D"), myEd = h.pi("myDim_edge2D"), nl = h.pi("nl");
    const double dt = h.pd("dt"), alpha = h.pd("alpha"), theta = h.pd("theta");
    const int* edges = h.i("edges"); const int* edge_tri = h.i("edge_tri"); const int* en = h.i("elem2D_nodes");
    const int* uln = h.i("ulevels_nod2D");
    const double* gs = h.d("gradient_sca"); const double* ecd = h.d("edge_cross_dxdy");
    const double* zeb = h.d("zbar_e_bot"); const double* zes = h.d("zbar_e_srf"); const double* areasvol = h.d("areasvol");
    std::vector<int> n_num(N + 1, 0);
    const int MAXNB = 32;
    std::vector<int> n_pos(size_t(MAXNB) * myN, 0);
    for (int n = 1; n <= myN; n++) { n_num[n] = 1; n_pos[size_t(n - 1) * MAXNB] = n; }
    for (int ed = 1; ed <= myEd; ed++) {
        int n1 = A2(edges, 2, 1, ed), n2 = A2(edges, 2, 2, ed);
        if (n1 <= myN) { if (n_num[n1] >= MAXNB) throw std::runtime_error("too many neighbours"); n_pos[size_t(n1 - 1) * MAXNB + n_num[n1]] = n2; n_num[n1]++; }
        if (n2 <= myN) { if (n_num[n2] >= MAXNB) throw std::runtime_error("too many neighbours"); n_pos[size_t(n2 - 1) * MAXNB + n_num[n2]] = n1; n_num[n2]++; }
    }
    int* rowptr = h.I("ssh_rowptr", {myN + 1});
    rowptr[0] = 1;
    int maxrow = 0;
    for (int n = 1; n <= myN; n++) { rowptr[n] = rowptr[n - 1] + n_num[n]; maxrow = std::max(maxrow, n_num[n]); }
    const int nza = rowptr[myN] - 1;
    h.p["ssh_nza"] = nza; h.p["nn_max"] = maxrow;
    int* colind_loc = h.I("ssh_colind_loc", {nza}); int* colind = h.I("ssh_colind", {nza});
    double* values = h.D("ssh_values", {nza});
    for (int n = 1; n <= myN; n++)
        for (int k = 0; k < n_num[n]; k++) colind_loc[rowptr[n - 1] - 1 + k] = n_pos[size_t(n - 1) * MAXNB + k];
    // nn_num / nn_pos of muscl_adv_init (oce_muscl_adv.F90:63-90) are the same table
    int* nn_num = h.I("nn_num", {myN}); int* nn_pos = h.I("nn_pos", {maxrow, myN});
    for (int n = 1; n <= myN; n++) { nn_num[n - 1] = n_num[n]; for (int k = 0; k < n_num[n]; k++) A2(nn_pos, maxrow, k + 1, n) = n_pos[size_t(n - 1) * MAXNB + k]; }
    const double factor = G_ACC * dt * alpha * theta;
    std::vector<int> where(N + 1, 0);
    for (int ed = 1; ed <= myEd; ed++) {
        for (int i = 1; i <= 2; i++) {
            int el = A2(edge_tri, 2, i, ed);
            if (el < 1) continue;
            int eln[3] = {A2(en, 3, 1, el), A2(en, 3, 2, el), A2(en, 3, 3, el)};
            double fy[3];
            for (int k = 0; k < 3; k++)
                fy[k] = (zeb[el - 1] - zes[el - 1]) * (A2(gs, 6, k + 1, el) * A2(ecd, 4, 2 * i, ed) - A2(gs, 6, k + 4, el) * A2(ecd, 4, 2 * i - 1, ed));
            if (i == 2) for (int k = 0; k < 3; k++) fy[k] = -fy[k];
            for (int j = 1; j <= 2; j++) {
                int row = A2(edges, 2, j, ed);
                if (row > myN) continue;
                for (int q = rowptr[row - 1]; q <= rowptr[row] - 1; q++) where[colind_loc[q - 1]] = q;
                for (int k = 0; k < 3; k++) {
                    int pos = where[eln[k]];
                    if (j == 1) values[pos - 1] = values[pos - 1] + fy[k] * factor;
                    else values[pos - 1] = values[pos - 1] - fy[k] * factor;
                }
            }
        }
    }
    for (int row = 1; row <= myN; row++) {
        if (uln[row - 1] > 1) continue;
        int off = rowptr[row - 1];
        values[off - 1] = values[off - 1] + A2(areasvol, nl, uln[row - 1], row) / dt;
    }
    // global PE-contiguous numbering: for one partition the identity; partitions go through "pe_contig_map"
    if (h.has("pe_contig_map")) {
        const int* map = h.i("pe_contig_map"); const int* ml = h.i("myList_nod2D");
        for (int q = 0; q < nza; q++) colind[q] = map[ml[colind_loc[q] - 1] - 1];
    } else
        for (int q = 0; q < nza; q++) colind[q] = colind_loc[q];
}

// update_stiff_mat_ale (oce_ale.F90:1371-1470)
void stiff_mat_update(Host& h) {
    const int myN = h.pi("myDim_nod2D"), N = myN + h.pi("eDim_nod2D"), myEd = h.pi("myDim_edge2D");
    const double dt = h.pd("dt"), alpha = h.pd("alpha"), theta = h.pd("theta");
    const int* edges = h.i("edges"); const int* edge_tri = h.i("edge_tri"); const int* en = h.i("elem2D_nodes");
    const double* gs = h.d("gradient_sca"); const double* ecd = h.d("edge_cross_dxdy"); const double* dhe = h.d("dhe");
    const int* rowptr = h.i("ssh_rowptr"); const int* nn_pos = h.i("nn_pos"); const int maxrow = h.pi("nn_max");
    double* values = h.d("ssh_values");
    const double factor = G_ACC * dt * alpha * theta;
    std::vector<int> n_num(N + 1, 0);
    for (int ed = 1; ed <= myEd; ed++) {
        for (int j = 1; j <= 2; j++) {
            int row = A2(edges, 2, j, ed);
            if (row > myN) continue;
            int offset = rowptr[row - 1] - rowptr[0];
            for (int n = 1; n <= rowptr[row] - rowptr[row - 1]; n++) n_num[A2(nn_pos, maxrow, n, row)] = offset + n;
            for (int i = 1; i <= 2; i++) {
                int elem = A2(edge_tri, 2, i, ed);
                if (elem < 1) continue;
                for (int k = 0; k < 3; k++) {
                    double fy = -dhe[elem - 1] * (A2(gs, 6, k + 1, elem) * A2(ecd, 4, 2 * i, ed) - A2(gs, 6, k + 4, elem) * A2(ecd, 4, 2 * i - 1, ed));
                    if (i == 2) fy = -fy;
                    if (j == 2) fy = -fy;
                    int pos = n_num[A2(en, 3, k + 1, elem)];
                    values[pos - 1] = values[pos - 1] + fy * factor;
                }
            }
        }
    }
}

// compute_ssh_rhs_ale (oce_ale.F90:1478-1572)
void compute_ssh_rhs(Host& h) {
    const int myN = h.pi("myDim_nod2D"), N = myN + h.pi("eDim_nod2D"), myEd = h.pi("myDim_edge2D"), nl = h.pi("nl");
    const int L = nl - 1, ale = h.pi("which_ale");
    const double alpha = h.pd("alpha");
    const int* edges = h.i("edges"); const int* edge_tri = h.i("edge_tri");
    const int* nle = h.i("nlevels"); const int* ule = h.i("ulevels"); const int* uln = h.i("ulevels_nod2D");
    const double* ecd = h.d("edge_cross_dxdy"); const double* UV = h.d("UV"); const double* UVr = h.d("UV_rhs");
    const double* helem = h.d("helem"); const double* areasvol = h.d("areasvol");
    const double* water_flux = h.d("water_flux"); const double* sro = h.d("ssh_rhs_old");
    double* ssh_rhs = h.d("ssh_rhs");
    for (int n = 0; n < N; n++) ssh_rhs[n] = 0.0;
    for (int ed = 1; ed <= myEd; ed++) {
        int e1 = A2(edge_tri, 2, 1, ed), e2 = A2(edge_tri, 2, 2, ed);
        double c1 = 0.0, c2 = 0.0;
        double dX1 = A2(ecd, 4, 1, ed), dY1 = A2(ecd, 4, 2, ed);
        for (int nz = ule[e1 - 1]; nz <= nle[e1 - 1] - 1; nz++)
            c1 = c1 + alpha * ((A3(UV, 2, L, 2, nz, e1) + A3(UVr, 2, L, 2, nz, e1)) * dX1 -
                               (A3(UV, 2, L, 1, nz, e1) + A3(UVr, 2, L, 1, nz, e1)) * dY1) * A2(helem, L, nz, e1);
        if (e2 > 0) {
            double dX2 = A2(ecd, 4, 3, ed), dY2 = A2(ecd, 4, 4, ed);
            for (int nz = ule[e2 - 1]; nz <= nle[e2 - 1] - 1; nz++)
                c2 = c2 - alpha * ((A3(UV, 2, L, 2, nz, e2) + A3(UVr, 2, L, 2, nz, e2)) * dX2 -
                                   (A3(UV, 2, L, 1, nz, e2) + A3(UVr, 2, L, 1, nz, e2)) * dY2) * A2(helem, L, nz, e2);
        }
        ssh_rhs[A2(edges, 2, 1, ed) - 1] = ssh_rhs[A2(edges, 2, 1, ed) - 1] + (c1 + c2);
        ssh_rhs[A2(edges, 2, 2, ed) - 1] = ssh_rhs[A2(edges, 2, 2, ed) - 1] - (c1 + c2);
    }
    if (ale != 0) {
        for (int n = 1; n <= myN; n++) {
            int nzmin = uln[n - 1];
            if (nzmin > 1) ssh_rhs[n - 1] = ssh_rhs[n - 1] - alpha * water_flux[n - 1] * A2(areasvol, nl, nzmin, n);
            else ssh_rhs[n - 1] = ssh_rhs[n - 1] - alpha * water_flux[n - 1] * A2(areasvol, nl, nzmin, n) + (1.0 - alpha) * sro[n - 1];
        }
    } else {
        for (int n = 1; n <= myN; n++) { if (uln[n - 1] > 1) continue; ssh_rhs[n - 1] = ssh_rhs[n - 1] + (1.0 - alpha) * sro[n - 1]; }
    }
}

// muscl_adv_init + find_up_downwind_triangles (oce_muscl_adv.F90:27-281)
void muscl_adv_init(Host& h) {
    const int myN = h.pi("myDim_nod2D"), N = myN + h.pi("eDim_nod2D"), myEd = h.pi("myDim_edge2D"), nl = h.pi("nl");
    const double cyc = h.pd("cyclic_length");
    const int* edges = h.i("edges"); const int* edge_tri = h.i("edge_tri"); const int* en = h.i("elem2D_nodes");
    const int* nle = h.i("nlevels");
    const int* nien = h.i("nod_in_elem2D_num"); const int* nie = h.i("nod_in_elem2D"); const int maxn = h.pi("nod_in_elem2D_max");
    const double* coord = h.d("coord_nod2D");
    int* updn = h.I("edge_up_dn_tri", {2, myEd});
    for (int n = 1; n <= myEd; n++) {
        int edn[2] = {A2(edges, 2, 1, n), A2(edges, 2, 2, n)};
        double x[2] = {A2(coord, 2, 1, edn[1]) - A2(coord, 2, 1, edn[0]), A2(coord, 2, 2, edn[1]) - A2(coord, 2, 2, edn[0])};
        if (x[0] > cyc / 2.0) x[0] -= cyc;
        if (x[0] < -cyc / 2.0) x[0] += cyc;
        for (int side = 0; side < 2; side++) {
            x[0] = -x[0]; x[1] = -x[1];           // side 0: looks for the triangle containing -x at node 1; side 1: +x at node 2
            int node = edn[side];
            for (int k = 1; k <= nien[node - 1]; k++) {
                int elem = A2(nie, maxn, k, node);
                int p1 = A2(en, 3, 1, elem), p2 = A2(en, 3, 2, elem), p3 = A2(en, 3, 3, elem);
                double b[2], c[2];
                auto cx = [&](int p) { return A2(coord, 2, 1, p); };
                auto cy = [&](int p) { return A2(coord, 2, 2, p); };
                if (p1 == node) { b[0] = cx(p2) - cx(p1); b[1] = cy(p2) - cy(p1); c[0] = cx(p3) - cx(p1); c[1] = cy(p3) - cy(p1); }
                else if (p2 == node) { b[0] = cx(p1) - cx(p2); b[1] = cy(p1) - cy(p2); c[0] = cx(p3) - cx(p2); c[1] = cy(p3) - cy(p2); }
                else { b[0] = cx(p1) - cx(p3); b[1] = cy(p1) - cy(p3); c[0] = cx(p2) - cx(p3); c[1] = cy(p2) - cy(p3); }
                if (b[0] > cyc / 2.0) b[0] -= cyc;
                if (b[0] < -cyc / 2.0) b[0] += cyc;
                if (c[0] > cyc / 2.0) c[0] -= cyc;
                if (c[0] < -cyc / 2.0) c[0] += cyc;
                double cr = c[0] * c[0] + c[1] * c[1];
                double bx = (b[0] * c[0] + b[1] * c[1]) / cr;
                double by = (-b[0] * c[1] + b[1] * c[0]) / cr;
                double xx = (x[0] * c[0] + x[1] * c[1]) / cr;
                double xy = (-x[0] * c[1] + x[1] * c[0]) / cr;
                double ab = atan2(by, bx), ax = atan2(xy, xx);
                // the Fortran 'cycle's the k loop after a hit, i.e. a later matching element overwrites an earlier one
                if (ab > 0.0 && ax > 0.0 && ax < ab) { A2(updn, 2, side + 1, n) = elem; continue; }
                if (ab < 0.0 && ax < 0.0 && ax > ab) { A2(updn, 2, side + 1, n) = elem; continue; }
                if (ab == ax || ax == 0.0) { A2(updn, 2, side + 1, n) = elem; continue; }
            }
        }
    }
    int* nbl = h.I("nboundary_lay", {N});
    for (int n = 0; n < N; n++) nbl[n] = nl - 1;
    for (int n = 1; n <= myEd; n++) {                 // :76-104
        int n1 = A2(edges, 2, 1, n), n2 = A2(edges, 2, 2, n);
        int t1 = A2(edge_tri, 2, 1, n), t2 = A2(edge_tri, 2, 2, n);
        if (t1 <= 0 || t2 <= 0) { nbl[n1 - 1] = 0; nbl[n2 - 1] = 0; }
        else {
            int mn = std::min(nle[t1 - 1], nle[t2 - 1]) - 1;
            nbl[n1 - 1] = std::min(nbl[n1 - 1], mn);
            nbl[n2 - 1] = std::min(nbl[n2 - 1], mn);
        }
    }
}

}  // namespace fh
