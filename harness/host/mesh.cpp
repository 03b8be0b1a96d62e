// Mesh ingestion and derived geometry — the caller-side mirror of the reference's mesh_setup.
// Every block cites the Fortran it follows; loop order and operation order are kept so that the
// arrays handed to the C ABI are the ones the Fortran driver would hand over.
#include "host.hpp"

namespace fh {

// 1-based accessors over Fortran column-major storage
#define A2(p, ld, i, j) (p)[(int64_t)((j) - 1) * (ld) + ((i) - 1)]

static void rot_matrix(double al, double be, double ga, double m[3][3]) {   // gen_modules_rotate_grid.F90 set_mesh_transform_matrix
    m[0][0] = cos(ga) * cos(al) - sin(ga) * cos(be) * sin(al);
    m[0][1] = cos(ga) * sin(al) + sin(ga) * cos(be) * cos(al);
    m[0][2] = sin(ga) * sin(be);
    m[1][0] = -sin(ga) * cos(al) - cos(ga) * cos(be) * sin(al);
    m[1][1] = -sin(ga) * sin(al) + cos(ga) * cos(be) * cos(al);
    m[1][2] = cos(ga) * sin(be);
    m[2][0] = sin(be) * sin(al);
    m[2][1] = -sin(be) * cos(al);
    m[2][2] = cos(be);
}
static void g2r(const double m[3][3], double glon, double glat, double& rlon, double& rlat) {
    double xg = cos(glat) * cos(glon), yg = cos(glat) * sin(glon), zg = sin(glat);
    double xr = m[0][0] * xg + m[0][1] * yg + m[0][2] * zg;
    double yr = m[1][0] * xg + m[1][1] * yg + m[1][2] * zg;
    double zr = m[2][0] * xg + m[2][1] * yg + m[2][2] * zg;
    rlat = asin(zr);
    rlon = (yr == 0.0 && xr == 0.0) ? 0.0 : atan2(yr, xr);
}
static void r2g(const double m[3][3], double& glon, double& glat, double rlon, double rlat) {
    double xr = cos(rlat) * cos(rlon), yr = cos(rlat) * sin(rlon), zr = sin(rlat);
    double xg = m[0][0] * xr + m[1][0] * yr + m[2][0] * zr;
    double yg = m[0][1] * xr + m[1][1] * yr + m[2][1] * zr;
    double zg = m[0][2] * xr + m[1][2] * yr + m[2][2] * zr;
    glat = asin(zg);
    glon = (yg == 0.0 && xg == 0.0) ? 0.0 : atan2(yg, xg);
}

// read_mesh (oce_mesh.F90:147-692) for the one-partition case: every node/element/edge is "own".
void mesh_set_global(Host& h, int nod2D, const double* lon_deg, const double* lat_deg, int elem2D, const int* tri,
                     int nl, const double* zbar_in, const double* depth_in, const int* nlevels_nod, const int* nlevels_elem,
                     int edge2D, int edge2D_in, const int* edges_in, const int* edge_tri_in,
                     int force_rotation, double cyclic_deg, double alphaE, double betaE, double gammaE) {
    h.p["nod2D"] = nod2D; h.p["elem2D"] = elem2D; h.p["nl"] = nl;
    h.p["npes"] = 1; h.p["mype"] = 0;
    h.p["myDim_nod2D"] = nod2D; h.p["eDim_nod2D"] = 0;
    h.p["myDim_elem2D"] = elem2D; h.p["eDim_elem2D"] = 0; h.p["eXDim_elem2D"] = 0;
    h.p["cyclic_length"] = cyclic_deg * RAD;                  // gen_model_setup.F90:49
    h.p["force_rotation"] = force_rotation;
    h.p["alphaEuler"] = alphaE * RAD; h.p["betaEuler"] = betaE * RAD; h.p["gammaEuler"] = gammaE * RAD;
    double m[3][3];
    rot_matrix(alphaE * RAD, betaE * RAD, gammaE * RAD, m);

    double* coord = h.D("coord_nod2D", {2, nod2D});
    for (int n = 1; n <= nod2D; n++) {                        // oce_mesh.F90:337-353
        double x = lon_deg[n - 1] * RAD, y = lat_deg[n - 1] * RAD;
        if (force_rotation) { double rx = x, ry = y; g2r(m, rx, ry, x, y); }
        A2(coord, 2, 1, n) = x; A2(coord, 2, 2, n) = y;
    }
    int* en = h.I("elem2D_nodes", {3, elem2D});
    memcpy(en, tri, sizeof(int) * 3 * size_t(elem2D));
    double* zbar = h.D("zbar", {nl});
    for (int k = 0; k < nl; k++) zbar[k] = zbar_in[k];
    if (zbar[1] > 0) for (int k = 0; k < nl; k++) zbar[k] = -zbar[k];   // :496
    double* Z = h.D("Z", {nl - 1});
    for (int k = 0; k < nl - 1; k++) { Z[k] = zbar[k] + zbar[k + 1]; Z[k] = 0.5 * Z[k]; }   // :498-499
    double* depth = h.D("depth", {nod2D});
    for (int n = 0; n < nod2D; n++) {                         // :525-533
        double x = depth_in[n];
        if (x > 0) x = -x;
        if (x > zbar[4]) x = zbar[4];
        depth[n] = x;
    }
    int* ml = h.I("myList_nod2D", {nod2D}); for (int n = 0; n < nod2D; n++) ml[n] = n + 1;
    int* me = h.I("myList_elem2D", {elem2D}); for (int n = 0; n < elem2D; n++) me[n] = n + 1;
    int* part = h.I("part", {2}); part[0] = 1; part[1] = nod2D + 1;
    if (nlevels_nod && nlevels_elem) {                        // find_levels, oce_mesh.F90:699-893
        int* nle = h.I("nlevels", {elem2D}); memcpy(nle, nlevels_elem, sizeof(int) * size_t(elem2D));
        int* nln = h.I("nlevels_nod2D", {nod2D}); memcpy(nln, nlevels_nod, sizeof(int) * size_t(nod2D));
        int* ule = h.I("ulevels", {elem2D}); for (int n = 0; n < elem2D; n++) ule[n] = 1;
        int* uln = h.I("ulevels_nod2D", {nod2D}); for (int n = 0; n < nod2D; n++) uln[n] = 1;
    }
    if (edges_in && edge_tri_in) {                            // load_edges (a), oce_mesh.F90:1456-1599
        h.p["edge2D"] = edge2D; h.p["edge2D_in"] = edge2D_in;
        h.p["myDim_edge2D"] = edge2D; h.p["eDim_edge2D"] = 0;
        int* ed = h.I("edges", {2, edge2D}); memcpy(ed, edges_in, sizeof(int) * 2 * size_t(edge2D));
        int* et = h.I("edge_tri", {2, edge2D});
        for (int64_t q = 0; q < 2 * int64_t(edge2D); q++) et[q] = edge_tri_in[q] < 0 ? 0 : edge_tri_in[q];   // :1566-1568
        int* mle = h.I("myList_edge2D", {edge2D}); for (int n = 0; n < edge2D; n++) mle[n] = n + 1;
    }
}

static void elem_center(Host& h, int elem, double& x, double& y) {     // oce_mesh.F90:1818-1838
    const int* en = h.i("elem2D_nodes"); const double* coord = h.d("coord_nod2D");
    const double cyc = h.pd("cyclic_length");
    double ax[3];
    for (int k = 0; k < 3; k++) ax[k] = A2(coord, 2, 1, A2(en, 3, k + 1, elem));
    double amin = std::min(ax[0], std::min(ax[1], ax[2]));
    for (int k = 0; k < 3; k++) {
        if (ax[k] - amin >= cyc / 2.0) ax[k] -= cyc;
        if (ax[k] - amin < -cyc / 2.0) ax[k] += cyc;
    }
    x = (ax[0] + ax[1] + ax[2]) / 3.0;
    y = (A2(coord, 2, 2, A2(en, 3, 1, elem)) + A2(coord, 2, 2, A2(en, 3, 2, elem)) + A2(coord, 2, 2, A2(en, 3, 3, elem))) / 3.0;
}

static void test_tri(Host& h) {                                        // oce_mesh.F90:1373-1417
    int* en = h.i("elem2D_nodes"); const double* coord = h.d("coord_nod2D");
    const double cyc = h.pd("cyclic_length");
    const int ne = h.pi("myDim_elem2D") + h.pi("eDim_elem2D") + h.pi("eXDim_elem2D");
    for (int n = 1; n <= ne; n++) {
        int e1 = A2(en, 3, 1, n), e2 = A2(en, 3, 2, n), e3 = A2(en, 3, 3, n);
        double a[2] = {A2(coord, 2, 1, e1), A2(coord, 2, 2, e1)};
        double b[2] = {A2(coord, 2, 1, e2) - a[0], A2(coord, 2, 2, e2) - a[1]};
        double c[2] = {A2(coord, 2, 1, e3) - a[0], A2(coord, 2, 2, e3) - a[1]};
        trim_cyclic(b[0], cyc); trim_cyclic(c[0], cyc);
        double r = b[0] * c[1] - b[1] * c[0];
        if (r > 0.0) { A2(en, 3, 2, n) = e3; A2(en, 3, 3, n) = e2; }
    }
}

// find_edges_ini (fvom_init.F90:315-650), triangles only: builds edges / edge_tri in the reference's
// order (interior edges first, then boundary edges; first triangle on the left of the edge).
void mesh_find_edges(Host& h) {
    const int nod2D = h.pi("nod2D"), elem2D = h.pi("elem2D");
    test_tri(h);   // fesom_ini runs test_tri_ini before find_edges_ini (fvom_init.F90:118-119)
    const int* en = h.i("elem2D_nodes"); const double* coord = h.d("coord_nod2D");
    const double cyc = h.pd("cyclic_length");
    std::vector<int> ne_ptr(nod2D + 2, 0);
    for (int e = 1; e <= elem2D; e++) for (int q = 1; q <= 3; q++) ne_ptr[A2(en, 3, q, e) + 1]++;
    for (int n = 1; n <= nod2D; n++) ne_ptr[n + 1] += ne_ptr[n];
    std::vector<int> ne_pos(ne_ptr[nod2D + 1]), fill(nod2D + 2, 0);
    for (int e = 1; e <= elem2D; e++) for (int q = 1; q <= 3; q++) { int n = A2(en, 3, q, e); ne_pos[ne_ptr[n] + fill[n]++] = e; }
    // node neighbours in discovery order
    std::vector<int> nn_ptr(nod2D + 2, 0), nn_pos; nn_pos.reserve(size_t(nod2D) * 7);
    std::vector<char> aux(nod2D + 1, 0);
    for (int n = 1; n <= nod2D; n++) {
        nn_ptr[n] = int(nn_pos.size());
        int start = nn_ptr[n];
        for (int k = ne_ptr[n]; k < ne_ptr[n + 1]; k++) {
            int e = ne_pos[k];
            for (int q = 1; q <= 3; q++) {
                int m = A2(en, 3, q, e);
                if (m == n || aux[m]) continue;
                aux[m] = 1; nn_pos.push_back(m);
            }
        }
        for (size_t k = start; k < nn_pos.size(); k++) aux[nn_pos[k]] = 0;
    }
    nn_ptr[nod2D + 1] = int(nn_pos.size());
    std::vector<int> ed, et;
    int edge2D_in = 0;
    for (int pass = 2; pass >= 1; pass--) {
        for (int n = 1; n <= nod2D; n++)
            for (int q = nn_ptr[n]; q < nn_ptr[n + 1]; q++) {
                int node = nn_pos[q];
                if (node < n) continue;
                int flag = 0, elems[2] = {0, 0};
                for (int k = ne_ptr[n]; k < ne_ptr[n + 1]; k++) {
                    int e = ne_pos[k];
                    for (int q1 = 1; q1 <= 3; q1++)
                        if (A2(en, 3, q1, e) == node) { if (flag < 2) elems[flag] = e; flag++; break; }
                }
                if (flag == pass) {
                    ed.push_back(n); ed.push_back(node);
                    et.push_back(elems[0]); et.push_back(pass == 2 ? elems[1] : -999);
                }
            }
        if (pass == 2) edge2D_in = int(ed.size() / 2);
    }
    const int edge2D = int(ed.size() / 2);
    for (int n = 1; n <= edge2D; n++) {                                // orientation, fvom_init.F90:559-587
        int* e2 = &ed[2 * (n - 1)]; int* t2 = &et[2 * (n - 1)];
        if (t2[0] <= 0) std::swap(t2[0], t2[1]);
        double xc[2], xe[2];
        elem_center(h, t2[0], xc[0], xc[1]);
        xc[0] -= A2(coord, 2, 1, e2[0]); xc[1] -= A2(coord, 2, 2, e2[0]);
        xe[0] = A2(coord, 2, 1, e2[1]) - A2(coord, 2, 1, e2[0]);
        xe[1] = A2(coord, 2, 2, e2[1]) - A2(coord, 2, 2, e2[0]);
        trim_cyclic(xe[0], cyc); trim_cyclic(xc[0], cyc);
        if (xc[0] * xe[1] - xc[1] * xe[0] > 0.0) {
            if (t2[1] > 0) std::swap(t2[0], t2[1]);
            else std::swap(e2[0], e2[1]);
        }
    }
    h.p["edge2D"] = edge2D; h.p["edge2D_in"] = edge2D_in;
    h.p["myDim_edge2D"] = edge2D; h.p["eDim_edge2D"] = 0;
    int* edges = h.I("edges", {2, edge2D}); memcpy(edges, ed.data(), sizeof(int) * ed.size());
    int* edge_tri = h.I("edge_tri", {2, edge2D});
    for (size_t q = 0; q < et.size(); q++) edge_tri[q] = et[q] < 0 ? 0 : et[q];
    int* mle = h.I("myList_edge2D", {edge2D}); for (int n = 0; n < edge2D; n++) mle[n] = n + 1;
}

// find_levels of the partitioner (fvom_init.F90:657-871), reduced to its core rule: an element's bottom
// level is the first zbar interface at or below the mean of its node depths (using the mid-depth Z as
// threshold), elements need >= 3 levels, and a node's depth index is the max over its elements.
// Used only for synthetic meshes; real meshes ship nlvls.out / elvls.out.
void mesh_find_levels(Host& h) {
    const int nod2D = h.pi("nod2D"), elem2D = h.pi("elem2D"), nl = h.pi("nl");
    const int* en = h.i("elem2D_nodes"); const double* depth = h.d("depth");
    const double* zbar = h.d("zbar"); const double* Z = h.d("Z");
    int* nle = h.I("nlevels", {elem2D}); int* nln = h.I("nlevels_nod2D", {nod2D});
    int* ule = h.I("ulevels", {elem2D}); int* uln = h.I("ulevels_nod2D", {nod2D});
    const int thers = 5;   // thers_zbar_lev, gen_modules_config.F90
    (void)zbar;
    for (int e = 1; e <= elem2D; e++) {
        double dmean = (depth[A2(en, 3, 1, e) - 1] + depth[A2(en, 3, 2, e) - 1] + depth[A2(en, 3, 3, e) - 1]) / 3.0;   // which_depth_n2e='mean'
        int lev = 0, found = 0;
        for (int nz = 1; nz <= nl - 1; nz++) if (Z[nz - 1] < dmean) { lev = nz; found = 1; break; }
        if (!found && dmean < 0) lev = nl;
        if (dmean >= 0) lev = thers;
        if (lev < thers) lev = thers;
        nle[e - 1] = lev; ule[e - 1] = 1;
    }
    // every element keeps >= 2 open neighbours at each level (fvom_init.F90:772-818); needs edges
    {
        const int edge2D = h.pi("edge2D");
        const int* edge_tri = h.i("edge_tri");
        std::vector<int> nb(size_t(elem2D) * 3, 0), cnt(elem2D, 0);
        for (int n = 1; n <= edge2D; n++) {
            int t1 = A2(edge_tri, 2, 1, n), t2 = A2(edge_tri, 2, 2, n);
            if (t1 > 0 && t2 > 0) { nb[size_t(t1 - 1) * 3 + cnt[t1 - 1]++] = t2; nb[size_t(t2 - 1) * 3 + cnt[t2 - 1]++] = t1; }
        }
        for (int nz = thers + 1; nz <= nl; nz++) {
            int exit_flag = 0, it = 0;
            while (!exit_flag && it < 1000) {
                exit_flag = 1; it++;
                for (int n = 1; n <= elem2D; n++) {
                    if (nle[n - 1] < nz) continue;
                    int open = 0;
                    for (int i = 0; i < 3; i++) { int e = nb[size_t(n - 1) * 3 + i]; if (e > 1 && nle[e - 1] >= nz) open++; }
                    if (open < 2) {
                        if (nz - 1 < thers) { for (int i = 0; i < 3; i++) { int e = nb[size_t(n - 1) * 3 + i]; if (e > 0) nle[e - 1] = std::max(nle[e - 1], nz); } }
                        else nle[n - 1] = nz - 1;
                        exit_flag = 0;
                    }
                }
            }
        }
    }
    for (int n = 0; n < nod2D; n++) { nln[n] = 0; uln[n] = 1; }
    for (int e = 1; e <= elem2D; e++) for (int q = 1; q <= 3; q++) {
        int n = A2(en, 3, q, e); nln[n - 1] = std::max(nln[n - 1], nle[e - 1]);
    }
}

void mesh_derive(Host& h) {
    const int myN = h.pi("myDim_nod2D"), N = myN + h.pi("eDim_nod2D");
    const int myE = h.pi("myDim_elem2D"), Eall = myE + h.pi("eDim_elem2D") + h.pi("eXDim_elem2D");
    const int Ed = h.pi("myDim_edge2D") + h.pi("eDim_edge2D");
    const int nl = h.pi("nl");
    const double cyc = h.pd("cyclic_length");
    const bool cartesian = h.p.count("cartesian") && h.pi("cartesian");
    test_tri(h);
    int* en = h.i("elem2D_nodes"); const double* coord = h.d("coord_nod2D");
    const int* edges = h.i("edges"); const int* edge_tri = h.i("edge_tri");
    const int* nle = h.i("nlevels"); const int* ule = h.i("ulevels");
    const int* nln = h.i("nlevels_nod2D"); const int* uln = h.i("ulevels_nod2D");

    // load_edges (b): elem_edges, oce_mesh.F90:1606-1635
    int* elem_edges = h.I("elem_edges", {3, myE});
    {
        std::vector<int> aux(myE + 1, 0);
        for (int n = 1; n <= Ed; n++) for (int k = 1; k <= 2; k++) {
            int q = A2(edge_tri, 2, k, n);
            if (q > 0 && q <= myE) { aux[q]++; A2(elem_edges, 3, aux[q], q) = n; }
        }
        for (int e = 1; e <= myE; e++) {
            int eled[3] = {A2(elem_edges, 3, 1, e), A2(elem_edges, 3, 2, e), A2(elem_edges, 3, 3, e)};
            for (int q = 1; q <= 3; q++) for (int k = 0; k < 3; k++)
                if (A2(edges, 2, 1, eled[k]) != A2(en, 3, q, e) && A2(edges, 2, 2, eled[k]) != A2(en, 3, q, e)) {
                    A2(elem_edges, 3, q, e) = eled[k]; break;
                }
        }
    }
    // find_neighbors, oce_mesh.F90:1685-1752
    int* elem_nb = h.I("elem_neighbors", {3, myE});
    for (int e = 1; e <= myE; e++) for (int j = 1; j <= 3; j++) {
        int ed = A2(elem_edges, 3, j, e);
        int e1 = A2(edge_tri, 2, 1, ed);
        if (e1 == e) e1 = A2(edge_tri, 2, 2, ed);
        A2(elem_nb, 3, j, e) = e1;
    }
    int* nien = h.I("nod_in_elem2D_num", {N});
    // all local elements contribute (own nodes see own elements; halo nodes see what the owner would send,
    // which in the restricted-global construction is every local element containing them)
    for (int e = 1; e <= Eall; e++) for (int j = 1; j <= 3; j++) { int n = A2(en, 3, j, e); if (n >= 1 && n <= N) nien[n - 1]++; }
    int maxn = 0; for (int n = 0; n < N; n++) maxn = std::max(maxn, nien[n]);
    h.p["nod_in_elem2D_max"] = maxn;
    int* nie = h.I("nod_in_elem2D", {maxn, N});
    for (int n = 0; n < N; n++) nien[n] = 0;
    for (int e = 1; e <= Eall; e++) for (int j = 1; j <= 3; j++) {
        int n = A2(en, 3, j, e); if (n < 1 || n > N) continue;
        nien[n - 1]++; A2(nie, maxn, nien[n - 1], n) = e;
    }
    // find_levels_min_e2n, oce_mesh.F90:1350-1359
    int* nlmin = h.I("nlevels_nod2D_min", {N}); int* ulmax = h.I("ulevels_nod2D_max", {N});
    for (int n = 1; n <= N; n++) {
        int mn = 1 << 30, mx = 0;
        for (int k = 1; k <= nien[n - 1]; k++) { int e = A2(nie, maxn, k, n); mn = std::min(mn, nle[e - 1]); mx = std::max(mx, ule[e - 1]); }
        nlmin[n - 1] = mn; ulmax[n - 1] = mx;
    }
    // mesh_areas, oce_mesh.F90:1882-1999
    double* elem_area = h.D("elem_area", {Eall});
    for (int n = 1; n <= Eall; n++) {
        int e1 = A2(en, 3, 1, n), e2 = A2(en, 3, 2, n), e3 = A2(en, 3, 3, n);
        double ay = (A2(coord, 2, 2, e1) + A2(coord, 2, 2, e2) + A2(coord, 2, 2, e3)) / 3.0;
        ay = cos(ay); if (cartesian) ay = 1.0;
        double a[2] = {A2(coord, 2, 1, e2) - A2(coord, 2, 1, e1), A2(coord, 2, 2, e2) - A2(coord, 2, 2, e1)};
        double b[2] = {A2(coord, 2, 1, e3) - A2(coord, 2, 1, e1), A2(coord, 2, 2, e3) - A2(coord, 2, 2, e1)};
        trim_cyclic(a[0], cyc); trim_cyclic(b[0], cyc);
        a[0] *= ay; b[0] *= ay;
        elem_area[n - 1] = 0.5 * fabs(a[0] * b[1] - b[0] * a[1]);
    }
    bool use_cavity = false;
    for (int e = 0; e < Eall; e++) if (ule[e] > 1) use_cavity = true;
    h.p["use_cavity"] = use_cavity;
    double* area = h.D("area", {nl, N}); double* areasvol = h.D("areasvol", {nl, N});
    std::vector<int> cav; if (use_cavity) cav.assign(size_t(nl) * N, 0);
    for (int n = 1; n <= N; n++) for (int j = 1; j <= nien[n - 1]; j++) {
        int e = A2(nie, maxn, j, n);
        int nzmin = ule[e - 1], nzmax = nle[e - 1] - 1;
        for (int nz = nzmin; nz <= nzmax; nz++) A2(area, nl, nz, n) = A2(area, nl, nz, n) + elem_area[e - 1] / 3.0;
        if (use_cavity) for (int nz = 1; nz <= nzmin - 1; nz++) A2(cav.data(), nl, nz, n)++;
    }
    for (int n = 1; n <= N; n++) {
        int nzmin = uln[n - 1], nzmax = nln[n - 1] - 1;
        for (int nz = nzmin; nz <= nzmax; nz++) {
            if (use_cavity && A2(cav.data(), nl, nz, n) > 0) A2(areasvol, nl, nz, n) = A2(area, nl, std::min(nz + 1, nzmax), n);
            else A2(areasvol, nl, nz, n) = A2(area, nl, nz, n);
        }
    }
    for (int e = 0; e < Eall; e++) elem_area[e] = elem_area[e] * R_EARTH * R_EARTH;
    for (int64_t q = 0; q < int64_t(nl) * N; q++) { area[q] = area[q] * R_EARTH * R_EARTH; areasvol[q] = areasvol[q] * R_EARTH * R_EARTH; }
    // mesh_resolution, oce_mesh.F90:2034-2057 (3 mass-matrix smoothing passes)
    double* res = h.D("mesh_resolution", {N});
    for (int n = 1; n <= N; n++) res[n - 1] = sqrt(A2(areasvol, nl, uln[n - 1], n) / PI) * 2.0;
    {
        std::vector<double> work(N);
        for (int q = 0; q < 3; q++) {
            for (int n = 1; n <= N; n++) {
                double vol = 0, w = 0;
                for (int j = 1; j <= nien[n - 1]; j++) {
                    int e = A2(nie, maxn, j, n);
                    w = w + (res[A2(en, 3, 1, e) - 1] + res[A2(en, 3, 2, e) - 1] + res[A2(en, 3, 3, e) - 1]) / 3.0 * elem_area[e - 1];
                    vol = vol + elem_area[e - 1];
                }
                work[n - 1] = w / vol;
            }
            for (int n = 0; n < N; n++) res[n] = work[n];
        }
    }
    // mesh_auxiliary_arrays, oce_mesh.F90:2178-2306
    double* elem_cos = h.D("elem_cos", {Eall}); double* metric = h.D("metric_factor", {Eall});
    std::vector<double> cx(Eall), cy(Eall);
    for (int n = 1; n <= Eall; n++) {
        double ax, ay; elem_center(h, n, ax, ay);
        cx[n - 1] = ax; cy[n - 1] = ay;
        elem_cos[n - 1] = cos(ay); metric[n - 1] = tan(ay) / R_EARTH;
        if (cartesian) { elem_cos[n - 1] = 1.0; metric[n - 1] = 0.0; }
    }
    double* edge_dxdy = h.D("edge_dxdy", {2, Ed}); double* ecd = h.D("edge_cross_dxdy", {4, Ed});
    for (int n = 1; n <= Ed; n++) {
        int e1 = A2(edges, 2, 1, n), e2 = A2(edges, 2, 2, n);
        double a[2] = {A2(coord, 2, 1, e2) - A2(coord, 2, 1, e1), A2(coord, 2, 2, e2) - A2(coord, 2, 2, e1)};
        trim_cyclic(a[0], cyc);
        A2(edge_dxdy, 2, 1, n) = a[0]; A2(edge_dxdy, 2, 2, n) = a[1];
        // edge_center, oce_mesh.F90:1799-1816
        double pa[2] = {A2(coord, 2, 1, e1), A2(coord, 2, 2, e1)}, pb[2] = {A2(coord, 2, 1, e2), A2(coord, 2, 2, e2)};
        if (pa[0] - pb[0] > cyc / 2.0) pa[0] -= cyc;
        if (pa[0] - pb[0] < -cyc / 2.0) pb[0] -= cyc;
        double ec[2] = {0.5 * (pa[0] + pb[0]), 0.5 * (pa[1] + pb[1])};
        for (int s = 0; s < 2; s++) {
            int el = A2(edge_tri, 2, s + 1, n);
            if (el > 0) {
                double b[2] = {cx[el - 1] - ec[0], cy[el - 1] - ec[1]};
                trim_cyclic(b[0], cyc);
                b[0] = b[0] * elem_cos[el - 1];
                b[0] = b[0] * R_EARTH; b[1] = b[1] * R_EARTH;
                A2(ecd, 4, 2 * s + 1, n) = b[0]; A2(ecd, 4, 2 * s + 2, n) = b[1];
            }
        }
    }
    double* gs = h.D("gradient_sca", {6, myE});
    for (int e = 1; e <= myE; e++) {
        int n1 = A2(en, 3, 1, e), n2 = A2(en, 3, 2, e), n3 = A2(en, 3, 3, e);
        double dX31 = A2(coord, 2, 1, n3) - A2(coord, 2, 1, n1); trim_cyclic(dX31, cyc); dX31 = elem_cos[e - 1] * dX31;
        double dX21 = A2(coord, 2, 1, n2) - A2(coord, 2, 1, n1); trim_cyclic(dX21, cyc); dX21 = elem_cos[e - 1] * dX21;
        double dY31 = A2(coord, 2, 2, n3) - A2(coord, 2, 2, n1);
        double dY21 = A2(coord, 2, 2, n2) - A2(coord, 2, 2, n1);
        double df = -0.5 * R_EARTH / elem_area[e - 1];
        A2(gs, 6, 1, e) = (-dY31 + dY21) * df; A2(gs, 6, 2, e) = dY31 * df; A2(gs, 6, 3, e) = -dY21 * df;
        A2(gs, 6, 4, e) = (dX31 - dX21) * df; A2(gs, 6, 5, e) = -dX31 * df; A2(gs, 6, 6, e) = dX21 * df;
    }
    double* geo = h.D("geo_coord_nod2D", {2, N});
    {
        double m[3][3]; rot_matrix(h.pd("alphaEuler"), h.pd("betaEuler"), h.pd("gammaEuler"), m);
        for (int n = 1; n <= N; n++) {
            double lon, lat;
            if (h.pi("force_rotation")) r2g(m, lon, lat, A2(coord, 2, 1, n), A2(coord, 2, 2, n));
            else { lon = A2(coord, 2, 1, n); lat = A2(coord, 2, 2, n); }
            A2(geo, 2, 1, n) = lon; A2(geo, 2, 2, n) = lat;
        }
    }
}

// Build the arrays of one METIS partition from the global mesh + the reference's dist_N lists
// (my_listNNNNN.out: own nodes first, then halo; own elements, eDim, eXDim; own edges, eDim).
// Global quantities are restricted through the lists, which is what the reference obtains by reading
// in chunks + halo exchange (oce_mesh.F90:240-254, 419-475, 1468-1599).
void mesh_restrict(Host& g, Host& loc, int npes, int mype, const int* mylist_nod, int myDim_nod, int eDim_nod,
                   const int* mylist_elem, int myDim_elem, int eDim_elem, int eXDim_elem,
                   const int* mylist_edge, int myDim_edge, int eDim_edge) {
    const int N = myDim_nod + eDim_nod, Eall = myDim_elem + eDim_elem + eXDim_elem, Ed = myDim_edge + eDim_edge;
    const int gN = g.pi("nod2D"), gE = g.pi("elem2D"), nl = g.pi("nl");
    loc.p = g.p;
    loc.p["npes"] = npes; loc.p["mype"] = mype;
    loc.p["myDim_nod2D"] = myDim_nod; loc.p["eDim_nod2D"] = eDim_nod;
    loc.p["myDim_elem2D"] = myDim_elem; loc.p["eDim_elem2D"] = eDim_elem; loc.p["eXDim_elem2D"] = eXDim_elem;
    loc.p["myDim_edge2D"] = myDim_edge; loc.p["eDim_edge2D"] = eDim_edge;
    std::vector<int> g2l_n(gN + 1, 0), g2l_e(gE + 1, 0);
    for (int n = 0; n < N; n++) g2l_n[mylist_nod[n]] = n + 1;
    for (int e = 0; e < Eall; e++) g2l_e[mylist_elem[e]] = e + 1;
    int* ml = loc.I("myList_nod2D", {N}); memcpy(ml, mylist_nod, sizeof(int) * size_t(N));
    int* mle = loc.I("myList_elem2D", {Eall}); memcpy(mle, mylist_elem, sizeof(int) * size_t(Eall));
    int* mled = loc.I("myList_edge2D", {Ed}); memcpy(mled, mylist_edge, sizeof(int) * size_t(Ed));
    const double* gc = g.d("coord_nod2D"); double* c = loc.D("coord_nod2D", {2, N});
    const double* gd = g.d("depth"); double* d = loc.D("depth", {N});
    const int* gnln = g.i("nlevels_nod2D"); int* nln = loc.I("nlevels_nod2D", {N});
    const int* guln = g.i("ulevels_nod2D"); int* uln = loc.I("ulevels_nod2D", {N});
    for (int n = 1; n <= N; n++) {
        int gn = mylist_nod[n - 1];
        A2(c, 2, 1, n) = A2(gc, 2, 1, gn); A2(c, 2, 2, n) = A2(gc, 2, 2, gn);
        d[n - 1] = gd[gn - 1]; nln[n - 1] = gnln[gn - 1]; uln[n - 1] = guln[gn - 1];
    }
    const int* gen = g.i("elem2D_nodes"); int* en = loc.I("elem2D_nodes", {3, Eall});
    const int* gnle = g.i("nlevels"); int* nle = loc.I("nlevels", {Eall});
    const int* gule = g.i("ulevels"); int* ule = loc.I("ulevels", {Eall});
    for (int e = 1; e <= Eall; e++) {
        int ge = mylist_elem[e - 1];
        for (int k = 1; k <= 3; k++) A2(en, 3, k, e) = g2l_n[A2(gen, 3, k, ge)];   // 0 if the node is not local (eXDim corner)
        nle[e - 1] = gnle[ge - 1]; ule[e - 1] = gule[ge - 1];
    }
    const int* ged = g.i("edges"); const int* get = g.i("edge_tri");
    int* ed = loc.I("edges", {2, Ed}); int* et = loc.I("edge_tri", {2, Ed});
    for (int n = 1; n <= Ed; n++) {
        int gn = mylist_edge[n - 1];
        for (int k = 1; k <= 2; k++) {
            A2(ed, 2, k, n) = g2l_n[A2(ged, 2, k, gn)];
            int t = A2(get, 2, k, gn);
            A2(et, 2, k, n) = t > 0 ? g2l_e[t] : 0;
        }
    }
    for (const char* nm : {"zbar", "Z"}) { Arr& s = g.a[nm]; loc.a[nm] = s; }
    int* part = loc.I("part", {npes + 1});
    (void)part;
}

}  // namespace fh
