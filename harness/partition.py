"""Domain decomposition of a HostModel: the reference's offline partitioner output (dist_N/rpart.out,
my_listNNNNN.out, com_infoNNNNN.out) and the per-rank model state built from it.

The node partition itself comes from METIS in the reference (src/fort_part.c, kept as-is); here it is either read
from the reference's own dist_N files or a contiguous split of the node numbering.  Everything downstream follows
src/gen_comm.F90:
  communication_nodn  :12-215   own nodes, halo nodes = other nodes of my elements, send/receive lists
  communication_elemn :218-515  my elements = at least one own node; small halo = edge neighbours, full halo = all
                                elements around my nodes; an element's "main owner" is the PE of its first node
  mymesh              :517-641  my edges = at least one own node
Local numbering keeps ascending global order inside every group, so a node-centric gather on a partition adds its
terms in the same order as on the undivided mesh.
"""
import os

import numpy as np

from .host import HostModel

NODE_ARRAYS = ["CFL_z", "Ki", "Kv", "Wvel", "Wvel_e", "Wvel_i", "Z_3d_n", "area", "areasvol", "bottom_node_thickness",
               "coord_nod2D", "d_eta", "del_ttf", "depth", "eta_n", "geo_coord_nod2D", "hbar", "hbar_old", "heat_flux",
               "hnode", "hnode_new", "mesh_resolution", "nboundary_lay", "nlevels_nod2D", "nlevels_nod2D_min",
               "real_salt_flux", "relax_salt", "slope_tapered", "ssh_rhs", "ssh_rhs_old", "sw_3d", "tr_z",
               "ulevels_nod2D", "ulevels_nod2D_max", "virtual_salt", "water_flux", "zbar_3d_n", "zbar_n_bot", "zbar_n_srf"]
ELEM_ARRAYS = ["UV", "UV_rhs", "bottom_elem_thickness", "dhe", "elem_area", "elem_cos", "gradient_sca", "helem",
               "metric_factor", "nlevels", "tr_xy", "ulevels", "zbar_e_bot", "zbar_e_srf"]
EDGE_ARRAYS = ["edge_cross_dxdy", "edge_dxdy"]
GLOBAL_ARRAYS = ["Z", "zbar"]


def contiguous_partition(hm, npes):
    """part[n] = PE (0-based) of global node n: npes contiguous chunks of the node numbering with equal node counts"""
    n = hm.ipar("myDim_nod2D")
    return np.minimum((np.arange(n, dtype=np.int64) * npes) // n, npes - 1).astype(np.int32)


def read_rpart(path, npes):
    """dist_N/rpart.out: npes, nodes per PE, then for every global node its index in the PE-contiguous numbering
    (oce_mesh.F90:199-213, oce_ale.F90:1317-1344).  Returns part (0-based PE per node)."""
    t = open(os.path.join(path, "rpart.out")).read().split()
    assert int(t[0]) == npes
    cnt = np.array(t[1:1 + npes], dtype=np.int64)
    pc = np.array(t[1 + npes:], dtype=np.int64)              # 1-based PE-contiguous index of every node
    off = np.concatenate([[0], np.cumsum(cnt)])
    return (np.searchsorted(off, pc - 1, side="right") - 1).astype(np.int32)


def read_my_list(path, mype):
    """dist_N/my_listNNNNN.out (oce_mesh.F90:234-256)"""
    t = iter(open(os.path.join(path, "my_list%05d.out" % mype)).read().split())
    nx = lambda: int(next(t))                                  # noqa: E731
    assert nx() == mype
    d = {}
    d["myDim_nod2D"], d["eDim_nod2D"] = nx(), nx()
    d["myList_nod2D"] = np.array([nx() for _ in range(d["myDim_nod2D"] + d["eDim_nod2D"])], dtype=np.int32)
    d["myDim_elem2D"], d["eDim_elem2D"], d["eXDim_elem2D"] = nx(), nx(), nx()
    d["myList_elem2D"] = np.array([nx() for _ in range(d["myDim_elem2D"] + d["eDim_elem2D"] + d["eXDim_elem2D"])], dtype=np.int32)
    d["myDim_edge2D"], d["eDim_edge2D"] = nx(), nx()
    d["myList_edge2D"] = np.array([nx() for _ in range(d["myDim_edge2D"] + d["eDim_edge2D"])], dtype=np.int32)
    return d


class RankMesh:
    """lists of one partition, all 1-based global ids like the reference's myList_* arrays"""

    def __init__(self, npes, mype):
        self.npes, self.mype = npes, mype
        self.nod = self.elem = self.edge = None
        self.myDim_nod2D = self.eDim_nod2D = 0
        self.myDim_elem2D = self.eDim_elem2D = self.eXDim_elem2D = 0
        self.myDim_edge2D = self.eDim_edge2D = 0
        self.com_nod = None          # dict(rPE, rptr, rlist, sPE, sptr, slist): com_struct, local 1-based indices
        self.com_elem_full = None


def _com(recv_owner, recv_local, send_map, npes, mype):
    """com_struct (gen_modules_partitioning.F90:17-29) from: owner PE and local index of every halo entity (already
    grouped by owner, ascending global id inside a group) and send_map[pe] = local indices to send to pe"""
    rPE = [int(p) for p in np.unique(recv_owner)]
    rptr = [1]
    for p in rPE:
        rptr.append(rptr[-1] + int((recv_owner == p).sum()))
    sPE = sorted(p for p in send_map if len(send_map[p]))
    sptr, slist = [1], []
    for p in sPE:
        slist.append(np.asarray(send_map[p], dtype=np.int32))
        sptr.append(sptr[-1] + len(send_map[p]))
    return dict(rPE=np.array(rPE, dtype=np.int32), rptr=np.array(rptr, dtype=np.int32),
                rlist=np.asarray(recv_local, dtype=np.int32),
                sPE=np.array(sPE, dtype=np.int32), sptr=np.array(sptr, dtype=np.int32),
                slist=(np.concatenate(slist) if slist else np.zeros(0, dtype=np.int32)))


def build_partition(hm, part, npes):
    """RankMesh for every PE from the global mesh of `hm` and the node partition `part` (0-based PE per node)."""
    part = np.asarray(part, dtype=np.int32)
    en = hm["elem2D_nodes"].astype(np.int64) - 1             # (E, 3)
    ed = hm["edges"].astype(np.int64) - 1                    # (Ed, 2)
    enb = hm["elem_neighbors"].astype(np.int64) - 1          # (E, 3), -1 = boundary
    N, E = part.size, en.shape[0]
    epart = part[en]                                          # PE of the three corners
    main_owner = epart[:, 0]
    ranks = []
    for me in range(npes):
        r = RankMesh(npes, me)
        own = np.flatnonzero(part == me)
        my_el = np.flatnonzero((epart == me).any(axis=1))
        # halo nodes: other corners of my elements, grouped by owner PE, ascending id inside
        cand = np.unique(en[my_el].ravel())
        halo = cand[part[cand] != me]
        halo = halo[np.lexsort((halo, part[halo]))]
        loc_nodes = np.concatenate([own, halo])
        is_loc = np.zeros(N, dtype=bool)
        is_loc[loc_nodes] = True
        is_my_el = np.zeros(E, dtype=bool)
        is_my_el[my_el] = True
        # small halo: edge neighbours of my elements without an own node; full halo: every other element around a local node
        nb = enb[my_el].ravel()
        nb = np.unique(nb[nb >= 0])
        small = nb[~is_my_el[nb]]
        around = np.flatnonzero(is_loc[en].any(axis=1))
        rest = around[~is_my_el[around]]
        is_small = np.zeros(E, dtype=bool)
        is_small[small] = True
        extra = rest[~is_small[rest]]
        small = small[np.lexsort((small, main_owner[small]))]
        extra = extra[np.lexsort((extra, main_owner[extra]))]
        my_ed = np.flatnonzero((part[ed] == me).any(axis=1))
        r.nod = (loc_nodes + 1).astype(np.int32)
        r.elem = (np.concatenate([my_el, small, extra]) + 1).astype(np.int32)
        r.edge = (my_ed + 1).astype(np.int32)
        r.myDim_nod2D, r.eDim_nod2D = own.size, halo.size
        r.myDim_elem2D, r.eDim_elem2D, r.eXDim_elem2D = my_el.size, small.size, extra.size
        r.myDim_edge2D, r.eDim_edge2D = my_ed.size, 0
        ranks.append(r)
    # communication lists
    for me, r in enumerate(ranks):
        g2l_n = np.zeros(N, dtype=np.int64)
        g2l_n[r.nod - 1] = np.arange(1, r.nod.size + 1)
        g2l_e = np.zeros(E, dtype=np.int64)
        g2l_e[r.elem - 1] = np.arange(1, r.elem.size + 1)
        halo = r.nod[r.myDim_nod2D:] - 1
        send_n, send_e = {}, {}
        for other, ro in enumerate(ranks):
            if other == me:
                continue
            oh = ro.nod[ro.myDim_nod2D:] - 1
            mine = oh[part[oh] == me]                         # in the receiver's order: ascending global id
            if mine.size:
                send_n[other] = g2l_n[mine]
            oe = ro.elem[ro.myDim_elem2D:] - 1
            # the receiver keeps two owner-sorted groups (small, then extra halo); its receive order is by (group, owner, id)
            mine_e = oe[main_owner[oe] == me]
            if mine_e.size:
                send_e[other] = g2l_e[mine_e]
        r.com_nod = _com(part[halo], np.arange(r.myDim_nod2D + 1, r.nod.size + 1), send_n, npes, me)
        he = r.elem[r.myDim_elem2D:] - 1
        # receive list of the full element halo grouped by main owner, keeping the (small, extra) storage order
        order = np.argsort(main_owner[he], kind="stable")
        r.com_elem_full = _com(main_owner[he][order], (r.myDim_elem2D + 1 + order), send_e, npes, me)
        # the sender lists elements in the receiver's storage order filtered by owner == sender; same filter, same order
    return ranks


def pe_contiguous_numbering(part, npes):
    """(offsets[npes+1] 1-based like the reference's `part` array in oce_mesh.F90:206-211, index[n] 1-based position
    of global node n when nodes are sorted by (PE, global id)) — the numbering of ssh_stiff%colind (oce_ale.F90:1317-1344)"""
    part = np.asarray(part)
    order = np.lexsort((np.arange(part.size), part))
    idx = np.empty(part.size, dtype=np.int64)
    idx[order] = np.arange(1, part.size + 1)
    cnt = np.bincount(part, minlength=npes)
    off = np.concatenate([[1], 1 + np.cumsum(cnt)]).astype(np.int32)
    return off, idx


PARAMS = ("nl", "dt", "alpha", "theta", "w_split", "w_max_cfl", "K_hor", "is_nonlinfs", "which_ale", "num_tracers",
          "cyclic_length", "nod_in_elem2D_max", "use_partial_cell", "cartesian", "force_rotation")


def local_arrays(hm, r, part):
    """(params, arrays) of partition `r` (a RankMesh): every array of the global model `hm` restricted to the local
    lists, connectivity renumbered to local 1-based ids (0 = not local), SSH matrix rows of the own nodes with columns
    in the PE-contiguous global numbering.  This is the state the reference's rank `r.mype` holds after mesh_setup +
    ocean_setup (src/oce_mesh.F90:108-143, src/oce_setup_step.F90:20).  Arrays are numpy, reversed-Fortran shapes."""
    names = set(hm.names())
    params = {}
    for k in PARAMS:
        try:
            params[k] = hm.par(k)
        except RuntimeError:
            pass
    nod, elem, edge = r.nod.astype(np.int64) - 1, r.elem.astype(np.int64) - 1, r.edge.astype(np.int64) - 1
    params.update(npes=r.npes, mype=r.mype, myDim_nod2D=r.myDim_nod2D, eDim_nod2D=r.eDim_nod2D, myDim_elem2D=r.myDim_elem2D,
                  eDim_elem2D=r.eDim_elem2D, eXDim_elem2D=r.eXDim_elem2D, myDim_edge2D=r.myDim_edge2D, eDim_edge2D=r.eDim_edge2D,
                  nod2D=hm.ipar("myDim_nod2D"), elem2D=hm.ipar("myDim_elem2D"), edge2D=hm.ipar("myDim_edge2D"))
    Ng, Eg = hm.ipar("myDim_nod2D"), hm.ipar("myDim_elem2D")
    g2l_n = np.zeros(Ng + 1, dtype=np.int32)
    g2l_n[nod + 1] = np.arange(1, nod.size + 1)
    g2l_e = np.zeros(Eg + 1, dtype=np.int32)
    g2l_e[elem + 1] = np.arange(1, elem.size + 1)
    A = {}
    for nm in NODE_ARRAYS:
        if nm in names:
            A[nm] = hm[nm][nod]
    for nm in ELEM_ARRAYS:
        if nm in names:
            A[nm] = hm[nm][elem]
    for nm in EDGE_ARRAYS:
        if nm in names:
            A[nm] = hm[nm][edge]
    for nm in GLOBAL_ARRAYS:
        if nm in names:
            A[nm] = hm[nm].copy()
    for nm in ("tr_arr", "tr_arr_old"):
        A[nm] = hm[nm][:, nod]
    A["myList_nod2D"], A["myList_elem2D"], A["myList_edge2D"] = r.nod, r.elem, r.edge
    A["edges"] = g2l_n[hm["edges"][edge]]
    A["edge_tri"] = g2l_e[np.maximum(hm["edge_tri"][edge], 0)]
    A["elem2D_nodes"] = g2l_n[hm["elem2D_nodes"][elem]]
    A["edge_up_dn_tri"] = g2l_e[np.maximum(hm["edge_up_dn_tri"][edge], 0)]
    A["nod_in_elem2D_num"] = hm["nod_in_elem2D_num"][nod]
    A["nod_in_elem2D"] = g2l_e[np.maximum(hm["nod_in_elem2D"][nod], 0)]
    # SSH system: own rows, PE-contiguous global columns (init_stiff_mat_ale, oce_ale.F90:1128-1344)
    off, pc = pe_contiguous_numbering(part, r.npes)
    own = nod[:r.myDim_nod2D]
    rp = hm["ssh_rowptr"].astype(np.int64)
    cnt = rp[own + 1] - rp[own]
    lrp = np.concatenate([[0], np.cumsum(cnt)])
    gather = np.repeat(rp[own] - 1 - lrp[:-1], cnt) + np.arange(lrp[-1])
    A["ssh_rowptr"] = (lrp + 1).astype(np.int32)             # local 1-based row pointers (rowptr_loc)
    A["ssh_colind"] = pc[hm["ssh_colind"][gather] - 1].astype(np.int32)
    A["ssh_values"] = hm["ssh_values"][gather]
    A["ssh_colind_loc"] = g2l_n[hm["ssh_colind_loc"][gather]]    # local 1-based column numbers (colind_loc, MOD_MESH.F90:15)
    A["part"] = off
    for pre, c in (("com_nod2D", r.com_nod), ("com_elem2D_full", r.com_elem_full)):
        for k in ("rPE", "rptr", "rlist", "sPE", "sptr", "slist"):
            A["%s_%s" % (pre, k)] = np.asarray(c[k], dtype=np.int32) if len(c[k]) else np.zeros(1, dtype=np.int32)
        params[pre + "_rPEnum"] = len(c["rPE"])
        params[pre + "_sPEnum"] = len(c["sPE"])
    return params, {k: np.ascontiguousarray(v) for k, v in A.items()}


def model_from_arrays(params, arrays):
    lm = HostModel()
    lm.set(**params)
    for name, a in arrays.items():
        dst = lm.new(name, a.shape[::-1], np.int32 if a.dtype == np.int32 else np.float64)
        dst[...] = a
    return lm


def local_model(hm, r, part):
    return model_from_arrays(*local_arrays(hm, r, part))


def scatter_local_models(hm, part, dist, group=None):
    """collective over `dist` (a CPU-capable torch.distributed group, e.g. gloo): rank 0 holds the global model `hm`,
    partitions it and ships every rank its own partition (the reference reaches the same state by every rank reading
    the mesh files + its dist_N lists, oce_mesh.F90:147-692).  Returns (local HostModel, own wet node-layers)."""
    import torch
    rank, world = dist.get_rank(), dist.get_world_size()
    if rank == 0:
        ranks = build_partition(hm, part, world)
        mine = None
        for dst in range(world):
            params, arrays = local_arrays(hm, ranks[dst], part)
            if dst == 0:
                mine = (params, arrays)
                continue
            meta = [(k, a.shape, str(a.dtype)) for k, a in arrays.items()]
            dist.send_object_list([params, meta], dst=dst, group=group)
            for k, a in arrays.items():
                dist.send(torch.from_numpy(a.reshape(-1)), dst=dst, group=group)
            del arrays
        params, arrays = mine
    else:
        box = [None, None]
        dist.recv_object_list(box, src=0, group=group)
        params, meta = box
        arrays = {}
        for k, shape, dt in meta:
            t = torch.empty(int(np.prod(shape)), dtype=torch.int32 if dt == "int32" else torch.float64)
            dist.recv(t, src=0, group=group)
            arrays[k] = t.numpy().reshape(shape)
    return model_from_arrays(params, arrays)


def com_arrays(c):
    """the six arrays of a com_struct in the order fb200_init_halo takes them"""
    return c["rPE"], c["rptr"], c["rlist"], c["sPE"], c["sptr"], c["slist"]


def exchange_host(ranks, fields, kind="nod"):
    """reference halo exchange (exchange_nod / exchange_elem, gen_halo_exchange.F90) on host arrays of ALL ranks held
    in one process: fields[rank] is an array whose first numpy axis is the local entity.  Used by the CPU tests to
    drive the oracle partition by partition."""
    for me, r in enumerate(ranks):
        c = r.com_nod if kind == "nod" else r.com_elem_full
        for k, pe in enumerate(c["sPE"]):
            sl = c["slist"][c["sptr"][k] - 1:c["sptr"][k + 1] - 1] - 1
            co = ranks[pe].com_nod if kind == "nod" else ranks[pe].com_elem_full
            j = list(co["rPE"]).index(me)
            rl = co["rlist"][co["rptr"][j] - 1:co["rptr"][j + 1] - 1] - 1
            assert rl.size == sl.size
            fields[pe][rl] = fields[me][sl]


def exchange_dist(lm, field, kind, dist, group=None):
    """exchange_nod / exchange_elem (gen_halo_exchange.F90:129-164, 317-363, 662-786) of one host array over
    torch.distributed point-to-point (gloo on CPU): `field`'s first numpy axis is the local entity; com lists from `lm`."""
    import torch
    pre = "com_nod2D" if kind == "nod" else "com_elem2D_full"
    nr, ns = lm.ipar(pre + "_rPEnum"), lm.ipar(pre + "_sPEnum")
    rPE, rptr, rlist = lm[pre + "_rPE"], lm[pre + "_rptr"], lm[pre + "_rlist"]
    sPE, sptr, slist = lm[pre + "_sPE"], lm[pre + "_sptr"], lm[pre + "_slist"]
    reqs, rbufs = [], []
    for k in range(nr):
        idx = rlist[rptr[k] - 1:rptr[k + 1] - 1] - 1
        t = torch.empty((idx.size,) + field.shape[1:], dtype=torch.float64)
        rbufs.append((idx, t))
        reqs.append(dist.irecv(t, src=int(rPE[k]), group=group))
    for k in range(ns):
        idx = slist[sptr[k] - 1:sptr[k + 1] - 1] - 1
        reqs.append(dist.isend(torch.from_numpy(np.ascontiguousarray(field[idx])), dst=int(sPE[k]), group=group))
    for q in reqs:
        q.wait()
    for idx, t in rbufs:
        field[idx] = t.numpy()


# ---------------------------------------------------------------------------------------------------
# The reference's own partitioner: stiff_mat_ini (the node graph), partit (src/fort_part.c -> METIS 5.1.0, compiled from
# /root/reference by oracle/Makefile.metis into oracle/_ref/libmetis_part.so) and check_partitioning.
def node_graph(elem2D_nodes, nod2D):
    """stiff_mat_ini (src/fvom_init.F90:1351-1412): CSR adjacency of the nodes, 1-based, no diagonal; a node's neighbours in the
    order in which the element loop first meets them (pairs (2,1), (3,1), (3,2) of every triangle)."""
    en = np.asarray(elem2D_nodes, dtype=np.int64)                    # (E, 3), 1-based
    a = np.stack([en[:, 1], en[:, 2], en[:, 2]], axis=1).ravel()     # nod(i)
    b = np.stack([en[:, 0], en[:, 0], en[:, 1]], axis=1).ravel()     # nod(j)
    key = np.minimum(a, b) * (nod2D + 1) + np.maximum(a, b)
    _, first = np.unique(key, return_index=True)
    first.sort()                                                     # pairs in the order of their first occurrence
    a, b = a[first], b[first]
    src = np.concatenate([a, b])
    dst = np.concatenate([b, a])
    t = np.concatenate([np.arange(a.size), np.arange(a.size)])       # both directions enter the lists at the same time
    order = np.lexsort((t, src))
    src, dst = src[order], dst[order]
    rowptr = np.ones(nod2D + 1, dtype=np.int32)
    rowptr[1:] = 1 + np.cumsum(np.bincount(src - 1, minlength=nod2D))
    return rowptr, dst.astype(np.int32)


def check_partitioning(part, rowptr, colind, nlevels_nod2D, npes, verbose=False):
    """check_partitioning (src/fvom_init.F90:1543-1690): nodes with at most one neighbour in their own partition move to an
    adjacent partition.  part: 0-based, modified in place; returns the number of moved nodes."""
    n = part.size
    rp, ci = rowptr.astype(np.int64) - 1, colind.astype(np.int64) - 1
    nl = nlevels_nod2D.astype(np.int64)
    npp = np.zeros((2, npes), dtype=np.int64)
    npp[0] = np.bincount(part, minlength=npes)
    npp[1] = np.bincount(part, weights=nl, minlength=npes).astype(np.int64)
    max1, max2 = npp[0].max(), npp[1].max()
    # candidates: the test reads the CURRENT partition vector, so every node is re-tested in order; nodes whose own and
    # neighbours' partitions never changed keep their initial verdict
    same0 = np.add.reduceat((part[ci] == np.repeat(part, np.diff(rp))).astype(np.int64), rp[:-1])
    dirty = np.zeros(n, dtype=bool)
    moved = 0
    for k in range(n):
        if same0[k] > 1 and not dirty[k]:
            continue
        nb = part[ci[rp[k]:rp[k + 1]]]
        if (nb == part[k]).sum() > 1:
            continue
        ne_part, ne_num, ne_load = [int(nb[0])], [1], [(npp[0, nb[0]] + 1, npp[1, nb[0]] + nl[k])]
        for p_ in nb:
            if p_ == part[k]:
                continue
            if p_ in ne_part:
                ne_num[ne_part.index(int(p_))] += 1
            else:
                ne_part.append(int(p_)); ne_num.append(1); ne_load.append((npp[0, p_] + 1, npp[1, p_] + nl[k]))
        kmax = int(np.argmax(ne_num))
        if ne_num[kmax] <= 1:
            continue
        if ne_load[kmax][0] <= max1 and ne_load[kmax][1] <= max2:
            kk = kmax
        else:
            kk = None
            for q in range(len(ne_part)):
                if ne_num[q] == 1 or q == kmax:
                    continue
                if ne_load[q][0] <= max1 and ne_load[q][1] <= max2:
                    kk = q
                    break
            if kk is None:
                kk = kmax
        new = ne_part[kk]
        npp[0, new] += 1; npp[1, new] += nl[k]; npp[0, part[k]] -= 1; npp[1, part[k]] -= nl[k]
        if verbose:
            print("node", k + 1, "moves from part", part[k], "to", new)
        part[k] = new
        dirty[ci[rp[k]:rp[k + 1]]] = True
        moved += 1
    return moved


_metis = {}


def metis_lib(plain=False):
    """oracle/_ref/libmetis_part.so (built here from /root/reference by oracle/Makefile.metis; it travels with the snapshot);
    plain: the copy compiled without -DPART_WEIGHTED / -DMETISRANDOMSEED"""
    global _metis
    if plain not in _metis:
        import ctypes as C
        root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
        so = os.path.join(root, "oracle", "_ref", "libmetis_part_plain.so" if plain else "libmetis_part.so")
        mk = os.path.join(root, "oracle", "Makefile.metis")
        if not os.path.exists(so) and os.path.isdir("/root/reference/lib/metis-5.1.0") and os.path.exists(mk):
            import subprocess
            subprocess.run(["make", "-s", "-f", mk, "-j8"], cwd=os.path.join(root, "oracle"), check=True)
        if not os.path.exists(so):
            return None
        _metis[plain] = C.CDLL(so)
    return _metis[plain]


def metis_partition(hm, npes, check=True, plain=False):
    """the reference's partitioning of the mesh in hm (an undivided HostModel) into npes parts: stiff_mat_ini ->
    partit(ssh_stiff%dim, rowptr, colind, nlevels_nod2D, np, part) (src/fvom_init.F90:1528) -> check_partitioning.
    Returns part (0-based PE per node), like the part vector behind dist_N/rpart.out."""
    import ctypes as C
    L = metis_lib(plain)
    if L is None:
        raise RuntimeError("oracle/_ref/libmetis_part.so has not been built (needs /root/reference)")
    nod2D = hm.ipar("myDim_nod2D")
    E = hm.ipar("myDim_elem2D")
    rowptr, colind = node_graph(hm["elem2D_nodes"][:E], nod2D)
    wgt = np.ascontiguousarray(hm["nlevels_nod2D"][:nod2D], dtype=np.int32)
    npv = np.zeros(10, dtype=np.int32)
    npv[0] = npes
    part = np.zeros(nod2D, dtype=np.int32)
    n = C.c_int32(nod2D)
    # fort_part.c reports on stdout (printf); bench.py owns stdout (ONE JSON line): send the C-level stdout to stderr meanwhile
    import sys
    sys.stdout.flush()
    keep = os.dup(1)
    os.dup2(2, 1)
    try:
        L.partit(C.byref(n), C.c_void_p(rowptr.ctypes.data), C.c_void_p(colind.ctypes.data), C.c_void_p(wgt.ctypes.data),
                 C.c_void_p(npv.ctypes.data), C.c_void_p(part.ctypes.data))
        C.CDLL(None).fflush(None)
    finally:
        os.dup2(keep, 1)
        os.close(keep)
    if check:
        check_partitioning(part, rowptr, colind, wgt, npes)
    return part
