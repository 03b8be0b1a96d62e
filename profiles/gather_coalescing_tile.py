"""Same sectors-per-request count as gather_coalescing.py for the experimental FB200_EDGE_ORDER=tile numbering of
fesom2_b200/csrc/context.cu (edges sorted by (tile of first node, rank among that node's edges, lane)); pure numpy on the mesh
numbering, no GPU.  Output: profiles/r01_gather_coalescing_tile.txt"""
import sys, numpy as np
sys.path.insert(0, "/root/repo")
from tests import util
hm = util.build_model("band", nlon=740, nlat=370, nl=8, dt=600.0)
N = hm.ipar("myDim_nod2D"); Ed = hm.ipar("myDim_edge2D")
ed = hm["edges"].reshape(-1, 2) - 1
et = hm["edge_tri"].reshape(-1, 2) - 1
updn = hm["edge_up_dn_tri"].reshape(-1, 2) - 1
n1 = ed[:, 0]
rank = np.zeros(Ed, dtype=np.int64)
o = np.argsort(n1, kind="stable")
st = np.searchsorted(n1[o], np.arange(N))
rank[o] = np.arange(Ed) - st[n1[o]]
order = np.lexsort((n1 & 31, rank, n1 >> 5))          # last key is primary
perm = np.empty(Ed, dtype=np.int64); perm[order] = np.arange(Ed)
def sectors(idx):
    sec = np.where(idx >= 0, (idx >> 5) * 1000003 + ((idx & 31) >> 2), -1)
    s = np.sort(sec, axis=1)
    distinct = (np.diff(s, axis=1) != 0).sum(axis=1) + 1 - (s[:, 0] == -1)
    act = (idx >= 0).sum(axis=1); m = act > 0
    return distinct[m].mean(), (act[m] / 4).mean()
for label, P in (("reference order", np.arange(Ed)), ("tile order", perm)):
    print("==", label)
    node_of = np.concatenate([ed[:, 0], ed[:, 1]]); edge_old = np.concatenate([np.arange(Ed), np.arange(Ed)]); other_of = np.concatenate([ed[:, 1], ed[:, 0]])
    o2 = np.argsort(node_of.astype(np.int64) * Ed + edge_old)          # per node in ascending REFERENCE edge id
    node_of, edge_new, other_of = node_of[o2], P[edge_old[o2]], other_of[o2]
    start = np.searchsorted(node_of, np.arange(N)); deg = np.bincount(node_of, minlength=N); maxdeg = deg.max()
    Nt = (N + 31) // 32
    tot = 0.0; wsum = 0.0
    for j in range(maxdeg):
        m = deg > j
        se = -np.ones(Nt * 32, dtype=np.int64); se[:N][m] = edge_new[start[m] + j]
        s, ideal = sectors(se.reshape(Nt, 32)); w = m.sum()
        tot += s * w; wsum += ideal * w
        print("slot %d edge-field gather sectors/request %.1f (ideal %.1f)" % (j, s, ideal))
    print("weighted mean over slots: %.1f sectors/request, ideal %.1f" % (tot / deg.sum(), wsum / deg.sum()))
    Et = (Ed + 31) // 32
    inv = np.empty(Ed, dtype=np.int64); inv[P] = np.arange(Ed)          # reference edge stored at internal position k
    def pad(a):
        b = -np.ones(Et * 32, dtype=np.int64); b[:Ed] = a[inv]; return b.reshape(Et, 32)
    for nm, a in [("bn1", ed[:, 0]), ("bn2", ed[:, 1]), ("be1", et[:, 0]), ("be2", et[:, 1]), ("bup", updn[:, 0]), ("bdn", updn[:, 1])]:
        print("edge kernel %s sectors/request %.1f (ideal %.1f)" % ((nm,) + sectors(pad(a))))
