"""Model of a second candidate numbering (not implemented): groups of 4 nodes, inside a group edges ordered by (rank, node), padded to
a multiple of 4 so that 4 neighbouring lanes read one 32-byte sector.  Result on the 274 k-node band mesh: 25 % padding, owned slots ideal
(8.0-8.6 sectors), the other slots 17, weighted mean 12.5 (reference order 22.7, tile order 11.6); edge kernels keep their locality."""
import sys, numpy as np
sys.path.insert(0, "/root/repo")
from tests import util
hm = util.build_model("band", nlon=740, nlat=370, nl=8, dt=600.0)
N = hm.ipar("myDim_nod2D"); Ed = hm.ipar("myDim_edge2D")
ed = hm["edges"].reshape(-1, 2) - 1; et = hm["edge_tri"].reshape(-1, 2) - 1; updn = hm["edge_up_dn_tri"].reshape(-1, 2) - 1
n1 = ed[:, 0]
rank = np.zeros(Ed, dtype=np.int64); o = np.argsort(n1, kind="stable"); st = np.searchsorted(n1[o], np.arange(N)); rank[o] = np.arange(Ed) - st[n1[o]]
own = np.bincount(n1, minlength=N)
G = (N + 3) // 4
ownp = np.zeros(G * 4, dtype=np.int64); ownp[:N] = own
gmax = ownp.reshape(G, 4).max(axis=1)
gbase = np.concatenate([[0], np.cumsum(4 * gmax)])
P = gbase[n1 >> 2] + 4 * rank + (n1 & 3)
nint = gbase[-1]
print("edges", Ed, "internal slots", nint, "holes %.2f%%" % (100.0 * (nint - Ed) / nint))
def sectors(idx):
    sec = np.where(idx >= 0, idx >> 2, -1)
    s = np.sort(sec, axis=1)
    distinct = (np.diff(s, axis=1) != 0).sum(axis=1) + 1 - (s[:, 0] == -1)
    act = (idx >= 0).sum(axis=1); m = act > 0
    return distinct[m].mean(), (act[m] / 4).mean()
node_of = np.concatenate([ed[:, 0], ed[:, 1]]); edge_old = np.concatenate([np.arange(Ed), np.arange(Ed)])
o2 = np.argsort(node_of.astype(np.int64) * Ed + edge_old)
node_of, edge_new = node_of[o2], P[edge_old[o2]]
start = np.searchsorted(node_of, np.arange(N)); deg = np.bincount(node_of, minlength=N); maxdeg = deg.max(); Nt = (N + 31) // 32
tot = 0.0
for j in range(maxdeg):
    m = deg > j
    se = -np.ones(Nt * 32, dtype=np.int64); se[:N][m] = edge_new[start[m] + j]
    s, ideal = sectors(se.reshape(Nt, 32)); tot += s * m.sum()
    print("slot %d edge-field gather sectors/request %.1f (ideal %.1f)" % (j, s, ideal))
print("weighted mean %.1f" % (tot / deg.sum()))
Et = (nint + 31) // 32
def sectors_ent(idx):
    sec = np.where(idx >= 0, (idx >> 5) * 1000003 + ((idx & 31) >> 2), -1)
    s = np.sort(sec, axis=1)
    distinct = (np.diff(s, axis=1) != 0).sum(axis=1) + 1 - (s[:, 0] == -1)
    act = (idx >= 0).sum(axis=1); m = act > 0
    return distinct[m].mean(), (act[m] / 4).mean()
def pad(a):
    b = -np.ones(Et * 32, dtype=np.int64); b[P] = a; return b.reshape(Et, 32)
for nm, a in [("bn1", ed[:, 0]), ("bn2", ed[:, 1]), ("be1", et[:, 0]), ("be2", et[:, 1]), ("bup", updn[:, 0]), ("bdn", updn[:, 1])]:
    print("edge kernel %s sectors/request %.1f (ideal %.1f)" % ((nm,) + sectors_ent(pad(a))))
