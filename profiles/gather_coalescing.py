import sys, numpy as np
sys.path.insert(0,"/root/repo")
from tests import util
hm = util.build_model("band", nlon=740, nlat=370, nl=8, dt=600.0)
N = hm.ipar("myDim_nod2D"); Ed = hm.ipar("myDim_edge2D")
edges = hm["edges"] ; print(edges.shape)
ed = (edges.reshape(-1,2) if edges.shape[-1]==2 else edges.T) - 1
et = hm["edge_tri"]; et = (et.reshape(-1,2) if et.shape[-1]==2 else et.T) - 1
updn = hm["edge_up_dn_tri"]; updn = (updn.reshape(-1,2) if updn.shape[-1]==2 else updn.T) - 1
print("edges", ed[:5], et[:5])
def sectors(idx):
    """idx: (ntile, 32) entity ids per lane (-1 = inactive) -> mean distinct 32B sectors per request"""
    tot=0; cnt=0
    sec = np.where(idx>=0, (idx>>5)*1000003 + ((idx&31)>>2), -1)
    s = np.sort(sec, axis=1)
    distinct = (np.diff(s, axis=1)!=0).sum(axis=1) + 1 - (s[:,0]==-1)
    act = (idx>=0).sum(axis=1)
    m = act>0
    return distinct[m].mean(), (act[m]/4).mean()
# node kernels: slot tables
order = np.argsort(np.concatenate([ed[:,0], ed[:,1]]), kind="stable")
node_of = np.concatenate([ed[:,0], ed[:,1]])[order]
edge_of = np.concatenate([np.arange(Ed), np.arange(Ed)])[order]
other_of = np.concatenate([ed[:,1], ed[:,0]])[order]
# sort by (node, edge id)
key = node_of.astype(np.int64)*Ed + edge_of
o2 = np.argsort(key); node_of, edge_of, other_of = node_of[o2], edge_of[o2], other_of[o2]
start = np.searchsorted(node_of, np.arange(N)); deg = np.bincount(node_of, minlength=N)
maxdeg = deg.max(); print("maxdeg", maxdeg, "mean deg", deg.mean())
Nt = (N+31)//32
slot_e = -np.ones((Nt*32, maxdeg), dtype=np.int64); slot_o = -np.ones((Nt*32, maxdeg), dtype=np.int64)
for j in range(maxdeg):
    m = deg>j
    slot_e[:N][m, j] = edge_of[start[m]+j]; slot_o[:N][m, j] = other_of[start[m]+j]
for j in range(maxdeg):
    se, ideal = sectors(slot_e[:, j].reshape(Nt,32)); so,_ = sectors(slot_o[:, j].reshape(Nt,32))
    print("slot", j, "edge-gather sectors/request %.1f  other-node %.1f (ideal %.1f)"%(se, so, ideal))
# edge kernels
Et = (Ed+31)//32
def pad(a):
    b = -np.ones(Et*32, dtype=np.int64); b[:Ed]=a; return b.reshape(Et,32)
for nm, a in [("bn1", ed[:,0]), ("bn2", ed[:,1]), ("be1", et[:,0]), ("be2", et[:,1]), ("bup", updn[:,0]), ("bdn", updn[:,1])]:
    print("edge kernel", nm, "sectors/request %.1f (ideal %.1f)"%sectors(pad(a)))
