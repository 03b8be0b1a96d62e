"""One SSH solve on a band mesh of --nodes surface nodes, plain launches (no graph), for an ncu launch list:
    ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file out.csv python profiles/ssh_timeline.py --nodes 500000
Not a bench: the numbers under ncu are per-kernel durations (cold caches, serialised)."""
import argparse, ctypes as C, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
ap = argparse.ArgumentParser()
ap.add_argument("--nodes", type=int, default=500000)
ap.add_argument("--solves", type=int, default=2)
ap.add_argument("--graph", type=int, default=0)
a = ap.parse_args()
import bench
from fesom2_b200 import api
hm, info = bench.build_workload(a.nodes, 49)
api.lib().fb200_psolver_use_graph(C.byref(C.c_int(a.graph)))
s = api.SSHSolver(hm, pc="amg")
sol = np.zeros_like(hm["d_eta"])
ms = []
for k in range(a.solves):
    sol[:] = 0
    s.solve(new_values=1, sol=sol)
    ms.append(round(s.last_ms(), 3))
print("ms per solve", ms)
print("its, res", s.info(), "levels", s.amg_levels(), "ssh_ms", s.last_ms())
