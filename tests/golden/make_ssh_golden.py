"""Golden vectors for the SSH solve, produced by the reference's OWN code: src/psolve.c + lib/parms compiled by
oracle/Makefile.ref from /root/reference.  Run in the authoring container:  python tests/golden/make_ssh_golden.py
Each file holds the CSR system (init_stiff_mat_ale + compute_ssh_rhs_ale of the host mirror on the reference's
test mesh) and x_ref = psolve() output for a zero initial guess."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
from oracle import binding as ob  # noqa: E402
from tests import util  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
for name in ("pi", "soufflet"):
    hm = util.build_model(name)
    hm.compute_ssh_rhs()
    rp = hm["ssh_rowptr"]
    n = hm.ipar("myDim_nod2D")
    rp0, ci0 = (rp - rp[0]).astype(np.int32), (hm["ssh_colind"] - 1).astype(np.int32)
    vals, rhs = hm["ssh_values"].copy(), hm["ssh_rhs"][:n].copy()
    x = ob.RefSolver(rp0, ci0, vals).solve(rhs, vals, np.zeros(n), new_values=0)
    np.savez_compressed(os.path.join(HERE, "ssh_%s.npz" % name), rowptr0=rp0, colind0=ci0, values=vals, rhs=rhs, x_ref=x)
    print(name, n, rp0[-1], float(np.abs(x).max()))
