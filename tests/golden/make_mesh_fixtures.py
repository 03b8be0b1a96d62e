"""Pack the reference's own test meshes (test/meshes/pi, test/meshes/soufflet: plain-text files, SURVEY.md §4
fixtures) into compressed .npz inputs for the tests, because /root/reference does not exist on the GPU box.
Run in the authoring container:  python tests/golden/make_mesh_fixtures.py
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
from harness.host import read_mesh_files  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
for name in ("pi", "soufflet"):
    m = read_mesh_files(os.path.join("/root/reference/test/meshes", name))
    np.savez_compressed(os.path.join(HERE, name + "_mesh.npz"), **m)
    print(name, {k: (v.shape if hasattr(v, "shape") else v) for k, v in m.items()})

# the reference's own 2-way partition of the pi mesh (test/meshes/pi/dist_2: rpart.out, my_listNNNNN.out)
from harness import partition as P  # noqa: E402
d = {"part": P.read_rpart("/root/reference/test/meshes/pi/dist_2", 2)}
for me in range(2):
    for k, v in P.read_my_list("/root/reference/test/meshes/pi/dist_2", me).items():
        d["%s_%d" % (k, me)] = v
np.savez_compressed(os.path.join(HERE, "pi_dist2.npz"), **d)
print("pi_dist2", {k: (v.shape if hasattr(v, "shape") and v.shape else v) for k, v in d.items()})
