"""TEST / BENCH INFRASTRUCTURE — not part of the product.  The host side a FESOM driver would own (mesh arrays, ALE
producers, SSH matrix: harness/host, restated from the Fortran), the reference's partitioning rules (partition.py), synthetic
meshes (synth.py) and the builds of the oracle and of the reference's own C code (build.py).  The product is fesom2_b200/."""
