"""In-tree builds of the TEST INFRASTRUCTURE (nothing here is part of the product library):

  harness/host/libfesom_host.so   harness/host/*.cpp          g++   host-side mirror of the Fortran setup code (mesh reading, ALE, SSH matrix)
  oracle/liboracle.so             oracle/*.c                  gcc   CPU restatement of the reference, -O2 -ffp-contract=off (parity)
  oracle/liboracle_timing.so      oracle/*.c                  gcc   the same sources with the reference's flags (bench.py's CPU arm only)
  oracle/_ref/...                 /root/reference pARMS + psolve.c, fort_part.c + METIS   gcc (only where /root/reference exists)
"""
import glob
import os

from fesom2_b200.build import ROOT, _BuildLock, _newer, _run

HOST_DIR = os.path.join(ROOT, "harness", "host")
HOST_SO = os.path.join(HOST_DIR, "libfesom_host.so")
ORACLE_SO = os.path.join(ROOT, "oracle", "liboracle.so")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libpsolve_ref.so")
REF_TIMING_SO = os.path.join(ROOT, "oracle", "_ref", "libpsolve_ref_timing.so")
ORACLE_TIMING_SO = os.path.join(ROOT, "oracle", "liboracle_timing.so")
ORACLE_TIMING_FLAGS = ["-O3", "-march=native", "-finline-functions"]      # the reference's own (src/CMakeLists.txt:73); FMA contraction allowed


def build_host(force=False):
    src = sorted(glob.glob(os.path.join(HOST_DIR, "*.cpp")))
    dep = src + glob.glob(os.path.join(HOST_DIR, "*.hpp"))
    if force or _newer(HOST_SO, dep):
        with _BuildLock(HOST_DIR):                         # the ranks of one job all get here at once
            if force or _newer(HOST_SO, dep):
                tmp = HOST_SO + ".%d.tmp" % os.getpid()
                _run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off", "-o", tmp] + src)
                os.replace(tmp, HOST_SO)
    return HOST_SO


def build_oracle(force=False):
    src = sorted(glob.glob(os.path.join(ROOT, "oracle", "*.c")))
    dep = src + glob.glob(os.path.join(ROOT, "oracle", "*.h"))
    if force or _newer(ORACLE_SO, dep):
        tmp = ORACLE_SO + ".%d.tmp" % os.getpid()
        _run(["gcc", "-O2", "-std=c99", "-fPIC", "-shared", "-ffp-contract=off", "-fopenmp", "-o", tmp] + src + ["-lm"])
        os.replace(tmp, ORACLE_SO)
    return ORACLE_SO


def build_oracle_timing(force=False):
    """A second copy of the oracle for bench.py's CPU arm only, compiled like the reference compiles its Fortran
    (-O3 -march=native -finline-functions, contraction allowed).  Built on the machine that runs it (the sources travel)."""
    src = sorted(glob.glob(os.path.join(ROOT, "oracle", "*.c")))
    dep = src + glob.glob(os.path.join(ROOT, "oracle", "*.h"))
    if force or _newer(ORACLE_TIMING_SO, dep):
        tmp = ORACLE_TIMING_SO + ".%d.tmp" % os.getpid()
        _run(["gcc"] + ORACLE_TIMING_FLAGS + ["-std=c99", "-fPIC", "-shared", "-fopenmp", "-o", tmp] + src + ["-lm"])
        os.replace(tmp, ORACLE_TIMING_SO)                  # atomic: several processes may get here at the same time
    return ORACLE_TIMING_SO


def build_ref(force=False):
    """Compile the reference's own pARMS + psolve.c from /root/reference (read in place) into oracle/_ref."""
    mk = os.path.join(ROOT, "oracle", "Makefile.ref")
    if not os.path.isdir("/root/reference/lib/parms") or not os.path.exists(mk):
        return REF_SO if os.path.exists(REF_SO) else None
    if force or not os.path.exists(REF_SO) or not os.path.exists(REF_TIMING_SO):
        with _BuildLock(os.path.join(ROOT, "oracle", "_ref")):
            if force or not os.path.exists(REF_SO) or not os.path.exists(REF_TIMING_SO):
                _run(["make", "-s", "-f", mk, "-j8", "_ref/libpsolve_ref.so", "_ref/libpsolve_ref_timing.so"], cwd=os.path.join(ROOT, "oracle"))
    return REF_SO


def build_all(force=False):
    from fesom2_b200.build import build_cuda
    build_host(force)
    build_oracle(force)
    build_ref(force)
    build_cuda(force)
