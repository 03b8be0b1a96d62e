"""In-tree builds of the native pieces (no JIT cache: the .so files travel with the repo snapshot).

  libfesom_host.so   fesom2_b200/host/*.cpp       g++      host-side mirror of the Fortran setup code
  libfesom_b200.so   fesom2_b200/csrc/*.cu        nvcc     the product: CUDA kernels + C ABI (sm_100a)
  liboracle.so       oracle/*.c                   gcc      CPU restatement of the reference (test infrastructure)
  oracle/_ref/...    /root/reference pARMS + psolve.c  gcc (only when /root/reference exists)
"""
import glob
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "fesom2_b200")
HOST_SO = os.path.join(PKG, "host", "libfesom_host.so")
CUDA_SO = os.path.join(PKG, "csrc", "libfesom_b200.so")
ORACLE_SO = os.path.join(ROOT, "oracle", "liboracle.so")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libpsolve_ref.so")
REF_TIMING_SO = os.path.join(ROOT, "oracle", "_ref", "libpsolve_ref_timing.so")
ORACLE_TIMING_SO = os.path.join(ROOT, "oracle", "liboracle_timing.so")
ORACLE_TIMING_FLAGS = ["-O3", "-march=native", "-finline-functions"]      # the reference's own (src/CMakeLists.txt:73); FMA contraction allowed


def _newer(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def _run(cmd, cwd=None):
    r = subprocess.run(cmd, cwd=cwd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        sys.stderr.write(" ".join(cmd) + "\n" + r.stdout)
        raise RuntimeError("build failed: " + " ".join(cmd[:3]))
    return r.stdout


def build_host(force=False):
    src = sorted(glob.glob(os.path.join(PKG, "host", "*.cpp")))
    dep = src + glob.glob(os.path.join(PKG, "host", "*.hpp"))
    if force or _newer(HOST_SO, dep):
        _run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off", "-o", HOST_SO] + src)
    return HOST_SO


def nvcc_path():
    for p in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if p and os.path.exists(p):
            return p
    return None


NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "--fmad=false", "-Xcompiler", "-fPIC", "-shared"]


def nccl_flags():
    """NCCL to link against: the copy bundled with PyTorch when there is one (both libraries have the soname
    libnccl.so.2, and a process that also imports torch must not end up with the older system copy loaded first),
    else the system library.  A Fortran-only host links the system NCCL."""
    for sp in sys.path:
        lib = os.path.join(sp, "nvidia", "nccl", "lib")
        inc = os.path.join(sp, "nvidia", "nccl", "include")
        if os.path.exists(os.path.join(lib, "libnccl.so.2")) and os.path.exists(os.path.join(inc, "nccl.h")):
            return ["-I", inc, "-L", lib, "-l:libnccl.so.2", "-Xlinker", "-rpath", "-Xlinker", lib]
    return ["-lnccl"]


def build_cuda(force=False, verbose=False, out=None, defines=(), fmad=False):
    """every .cu -> its own object (in parallel, rebuilt only when it or a header changed), then one link.
    out / defines: a variant library for A/B measurements (FB200_SO selects it at run time, see api.lib)"""
    import concurrent.futures as cf
    so = out or CUDA_SO
    tag = "" if not out else "." + os.path.splitext(os.path.basename(out))[0]
    src = sorted(glob.glob(os.path.join(PKG, "csrc", "*.cu")))
    hdr = glob.glob(os.path.join(PKG, "csrc", "*.cuh")) + glob.glob(os.path.join(ROOT, "include", "*.h"))
    objdir = os.path.join(PKG, "csrc", "build")
    nvcc = nvcc_path()
    objs = [os.path.join(objdir, os.path.basename(f)[:-3] + tag + ".o") for f in src]
    todo = [(f, o) for f, o in zip(src, objs) if force or _newer(o, [f] + hdr)]
    if todo or force or _newer(so, objs):
        if nvcc is None:
            if os.path.exists(so) and not force:
                return so                  # a box without nvcc uses the library that travelled with the snapshot
            raise RuntimeError("nvcc not found: cannot build libfesom_b200.so")
        os.makedirs(objdir, exist_ok=True)
        cflags = [x for x in NVCC_FLAGS if x != "-shared" and not (fmad and x == "--fmad=false")] + ["-c"] + ["-D" + d for d in defines]
        if verbose:
            cflags += ["-Xptxas", "-v"]
        nf = nccl_flags()
        inc = ["-I", os.path.join(ROOT, "include")] + (nf[nf.index("-I"):nf.index("-I") + 2] if "-I" in nf else [])

        def one(fo):
            return _run([nvcc] + cflags + inc + ["-o", fo[1], fo[0]])
        with cf.ThreadPoolExecutor(max_workers=8) as ex:
            outs = list(ex.map(one, todo))
        if verbose:
            print("\n".join(outs))
        _run([nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-Xcompiler", "-fPIC", "-shared", "-o", so] + objs + nccl_flags())
    return so


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
        _run(["make", "-s", "-f", mk, "-j8", "_ref/libpsolve_ref.so", "_ref/libpsolve_ref_timing.so"], cwd=os.path.join(ROOT, "oracle"))
    return REF_SO


def build_all(force=False):
    build_host(force)
    build_oracle(force)
    build_ref(force)
    build_cuda(force)
