"""In-tree build of the product (no JIT cache: the .so travels with the repo snapshot).

  libfesom_b200.so   fesom2_b200/csrc/*.cu        nvcc     CUDA kernels + C ABI (sm_100a)

The builds of the test infrastructure (host-side mirror of the Fortran setup code, oracle, the reference's own psolve.c + pARMS
and partitioner) live in harness/build.py.
"""
import glob
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "fesom2_b200")
CUDA_SO = os.path.join(PKG, "csrc", "libfesom_b200.so")


def _newer(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


class _BuildLock:
    """One builder at a time per output directory (flock on a lock file): several ranks of one job — or pytest-xdist workers — that
    find the same library stale must not compile and link into the same files at once; the ones that waited find it fresh."""

    def __init__(self, directory):
        os.makedirs(directory, exist_ok=True)
        self.path = os.path.join(directory, ".build.lock")
        self.fd = None

    def __enter__(self):
        import fcntl
        self.fd = open(self.path, "w")
        fcntl.flock(self.fd, fcntl.LOCK_EX)
        return self

    def __exit__(self, *exc):
        import fcntl
        fcntl.flock(self.fd, fcntl.LOCK_UN)
        self.fd.close()
        return False


def _run(cmd, cwd=None):
    r = subprocess.run(cmd, cwd=cwd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        sys.stderr.write(" ".join(cmd) + "\n" + r.stdout)
        raise RuntimeError("build failed: " + " ".join(cmd[:3]))
    return r.stdout


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

    def stale():
        return [(f, o) for f, o in zip(src, objs) if force or _newer(o, [f] + hdr)]
    if not (stale() or force or _newer(so, objs)):
        return so
    if nvcc is None:
        if os.path.exists(so) and not force:
            return so                      # a box without nvcc uses the library that travelled with the snapshot
        raise RuntimeError("nvcc not found: cannot build libfesom_b200.so")
    with _BuildLock(objdir):
        todo = stale()                     # again under the lock: another process may just have built it
        if not (todo or _newer(so, objs)):
            return so
        cflags = [x for x in NVCC_FLAGS if x != "-shared" and not (fmad and x == "--fmad=false")] + ["-c"] + ["-D" + d for d in defines]
        if verbose:
            cflags += ["-Xptxas", "-v"]
        nf = nccl_flags()
        inc = ["-I", os.path.join(ROOT, "include")] + (nf[nf.index("-I"):nf.index("-I") + 2] if "-I" in nf else [])

        def one(fo):
            tmp = fo[1] + ".%d.tmp" % os.getpid()
            out_ = _run([nvcc] + cflags + inc + ["-o", tmp, fo[0]])
            os.replace(tmp, fo[1])
            return out_
        with cf.ThreadPoolExecutor(max_workers=8) as ex:
            outs = list(ex.map(one, todo))
        if verbose:
            print("\n".join(outs))
        tmp_so = so + ".%d.tmp" % os.getpid()
        _run([nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-Xcompiler", "-fPIC", "-shared", "-o", tmp_so] + objs + nccl_flags())
        os.replace(tmp_so, so)             # a process that already maps the old file keeps it; nobody ever sees a half-written one
    return so


