"""TEST INFRASTRUCTURE — CUDA-on-CPU emulation of libfesom_b200 (see cuda_emu.h).

build() rewrites fesom2_b200/csrc/*.cu into plain C++ (kernel launches become calls of emu_launch, dynamic shared memory a
per-launch buffer) and compiles them with g++ -ffp-contract=off (the counterpart of nvcc --fmad=false) together with the
fiber scheduler (cuda_emu.cpp) and an NCCL-over-files stand-in (nccl.h, nccl_emu.cpp) into
tests/emu/_build/libfesom_b200_emu.so.  load() returns it as a ctypes library set up like fesom2_b200.api.lib().
Used by tests/test_emulated*.py, by tests/mp_check.py --emulate and, with FB200_EMULATE=1, by tests/conftest.py; never by
the product."""
import ctypes as C
import os
import re
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
CSRC = os.path.join(ROOT, "fesom2_b200", "csrc")
OUT = os.path.join(HERE, "_build")
SO = os.path.join(OUT, "libfesom_b200_emu.so")
SOURCES = ["tracer_kernels.cu", "tracer_kernels2.cu", "context.cu", "capi.cu", "comm.cu", "ssh_solver.cu", "ssh_amg.cu"]


def _match_paren(s, i):
    """s[i] == '(' -> index of the matching ')'"""
    d = 0
    for j in range(i, len(s)):
        if s[j] == "(":
            d += 1
        elif s[j] == ")":
            d -= 1
            if d == 0:
                return j
    raise ValueError("unbalanced parentheses")


def transform(src):
    """kernel<<<cfg>>>(args)  ->  emu_launch(emu_cfg(cfg), [&]() { kernel(args); });  extern __shared__ T n[] -> pointer"""
    src = re.sub(r"extern\s+__shared__\s+(\w+)\s+(\w+)\[\];", r"\1* \2 = (\1*)emu_dyn_smem();", src)
    out, pos = [], 0
    while True:
        k = src.find("<<<", pos)
        if k < 0:
            out.append(src[pos:])
            break
        # kernel expression: identifier, optionally followed by <template args>, right before <<<
        b = k
        if src[b - 1] == ">":
            d = 0
            while True:
                b -= 1
                if src[b] == ">":
                    d += 1
                elif src[b] == "<":
                    d -= 1
                    if d == 0:
                        break
        while b > 0 and (src[b - 1].isalnum() or src[b - 1] == "_"):
            b -= 1
        kern = src[b:k]
        e = src.find(">>>", k)
        cfg = src[k + 3:e]
        p0 = e + 3
        assert src[p0] == "(", src[k - 40:k + 80]
        p1 = _match_paren(src, p0)
        args = src[p0 + 1:p1]
        out.append(src[pos:b])
        out.append("emu_launch(emu_cfg(%s), [&]() { %s(%s); })" % (cfg, kern, args))
        pos = p1 + 1
    return "".join(out)


def build(force=False, asan=False):
    """asan=True: a second library with -fsanitize=address (heap redzones around every emulated device allocation: run it
    with LD_PRELOAD=$(g++ -print-file-name=libasan.so) to find out-of-bounds indexing in the kernels)"""
    if asan:
        return _build(force, os.path.join(OUT, "libfesom_b200_emu_asan.so"), ["-O1", "-g", "-fsanitize=address", "-fno-omit-frame-pointer"])
    return _build(force, SO, ["-O2"])


def _build(force, SO, optflags):
    os.makedirs(OUT, exist_ok=True)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh"))] + \
           [os.path.join(HERE, f) for f in os.listdir(HERE) if f.endswith((".h", ".cpp", ".py"))]
    if not force and os.path.exists(SO) and all(os.path.getmtime(SO) >= os.path.getmtime(d) for d in deps):
        return SO
    cpp = []
    for f in SOURCES:
        t = transform(open(os.path.join(CSRC, f)).read())
        dst = os.path.join(OUT, f.replace(".cu", "_emu.cpp"))
        with open(dst, "w") as fh:
            fh.write(t)
        cpp.append(dst)
    cpp += [os.path.join(HERE, "cuda_emu.cpp"), os.path.join(HERE, "nccl_emu.cpp")]
    cmd = ["g++", "-std=c++17", "-DFB200_EMULATION"] + optflags + ["-ffp-contract=off", "-fPIC", "-shared", "-include", os.path.join(HERE, "cuda_emu.h"),
           "-I", HERE, "-I", CSRC, "-I", os.path.join(ROOT, "include"), "-o", SO + ".tmp"] + cpp
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("emulation build failed:\n" + r.stdout[-6000:])
    os.replace(SO + ".tmp", SO)                            # a process that has the old library mapped keeps its copy
    return SO


def load():
    L = C.CDLL(build(asan=bool(os.environ.get("FB200_EMU_ASAN"))))
    L.fb200_status.restype = C.c_int
    L.fb200_status_message.restype = C.c_char_p
    L.fb200_kernel_launches.restype = C.c_longlong
    L.fb200_last_ms.restype = C.c_double
    L.fb200_last_ms.argtypes = [C.c_char_p]
    L.fb200_debug_fetch.argtypes = [C.c_char_p, C.c_void_p]
    L.fb200_set_stream.argtypes = [C.c_void_p]
    L.fb200_psolver_use_graph(C.byref(C.c_int(0)))      # CUDA graphs are not emulated: plain launches (the same kernels)
    return L
