"""In-tree build of the native libraries (nvcc cross-compiles sm_100a without a GPU).

    libhelfem_b200.so   C ABI (include/helfem_b200.h): CUDA kernels + context, plus the host-side
                        mirror classes (host/*.cpp) and their flat entry points (host/hostapi.cpp)

Objects are cached under helfem_b200/_build keyed on source mtimes, so a rebuild after
touching one file recompiles only that file.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
BUILD = os.path.join(HERE, "_build")
LIB = os.path.join(HERE, "libhelfem_b200.so")

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
NVCC_FLAGS = ARCH + ["-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC,-fopenmp,-O2"]
NVCC_FLAGS += os.environ.get("HFB200_NVCC_EXTRA", "").split()  # experiments: extra -D switches (rebuild with --force)

CU_SOURCES = ["csrc/capi_common.cu", "csrc/capi_sadatom.cu", "csrc/sadatom_grid_lda.cu", "csrc/sadatom_grid_gga.cu", "csrc/sadatom_grid_mgga.cu",
              "csrc/capi_sadatom_exchange.cu", "csrc/capi_xc.cu",
              "csrc/chan_engine.cu", "csrc/comm.cu", "csrc/capi_atomic.cu", "csrc/capi_diatomic.cu", "csrc/grid2d.cu", "csrc/capi_grid2d.cu"]
CPP_SOURCES = ["host/fem.cpp", "host/angular.cpp", "host/sadatom.cpp", "host/atomic.cpp", "host/legendre.cpp", "host/diatomic.cpp", "host/dftgrid.cpp", "host/hostapi.cpp", "host/blockscan.cpp"]


def _headers():
    out = []
    for d in ("csrc", "host"):
        for f in os.listdir(os.path.join(HERE, d)):
            if f.endswith((".h", ".cuh", ".hpp")):
                out.append(os.path.join(HERE, d, f))
    out.append(os.path.join(ROOT, "include", "helfem_b200.h"))
    return out


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(verbose=False, force=False):
    os.makedirs(BUILD, exist_ok=True)
    hdrs = _headers()
    objs = []
    procs = []
    for src in CU_SOURCES + CPP_SOURCES:
        path = os.path.join(HERE, src)
        obj = os.path.join(BUILD, src.replace("/", "_") + ".o")
        objs.append(obj)
        if force or _stale(obj, [path] + hdrs):
            cmd = [NVCC] + NVCC_FLAGS + (["-x", "cu"] if src.endswith(".cu") else []) + ["-c", path, "-o", obj]
            if verbose:
                print(" ".join(cmd), flush=True)
            procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
    for src, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            sys.stderr.write(out.decode())
            raise RuntimeError("compilation of %s failed" % src)
        if verbose and out:
            print(out.decode())
    if force or procs or _stale(LIB, objs):
        cmd = [NVCC] + ARCH + ["-shared", "-o", LIB] + objs + ["-Xcompiler", "-fopenmp", "-lgomp", "-ldl"]
        if verbose:
            print(" ".join(cmd), flush=True)
        subprocess.check_call(cmd)
    return LIB


if __name__ == "__main__":
    print(build(verbose=True, force="--force" in sys.argv))
