"""In-tree build of the native libraries (no JIT cache: the built .so files travel to the GPU box).

  lib/libbsnls.so   CUDA kernels + C-ABI (nvcc, sm_100a only)
  lib/libbsfuse.so  C++ host mirror of the fuse graph interface + flattener + synthetic generator (g++)
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "lib")
CSRC = os.path.join(HERE, "csrc")
HOST = os.path.join(HERE, "host")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC",
              "-shared", "-cudart", "static"]
HOST_SOURCES = ["fuse_mirror.cpp", "flatten.cpp", "gpu_hash_graph.cpp", "preintegrator.cpp", "synthetic.cpp", "smoother_replay.cpp", "marginalize.cpp", "capi.cpp"]


def _newer(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def _deps(d, exts):
    return [os.path.join(d, f) for f in os.listdir(d) if f.endswith(exts)]


def build_cuda(force=False, verbose=False):
    os.makedirs(LIB, exist_ok=True)
    out = os.path.join(LIB, "libbsnls.so")
    deps = _deps(CSRC, (".cu", ".cuh", ".cpp")) + [os.path.join(HERE, "..", "include", "bsnls.h")]
    if force or _newer(out, deps):
        extra = os.environ.get("BSNLS_NVCC_EXTRA", "").split()   # A/B tuning of compile-time knobs (-DNAME=value)
        cmd = [NVCC] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-o", out, os.path.join(CSRC, "bsnls.cu"), os.path.join(CSRC, "noalloc.cu"), os.path.join(CSRC, "snapshot.cpp")]
        subprocess.check_call(cmd)
    return out


def build_host(force=False):
    os.makedirs(LIB, exist_ok=True)
    out = os.path.join(LIB, "libbsfuse.so")
    deps = _deps(HOST, (".cpp", ".h")) + [os.path.join(LIB, "libbsnls.so")]
    if force or _newer(out, deps):
        cmd = ["g++", "-std=c++17", "-O2", "-march=x86-64-v3", "-fPIC", "-shared", "-pthread", "-Wall", "-o", out] + \
              [os.path.join(HOST, s) for s in HOST_SOURCES] + ["-L" + LIB, "-lbsnls", "-Wl,-rpath,$ORIGIN"]
        subprocess.check_call(cmd)
    return out


def build_all(force=False, verbose=False):
    return build_cuda(force, verbose), build_host(force)


if __name__ == "__main__":
    print(build_all(force="--force" in sys.argv, verbose="-v" in sys.argv))
