"""Builds the in-tree native libraries.

  libbrotligpu.so   nvcc, sm_100a only (the product: kernels + C ABI)
  libbgpu_hosttest.so  g++ build of the __host__ __device__ format code for the CPU test tier
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libbrotligpu.so")
HOSTTEST = os.path.join(HERE, "libbgpu_hosttest.so")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-shared",
              "-Xcompiler", "-fPIC", "-cudart", "static"]


def _stale(target, srcs):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in srcs)


def _sources():
    return [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC))] + [os.path.join(HERE, "..", "include", "brotligpu.h")]


def build_gpu(force=False, verbose=False):
    srcs = _sources()
    if force or _stale(LIB, srcs):
        nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
        cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB] + \
              [os.path.join(CSRC, "bgpu.cu"), os.path.join(CSRC, "zopfli.cu")]
        cmd = [c for c in cmd if not c.endswith("zopfli.cu") or os.path.exists(c)]
        subprocess.check_call(cmd)
    return LIB


def build_hosttest(force=False):
    srcs = _sources()
    if force or _stale(HOSTTEST, srcs):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-x", "c++", "-o", HOSTTEST,
                               os.path.join(CSRC, "hosttest.cc")])
    return HOSTTEST


if __name__ == "__main__":
    build_gpu(force="--force" in sys.argv, verbose="-v" in sys.argv)
    build_hosttest(force="--force" in sys.argv)
    print(LIB)
