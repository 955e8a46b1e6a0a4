"""Builds dash_b200/libdash_b200.so (sm_100a only) with nvcc, in-tree."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libdash_b200.so")
SOURCES = ["kernels.cu", "dense_tc.cu", "dashb_api.cu", "garbler.cu"]
HEADERS = ["kernels.h", "gc_common.cuh", "garble_kernels.cuh", "shard_kernels.cuh", "../../include/dash_b200.h",
           "../../include/dash_b200_gcm.h", "../../include/dash_gcm_format.h", "../../include/dash_b200_garble.h"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "--extended-lambda",
         "-Xcompiler", "-fPIC", "-shared", "-Xptxas", "-v", "-diag-suppress", "20013"]


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(os.path.join(CSRC, f)) > t for f in SOURCES + HEADERS)


def build(force=False, verbose=False):
    if not force and not needs_build():
        return LIB
    extra = os.environ.get("DASHB_NVCC_EXTRA", "").split()
    cmd = [NVCC] + FLAGS + extra + ["-o", LIB] + [os.path.join(CSRC, s) for s in SOURCES] + ["-lcudart"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    with open(os.path.join(HERE, "build.log"), "w") as f:
        f.write(" ".join(cmd) + "\n" + res.stdout + res.stderr)
    if res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
        raise RuntimeError("nvcc failed building libdash_b200.so")
    if verbose:
        print(res.stderr)
    build_examples()
    return LIB


def build_examples():
    """integration/example_sharded.cpp: plain C++ against the C ABI (no CUDA, no reference headers)."""
    root = os.path.dirname(HERE)
    out_dir = os.path.join(root, "integration", "bin")
    os.makedirs(out_dir, exist_ok=True)
    cmd = ["g++", "-std=c++17", "-O2", "-I" + os.path.join(root, "include"), os.path.join(root, "integration", "example_sharded.cpp"),
           "-L" + HERE, "-ldash_b200", "-Wl,-rpath," + HERE, "-Wl,-rpath,$ORIGIN/../../dash_b200", "-pthread",
           "-o", os.path.join(out_dir, "example_sharded")]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
        raise RuntimeError("g++ failed building integration/bin/example_sharded")


if __name__ == "__main__":
    build(force=True, verbose=True)
