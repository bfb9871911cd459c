#!/usr/bin/env python3
"""Build libpdnn_b200.so (all sm_100a kernels) in-tree with nvcc.

    python csrc/build.py [--force] [--verbose]

The shared object lands in pytorch_distributed_nn_b200/lib/ so that it travels with the source
snapshot to the GPU box (a JIT cache would not). nvcc cross-compiles without a GPU.
"""
import argparse
import concurrent.futures as cf
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "csrc")
OUT_DIR = os.path.join(ROOT, "pytorch_distributed_nn_b200", "lib")
OBJ_DIR = os.path.join(ROOT, "build", "obj")
LIB = os.path.join(OUT_DIR, "libpdnn_b200.so")
SOURCES = ["core.cu", "gemm_tcgen05.cu", "conv_halo.cu", "conv_stem.cu", "nn_kernels.cu", "ps_kernels.cu", "fc_ps_native.cu"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=default", "--expt-relaxed-constexpr",
]


def _newer(src_list, target):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in src_list)


def build(force=False, verbose=False):
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    os.makedirs(OUT_DIR, exist_ok=True)
    os.makedirs(OBJ_DIR, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    headers.append(os.path.join(ROOT, "include", "pdnn_b200.h"))
    sources = [s for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    jobs = []
    for s in sources:
        src = os.path.join(CSRC, s)
        obj = os.path.join(OBJ_DIR, s.replace(".cu", ".o"))
        if force or _newer([src] + headers, obj):
            cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", src, "-o", obj]
            jobs.append((s, cmd))
    def run(job):
        name, cmd = job
        r = subprocess.run(cmd, capture_output=True, text=True)
        return name, r.returncode, r.stdout + r.stderr
    if jobs:
        with cf.ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            for name, rc, out in ex.map(run, jobs):
                if verbose or rc != 0:
                    sys.stderr.write(f"--- {name} ---\n{out}\n")
                if rc != 0:
                    raise RuntimeError(f"nvcc failed on {name}")
    objs = [os.path.join(OBJ_DIR, s.replace(".cu", ".o")) for s in sources]
    if force or jobs or _newer(objs, LIB):
        cmd = [nvcc, "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-lpthread"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            sys.stderr.write(r.stdout + r.stderr)
            raise RuntimeError("link failed")
    return LIB


def build_native_driver(force=False):
    """build/distributed_nn_native: the stand-alone FC parameter-server program (counterpart of MPI_code's `make distributed`)."""
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    src = os.path.join(CSRC, "fc_ps_main.cpp")
    out = os.path.join(ROOT, "build", "distributed_nn_native")
    if force or _newer([src, LIB, os.path.join(ROOT, "include", "pdnn_b200.h")], out):
        cmd = [nvcc] + NVCC_FLAGS + [src, "-o", out, "-L" + OUT_DIR, "-lpdnn_b200", "-Xlinker", "-rpath", "-Xlinker",
                                    "$ORIGIN/../pytorch_distributed_nn_b200/lib"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            sys.stderr.write(r.stdout + r.stderr)
            raise RuntimeError("native driver build failed")
    return out


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    a = ap.parse_args()
    print(build(a.force, a.verbose))
    print(build_native_driver(a.force))
