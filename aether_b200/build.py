"""Build the CUDA library in-tree with nvcc for sm_100a.

    python -m aether_b200.build          # libaether_b200.so (+ the CPU-side checkers)

Outputs go to ``aether_b200/lib/`` (git-ignored, shipped to the GPU box by gpurun).  nvcc
cross-compiles without a GPU.  Objects are rebuilt only when a source or header is newer.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "lib")
OBJ = os.path.join(LIB, "obj")
SO = os.path.join(LIB, "libaether_b200.so")
HOSTCHECK_SO = os.path.join(LIB, "libaether_hostcheck.so")

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-I", os.path.join(ROOT, "include")]
# extra nvcc flags for experiments (e.g. AETHER_NVCC_FLAGS=-DAETHER_TC_REGSPLIT with `-f`); empty for the shipped build
COMMON += os.environ.get("AETHER_NVCC_FLAGS", "").split()

# (source, extra flags).  The geometry files are built without FMA contraction so that products and
# sums round separately, like the reference's eager ATen ops.
SOURCES = [
    ("gemm_f32.cu", []),
    ("thin.cu", []),
    ("segment.cu", []),
    ("graph.cu", []),
    ("step.cu", []),
    ("tc_gemm.cu", []),
    ("tc_wgrad.cu", []),
    ("tc_chain.cu", []),
    ("tc_chain_bwd.cu", []),
    ("tc_abi.cu", []),
    ("edge_bf16.cu", []),
    ("prof.cu", []),
    ("optim.cu", []),
    ("peer.cu", []),
    ("field.cu", []),
    ("dynfield.cu", []),
    ("geom.cu", ["-fmad=false"]),
    ("geom_bwd.cu", ["-fmad=false"]),
    ("small.cu", ["-fmad=false"]),
    ("seqdec.cu", ["-fmad=false"]),
]


def _headers():
    hs = [os.path.join(ROOT, "include", "aether_b200.h")]
    for f in os.listdir(CSRC):
        if f.endswith((".cuh", ".inl", ".h")):
            hs.append(os.path.join(CSRC, f))
    return hs


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _run(cmd):
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
        raise RuntimeError("nvcc failed: " + " ".join(cmd[:6]))
    return r.stdout + r.stderr


def build(verbose: bool = False, force: bool = False) -> str:
    os.makedirs(OBJ, exist_ok=True)
    hdrs = _headers()
    jobs = []
    objs = []
    for src, extra in SOURCES:
        s = os.path.join(CSRC, src)
        o = os.path.join(OBJ, src.replace(".cu", ".o"))
        objs.append(o)
        if force or _stale(o, [s] + hdrs):
            jobs.append([NVCC] + ARCH + COMMON + extra + ["-c", s, "-o", o])
    if jobs:
        with ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            for out in ex.map(_run, jobs):
                if verbose and out.strip():
                    print(out)
    if force or jobs or _stale(SO, objs):
        _run([NVCC] + ARCH + ["-shared", "-o", SO] + objs + ["-lcudart"])
    return SO


def build_hostcheck(force: bool = False) -> str:
    """CPU-only build of the geometry bodies (the same __host__ __device__ code the kernels run),
    used by the not-gpu tests to check the hand-derived backward against autograd."""
    os.makedirs(LIB, exist_ok=True)
    src = os.path.join(ROOT, "tests", "host_geom_check.cu")
    deps = [src] + _headers() + [os.path.join(CSRC, "geom.cu"), os.path.join(CSRC, "geom_bwd.cu")]
    if force or _stale(HOSTCHECK_SO, deps):
        _run([NVCC] + ARCH + COMMON + ["-fmad=false", "-shared", "-o", HOSTCHECK_SO, src, "-lcudart"])
    return HOSTCHECK_SO


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv, force="-f" in sys.argv))
    if os.path.exists(os.path.join(ROOT, "tests", "host_geom_check.cu")):
        print(build_hostcheck(force="-f" in sys.argv))
