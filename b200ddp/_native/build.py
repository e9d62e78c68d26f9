"""Builds ``libb200ddp.so`` in-tree with nvcc for sm_100a.

    python -m b200ddp._native.build [--force] [--verbose]

Each ``csrc/*.cu`` is compiled to an object in ``build/`` (in parallel, skipped
when up to date) and linked into ``b200ddp/_native/libb200ddp.so``.  No torch
headers are involved, so a full build takes seconds and works on a box without
a GPU (nvcc cross-compiles).
"""
from __future__ import annotations

import argparse
import concurrent.futures as cf
import glob
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
CSRC = os.path.join(ROOT, "csrc")
OBJ = os.path.join(ROOT, "build", "obj")
LIB = os.path.join(HERE, "libb200ddp.so")
ORACLE_SRC = os.path.join(ROOT, "oracle", "oracle.cpp")
ORACLE_LIB = os.path.join(HERE, "liboracle.so")
CXX = os.environ.get("CXX", "g++")

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
CFLAGS = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=default",
          "--expt-relaxed-constexpr", "-Xptxas", "-v"]


def _deps_mtime() -> float:
    hdrs = glob.glob(os.path.join(CSRC, "*.cuh")) + glob.glob(os.path.join(ROOT, "include", "*.h"))
    return max(os.path.getmtime(h) for h in hdrs) if hdrs else 0.0


def _compile(src: str, force: bool, verbose: bool):
    obj = os.path.join(OBJ, os.path.basename(src).rsplit(".", 1)[0] + ".o")
    if not force and os.path.exists(obj) and os.path.getmtime(obj) >= max(os.path.getmtime(src), _deps_mtime()):
        return obj, ""
    cmd = [NVCC, *ARCH, *CFLAGS, "-c", src, "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
    log = r.stderr
    with open(obj + ".ptxas.log", "w") as f:
        f.write(log)
    if verbose:
        print(log)
    return obj, log


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OBJ, exist_ok=True)
    srcs = sorted(glob.glob(os.path.join(CSRC, "*.cu")))
    if not srcs:
        raise RuntimeError(f"no sources under {CSRC}")
    with cf.ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        results = list(ex.map(lambda s: _compile(s, force, verbose), srcs))
    objs = [o for o, _ in results]
    if force or not os.path.exists(LIB) or any(os.path.getmtime(o) > os.path.getmtime(LIB) for o in objs):
        cmd = [NVCC, *ARCH, "-shared", "-o", LIB, *objs, "-lcudart"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    spills = [l for _, log in results for l in log.splitlines() if "spill" in l and "0 bytes spill stores, 0 bytes spill loads" not in l and "bytes spill" in l]
    if spills and verbose:
        print("register spills:\n" + "\n".join(spills))
    return LIB


def build_oracle(force: bool = False) -> str:
    """Host-only numerics oracle (oracle/oracle.cpp): plain g++, no CUDA.  -ffp-contract=off keeps
    the optimizer math shared with the kernels (csrc/optim_math.h) bit-exact."""
    deps = [ORACLE_SRC, os.path.join(CSRC, "optim_math.h"), os.path.join(ROOT, "include", "b200ddp.h")]
    if not force and os.path.exists(ORACLE_LIB) and os.path.getmtime(ORACLE_LIB) >= max(os.path.getmtime(d) for d in deps):
        return ORACLE_LIB
    cmd = [CXX, "-O2", "-std=c++17", "-ffp-contract=off", "-shared", "-fPIC", "-Wall", "-o", ORACLE_LIB, ORACLE_SRC]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"oracle build failed:\n{r.stdout}\n{r.stderr}")
    return ORACLE_LIB


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    a = ap.parse_args()
    print(build(a.force, a.verbose))
    print(build_oracle(a.force))
