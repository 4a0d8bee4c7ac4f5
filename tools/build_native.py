#!/usr/bin/env python
"""Build libmlt_b200.so in-tree with nvcc for sm_100a (and nothing else).

Each ``csrc/*.cu|*.cpp`` is compiled to ``build/obj/*.o`` (in parallel, skipped when up to date) and linked
into ``ml_trainer_b200/lib/libmlt_b200.so``.  cudart is linked statically and the driver API is resolved at
run time, so the library also loads on a CPU-only box (host-side entry points still work there).
"""
from __future__ import annotations

import concurrent.futures as cf
import glob
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "csrc")
OBJ = os.path.join(ROOT, "build", "obj")
OUT_DIR = os.path.join(ROOT, "ml_trainer_b200", "lib")
OUT = os.path.join(OUT_DIR, "libmlt_b200.so")

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-Xcompiler", "-Wall",
          "--expt-relaxed-constexpr", "-I", os.path.join(ROOT, "include"), "-I", CSRC]


def nvcc() -> str:
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found")
    return exe


def _deps_mtime() -> float:
    hdrs = glob.glob(os.path.join(CSRC, "*.cuh")) + glob.glob(os.path.join(CSRC, "*.h")) + \
        glob.glob(os.path.join(ROOT, "include", "*.h"))
    return max((os.path.getmtime(h) for h in hdrs), default=0.0)


def compile_one(src: str, verbose: bool, ptxas_v: bool) -> str:
    obj = os.path.join(OBJ, os.path.basename(src) + ".o")
    newest = max(os.path.getmtime(src), _deps_mtime())
    if os.path.exists(obj) and os.path.getmtime(obj) >= newest:
        return obj
    cmd = [nvcc(), *ARCH, *COMMON, "-c", src, "-o", obj]
    if ptxas_v:
        cmd += ["-Xptxas", "-v"]
    if verbose:
        print(" ".join(cmd), flush=True)
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
    if (ptxas_v or verbose) and r.stderr.strip():
        print(r.stderr, flush=True)
    return obj


def build(verbose: bool = False, ptxas_v: bool = False, jobs: int | None = None) -> str:
    os.makedirs(OBJ, exist_ok=True)
    os.makedirs(OUT_DIR, exist_ok=True)
    srcs = sorted(glob.glob(os.path.join(CSRC, "*.cu")) + glob.glob(os.path.join(CSRC, "*.cpp")))
    if not srcs:
        raise RuntimeError("no sources under csrc/")
    with cf.ThreadPoolExecutor(max_workers=jobs or min(8, os.cpu_count() or 4)) as ex:
        objs = list(ex.map(lambda s: compile_one(s, verbose, ptxas_v), srcs))
    if os.path.exists(OUT) and all(os.path.getmtime(OUT) >= os.path.getmtime(o) for o in objs):
        return OUT
    cmd = [nvcc(), *ARCH, "-shared", "-o", OUT, *objs, "-cudart", "static", "-lpthread", "-ldl", "-lrt"]
    if verbose:
        print(" ".join(cmd), flush=True)
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return OUT


if __name__ == "__main__":
    v = "-v" in sys.argv
    print(build(verbose=v, ptxas_v="--ptxas" in sys.argv))
