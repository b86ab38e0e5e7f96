#!/usr/bin/env python
"""Build libfedb200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

    python build.py            # incremental: only sources newer than their object are rebuilt
    python build.py --force    # rebuild everything
    python build.py --ptxas-v  # print registers / spills / smem per kernel
"""
from __future__ import annotations

import argparse
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(ROOT, "csrc")
OUT_DIR = os.path.join(ROOT, "fedb200", "lib")
OBJ_DIR = os.path.join(ROOT, "build", "obj")
LIB = os.path.join(OUT_DIR, "libfedb200.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "--use_fast_math",
         "-Xcompiler", "-fPIC", "-Xcompiler", "-Wall", "-Xcudafe", "--diag_suppress=177",
         "-I", os.path.join(ROOT, "include")]


def sources():
    return sorted(f for f in os.listdir(SRC) if f.endswith(".cu"))


def headers_mtime():
    hs = [os.path.join(SRC, f) for f in os.listdir(SRC) if f.endswith((".cuh", ".h"))]
    hs += [os.path.join(ROOT, "include", f) for f in os.listdir(os.path.join(ROOT, "include"))]
    return max(os.path.getmtime(h) for h in hs)


# IEEE division / sqrt and denormals are kept where fp32 master state is updated (AdamW: sqrt(v) + eps on v ~ 1e-40 for
# tiny gradients) and in the fp64/fp32 PageRank scorer; everything else is bf16-rounded anyway and uses --use_fast_math.
PRECISE = {"elementwise.cu", "pagerank.cu"}


def compile_one(name: str, force: bool, verbose_ptxas: bool):
    src, obj = os.path.join(SRC, name), os.path.join(OBJ_DIR, name[:-3] + ".o")
    if not force and os.path.exists(obj) and os.path.getmtime(obj) > max(os.path.getmtime(src), headers_mtime()):
        return name, 0, "", False
    flags = [f for f in FLAGS if not (name in PRECISE and f == "--use_fast_math")]
    cmd = [NVCC, *flags, "-c", src, "-o", obj]
    if verbose_ptxas:
        cmd[1:1] = ["-Xptxas", "-v"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    return name, r.returncode, r.stdout + r.stderr, True


def build(force: bool = False, verbose_ptxas: bool = False, quiet: bool = False) -> str:
    os.makedirs(OUT_DIR, exist_ok=True)
    os.makedirs(OBJ_DIR, exist_ok=True)
    srcs = sources()
    rebuilt = False
    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        for name, rc, log, did in ex.map(lambda n: compile_one(n, force, verbose_ptxas), srcs):
            rebuilt |= did
            if log.strip() and (rc != 0 or verbose_ptxas or not quiet):
                print(f"--- {name}\n{log.strip()}", file=sys.stderr)
            if rc != 0:
                raise RuntimeError(f"nvcc failed on {name}")
    objs = [os.path.join(OBJ_DIR, n[:-3] + ".o") for n in srcs]
    if rebuilt or not os.path.exists(LIB):
        cmd = [NVCC, "-shared", "-o", LIB, *objs, "-gencode", "arch=compute_100a,code=sm_100a", "-lcudart"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            print(r.stdout + r.stderr, file=sys.stderr)
            raise RuntimeError("link failed")
    return LIB


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--ptxas-v", action="store_true")
    a = ap.parse_args()
    print(build(a.force, a.ptxas_v))
