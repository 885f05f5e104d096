#!/usr/bin/env python
"""Build libpicd_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

    python build.py [--force] [--verbose]

Output: pytorch-imageclassification-distributed_b200/lib/libpicd_b200.so (git-ignored, travels with gpurun).
Objects are rebuilt only when a source or header is newer.
"""
from __future__ import annotations

import argparse
import concurrent.futures as cf
import glob
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "pytorch-imageclassification-distributed_b200")
OUT = os.path.join(PKG, "lib", "libpicd_b200.so")
OBJ_DIR = os.path.join(ROOT, "build", "obj")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
FLAGS = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC"]


def sources():
    return sorted(glob.glob(os.path.join(ROOT, "csrc", "**", "*.cu"), recursive=True) +
                  glob.glob(os.path.join(ROOT, "csrc", "**", "*.cpp"), recursive=True))


def headers():
    return glob.glob(os.path.join(ROOT, "csrc", "**", "*.cuh"), recursive=True) + \
        glob.glob(os.path.join(ROOT, "include", "*.h"))


def _compile(src: str, obj: str, verbose: bool) -> None:
    cmd = [NVCC, *ARCH, *FLAGS, "-I", os.path.join(ROOT, "include"), "-c", src, "-o", obj]
    if src.endswith(".cpp"):
        cmd = [NVCC, "-O3", "-std=c++17", "-Xcompiler", "-fPIC", "-x", "c++", "-I", os.path.join(ROOT, "include"), "-c", src, "-o", obj]
    if verbose:
        print(" ".join(cmd), flush=True)
    subprocess.run(cmd, check=True)


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OBJ_DIR, exist_ok=True)
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    hdr_time = max([os.path.getmtime(h) for h in headers()] + [os.path.getmtime(__file__)])
    jobs, objs = [], []
    for src in sources():
        obj = os.path.join(OBJ_DIR, os.path.relpath(src, ROOT).replace(os.sep, "_") + ".o")
        objs.append(obj)
        stale = force or not os.path.exists(obj) or os.path.getmtime(obj) < max(os.path.getmtime(src), hdr_time)
        if stale:
            jobs.append((src, obj))
    with cf.ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        for f in [ex.submit(_compile, s, o, verbose) for s, o in jobs]:
            f.result()
    if jobs or not os.path.exists(OUT):
        cmd = [NVCC, *ARCH, "-shared", "-o", OUT, *objs, "-cudart", "static"]
        if verbose:
            print(" ".join(cmd), flush=True)
        subprocess.run(cmd, check=True)
    return OUT


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    a = ap.parse_args()
    print(build(a.force, a.verbose))
