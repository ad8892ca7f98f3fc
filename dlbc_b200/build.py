"""Builds libdlbc_cuda.so in-tree with nvcc for sm_100a (no GPU needed: nvcc cross-compiles).

    python -m dlbc_b200.build [--force] [--jobs N]

Translation units: one per (component count, arithmetic mode) for the step kernels -- the
EXACT units are compiled with -fmad=false so that every fp64 operation is rounded separately,
exactly as the reference's CPU code does; the FAST units keep FMA contraction -- plus the
layout-conversion/output kernels and the C ABI itself.
"""
from __future__ import annotations

import argparse
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(CSRC, "build")
SO = os.path.join(HERE, "libdlbc_cuda.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC,-O2,-Wall", "-I", os.path.join(HERE, "..", "include")]


def units():
    u = [("dlbc_cuda.cu", []), ("field_io.cu", ["-fmad=false"]), ("no_alloc.cu", ["-fmad=false"]),
         ("d2q9.cu", ["-fmad=false"])]
    for n in (1, 2, 3, 4):
        u.append((f"step_c{n}_exact.cu", ["-fmad=false"]))
        u.append((f"step_c{n}_fast.cu", ["-fmad=true"]))
    return u


def _deps():
    return [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh", ".inl"))] + [
        os.path.join(HERE, "..", "include", "dlbc_cuda.h"), os.path.abspath(__file__)]


def _stale(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def _compile(src, flags, force, verbose, objdir=None):
    obj = os.path.join(objdir or OBJ, src.replace(".cu", ".o"))
    if not force and not _stale(obj, [os.path.join(CSRC, src)] + _deps()):
        return obj
    cmd = [NVCC] + ARCH + COMMON + flags + ["-Xptxas", "-v" if verbose else "-warn-spills", "-c",
                                            os.path.join(CSRC, src), "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if verbose:
        with open(obj.replace(".o", ".ptxas.txt"), "w") as f:
            f.write(r.stderr)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
    return obj


def build(force=False, jobs=None, verbose=False, out=None, defines=()):
    """out / defines: an A/B build (scripts/gpu_ab.sh) -- its own object directory, -D flags, another .so."""
    so = os.path.abspath(out) if out else SO
    objdir = OBJ if not out else os.path.join(CSRC, "build_" + os.path.splitext(os.path.basename(so))[0])
    os.makedirs(objdir, exist_ok=True)
    os.makedirs(os.path.dirname(so), exist_ok=True)
    jobs = jobs or min(10, os.cpu_count() or 4)
    us = units()
    dflags = ["-D" + d for d in defines]
    with ThreadPoolExecutor(jobs) as ex:
        objs = list(ex.map(lambda u: _compile(u[0], u[1] + dflags, force, verbose, objdir), us))
    if force or _stale(so, objs):
        cmd = [NVCC] + ARCH + ["-shared", "-o", so] + objs + ["-lcudart_static", "-ldl", "-lpthread", "-lrt"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return so


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--jobs", type=int, default=None)
    ap.add_argument("--verbose", action="store_true", help="keep ptxas -v output next to the objects")
    ap.add_argument("--out", default=None, help="A/B build: path of the .so (objects go to csrc/build_<name>)")
    ap.add_argument("--define", action="append", default=[], help="-D macro for an A/B build, e.g. DLBC_TMA_CWARPS=8")
    a = ap.parse_args()
    print(build(a.force, a.jobs, a.verbose, a.out, a.define))
