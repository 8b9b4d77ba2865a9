"""Build csrc/*.cu into dba_mod_b200/ops/libdba_b200.so for sm_100a (cross-compiles without a GPU).

    python tools/build_native.py [--force] [--verbose]

Objects are compiled in parallel with ``nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo`` and linked into one
shared library kept IN-TREE (git-ignored) so that it travels to the GPU box with the source snapshot.
"""
from __future__ import annotations

import argparse
import concurrent.futures as cf
import glob
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "csrc")
OUT = os.path.join(ROOT, "dba_mod_b200", "ops", "libdba_b200.so")
OBJ = os.path.join(ROOT, "build", "obj")
NVCC = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
# IEEE division / sqrt / expf / logf everywhere (no --use_fast_math): cross-entropy, BatchNorm 1/sqrt(var + eps) and the
# FoolsGold logit follow the reference's fp32 arithmetic; none of these is on a throughput-critical path.
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC",
         "-I", os.path.join(ROOT, "include"), "-I", SRC]


def _stale(target: str, deps) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OBJ, exist_ok=True)
    srcs = sorted(glob.glob(os.path.join(SRC, "*.cu")))
    hdrs = glob.glob(os.path.join(SRC, "*.cuh")) + glob.glob(os.path.join(ROOT, "include", "*.h"))
    jobs = []
    for s in srcs:
        o = os.path.join(OBJ, os.path.basename(s)[:-3] + ".o")
        if force or _stale(o, [s] + hdrs):
            jobs.append((s, o))

    def _cc(job):
        s, o = job
        cmd = [NVCC] + FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", s, "-o", o]
        r = subprocess.run(cmd, capture_output=True, text=True)
        return s, r.returncode, r.stdout + r.stderr

    with cf.ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        for s, rc, out in ex.map(_cc, jobs):
            if verbose or rc:
                print(f"--- {os.path.basename(s)}\n{out}")
            if rc:
                raise RuntimeError(f"nvcc failed on {s}")
    objs = [os.path.join(OBJ, os.path.basename(s)[:-3] + ".o") for s in srcs]
    if force or jobs or _stale(OUT, objs):
        cmd = [NVCC, "-shared", "-o", OUT] + objs + ["-lcudart"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode:
            print(r.stdout + r.stderr)
            raise RuntimeError("link failed")
    return OUT


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    a = ap.parse_args()
    print(build(a.force, a.verbose))
