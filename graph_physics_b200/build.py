"""Builds graph_physics_b200/libmgn_b200.so in-tree with nvcc for sm_100a.

    python -m graph_physics_b200.build [--force]

nvcc cross-compiles without a GPU.  One object per .cu (compiled in parallel),
then one shared library with a plain C ABI (include/mgn_b200.h).
"""
from __future__ import annotations

import concurrent.futures as cf
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
# MGN_BUILD_TAG=<tag>: a developer variant (with MGN_NVCC_EXTRA, e.g. -DMGN_TIMELINE) built next to the production
# library as libmgn_b200_<tag>.so from its own object directory; MGN_LIB_PATH selects it at run time
TAG = os.environ.get("MGN_BUILD_TAG", "")
OUT = os.path.join(HERE, f"libmgn_b200_{TAG}.so" if TAG else "libmgn_b200.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
         "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr", "-Xptxas", "-v"]
FLAGS += os.environ.get("MGN_NVCC_EXTRA", "").split()   # developer experiments, e.g. -DMGN_FINE_TL


def _sources():
    return sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    hdrs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    objdir = os.path.join(CSRC, f"build_{TAG}" if TAG else "build")
    os.makedirs(objdir, exist_ok=True)
    jobs = []
    for src in _sources():
        s = os.path.join(CSRC, src)
        o = os.path.join(objdir, src[:-3] + ".o")
        if force or _stale(o, [s] + hdrs):
            jobs.append((s, o))

    def cc(job):
        s, o = job
        r = subprocess.run([NVCC, *FLAGS, "-c", s, "-o", o], capture_output=True, text=True)
        return s, r

    with cf.ThreadPoolExecutor(max_workers=os.cpu_count() or 4) as ex:
        for s, r in ex.map(cc, jobs):
            if verbose or r.returncode != 0:
                sys.stderr.write(r.stdout + r.stderr)
            if r.returncode != 0:
                raise RuntimeError(f"nvcc failed on {s}")
            with open(os.path.join(objdir, os.path.basename(s)[:-3] + ".ptxas.log"), "w") as f:
                f.write(r.stderr)
    objs = [os.path.join(objdir, src[:-3] + ".o") for src in _sources()]
    if force or jobs or _stale(OUT, objs):
        r = subprocess.run([NVCC, "-shared", "-o", OUT, *objs, "-lcudart"], capture_output=True, text=True)
        if r.returncode != 0:
            sys.stderr.write(r.stdout + r.stderr)
            raise RuntimeError("link failed")
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
