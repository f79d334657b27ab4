#!/usr/bin/env python
"""Build libcelerite2_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

    python -m celerite2_b200.build [--force] [--verbose-ptxas]

One object per translation unit, compiled in parallel, linked into celerite2_b200/libcelerite2_b200.so.
The .so is git-ignored but travels to the GPU box with the gpurun snapshot.
"""
import os
import subprocess
import sys
import time
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "_build")
LIB = os.path.join(HERE, "libcelerite2_b200.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")

# (source, width or None).  The width lists must match C2B_DISPATCH_J_* in csrc/host_common.cuh.
WIDTHS_SWEEP = list(range(0, 17))
WIDTHS_FACTOR = list(range(0, 17))
UNITS = (
    [("api_misc.cu", None), ("api_sweep.cu", None), ("api_factor.cu", None), ("api_noalloc.cu", None), ("api_predict.cu", None), ("api_terms.cu", None)]
    + [("inst_sweep.cu", w) for w in WIDTHS_SWEEP]
    + [("inst_factor.cu", w) for w in WIDTHS_FACTOR]
)
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
FLAGS = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr"]


_COMMON = ["common.cuh", "monoids.cuh", "coop.cuh", "scan.cuh", "fold.cuh", "args.cuh", "host_common.cuh"]
_EXTRA = {
    "api_misc.cu": ["misc.cuh"],
    "inst_sweep.cu": ["sweep.cuh", "group.cuh", "gsweep.cuh", "csweep_rev.cuh"],
    "inst_factor.cu": ["factor.cuh", "group.cuh", "gfactor.cuh", "gfactor_rev.cuh"],
}


def _deps(unit):
    files = [unit] + _COMMON + _EXTRA.get(unit, [])
    return [os.path.join(CSRC, f) for f in files] + [os.path.join(HERE, "..", "include", "celerite2_b200.h")]


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def build(force=False, verbose_ptxas=False):
    os.makedirs(OBJ, exist_ok=True)
    # extra compile flags (tuning macros such as -DC2B_MINB_SWEEP=10) apply to EVERY translation unit; a change of flags
    # rebuilds everything
    extra = os.environ.get("C2B_NVCC_FLAGS", "").split()
    stamp = os.path.join(OBJ, "flags.txt")
    old = open(stamp).read() if os.path.exists(stamp) else ""
    if old != " ".join(extra):
        force = True
    with open(stamp, "w") as f:
        f.write(" ".join(extra))
    units = list(UNITS)
    missing = [u for (u, _) in units if not os.path.exists(os.path.join(CSRC, u))]
    if missing:
        raise RuntimeError("celerite2_b200.build: listed sources are missing: %s" % ", ".join(sorted(set(missing))))

    def objname(u, w):
        return os.path.join(OBJ, u.replace(".cu", "") + ("" if w is None else "_j%d" % w) + ".o")

    jobs = []
    for u, w in units:
        obj = objname(u, w)
        if force or _stale(obj, _deps(u)):
            cmd = [NVCC] + ARCH + FLAGS + extra + (["-Xptxas", "-v"] if verbose_ptxas else [])
            if w is not None:
                cmd += ["-DC2B_JT=%d" % w]
            cmd += ["-c", os.path.join(CSRC, u), "-o", obj]
            jobs.append((os.path.basename(obj), cmd))

    def run(job):
        u, cmd = job
        t0 = time.time()
        r = subprocess.run(cmd, capture_output=True, text=True)
        print("[build] %-22s %6.1f s" % (u, time.time() - t0), flush=True)
        return u, r

    failed = False
    with ThreadPoolExecutor(max_workers=max(1, min(len(jobs), os.cpu_count() or 1))) as ex:
        for u, r in ex.map(run, jobs):
            if verbose_ptxas or r.returncode != 0:
                sys.stderr.write(r.stdout + r.stderr)
            if r.returncode != 0:
                failed = True
    if failed:
        raise RuntimeError("nvcc failed")
    objs = [objname(u, w) for (u, w) in units]
    if force or jobs or _stale(LIB, objs):
        cmd = [NVCC] + ARCH + ["-shared", "-o", LIB] + objs
        print("[build] " + " ".join(cmd), flush=True)
        subprocess.check_call(cmd)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose_ptxas="--verbose-ptxas" in sys.argv))
