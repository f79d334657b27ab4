#!/usr/bin/env python
"""Build the checkers under oracle/ (TEST INFRASTRUCTURE ONLY, never on the product path).

  oracle/liboracle.so          plain-C restatement of the reference algorithm (oracle/c2_oracle.c)
  oracle/_ref/libc2ref.so      the reference's own C++ core, compiled from the headers where they lie
                               under /root/reference (strict build: no FMA contraction, stock J<=4 dispatch)
  oracle/_ref/libc2ref_fast.so same sources, -O3 -march=x86-64-v3, fixed-size dispatch up to J=8 and 16
                               (the reference's CELERITE2_FAST_RUN idea) -- used as the timed CPU baseline

The _ref libraries can only be (re)built where /root/reference exists (this container); the GPU box uses
the prebuilt files that travel with the snapshot.  Nothing from /root/reference is copied into the repo.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("C2B_REFERENCE_ROOT", "/root/reference")
REF_INC = [os.path.join(REF, "c++", "include"), os.path.join(REF, "c++", "vendor", "eigen")]


def _newer(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources if os.path.exists(s))


def _run(cmd):
    print("[oracle] " + " ".join(cmd), flush=True)
    subprocess.check_call(cmd)


def build_port(force=False):
    src = os.path.join(HERE, "c2_oracle.c")
    out = os.path.join(HERE, "liboracle.so")
    if force or _newer(out, [src]):
        _run(["gcc", "-O2", "-std=c11", "-fPIC", "-shared", "-ffp-contract=off", "-Wall", "-Wextra", "-o", out, src, "-lm"])
    return out


def have_reference():
    return all(os.path.isdir(p) for p in REF_INC)


def build_ref(force=False, fast=True):
    """Returns the list of built libraries, [] when /root/reference is absent and nothing is prebuilt."""
    src = os.path.join(HERE, "ref_shim.cpp")
    outdir = os.path.join(HERE, "_ref")
    os.makedirs(outdir, exist_ok=True)
    strict = os.path.join(outdir, "libc2ref.so")
    fastlib = os.path.join(outdir, "libc2ref_fast.so")
    if not have_reference():
        return [p for p in (strict, fastlib) if os.path.exists(p)]
    inc = []
    for p in REF_INC:
        inc += ["-I", p]
    common = ["g++", "-std=c++17", "-fPIC", "-shared", "-DNDEBUG", "-w"] + inc
    procs = []
    if force or _newer(strict, [src]):
        cmd = common + ["-O2", "-ffp-contract=off", "-o", strict, src]
        print("[oracle] " + " ".join(cmd), flush=True)
        procs.append(subprocess.Popen(cmd))
    if fast and (force or _newer(fastlib, [src])):
        cmd = common + ["-O3", "-march=x86-64-v3", "-DC2REF_FAST", "-o", fastlib, src]
        print("[oracle] " + " ".join(cmd), flush=True)
        procs.append(subprocess.Popen(cmd))
    for p in procs:
        if p.wait() != 0:
            raise RuntimeError("reference build failed")
    return [p for p in (strict, fastlib) if os.path.exists(p)]


if __name__ == "__main__":
    force = "--force" in sys.argv
    print(build_port(force))
    print(build_ref(force))
