#!/usr/bin/env python
"""Compile tests/cpp/test_dropin.cpp (the drop-in C++ header exercised like the reference's own C++ tests) into
tests/cpp/_bin/test_dropin.  Needs an Eigen include tree: the reference's vendored copy when /root/reference is
present (this container), or $EIGEN_INCLUDE.  The binary is git-ignored but travels to the GPU box."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
OUT = os.path.join(HERE, "_bin", "test_dropin")


def eigen_include():
    for p in (os.environ.get("EIGEN_INCLUDE"), "/root/reference/c++/vendor/eigen", "/usr/include/eigen3"):
        if p and os.path.isdir(os.path.join(p, "Eigen")):
            return p
    return None


def build(force=False):
    inc = eigen_include()
    if inc is None:
        return OUT if os.path.exists(OUT) else None
    src = os.path.join(HERE, "test_dropin.cpp")
    deps = [src, os.path.join(ROOT, "include", "celerite2_b200", "core.hpp"), os.path.join(ROOT, "include", "celerite2_b200.h")]
    if not force and os.path.exists(OUT) and all(os.path.getmtime(d) <= os.path.getmtime(OUT) for d in deps):
        return OUT
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    cmd = ["g++", "-std=c++17", "-O1", "-I", os.path.join(ROOT, "include"), "-I", inc, src, "-L", os.path.join(ROOT, "celerite2_b200"),
           "-lcelerite2_b200", "-Wl,-rpath,$ORIGIN/../../../celerite2_b200", "-o", OUT]
    print("[cpp] " + " ".join(cmd), flush=True)
    subprocess.check_call(cmd)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
