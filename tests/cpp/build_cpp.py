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
PROGRAMS = ("test_dropin", "test_names")  # test_names includes <celerite2/celerite2.h>: the shim headers under the reference's names


def eigen_include():
    for p in (os.environ.get("EIGEN_INCLUDE"), "/root/reference/c++/vendor/eigen", "/usr/include/eigen3"):
        if p and os.path.isdir(os.path.join(p, "Eigen")):
            return p
    return None


def build(force=False, name="test_dropin"):
    out = os.path.join(HERE, "_bin", name)
    inc = eigen_include()
    if inc is None:
        return out if os.path.exists(out) else None
    src = os.path.join(HERE, name + ".cpp")
    deps = [src, os.path.join(ROOT, "include", "celerite2_b200", "core.hpp"), os.path.join(ROOT, "include", "celerite2_b200.h"),
            os.path.join(ROOT, "include", "celerite2", "terms.hpp")]
    if not force and os.path.exists(out) and all(os.path.getmtime(d) <= os.path.getmtime(out) for d in deps):
        return out
    os.makedirs(os.path.dirname(out), exist_ok=True)
    cmd = ["g++", "-std=c++17", "-O1", "-I", os.path.join(ROOT, "include"), "-I", inc, src, "-L", os.path.join(ROOT, "celerite2_b200"),
           "-lcelerite2_b200", "-Wl,-rpath,$ORIGIN/../../../celerite2_b200", "-o", out]
    print("[cpp] " + " ".join(cmd), flush=True)
    subprocess.check_call(cmd)
    return out


def build_all(force=False):
    return [build(force, name) for name in PROGRAMS]


if __name__ == "__main__":
    print(build_all(force="--force" in sys.argv))
