#!/usr/bin/env python
"""TEST INFRASTRUCTURE ONLY: lay the reference's own Python front-end out twice under oracle/_ref/ (git-ignored,
travels to the GPU box), so that the SAME unmodified `celerite2` package -- core.py, numpy.py, terms.py, testing.py and the
reference's own tests -- runs once on top of this repository's GPU driver and once on top of the reference's CPU driver:

  oracle/_ref/pyref/celerite2/       core.py numpy.py terms.py testing.py citation.py __init__.py definitions.json copied
                                     from /root/reference/python/celerite2 at build time (nothing enters the repository),
                                     celerite2_version.py (stub), and driver.py / backprop.py written by
                                     celerite2_b200.install_into(): the names `celerite2.driver` / `celerite2.backprop`
                                     resolve to the CUDA path (what python/celerite2/driver.cpp:482-499 and
                                     backprop.cpp:906-919 register for the CPU).
  oracle/_ref/pyref/tests/           python/test/test_driver.py, test_backprop.py, test_celerite2.py (the reference's own)
  oracle/_ref/pyref_cpu/celerite2/   the same Python files plus the reference's pybind11 modules compiled from
                                     python/celerite2/{driver,backprop}.cpp where they lie (g++, pybind11 and Eigen headers;
                                     no cmake): the A/B partner of tests/test_gpu_reference_frontend.py.

Only runs where /root/reference exists (this container); elsewhere it reports what is already there.
"""
import os
import shutil
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = os.environ.get("C2B_REFERENCE_ROOT", "/root/reference")
PY_SRC = os.path.join(REF, "python", "celerite2")
PY_FILES = ["__init__.py", "core.py", "numpy.py", "terms.py", "testing.py", "citation.py", "definitions.json"]
TESTS = ["test_driver.py", "test_backprop.py", "test_celerite2.py"]
GPU_DIR = os.path.join(HERE, "_ref", "pyref")
CPU_DIR = os.path.join(HERE, "_ref", "pyref_cpu")


def have_reference():
    return os.path.isdir(PY_SRC)


def _copy_front_end(dst_pkg):
    os.makedirs(dst_pkg, exist_ok=True)
    for f in PY_FILES:
        shutil.copyfile(os.path.join(PY_SRC, f), os.path.join(dst_pkg, f))
    with open(os.path.join(dst_pkg, "celerite2_version.py"), "w") as f:
        f.write('__version__ = "0.0.0+reference.49d7b2416ae8"\n')


def _ext_suffix():
    return sysconfig.get_config_var("EXT_SUFFIX") or ".so"


def build(force=False):
    """-> dict(gpu=path or None, cpu=path or None)"""
    out = {"gpu": GPU_DIR if os.path.isdir(os.path.join(GPU_DIR, "celerite2")) else None,
           "cpu": CPU_DIR if os.path.exists(os.path.join(CPU_DIR, "celerite2", "driver" + _ext_suffix())) else None}
    if not have_reference():
        return out
    if ROOT not in sys.path:
        sys.path.insert(0, ROOT)
    # ---- the front-end over the GPU driver -------------------------------------------------------------------------
    pkg = os.path.join(GPU_DIR, "celerite2")
    _copy_front_end(pkg)
    import celerite2_b200

    celerite2_b200.install_into(pkg)
    tdir = os.path.join(GPU_DIR, "tests")
    os.makedirs(tdir, exist_ok=True)
    for f in TESTS:
        shutil.copyfile(os.path.join(REF, "python", "test", f), os.path.join(tdir, f))
    out["gpu"] = GPU_DIR
    # ---- the same front-end over the reference's own pybind11 modules ----------------------------------------------------
    cpkg = os.path.join(CPU_DIR, "celerite2")
    _copy_front_end(cpkg)
    try:
        import pybind11
    except ImportError:
        print("[pyref] pybind11 missing: the CPU partner is not built", flush=True)
        return out
    inc = ["-I", pybind11.get_include(), "-I", sysconfig.get_paths()["include"], "-I", os.path.join(REF, "c++", "include"),
           "-I", os.path.join(REF, "c++", "vendor", "eigen"), "-I", PY_SRC]
    procs = []
    for mod in ("driver", "backprop"):
        target = os.path.join(cpkg, mod + _ext_suffix())
        src = os.path.join(PY_SRC, mod + ".cpp")
        if force or not os.path.exists(target) or os.path.getmtime(src) > os.path.getmtime(target):
            cmd = ["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-fvisibility=hidden", "-DNDEBUG", "-w"] + inc + [src, "-o", target]
            print("[pyref] " + " ".join(cmd), flush=True)
            procs.append(subprocess.Popen(cmd))
    for p in procs:
        if p.wait() != 0:
            raise RuntimeError("building the reference's pybind11 modules failed")
    out["cpu"] = CPU_DIR
    return out


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
