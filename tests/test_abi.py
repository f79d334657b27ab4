"""CPU: the C-ABI library loads and exports every symbol include/celerite2_b200.h declares (no compute calls),
argument validation works without a GPU, and the product package never touches the oracle."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "celerite2_b200.h")
LIB = os.path.join(ROOT, "celerite2_b200", "libcelerite2_b200.so")


def _declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(c2b_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    if not os.path.exists(LIB):
        pytest.fail("libcelerite2_b200.so is missing: run `python -m celerite2_b200.build`")
    lib = ctypes.CDLL(LIB)
    names = _declared()
    assert len(names) >= 25
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing


def test_python_binding_covers_the_header():
    from celerite2_b200 import _lib

    assert sorted(_lib.EXPORTS) == _declared()
    assert "sm_100a" in _lib.version()


def test_argument_validation_needs_no_gpu():
    from celerite2_b200 import _lib

    lib = _lib.lib
    rc = lib.c2b_factor(0, 2, 1, None, None, None, None, None, None, None, None, None, 0, None, 0, None)
    assert rc == -1 and "N must be" in lib.c2b_last_error().decode()
    rc = lib.c2b_factor(10, 33, 1, None, None, None, None, None, None, None, None, None, 0, None, 0, None)
    assert rc == -1 and "J must be" in lib.c2b_last_error().decode()
    rc = lib.c2b_solve_lower(10, 2, 0, 1, None, None, None, None, None, None, None, 0, None, 0, None)
    assert rc == -1
    # scratch sizes are pure host arithmetic
    assert lib.c2b_scratch_bytes(_lib.OP_IDS["factor"], 1_000_000, 0, 6, 1, 1) > 0
    assert lib.c2b_scratch_bytes(_lib.OP_IDS["factor"], 1000, 0, 6, 1, 65536) == 0  # a full batch runs un-chunked


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "celerite2_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in text and "from oracle" not in text and "liboracle" not in text, f
