"""GPU: the drop-in C++ header include/celerite2_b200/core.hpp, driven by a C++ program written like the reference's
own C++ tests (tests/cpp/test_dropin.cpp).  The binary is built where an Eigen tree exists (tests/cpp/build_cpp.py,
also run by __graft_entry__.build()) and travels to the GPU box."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.parametrize("name", ["test_dropin", "test_names"])
def test_cpp_dropin_program(name):
    """test_dropin: celerite2::core::* through <celerite2_b200/core.hpp>; test_names: the same through the headers that carry
    the reference's own file names (<celerite2/celerite2.h>), incl. terms.hpp, to_dense and the failure semantics of factor"""
    sys.path.insert(0, os.path.join(HERE, "cpp"))
    import build_cpp

    exe = build_cpp.build(name=name)
    if exe is None or not os.path.exists(exe):
        pytest.skip("no Eigen include tree and no prebuilt tests/cpp/_bin/%s" % name)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "OK" in r.stdout, r.stdout + r.stderr
