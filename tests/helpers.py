"""Shared fixtures: the reference's own test inputs, restated.

  cpp_data / CPP_KERNELS   c++/test/helpers.hpp:14-62   (get_data, the eight TestKernels)
  py_matrices              python/celerite2/testing.py:10-52 (default_rng(721) fixture, SHOTerm J=2)
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from celerite2_b200 import workloads as wl  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")


def cpp_data(N=50, nrhs=5):
    """c++/test/helpers.hpp:14-24"""
    delta = np.arange(N) / (N - 1)
    x = 10 * delta + delta * delta
    diag = np.full(N, 0.5)
    Y = np.sin(x[:, None] + np.arange(nrhs)[None, :] / nrhs)
    return x, diag, np.ascontiguousarray(Y)


# c++/test/helpers.hpp:27-62.  NOTE the reference's C++ Term::operator+ has a defect (terms.hpp:160-162 copies
# ac_ into bc/cc/dc); these sums use the *intended* concatenation, as the Python terms do.
_real = wl.real_coeffs(1.0, 0.1)
_complex = wl.complex_coeffs(0.8, 0.03, 1.0, 0.1)
_sho1 = wl.sho_coeffs(1.2, 0.3, 0.1)
_sho2 = wl.sho_coeffs(0.1, 1.3, 5.3)
CPP_KERNELS = {
    "real": _real,  # J=1
    "complex": _complex,  # J=2
    "sho_over": _sho1,  # J=2
    "sho_under": _sho2,  # J=2
    "sum1": wl.sum_coeffs(_real, _complex),  # J=3
    "sum2": wl.sum_coeffs(_real, _complex, _sho1),  # J=5
    "sum3": wl.sum_coeffs(_real, _complex, _sho1, _sho2),  # J=7
    "sum4": wl.sum_coeffs(_sho1, _sho2),  # J=4
}
# extra widths the reference's tests do not reach (J=6 is the bench kernel, 8/16 exercise the wide paths)
EXTRA_KERNELS = {
    "c2_J6": wl.config_c2()["coeffs"],
    "sho4_J8": wl.sum_coeffs(*[wl.sho_coeffs(1.0 / (i + 1), 0.5 + 0.7 * i, 1.0 + 0.5 * i) for i in range(4)]),
    "c4_J16": wl.config_c4()["coeffs"],
    "mixed_J9": wl.sum_coeffs(_real, _sho1, _sho2, _complex, wl.matern32_coeffs(1.0, 2.0)),
}
ALL_KERNELS = dict(CPP_KERNELS, **EXTRA_KERNELS)


def assemble(coeffs, x, diag, layout="blocked"):
    """a/U/V via the oracle port (test-side only)."""
    import oracle

    return oracle.port().get_celerite_matrices(*coeffs, x, diag, layout=layout)


def py_matrices(size=100, coeffs=None, vector=False, conditional=False, no_diag=False):
    """python/celerite2/testing.py:10-52 (interleaved layout, as the numpy front-end builds it)."""
    random = np.random.default_rng(721)
    x = np.sort(random.uniform(0, 10, size))
    if vector:
        Y = np.sin(x)
    else:
        Y = np.ascontiguousarray(np.vstack([np.sin(x), np.cos(x), x**2]).T, dtype=np.float64)
    diag = np.zeros_like(x) if no_diag else random.uniform(0.1, 0.3, len(x))
    coeffs = coeffs if coeffs is not None else wl.sho_coeffs(5.0, 0.1, 3.45)
    c, a, U, V = assemble(coeffs, x, diag, layout="interleaved")
    if not conditional:
        return x, c, a, U, V, Y
    t = np.sort(random.uniform(-1, 12, 200))
    _, _, U2, V2 = assemble(coeffs, t, np.zeros_like(t), layout="interleaved")
    return x, c, a, U, V, Y, t, U2, V2


def kernel_value(coeffs, tau):
    """k(tau) of a celerite kernel, python/celerite2/terms.py:64-85"""
    ar, cr, ac, bc, cc, dc = coeffs
    tau = np.abs(np.asarray(tau, dtype=np.float64))
    k = np.zeros_like(tau)
    for i in range(len(ar)):
        k += ar[i] * np.exp(-cr[i] * tau)
    for i in range(len(ac)):
        arg = dc[i] * tau
        k += np.exp(-cc[i] * tau) * (ac[i] * np.cos(arg) + bc[i] * np.sin(arg))
    return k


def relerr(x, ref):
    x = np.asarray(x, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    scale = np.max(np.abs(ref)) if ref.size else 1.0
    if scale == 0.0:
        scale = 1.0
    return float(np.max(np.abs(x - ref)) / scale) if ref.size else 0.0
