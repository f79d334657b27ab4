#!/usr/bin/env python
"""Generate the golden vectors under tests/golden/ from the REFERENCE ITSELF.

Runs only where oracle/_ref/libc2ref.so exists, i.e. in the build container after
`python oracle/build_oracle.py` compiled the reference's C++ core from /root/reference (strict flags:
-O2 -ffp-contract=off, generic x86-64).  The inputs are the reference's own test fixtures
(c++/test/helpers.hpp:14-62 and python/celerite2/testing.py:10-52) plus wider kernels (J = 6, 8, 9, 16).

Each golden_<kernel>.npz holds the inputs and every output of every op on the path, so the tests can
pin (a) the plain-C oracle on CPU and (b) the CUDA path on the GPU box, where /root/reference is absent.
"""
import os
import sys
import zlib

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import oracle  # noqa: E402
from helpers import ALL_KERNELS, cpp_data, py_matrices  # noqa: E402


def run_all(R, co, x, diag, Y, tag):
    out = {"x": x, "diag": diag, "Y": Y}
    for i, nm in enumerate(("ar", "cr", "ac", "bc", "cc", "dc")):
        out["co_" + nm] = np.asarray(co[i], dtype=np.float64)
    c, a, U, V = R.get_celerite_matrices(*co, x, diag)  # blocked layout, terms.hpp:112-139
    out.update(c=c, a=a, U=U, V=V)
    N, J = U.shape
    flag, d, W, S = R.factor(x, c, a, U, V)
    assert flag == 0, tag
    out.update(d=d, W=W, S=S)
    rng = np.random.default_rng(zlib.crc32(tag.encode()))
    bd, bW = rng.standard_normal(N), rng.standard_normal((N, J))
    out.update(bd=bd, bW=bW)
    for nm, v in zip(("bt", "bc", "ba", "bU", "bV"), R.factor_rev(x, c, a, U, V, d, W, S, bd, bW)):
        out["factor_rev_" + nm] = v
    bZ = rng.standard_normal(Y.shape)
    Z0 = rng.standard_normal(Y.shape)
    out.update(bZ=bZ, Z0=Z0)
    for op in oracle.OPS:
        z0 = None if op.startswith("solve") else Z0
        Z, F = R.sweep(op, x, c, U, W, Y, Z0=z0)
        Zi, _ = R.sweep(op, x, c, U, W, Y, inplace=True, workspace=False)
        out[op + "_Z"], out[op + "_F"], out[op + "_Zinplace"] = Z, F, Zi
        for nm, v in zip(("bt", "bc", "bU", "bW", "bY"), R.sweep_rev(op, x, c, U, W, Y, Z, F, bZ)):
            out[op + "_rev_" + nm] = v
    t1 = np.sort(rng.uniform(-1, 12, 77))
    P = oracle.port()
    _, _, U2, V2 = R.get_celerite_matrices(*co, t1, np.zeros_like(t1))
    out.update(t1=t1, U2=U2, V2=V2)
    for up, nm in ((0, "general_matmul_lower"), (1, "general_matmul_upper")):
        Z, F = R.general_matmul(up, t1, x, c, U2 if up == 0 else V2, V if up == 0 else U, Y)
        out[nm + "_Z"], out[nm + "_F"] = Z, F
    out["dense"] = R.to_dense(x, c, a, U, V)
    return out


def main():
    R = oracle.ref()
    if R is None:
        raise SystemExit("oracle/_ref/libc2ref.so missing: run python oracle/build_oracle.py where /root/reference exists")
    x, diag, Y = cpp_data(50, 5)
    for name, co in ALL_KERNELS.items():
        out = run_all(R, co, x, diag, Y, name)
        np.savez_compressed(os.path.join(HERE, "golden_%s.npz" % name), **out)
        print(name, "J=%d" % out["U"].shape[1], "ok")
    # the Python-side fixture (seed 721), vector and matrix right-hand sides, N=100
    px, pc, pa, pU, pV, pY = py_matrices()
    from celerite2_b200 import workloads as wl

    co = wl.sho_coeffs(5.0, 0.1, 3.45)
    out = run_all(R, co, px, pa - (co[2].sum() + co[0].sum()), pY, "py721")
    np.savez_compressed(os.path.join(HERE, "golden_py721.npz"), **out)
    print("py721 ok")


if __name__ == "__main__":
    main()
