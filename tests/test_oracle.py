"""CPU tests that PIN the oracle (oracle/c2_oracle.c):

  1. against the committed golden vectors generated from the reference's own C++ core
     (tests/golden/make_golden.py), every op, J = 1..9 and 16;
  2. against the compiled reference itself when oracle/_ref is present (build container);
  3. against the dense-algebra identities the reference's tests use
     (c++/test/test_factor.cpp:10-78, test_solve_lower.cpp:11-43, test_matmul_lower.cpp:10-32,
      test_general_matmul_lower.cpp:10-32, python/test/test_driver.py);
  4. reverse ops against forward differences exactly as c++/test/helpers.hpp:230-244 does
     (eps = 1.234e-8, tol = 500 eps).
"""
import glob
import os

import numpy as np
import pytest

import oracle
from helpers import ALL_KERNELS, CPP_KERNELS, GOLDEN, cpp_data, kernel_value, py_matrices, relerr

TOL = 1e-12  # port vs reference: same algorithm, same operation order, libm differences only

GOLDEN_FILES = sorted(glob.glob(os.path.join(GOLDEN, "golden_*.npz")))


def _co(g):
    return tuple(g["co_" + k] for k in ("ar", "cr", "ac", "bc", "cc", "dc"))


@pytest.mark.parametrize("path", GOLDEN_FILES, ids=[os.path.basename(p)[7:-4] for p in GOLDEN_FILES])
def test_port_matches_golden(path):
    g = np.load(path)
    P = oracle.port()
    x, diag, Y = g["x"], g["diag"], g["Y"]
    c, a, U, V = P.get_celerite_matrices(*_co(g), x, diag)
    for nm, v in zip("caUV", (c, a, U, V)):
        assert relerr(v, g[nm]) < TOL, nm
    c, a, U, V = g["c"], g["a"], g["U"], g["V"]
    flag, d, W, S = P.factor(x, c, a, U, V)
    assert flag == 0
    assert relerr(d, g["d"]) < TOL and relerr(W, g["W"]) < TOL and relerr(S, g["S"]) < TOL
    d, W, S = g["d"], g["W"], g["S"]
    for nm, v in zip(("bt", "bc", "ba", "bU", "bV"), P.factor_rev(x, c, a, U, V, d, W, S, g["bd"], g["bW"])):
        assert relerr(v, g["factor_rev_" + nm]) < TOL, nm
    for op in oracle.OPS:
        z0 = None if op.startswith("solve") else g["Z0"]
        Z, F = P.sweep(op, x, c, U, W, Y, Z0=z0)
        Zi, _ = P.sweep(op, x, c, U, W, Y, inplace=True, workspace=False)
        assert relerr(Z, g[op + "_Z"]) < TOL, op
        assert relerr(F, g[op + "_F"]) < TOL, op
        assert relerr(Zi, g[op + "_Zinplace"]) < TOL, op
        outs = P.sweep_rev(op, x, c, U, W, Y, g[op + "_Z"], g[op + "_F"], g["bZ"])
        for nm, v in zip(("bt", "bc", "bU", "bW", "bY"), outs):
            assert relerr(v, g[op + "_rev_" + nm]) < TOL, (op, nm)
    for up, nm in ((0, "general_matmul_lower"), (1, "general_matmul_upper")):
        Z, F = P.general_matmul(up, g["t1"], x, c, g["U2"] if up == 0 else g["V2"], V if up == 0 else U, Y)
        assert relerr(Z, g[nm + "_Z"]) < TOL
        assert relerr(np.nan_to_num(F, nan=-7.0), np.nan_to_num(g[nm + "_F"], nan=-7.0)) < TOL
    assert relerr(P.to_dense(x, c, a, U, V), g["dense"]) < TOL


@pytest.mark.skipif(oracle.ref() is None, reason="oracle/_ref not built (needs /root/reference)")
@pytest.mark.parametrize("name", list(ALL_KERNELS))
def test_port_matches_compiled_reference(name):
    P, R = oracle.port(), oracle.ref()
    x, diag, Y = cpp_data(64, 3)
    co = ALL_KERNELS[name]
    c, a, U, V = R.get_celerite_matrices(*co, x, diag)
    fp, fr = P.factor(x, c, a, U, V), R.factor(x, c, a, U, V)
    assert fp[0] == fr[0] == 0
    for u, v in zip(fp[1:], fr[1:]):
        assert relerr(u, v) < TOL
    _, d, W, S = fr
    bd, bW = np.cos(x), np.sin(U * 3)
    for u, v in zip(P.factor_rev(x, c, a, U, V, d, W, S, bd, bW), R.factor_rev(x, c, a, U, V, d, W, S, bd, bW)):
        assert relerr(u, v) < TOL
    for op in oracle.OPS:
        zp, zr = P.sweep(op, x, c, U, W, Y), R.sweep(op, x, c, U, W, Y)
        assert relerr(zp[0], zr[0]) < TOL and relerr(zp[1], zr[1]) < TOL
        bZ = np.cos(3 * Y)
        for u, v in zip(P.sweep_rev(op, x, c, U, W, Y, zr[0], zr[1], bZ), R.sweep_rev(op, x, c, U, W, Y, zr[0], zr[1], bZ)):
            assert relerr(u, v) < TOL


def test_factor_failure_flag_matches_reference_semantics():
    """forward.hpp:128: first n >= 1 with d[n] <= 0 is returned and later rows are untouched."""
    P = oracle.port()
    x, diag, _ = cpp_data(20, 1)
    c, a, U, V = P.get_celerite_matrices(*CPP_KERNELS["sum4"], x, diag)
    a = a.copy()
    a[7] = -5.0
    flag, d, W, _ = P.factor(x, c, a, U, V)
    assert flag == 7
    assert np.all(d[8:] == a[8:]) and np.all(W[8:] == V[8:])  # untouched (copies of the inputs here)
    if oracle.ref() is not None:
        assert oracle.ref().factor(x, c, a, U, V)[0] == 7


# ---- dense identities (the reference's own bar: c++ tests use |resid| < 1e-12 at N=50) -------------

@pytest.mark.parametrize("name", list(CPP_KERNELS))
def test_dense_identities(name):
    P = oracle.port()
    co = CPP_KERNELS[name]
    x, diag, Y = cpp_data(50, 5)
    c, a, U, V = P.get_celerite_matrices(*co, x, diag)
    K = kernel_value(co, x[:, None] - x[None, :])
    K[np.diag_indices_from(K)] += diag
    assert np.max(np.abs(P.to_dense(x, c, a, U, V) - K)) < 1e-10  # test_to_dense.cpp uses 1e-8
    flag, d, W, S = P.factor(x, c, a, U, V)
    assert flag == 0
    # LDL^T of K: L = chol(K) / sqrt(d)
    Lc = np.linalg.cholesky(K)
    assert np.max(np.abs(np.diag(Lc) ** 2 - d)) < 1e-12
    L = Lc / np.sqrt(d)[None, :]
    Z, _ = P.sweep("solve_lower", x, c, U, W, Y)
    assert np.max(np.abs(Z - np.linalg.solve(L, Y))) < 1e-11
    Z, _ = P.sweep("solve_upper", x, c, U, W, Y)
    assert np.max(np.abs(Z - np.linalg.solve(L.T, Y))) < 1e-11
    Z, _ = P.sweep("matmul_lower", x, c, U, V, Y)
    assert np.max(np.abs(Z - np.tril(K, -1) @ Y)) < 1e-11
    Z, _ = P.sweep("matmul_upper", x, c, U, V, Y)
    assert np.max(np.abs(Z - np.triu(K, 1) @ Y)) < 1e-11
    # in-place matmul_lower with W gives L @ Y (python/celerite2/numpy.py:100-102)
    Z, _ = P.sweep("matmul_lower", x, c, U, W, Y, inplace=True, workspace=False)
    assert np.max(np.abs(Z - L @ Y)) < 1e-11


def test_general_matmul_matches_dense_kstar():
    """python/test/test_driver.py:100-135"""
    P = oracle.port()
    from celerite2_b200 import workloads as wl

    co = wl.sho_coeffs(5.0, 0.1, 3.45)
    x, c, a, U, V, Y, t, U2, V2 = py_matrices(conditional=True)
    Kstar = kernel_value(co, t[:, None] - x[None, :])
    Z, _ = P.general_matmul(0, t, x, c, U2, V, Y)
    Z, _ = P.general_matmul(1, t, x, c, V2, U, Y, Z0=Z)
    assert np.max(np.abs(Z - Kstar @ Y)) < 1e-10


# ---- reverse ops vs forward differences (c++/test/helpers.hpp:230-244) ------------------------------

EPS = 1.234e-8
FD_TOL = 500 * EPS


def _fd_check(fwd, rev, args, n_out):
    """fwd(*args) -> tuple of outputs; rev(args, outputs_cotangents) -> tuple of input cotangents."""
    base = fwd(*args)
    rng = np.random.default_rng(0)
    cots = [rng.standard_normal(np.shape(b)) for b in base[:n_out]]
    grads = rev(args, cots)
    for ia, (arg, g) in enumerate(zip(args, grads)):
        flat = arg.reshape(-1)
        num = np.zeros_like(flat)
        for i in range(flat.size):
            old = flat[i]
            flat[i] = old + EPS
            pert = fwd(*args)
            flat[i] = old
            num[i] = sum(np.sum((p - b) * ct) for p, b, ct in zip(pert[:n_out], base[:n_out], cots)) / EPS
        scale = max(1.0, np.max(np.abs(num)))
        assert np.max(np.abs(num - g.reshape(-1))) < FD_TOL * scale * 10, ia


@pytest.mark.parametrize("name", ["real", "sho_under", "sum1", "sum4", "c2_J6"])
def test_factor_rev_fd(name):
    P = oracle.port()
    x, diag, _ = cpp_data(10, 1)
    c, a, U, V = P.get_celerite_matrices(*ALL_KERNELS[name], x, diag)

    def fwd(x, c, a, U, V):
        _, d, W, S = P.factor(x, c, a, U, V)
        return d, W, S

    def rev(args, cots):
        d, W, S = fwd(*args)
        return P.factor_rev(*args, d, W, S, cots[0], cots[1])

    _fd_check(fwd, rev, [x.copy(), c.copy(), a.copy(), U.copy(), V.copy()], 2)


@pytest.mark.parametrize("op", list(oracle.OPS))
@pytest.mark.parametrize("name", ["sho_under", "sum1", "c2_J6"])
def test_sweep_rev_fd(op, name):
    P = oracle.port()
    x, diag, Y = cpp_data(10, 2)
    c, a, U, V = P.get_celerite_matrices(*ALL_KERNELS[name], x, diag)
    _, _, W, _ = P.factor(x, c, a, U, V)
    B = W if op.startswith("solve") else V

    def fwd(x, c, U, B, Y):
        return P.sweep(op, x, c, U, B, Y)

    def rev(args, cots):
        Z, F = fwd(*args)
        return P.sweep_rev(op, *args, Z, F, cots[0])

    _fd_check(fwd, rev, [x.copy(), c.copy(), U.copy(), B.copy(), Y.copy()], 1)


# ---- round 2: the restated assembly adjoint and the preallocated pipeline -------------------------------------------
@pytest.mark.parametrize("name", ["sum3", "c2_J6", "mixed_J9"])
def test_assembly_adjoint_restatement_against_central_differences(name):
    """oracle.get_celerite_matrices_rev (numpy) is the derivative of the blocked assembly (terms.hpp:112-139)."""
    import oracle
    from helpers import ALL_KERNELS

    co = [np.array(v, dtype=np.float64) for v in ALL_KERNELS[name]]
    rng = np.random.default_rng(3)
    N = 40
    x = np.sort(rng.uniform(0, 10, N))
    diag = rng.uniform(0.1, 0.3, N)
    P = oracle.port()
    c, a, U, V = P.get_celerite_matrices(*co, x, diag)
    b_c, b_a, b_U, b_V = (rng.standard_normal(v.shape) for v in (c, a, U, V))

    def f(co_, x_, diag_):
        c_, a_, U_, V_ = P.get_celerite_matrices(*co_, x_, diag_)
        return np.sum(c_ * b_c) + np.sum(a_ * b_a) + np.sum(U_ * b_U) + np.sum(V_ * b_V)

    g = oracle.get_celerite_matrices_rev(*co, x, b_c, b_a, b_U, b_V)
    eps = 1e-6
    for i in range(6):
        for j in range(len(co[i])):
            cp, cm = [v.copy() for v in co], [v.copy() for v in co]
            cp[i][j] += eps
            cm[i][j] -= eps
            fd = (f(cp, x, diag) - f(cm, x, diag)) / (2 * eps)
            assert abs(fd - g[i][j]) <= 1e-7 * max(1.0, abs(fd)), (i, j)
    for n in (0, 7, N - 1):
        xp, xm = x.copy(), x.copy()
        xp[n] += eps
        xm[n] -= eps
        fd = (f(co, xp, diag) - f(co, xm, diag)) / (2 * eps)
        assert abs(fd - g[6][n]) <= 1e-7 * max(1.0, abs(fd))
        dp, dm = diag.copy(), diag.copy()
        dp[n] += eps
        dm[n] -= eps
        fd = (f(co, x, dp) - f(co, x, dm)) / (2 * eps)
        assert abs(fd - g[7][n]) <= 1e-7 * max(1.0, abs(fd))


def test_ref_pipeline_equals_op_by_op_calls():
    """oracle.RefPipeline (preallocated buffers, raw ctypes calls) returns what the allocating wrappers return."""
    import oracle
    from celerite2_b200 import workloads as wl

    cfg = wl.config_c2(N=3000)
    t, y, diag = wl.light_curve(3000, 9)
    for lib in (oracle.port(), oracle.ref(), oracle.ref_fast()):
        if lib is None:
            continue
        pipe = oracle.RefPipeline(lib, 3000, 0, 3)
        for _ in range(2):  # the second call reuses every buffer
            ll, bc, flag = pipe.value_and_grad(cfg["coeffs"], t, y, diag)
        c, a, U, V = lib.get_celerite_matrices(*cfg["coeffs"], t, diag)
        flag2, d, W, S = lib.factor(t, c, a, U, V)
        Z, F = lib.sweep("solve_lower", t, c, U, W, y[:, None])
        z = Z[:, 0]
        ll2 = -0.5 * (np.sum(z * z / d) + np.sum(np.log(d)) + len(t) * np.log(2 * np.pi))
        bt, bc1, bU, bW, bY = lib.sweep_rev("solve_lower", t, c, U, W, y[:, None], Z, F, (-z / d)[:, None])
        bt2, bc2, ba, bU2, bV = lib.factor_rev(t, c, a, U, V, d, W, S, 0.5 * z * z / d**2 - 0.5 / d, bW)
        assert flag == flag2 == 0 and ll == ll2
        # (the seed bd is rounded differently by the two formulations: 0.5 (z/d)^2 vs 0.5 z^2 / d^2)
        assert np.allclose(bc, bc1 + bc2, rtol=1e-11, atol=0) and relerr(pipe.bV, bV) < 1e-12 and np.array_equal(pipe.S, S)
        llv, _ = oracle.RefPipeline(lib, 3000, 0, 3, grad=False).value(cfg["coeffs"], t, y, diag)
        assert abs(llv - ll) <= 1e-13 * abs(ll)
