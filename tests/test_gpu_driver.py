"""GPU: the host mirrors of `celerite2.driver` / `celerite2.backprop`, written the way the reference's own tests are
(python/test/test_driver.py, python/test/test_backprop.py): dense numpy identities on the seed-721 fixture,
in-place behaviour, error behaviour."""
import numpy as np
import pytest

from helpers import kernel_value, py_matrices

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mods():
    from celerite2_b200 import backprop, driver, ops

    ops.set_chunks_per_series(0)
    return driver, backprop


def test_factor_and_solves_match_dense(mods):
    """test_driver.py:19-63"""
    driver, _ = mods
    x, c, a, U, V, Y = py_matrices()
    from celerite2_b200 import workloads as wl

    co = wl.sho_coeffs(5.0, 0.1, 3.45)
    K = kernel_value(co, x[:, None] - x[None, :])
    K[np.diag_indices_from(K)] = a
    d, W = driver.factor(x, c, a.copy(), U, V, a.copy(), V.copy())
    Lc = np.linalg.cholesky(K)
    np.testing.assert_allclose(np.diag(Lc) ** 2, d, rtol=1e-9)
    L = Lc / np.sqrt(d)[None, :]
    np.testing.assert_allclose(driver.solve_lower(x, c, U, W, Y, np.zeros_like(Y)), np.linalg.solve(L, Y), rtol=1e-7, atol=1e-9)
    np.testing.assert_allclose(driver.solve_upper(x, c, U, W, Y, np.zeros_like(Y)), np.linalg.solve(L.T, Y), rtol=1e-7, atol=1e-9)
    np.testing.assert_allclose(driver.matmul_lower(x, c, U, V, Y, np.zeros_like(Y)), np.tril(K, -1) @ Y, rtol=1e-7, atol=1e-9)
    np.testing.assert_allclose(driver.matmul_upper(x, c, U, V, Y, np.zeros_like(Y)), np.triu(K, 1) @ Y, rtol=1e-7, atol=1e-9)


def test_inplace_and_return_identity(mods):
    """test_driver.py:9-16,33-40: outputs are the caller's arrays"""
    driver, _ = mods
    x, c, a, U, V, Y = py_matrices()
    d, W = a.copy(), V.copy()
    d2, W2 = driver.factor(x, c, d, U, W, d, W)
    assert d2 is d and W2 is W
    Z = Y.copy()
    Z2 = driver.solve_lower(x, c, U, W, Z, Z)
    assert Z2 is Z
    np.testing.assert_allclose(Z, driver.solve_lower(x, c, U, W, Y, np.zeros_like(Y)), rtol=1e-12)


def test_general_matmul_kstar(mods):
    """test_driver.py:100-135"""
    driver, _ = mods
    from celerite2_b200 import workloads as wl

    co = wl.sho_coeffs(5.0, 0.1, 3.45)
    x, c, a, U, V, Y, t, U2, V2 = py_matrices(conditional=True)
    Kstar = kernel_value(co, t[:, None] - x[None, :])
    Z = np.zeros((len(t), Y.shape[1]))
    Z = driver.general_matmul_lower(t, x, c, U2, V, Y, Z)
    Z = driver.general_matmul_upper(t, x, c, V2, U, Y, Z)
    np.testing.assert_allclose(Z, Kstar @ Y, rtol=1e-7, atol=1e-9)


def test_get_celerite_matrices_interleaved(mods):
    driver, _ = mods
    import oracle
    from celerite2_b200 import workloads as wl

    co = wl.sum_coeffs(wl.real_coeffs(1.0, 0.1), wl.sho_coeffs(5.0, 0.1, 3.45), wl.sho_coeffs(0.1, 1.3, 5.3))
    x = np.sort(np.random.default_rng(0).uniform(0, 10, 64))
    diag = np.full(64, 0.3)
    J = wl.width(co)
    a, U, V = driver.get_celerite_matrices(co[0], co[2], co[3], co[5], x, diag, np.empty(64), np.empty((64, J)), np.empty((64, J)))
    _, a_r, U_r, V_r = oracle.port().get_celerite_matrices(*co, x, diag, layout="interleaved")
    np.testing.assert_allclose(a, a_r, rtol=1e-13)
    np.testing.assert_allclose(U, U_r, rtol=1e-12, atol=1e-14)
    np.testing.assert_allclose(V, V_r, rtol=1e-12, atol=1e-14)


def test_errors(mods):
    """shape errors -> ValueError (driver.cpp:32-46); non-positive-definite -> LinAlgError (driver.cpp:62)"""
    driver, backprop = mods
    x, c, a, U, V, Y = py_matrices()
    with pytest.raises(ValueError):
        driver.factor(x, c, a[:-1], U, V, a.copy(), V.copy())
    with pytest.raises(ValueError):
        driver.solve_lower(x, c, U, V, Y[:, 0], Y[:, 0])
    bad = a.copy()
    bad[10] = -100.0
    with pytest.raises(driver.LinAlgError):
        driver.factor(x, c, bad, U, V, bad.copy(), V.copy())
    with pytest.raises(backprop.LinAlgError):
        backprop.factor_fwd(x, c, bad, U, V, np.empty_like(a), np.empty_like(V), np.empty((len(x), 2, 2)))


def test_backprop_fwd_equals_driver_and_rev_matches_fd(mods):
    """test_backprop.py:9-42: *_fwd equals the driver op; *_rev vs forward differences (eps 1.234e-8, atol 1e-3)"""
    driver, backprop = mods
    x, c, a, U, V, Y = py_matrices(size=30)
    N, J = U.shape
    K = Y.shape[1]
    d, W, S = backprop.factor_fwd(x, c, a, U, V, np.empty(N), np.empty((N, J)), np.empty((N, J, J)))
    d0, W0 = driver.factor(x, c, a.copy(), U, V, a.copy(), V.copy())
    np.testing.assert_allclose(d, d0, rtol=1e-13)
    np.testing.assert_allclose(W, W0, rtol=1e-12, atol=1e-14)
    Z, F = backprop.solve_lower_fwd(x, c, U, W, Y, np.empty_like(Y), np.empty((N, J, K)))
    np.testing.assert_allclose(Z, driver.solve_lower(x, c, U, W, Y, np.zeros_like(Y)), rtol=1e-12, atol=1e-14)
    rng = np.random.default_rng(1)
    bZ = rng.standard_normal(Y.shape)
    bt, bc, bU, bW, bY = backprop.solve_lower_rev(x, c, U, W, Y, Z, F, bZ, np.empty(N), np.empty(J), np.empty((N, J)), np.empty((N, J)),
                                                  np.empty_like(Y))
    eps = 1.234e-8

    def f(x_, c_, U_, W_, Y_):
        return np.sum(driver.solve_lower(x_, c_, U_, W_, Y_, np.zeros_like(Y_)) * bZ)

    base = f(x, c, U, W, Y)
    for arr, grad, idx in ((c, bc, (1,)), (U, bU, (7, 1)), (W, bW, (3, 0)), (Y, bY, (5, 2)), (x, bt, (11,))):
        pert = arr.copy()
        pert[idx] += eps
        args = [pert if v is arr else v for v in (x, c, U, W, Y)]
        np.testing.assert_allclose((f(*args) - base) / eps, grad[idx], atol=1e-3, rtol=1e-4)
