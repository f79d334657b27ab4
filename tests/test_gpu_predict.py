"""GPU: the predictive distribution (celerite2_b200.predict) against the reference's own formulas evaluated densely in numpy
(python/celerite2/core.py:115-150: mean, variance, covariance of the conditional distribution; kernel function
terms.py:64-85), rel 1e-9 on quantities that are differences of O(1) numbers, 1e-10 otherwise."""
import numpy as np
import pytest

from helpers import ALL_KERNELS, kernel_value, relerr

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name,N,M", [("sho_under", 300, 120), ("c2_J6", 500, 77), ("mixed_J9", 257, 300), ("real", 100, 40)])
def test_predict_mean_variance_covariance(name, N, M):
    import torch

    from celerite2_b200 import predict

    dev = torch.device("cuda:0")
    T = lambda x: torch.as_tensor(np.ascontiguousarray(x, dtype=np.float64), device=dev)
    co = [np.array(v, dtype=np.float64) for v in ALL_KERNELS[name]]
    rng = np.random.default_rng(51)
    x = np.sort(rng.uniform(0, 10, N))
    xs = np.sort(rng.uniform(-1, 12, M))
    diag = rng.uniform(0.1, 0.3, N)
    y = np.sin(x) + 0.1 * rng.standard_normal(N)
    K = kernel_value(co, x[:, None] - x[None, :]) + np.diag(diag)
    Ks = kernel_value(co, x[:, None] - xs[None, :])  # (N, M)
    Kss = kernel_value(co, xs[:, None] - xs[None, :])
    alpha = np.linalg.solve(K, y)
    KinvKs = np.linalg.solve(K, Ks)
    cod = [T(v) for v in co]
    fac = predict.factorize(cod, T(x), T(diag))
    assert int(fac["status"].item()) == 0
    H = lambda v: v.detach().cpu().numpy()
    assert relerr(H(predict.kernel_matrix(cod, T(x), T(xs))), Ks) < 1e-12
    assert relerr(H(predict.mean(cod, fac, T(y), T(xs))), Ks.T @ alpha) < 1e-9
    assert relerr(H(predict.mean(cod, fac, T(y))), y - diag * alpha) < 1e-9
    k0 = float(kernel_value(co, np.zeros(1))[0])
    assert relerr(H(predict.variance(cod, fac, T(xs))), k0 - np.einsum("ij,ij->j", Ks, KinvKs)) < 1e-8
    assert relerr(H(predict.covariance(cod, fac, T(xs))), Kss - Ks.T @ KinvKs) < 1e-8
