"""Predictive (conditional) distribution on the device: the reference's `ConditionalDistribution`
(python/celerite2/core.py:20-150, numpy.py:14-31) for a factorised GP, with the O(N M) pieces evaluated by kernels instead
of numpy broadcasting:

    mean        alpha = K^-1 (y - mean);  mu = K(xs, x) alpha   by the two rectangular sweeps (core.py:115-132)
    variance    k(0) - diag(K*^T K^-1 K*)                        (core.py:134-140): K* from c2b_kernel_matrix, K^-1 K* from
                                                                 c2b_apply_inverse (M right-hand sides), the reduction
                                                                 re-evaluates K* instead of reading it (c2b_predict_variance)
    covariance  K(xs, xs) - K*^T K^-1 K*                         (core.py:142-150)

`coeffs` are the six coefficient tensors (ar, cr, ac, bc, cc, dc) on the device; `fac` is what `factorize` returns.
"""
import torch

from . import _lib, ops
from ._lib import check, lib


def _p(x):
    return None if x is None else x.data_ptr()


def kernel_matrix(coeffs, x1, x2):
    """K[n, m] = k(x1[n] - x2[m]) -> (N, M) (python/celerite2/terms.py:64-85 evaluated on a grid)"""
    ar, cr, ac, bc, cc, dc = [ops._f64(v, "coeff") for v in coeffs]
    x1, x2 = ops._f64(x1, "x1"), ops._f64(x2, "x2")
    N, M = x1.shape[0], x2.shape[0]
    out = torch.empty((N, M), dtype=torch.float64, device=x1.device)
    check(lib.c2b_kernel_matrix(N, M, ar.shape[0], ac.shape[0], _p(ar), _p(cr), _p(ac), _p(bc), _p(cc), _p(dc), _p(x1), _p(x2), _p(out),
                                ops._stream()))
    return out


def factorize(coeffs, t, diag):
    """assembly + factorisation (GaussianProcess.compute, core.py:262-317) -> dict(t, c, U, V, W, d, diag, status)"""
    c, a, U, V = ops.get_celerite_matrices(*coeffs, t, diag)
    d, W, _, status = ops.factor(t, c, a, U, V, workspace=False)
    return dict(t=t, c=c, U=U, V=V, W=W, d=d, diag=diag, status=status)


def mean(coeffs, fac, resid, xs=None):
    """resid = y - mean(x).  xs None: the prediction at the observed times, y - diag * alpha up to the mean (core.py:121-125);
    otherwise K(xs, x) alpha through general_matmul_lower / upper (core.py:74-113).  -> (M,) without the mean function"""
    t, c, U, V, W, d = (fac[k] for k in ("t", "c", "U", "V", "W", "d"))
    alpha = ops.apply_inverse(t, c, U, W, d, resid.unsqueeze(-1))
    if xs is None:
        return resid - fac["diag"] * alpha.squeeze(-1)
    _, _, U2, V2 = ops.get_celerite_matrices(*coeffs, xs, torch.zeros_like(xs))
    mu = ops.general_matmul_lower(xs, t, c, U2, V, alpha)
    mu = ops.general_matmul_upper(xs, t, c, V2, U, alpha, Z=mu)
    return mu.squeeze(-1)


def _kinv_kstar(coeffs, fac, xs):
    Z = kernel_matrix(coeffs, fac["t"], xs)
    return ops.apply_inverse(fac["t"], fac["c"], fac["U"], fac["W"], fac["d"], Z, Z=Z)  # in place: K* itself is not kept


def variance(coeffs, fac, xs=None):
    """k(0) - diag(K*^T K^-1 K*) -> (M,)"""
    xs = fac["t"] if xs is None else ops._f64(xs, "xs")
    ar, cr, ac, bc, cc, dc = [ops._f64(v, "coeff") for v in coeffs]
    Z = _kinv_kstar(coeffs, fac, xs)
    N, M = Z.shape
    var = torch.empty(M, dtype=torch.float64, device=Z.device)
    need = lib.c2b_scratch_bytes(_lib.OP_IDS["predict_variance"], N, M, 1, 1, 1)
    buf = ops._scratch_buffer("predict", need, Z.device)
    check(lib.c2b_predict_variance(N, M, ar.shape[0], ac.shape[0], _p(ar), _p(cr), _p(ac), _p(bc), _p(cc), _p(dc), _p(fac["t"]), _p(xs),
                                   _p(Z), _p(var), buf.data_ptr(), buf.numel(), ops._stream()))
    return var


def covariance(coeffs, fac, xs=None):
    """K(xs, xs) - K*^T K^-1 K* -> (M, M)"""
    xs = fac["t"] if xs is None else ops._f64(xs, "xs")
    t, c, U, V = fac["t"], fac["c"], fac["U"], fac["V"]
    Z = _kinv_kstar(coeffs, fac, xs)
    _, _, U2, V2 = ops.get_celerite_matrices(*coeffs, xs, torch.zeros_like(xs))
    neg = -kernel_matrix(coeffs, xs, xs)
    neg = ops.general_matmul_lower(xs, t, c, U2, V, Z, Z=neg)
    neg = ops.general_matmul_upper(xs, t, c, V2, U, Z, Z=neg)
    return -neg
