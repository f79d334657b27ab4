"""Log-likelihood (+ gradient) pipelines assembled from the device operators.

This is the reference's `GaussianProcess.compute` + `log_likelihood` call chain
(python/celerite2/core.py:262-317,407-428; numpy.py:66-87,104-109) and, for the gradient, the op sequence JAX
runs under `value_and_grad` (python/celerite2/jax/celerite2.py:34-69, jax/ops.py:140-151):

    assemble -> factor (S saved) -> solve_lower (F saved) -> logdet / chi2
             -> solve_lower_rev -> factor_rev -> adjoint of the assembly

Everything stays on the device; only the log-likelihood and the coefficient gradients (a few doubles per
series) need to leave it.
"""
import math

import torch

from . import ops

COEFF_NAMES = ("ar", "cr", "ac", "bc", "cc", "dc")


def log_likelihood(coeffs, t, y, diag, layout="blocked"):
    """coeffs: six CUDA tensors (Jr,)/(Jc,) or (B,Jr)/(B,Jc); t (N) or (B,N); y (N) or (B,N); diag (N) or (B,N).
    -> (loglike (B,) or scalar tensor, status)"""
    c, a, U, V = ops.get_celerite_matrices(*coeffs, t, diag, layout=layout)
    d, W, _, status = ops.factor(t, c, a, U, V, d=a, W=V, workspace=False)
    Y = y.unsqueeze(-1)
    Z, _ = ops.solve_lower(t, c, U, W, Y)
    out, _, _ = ops.norm_logdet(d, Z.squeeze(-1))
    N = diag.shape[-1]
    ll = -0.5 * (out[..., 0] + out[..., 1] + N * math.log(2 * math.pi))
    return ll, status


def log_likelihood_and_grad(coeffs, t, y, diag, layout="blocked", want_data_grads=False):
    """-> dict(loglike, status, grads = {ar, cr, ac, bc, cc, dc} [, t, y, diag])."""
    c, a, U, V = ops.get_celerite_matrices(*coeffs, t, diag, layout=layout)
    d, W, S, status = ops.factor(t, c, a, U, V)
    Y = y.unsqueeze(-1)
    if Y.dim() < U.dim():  # shared y over a batch of coefficient sets
        pass
    Z, F = ops.solve_lower(t, c, U, W, Y, workspace=True)
    out, bz, bd = ops.norm_logdet(d, Z.squeeze(-1), want_grad=True)
    N = diag.shape[-1]
    ll = -0.5 * (out[..., 0] + out[..., 1] + N * math.log(2 * math.pi))
    bt, bc, bU, bW, bY = ops.solve_lower_rev(t, c, U, W, Y, Z, F, bz.unsqueeze(-1))
    _, _, ba, _, bV = ops.factor_rev(t, c, a, U, V, d, W, S, bd, bW, accumulate_into=(bt, bc, bU))
    g = ops.get_celerite_matrices_rev(*coeffs, t, bc, ba, bU, bV, b_x=bt, layout=layout, want_diag=want_data_grads)
    res = dict(loglike=ll, status=status, grads=dict(zip(COEFF_NAMES, g[:6])))
    if want_data_grads:
        res["grads"].update(t=g[6], y=bY.squeeze(-1), diag=g[7])
    return res
