"""Host mirror of the reference's `celerite2.driver` module (python/celerite2/driver.cpp:484-492).

Same function names, positional signatures, in-place conventions and error behaviour; numpy arrays in, numpy
arrays out.  Inputs are staged to the GPU (pinned H2D), the C ABI does the work, results are copied back into
the caller's output arrays.  There is no CPU fallback.

Like the reference (`py::array_t<double, py::array::c_style>`), arguments that are not C-contiguous float64
are converted to temporaries, so in-place semantics only hold for conforming arrays (driver.cpp:13-21).
"""
import numpy as np
import torch

from . import ops


class LinAlgError(Exception):
    """driver.LinAlgError (driver.cpp:483): "failed to factorize or solve matrix"."""


def _dev():
    if not torch.cuda.is_available():
        raise RuntimeError("celerite2_b200.driver: no CUDA device; there is no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


def _in(x, name, ndim=None):
    a = np.ascontiguousarray(x, dtype=np.float64)
    if ndim is not None and a.ndim != ndim:
        raise ValueError("Invalid number of dimensions: %s" % name)
    return a


def _up(a):
    return torch.from_numpy(a).to(_dev(), non_blocking=False)


def _down(t, out):
    """copy a device tensor into the caller's numpy array (or return a new array)"""
    res = t.detach().cpu().numpy()
    if isinstance(out, np.ndarray) and out.dtype == np.float64 and out.flags.c_contiguous and out.flags.writeable and out.shape == res.shape:
        out[...] = res
        return out
    return res


def _check(cond, name):
    if not cond:
        raise ValueError("Invalid shape: %s" % name)


def factor(t, c, a, U, V, d, W):
    """driver.factor (driver.cpp:13-64) -> (d, W); d may be the same array as a, W the same as V."""
    t_, c_, a_, U_, V_ = _in(t, "t"), _in(c, "c"), _in(a, "a"), _in(U, "U"), _in(V, "V")
    if t_.ndim <= 0:
        raise ValueError("Invalid number of dimensions: t")
    if c_.ndim <= 0:
        raise ValueError("Invalid number of dimensions: c")
    N, J = t_.shape[0], c_.shape[0]
    _check(t_.ndim == 1, "t")
    _check(c_.ndim == 1, "c")
    _check(a_.shape == (N,), "a")
    _check(U_.shape == (N, J), "U")
    _check(V_.shape == (N, J), "V")
    _check(np.shape(d) == (N,), "d")
    _check(np.shape(W) == (N, J), "W")
    a_d, V_d = _up(a_), _up(V_)
    d_d, W_d, _, status = ops.factor(_up(t_), _up(c_), a_d, _up(U_), V_d, d=a_d, W=V_d, workspace=False)
    if int(status.item()) != 0:
        # the reference leaves rows after the failure untouched in the caller's arrays; mirror that
        flag = int(status.item())
        dd, WW = d_d.cpu().numpy(), W_d.cpu().numpy()
        if isinstance(d, np.ndarray):
            d[: flag + 1] = dd[: flag + 1]
        if isinstance(W, np.ndarray):
            W[:flag] = WW[:flag]
        raise LinAlgError("failed to factorize or solve matrix")
    return _down(d_d, d), _down(W_d, W)


def _sweep(fn, t, c, U, W, Y, Z):
    t_, c_, U_, W_, Y_ = _in(t, "t"), _in(c, "c"), _in(U, "U"), _in(W, "W"), _in(Y, "Y")
    if Y_.ndim <= 1:
        raise ValueError("Invalid number of dimensions: Y")
    N, J, K = t_.shape[0], c_.shape[0], Y_.shape[1]
    _check(t_.ndim == 1, "t")
    _check(c_.ndim == 1, "c")
    _check(U_.shape == (N, J), "U")
    _check(W_.shape == (N, J), "W")
    _check(Y_.shape == (N, K), "Y")
    _check(np.shape(Z) == (N, K), "Z")
    Z_d = _up(_in(Z, "Z"))
    Y_d = Z_d if Z is Y else _up(Y_)
    out, _ = fn(_up(t_), _up(c_), _up(U_), _up(W_), Y_d, Z=Z_d)
    return _down(out, Z)


def solve_lower(t, c, U, W, Y, Z):
    """driver.solve_lower (driver.cpp:66-121): Z = L^-1 Y -> Z"""
    return _sweep(ops.solve_lower, t, c, U, W, Y, Z)


def solve_upper(t, c, U, W, Y, Z):
    return _sweep(ops.solve_upper, t, c, U, W, Y, Z)


def matmul_lower(t, c, U, V, Y, Z):
    """driver.matmul_lower: Z += tril(U V^T) Y -> Z (Z is accumulated into, as in the reference)"""
    return _sweep(ops.matmul_lower, t, c, U, V, Y, Z)


def matmul_upper(t, c, U, V, Y, Z):
    return _sweep(ops.matmul_upper, t, c, U, V, Y, Z)


def _general(fn, t1, t2, c, U, V, Y, Z):
    t1_, t2_, c_, U_, V_, Y_ = _in(t1, "t1"), _in(t2, "t2"), _in(c, "c"), _in(U, "U"), _in(V, "V"), _in(Y, "Y")
    if Y_.ndim <= 1:
        raise ValueError("Invalid number of dimensions: Y")
    N, M, J, K = t1_.shape[0], t2_.shape[0], c_.shape[0], Y_.shape[1]
    _check(U_.shape == (N, J), "U")
    _check(V_.shape == (M, J), "V")
    _check(Y_.shape == (M, K), "Y")
    _check(np.shape(Z) == (N, K), "Z")
    out = fn(_up(t1_), _up(t2_), _up(c_), _up(U_), _up(V_), _up(Y_), Z=_up(_in(Z, "Z")))
    return _down(out, Z)


def general_matmul_lower(t1, t2, c, U, V, Y, Z):
    return _general(ops.general_matmul_lower, t1, t2, c, U, V, Y, Z)


def general_matmul_upper(t1, t2, c, U, V, Y, Z):
    return _general(ops.general_matmul_upper, t1, t2, c, U, V, Y, Z)


def get_celerite_matrices(ar, ac, bc, dc, x, diag, a, U, V):
    """driver.get_celerite_matrices (driver.cpp:422-477): interleaved column order -> (a, U, V)"""
    ar_, ac_, bc_, dc_, x_, diag_ = [_in(v, n) for v, n in zip((ar, ac, bc, dc, x, diag), ("ar", "ac", "bc", "dc", "x", "diag"))]
    N, Jr, Jc = x_.shape[0], ar_.shape[0], ac_.shape[0]
    J = Jr + 2 * Jc
    if bc_.shape[0] != Jc:
        raise ValueError("dimension mismatch: bc")
    if dc_.shape[0] != Jc:
        raise ValueError("dimension mismatch: dc")
    if diag_.shape[0] != N:
        raise ValueError("dimension mismatch: diag")
    if np.shape(a) != (N,):
        raise ValueError("dimension mismatch: a")
    if np.shape(U) != (N, J):
        raise ValueError("dimension mismatch: U")
    if np.shape(V) != (N, J):
        raise ValueError("dimension mismatch: V")
    z = torch.zeros(Jr, dtype=torch.float64, device=_dev())
    zc = torch.zeros(Jc, dtype=torch.float64, device=_dev())
    _, a_d, U_d, V_d = ops.get_celerite_matrices(_up(ar_), z, _up(ac_), _up(bc_), zc, _up(dc_), _up(x_), _up(diag_), layout="interleaved")
    return _down(a_d, a), _down(U_d, U), _down(V_d, V)


__version__ = "0.1.0"
