"""Host mirror of the reference's `celerite2.backprop` module (python/celerite2/backprop.cpp:906-919).

`*_fwd(inputs..., outputs..., extra_outputs...)` zero the outputs first and return `(outputs..., extra...)`;
`*_rev(inputs, outputs, extra, b_outputs, b_inputs...)` return the `b_inputs` tuple.  Workspace shapes follow the
reference: S is (N, J, J), F is (N, J, nrhs) -- both are views of flat rows.
Unlike the reference's generated bindings (backprop.cpp:60,136) this is correct for J > 4.
"""
import numpy as np
import torch

from . import ops
from .driver import LinAlgError as _DriverLinAlgError
from .driver import _dev, _down, _in, _up


class LinAlgError(Exception):
    """backprop.LinAlgError (backprop.cpp:907)"""


def factor_fwd(t, c, a, U, V, d, W, S):
    N, J = np.shape(U)
    d_d, W_d, S_d, status = ops.factor(_up(_in(t, "t")), _up(_in(c, "c")), _up(_in(a, "a")), _up(_in(U, "U")), _up(_in(V, "V")))
    flag = int(status.item())
    if flag != 0:
        # backprop.cpp:48-65: the outputs are zeroed, the core writes d up to and including the failing row, W up to the row
        # before it and S up to and including it (forward.hpp:107-128), then the exception is raised; rows computed on the
        # device after the failure (unspecified for a chunked series) never reach the caller
        dd, WW, SS = d_d.cpu().numpy(), W_d.cpu().numpy(), S_d.reshape(N, J, J).cpu().numpy()
        for dst, src, rows in ((d, dd, flag + 1), (W, WW, flag), (S, SS, flag + 1)):
            if isinstance(dst, np.ndarray) and dst.dtype == np.float64 and dst.flags.c_contiguous and dst.flags.writeable and dst.shape == src.shape:
                dst[...] = 0.0
                dst[:rows] = src[:rows]
        raise LinAlgError("failed to factorize or solve matrix")
    return _down(d_d, d), _down(W_d, W), _down(S_d.reshape(N, J, J), S)


def factor_rev(t, c, a, U, V, d, W, S, bd, bW, bt, bc, ba, bU, bV):
    N, J = np.shape(U)
    Sf = _in(S, "S").reshape(N, J * J)
    outs = ops.factor_rev(*[_up(_in(v, "arg")) for v in (t, c, a, U, V, d, W)], _up(Sf), _up(_in(bd, "bd")), _up(_in(bW, "bW")))
    return tuple(_down(o, dst) for o, dst in zip(outs, (bt, bc, ba, bU, bV)))


def _fwd(fn, t, c, U, W, Y, Z, F):
    N, J = np.shape(U)
    K = np.shape(Y)[1]
    Z_d, F_d = fn(_up(_in(t, "t")), _up(_in(c, "c")), _up(_in(U, "U")), _up(_in(W, "W")), _up(_in(Y, "Y")), Z=None, workspace=True)
    return _down(Z_d, Z), _down(F_d.reshape(N, J, K), F)


def _rev(fn, t, c, U, W, Y, Z, F, bZ, bt, bc, bU, bW, bY):
    N, J = np.shape(U)
    K = np.shape(Y)[1]
    Ff = _in(F, "F").reshape(N, J * K)
    outs = fn(*[_up(_in(v, "arg")) for v in (t, c, U, W, Y, Z)], _up(Ff), _up(_in(bZ, "bZ")))
    return tuple(_down(o, dst) for o, dst in zip(outs, (bt, bc, bU, bW, bY)))


def solve_lower_fwd(t, c, U, W, Y, Z, F):
    return _fwd(ops.solve_lower, t, c, U, W, Y, Z, F)


def solve_lower_rev(t, c, U, W, Y, Z, F, bZ, bt, bc, bU, bW, bY):
    return _rev(ops.solve_lower_rev, t, c, U, W, Y, Z, F, bZ, bt, bc, bU, bW, bY)


def solve_upper_fwd(t, c, U, W, Y, Z, F):
    return _fwd(ops.solve_upper, t, c, U, W, Y, Z, F)


def solve_upper_rev(t, c, U, W, Y, Z, F, bZ, bt, bc, bU, bW, bY):
    return _rev(ops.solve_upper_rev, t, c, U, W, Y, Z, F, bZ, bt, bc, bU, bW, bY)


def matmul_lower_fwd(t, c, U, V, Y, Z, F):
    return _fwd(ops.matmul_lower, t, c, U, V, Y, Z, F)


def matmul_lower_rev(t, c, U, V, Y, Z, F, bZ, bt, bc, bU, bV, bY):
    return _rev(ops.matmul_lower_rev, t, c, U, V, Y, Z, F, bZ, bt, bc, bU, bV, bY)


def matmul_upper_fwd(t, c, U, V, Y, Z, F):
    return _fwd(ops.matmul_upper, t, c, U, V, Y, Z, F)


def matmul_upper_rev(t, c, U, V, Y, Z, F, bZ, bt, bc, bU, bV, bY):
    return _rev(ops.matmul_upper_rev, t, c, U, V, Y, Z, F, bZ, bt, bc, bU, bV, bY)


def _gm_fwd(fn, t1, t2, c, U, V, Y, Z, F):
    M, J, K = np.shape(t2)[0], np.shape(c)[0], np.shape(Y)[1]
    F_d = torch.zeros((M, J * K), dtype=torch.float64, device=_dev())  # the reference zeroes F first
    Z_d = fn(*[_up(_in(v, "arg")) for v in (t1, t2, c, U, V, Y)], Z=None, F=F_d)
    return _down(Z_d, Z), _down(F_d.reshape(M, J, K), F)


def general_matmul_lower_fwd(t1, t2, c, U, V, Y, Z, F):
    return _gm_fwd(ops.general_matmul_lower, t1, t2, c, U, V, Y, Z, F)


def general_matmul_upper_fwd(t1, t2, c, U, V, Y, Z, F):
    return _gm_fwd(ops.general_matmul_upper, t1, t2, c, U, V, Y, Z, F)


__version__ = "0.1.0"
