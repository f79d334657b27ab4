"""TEST INFRASTRUCTURE ONLY: ctypes front-end for the checkers in this directory.

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline / --impl reference legs may import this
package.  The product package (celerite2_b200) never does.

`port()`  -> the plain-C restatement (oracle/c2_oracle.c, liboracle.so)
`ref()`   -> the reference's own C++ core built into oracle/_ref/libc2ref.so (strict flags), or None
`ref_fast()` -> same, -O3 -march=x86-64-v3 with wide fixed-size dispatch (timed CPU baseline), or None

All three expose the same numpy-level functions (see `_Lib`).
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_dp = ctypes.POINTER(ctypes.c_double)
_i64 = ctypes.c_int64
_i32 = ctypes.c_int32

OPS = {"solve_lower": 0, "solve_upper": 1, "matmul_lower": 2, "matmul_upper": 3}


def _p(x):
    if x is None:
        return ctypes.cast(None, _dp)
    assert x.dtype == np.float64 and x.flags.c_contiguous, "oracle wants C-contiguous float64"
    return x.ctypes.data_as(_dp)


def _f64(x):
    return np.ascontiguousarray(x, dtype=np.float64)


class _Lib:
    """numpy-level calls; inputs are converted to C-contiguous float64, outputs are freshly allocated
    unless the docstring says an argument is used in place."""

    def __init__(self, path, prefix, kind):
        self.path = path
        self.kind = kind
        self._l = ctypes.CDLL(path)
        g = lambda name: getattr(self._l, prefix + name)
        self._factor = g("factor")
        self._factor.restype = _i64
        self._factor.argtypes = [_i64, _i64] + [_dp] * 8
        self._factor_rev = g("factor_rev")
        self._factor_rev.restype = None
        self._factor_rev.argtypes = [_i64, _i64] + [_dp] * 15
        self._sweep = g("sweep")
        self._sweep.restype = None
        self._sweep.argtypes = [_i32, _i64, _i64, _i64] + [_dp] * 7
        self._sweep_rev = g("sweep_rev")
        self._sweep_rev.restype = None
        self._sweep_rev.argtypes = [_i32, _i64, _i64, _i64] + [_dp] * 13
        self._gm = g("general_matmul")
        self._gm.restype = None
        self._gm.argtypes = [_i32, _i64, _i64, _i64, _i64] + [_dp] * 8
        self._dense = g("to_dense")
        self._dense.restype = None
        self._dense.argtypes = [_i64, _i64] + [_dp] * 6
        self._mats = g("get_celerite_matrices")
        self._mats.restype = None
        if kind == "port":
            self._mats.argtypes = [_i32, _i64, _i64, _i64] + [_dp] * 12
        else:
            self._mats.argtypes = [_i64, _i64, _i64] + [_dp] * 12

    # ---- factor -------------------------------------------------------------------------------
    def factor(self, t, c, a, U, V, workspace=True):
        """-> (flag, d, W, S or None)"""
        t, c, a, U, V = map(_f64, (t, c, a, U, V))
        N, J = U.shape
        d = np.array(a, copy=True)
        W = np.array(V, copy=True)
        S = np.zeros((N, J * J)) if workspace else None
        flag = self._factor(N, J, _p(t), _p(c), _p(a), _p(U), _p(V), _p(d), _p(W), _p(S))
        return int(flag), d, W, S

    def factor_inplace(self, t, c, a, U, V):
        """d aliases a, W aliases V (numpy.py:69-77 usage); a and V are overwritten.  -> flag"""
        N, J = U.shape
        return int(self._factor(N, J, _p(t), _p(c), _p(a), _p(U), _p(V), _p(a), _p(V), _p(None)))

    def factor_rev(self, t, c, a, U, V, d, W, S, bd, bW):
        """-> (bt, bc, ba, bU, bV)"""
        t, c, a, U, V, d, W, S, bd, bW = map(_f64, (t, c, a, U, V, d, W, S, bd, bW))
        N, J = U.shape
        bt, bc, ba = np.zeros(N), np.zeros(J), np.zeros(N)
        bU, bV = np.zeros((N, J)), np.zeros((N, J))
        self._factor_rev(N, J, _p(t), _p(c), _p(a), _p(U), _p(V), _p(d), _p(W), _p(S), _p(bd), _p(bW),
                         _p(bt), _p(bc), _p(ba), _p(bU), _p(bV))
        return bt, bc, ba, bU, bV

    # ---- sweeps -------------------------------------------------------------------------------
    def sweep(self, op, t, c, U, W, Y, Z0=None, workspace=True, inplace=False):
        """op in OPS.  Y is (N,K).  For the matmuls Z0 is the caller's pre-filled Z (default zeros).
        inplace=True passes the same buffer as Y and Z.  -> (Z, F or None)"""
        t, c, U, W = map(_f64, (t, c, U, W))
        Y = _f64(Y)
        N, J = U.shape
        K = Y.shape[1]
        F = np.zeros((N, J * K)) if workspace else None
        if inplace:
            Z = np.array(Y, copy=True)
            self._sweep(OPS[op], N, J, K, _p(t), _p(c), _p(U), _p(W), _p(Z), _p(Z), _p(F))
            return Z, F
        Z = np.zeros_like(Y) if Z0 is None else np.array(_f64(Z0), copy=True)
        self._sweep(OPS[op], N, J, K, _p(t), _p(c), _p(U), _p(W), _p(Y), _p(Z), _p(F))
        return Z, F

    def sweep_rev(self, op, t, c, U, W, Y, Z, F, bZ):
        """-> (bt, bc, bU, bW, bY)"""
        t, c, U, W, Y, Z, F, bZ = map(_f64, (t, c, U, W, Y, Z, F, bZ))
        N, J = U.shape
        K = Y.shape[1]
        bt, bc = np.zeros(N), np.zeros(J)
        bU, bW, bY = np.zeros((N, J)), np.zeros((N, J)), np.zeros((N, K))
        self._sweep_rev(OPS[op], N, J, K, _p(t), _p(c), _p(U), _p(W), _p(Y), _p(Z), _p(F), _p(bZ),
                        _p(bt), _p(bc), _p(bU), _p(bW), _p(bY))
        return bt, bc, bU, bW, bY

    # ---- rectangular products -----------------------------------------------------------------
    def general_matmul(self, upper, t1, t2, c, U, V, Y, Z0=None, workspace=True):
        """-> (Z, F or None); F rows the reference never writes are left at the NaN fill used here."""
        t1, t2, c, U, V, Y = map(_f64, (t1, t2, c, U, V, Y))
        N, M = t1.shape[0], t2.shape[0]
        J, K = c.shape[0], Y.shape[1]
        Z = np.zeros((N, K)) if Z0 is None else np.array(_f64(Z0), copy=True)
        F = np.full((M, J * K), np.nan) if workspace else None
        self._gm(1 if upper else 0, N, M, J, K, _p(t1), _p(t2), _p(c), _p(U), _p(V), _p(Y), _p(Z), _p(F))
        return Z, F

    def to_dense(self, t, c, a, U, V):
        t, c, a, U, V = map(_f64, (t, c, a, U, V))
        N, J = U.shape
        K = np.zeros((N, N))
        self._dense(N, J, _p(t), _p(c), _p(a), _p(U), _p(V), _p(K))
        return K

    # ---- assembly -----------------------------------------------------------------------------
    def get_celerite_matrices(self, ar, cr, ac, bc, cc, dc, x, diag, layout="blocked"):
        """-> (c, a, U, V).  The _ref build only offers the blocked layout (terms.hpp:112-139)."""
        ar, cr, ac, bc, cc, dc, x, diag = map(_f64, (ar, cr, ac, bc, cc, dc, x, diag))
        N, Jr, Jc = x.shape[0], ar.shape[0], ac.shape[0]
        J = Jr + 2 * Jc
        c, a = np.zeros(J), np.zeros(N)
        U, V = np.zeros((N, J)), np.zeros((N, J))
        args = [_p(ar), _p(cr), _p(ac), _p(bc), _p(cc), _p(dc), _p(x), _p(diag), _p(c), _p(a), _p(U), _p(V)]
        if self.kind == "port":
            self._mats({"blocked": 0, "interleaved": 1}[layout], N, Jr, Jc, *args)
        else:
            if layout != "blocked":
                raise ValueError("the C++ reference core only has the blocked layout")
            self._mats(N, Jr, Jc, *args)
        return c, a, U, V


_cache = {}


def port():
    if "port" not in _cache:
        path = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(path):
            from . import build_oracle

            build_oracle.build_port()
        _cache["port"] = _Lib(path, "c2o_", "port")
    return _cache["port"]


def _ref(name):
    if name not in _cache:
        path = os.path.join(_HERE, "_ref", name)
        _cache[name] = _Lib(path, "c2ref_", "reference") if os.path.exists(path) else None
    return _cache[name]


def ref():
    return _ref("libc2ref.so")


def ref_fast():
    return _ref("libc2ref_fast.so")


class RefPipeline:
    """The log-likelihood (+ gradient) op chain of one series on ONE core, with every array allocated once
    (bench.py's `--impl reference` arm and `cpu_baseline` leg: first-touch page faults of fresh outputs are not
    charged to the reference).  `lib` is port() / ref() / ref_fast().  Not thread-safe: one instance per thread.

      value(...)          assemble -> factor (in place, no workspace) -> solve_lower -> log-det, chi^2
                          (python/celerite2/core.py:262-317,407-428; numpy.py:66-87,104-109)
      value_and_grad(...) assemble -> factor (S) -> solve_lower (F) -> reductions -> solve_lower_rev -> factor_rev
                          (the chain JAX runs, python/celerite2/jax/celerite2.py:34-69)
    """

    def __init__(self, lib, N, Jr, Jc, grad=True):
        self.lib, self.N, self.Jr, self.Jc = lib, int(N), int(Jr), int(Jc)
        J = self.J = Jr + 2 * Jc
        z = np.zeros
        self.c, self.a, self.U, self.V = z(J), z(N), z((N, J)), z((N, J))
        self.d, self.W = z(N), z((N, J))
        self.Z = z((N, 1))
        if grad:
            self.S, self.F = z((N, J * J)), z((N, J))
            self.bz, self.bd = z((N, 1)), z(N)
            self.bt, self.bc, self.bU, self.bW, self.bY = z(N), z(J), z((N, J)), z((N, J)), z((N, 1))
            self.bt2, self.bc2, self.ba, self.bU2, self.bV = z(N), z(J), z(N), z((N, J)), z((N, J))

    def _assemble(self, co, t, diag):
        L = self.lib
        args = [_p(_f64(v)) for v in co] + [_p(t), _p(diag), _p(self.c), _p(self.a), _p(self.U), _p(self.V)]
        if L.kind == "port":
            L._mats(0, self.N, self.Jr, self.Jc, *args)
        else:
            L._mats(self.N, self.Jr, self.Jc, *args)

    def value(self, co, t, y, diag):
        """-> (loglike, flag)"""
        L, N, J = self.lib, self.N, self.J
        self._assemble(co, t, diag)
        flag = L._factor(N, J, _p(t), _p(self.c), _p(self.a), _p(self.U), _p(self.V), _p(self.a), _p(self.V), _p(None))
        self.Z[:, 0] = y
        L._sweep(OPS["solve_lower"], N, J, 1, _p(t), _p(self.c), _p(self.U), _p(self.V), _p(self.Z), _p(self.Z), _p(None))
        z, d = self.Z[:, 0], self.a
        ll = -0.5 * (np.sum(z * z / d) + np.sum(np.log(d)) + N * np.log(2 * np.pi))
        return float(ll), int(flag)

    def value_and_grad(self, co, t, y, diag):
        """-> (loglike, bc of both reverse ops summed, flag); all per-row cotangents stay in the instance
        (bt + bt2, bU + bU2, bV, ba, bY)."""
        L, N, J = self.lib, self.N, self.J
        self._assemble(co, t, diag)
        flag = L._factor(N, J, _p(t), _p(self.c), _p(self.a), _p(self.U), _p(self.V), _p(self.d), _p(self.W), _p(self.S))
        Y = np.ascontiguousarray(y).reshape(N, 1)
        L._sweep(OPS["solve_lower"], N, J, 1, _p(t), _p(self.c), _p(self.U), _p(self.W), _p(Y), _p(self.Z), _p(self.F))
        z, d = self.Z[:, 0], self.d
        ll = -0.5 * (np.sum(z * z / d) + np.sum(np.log(d)) + N * np.log(2 * np.pi))
        np.divide(z, d, out=self.bz[:, 0])
        np.negative(self.bz[:, 0], out=self.bz[:, 0])
        np.multiply(self.bz[:, 0], self.bz[:, 0], out=self.bd)
        self.bd *= 0.5
        self.bd -= 0.5 / d
        L._sweep_rev(OPS["solve_lower"], N, J, 1, _p(t), _p(self.c), _p(self.U), _p(self.W), _p(Y), _p(self.Z), _p(self.F),
                     _p(self.bz), _p(self.bt), _p(self.bc), _p(self.bU), _p(self.bW), _p(self.bY))
        L._factor_rev(N, J, _p(t), _p(self.c), _p(self.a), _p(self.U), _p(self.V), _p(self.d), _p(self.W), _p(self.S),
                      _p(self.bd), _p(self.bW), _p(self.bt2), _p(self.bc2), _p(self.ba), _p(self.bU2), _p(self.bV))
        return float(ll), self.bc + self.bc2, int(flag)


def get_celerite_matrices_rev(ar, cr, ac, bc, cc, dc, x, b_c, b_a, b_U, b_V):
    """Adjoint of the blocked-layout assembly (c++/include/celerite2/terms.hpp:112-139; the derivative JAX takes of
    python/celerite2/jax/terms.py:119-135), restated in numpy: cotangents of (c, a, U, V) ->
    (b_ar, b_cr, b_ac, b_bc, b_cc, b_dc, b_x, b_diag).  Checked against central differences of the forward
    assembly in tests/test_oracle.py."""
    ar, cr, ac, bc, cc, dc, x, b_c, b_a, b_U, b_V = map(_f64, (ar, cr, ac, bc, cc, dc, x, b_c, b_a, b_U, b_V))
    Jr, Jc = ar.shape[0], ac.shape[0]
    sa = np.sum(b_a)
    b_ar = sa + np.sum(b_U[:, :Jr], axis=0)
    b_cr = b_c[:Jr].copy()
    arg = x[:, None] * dc[None, :]
    co, si = np.cos(arg), np.sin(arg)
    bU1, bU2 = b_U[:, Jr:Jr + Jc], b_U[:, Jr + Jc:]
    bV1, bV2 = b_V[:, Jr:Jr + Jc], b_V[:, Jr + Jc:]
    b_ac = sa + np.sum(bU1 * co + bU2 * si, axis=0)
    b_bc = np.sum(bU1 * si - bU2 * co, axis=0)
    b_cc = b_c[Jr:Jr + Jc] + b_c[Jr + Jc:]
    # d/d(arg) of the four columns of one complex term
    darg = bU1 * (-ac * si + bc * co) + bU2 * (ac * co + bc * si) - bV1 * si + bV2 * co
    b_dc = np.sum(darg * x[:, None], axis=0)
    b_x = darg @ dc
    return b_ar, b_cr, b_ac, b_bc, b_cc, b_dc, b_x, b_a.copy()
