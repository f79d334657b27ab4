"""Device-level operator API: torch CUDA tensors in, torch CUDA tensors out, zero-copy into the C ABI.

torch is plumbing here (device memory, streams); all arithmetic happens in libcelerite2_b200.so.  Each
function mirrors one reference operation (names, argument order and meaning as in
python/celerite2/definitions.json); a leading batch dimension on the per-series arrays selects the batched
kernels.  Shapes: t (N) or (B,N); c (J) or (B,J); a,d (..,N); U,V,W (..,N,J); Y,Z (..,N,K); S (..,N,J,J)
(S[n] is the reference's row reshaped, i.e. the TRANSPOSE of the column-major matrix); F (..,N,J*K) flat.
"""
import contextlib
import threading

import torch

from . import _lib
from ._lib import SHARED_C, SHARED_T, SHARED_Y, check, lib

# ---- scan scratch -------------------------------------------------------------------------------------------
# Eager calls share one buffer per (device, stream): calls on one stream are ordered, calls on different streams must
# not share scratch.  A buffer that was in use while a stream was capturing is baked into a CUDA graph by address and
# is therefore never released (it moves to _retired when a larger one replaces it).  Pipelines that are captured and
# replayed (gp.GraphedLogLikelihood) own their scratch outright: `with scratch_owner(obj_dict):` routes every request
# made inside the block to that dictionary, which lives as long as the object that holds the graphs.
_scratch = {}
_retired = []
_captured = set()
_tls = threading.local()


@contextlib.contextmanager
def scratch_owner(store):
    """Route the scratch requests of the calls made inside the block to `store` (a dict owned by the caller)."""
    prev = getattr(_tls, "store", None)
    _tls.store = store
    try:
        yield store
    finally:
        _tls.store = prev


def _ptr(x):
    return None if x is None else x.data_ptr()


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _scratch_buffer(tag, need, device):
    """-> uint8 tensor of at least `need` bytes for (tag, device, current stream), owned by the active store"""
    own = getattr(_tls, "store", None)
    store = _scratch if own is None else own
    idx = device.index if device.index is not None else torch.cuda.current_device()
    key = (tag, idx) if own is not None else (tag, idx, torch.cuda.current_stream(device).cuda_stream)
    buf = store.get(key)
    capturing = torch.cuda.is_current_stream_capturing()
    if buf is None or buf.numel() < need:
        if buf is not None and own is not None:
            own.setdefault("retired", []).append(buf)  # an earlier capture of the owner may hold its address
        elif buf is not None and key in _captured:
            _retired.append(buf)  # a graph the caller captured holds its address
        buf = torch.empty(int(need * 1.25) + 256, dtype=torch.uint8, device=device)
        store[key] = buf
        _captured.discard(key)
    if capturing and own is None:
        _captured.add(key)
    return buf


def _get_scratch(op, N, M, J, K, batch, device):
    need = lib.c2b_scratch_bytes(_lib.OP_IDS[op], N, M, J, K, batch)
    if need == 0:
        return None, 0
    buf = _scratch_buffer("scan", need, device)
    return buf.data_ptr(), buf.numel()


def _f64(x, name):
    if not (isinstance(x, torch.Tensor) and x.is_cuda):
        raise TypeError("%s must be a CUDA tensor" % name)
    if x.dtype != torch.float64:
        raise TypeError("%s must be float64 (got %s)" % (name, x.dtype))
    return x if x.is_contiguous() else x.contiguous()


def _batch_of(U):
    """(batch, N, J) from a (N,J) or (B,N,J) tensor"""
    if U.dim() == 2:
        return 1, U.shape[0], U.shape[1]
    if U.dim() == 3:
        return U.shape[0], U.shape[1], U.shape[2]
    raise ValueError("low-rank matrices must have shape (N,J) or (B,N,J)")


def _flags(batch, t, c, Y=None):
    f = 0
    if t.dim() == 1:
        f |= SHARED_T
    if c.dim() == 1:
        f |= SHARED_C
    if Y is not None and batch > 1 and Y.dim() == 2:
        f |= SHARED_Y
    return f


def set_chunks_per_series(chunks):
    """Force the number of chunks each series is cut into (0 = heuristic).  Testing / tuning knob."""
    lib.c2b_testing_set_chunks_per_series(int(chunks))


# ---- assembly ------------------------------------------------------------------------------------------------
def get_celerite_matrices(ar, cr, ac, bc, cc, dc, x, diag, layout="blocked", out=None):
    """-> (c, a, U, V).  Coefficient vectors (Jr,)/(Jc,) or (B,Jr)/(B,Jc); x (N) or (B,N); diag (N) or (B,N).
    out = (c, a, U, V): write into these tensors instead of fresh ones."""
    ar, cr, ac, bc, cc, dc, x, diag = [_f64(v, n) for v, n in zip((ar, cr, ac, bc, cc, dc, x, diag), "ar cr ac bc cc dc x diag".split())]
    batch = diag.shape[0] if diag.dim() == 2 else 1
    N = diag.shape[-1]
    Jr, Jc = ar.shape[-1], ac.shape[-1]
    J = Jr + 2 * Jc
    flags = (SHARED_T if x.dim() == 1 else 0) | (SHARED_C if ar.dim() == 1 and ac.dim() == 1 else 0)
    lead = (batch,) if diag.dim() == 2 else ()
    dev = diag.device
    if out is not None:
        c, a, U, V = out
    else:
        c = torch.empty(lead + (J,), dtype=torch.float64, device=dev)
        a = torch.empty(lead + (N,), dtype=torch.float64, device=dev)
        U = torch.empty(lead + (N, J), dtype=torch.float64, device=dev)
        V = torch.empty(lead + (N, J), dtype=torch.float64, device=dev)
    check(lib.c2b_get_celerite_matrices({"blocked": 0, "interleaved": 1}[layout], N, Jr, Jc, batch, _ptr(ar), _ptr(cr), _ptr(ac),
                                        _ptr(bc), _ptr(cc), _ptr(dc), _ptr(x), _ptr(diag), _ptr(c), _ptr(a), _ptr(U), _ptr(V), flags,
                                        _stream()))
    return c, a, U, V


def get_celerite_matrices_rev(ar, cr, ac, bc, cc, dc, x, b_c, b_a, b_U, b_V, b_x=None, layout="blocked", want_diag=True, out=None):
    """Adjoint of the assembly -> (b_ar, b_cr, b_ac, b_bc, b_cc, b_dc, b_x, b_diag).  b_x (if given) is accumulated into.
    out: optional six contiguous tensors that receive the coefficient cotangents (e.g. views of one packed result buffer)."""
    ar, cr, ac, bc, cc, dc, x, b_a, b_U, b_V = [_f64(v, "arg") for v in (ar, cr, ac, bc, cc, dc, x, b_a, b_U, b_V)]
    batch = b_a.shape[0] if b_a.dim() == 2 else 1
    N = b_a.shape[-1]
    Jr, Jc = ar.shape[-1], ac.shape[-1]
    flags = (SHARED_T if x.dim() == 1 else 0) | (SHARED_C if ar.dim() == 1 and ac.dim() == 1 else 0)
    lead = (batch,) if b_a.dim() == 2 else ()
    dev = b_a.device
    new = lambda n: torch.empty(lead + (n,), dtype=torch.float64, device=dev)
    if out is not None:
        b_ar, b_cr, b_ac, b_bc, b_cc, b_dc = out
        for v, n in zip(out, (Jr, Jr, Jc, Jc, Jc, Jc)):
            if v.numel() != batch * n or not v.is_contiguous() or v.dtype != torch.float64:
                raise ValueError("out: six contiguous float64 tensors shaped like the coefficient arrays")
    else:
        b_ar, b_cr, b_ac, b_bc, b_cc, b_dc = new(Jr), new(Jr), new(Jc), new(Jc), new(Jc), new(Jc)
    if b_x is None:
        b_x = torch.zeros(lead + (N,), dtype=torch.float64, device=dev)
    b_diag = new(N) if want_diag else None
    sp, sb = _get_scratch("assemble_rev", N, 0, Jr + 2 * Jc, 1, batch, dev)
    check(lib.c2b_get_celerite_matrices_rev({"blocked": 0, "interleaved": 1}[layout], N, Jr, Jc, batch, _ptr(ar), _ptr(cr), _ptr(ac),
                                            _ptr(bc), _ptr(cc), _ptr(dc), _ptr(x), _ptr(b_c), _ptr(b_a), _ptr(b_U), _ptr(b_V),
                                            _ptr(b_ar), _ptr(b_cr), _ptr(b_ac), _ptr(b_bc), _ptr(b_cc), _ptr(b_dc), _ptr(b_x),
                                            _ptr(b_diag), flags, sp, sb, _stream()))
    return b_ar, b_cr, b_ac, b_bc, b_cc, b_dc, b_x, b_diag


def to_dense(t, c, a, U, V):
    """-> K (N,N): the dense matrix of the representation (forward.hpp:19-42; test utility)"""
    t, c, a, U, V = [_f64(v, n) for v, n in zip((t, c, a, U, V), "tcaUV")]
    N, J = U.shape
    K = torch.empty((N, N), dtype=torch.float64, device=U.device)
    check(lib.c2b_to_dense(N, J, _ptr(t), _ptr(c), _ptr(a), _ptr(U), _ptr(V), _ptr(K), _stream()))
    return K


# ---- factor --------------------------------------------------------------------------------------------------
def factor(t, c, a, U, V, d=None, W=None, workspace=True, status=None):
    """-> (d, W, S, status).  d / W default to fresh tensors; pass d=a, W=V for the in-place form.
    status is an int64 tensor (batch): 0 or the first row with a non-positive pivot."""
    t, c, a, U, V = [_f64(v, n) for v, n in zip((t, c, a, U, V), "tcaUV")]
    batch, N, J = _batch_of(U)
    dev = U.device
    d = torch.empty_like(a) if d is None else d
    W = torch.empty_like(V) if W is None else W
    S = torch.empty(U.shape[:-1] + (J, J), dtype=torch.float64, device=dev) if workspace else None
    status = torch.empty(batch, dtype=torch.int64, device=dev) if status is None else status
    sp, sb = _get_scratch("factor", N, 0, J, 1, batch, dev)
    check(lib.c2b_factor(N, J, batch, _ptr(t), _ptr(c), _ptr(a), _ptr(U), _ptr(V), _ptr(d), _ptr(W), _ptr(S), _ptr(status),
                         _flags(batch, t, c), sp, sb, _stream()))
    return d, W, S, status


def factor_rev(t, c, a, U, V, d, W, S, bd, bW, accumulate_into=None):
    """-> (bt, bc, ba, bU, bV).  accumulate_into = (bt, bc, bU): add the cotangents of t, c and U to these arrays (those
    solve_*_rev filled) instead of returning fresh ones (C2B_ACCUMULATE)."""
    t, c, U, d, W, S, bd, bW = [_f64(v, "arg") for v in (t, c, U, d, W, S, bd, bW)]
    batch, N, J = _batch_of(U)
    dev = U.device
    lead = U.shape[:-2]
    extra = 0
    if accumulate_into is not None:
        bt, bc, bU = accumulate_into
        for v in (bt, bc, bU):
            if _f64(v, "accumulate_into") is not v:
                raise ValueError("accumulate_into: arrays must be contiguous (they are updated in place)")
        if bt.shape != lead + (N,) or bc.shape != lead + (J,) or bU.shape != U.shape:
            raise ValueError("accumulate_into: shapes must be those of bt, bc, bU")
        extra = _lib.ACCUMULATE
        bV = torch.empty_like(U)
    else:
        bt = torch.empty(lead + (N,), dtype=torch.float64, device=dev)
        bc = torch.empty(lead + (J,), dtype=torch.float64, device=dev)
        bU, bV = torch.empty_like(U), torch.empty_like(U)
    ba = torch.empty(lead + (N,), dtype=torch.float64, device=dev)
    sp, sb = _get_scratch("factor_rev", N, 0, J, 1, batch, dev)
    check(lib.c2b_factor_rev(N, J, batch, _ptr(t), _ptr(c), _ptr(a), _ptr(U), _ptr(V), _ptr(d), _ptr(W), _ptr(S), _ptr(bd), _ptr(bW),
                             _ptr(bt), _ptr(bc), _ptr(ba), _ptr(bU), _ptr(bV), _flags(batch, t, c) | extra, sp, sb, _stream()))
    return bt, bc, ba, bU, bV


# ---- sweeps --------------------------------------------------------------------------------------------------
def _sweep(op, t, c, U, W, Y, Z, workspace):
    t, c, U, W, Y = [_f64(v, n) for v, n in zip((t, c, U, W, Y), "tcUWY")]
    batch, N, J = _batch_of(U)
    K = Y.shape[-1]
    dev = U.device
    solve = op.startswith("solve")
    if Z is None:
        Z = torch.empty(U.shape[:-1] + (K,), dtype=torch.float64, device=dev) if solve else torch.zeros(
            U.shape[:-1] + (K,), dtype=torch.float64, device=dev)
    F = torch.empty(U.shape[:-1] + (J * K,), dtype=torch.float64, device=dev) if workspace else None
    sp, sb = _get_scratch(op, N, 0, J, K, batch, dev)
    fn = getattr(lib, "c2b_" + op)
    check(fn(N, J, K, batch, _ptr(t), _ptr(c), _ptr(U), _ptr(W), _ptr(Y), _ptr(Z), _ptr(F), _flags(batch, t, c, Y), sp, sb, _stream()))
    return Z, F


def solve_lower(t, c, U, W, Y, Z=None, workspace=False):
    """Z = L^-1 Y -> (Z, F).  Pass Z=Y for the in-place form."""
    return _sweep("solve_lower", t, c, U, W, Y, Z, workspace)


def solve_upper(t, c, U, W, Y, Z=None, workspace=False):
    return _sweep("solve_upper", t, c, U, W, Y, Z, workspace)


def matmul_lower(t, c, U, V, Y, Z=None, workspace=False):
    """Z += tril(U V^T) Y (Z defaults to zeros) -> (Z, F)"""
    return _sweep("matmul_lower", t, c, U, V, Y, Z, workspace)


def matmul_upper(t, c, U, V, Y, Z=None, workspace=False):
    return _sweep("matmul_upper", t, c, U, V, Y, Z, workspace)


def _composite(name, scratch_op, t, c, U, W, d, Y, Z):
    t, c, U, W, d, Y = [_f64(v, n) for v, n in zip((t, c, U, W, d, Y), "tcUWdY")]
    batch, N, J = _batch_of(U)
    K = Y.shape[-1]
    if Z is None:
        Z = torch.empty(U.shape[:-1] + (K,), dtype=torch.float64, device=U.device)
    sp, sb = _get_scratch(scratch_op, N, 0, J, K, batch, U.device)
    check(getattr(lib, name)(N, J, K, batch, _ptr(t), _ptr(c), _ptr(U), _ptr(W), _ptr(d), _ptr(Y), _ptr(Z), _flags(batch, t, c, Y), sp, sb,
                             _stream()))
    return Z


def apply_inverse(t, c, U, W, d, Y, Z=None):
    """Z = K^-1 Y: solve_lower, division by d (folded into the first sweep's store), solve_upper
    (python/celerite2/numpy.py:94-98).  Pass Z=Y for the in-place form."""
    return _composite("c2b_apply_inverse", "solve_lower", t, c, U, W, d, Y, Z)


def dot_tril(t, c, U, W, d, Y, Z=None):
    """Z = L (sqrt(d) o Y) (python/celerite2/numpy.py:100-102).  Pass Z=Y for the in-place form."""
    return _composite("c2b_dot_tril", "matmul_lower", t, c, U, W, d, Y, Z)


def _sweep_rev(op, t, c, U, W, Y, Z, F, bZ):
    t, c, U, W, Y, Z, F, bZ = [_f64(v, "arg") for v in (t, c, U, W, Y, Z, F, bZ)]
    batch, N, J = _batch_of(U)
    K = Z.shape[-1]
    dev = U.device
    lead = U.shape[:-2]
    bt = torch.empty(lead + (N,), dtype=torch.float64, device=dev)
    bc = torch.empty(lead + (J,), dtype=torch.float64, device=dev)
    bU, bW = torch.empty_like(U), torch.empty_like(U)
    bY = torch.empty_like(Z)
    sp, sb = _get_scratch(op, N, 0, J, K, batch, dev)
    fn = getattr(lib, "c2b_" + op)
    check(fn(N, J, K, batch, _ptr(t), _ptr(c), _ptr(U), _ptr(W), _ptr(Y), _ptr(Z), _ptr(F), _ptr(bZ), _ptr(bt), _ptr(bc), _ptr(bU),
             _ptr(bW), _ptr(bY), _flags(batch, t, c, Y), sp, sb, _stream()))
    return bt, bc, bU, bW, bY


def solve_lower_rev(t, c, U, W, Y, Z, F, bZ):
    """-> (bt, bc, bU, bW, bY)"""
    return _sweep_rev("solve_lower_rev", t, c, U, W, Y, Z, F, bZ)


def solve_upper_rev(t, c, U, W, Y, Z, F, bZ):
    return _sweep_rev("solve_upper_rev", t, c, U, W, Y, Z, F, bZ)


def matmul_lower_rev(t, c, U, V, Y, Z, F, bZ):
    return _sweep_rev("matmul_lower_rev", t, c, U, V, Y, Z, F, bZ)


def matmul_upper_rev(t, c, U, V, Y, Z, F, bZ):
    return _sweep_rev("matmul_upper_rev", t, c, U, V, Y, Z, F, bZ)


# ---- rectangular products ------------------------------------------------------------------------------------
def _general_matmul(upper, t1, t2, c, U, V, Y, Z, F):
    t1, t2, c, U, V, Y = [_f64(v, "arg") for v in (t1, t2, c, U, V, Y)]
    N, M, J, K = t1.shape[0], t2.shape[0], c.shape[0], Y.shape[1]
    dev = U.device
    if Z is None:
        Z = torch.zeros((N, K), dtype=torch.float64, device=dev)
    sp, sb = _get_scratch("general_matmul", N, M, J, K, 1, dev)
    fn = lib.c2b_general_matmul_upper if upper else lib.c2b_general_matmul_lower
    check(fn(N, M, J, K, _ptr(t1), _ptr(t2), _ptr(c), _ptr(U), _ptr(V), _ptr(Y), _ptr(Z), _ptr(F), sp, sb, _stream()))
    return Z


def general_matmul_lower(t1, t2, c, U, V, Y, Z=None, F=None):
    return _general_matmul(False, t1, t2, c, U, V, Y, Z, F)


def general_matmul_upper(t1, t2, c, U, V, Y, Z=None, F=None):
    return _general_matmul(True, t1, t2, c, U, V, Y, Z, F)


# ---- reductions ----------------------------------------------------------------------------------------------
def norm_logdet(d, z, want_grad=False, loglike=None):
    """-> (out, bz, bd): out[..., 0] = sum log d, out[..., 1] = sum z^2/d; bz, bd seed the backward pass of
    loglike = -0.5 (chi2 + logdet + N log 2 pi) when want_grad.  loglike: optional tensor of `batch` doubles that receives
    that value from the same launch (no follow-up arithmetic on the host side)."""
    d, z = _f64(d, "d"), _f64(z, "z")
    batch = d.shape[0] if d.dim() == 2 else 1
    N = d.shape[-1]
    dev = d.device
    out = torch.empty(d.shape[:-1] + (2,), dtype=torch.float64, device=dev)
    bz = torch.empty_like(z) if want_grad else None
    bd = torch.empty_like(d) if want_grad else None
    buf = _scratch_buffer("red", max(256 * 1024, batch * 592 * 16 + 256), dev)
    if loglike is not None:
        if loglike.numel() != batch or not loglike.is_contiguous() or loglike.dtype != torch.float64 or loglike.device != dev:
            raise ValueError("loglike must be a contiguous float64 tensor of batch elements on the device of d")
        check(lib.c2b_norm_logdet_loglike(N, batch, _ptr(d), _ptr(z), _ptr(out), _ptr(_f64(loglike, "loglike")), _ptr(bz), _ptr(bd),
                                          buf.data_ptr(), buf.numel(), _stream()))
    else:
        check(lib.c2b_norm_logdet(N, batch, _ptr(d), _ptr(z), _ptr(out), _ptr(bz), _ptr(bd), buf.data_ptr(), buf.numel(), _stream()))
    return out, bz, bd


def add_inplace(x, y):
    """y += x on the current stream (sum of two cotangents of the same array)."""
    x, y = _f64(x, "x"), _f64(y, "y")
    if x.numel() != y.numel():
        raise ValueError("size mismatch")
    check(lib.c2b_add_inplace(x.numel(), _ptr(x), _ptr(y), _stream()))
    return y
