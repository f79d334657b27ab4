"""GPU: edge cases of the operator boundary -- shortest series, widths at both ends of the supported range, ragged
chunking (N not a multiple of the chunk length, odd N), argument errors -- and full-size (BASELINE c2: N = 1e6, J = 6)
size-independent properties: L (L^-1 y) = y, L^T (L^-T y) = y, and value/gradient consistency of the whole pipeline."""
import math

import numpy as np
import pytest

from helpers import ALL_KERNELS, relerr

pytestmark = pytest.mark.gpu
TOL = 1e-10


@pytest.fixture(scope="module")
def env():
    import torch

    from celerite2_b200 import ops

    dev = torch.device("cuda:0")
    T = lambda x: torch.as_tensor(np.ascontiguousarray(x, dtype=np.float64), device=dev)
    H = lambda x: x.detach().cpu().numpy()
    yield ops, T, H, torch
    ops.set_chunks_per_series(0)


@pytest.mark.parametrize("N", [1, 2, 3])
@pytest.mark.parametrize("name", ["real", "sum4", "c2_J6", "c4_J16"])
def test_shortest_series(env, N, name):
    """N = 1: the loops do not run, d = a, W = V / a (forward.hpp:107-108); N = 2, 3: one / two recursion steps."""
    ops, T, H, torch = env
    import oracle

    P = oracle.port()
    x = np.linspace(0.3, 1.1, N)
    diag = np.full(N, 0.4)
    c, a, U, V = P.get_celerite_matrices(*ALL_KERNELS[name], x, diag)
    Y = np.cos(np.arange(N * 2).reshape(N, 2) + 0.5)
    flag, d_r, W_r, S_r = P.factor(x, c, a, U, V)
    d, W, S, st = ops.factor(T(x), T(c), T(a), T(U), T(V))
    assert int(st.item()) == 0 == flag
    assert relerr(H(d), d_r) < TOL and relerr(H(W), W_r) < TOL and relerr(H(S).reshape(N, -1), S_r) < TOL
    for op in ("solve_lower", "solve_upper", "matmul_lower", "matmul_upper"):
        Z_r, F_r = P.sweep(op, x, c, U, W_r, Y)
        Z, F = getattr(ops, op)(T(x), T(c), T(U), T(W_r), T(Y), workspace=True)
        assert relerr(H(Z), Z_r) < TOL and relerr(H(F), F_r) < TOL, op
        bZ = np.sin(Y + 1.0)
        want = P.sweep_rev(op, x, c, U, W_r, Y, Z_r, F_r, bZ)
        got = getattr(ops, op + "_rev")(T(x), T(c), T(U), T(W_r), T(Y), T(Z_r), T(F_r), T(bZ))
        for u, v in zip(got, want):
            assert relerr(H(u), v) < TOL, op
    bd, bW = np.ones(N), np.cos(W_r)
    want = P.factor_rev(x, c, a, U, V, d_r, W_r, S_r, bd, bW)
    got = ops.factor_rev(T(x), T(c), T(a), T(U), T(V), T(d_r), T(W_r), T(S_r), T(bd), T(bW))
    for u, v in zip(got, want):
        assert relerr(H(u), v) < TOL


@pytest.mark.parametrize("N,chunks", [(1001, 0), (999, 7), (4097, 64), (130, 65), (257, 256)])
def test_ragged_chunking(env, N, chunks):
    """odd N, chunk lengths that do not divide N, almost one row per chunk"""
    ops, T, H, torch = env
    import oracle
    from celerite2_b200 import workloads as wl

    P = oracle.port()
    t, y, diag = wl.light_curve(N, 17)
    c, a, U, V = P.get_celerite_matrices(*ALL_KERNELS["c2_J6"], t, diag)
    _, d_r, W_r, S_r = P.factor(t, c, a, U, V)
    Y = y[:, None]
    Z_r, F_r = P.sweep("solve_lower", t, c, U, W_r, Y)
    ops.set_chunks_per_series(chunks)
    d, W, S, st = ops.factor(T(t), T(c), T(a), T(U), T(V))
    Z, F = ops.solve_lower(T(t), T(c), T(U), T(W_r), T(Y), workspace=True)
    bz = -Z_r / d_r[:, None]
    want = P.sweep_rev("solve_lower", t, c, U, W_r, Y, Z_r, F_r, bz)
    got = ops.solve_lower_rev(T(t), T(c), T(U), T(W_r), T(Y), T(Z_r), T(F_r), T(bz))
    want2 = P.factor_rev(t, c, a, U, V, d_r, W_r, S_r, np.cos(t), want[3])
    got2 = ops.factor_rev(T(t), T(c), T(a), T(U), T(V), T(d_r), T(W_r), T(S_r), T(np.cos(t)), T(want[3]))
    ops.set_chunks_per_series(0)
    assert int(st.item()) == 0
    assert relerr(H(d), d_r) < TOL and relerr(H(W), W_r) < TOL and relerr(H(S).reshape(N, -1), S_r) < TOL
    assert relerr(H(Z), Z_r) < TOL and relerr(H(F), F_r) < TOL
    for u, v in zip(got, want):
        assert relerr(H(u), v) < TOL
    for u, v in zip(got2, want2):
        assert relerr(H(u), v) < TOL


def test_width_limits_and_argument_errors(env):
    ops, T, H, torch = env
    import oracle
    from celerite2_b200 import workloads as wl
    from celerite2_b200._lib import C2BError

    P = oracle.port()
    # J = 32, the widest supported (run-time-width path)
    co = wl.sum_coeffs(*[wl.sho_coeffs(1.0 / (i + 1), 0.5 + 0.3 * i, 1.0 + 0.2 * i) for i in range(16)])
    t, y, diag = wl.light_curve(300, 2)
    c, a, U, V = P.get_celerite_matrices(*co, t, diag)
    assert U.shape[1] == 32
    _, d_r, W_r, _ = P.factor(t, c, a, U, V, workspace=False)
    for chunks in (1, 4):
        ops.set_chunks_per_series(chunks)
        d, W, _, st = ops.factor(T(t), T(c), T(a), T(U), T(V), workspace=False)
        Z, _ = ops.solve_lower(T(t), T(c), T(U), T(W_r), T(y[:, None]))
        assert int(st.item()) == 0 and relerr(H(d), d_r) < TOL and relerr(H(W), W_r) < 1e-9
        assert relerr(H(Z), P.sweep("solve_lower", t, c, U, W_r, y[:, None], workspace=False)[0]) < TOL
    ops.set_chunks_per_series(0)
    # J = 33 is refused with a message, not silently mis-computed
    U33 = torch.zeros((10, 33), dtype=torch.float64, device="cuda:0")
    with pytest.raises(C2BError, match="J must be"):
        ops.factor(T(t[:10]), torch.ones(33, dtype=torch.float64, device="cuda:0"), T(a[:10]), U33, U33)
    with pytest.raises(TypeError):
        ops.factor(T(t).float(), T(c), T(a), T(U), T(V))


def test_full_size_round_trips_and_gradient_consistency(env):
    """BASELINE c2 at full size (N = 1,000,000, J = 6), default chunk plan: properties that need no CPU reference."""
    ops, T, H, torch = env
    from celerite2_b200 import gp
    from celerite2_b200 import workloads as wl

    cfg = wl.config_c2()
    N = cfg["N"]
    t, y, diag = wl.light_curve(N, cfg["seed"])
    co = [T(v) for v in cfg["coeffs"]]
    tt, yy, dd = T(t), T(y), T(diag)
    c, a, U, V = ops.get_celerite_matrices(*co, tt, dd)
    d, W, _, st = ops.factor(tt, c, a, U, V, workspace=False)
    assert int(st.item()) == 0 and bool((d > 0).all())
    Y = torch.stack([yy, torch.cos(tt)], dim=1).contiguous()
    # L (L^-1 Y) = Y : matmul_lower with W, in place, is the product with L (numpy.py:100-102)
    Z, _ = ops.solve_lower(tt, c, U, W, Y)
    LZ = Z.clone()
    ops.matmul_lower(tt, c, U, W, LZ, Z=LZ)
    assert float((LZ - Y).abs().max() / Y.abs().max()) < 1e-9
    # L^T (L^-T Y) = Y : matmul_upper(t, c, U, W) multiplies by triu(W U^T), i.e. by L^T - I
    Zu, _ = ops.solve_upper(tt, c, U, W, Y)
    LtZ = Zu.clone()
    ops.matmul_upper(tt, c, U, W, Zu, Z=LtZ)
    assert float((LtZ - Y).abs().max() / Y.abs().max()) < 1e-9
    # K alpha = y with alpha = L^-T D^-1 L^-1 y, checked through K alpha = a alpha + tril(UV^T) alpha + triu(VU^T) alpha
    z1, _ = ops.solve_lower(tt, c, U, W, yy[:, None])
    al, _ = ops.solve_upper(tt, c, U, W, z1 / d[:, None])
    Ka = a[:, None] * al
    ops.matmul_lower(tt, c, U, V, al, Z=Ka)
    ops.matmul_upper(tt, c, U, V, al, Z=Ka)
    assert float((Ka[:, 0] - yy).abs().max() / yy.abs().max()) < 1e-8
    # value / gradient consistency of the whole pipeline along a random direction in coefficient space
    r = gp.log_likelihood_and_grad(co, tt, yy, dd)
    rng = np.random.default_rng(0)
    dirs = [T(rng.standard_normal(v.shape) * np.abs(v)) for v in cfg["coeffs"]]
    h = 1e-6
    lp, _ = gp.log_likelihood([v + h * dv for v, dv in zip(co, dirs)], tt, yy, dd)
    lm, _ = gp.log_likelihood([v - h * dv for v, dv in zip(co, dirs)], tt, yy, dd)
    fd = float((lp - lm).item()) / (2 * h)
    an = sum(float((r["grads"][nm] * dv).sum().item()) for nm, dv in zip(gp.COEFF_NAMES, dirs))
    assert abs(fd - an) <= 1e-5 * max(1.0, abs(an)), (fd, an)


@pytest.mark.parametrize("name,chunks", [("c2_J6", 0), ("c2_J6", 37), ("sho4_J8", 1), ("sho4_J8", 9), ("sum4", 5)])
def test_factor_rev_accumulate_flag(env, name, chunks):
    """C2B_ACCUMULATE: bt, bc, bU are added to (lane-group and generic kernels, chunked and not); ba, bV overwritten."""
    ops, T, H, torch = env
    import oracle
    from celerite2_b200 import workloads as wl

    P = oracle.port()
    N = 1500
    t, y, diag = wl.light_curve(N, 23)
    c, a, U, V = P.get_celerite_matrices(*ALL_KERNELS[name], t, diag)
    J = U.shape[1]
    _, d_r, W_r, S_r = P.factor(t, c, a, U, V)
    rng = np.random.default_rng(5)
    bd, bW = rng.standard_normal(N), rng.standard_normal((N, J))
    want = P.factor_rev(t, c, a, U, V, d_r, W_r, S_r, bd, bW)
    pre = [rng.standard_normal(N), rng.standard_normal(J), rng.standard_normal((N, J))]
    acc = [T(v) for v in pre]
    ops.set_chunks_per_series(chunks)
    bt, bc, ba, bU, bV = ops.factor_rev(T(t), T(c), T(a), T(U), T(V), T(d_r), T(W_r), T(S_r), T(bd), T(bW), accumulate_into=acc)
    ops.set_chunks_per_series(0)
    assert bt is acc[0] and bc is acc[1] and bU is acc[2]
    assert relerr(H(bt), want[0] + pre[0]) < TOL
    assert relerr(H(bc), want[1] + pre[1]) < TOL
    assert relerr(H(ba), want[2]) < TOL
    assert relerr(H(bU), want[3] + pre[2]) < TOL
    assert relerr(H(bV), want[4]) < TOL


@pytest.mark.parametrize("chunks", [0, 5])
def test_pointers_aligned_to_8_bytes_only(env, chunks):
    """The tile kernels copy with 16-byte cp.async when the arrays allow it; arrays that start on an odd multiple of
    8 bytes must take the 8-byte path and give the same results."""
    ops, T, H, torch = env
    import oracle
    from celerite2_b200 import workloads as wl

    def off(x):  # same values, storage shifted by one double
        x = np.ascontiguousarray(x, dtype=np.float64)
        buf = torch.empty(x.size + 1, dtype=torch.float64, device="cuda:0")
        v = buf[1:].view(x.shape)
        v.copy_(torch.from_numpy(x))
        assert v.data_ptr() % 16 == 8
        return v

    P = oracle.port()
    N = 1200
    t, y, diag = wl.light_curve(N, 9)
    c, a, U, V = P.get_celerite_matrices(*ALL_KERNELS["c2_J6"], t, diag)
    _, d_r, W_r, S_r = P.factor(t, c, a, U, V)
    Y = y[:, None]
    Z_r, F_r = P.sweep("solve_lower", t, c, U, W_r, Y)
    bz = np.cos(Z_r)
    ops.set_chunks_per_series(chunks)
    try:
        d, W, S, st = ops.factor(off(t), off(c), off(a), off(U), off(V), d=off(a), W=off(V))
        Z, F = ops.solve_lower(off(t), off(c), off(U), off(W_r), off(Y), workspace=True)
        M, _ = ops.matmul_lower(off(t), off(c), off(U), off(V), off(Y))
        got = ops.solve_lower_rev(off(t), off(c), off(U), off(W_r), off(Y), off(Z_r), off(F_r), off(bz))
        got2 = ops.factor_rev(off(t), off(c), off(a), off(U), off(V), off(d_r), off(W_r), off(S_r.reshape(N, 6, 6)), off(np.sin(t)), off(np.cos(W_r)))
    finally:
        ops.set_chunks_per_series(0)
    assert int(st.item()) == 0
    assert relerr(H(d), d_r) < TOL and relerr(H(W), W_r) < TOL and relerr(H(S).reshape(N, -1), S_r) < TOL
    assert relerr(H(Z), Z_r) < TOL and relerr(H(F), F_r) < TOL
    assert relerr(H(M), P.sweep("matmul_lower", t, c, U, V, Y, workspace=False)[0]) < TOL
    for u, v in zip(got, P.sweep_rev("solve_lower", t, c, U, W_r, Y, Z_r, F_r, bz)):
        assert relerr(H(u), v) < TOL
    for u, v in zip(got2, P.factor_rev(t, c, a, U, V, d_r, W_r, S_r, np.sin(t), np.cos(W_r))):
        assert relerr(H(u), v) < TOL
