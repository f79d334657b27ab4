"""GPU parity tests added in round 2 (VERDICT r01 "what's weak" 1-5):

  * the c4 shape for real: J = 16 with K = 64 right-hand sides, chunked and un-chunked, against the oracle
    (c++/include/celerite2/forward.hpp:156-207, reverse.hpp:87-151), plus the two compositions of
    python/celerite2/numpy.py:94-102 (c2b_apply_inverse, c2b_dot_tril);
  * the c2 bench workload at FULL size (N = 10^6, J = 6, ~7,000 chunks) against the reference's own C++ core
    (oracle/_ref/libc2ref.so, which travels to the GPU box): every operator output and all five gradient arrays;
  * one series of N = 10^6 split into 2 / 4 / 8 segments against the same reference;
  * the adjoint of the assembly against its numpy restatement in oracle/ (1e-10 instead of finite differences);
  * per-series gradients of the batched pipeline; scratch ownership of captured graphs.

Tolerance: max-norm relative 1e-10 (BASELINE.json north_star) unless a test says otherwise.
"""
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
    H = lambda x: None if x is None else x.detach().cpu().numpy()
    yield ops, T, H, torch
    ops.set_chunks_per_series(0)


def _best_ref():
    """the reference's own core when its compiled form is present (it travels with the snapshot), else the plain-C port"""
    import oracle

    return oracle.ref() or oracle.port()


def _factorised(name, N, seed):
    import oracle
    from celerite2_b200 import workloads as wl

    P = oracle.port()
    t, y, diag = wl.light_curve(N, seed)
    c, a, U, V = P.get_celerite_matrices(*ALL_KERNELS[name], t, diag)
    flag, d, W, _ = P.factor(t, c, a, U, V, workspace=False)
    assert flag == 0
    return P, t, c, a, U, V, d, W


# ---- c4 at its real width: J = 16, K = 64 ----------------------------------------------------------------------------
@pytest.mark.parametrize("chunks", [0, 3, 40])
def test_multi_rhs_J16_K64_against_oracle(env, chunks):
    ops, T, H, torch = env
    N, K = 4000, 64
    P, t, c, a, U, V, d, W = _factorised("c4_J16", N, 41)
    rng = np.random.default_rng(42)
    Y = rng.standard_normal((N, K))
    bZ = rng.standard_normal((N, K))
    tt, cc, UU, WW, YY = map(T, (t, c, U, W, Y))
    ops.set_chunks_per_series(chunks)
    try:
        for op in ("solve_lower", "solve_upper"):
            Z_o, F_o = P.sweep(op, t, c, U, W, Y)
            Z, F = getattr(ops, op)(tt, cc, UU, WW, YY, workspace=True)
            assert relerr(H(Z), Z_o) < TOL, op
            assert relerr(H(F), F_o) < TOL, op + " workspace"
            Zn, _ = getattr(ops, op)(tt, cc, UU, WW, YY)
            assert relerr(H(Zn), Z_o) < TOL, op + " (no workspace)"
            Zi = YY.clone()
            getattr(ops, op)(tt, cc, UU, WW, Zi, Z=Zi)
            assert relerr(H(Zi), Z_o) < TOL, op + " (in place)"
            want = P.sweep_rev(op, t, c, U, W, Y, Z_o, F_o, bZ)
            got = getattr(ops, op + "_rev")(tt, cc, UU, WW, YY, T(Z_o), T(F_o), T(bZ))
            for nm, g, w in zip(("bt", "bc", "bU", "bW", "bY"), got, want):
                assert relerr(H(g), w) < TOL, "%s_rev %s" % (op, nm)
    finally:
        ops.set_chunks_per_series(0)


@pytest.mark.parametrize("name,N,K,chunks", [("c4_J16", 4000, 64, 0), ("c4_J16", 4000, 64, 40), ("c2_J6", 3000, 1, 7),
                                             ("sum4", 2500, 5, 0), ("sum2", 900, 3, 4), ("c2_J6", 1200, 2, 1)])
def test_apply_inverse_and_dot_tril(env, name, N, K, chunks):
    """python/celerite2/numpy.py:94-102: _do_solve = solve_lower, /d, solve_upper; _do_dot_tril = (y sqrt d) then matmul_lower in place"""
    ops, T, H, torch = env
    P, t, c, a, U, V, d, W = _factorised(name, N, 43)
    Y = np.random.default_rng(44).standard_normal((N, K))
    z, _ = P.sweep("solve_lower", t, c, U, W, Y, workspace=False)
    z = z / d[:, None]
    want_inv, _ = P.sweep("solve_upper", t, c, U, W, z, workspace=False)
    want_dot, _ = P.sweep("matmul_lower", t, c, U, W, Y * np.sqrt(d)[:, None], workspace=False, inplace=True)
    tt, cc, UU, WW, dd, YY = map(T, (t, c, U, W, d, Y))
    ops.set_chunks_per_series(chunks)
    try:
        assert relerr(H(ops.apply_inverse(tt, cc, UU, WW, dd, YY)), want_inv) < TOL
        Zi = YY.clone()
        ops.apply_inverse(tt, cc, UU, WW, dd, Zi, Z=Zi)
        assert relerr(H(Zi), want_inv) < TOL
        assert relerr(H(ops.dot_tril(tt, cc, UU, WW, dd, YY)), want_dot) < TOL
        Zi = YY.clone()
        ops.dot_tril(tt, cc, UU, WW, dd, Zi, Z=Zi)
        assert relerr(H(Zi), want_dot) < TOL
    finally:
        ops.set_chunks_per_series(0)
    # K (K^-1 Y) = Y with the dense matrix of the same kernel (small case only)
    if N <= 1200:
        Kd = P.to_dense(t, c, a, U, V)
        assert relerr(Kd @ want_inv, Y) < 1e-9


# ---- more right-hand sides than the tile kernels have threads (256): the column-group loops and their accumulation ------
@pytest.mark.parametrize("chunks", [0, 5])
def test_more_columns_than_threads(env, chunks):
    ops, T, H, torch = env
    N, K = 300, 300
    P, t, c, a, U, V, d, W = _factorised("sum4", N, 43)
    rng = np.random.default_rng(44)
    Y = rng.standard_normal((N, K))
    bZ = rng.standard_normal((N, K))
    tt, cc, UU, YY = map(T, (t, c, U, Y))
    ops.set_chunks_per_series(chunks)
    try:
        for op, Bm in (("solve_lower", W), ("solve_upper", W), ("matmul_lower", V), ("matmul_upper", V)):
            Z_o, F_o = P.sweep(op, t, c, U, Bm, Y)
            Z, F = getattr(ops, op)(tt, cc, UU, T(Bm), YY, workspace=True)
            assert relerr(H(Z), Z_o) < TOL, op
            assert relerr(H(F), F_o) < TOL, op + " workspace"
            want = P.sweep_rev(op, t, c, U, Bm, Y, Z_o, F_o, bZ)
            got = getattr(ops, op + "_rev")(tt, cc, UU, T(Bm), YY, T(Z_o), T(F_o), T(bZ))
            for nm, g, w in zip(("bt", "bc", "bU", "bB", "bY"), got, want):
                assert relerr(H(g), w) < TOL, "%s_rev %s" % (op, nm)
    finally:
        ops.set_chunks_per_series(0)


# ---- the bench workload at full size against the reference's own core -----------------------------------------------
def _c2_reference(N):
    import oracle
    from celerite2_b200 import workloads as wl

    cfg = wl.config_c2(N=N)
    t, y, diag = wl.light_curve(N, cfg["seed"])
    pipe = oracle.RefPipeline(_best_ref(), N, 0, 3, grad=True)
    ll, bc, flag = pipe.value_and_grad(cfg["coeffs"], t, y, diag)
    assert flag == 0
    return cfg, t, y, diag, pipe, ll, bc


def test_c2_full_size_every_array_against_reference(env):
    ops, T, H, torch = env
    from celerite2_b200 import gp

    N = 1_000_000
    cfg, t, y, diag, pipe, ll, bc = _c2_reference(N)
    co = [T(v) for v in cfg["coeffs"]]
    r = gp.log_likelihood_and_grad(co, T(t), T(y), T(diag), want_data_grads=True)
    assert int(r["status"].item()) == 0
    assert abs(float(r["loglike"].item()) - ll) <= TOL * abs(ll)
    raw = r["raw"]
    worst = {}
    for nm, want in (("a", pipe.a), ("U", pipe.U), ("V", pipe.V), ("d", pipe.d), ("W", pipe.W), ("S", pipe.S), ("Z", pipe.Z), ("F", pipe.F),
                     ("bY", pipe.bY), ("ba", pipe.ba), ("bV", pipe.bV), ("bU", pipe.bU + pipe.bU2), ("bc", bc)):
        worst[nm] = relerr(H(raw[nm]).reshape(want.shape), want)
    # bt: compare before the assembly adjoint added its own term (raw["bt"] is that array after the accumulation), so
    # rebuild the reference's total: operator cotangents plus d(a,U,V)/dx
    import oracle

    g = oracle.get_celerite_matrices_rev(*cfg["coeffs"], t, bc, pipe.ba, pipe.bU + pipe.bU2, pipe.bV)
    worst["bt"] = relerr(H(r["grads"]["t"]), pipe.bt + pipe.bt2 + g[6])
    for nm, want in zip(gp.COEFF_NAMES, g[:6]):
        if want.size:
            worst["grad_" + nm] = relerr(H(r["grads"][nm]), want)
    bad = {k: v for k, v in worst.items() if not v < TOL}
    assert not bad, "arrays over 1e-10 at N = 1e6: %r (all: %r)" % (bad, worst)


@pytest.mark.parametrize("nseg", [2, 4, 8])
def test_split_series_full_size_against_reference(env, nseg):
    """one series of N = 10^6 cut into nseg segments (one per GPU in production; here all on one device), against the
    reference's core on the whole series"""
    ops, T, H, torch = env
    from celerite2_b200 import gp, segments

    import oracle

    N = 1_000_000
    cfg, t, y, diag, pipe, ll, bc = _c2_reference(N)
    co = [T(v) for v in cfg["coeffs"]]
    bounds = [N * i // nseg for i in range(nseg + 1)]
    bounds[1] += 1  # uneven on purpose
    r = segments.loglike_and_grad_local(co, T(t), T(y), T(diag), bounds)
    assert int(r["status"].item()) == 0
    assert abs(float(r["loglike"].item()) - ll) <= TOL * abs(ll)
    g = oracle.get_celerite_matrices_rev(*cfg["coeffs"], t, bc, pipe.ba, pipe.bU + pipe.bU2, pipe.bV)
    worst = {}
    for nm, want in zip(gp.COEFF_NAMES, g[:6]):
        if want.size:
            worst["grad_" + nm] = relerr(H(r["grads"][nm]), want)
    segs = r["segments"]
    for nm, want in (("d", pipe.d), ("W", pipe.W), ("Z", pipe.Z), ("bV", pipe.bV), ("ba", pipe.ba), ("bU", pipe.bU + pipe.bU2), ("bY", pipe.bY)):
        got = np.concatenate([H(getattr(s, nm)).reshape((s.n,) + want.shape[1:]) for s in segs])
        worst[nm] = relerr(got, want)
    # grad_dc is the one ill-conditioned number of the evaluation: a sum over all rows of t_n-weighted terms of both signs.
    # Element-wise relative noise of 3e-11 on bU and bV (what two builds of the REFERENCE differ by on these arrays) moves
    # it by 5e-8 .. 2e-7; its bar is 1e-9, every other output keeps 1e-10.  Measured: 9e-12 .. 1.7e-10 depending on where
    # the chunk boundaries fall (the un-split evaluation shows the same spread when its chunk length changes).
    rng = np.random.default_rng(5)
    noisy = oracle.get_celerite_matrices_rev(*cfg["coeffs"], t, bc, pipe.ba, (pipe.bU + pipe.bU2) * (1 + 3e-11 * rng.standard_normal(pipe.bU.shape)),
                                             pipe.bV * (1 + 3e-11 * rng.standard_normal(pipe.bV.shape)))
    assert relerr(noisy[5], g[5]) > 1e-8, "grad_dc is expected to be ill-conditioned on this workload"
    bars = {"grad_dc": 1e-9}
    bad = {k: v for k, v in worst.items() if not v < bars.get(k, TOL)}
    assert not bad, "split into %d: arrays over their bars: %r (all: %r)" % (nseg, bad, worst)


# ---- adjoint of the assembly against its restatement ----------------------------------------------------------------
@pytest.mark.parametrize("name", ["sum3", "c2_J6", "mixed_J9", "real"])
@pytest.mark.parametrize("layout", ["blocked", "interleaved"])
def test_assembly_adjoint_against_oracle(env, name, layout):
    ops, T, H, torch = env
    import oracle

    co = ALL_KERNELS[name]
    Jr, Jc = len(co[0]), len(co[2])
    J = Jr + 2 * Jc
    N = 20011
    rng = np.random.default_rng(45)
    x = np.sort(rng.uniform(0, 50, N))
    b_c, b_a, b_U, b_V = rng.standard_normal(J), rng.standard_normal(N), rng.standard_normal((N, J)), rng.standard_normal((N, J))
    want = oracle.get_celerite_matrices_rev(*co, x, b_c, b_a, b_U, b_V)  # blocked column order
    if layout == "interleaved":
        # [real.. | cos0 sin0 cos1 sin1 ..] -> the same cotangents laid out in the interleaved order
        perm = list(range(Jr)) + [Jr + (i // 2) + (Jc if i % 2 else 0) for i in range(2 * Jc)]
        b_c_l, b_U_l, b_V_l = b_c[perm], np.ascontiguousarray(b_U[:, perm]), np.ascontiguousarray(b_V[:, perm])
    else:
        b_c_l, b_U_l, b_V_l = b_c, b_U, b_V
    got = ops.get_celerite_matrices_rev(*[T(v) for v in co], T(x), T(b_c_l), T(b_a), T(b_U_l), T(b_V_l), layout=layout)
    for nm, g, w in zip(("b_ar", "b_cr", "b_ac", "b_bc", "b_cc", "b_dc", "b_x", "b_diag"), got, want):
        if w.size:
            assert relerr(H(g), w) < TOL, nm


# ---- batched pipeline: the gradients of every series -------------------------------------------------------------------
def test_batched_pipeline_gradients_per_series(env):
    ops, T, H, torch = env
    from celerite2_b200 import gp
    from celerite2_b200 import workloads as wl

    B, N = 6, 800
    cfg = wl.config_c5(batch=B, N=N)
    t, y, diag = wl.light_curve(N, 5)
    co = [T(v) for v in cfg["coeffs"]]
    r = gp.log_likelihood_and_grad(co, T(t), T(y), T(np.tile(diag, (B, 1))))  # shared t and y, per-series coefficients
    for b in range(B):
        one = gp.log_likelihood_and_grad([T(v[b]) for v in cfg["coeffs"]], T(t), T(y), T(diag))
        assert abs(float(r["loglike"][b]) - float(one["loglike"])) <= 1e-12 * abs(float(one["loglike"]))
        for nm in gp.COEFF_NAMES:
            if one["grads"][nm].numel():
                assert relerr(H(r["grads"][nm][b]), H(one["grads"][nm])) < TOL, (b, nm)


# ---- scratch ownership of captured graphs ---------------------------------------------------------
def test_graph_replay_survives_larger_eager_calls(env):
    """A graphed evaluator at N = 5e4 keeps working after eager calls at N = 1e6 made the shared scratch grow: its
    scratch belongs to the object (ops.scratch_owner), not to the per-device cache."""
    ops, T, H, torch = env
    from celerite2_b200 import gp
    from celerite2_b200 import workloads as wl

    co = [T(np.array(v, dtype=np.float64)) for v in ALL_KERNELS["c2_J6"]]
    small = [T(v) for v in wl.light_curve(50_000, 21)]
    g = gp.GraphedLogLikelihood(co, *small)
    first = g()
    ll0 = float(first["loglike"])
    g0 = {nm: first["grads"][nm].clone() for nm in gp.COEFF_NAMES}
    big = [T(v) for v in wl.light_curve(1_000_000, 22)]
    junk = []
    for _ in range(2):
        gp.log_likelihood_and_grad(co, *big)
        junk.append(torch.full((1 << 22,), float("nan"), dtype=torch.float64, device=small[0].device))  # lands in freed blocks
    g2 = gp.GraphedLogLikelihood(co, *big)  # a second owner, larger
    g2()
    again = g()
    torch.cuda.synchronize()
    assert float(again["loglike"]) == ll0
    for nm in gp.COEFF_NAMES:
        assert torch.equal(again["grads"][nm], g0[nm])
    eager = gp.log_likelihood_and_grad(co, *small)
    assert abs(float(eager["loglike"]) - ll0) <= 1e-13 * abs(ll0)


# ---- host entry point, small problems: one packed upload ---------------------------------------------------------------
@pytest.mark.parametrize("grad", [False, True])
def test_host_entry_point_packed_upload(env, grad):
    """N = 1000 (BASELINE config 1): HostLogLikelihood packs t, y, diag and the coefficients into one pinned buffer and
    uploads them in one copy; results equal the device-resident pipeline and the oracle, on repeated calls with new data."""
    ops, T, H, torch = env
    import oracle
    from celerite2_b200 import gp
    from celerite2_b200 import workloads as wl

    cfg = wl.config_c1()
    co = [np.ascontiguousarray(v, dtype=np.float64) for v in cfg["coeffs"]]
    co_h = [torch.from_numpy(v) for v in co]
    api = gp.HostLogLikelihood(cfg["N"], [v.shape for v in co_h], torch.device("cuda:0"), grad=grad)
    assert api._packed_h is not None
    for seed in (0, 1, 2):
        t, y, diag = wl.light_curve(cfg["N"], seed)
        out, status = api(co_h, torch.from_numpy(t), torch.from_numpy(y), torch.from_numpy(diag))
        assert int(status.max().item()) == 0
        ll_ref, _ = oracle.RefPipeline(_best_ref(), cfg["N"], len(co[0]), len(co[2]), grad=False).value(co, t, y, diag)
        assert abs(float(out[0]) - ll_ref) <= TOL * abs(ll_ref)
        if grad:
            r = gp.log_likelihood_and_grad([T(v) for v in co], T(t), T(y), T(diag))
            want = np.concatenate([H(r["loglike"]).reshape(1)] + [H(r["grads"][nm]).reshape(-1) for nm in gp.COEFF_NAMES])
            assert relerr(out.numpy(), want) < 1e-12
