"""GPU: seeded differential fuzz of the C ABI against the oracle — random series lengths (so that chunks and tiles end
at every possible offset), chunk counts, widths (lane-group, multi-right-hand-side and generic kernels), right-hand-side
counts and batch sizes."""
import numpy as np
import pytest

from helpers import ALL_KERNELS, relerr

pytestmark = pytest.mark.gpu
TOL = 1e-10
NAMES = ["sho_under", "sum4", "c2_J6", "c4_J16", "sum1", "sho4_J8", "real"]  # J = 2, 4, 6, 16 (compiled), 3, 8, 1 (generic)


def _case(seed):
    rng = np.random.default_rng(1000 + seed)
    name = NAMES[seed % len(NAMES)]
    N = int(rng.integers(1, 900)) if seed % 5 else int(rng.integers(1, 40))
    K = int(rng.choice([1, 1, 1, 2, 3, 7]))
    chunks = int(rng.choice([0, 1, 2, 3, 5, 13, 64]))
    batch = int(rng.choice([1, 1, 2, 3]))
    return name, N, K, chunks, batch


@pytest.mark.parametrize("seed", range(42))
def test_fuzz_all_operators(seed):
    name, N, K, chunks, batch = _case(seed)
    _check(ALL_KERNELS[name], name, N, K, chunks, batch, seed)


def _width_kernel(J):
    """J columns: SHO terms (two columns each) plus one real term when J is odd."""
    from celerite2_b200 import workloads as wl

    terms = [wl.sho_coeffs(1.0 / (i + 1), 0.5 + 0.7 * i, 1.0 + 0.5 * i) for i in range(J // 2)]
    if J % 2:
        terms.append(wl.real_coeffs(0.7, 0.3))
    return wl.sum_coeffs(*terms)


@pytest.mark.parametrize("J", list(range(1, 18)))
def test_every_width_all_operators(J):
    """Every compiled width (1..16) and one run-time width (17), one and three right-hand sides, whole series and chunks."""
    for K, chunks in ((1, 0), (1, 6), (3, 6)):
        _check(_width_kernel(J), "J=%d" % J, 333 + 7 * J, K, chunks, 1, 100 + J)


def _check(coeffs, name, N, K, chunks, batch, seed):
    import torch

    import oracle
    from celerite2_b200 import ops

    P = oracle.port()
    dev = torch.device("cuda:0")
    T = lambda x: torch.as_tensor(np.ascontiguousarray(x, dtype=np.float64), device=dev)
    H = lambda x: x.detach().cpu().numpy()
    rng = np.random.default_rng(seed)
    ts = [np.sort(rng.uniform(0, 0.1 * max(N, 2), N)) + 1e-3 * np.arange(N) for _ in range(batch)]
    diags = [rng.uniform(0.1, 0.3, N) ** 2 for _ in range(batch)]
    Ys = [rng.standard_normal((N, K)) for _ in range(batch)]
    mats = [P.get_celerite_matrices(*coeffs, t, dg) for t, dg in zip(ts, diags)]
    c = mats[0][0]
    J = mats[0][2].shape[1]
    st = lambda xs: T(np.stack(xs)) if batch > 1 else T(xs[0])
    t, a, U, V, Y = st(ts), st([m[1] for m in mats]), st([m[2] for m in mats]), st([m[3] for m in mats]), st(Ys)
    cc = T(c)
    ref = []
    for b in range(batch):
        flag, d_r, W_r, S_r = P.factor(ts[b], c, mats[b][1], mats[b][2], mats[b][3])
        assert flag == 0
        ref.append((d_r, W_r, S_r))
    W_ref = st([r[1] for r in ref])
    ops.set_chunks_per_series(chunks)
    try:
        d, W, S, status = ops.factor(t, cc, a, U, V)
        outs = {}
        for op in ("solve_lower", "solve_upper", "matmul_lower", "matmul_upper"):
            B = W_ref if op.startswith("solve") else V
            Z, F = getattr(ops, op)(t, cc, U, B, Y, workspace=True)
            bZ = torch.cos(Z)
            outs[op] = (Z, F, bZ, getattr(ops, op + "_rev")(t, cc, U, B, Y, Z, F, bZ))
        bd, bW = torch.sin(d), torch.cos(W)
        frev = ops.factor_rev(t, cc, a, U, V, st([r[0] for r in ref]), W_ref, st([r[2].reshape(N, J, J) for r in ref]), bd, bW)
    finally:
        ops.set_chunks_per_series(0)
    assert int(status.max().item()) == 0
    tag = "%s N=%d K=%d chunks=%d batch=%d" % (name, N, K, chunks, batch)
    sel = (lambda x, b: H(x)[b]) if batch > 1 else (lambda x, b: H(x))
    for b in range(batch):
        d_r, W_r, S_r = ref[b]
        assert relerr(sel(d, b), d_r) < TOL and relerr(sel(W, b), W_r) < TOL, tag
        assert relerr(sel(S, b).reshape(N, -1), S_r) < TOL, tag
        for op, (Z, F, bZ, rev) in outs.items():
            B_r = W_r if op.startswith("solve") else mats[b][3]
            Z_r, F_r = P.sweep(op, ts[b], c, mats[b][2], B_r, Ys[b])
            assert relerr(sel(Z, b), Z_r) < TOL and relerr(sel(F, b), F_r) < TOL, tag + " " + op
            want = P.sweep_rev(op, ts[b], c, mats[b][2], B_r, Ys[b], sel(Z, b), sel(F, b), sel(bZ, b))
            for u, v in zip(rev, want):
                assert relerr(sel(u, b), v) < 1e-9, tag + " " + op + "_rev"
        want = P.factor_rev(ts[b], c, mats[b][1], mats[b][2], mats[b][3], d_r, W_r, S_r, sel(bd, b), sel(bW, b))
        for u, v in zip(frev, want):
            assert relerr(sel(u, b), v) < 1e-9, tag + " factor_rev"
