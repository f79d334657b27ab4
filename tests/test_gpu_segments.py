"""GPU: one series split into contiguous segments (celerite2_b200/segments.py, the c2b_*_segment entry points):
summaries exchanged between segments, carried states computed on the device.  All segments run from one process on
one GPU here; the torch.distributed driver is exercised by dev/run_segments_2gpu.py under gpurun --gpus 2."""
import numpy as np
import pytest

from helpers import ALL_KERNELS, relerr

pytestmark = pytest.mark.gpu
TOL = 1e-10


@pytest.mark.parametrize("name", ["c2_J6", "sum4", "sho_under", "c4_J16"])
@pytest.mark.parametrize("bounds,chunks", [((0, 1700, 5000), 0), ((0, 1, 2000, 4999, 5000), 0), ((0, 40, 2100, 5000), 7),
                                           ((0, 1250, 2500, 3750, 5000), 3)])
def test_segments_reproduce_the_whole_series(name, bounds, chunks):
    import torch

    from celerite2_b200 import gp, ops, segments
    from celerite2_b200 import workloads as wl

    dev = torch.device("cuda:0")
    T = lambda x: torch.as_tensor(np.ascontiguousarray(x, dtype=np.float64), device=dev)
    H = lambda x: x.detach().cpu().numpy()
    N = bounds[-1]
    t, y, diag = wl.light_curve(N, 31)
    co = [T(np.array(v, dtype=np.float64)) for v in ALL_KERNELS[name]]
    tt, yy, dd = T(t), T(y), T(diag)
    ops.set_chunks_per_series(chunks)
    try:
        r = segments.loglike_and_grad_local(co, tt, yy, dd, list(bounds))
        ops.set_chunks_per_series(0)
        # reference: the same pipeline on the whole series, operator by operator
        c, a, U, V = ops.get_celerite_matrices(*co, tt, dd)
        d, W, S, st = ops.factor(tt, c, a, U, V)
        Z, F = ops.solve_lower(tt, c, U, W, yy[:, None], workspace=True)
        out, bz, bd = ops.norm_logdet(d, Z[:, 0], want_grad=True)
        bt, bc, bU, bW, bY = ops.solve_lower_rev(tt, c, U, W, yy[:, None], Z, F, bz[:, None])
        bt2, bc2, ba, bU2, bV = ops.factor_rev(tt, c, a, U, V, d, W, S, bd, bW)
        whole = gp.log_likelihood_and_grad(co, tt, yy, dd)
    finally:
        ops.set_chunks_per_series(0)
    segs = r["segments"]
    assert int(r["status"].item()) == 0 and int(st.item()) == 0
    cat = lambda nm: np.concatenate([H(getattr(s, nm)) for s in segs])
    assert relerr(cat("d"), H(d)) < TOL and relerr(cat("W"), H(W)) < TOL and relerr(cat("S"), H(S)) < TOL
    assert relerr(cat("Z"), H(Z)) < TOL and relerr(cat("F"), H(F)) < TOL
    assert relerr(cat("bW"), H(bW)) < TOL and relerr(cat("bY"), H(bY)) < TOL
    assert relerr(cat("ba"), H(ba)) < TOL and relerr(cat("bV"), H(bV)) < TOL
    assert relerr(cat("bU"), H(bU) + H(bU2)) < TOL
    assert relerr(sum(H(s.bc) for s in segs), H(bc) + H(bc2)) < TOL
    # bt: every segment owes one term to the row just below it.  coefficient_gradients() has already added the adjoint
    # of the assembly to bt, so compare against the whole-series pipeline with data gradients
    full = gp.log_likelihood_and_grad(co, tt, yy, dd, want_data_grads=True)
    bt_seg = [H(s.bt).copy() for s in segs]
    for i in range(1, len(segs)):
        bt_seg[i - 1][-1] += float(H(segs[i].edge_terms())[0])
    assert relerr(np.concatenate(bt_seg), H(full["grads"]["t"])) < 1e-9
    assert abs(float(r["loglike"]) - float(whole["loglike"])) < 1e-12 * abs(float(whole["loglike"]))
    for nm in gp.COEFF_NAMES:
        if whole["grads"][nm].numel():
            assert relerr(H(r["grads"][nm]), H(whole["grads"][nm])) < 1e-9, nm


def test_segment_argument_errors():
    import ctypes

    import torch

    from celerite2_b200 import _lib

    lib = _lib.lib
    dev = torch.device("cuda:0")
    x = torch.zeros(64 * 17, dtype=torch.float64, device=dev)
    st = torch.zeros(1, dtype=torch.int64, device=dev)
    seg = _lib.Segment(_lib.SEG_SUMMARISE, 0, 0, 0, None, x.data_ptr(), None)
    p = x.data_ptr()
    # no scratch buffer: refused (the tree has to survive between the two phases)
    rc = lib.c2b_factor_segment(64, 2, 1, p, p, p, p, p, p, p, None, st.data_ptr(), 3, None, 0, None, ctypes.byref(seg))
    assert rc != 0 and b"scratch" in lib.c2b_last_error()
    # a width without compiled lane-group kernels: refused
    buf = torch.empty(1 << 20, dtype=torch.uint8, device=dev)
    rc = lib.c2b_factor_segment(64, 17, 1, p, p, p, p, p, p, p, None, st.data_ptr(), 3, buf.data_ptr(), buf.numel(), None, ctypes.byref(seg))
    assert rc != 0 and b"compiled width" in lib.c2b_last_error()
    e, s = ctypes.c_int64(), ctypes.c_int64()
    assert lib.c2b_segment_sizes(_lib.OP_IDS["factor"], 6, 1, ctypes.byref(e), ctypes.byref(s)) == 0 and (e.value, s.value) == (108, 36)
    assert lib.c2b_segment_sizes(_lib.OP_IDS["matmul_lower"], 6, 1, ctypes.byref(e), ctypes.byref(s)) != 0


@pytest.mark.parametrize("name", ["c2_J6", "sum4"])
def test_solve_upper_segments(name):
    """solve_upper walks downwards (earlier segments = higher rows); its reverse operator walks upwards and owes its
    edge term to the row just above the segment."""
    import ctypes

    import torch

    import oracle
    from celerite2_b200 import _lib, segments
    from celerite2_b200 import workloads as wl

    lib = _lib.lib
    dev = torch.device("cuda:0")
    T = lambda x: torch.as_tensor(np.ascontiguousarray(x, dtype=np.float64), device=dev)
    H = lambda x: x.detach().cpu().numpy()
    P = oracle.port()
    N, bounds = 3000, [0, 700, 701, 2200, 3000]
    t, y, diag = wl.light_curve(N, 41)
    c, a, U, V = P.get_celerite_matrices(*ALL_KERNELS[name], t, diag)
    J = U.shape[1]
    _, d_r, W_r, _ = P.factor(t, c, a, U, V, workspace=False)
    Y = y[:, None]
    Z_r, F_r = P.sweep("solve_upper", t, c, U, W_r, Y)
    bZ = np.cos(Z_r)
    want = P.sweep_rev("solve_upper", t, c, U, W_r, Y, Z_r, F_r, bZ)
    tt, cc = T(t), T(c)
    nseg = len(bounds) - 1
    flags = _lib.SHARED_T | _lib.SHARED_C
    segs = []
    for r0, r1 in zip(bounds[:-1], bounds[1:]):
        n = r1 - r0
        s = dict(n=n, r0=r0, t=tt[r0:r1], U=T(U[r0:r1]), W=T(W_r[r0:r1]), Y=T(Y[r0:r1]), Z=torch.empty((n, 1), dtype=torch.float64, device=dev),
                 F=torch.empty((n, J), dtype=torch.float64, device=dev), bZ=T(bZ[r0:r1]), left=int(r0 > 0), right=int(r1 < N))
        for nm, shape in (("bt", (n,)), ("bc", (J,)), ("bU", (n, J)), ("bW", (n, J)), ("bY", (n, 1))):
            s[nm] = torch.empty(shape, dtype=torch.float64, device=dev)
        s["edge"] = torch.zeros(1, dtype=torch.float64, device=dev)
        for op in ("solve_upper", "solve_upper_rev"):
            need = lib.c2b_scratch_bytes(_lib.OP_IDS[op], n, 0, J, 1, 1)
            s["scr_" + op] = torch.empty(int(need) + 256, dtype=torch.uint8, device=dev)
        segs.append(s)
    stream = torch.cuda.current_stream().cuda_stream

    def fwd(s, phase, state=None, out=None):
        seg = _lib.Segment(phase, s["left"], s["right"], 0, None if state is None else state.data_ptr(), None if out is None else out.data_ptr(), None)
        b = s["scr_solve_upper"]
        _lib.check(lib.c2b_solve_upper_segment(s["n"], J, 1, 1, s["t"].data_ptr(), cc.data_ptr(), s["U"].data_ptr(), s["W"].data_ptr(),
                                               s["Y"].data_ptr(), s["Z"].data_ptr(), s["F"].data_ptr(), flags, b.data_ptr(), b.numel(), stream,
                                               ctypes.byref(seg)))

    def rev(s, phase, state=None, out=None):
        seg = _lib.Segment(phase, s["left"], s["right"], 0, None if state is None else state.data_ptr(), None if out is None else out.data_ptr(),
                           s["edge"].data_ptr())
        b = s["scr_solve_upper_rev"]
        _lib.check(lib.c2b_solve_upper_rev_segment(s["n"], J, 1, 1, s["t"].data_ptr(), cc.data_ptr(), s["U"].data_ptr(), s["W"].data_ptr(),
                                                   s["Y"].data_ptr(), s["Z"].data_ptr(), s["F"].data_ptr(), s["bZ"].data_ptr(), s["bt"].data_ptr(),
                                                   s["bc"].data_ptr(), s["bU"].data_ptr(), s["bW"].data_ptr(), s["bY"].data_ptr(), flags,
                                                   b.data_ptr(), b.numel(), stream, ctypes.byref(seg)))

    for op, call in (("solve_upper", fwd), ("solve_upper_rev", rev)):
        es, _ = segments.segment_sizes("solve_lower", J)
        summ = torch.empty((nseg, es), dtype=torch.float64, device=dev)
        for i, s in enumerate(segs):
            call(s, _lib.SEG_SUMMARISE, out=summ[i])
        for i, s in enumerate(segs):
            call(s, _lib.SEG_APPLY, state=segments.segment_carry(op, J, summ, i))
    cat = lambda nm: np.concatenate([H(s[nm]) for s in segs])
    assert relerr(cat("Z"), Z_r) < TOL and relerr(cat("F"), F_r) < TOL
    bt = [H(s["bt"]).copy() for s in segs]
    for i in range(nseg - 1):  # the edge term of segment i belongs to the first row of segment i + 1
        bt[i + 1][0] += float(H(segs[i]["edge"])[0])
    assert relerr(np.concatenate(bt), want[0]) < TOL
    assert relerr(sum(H(s["bc"]) for s in segs), want[1]) < TOL
    assert relerr(cat("bU"), want[2]) < TOL and relerr(cat("bW"), want[3]) < TOL and relerr(cat("bY"), want[4]) < TOL
