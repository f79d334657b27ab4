"""GPU: the log-likelihood + gradient pipeline (celerite2_b200.gp) against the oracle and finite differences."""
import math

import numpy as np
import pytest

from helpers import ALL_KERNELS, relerr

pytestmark = pytest.mark.gpu


def _oracle_loglike(co, t, y, diag):
    import oracle

    P = oracle.port()
    c, a, U, V = P.get_celerite_matrices(*co, t, diag)
    flag, d, W, _ = P.factor(t, c, a, U, V, workspace=False)
    assert flag == 0
    Z, _ = P.sweep("solve_lower", t, c, U, W, y[:, None], workspace=False)
    z = Z[:, 0]
    return -0.5 * (np.sum(z * z / d) + np.sum(np.log(d)) + len(t) * math.log(2 * math.pi))


@pytest.mark.parametrize("layout", ["blocked", "interleaved"])
@pytest.mark.parametrize("name,chunks", [("sum2", 1), ("c2_J6", 9), ("sum4", 50)])
def test_loglike_and_coefficient_gradients(name, chunks, layout):
    import torch

    from celerite2_b200 import gp, ops
    from celerite2_b200 import workloads as wl

    dev = torch.device("cuda:0")
    co = [np.array(v, dtype=np.float64) for v in ALL_KERNELS[name]]
    N = 1500
    t, y, diag = wl.light_curve(N, 21)
    T = lambda x: torch.as_tensor(np.ascontiguousarray(x, dtype=np.float64), device=dev)
    ops.set_chunks_per_series(chunks)
    try:
        r = gp.log_likelihood_and_grad([T(v) for v in co], T(t), T(y), T(diag), layout=layout, want_data_grads=True)
        ll_only, status = gp.log_likelihood([T(v) for v in co], T(t), T(y), T(diag), layout=layout)
    finally:
        ops.set_chunks_per_series(0)
    ll = _oracle_loglike(co, t, y, diag)
    assert abs(float(r["loglike"].item()) - ll) <= 1e-10 * abs(ll)
    assert abs(float(ll_only.item()) - ll) <= 1e-10 * abs(ll)
    assert int(status.item()) == 0
    # coefficient gradients vs central differences of the oracle's log-likelihood
    for i, nm in enumerate(gp.COEFF_NAMES):
        g = r["grads"][nm].cpu().numpy()
        for j in range(len(co[i])):
            h = 1e-6 * max(1.0, abs(co[i][j]))
            cp = [v.copy() for v in co]
            cm = [v.copy() for v in co]
            cp[i][j] += h
            cm[i][j] -= h
            fd = (_oracle_loglike(cp, t, y, diag) - _oracle_loglike(cm, t, y, diag)) / (2 * h)
            assert abs(g[j] - fd) <= 2e-5 * max(1.0, abs(fd)), (nm, j, g[j], fd)
    # data gradients: d loglike / d y = -alpha (alpha = K^-1 y), checked by a directional difference
    rng = np.random.default_rng(0)
    for key, base in (("y", y), ("diag", diag), ("t", t)):
        v = rng.standard_normal(N)
        if key == "t":
            v = np.sort(v)  # t + h v stays sorted
        h = 1e-6
        args = {"t": t, "y": y, "diag": diag}
        ap, am = dict(args), dict(args)
        ap[key] = base + h * v
        am[key] = base - h * v
        fd = (_oracle_loglike(co, ap["t"], ap["y"], ap["diag"]) - _oracle_loglike(co, am["t"], am["y"], am["diag"])) / (2 * h)
        got = float(np.dot(r["grads"][key].cpu().numpy(), v))
        assert abs(got - fd) <= 2e-5 * max(1.0, abs(fd)), (key, got, fd)


def test_batched_pipeline_shared_data():
    """c5-style: many coefficient sets, shared t / y / diag."""
    import torch

    from celerite2_b200 import gp
    from celerite2_b200 import workloads as wl

    dev = torch.device("cuda:0")
    B, N = 6, 800
    cfg = wl.config_c5(batch=B, N=N)
    t, y, diag = wl.light_curve(N, 5)
    T = lambda x: torch.as_tensor(np.ascontiguousarray(x, dtype=np.float64), device=dev)
    co = [T(v) for v in cfg["coeffs"]]
    r = gp.log_likelihood_and_grad(co, T(t), T(np.tile(y, (B, 1))), T(np.tile(diag, (B, 1))))
    for b in range(B):
        cb = [v[b] for v in cfg["coeffs"]]
        ll = _oracle_loglike(cb, t, y, diag)
        assert abs(float(r["loglike"][b].item()) - ll) <= 1e-10 * abs(ll)
    assert relerr(r["grads"]["ac"].shape, (B, 2)) == 0.0


def test_host_buffer_entry_point_matches_device_pipeline():
    """HostLogLikelihood (pinned host arrays in, y uploaded on a second stream) == the device-resident pipeline, on
    repeated calls with changing data (the staging buffers are reused)."""
    import torch

    from celerite2_b200 import gp
    from celerite2_b200 import workloads as wl

    dev = torch.device("cuda:0")
    co = [np.array(v, dtype=np.float64) for v in ALL_KERNELS["c2_J6"]]
    N = 50000
    co_h = [torch.from_numpy(np.ascontiguousarray(v)).pin_memory() for v in co]
    api = gp.HostLogLikelihood(N, [v.shape for v in co_h], dev)
    for seed in (1, 2, 3):
        t, y, diag = wl.light_curve(N, seed)
        hs = [torch.from_numpy(v).pin_memory() for v in (t, y, diag)]
        out, status = api(co_h, *hs)
        assert int(status.item()) == 0
        got = out.numpy().copy()
        r = gp.log_likelihood_and_grad([v.to(dev) for v in co_h], *[v.to(dev) for v in hs])
        want = np.concatenate([r["loglike"].reshape(1).cpu().numpy()] + [r["grads"][nm].reshape(-1).cpu().numpy() for nm in gp.COEFF_NAMES])
        assert got.shape == want.shape
        assert np.max(np.abs(got - want) / np.maximum(1.0, np.abs(want))) < 1e-12
        assert abs(got[0] - _oracle_loglike(co, t, y, diag)) < 1e-10 * abs(got[0])


def test_streaming_host_entry_point():
    """HostLogLikelihoodStream: results in submission order, equal to the synchronous entry point, slots reused."""
    import torch

    from celerite2_b200 import gp
    from celerite2_b200 import workloads as wl

    dev = torch.device("cuda:0")
    co = [np.array(v, dtype=np.float64) for v in ALL_KERNELS["c2_J6"]]
    N = 40000
    co_h = [torch.from_numpy(np.ascontiguousarray(v)).pin_memory() for v in co]
    sync = gp.HostLogLikelihood(N, [v.shape for v in co_h], dev)
    stream = gp.HostLogLikelihoodStream(N, [v.shape for v in co_h], dev)
    data = [[torch.from_numpy(v).pin_memory() for v in wl.light_curve(N, seed)] for seed in (4, 5, 6, 7, 8)]
    want = [sync(co_h, *hs)[0].numpy().copy() for hs in data]
    got = []
    stream.submit(co_h, *data[0])
    for hs in data[1:]:
        stream.submit(co_h, *hs)
        got.append(stream.collect()[0].numpy().copy())
    got.append(stream.collect()[0].numpy().copy())
    with pytest.raises(IndexError):
        stream.collect()
    for g, w in zip(got, want):
        assert np.max(np.abs(g - w) / np.maximum(1.0, np.abs(w))) < 1e-12
    stream.submit(co_h, *data[0])
    stream.submit(co_h, *data[1])
    with pytest.raises(RuntimeError):
        stream.submit(co_h, *data[2])
    stream.collect(), stream.collect()


def test_graphed_pipeline_matches_eager():
    """GraphedLogLikelihood: replays equal the eager pipeline, also after the input buffers were overwritten."""
    import torch

    from celerite2_b200 import gp
    from celerite2_b200 import workloads as wl

    dev = torch.device("cuda:0")
    T = lambda x: torch.as_tensor(np.ascontiguousarray(x, dtype=np.float64), device=dev)
    co = [T(np.array(v, dtype=np.float64)) for v in ALL_KERNELS["c2_J6"]]
    N = 30000
    t, y, diag = [T(v) for v in wl.light_curve(N, 12)]
    g = gp.GraphedLogLikelihood(co, t, y, diag)
    for seed in (12, 13):
        tn, yn, dn = [T(v) for v in wl.light_curve(N, seed)]
        r = g(t=tn, y=yn, diag=dn)
        want = gp.log_likelihood_and_grad(co, tn, yn, dn)
        assert int(r["status"].item()) == 0
        assert abs(float(r["loglike"]) - float(want["loglike"])) <= 1e-13 * abs(float(want["loglike"]))
        for nm in gp.COEFF_NAMES:
            if want["grads"][nm].numel():
                assert torch.equal(r["grads"][nm], want["grads"][nm]) or relerr(r["grads"][nm].cpu().numpy(), want["grads"][nm].cpu().numpy()) < 1e-12
