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


def log_likelihood_and_grad(coeffs, t, y, diag, layout="blocked", want_data_grads=False, y_ready=None):
    """-> dict(loglike, status, grads = {ar, cr, ac, bc, cc, dc} [, t, y, diag]).
    y_ready: optional CUDA event after which y is valid (y is first read by solve_lower, so its upload can overlap the
    assembly and the factorisation; see HostLogLikelihood)."""
    c, a, U, V = ops.get_celerite_matrices(*coeffs, t, diag, layout=layout)
    d, W, S, status = ops.factor(t, c, a, U, V)
    if y_ready is not None:
        torch.cuda.current_stream().wait_event(y_ready)
    Y = y.unsqueeze(-1)
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


class HostLogLikelihood:
    """Host-buffer entry point: like the reference's GaussianProcess.compute + log_likelihood (core.py:262-317,407-428),
    which take host arrays, this takes HOST tensors (pinned for speed) of one series, uploads them, runs the device
    pipeline and returns the log-likelihood and the coefficient gradients in a pinned host buffer
    [loglike | d/d ar | d/d cr | d/d ac | d/d bc | d/d cc | d/d dc].

    t and diag are uploaded on the caller's stream; y goes up on a second stream and the pipeline waits for it only
    where y is first read (solve_lower), so its transfer overlaps the assembly and the factorisation.
    """

    def __init__(self, N, coeff_shapes, device):
        self.dev = device
        f8 = dict(dtype=torch.float64, device=device)
        self.t, self.y, self.diag = (torch.empty(N, **f8) for _ in range(3))
        self.co = [torch.empty(tuple(sh), **f8) for sh in coeff_shapes]
        self.n_out = 1 + sum(v.numel() for v in self.co)
        self.out_d = torch.empty(self.n_out, **f8)
        self.out_h = torch.empty(self.n_out, dtype=torch.float64).pin_memory()
        self.side = torch.cuda.Stream(device=device)
        self.y_ready = torch.cuda.Event()
        self.h2d_bytes = 8 * (3 * N + sum(v.numel() for v in self.co))
        self.d2h_bytes = 8 * self.n_out

    def __call__(self, coeffs_h, t_h, y_h, diag_h):
        main = torch.cuda.current_stream()
        for s, h in zip(self.co, coeffs_h):
            s.copy_(h, non_blocking=True)
        self.t.copy_(t_h, non_blocking=True)
        self.diag.copy_(diag_h, non_blocking=True)
        # y follows t and diag over the link (it would only share their bandwidth if started with them) and overlaps
        # the kernels that do not read it; waiting on `main` also orders it after the previous call's readers of y
        self.side.wait_stream(main)
        with torch.cuda.stream(self.side):
            self.y.copy_(y_h, non_blocking=True)
            self.y_ready.record(self.side)
        r = log_likelihood_and_grad(self.co, self.t, self.y, self.diag, y_ready=self.y_ready)
        self.out_d[0:1].copy_(r["loglike"].reshape(1))
        o = 1
        for nm in COEFF_NAMES:
            g = r["grads"][nm]
            self.out_d[o:o + g.numel()].copy_(g.reshape(-1))
            o += g.numel()
        self.out_h.copy_(self.out_d, non_blocking=True)
        main.synchronize()  # the caller needs the values
        return self.out_h, r["status"]


class HostLogLikelihoodStream:
    """Streaming form of HostLogLikelihood for callers that evaluate many series (or many parameter sets) in a row:
    `submit` enqueues the uploads of one evaluation on a copy stream and its kernels on the caller's stream and returns
    at once; `collect` waits for that evaluation's result.  With two slots the uploads of evaluation i+1 overlap the
    kernels of evaluation i.  Results come back in submission order.
    """

    def __init__(self, N, coeff_shapes, device, slots=2):
        self.dev = device
        f8 = dict(dtype=torch.float64, device=device)
        self.n_out = 1 + sum(int(torch.Size(sh).numel()) for sh in coeff_shapes)
        self.copy = torch.cuda.Stream(device=device)
        self.slots = []
        for _ in range(slots):
            self.slots.append(dict(
                t=torch.empty(N, **f8), y=torch.empty(N, **f8), diag=torch.empty(N, **f8),
                co=[torch.empty(tuple(sh), **f8) for sh in coeff_shapes],
                out_d=torch.empty(self.n_out, **f8), out_h=torch.empty(self.n_out, dtype=torch.float64).pin_memory(),
                td_ready=torch.cuda.Event(), y_ready=torch.cuda.Event(), done=torch.cuda.Event(), busy=False, status=None))
        self.next, self.pending = 0, []

    def submit(self, coeffs_h, t_h, y_h, diag_h):
        sl = self.slots[self.next]
        if sl["busy"]:
            raise RuntimeError("all slots in flight: collect() a result first")
        main = torch.cuda.current_stream()
        # the slot's previous evaluation (kernels on `main`) has been collected, hence finished: its buffers are free
        with torch.cuda.stream(self.copy):
            for s, h in zip(sl["co"], coeffs_h):
                s.copy_(h, non_blocking=True)
            sl["t"].copy_(t_h, non_blocking=True)
            sl["diag"].copy_(diag_h, non_blocking=True)
            sl["td_ready"].record(self.copy)
            sl["y"].copy_(y_h, non_blocking=True)
            sl["y_ready"].record(self.copy)
        main.wait_event(sl["td_ready"])
        r = log_likelihood_and_grad(sl["co"], sl["t"], sl["y"], sl["diag"], y_ready=sl["y_ready"])
        sl["out_d"][0:1].copy_(r["loglike"].reshape(1))
        o = 1
        for nm in COEFF_NAMES:
            g = r["grads"][nm]
            sl["out_d"][o:o + g.numel()].copy_(g.reshape(-1))
            o += g.numel()
        sl["out_h"].copy_(sl["out_d"], non_blocking=True)
        sl["done"].record(main)
        sl["busy"], sl["status"] = True, r["status"]
        self.pending.append(self.next)
        self.next = (self.next + 1) % len(self.slots)

    def collect(self):
        """-> (pinned host result of the oldest evaluation in flight, its status tensor); valid until the slot is reused"""
        sl = self.slots[self.pending.pop(0)]
        sl["done"].synchronize()
        sl["busy"] = False
        return sl["out_h"], sl["status"]
