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
    r = _back_value(y, diag, _front_value(coeffs, t, diag, layout))
    return r["loglike"], r["status"]


def _front_value(coeffs, t, diag, layout):
    """value only: assembly + in-place factorisation without the workspace (numpy.py:66-87)"""
    c, a, U, V = ops.get_celerite_matrices(*coeffs, t, diag, layout=layout)
    d, W, _, status = ops.factor(t, c, a, U, V, d=a, W=V, workspace=False)
    return t, c, U, W, d, status


def _back_value(y, diag, front):
    t, c, U, W, d, status = front
    Z, _ = ops.solve_lower(t, c, U, W, y.unsqueeze(-1))
    ll = torch.empty(diag.shape[:-1], dtype=torch.float64, device=diag.device)
    ops.norm_logdet(d, Z.squeeze(-1), loglike=ll.reshape(-1))  # (the value comes out of the reduction's own launch)
    return dict(loglike=ll, status=status)


def _front(coeffs, t, diag, layout):
    """the part of the pipeline that does not read y: assembly + factorisation"""
    c, a, U, V = ops.get_celerite_matrices(*coeffs, t, diag, layout=layout)
    d, W, S, status = ops.factor(t, c, a, U, V)
    return c, a, U, V, d, W, S, status


def _packed_views(coeffs, diag):
    """one result buffer [loglike | d/d ar | d/d cr | d/d ac | d/d bc | d/d cc | d/d dc] (series-major blocks for a batch) and
    views of its seven parts shaped like the quantities they hold"""
    lead = tuple(diag.shape[:-1])
    nb = int(torch.Size(lead).numel()) if lead else 1
    sizes = [nb] + [v.numel() for v in coeffs]
    packed = torch.empty(sum(sizes), dtype=torch.float64, device=diag.device)
    offs = [0]
    for n in sizes:
        offs.append(offs[-1] + n)
    views = [packed[offs[0]:offs[1]].view(lead)] + [packed[offs[i + 1]:offs[i + 2]].view(v.shape) for i, v in enumerate(coeffs)]
    return packed, views


def _back(coeffs, t, y, diag, front, layout, want_data_grads, packed_views=None):
    """packed_views (from _packed_views): the log-likelihood and the coefficient gradients are written straight into them by
    the reduction and assembly-adjoint kernels (no host-side arithmetic, no concatenation)."""
    c, a, U, V, d, W, S, status = front
    Y = y.unsqueeze(-1)
    Z, F = ops.solve_lower(t, c, U, W, Y, workspace=True)
    N = diag.shape[-1]
    if packed_views is not None:
        ll = packed_views[0]
        out, bz, bd = ops.norm_logdet(d, Z.squeeze(-1), want_grad=True, loglike=ll.reshape(-1))
    else:
        out, bz, bd = ops.norm_logdet(d, Z.squeeze(-1), want_grad=True)
        ll = -0.5 * (out[..., 0] + out[..., 1] + N * math.log(2 * math.pi))
    bt, bc, bU, bW, bY = ops.solve_lower_rev(t, c, U, W, Y, Z, F, bz.unsqueeze(-1))
    _, _, ba, _, bV = ops.factor_rev(t, c, a, U, V, d, W, S, bd, bW, accumulate_into=(bt, bc, bU))
    g = ops.get_celerite_matrices_rev(*coeffs, t, bc, ba, bU, bV, b_x=bt, layout=layout, want_diag=want_data_grads,
                                      out=packed_views[1:] if packed_views is not None else None)
    res = dict(loglike=ll, status=status, grads=dict(zip(COEFF_NAMES, g[:6])))
    # the operator-level outputs (views, nothing copied): what the parity checks compare with the reference's arrays
    res["raw"] = dict(c=c, a=a, U=U, V=V, d=d, W=W, S=S, Z=Z, F=F, bt=bt, bc=bc, ba=ba, bU=bU, bV=bV, bY=bY)
    if want_data_grads:
        res["grads"].update(t=g[6], y=bY.squeeze(-1), diag=g[7])
    return res


def log_likelihood_and_grad(coeffs, t, y, diag, layout="blocked", want_data_grads=False, y_ready=None):
    """-> dict(loglike, status, grads = {ar, cr, ac, bc, cc, dc} [, t, y, diag]).
    y_ready: optional CUDA event after which y is valid (y is first read by solve_lower, so its upload can overlap the
    assembly and the factorisation; see HostLogLikelihood)."""
    front = _front(coeffs, t, diag, layout)
    if y_ready is not None:
        torch.cuda.current_stream().wait_event(y_ready)
    return _back(coeffs, t, y, diag, front, layout, want_data_grads)


class GraphedLogLikelihood:
    """log_likelihood_and_grad for FIXED input buffers, captured once in two CUDA graphs (the part that does not read y,
    and the rest) and replayed: the ~40 kernel launches of an evaluation become two graph launches, which takes the
    launch gaps out of the step.  Write new data into the tensors given here (or pass them to __call__), then replay.
    The results live in fixed tensors that every replay overwrites.
    """

    def __init__(self, coeffs, t, y, diag, layout="blocked", warmup=2, grad=True):
        self.coeffs, self.t, self.y, self.diag = list(coeffs), t, y, diag
        self.grad = grad
        # The graphs hold the scan scratch by address: this object owns it (ops.scratch_owner), so no later call, of
        # any size, on any stream, can replace or release memory a replay still writes to.
        self._scratch = {}
        side = torch.cuda.Stream(device=t.device)
        side.wait_stream(torch.cuda.current_stream())
        with ops.scratch_owner(self._scratch):
            with torch.cuda.stream(side):  # warm-up off the capturing stream: scratch buffers, function attributes
                for _ in range(warmup):
                    if grad:
                        log_likelihood_and_grad(self.coeffs, t, y, diag, layout=layout)
                    else:
                        _back_value(y, diag, _front_value(self.coeffs, t, diag, layout))
            torch.cuda.current_stream().wait_stream(side)
            self.g_front, self.g_back = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.g_front):
                self._front = _front(self.coeffs, t, diag, layout) if grad else _front_value(self.coeffs, t, diag, layout)
            with torch.cuda.graph(self.g_back, pool=self.g_front.pool()):
                if grad:
                    # [loglike | d/d ar | d/d cr | d/d ac | d/d bc | d/d cc | d/d dc] in one tensor (series-major blocks
                    # for a batch: all log-likelihoods, then all d/d ar, ...), written in place by the kernels
                    self.packed, views = _packed_views(self.coeffs, diag)
                    self.result = _back(self.coeffs, t, y, diag, self._front, layout, False, packed_views=views)
                else:
                    self.result = _back_value(y, diag, self._front)
                    self.packed = self.result["loglike"].reshape(-1)

    def replay_front(self):
        self.g_front.replay()

    def replay_back(self):
        self.g_back.replay()
        return self.result

    def __call__(self, coeffs=None, t=None, y=None, diag=None):
        for dst, src in list(zip(self.coeffs, coeffs or [])) + [(self.t, t), (self.y, y), (self.diag, diag)]:
            if src is not None and src is not dst:
                dst.copy_(src, non_blocking=True)
        self.g_front.replay()
        self.g_back.replay()
        return self.result


class HostLogLikelihood:
    """Host-buffer entry point: like the reference's GaussianProcess.compute + log_likelihood (core.py:262-317,407-428),
    which take host arrays, this takes HOST tensors (pinned for speed) of one series, uploads them, runs the device
    pipeline and returns the log-likelihood and the coefficient gradients in a pinned host buffer
    [loglike | d/d ar | d/d cr | d/d ac | d/d bc | d/d cc | d/d dc].

    t and diag are uploaded on the caller's stream; y goes up on a second stream and the pipeline waits for it only
    where y is first read (solve_lower), so its transfer overlaps the assembly and the factorisation.
    """

    def __init__(self, N, coeff_shapes, device, batch=None, shared_ty=False, grad=True):
        """batch: number of independent series (None = one series, the reference's call).  shared_ty: the series of a batch
        share t and y (hyper-parameter samples on one data set).  grad=False: log-likelihood only."""
        self.dev = device
        self.grad = grad
        f8 = dict(dtype=torch.float64, device=device)
        lead = () if batch is None else (int(batch),)
        ty = () if (batch is None or shared_ty) else lead
        # one device buffer, the inputs are views of it: small problems (launch-latency bound: c1) are packed on the host
        # and go up in ONE copy instead of nine
        shapes = [ty + (N,), ty + (N,), lead + (N,)] + [tuple(sh) for sh in coeff_shapes]
        sizes = [int(torch.Size(sh).numel()) for sh in shapes]
        offs = [0]
        for n in sizes:
            offs.append(offs[-1] + (n + 1) // 2 * 2)  # 16-byte aligned views
        self._buf = torch.empty(offs[-1], **f8)
        views = [self._buf[o:o + n].view(sh) for o, n, sh in zip(offs, sizes, shapes)]
        self.t, self.y, self.diag, self.co = views[0], views[1], views[2], views[3:]
        self._offs, self._sizes = offs, sizes
        self._packed_h = torch.empty(offs[-1], dtype=torch.float64).pin_memory() if offs[-1] * 8 <= (1 << 20) else None
        nb = 1 if batch is None else int(batch)
        self.n_out = nb + (sum(v.numel() for v in self.co) if grad else 0)
        self.out_h = torch.empty(self.n_out, dtype=torch.float64).pin_memory()
        self.side = torch.cuda.Stream(device=device)
        self.y_ready = torch.cuda.Event()
        self.h2d_bytes = 8 * (self.t.numel() + self.y.numel() + self.diag.numel() + sum(v.numel() for v in self.co))
        self.d2h_bytes = 8 * self.n_out
        self.graph = None  # captured at the first call (the staging buffers then hold real data)

    def __call__(self, coeffs_h, t_h, y_h, diag_h):
        main = torch.cuda.current_stream()
        if self._packed_h is not None:
            return self._call_packed(coeffs_h, t_h, y_h, diag_h, main)
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
        if self.graph is None:
            main.wait_event(self.y_ready)
            self.graph = GraphedLogLikelihood(self.co, self.t, self.y, self.diag, grad=self.grad)
        self.graph.replay_front()
        main.wait_event(self.y_ready)
        r = self.graph.replay_back()
        self.out_h.copy_(self.graph.packed, non_blocking=True)
        main.synchronize()  # the caller needs the values
        return self.out_h, r["status"]


    def _call_packed(self, coeffs_h, t_h, y_h, diag_h, main):
        """small problems: pack on the host (a few microseconds of memcpy), one upload, both graphs, one download"""
        for o, n, h in zip(self._offs, self._sizes, [t_h, y_h, diag_h] + list(coeffs_h)):
            self._packed_h[o:o + n].copy_(h.reshape(-1))
        self._buf.copy_(self._packed_h, non_blocking=True)
        if self.graph is None:
            self.graph = GraphedLogLikelihood(self.co, self.t, self.y, self.diag, grad=self.grad)
        self.graph.replay_front()
        r = self.graph.replay_back()
        self.out_h.copy_(self.graph.packed, non_blocking=True)
        main.synchronize()
        return self.out_h, r["status"]

    # (the per-array path below: large inputs, y on its own stream)
    __call__.__doc__ = "coeffs_h: six host tensors; t_h, y_h, diag_h: host tensors (pinned for speed) -> (pinned result, status)"


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
                td_ready=torch.cuda.Event(), y_ready=torch.cuda.Event(), done=torch.cuda.Event(), busy=False, status=None,
                graph=None))
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
        if sl["graph"] is None:
            main.wait_event(sl["y_ready"])
            sl["graph"] = GraphedLogLikelihood(sl["co"], sl["t"], sl["y"], sl["diag"])
        sl["graph"].replay_front()
        main.wait_event(sl["y_ready"])
        r = sl["graph"].replay_back()
        sl["out_h"].copy_(sl["graph"].packed, non_blocking=True)
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
