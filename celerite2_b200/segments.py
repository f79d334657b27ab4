"""One long series split into contiguous segments, one per GPU ("one long series").

Every operator of the log-likelihood + gradient pipeline that carries a state along the series (factor,
solve_lower, solve_lower_rev, factor_rev) runs in two calls per segment (include/celerite2_b200.h, c2b_segment_t):
`summarise` (summary pass + up-sweep of the segment's own tree -> one summary element), then, after the summaries of
all segments have been exchanged and folded into the state entering this segment (c2b_segment_carry, on the device),
`apply` (down-sweep + apply pass).  The exchange is the only communication: one all-gather of 3 J^2 (factor), J^2 + J
(sweeps) or 2 J^2 (factor_rev) doubles per operator, plus one all-reduce of the partial log-likelihood and coefficient
gradients at the end.

`SeriesSegment` is one segment on the current device.  `loglike_and_grad_local` drives several segments from ONE
process on one device (the test of the algebra); `loglike_and_grad_distributed` is the torch.distributed version, one
rank per segment.  The time stamps are replicated (every rank holds all of t: 8 bytes per point), everything else is
local to the segment.  The gradient with respect to the time stamps is not assembled across segments here (the
operators do return the one term that crosses each boundary, `edge`).
"""
import ctypes
import math

import torch

from . import _lib, ops
from ._lib import check, lib

_OPS = _lib.OP_IDS
COEFF_NAMES = ("ar", "cr", "ac", "bc", "cc", "dc")


def _ptr(x):
    return None if x is None else x.data_ptr()


def _stream():
    return torch.cuda.current_stream().cuda_stream


def segment_sizes(op, J, K=1):
    e, s = ctypes.c_int64(), ctypes.c_int64()
    check(lib.c2b_segment_sizes(_OPS[op], J, K, ctypes.byref(e), ctypes.byref(s)))
    return e.value, s.value


def segment_carry(op, J, summaries, which):
    """summaries: (nseg, esize) device tensor in series order -> state entering segment `which` (device tensor)."""
    _, ss = segment_sizes(op, J)
    out = torch.empty(ss, dtype=torch.float64, device=summaries.device)
    check(lib.c2b_segment_carry(_OPS[op], J, 1, summaries.shape[0], which, _ptr(summaries), _ptr(out), _stream()))
    return out


class SeriesSegment:
    """Rows [r0, r1) of a series of N points.  t_full: all N time stamps on this device."""

    def __init__(self, coeffs, t_full, y_seg, diag_seg, r0, r1):
        self.dev = t_full.device
        self.N_total = t_full.shape[0]
        self.r0, self.r1, self.n = r0, r1, r1 - r0
        if y_seg.shape[0] != self.n or diag_seg.shape[0] != self.n:
            raise ValueError("y / diag must hold the segment's rows")
        self.coeffs = coeffs
        self.t = t_full[r0:r1]  # a view: t[-1] and t[n] are addressable through its pointer
        self.y, self.diag = y_seg.contiguous(), diag_seg.contiguous()
        self.has_left, self.has_right = int(r0 > 0), int(r1 < self.N_total)
        self.c, self.a, self.U, self.V = ops.get_celerite_matrices(*coeffs, self.t, self.diag)
        self.J = self.U.shape[1]
        self._scratch = {}
        self._edge = {}
        self._alloc = {}

    def reassemble(self):
        """a/U/V again from the current contents of coeffs, t, diag (new data in the same buffers)"""
        ops.get_celerite_matrices(*self.coeffs, self.t, self.diag, out=(self.c, self.a, self.U, self.V))

    def _once(self, key, make):
        """outputs and summaries are allocated at the first evaluation and reused by every later one"""
        v = self._alloc.get(key)
        if v is None:
            v = self._alloc[key] = make()
        return v

    # -- plumbing --------------------------------------------------------------------------------------------------
    def _scr(self, op):
        buf = self._scratch.get(op)
        if buf is None:
            need = lib.c2b_scratch_bytes(_OPS[op], self.n, 0, self.J, 1, 1)
            buf = torch.empty(int(need) + 256, dtype=torch.uint8, device=self.dev)
            self._scratch[op] = buf
        return buf.data_ptr(), buf.numel()

    def _seg(self, phase, state_in=None, summary_out=None, edge_out=None):
        return _lib.Segment(phase, self.has_left, self.has_right, 0, _ptr(state_in), _ptr(summary_out), _ptr(edge_out))

    def _new(self, *shape):
        return torch.empty(shape, dtype=torch.float64, device=self.dev)

    # -- factor ----------------------------------------------------------------------------------------------------
    def _factor(self, phase, state_in=None):
        n, J = self.n, self.J
        if phase == _lib.SEG_SUMMARISE:
            self.d, self.W, self.S, self.status = self._once("factor", lambda: (
                self._new(n), self._new(n, J), self._new(n, J, J), torch.zeros(1, dtype=torch.int64, device=self.dev)))
            out = self._once("factor_sum", lambda: self._new(segment_sizes("factor", J)[0]))
        else:
            out = None
        sp, sb = self._scr("factor")
        seg = self._seg(phase, state_in, out)
        check(lib.c2b_factor_segment(n, J, 1, _ptr(self.t), _ptr(self.c), _ptr(self.a), _ptr(self.U), _ptr(self.V), _ptr(self.d),
                                     _ptr(self.W), _ptr(self.S), _ptr(self.status), _lib.SHARED_T | _lib.SHARED_C, sp, sb, _stream(),
                                     ctypes.byref(seg)))
        return out

    def factor_summarise(self):
        return self._factor(_lib.SEG_SUMMARISE)

    def factor_apply(self, state_in):
        self._factor(_lib.SEG_APPLY, state_in)

    # -- solve_lower -----------------------------------------------------------------------------------------------
    def _solve(self, phase, state_in=None):
        n, J = self.n, self.J
        if phase == _lib.SEG_SUMMARISE:
            self.Y = self.y.unsqueeze(-1)
            self.Z, self.F = self._once("solve", lambda: (self._new(n, 1), self._new(n, J)))
            out = self._once("solve_sum", lambda: self._new(segment_sizes("solve_lower", J)[0]))
        else:
            out = None
        sp, sb = self._scr("solve_lower")
        seg = self._seg(phase, state_in, out)
        check(lib.c2b_solve_lower_segment(n, J, 1, 1, _ptr(self.t), _ptr(self.c), _ptr(self.U), _ptr(self.W), _ptr(self.Y), _ptr(self.Z),
                                          _ptr(self.F), _lib.SHARED_T | _lib.SHARED_C, sp, sb, _stream(), ctypes.byref(seg)))
        return out

    def solve_summarise(self):
        return self._solve(_lib.SEG_SUMMARISE)

    def solve_apply(self, state_in):
        self._solve(_lib.SEG_APPLY, state_in)

    def reductions(self):
        """-> (sum log d, sum z^2 / d) of this segment; seeds bz, bd of the backward pass."""
        out, self.bz, self.bd = ops.norm_logdet(self.d, self.Z.squeeze(-1), want_grad=True)
        return out

    # -- solve_lower_rev -------------------------------------------------------------------------------------------
    def _solve_rev(self, phase, state_in=None):
        n, J = self.n, self.J
        if phase == _lib.SEG_SUMMARISE:
            self.bt, self.bc, self.bU, self.bW, self.bY = self._once("solve_rev", lambda: (
                self._new(n), self._new(J), self._new(n, J), self._new(n, J), self._new(n, 1)))
            self._edge["solve_lower_rev"] = self._once("solve_rev_edge", lambda: torch.zeros(1, dtype=torch.float64, device=self.dev))
            out = self._once("solve_rev_sum", lambda: self._new(segment_sizes("solve_lower_rev", J)[0]))
        else:
            out = None
        sp, sb = self._scr("solve_lower_rev")
        seg = self._seg(phase, state_in, out, self._edge["solve_lower_rev"])
        check(lib.c2b_solve_lower_rev_segment(n, J, 1, 1, _ptr(self.t), _ptr(self.c), _ptr(self.U), _ptr(self.W), _ptr(self.Y),
                                              _ptr(self.Z), _ptr(self.F), _ptr(self.bz.unsqueeze(-1)), _ptr(self.bt), _ptr(self.bc),
                                              _ptr(self.bU), _ptr(self.bW), _ptr(self.bY), _lib.SHARED_T | _lib.SHARED_C, sp, sb,
                                              _stream(), ctypes.byref(seg)))
        return out

    def solve_rev_summarise(self):
        return self._solve_rev(_lib.SEG_SUMMARISE)

    def solve_rev_apply(self, state_in):
        self._solve_rev(_lib.SEG_APPLY, state_in)

    # -- factor_rev (accumulates into bt, bc, bU of solve_lower_rev) -----------------------------------------------------
    def _factor_rev(self, phase, state_in=None):
        n, J = self.n, self.J
        if phase == _lib.SEG_SUMMARISE:
            self.ba, self.bV = self._once("factor_rev", lambda: (self._new(n), self._new(n, J)))
            self._edge["factor_rev"] = self._once("factor_rev_edge", lambda: torch.zeros(1, dtype=torch.float64, device=self.dev))
            out = self._once("factor_rev_sum", lambda: self._new(segment_sizes("factor_rev", J)[0]))
        else:
            out = None
        sp, sb = self._scr("factor_rev")
        seg = self._seg(phase, state_in, out, self._edge["factor_rev"])
        check(lib.c2b_factor_rev_segment(n, J, 1, _ptr(self.t), _ptr(self.c), _ptr(self.a), _ptr(self.U), _ptr(self.V), _ptr(self.d),
                                         _ptr(self.W), _ptr(self.S), _ptr(self.bd), _ptr(self.bW), _ptr(self.bt), _ptr(self.bc),
                                         _ptr(self.ba), _ptr(self.bU), _ptr(self.bV),
                                         _lib.SHARED_T | _lib.SHARED_C | _lib.ACCUMULATE, sp, sb, _stream(), ctypes.byref(seg)))
        return out

    def factor_rev_summarise(self):
        return self._factor_rev(_lib.SEG_SUMMARISE)

    def factor_rev_apply(self, state_in):
        self._factor_rev(_lib.SEG_APPLY, state_in)

    def edge_terms(self):
        """the bt contributions this segment owes to row r0 - 1 (device scalar)"""
        return self._edge["solve_lower_rev"] + self._edge["factor_rev"]

    def coefficient_gradients(self):
        """partial gradients of this segment's rows, one flat tensor [ar | cr | ac | bc | cc | dc]"""
        g = ops.get_celerite_matrices_rev(*self.coeffs, self.t, self.bc, self.ba, self.bU, self.bV, b_x=self.bt, want_diag=False)
        return torch.cat([v.reshape(-1) for v in g[:6]])


_STEPS = (("factor", "factor_summarise", "factor_apply"),
          ("solve_lower", "solve_summarise", "solve_apply"),
          ("solve_lower_rev", "solve_rev_summarise", "solve_rev_apply"),
          ("factor_rev", "factor_rev_summarise", "factor_rev_apply"))


def _split_grads(flat, coeffs):
    out, o = {}, 0
    for nm, v in zip(COEFF_NAMES, coeffs):
        out[nm] = flat[o:o + v.numel()].reshape(v.shape)
        o += v.numel()
    return out


def loglike_and_grad_local(coeffs, t, y, diag, bounds):
    """All segments driven from one process on one device.  bounds: [0 = b0 < b1 < ... < bS = N].
    -> dict(loglike, status, grads, segments)"""
    segs = [SeriesSegment(coeffs, t, y[a:b], diag[a:b], a, b) for a, b in zip(bounds[:-1], bounds[1:])]
    J = segs[0].J
    red = None
    for op, summarise, apply in _STEPS:
        summaries = torch.stack([getattr(s, summarise)() for s in segs])
        for i, s in enumerate(segs):
            getattr(s, apply)(segment_carry(op, J, summaries, i))
        if op == "solve_lower":
            red = sum(s.reductions() for s in segs)
    N = t.shape[0]
    ll = -0.5 * (red[0] + red[1] + N * math.log(2 * math.pi))
    flat = sum(s.coefficient_gradients() for s in segs)
    status = torch.stack([s.status for s in segs]).max()
    return dict(loglike=ll, status=status, grads=_split_grads(flat, coeffs), segments=segs)


def loglike_and_grad_distributed(coeffs, t_full, y_seg, diag_seg, r0, r1, group=None):
    """torch.distributed version: this rank owns rows [r0, r1); ranks are ordered along the series.
    -> dict(loglike, status, grads), identical on every rank."""
    import torch.distributed as dist

    world, rank = dist.get_world_size(group), dist.get_rank(group)
    seg = SeriesSegment(coeffs, t_full, y_seg, diag_seg, r0, r1)
    red = None
    for op, summarise, apply in _STEPS:
        mine = getattr(seg, summarise)()
        allsum = torch.empty((world, mine.numel()), dtype=torch.float64, device=mine.device)
        dist.all_gather_into_tensor(allsum, mine, group=group)
        getattr(seg, apply)(segment_carry(op, seg.J, allsum, rank))
        if op == "solve_lower":
            red = seg.reductions()
    flat = seg.coefficient_gradients()
    pack = torch.cat([red.reshape(-1), flat, seg.status.to(torch.float64)])
    dist.all_reduce(pack[:-1], op=dist.ReduceOp.SUM, group=group)
    dist.all_reduce(pack[-1:], op=dist.ReduceOp.MAX, group=group)
    N = t_full.shape[0]
    ll = -0.5 * (pack[0] + pack[1] + N * math.log(2 * math.pi))
    return dict(loglike=ll, status=pack[-1].to(torch.int64), grads=_split_grads(pack[2:-1], coeffs))


class SplitSeries:
    """Rank-resident evaluator of ONE series split over the ranks of a torch.distributed group:
    the segment's buffers, scratch and exchange buffers are allocated once; `step()` re-evaluates the log-likelihood and
    its gradient from the current contents of `coeffs`, `t_full`, `y`, `diag` (write new data into them between steps).

    Per operator: summarise -> all_gather of one summary per rank -> device-side fold into the state entering this rank
    (c2b_segment_carry) -> apply.  One all_reduce at the end sums the partial log-determinant, chi^2 and coefficient
    gradients and carries, in a slot per rank, the bt term every segment owes the last row of its left neighbour
    (internal.hpp:226-245, the `bt[n-1] +=` of the row at a segment's lower edge), so that the gradient with respect to the
    time stamps is complete across segment boundaries as well.

    packed = [sum log d, chi^2, d/d ar .. d/d dc, status] after the reduction, identical on every rank.
    """

    def __init__(self, coeffs, t_full, y_seg, diag_seg, r0, r1, group=None):
        import torch.distributed as dist

        self.dist, self.group = dist, group
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.coeffs, self.t_full, self.y, self.diag = list(coeffs), t_full, y_seg, diag_seg
        self.seg = SeriesSegment(self.coeffs, t_full, y_seg, diag_seg, r0, r1)
        self.y, self.diag = self.seg.y, self.seg.diag  # the buffers the segment reads (contiguous)
        self.N = t_full.shape[0]
        J = self.seg.J
        dev = t_full.device
        self._gather = {op: torch.empty((self.world, segment_sizes(op, J)[0]), dtype=torch.float64, device=dev) for op, _, _ in _STEPS}
        self._state = {op: torch.empty(segment_sizes(op, J)[1], dtype=torch.float64, device=dev) for op, _, _ in _STEPS}
        ncoef = sum(v.numel() for v in self.coeffs)
        self.n_out = 2 + ncoef + 1
        self.packed = torch.zeros(2 + ncoef + self.world + 1, dtype=torch.float64, device=dev)
        self._graph, self._result = None, None

    def capture(self, warmup=3):
        """Record one evaluation -- kernels AND the NCCL exchanges between them -- in a CUDA graph (every rank of the group
        must call this together); `step()` then replays it: the host no longer sits between summarise, exchange, carry and
        apply of the four operators.  Returns False (and stays eager) if the capture fails."""
        for _ in range(warmup):
            self._step_eager()
        torch.cuda.synchronize()
        try:
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                res = self._step_eager()
            torch.cuda.synchronize()
        except Exception as e:  # pragma: no cover  (depends on the NCCL / driver combination)
            self._graph, self.capture_error = None, repr(e)
            torch.cuda.synchronize()
            return False
        self._graph, self._result = g, res
        return True

    def step(self):
        if self._graph is not None:
            self._graph.replay()
            return self._result
        return self._step_eager()

    def _step_eager(self):
        seg, dist = self.seg, self.dist
        seg.reassemble()
        red = None
        for op, summarise, apply in _STEPS:
            mine = getattr(seg, summarise)()
            buf = self._gather[op]
            dist.all_gather_into_tensor(buf, mine, group=self.group)
            st = self._state[op]
            check(lib.c2b_segment_carry(_OPS[op], seg.J, 1, self.world, self.rank, _ptr(buf), _ptr(st), _stream()))
            getattr(seg, apply)(st)
            if op == "solve_lower":
                red = seg.reductions()
        flat = seg.coefficient_gradients()
        p = self.packed
        nc = flat.numel()
        p[0:2] = red.reshape(-1)
        p[2:2 + nc] = flat
        p[2 + nc:2 + nc + self.world] = 0.0
        p[2 + nc + self.rank] = seg.edge_terms()[0]
        p[-1] = seg.status[0].to(torch.float64)
        dist.all_reduce(p[:-1], op=dist.ReduceOp.SUM, group=self.group)
        dist.all_reduce(p[-1:], op=dist.ReduceOp.MAX, group=self.group)
        if self.rank + 1 < self.world:  # the term the right neighbour owes my last row
            seg.bt[-1] += p[2 + nc + self.rank + 1]
        ll = -0.5 * (p[0] + p[1] + self.N * math.log(2 * math.pi))
        return dict(loglike=ll, status=p[-1].to(torch.int64), grads=_split_grads(p[2:2 + nc], self.coeffs))
