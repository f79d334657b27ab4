"""Kernel terms on the device: physical hyper-parameters -> coefficient tensors for a BATCH of
parameter sets that live on the GPU, with the adjoint maps, so that log-likelihood + gradient with respect to the physical
parameters of a batch of models never leaves the device.  Mirrors `Term.get_coefficients()` of
python/celerite2/terms.py (RealTerm 485-521, ComplexTerm 524-569, SHOTerm 604-691, Matern32Term 694-745,
RotationTerm 748-814, TermSum 210-235, TermProduct 238-300); the arithmetic is in csrc/api_terms.cu.

    k = Sum(Rotation(params_rot), Matern32(params_mat))          # params_*: (B, P) CUDA tensors
    coeffs = k.coefficients()                                     # six tensors (B, Jr) / (B, Jc) for gp.* / ops.*
    grads = k.backward(b_coeffs)                                  # [d/d params_rot, d/d params_mat]

TermDiff and TermConvolution (terms.py:304-410) are not provided.
"""
import torch

from ._lib import check, lib

KINDS = {"real": (0, 2), "complex": (1, 4), "sho": (2, 3), "matern32": (3, 2), "rotation": (4, 5)}


def _p(x):
    return None if x is None else x.data_ptr()


def _stream():
    return torch.cuda.current_stream().cuda_stream


class Term:
    """base: a term knows its widths (Jr real, Jc complex terms) and writes its slots of the kernel's coefficient arrays"""

    Jr = Jc = 0

    def coefficients(self):
        B, dev = self.batch, self.device
        out = [torch.empty((B, self.Jr), dtype=torch.float64, device=dev) for _ in range(2)] + \
              [torch.empty((B, self.Jc), dtype=torch.float64, device=dev) for _ in range(4)]
        self._write(out, self.Jr, self.Jc, 0, 0)
        return tuple(out)

    def backward(self, b_coeffs):
        """cotangents of the six coefficient tensors -> list of parameter cotangents, one (B, P) tensor per primitive term
        (in the order the primitive terms appear)"""
        b_coeffs = [v.contiguous() for v in b_coeffs]
        return self._adjoint(b_coeffs, self.Jr, self.Jc, 0, 0)

    def __add__(self, other):
        return Sum(self, other)

    def __mul__(self, other):
        return Product(self, other)


class _Primitive(Term):
    kind = None

    def __init__(self, params, eps=None, underdamped=True):
        code, P = KINDS[self.kind]
        if not (isinstance(params, torch.Tensor) and params.is_cuda and params.dtype == torch.float64 and params.dim() == 2
                and params.shape[1] == P):
            raise TypeError("%s: params must be a float64 CUDA tensor of shape (batch, %d)" % (self.kind, P))
        self.params = params.contiguous()
        self.batch, self.device = params.shape[0], params.device
        self.code, self.eps, self.under = code, float(self.default_eps if eps is None else eps), int(bool(underdamped))
        self.bad = torch.zeros(1, dtype=torch.int32, device=self.device)

    default_eps = 0.0

    def _write(self, out, Jr_tot, Jc_tot, r_off, c_off):
        check(lib.c2b_term_coefficients(self.code, self.batch, _p(self.params), self.eps, self.under, Jr_tot, Jc_tot, r_off, c_off,
                                        *[_p(v) for v in out], _p(self.bad), _stream()))

    def _adjoint(self, b, Jr_tot, Jc_tot, r_off, c_off):
        g = torch.empty_like(self.params)
        check(lib.c2b_term_coefficients_rev(self.code, self.batch, _p(self.params), self.eps, self.under, Jr_tot, Jc_tot, r_off, c_off,
                                            *[_p(v) for v in b], _p(g), _stream()))
        return [g]


class RealTerm(_Primitive):
    """params (B, 2) = (a, c): k = a exp(-c tau)"""
    kind, Jr, Jc = "real", 1, 0


class ComplexTerm(_Primitive):
    """params (B, 4) = (a, b, c, d)"""
    kind, Jr, Jc = "complex", 0, 1


class SHOTerm(_Primitive):
    """params (B, 3) = (S0, w0, Q).  underdamped=True: every Q >= 1/2 (one complex term); False: every Q < 1/2 (two real
    terms).  `bad` is set to 1 on the device if a parameter set is on the wrong side."""
    kind, default_eps = "sho", 1e-5

    def __init__(self, params, eps=None, underdamped=True):
        super().__init__(params, eps, underdamped)
        self.Jr, self.Jc = (0, 1) if underdamped else (2, 0)


class Matern32Term(_Primitive):
    """params (B, 2) = (sigma, rho)"""
    kind, Jr, Jc, default_eps = "matern32", 0, 1, 0.01


class RotationTerm(_Primitive):
    """params (B, 5) = (sigma, period, Q0, dQ, f)"""
    kind, Jr, Jc = "rotation", 0, 2


class Sum(Term):
    """TermSum (terms.py:210-235): the coefficient vectors are concatenated -- here every term writes its own slots"""

    def __init__(self, *terms):
        self.terms = []
        for t in terms:
            self.terms.extend(t.terms if isinstance(t, Sum) else [t])
        self.batch, self.device = self.terms[0].batch, self.terms[0].device
        self.Jr, self.Jc = sum(t.Jr for t in self.terms), sum(t.Jc for t in self.terms)

    def _write(self, out, Jr_tot, Jc_tot, r_off, c_off):
        for t in self.terms:
            t._write(out, Jr_tot, Jc_tot, r_off, c_off)
            r_off, c_off = r_off + t.Jr, c_off + t.Jc

    def _adjoint(self, b, Jr_tot, Jc_tot, r_off, c_off):
        g = []
        for t in self.terms:
            g += t._adjoint(b, Jr_tot, Jc_tot, r_off, c_off)
            r_off, c_off = r_off + t.Jr, c_off + t.Jc
        return g


class Product(Term):
    """TermProduct (terms.py:238-300)"""

    def __init__(self, t1, t2):
        self.t1, self.t2 = t1, t2
        self.batch, self.device = t1.batch, t1.device
        self.Jr = t1.Jr * t2.Jr
        self.Jc = t1.Jr * t2.Jc + t2.Jr * t1.Jc + 2 * t1.Jc * t2.Jc

    def _factors(self):
        self._c1, self._c2 = self.t1.coefficients(), self.t2.coefficients()
        return self._c1, self._c2

    def _write(self, out, Jr_tot, Jc_tot, r_off, c_off):
        c1, c2 = self._factors()
        mine = out
        if (Jr_tot, Jc_tot, r_off, c_off) != (self.Jr, self.Jc, 0, 0):  # inside a sum: compute apart, then copy into the slots
            mine = [torch.empty((self.batch, self.Jr), dtype=torch.float64, device=self.device) for _ in range(2)] + \
                   [torch.empty((self.batch, self.Jc), dtype=torch.float64, device=self.device) for _ in range(4)]
        check(lib.c2b_term_product(self.batch, self.t1.Jr, self.t1.Jc, *[_p(v) for v in c1], self.t2.Jr, self.t2.Jc, *[_p(v) for v in c2],
                                   *[_p(v) for v in mine], _stream()))
        if mine is not out:
            for i in range(2):
                out[i][:, r_off:r_off + self.Jr] = mine[i]
            for i in range(2, 6):
                out[i][:, c_off:c_off + self.Jc] = mine[i]

    def _adjoint(self, b, Jr_tot, Jc_tot, r_off, c_off):
        c1, c2 = self._factors()
        mine = [b[i][:, r_off:r_off + self.Jr].contiguous() for i in range(2)] + [b[i][:, c_off:c_off + self.Jc].contiguous() for i in range(2, 6)]
        g1 = [torch.empty_like(v) for v in c1]
        g2 = [torch.empty_like(v) for v in c2]
        check(lib.c2b_term_product_rev(self.batch, self.t1.Jr, self.t1.Jc, *[_p(v) for v in c1], self.t2.Jr, self.t2.Jc,
                                       *[_p(v) for v in c2], *[_p(v) for v in mine], *[_p(v) for v in g1], *[_p(v) for v in g2], _stream()))
        return self.t1.backward(g1) + self.t2.backward(g2)
