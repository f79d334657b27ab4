"""GPU: the device-side term algebra (celerite2_b200.terms -> c2b_term_coefficients / c2b_term_product) against golden
coefficient vectors produced by the reference's own python/celerite2/terms.py (tests/golden/make_golden_terms.py), and the
adjoint maps against central differences of the device forward map (FP64, relative step 1e-6)."""
import os

import numpy as np
import pytest

from helpers import GOLDEN

pytestmark = pytest.mark.gpu
NAMES = ("ar", "cr", "ac", "bc", "cc", "dc")


@pytest.fixture(scope="module")
def env():
    import torch

    from celerite2_b200 import terms

    dev = torch.device("cuda:0")
    T = lambda x: torch.as_tensor(np.ascontiguousarray(x, dtype=np.float64), device=dev)
    g = np.load(os.path.join(GOLDEN, "terms_golden.npz"))
    return terms, T, g, torch


def _build(terms, T, case, P):
    """-> (term, list of parameter tensors of its primitive terms in order)"""
    if case == "real":
        return terms.RealTerm(T(P))
    if case == "complex":
        return terms.ComplexTerm(T(P))
    if case == "sho_under":
        return terms.SHOTerm(T(P))
    if case == "sho_over":
        return terms.SHOTerm(T(P), underdamped=False)
    if case == "matern32":
        return terms.Matern32Term(T(P))
    if case == "rotation":
        return terms.RotationTerm(T(P))
    if case == "rot_plus_mat":
        return terms.RotationTerm(T(P[:, :5])) + terms.Matern32Term(T(P[:, 5:]))
    if case == "sho_times_real":
        return terms.SHOTerm(T(P[:, :3])) * terms.RealTerm(T(P[:, 3:]))
    if case == "rot_times_sho":
        return terms.RotationTerm(T(P[:, :5])) * terms.SHOTerm(T(P[:, 5:]))
    if case == "over_times_under_plus_real":
        return terms.SHOTerm(T(P[:, :3]), underdamped=False) * terms.SHOTerm(T(P[:, 3:6])) + terms.RealTerm(T(P[:, 6:]))
    raise KeyError(case)


CASES = ["real", "complex", "sho_under", "sho_over", "matern32", "rotation", "rot_plus_mat", "sho_times_real", "rot_times_sho",
         "over_times_under_plus_real"]


@pytest.mark.parametrize("case", CASES)
def test_coefficients_against_reference_terms(env, case):
    terms, T, g, torch = env
    P = g[case + "_params"]
    co = _build(terms, T, case, P).coefficients()
    for nm, got in zip(NAMES, co):
        want = g["%s_%s" % (case, nm)]
        got = got.cpu().numpy()
        assert got.shape == want.shape, (nm, got.shape, want.shape)
        if want.size:
            assert np.max(np.abs(got - want) / np.maximum(np.abs(want), 1e-300)) < 1e-12, nm


@pytest.mark.parametrize("case", CASES)
def test_adjoint_against_central_differences(env, case):
    terms, T, g, torch = env
    P = g[case + "_params"].copy()
    if case == "sho_under":
        P = P[1:]  # the first set sits inside the eps floor of f(Q), where the map has a kink
    term = _build(terms, T, case, P)
    co = term.coefficients()
    rng = np.random.default_rng(8)
    bco = [T(rng.standard_normal(tuple(v.shape))) for v in co]
    grads = torch.cat(term.backward(bco), dim=1).cpu().numpy()  # primitive terms in order = columns of P in order
    assert grads.shape == P.shape

    def f(Pn):
        """per parameter set: <coefficients(Pn), bco>"""
        c = _build(terms, T, case, Pn).coefficients()
        tot = torch.zeros(Pn.shape[0], dtype=torch.float64, device=bco[0].device)
        for a, b in zip(c, bco):
            if a.shape[1]:
                tot = tot + (a * b).sum(dim=1)
        return tot.cpu().numpy()

    for j in range(P.shape[1]):
        h = 1e-6 * np.maximum(np.abs(P[:, j]), 1e-3)
        Pp, Pm = P.copy(), P.copy()
        Pp[:, j] += h
        Pm[:, j] -= h
        fd = (f(Pp) - f(Pm)) / (2 * h)
        scale = np.maximum(np.abs(fd), np.max(np.abs(grads), axis=1))
        assert np.max(np.abs(fd - grads[:, j]) / scale) < 2e-6, (case, j, fd, grads[:, j])


def test_wrong_side_of_critical_damping_is_flagged(env):
    terms, T, g, torch = env
    P = g["sho_over_params"]
    t = terms.SHOTerm(T(P), underdamped=True)  # Q < 1/2 declared underdamped
    t.coefficients()
    torch.cuda.synchronize()
    assert int(t.bad.item()) == 1
