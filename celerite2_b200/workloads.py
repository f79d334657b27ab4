"""Synthetic workloads: kernel coefficient sets and light-curve-shaped data.

Host-side O(J) algebra only -- the reference does this part in Python as well
(python/celerite2/terms.py).  Each helper returns the six coefficient vectors
``(ar, cr, ac, bc, cc, dc)`` the reference's ``Term.get_coefficients()`` returns, which is the input of
the a/U/V assembly (python/celerite2/terms.py:117-177, c++/include/celerite2/terms.hpp:112-139).

The data generators follow SURVEY.md section 8(d) (which mirrors python/celerite2/testing.py:18-31):
t = sort(U(0, 0.1 N)), yerr ~ U(0.1, 0.3), diag = yerr^2, y = sin(t) + yerr * N(0, 1).
"""
import numpy as np

_E = np.empty(0)


def real_coeffs(a, c):
    """RealTerm, c++/include/celerite2/terms.hpp:181-198"""
    return (np.array([float(a)]), np.array([float(c)]), _E, _E, _E, _E)


def complex_coeffs(a, b, c, d):
    """ComplexTerm, c++/include/celerite2/terms.hpp:209-228"""
    return (_E, _E, np.array([float(a)]), np.array([float(b)]), np.array([float(c)]), np.array([float(d)]))


def sho_coeffs(S0, w0, Q, eps=1e-5):
    """SHOTerm, python/celerite2/terms.py:658-691 (same formulas as terms.hpp:239-277)"""
    if Q < 0.5:
        f = np.sqrt(max(1.0 - 4.0 * Q**2, eps))
        return (
            0.5 * S0 * w0 * Q * np.array([1.0 + 1.0 / f, 1.0 - 1.0 / f]),
            0.5 * w0 / Q * np.array([1.0 - f, 1.0 + f]),
            _E, _E, _E, _E,
        )
    f = np.sqrt(max(4.0 * Q**2 - 1.0, eps))
    a = S0 * w0 * Q
    c = 0.5 * w0 / Q
    return (_E, _E, np.array([a]), np.array([a / f]), np.array([c]), np.array([c * f]))


def matern32_coeffs(sigma, rho, eps=0.01):
    """Matern32Term, python/celerite2/terms.py:733-745"""
    w0 = np.sqrt(3.0) / rho
    S0 = sigma**2 / w0
    return (_E, _E, np.array([w0 * S0]), np.array([w0**2 * S0 / eps]), np.array([w0]), np.array([eps]))


def rotation_coeffs(sigma, period, Q0, dQ, f):
    """RotationTerm = sum of two SHO terms, python/celerite2/terms.py:792-814"""
    amp = sigma**2 / (1 + f)
    Q1 = 0.5 + Q0 + dQ
    w1 = 4 * np.pi * Q1 / (period * np.sqrt(4 * Q1**2 - 1))
    S1 = amp / (w1 * Q1)
    Q2 = 0.5 + Q0
    w2 = 8 * np.pi * Q2 / (period * np.sqrt(4 * Q2**2 - 1))
    S2 = f * amp / (w2 * Q2)
    return sum_coeffs(sho_coeffs(S1, w1, Q1), sho_coeffs(S2, w2, Q2))


def sum_coeffs(*terms):
    """TermSum: concatenation of the coefficient vectors, python/celerite2/terms.py:229-235"""
    return tuple(np.concatenate([np.atleast_1d(t[i]) for t in terms]) for i in range(6))


def width(coeffs):
    return len(coeffs[0]) + 2 * len(coeffs[2])


def light_curve(N, seed):
    """-> (t, y, diag) as described in the module docstring."""
    rng = np.random.default_rng(seed)
    t = np.sort(rng.uniform(0.0, 0.1 * N, N))
    yerr = rng.uniform(0.1, 0.3, N)
    y = np.sin(t) + yerr * rng.standard_normal(N)
    return t, y, yerr**2


# ---- the five BASELINE.json configurations ---------------------------------------

def config_c1():
    """SHOTerm J=2, N=1,000, single series"""
    return dict(name="c1", coeffs=sho_coeffs(1.0, 1.0, 2.0), N=1000, seed=0, batch=1, K=1)


def config_c2(N=1_000_000):
    """RotationTerm + Matern32Term, J=6, N=1e6, single long series, log-likelihood + gradient"""
    co = sum_coeffs(rotation_coeffs(1.5, 3.45, 1.3, 1.05, 0.5), matern32_coeffs(1.0, 2.0))
    return dict(name="c2", coeffs=co, N=N, seed=1, batch=1, K=1)


def config_c3(batch=65536, N=10_000):
    """2 x underdamped SHOTerm per series with log-normal (sigma 0.1) parameters, J=4"""
    rng = np.random.default_rng(3)
    base = np.array([[1.0, 1.0, 2.0], [0.5, 3.0, 5.0]])
    pars = base[None] * np.exp(0.1 * rng.standard_normal((batch, 2, 3)))
    co = [sum_coeffs(sho_coeffs(*pars[b, 0]), sho_coeffs(*pars[b, 1])) for b in range(batch)]
    stacked = tuple(np.stack([c[i] for c in co]) for i in range(6))
    return dict(name="c3", coeffs=stacked, N=N, seed=3, batch=batch, K=1)


def config_c4(N=200_000, K=64):
    """8 x SHOTerm(S0=1/(i+1), w0=0.5+0.7i, Q=1+0.5i), J=16, multi-RHS solves"""
    co = sum_coeffs(*[sho_coeffs(1.0 / (i + 1), 0.5 + 0.7 * i, 1.0 + 0.5 * i) for i in range(8)])
    return dict(name="c4", coeffs=co, N=N, seed=4, batch=1, K=K)


def config_c5(batch=4096, N=50_000):
    """RotationTerm with log-normal(0.1) draws about the reference's test parameters, shared t,y, J=4"""
    rng = np.random.default_rng(5)
    base = np.array([1.5, 3.45, 1.3, 1.05, 0.5])
    pars = base[None] * np.exp(0.1 * rng.standard_normal((batch, 5)))
    co = [rotation_coeffs(*pars[b]) for b in range(batch)]
    stacked = tuple(np.stack([c[i] for c in co]) for i in range(6))
    return dict(name="c5", coeffs=stacked, N=N, seed=5, batch=batch, K=1)
