"""Test utility: numpy restatement of the chunk-local recursions and their summaries (the algorithm of
celerite2_b200/csrc), used to emulate "one rank = one row range" on CPU.  Not product code."""
import numpy as np


def pvec(t, c, n):
    return np.exp(c * (t[n - 1] - t[n])) if n > 0 else np.ones_like(c)


def factor_chunk(t, c, a, U, V, n0, n1, Sig):
    """rows n0..n1-1 of factor from the carried-in Sigma -> (d, W, S rows, Sigma_out)"""
    J = len(c)
    d = np.zeros(n1 - n0)
    W = np.zeros((n1 - n0, J))
    S = np.zeros((n1 - n0, J * J))
    Sig = Sig.copy()
    for n in range(n0, n1):
        p = pvec(t, c, n)
        St = p[:, None] * Sig
        S[n - n0] = St.T.reshape(-1)
        Sn = St * p[None, :]
        tmp = U[n] @ Sn
        dn = a[n] - tmp @ U[n]
        w = (V[n] - tmp) / dn
        Sig = Sn + dn * np.outer(w, w)
        d[n - n0] = dn
        W[n - n0] = w
    return d, W, S, Sig


def factor_summary(t, c, a, U, V, n0, n1):
    """(A, C, Q) of rows n0..n1-1, flat [A | C | Q] row-major"""
    J = len(c)
    A, C, Q = np.eye(J), np.zeros((J, J)), np.zeros((J, J))
    for n in range(n0, n1):
        p, u, v = pvec(t, c, n), U[n], V[n]
        A = p[:, None] * A
        C = p[:, None] * C * p[None, :]
        q = C @ u
        gam = a[n] + u @ q
        r = A.T @ u
        g = (q + v) / gam
        A = A - np.outer(g, r)
        C = C - gam * np.outer(g, g)
        Q = Q + np.outer(r, r) / gam
    return np.concatenate([A.reshape(-1), C.reshape(-1), Q.reshape(-1)])


def solve_lower_chunk(t, c, U, W, Y, n0, n1, G):
    """rows n0..n1-1 of solve_lower from the carried-in state G (J,K) -> (Z rows, F rows, G_out)"""
    J, K = len(c), Y.shape[1]
    G = G.copy()
    Z = np.zeros((n1 - n0, K))
    F = np.zeros((n1 - n0, J * K))
    for n in range(n0, n1):
        p = pvec(t, c, n)
        F[n - n0] = G.T.reshape(-1)
        Fp = p[:, None] * G
        z = Y[n] - U[n] @ Fp
        G = Fp + np.outer(W[n], z)
        Z[n - n0] = z
    return Z, F, G


def solve_lower_summary(t, c, U, W, Y, n0, n1):
    """flat [M | b] of rows n0..n1-1"""
    J, K = len(c), Y.shape[1]
    _, _, M = solve_lower_chunk(t, c, U, W, np.zeros((len(t), J)), n0, n1, np.eye(J))
    _, _, b = solve_lower_chunk(t, c, U, W, Y, n0, n1, np.zeros((J, K)))
    return np.concatenate([M.reshape(-1), b.T.reshape(-1)])
