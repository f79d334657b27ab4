// The four associative "chunk summary" algebras used to carry state across chunk boundaries.
//
// Every recursion on the hot path maps the state entering a chunk to the state leaving it by a map from a
// family that is closed under composition:
//
//   AffV  vector-affine      G -> M G + b            solves (J x J transition), state J x K
//   DiagV diagonal-affine    G -> m o G + b          matmuls (diagonal transition), state J x K
//   Cong  congruence-affine  G -> M^T G M + B        factor_rev, state J x J
//   Lft   linear-fractional  X -> A X (I + Q X)^-1 A^T + C    factor (Riccati step), state J x J (X = -Sigma)
//
// Each struct gives the element size, the state size, combine(first, second) = "first then second" and
// apply(element, state).  All routines are thread-serial on pointers into global scratch; J is a template
// constant (JT) or run-time (JT == 0).
#pragma once
#include "common.cuh"

// ---- tiny dense helpers (row-major n x n unless noted) ------------------------------------------------------
template <int JT>
struct Mat {
  static constexpr int MJ = Width<JT>::MAXJ;
};

// out = A * B  (all J x J row-major); out must not alias A or B
template <int JT>
C2B_DEV void mm(int J, const double *A, const double *B, double *out) {
  C2B_UNROLL_J
  for (int i = 0; i < J; ++i) {
    C2B_UNROLL_J
    for (int j = 0; j < J; ++j) {
      double s = 0.0;
      C2B_UNROLL_J
      for (int k = 0; k < J; ++k) s = fma(A[i * J + k], B[k * J + j], s);
      out[i * J + j] = s;
    }
  }
}
// out = A^T * B
template <int JT>
C2B_DEV void mtm(int J, const double *A, const double *B, double *out) {
  C2B_UNROLL_J
  for (int i = 0; i < J; ++i) {
    C2B_UNROLL_J
    for (int j = 0; j < J; ++j) {
      double s = 0.0;
      C2B_UNROLL_J
      for (int k = 0; k < J; ++k) s = fma(A[k * J + i], B[k * J + j], s);
      out[i * J + j] = s;
    }
  }
}
// out = A * B^T
template <int JT>
C2B_DEV void mmt(int J, const double *A, const double *B, double *out) {
  C2B_UNROLL_J
  for (int i = 0; i < J; ++i) {
    C2B_UNROLL_J
    for (int j = 0; j < J; ++j) {
      double s = 0.0;
      C2B_UNROLL_J
      for (int k = 0; k < J; ++k) s = fma(A[i * J + k], B[j * J + k], s);
      out[i * J + j] = s;
    }
  }
}

// Solve Mx * X = R in place (R is J x NR row-major, NR = NRJ * J) by Gaussian elimination with partial
// pivoting.  Mx is destroyed.  Row swaps are done with selects so that register arrays keep static indices.
template <int JT, int NRJ>
C2B_DEV void ge_solve(int J, double *Mx, double *R) {
  const int NR = NRJ * J;
  constexpr int NRU = (JT > 0 && JT <= 8) ? NRJ * JT : 1;  // unroll factor for loops over the right-hand sides
  C2B_UNROLL_J
  for (int k = 0; k < J; ++k) {
    // pivot search
    int piv     = k;
    double best = fabs(Mx[k * J + k]);
    C2B_UNROLL_J
    for (int i = 0; i < J; ++i) {
      if (i > k) {
        double v = fabs(Mx[i * J + k]);
        if (v > best) {
          best = v;
          piv  = i;
        }
      }
    }
    // swap rows k and piv (predicated, static indices)
    C2B_UNROLL_J
    for (int i = 0; i < J; ++i) {
      if (i > k && i == piv) {
        C2B_UNROLL_J
        for (int j = 0; j < J; ++j) {
          double tmp    = Mx[k * J + j];
          Mx[k * J + j] = Mx[i * J + j];
          Mx[i * J + j] = tmp;
        }
C2B_PRAGMA(unroll NRU)
        for (int j = 0; j < NR; ++j) {
          double tmp   = R[k * NR + j];
          R[k * NR + j] = R[i * NR + j];
          R[i * NR + j] = tmp;
        }
      }
    }
    const double inv = 1.0 / Mx[k * J + k];
    C2B_UNROLL_J
    for (int j = 0; j < J; ++j) Mx[k * J + j] *= inv;
C2B_PRAGMA(unroll NRU)
    for (int j = 0; j < NR; ++j) R[k * NR + j] *= inv;
    C2B_UNROLL_J
    for (int i = 0; i < J; ++i) {
      if (i != k) {
        const double f = Mx[i * J + k];
        C2B_UNROLL_J
        for (int j = 0; j < J; ++j) Mx[i * J + j] = fma(-f, Mx[k * J + j], Mx[i * J + j]);
C2B_PRAGMA(unroll NRU)
        for (int j = 0; j < NR; ++j) R[i * NR + j] = fma(-f, R[k * NR + j], R[i * NR + j]);
      }
    }
  }
}

// ---- AffV ----------------------------------------------------------------------------------------------------
// element: M (J*J, row-major) then b (J*K, index j + J*k).  state: G (J*K, index j + J*k).
template <int JT>
struct AffV {
  static constexpr bool kCoop = JT > 0;  // warp-cooperative scan (coop.cuh) for compile-time widths
  static constexpr int kJT    = JT;
  static constexpr int MJ = Width<JT>::MAXJ;
  __host__ __device__ static int esize(int J, int K) { return J * J + J * K; }
  __host__ __device__ static int ssize(int J, int K) { return J * K; }
  // element = head (shared by all right-hand sides) + ecol doubles per column; state = shead + scol per column
  __host__ __device__ static int ehead(int J) { return J * J; }
  __host__ __device__ static int ecol(int J) { return J; }
  __host__ __device__ static int shead(int) { return 0; }
  __host__ __device__ static int scol(int J) { return J; }
  C2B_DEV static void combine(int J, int K, const double *e1, const double *e2, double *out) {
    double M1[MJ * MJ], M2[MJ * MJ], M[MJ * MJ];
    for (int i = 0; i < J * J; ++i) {
      M1[i] = e1[i];
      M2[i] = e2[i];
    }
    mm<JT>(J, M2, M1, M);
    for (int i = 0; i < J * J; ++i) out[i] = M[i];
    const double *b1 = e1 + J * J, *b2 = e2 + J * J;
    double *b = out + J * J;
    for (int k = 0; k < K; ++k) {
      double x[MJ];
      C2B_UNROLL_J
      for (int j = 0; j < J; ++j) x[j] = b1[j + J * k];
      C2B_UNROLL_J
      for (int i = 0; i < J; ++i) {
        double s = b2[i + J * k];
        C2B_UNROLL_J
        for (int j = 0; j < J; ++j) s = fma(M2[i * J + j], x[j], s);
        b[i + J * k] = s;
      }
    }
  }
  C2B_DEV static void apply(int J, int K, const double *e, const double *sin, double *sout) {
    double M[MJ * MJ];
    for (int i = 0; i < J * J; ++i) M[i] = e[i];
    const double *b = e + J * J;
    for (int k = 0; k < K; ++k) {
      double x[MJ];
      C2B_UNROLL_J
      for (int j = 0; j < J; ++j) x[j] = sin[j + J * k];
      C2B_UNROLL_J
      for (int i = 0; i < J; ++i) {
        double s = b[i + J * k];
        C2B_UNROLL_J
        for (int j = 0; j < J; ++j) s = fma(M[i * J + j], x[j], s);
        sout[i + J * k] = s;
      }
    }
  }
};

// ---- DiagV ---------------------------------------------------------------------------------------------------
// element: m (J) then b (J*K).  state: G (J*K).
template <int JT>
struct DiagV {
  static constexpr bool kCoop = JT > 0;  // warp-cooperative scan (coop.cuh) for compile-time widths
  static constexpr int kJT    = JT;
  __host__ __device__ static int esize(int J, int K) { return J + J * K; }
  __host__ __device__ static int ssize(int J, int K) { return J * K; }
  __host__ __device__ static int ehead(int J) { return J; }
  __host__ __device__ static int ecol(int J) { return J; }
  __host__ __device__ static int shead(int) { return 0; }
  __host__ __device__ static int scol(int J) { return J; }
  C2B_DEV static void combine(int J, int K, const double *e1, const double *e2, double *out) {
    for (int j = 0; j < J; ++j) out[j] = e2[j] * e1[j];
    for (int k = 0; k < K; ++k)
      for (int j = 0; j < J; ++j) out[J + j + J * k] = fma(e2[j], e1[J + j + J * k], e2[J + j + J * k]);
  }
  C2B_DEV static void apply(int J, int K, const double *e, const double *sin, double *sout) {
    for (int k = 0; k < K; ++k)
      for (int j = 0; j < J; ++j) sout[j + J * k] = fma(e[j], sin[j + J * k], e[J + j + J * k]);
  }
};

// ---- Cong ----------------------------------------------------------------------------------------------------
// element: M (J*J row-major) then B (J*J row-major, symmetric).  state: G (J*J row-major, symmetric).
// apply: M^T G M + B.   combine(first, second): (M1 M2, M2^T B1 M2 + B2).
template <int JT>
struct Cong {
  static constexpr bool kCoop = JT > 0;  // warp-cooperative scan (coop.cuh) for compile-time widths
  static constexpr int kJT    = JT;
  static constexpr int MJ = Width<JT>::MAXJ;
  __host__ __device__ static int esize(int J, int) { return 2 * J * J; }
  __host__ __device__ static int ssize(int J, int) { return J * J; }
  __host__ __device__ static int ehead(int J) { return 2 * J * J; }
  __host__ __device__ static int ecol(int) { return 0; }
  __host__ __device__ static int shead(int J) { return J * J; }
  __host__ __device__ static int scol(int) { return 0; }
  C2B_DEV static void sandwich(int J, const double *M, const double *G, const double *B, double *out) {
    double T[MJ * MJ], R[MJ * MJ];
    mm<JT>(J, G, M, T);    // G M
    mtm<JT>(J, M, T, R);   // M^T G M
    C2B_UNROLL_J
    for (int i = 0; i < J; ++i) {
      C2B_UNROLL_J
      for (int j = 0; j < J; ++j) {
        if (j >= i) {
          const double v = 0.5 * (R[i * J + j] + R[j * J + i]) + B[i * J + j];
          out[i * J + j]  = v;
          out[j * J + i]  = v;
        }
      }
    }
  }
  C2B_DEV static void combine(int J, int, const double *e1, const double *e2, double *out) {
    double M1[MJ * MJ], M2[MJ * MJ], X[MJ * MJ], B2[MJ * MJ];
    for (int i = 0; i < J * J; ++i) {
      M1[i] = e1[i];
      M2[i] = e2[i];
      X[i]  = e1[J * J + i];
      B2[i] = e2[J * J + i];
    }
    double Bn[MJ * MJ];
    sandwich(J, M2, X, B2, Bn);
    mm<JT>(J, M1, M2, X);
    for (int i = 0; i < J * J; ++i) {
      out[i]         = X[i];
      out[J * J + i] = Bn[i];
    }
  }
  C2B_DEV static void apply(int J, int, const double *e, const double *sin, double *sout) {
    double M[MJ * MJ], G[MJ * MJ], B[MJ * MJ], R[MJ * MJ];
    for (int i = 0; i < J * J; ++i) {
      M[i] = e[i];
      B[i] = e[J * J + i];
      G[i] = sin[i];
    }
    sandwich(J, M, G, B, R);
    for (int i = 0; i < J * J; ++i) sout[i] = R[i];
  }
};

// ---- Lft -----------------------------------------------------------------------------------------------------
// element: A, C, Q (each J*J row-major; C and Q symmetric).  state: X (J*J row-major, symmetric), X = -Sigma.
// apply:   X -> A Z A^T + C,  Z = (I + X Q)^-1 X
// combine: Minv = (I + C1 Q2)^-1;  A = A2 Minv A1;  C = A2 Minv C1 A2^T + C2;  Q = A1^T Q2 Minv A1 + Q1
template <int JT>
struct Lft {
  static constexpr bool kCoop = JT > 0;  // warp-cooperative scan (coop.cuh) for compile-time widths
  static constexpr int kJT    = JT;
  static constexpr int MJ = Width<JT>::MAXJ;
  __host__ __device__ static int esize(int J, int) { return 3 * J * J; }
  __host__ __device__ static int ssize(int J, int) { return J * J; }
  __host__ __device__ static int ehead(int J) { return 3 * J * J; }
  __host__ __device__ static int ecol(int) { return 0; }
  __host__ __device__ static int shead(int J) { return J * J; }
  __host__ __device__ static int scol(int) { return 0; }
  C2B_DEV static void combine(int J, int, const double *e1, const double *e2, double *out) {
    const int JJ = J * J;
    double Mx[MJ * MJ], R[2 * MJ * MJ], T[MJ * MJ], T2[MJ * MJ], A2[MJ * MJ];
    // Mx = I + C1 Q2
    {
      double C1[MJ * MJ], Q2[MJ * MJ];
      for (int i = 0; i < JJ; ++i) {
        C1[i] = e1[JJ + i];
        Q2[i] = e2[2 * JJ + i];
      }
      mm<JT>(J, C1, Q2, Mx);
      C2B_UNROLL_J
      for (int i = 0; i < J; ++i) Mx[i * J + i] += 1.0;
      // R = [A1 | C1]  (J x 2J)
      C2B_UNROLL_J
      for (int i = 0; i < J; ++i) {
        C2B_UNROLL_J
        for (int j = 0; j < J; ++j) {
          R[i * 2 * J + j]     = e1[i * J + j];
          R[i * 2 * J + J + j] = C1[i * J + j];
        }
      }
    }
    ge_solve<JT, 2>(J, Mx, R);  // R = [X1 | X2] = Minv [A1 | C1]
    double X1[MJ * MJ], X2[MJ * MJ];
    C2B_UNROLL_J
    for (int i = 0; i < J; ++i) {
      C2B_UNROLL_J
      for (int j = 0; j < J; ++j) {
        X1[i * J + j] = R[i * 2 * J + j];
        X2[i * J + j] = R[i * 2 * J + J + j];
      }
    }
    for (int i = 0; i < JJ; ++i) A2[i] = e2[i];
    // A = A2 X1
    mm<JT>(J, A2, X1, T);
    for (int i = 0; i < JJ; ++i) out[i] = T[i];
    // C = A2 X2 A2^T + C2  (symmetrised)
    mm<JT>(J, A2, X2, T);
    mmt<JT>(J, T, A2, T2);
    C2B_UNROLL_J
    for (int i = 0; i < J; ++i) {
      C2B_UNROLL_J
      for (int j = 0; j < J; ++j) {
        if (j >= i) {
          const double v    = 0.5 * (T2[i * J + j] + T2[j * J + i]) + e2[JJ + i * J + j];
          out[JJ + i * J + j] = v;
          out[JJ + j * J + i] = v;
        }
      }
    }
    // Q = A1^T (Q2 X1) + Q1  (symmetrised)
    {
      double Q2[MJ * MJ], A1[MJ * MJ];
      for (int i = 0; i < JJ; ++i) {
        Q2[i] = e2[2 * JJ + i];
        A1[i] = e1[i];
      }
      mm<JT>(J, Q2, X1, T);
      mtm<JT>(J, A1, T, T2);
    }
    C2B_UNROLL_J
    for (int i = 0; i < J; ++i) {
      C2B_UNROLL_J
      for (int j = 0; j < J; ++j) {
        if (j >= i) {
          const double v        = 0.5 * (T2[i * J + j] + T2[j * J + i]) + e1[2 * JJ + i * J + j];
          out[2 * JJ + i * J + j] = v;
          out[2 * JJ + j * J + i] = v;
        }
      }
    }
  }
  C2B_DEV static void apply(int J, int, const double *e, const double *sin, double *sout) {
    const int JJ = J * J;
    double Mx[MJ * MJ], Z[MJ * MJ], T[MJ * MJ], T2[MJ * MJ], A[MJ * MJ];
    {
      double X[MJ * MJ], Q[MJ * MJ];
      for (int i = 0; i < JJ; ++i) {
        X[i] = sin[i];
        Q[i] = e[2 * JJ + i];
        Z[i] = sin[i];
      }
      mm<JT>(J, X, Q, Mx);
      C2B_UNROLL_J
      for (int i = 0; i < J; ++i) Mx[i * J + i] += 1.0;
    }
    ge_solve<JT, 1>(J, Mx, Z);  // Z = (I + X Q)^-1 X
    for (int i = 0; i < JJ; ++i) A[i] = e[i];
    mm<JT>(J, A, Z, T);
    mmt<JT>(J, T, A, T2);
    C2B_UNROLL_J
    for (int i = 0; i < J; ++i) {
      C2B_UNROLL_J
      for (int j = 0; j < J; ++j) {
        if (j >= i) {
          const double v = 0.5 * (T2[i * J + j] + T2[j * J + i]) + e[JJ + i * J + j];
          sout[i * J + j] = v;
          sout[j * J + i] = v;
        }
      }
    }
  }
};
