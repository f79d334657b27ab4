// Warp-cooperative versions of the summary algebras (monoids.cuh): one warp works on one combine / apply at a
// time, matrices staged in shared memory, lanes own output elements.  Used for compile-time widths (JT > 0);
// the thread-serial routines in monoids.cuh remain the run-time-width fallback.
//
// Why: the thread-serial J x J combines keep ~10 matrices per thread in local memory; with 32 lanes that is
// far more than L1 holds, so every access pays L2 latency (r01 profile: 216k cycles per tree level).  Here a
// combine is a handful of shared-memory phases (~1k cycles).
#pragma once
#include "monoids.cuh"

namespace c2b {

// out[i*ldo + j] = (Cadd ? Cadd[i*ldc + j] : 0) + sum_k opA(i,k) * opB(k,j),  i < n, j < m, k < kk
// opA(i,k) = TA ? A[k*lda + i] : A[i*lda + k];  opB(k,j) = TB ? B[j*ldb + k] : B[k*ldb + j]
template <bool TA, bool TB>
C2B_DEV void wmm(int lane, int n, int m, int kk, const double *A, int lda, const double *B, int ldb, const double *Cadd, int ldc,
                 double *out, int ldo) {
  for (int e = lane; e < n * m; e += 32) {
    const int i = e / m, j = e - i * m;
    double s    = Cadd ? Cadd[i * ldc + j] : 0.0;
    for (int k = 0; k < kk; ++k) {
      const double a = TA ? A[k * lda + i] : A[i * lda + k];
      const double b = TB ? B[j * ldb + k] : B[k * ldb + j];
      s              = fma(a, b, s);
    }
    out[i * ldo + j] = s;
  }
}

C2B_DEV void wcopy(int lane, int n, const double *src, double *dst) {
  for (int e = lane; e < n; e += 32) dst[e] = src[e];
}

// out = 0.5 (R + R^T) + Cadd   (J x J, all row-major with leading dimension J; out may be global memory)
C2B_DEV void wsym_add(int lane, int J, const double *R, const double *Cadd, double *out) {
  for (int e = lane; e < J * J; e += 32) {
    const int i = e / J, j = e - i * J;
    out[e]      = 0.5 * (R[i * J + j] + R[j * J + i]) + Cadd[e];
  }
}

// Compile-time-width versions: the passes over the J*J outputs (two for J = 6) are unrolled with a predicate, each
// output's 2 J operands are loaded before its multiply-add chain starts, so the passes and the loads overlap.  With
// one warp per scheduler at the upper tree levels every dependent instruction costs its full latency (r01i profile:
// 15-17 cycles per instruction, 2 us per level for a 6 x 6 product); this is what bounds the scan.
template <int J, bool TA, bool TB>
C2B_DEV void wmm_ct(int lane, const double *A, int lda, const double *B, int ldb, double *out, int ldo) {
  constexpr int NE = J * J, NP = (NE + 31) / 32;
  if constexpr (J <= 8) {
    double a[NP][J], b[NP][J];
#pragma unroll
    for (int p = 0; p < NP; ++p) {
      const int e = lane + 32 * p;
      const int i = e < NE ? e / J : 0, j = e < NE ? e % J : 0;
#pragma unroll
      for (int k = 0; k < J; ++k) {
        a[p][k] = TA ? A[k * lda + i] : A[i * lda + k];
        b[p][k] = TB ? B[j * ldb + k] : B[k * ldb + j];
      }
    }
#pragma unroll
    for (int p = 0; p < NP; ++p) {
      const int e = lane + 32 * p;
      double s    = 0.0;
#pragma unroll
      for (int k = 0; k < J; ++k) s = fma(a[p][k], b[p][k], s);
      if (e < NE) out[(e / J) * ldo + (e % J)] = s;
    }
  } else {
    // wide matrices: one pass at a time (the operands of all passes would not fit the registers), two accumulators
    for (int p = 0; p < NP; ++p) {
      const int e = lane + 32 * p;
      const int i = e < NE ? e / J : 0, j = e < NE ? e % J : 0;
      double s0 = 0.0, s1 = 0.0;
#pragma unroll
      for (int k = 0; k < J; k += 2) {
        s0 = fma(TA ? A[k * lda + i] : A[i * lda + k], TB ? B[j * ldb + k] : B[k * ldb + j], s0);
        if (k + 1 < J) s1 = fma(TA ? A[(k + 1) * lda + i] : A[i * lda + k + 1], TB ? B[j * ldb + k + 1] : B[(k + 1) * ldb + j], s1);
      }
      if (e < NE) out[i * ldo + j] = s0 + s1;
    }
  }
}

// out = 0.5 (R + R^T) + Cadd
template <int J>
C2B_DEV void wsym_add_ct(int lane, const double *R, const double *Cadd, double *out) {
  constexpr int NE = J * J, NP = (NE + 31) / 32;
#pragma unroll
  for (int p = 0; p < NP; ++p) {
    const int e = lane + 32 * p;
    if (e < NE) {
      const int i = e / J, j = e % J;
      out[e]      = 0.5 * (R[i * J + j] + R[j * J + i]) + Cadd[e];
    }
  }
}

// Solve Mx X = R in place (Mx: J x J, R: J x NR, both row-major in shared memory; Mx is destroyed) by Gauss-Jordan
// elimination with partial pivoting, the whole warp working on one system: per pivot step the lanes find the
// pivot row (shuffle arg-max), swap, scale the pivot row and eliminate it from all other rows, each lane owning
// elements of the augmented matrix.  Four short shared-memory phases per step instead of a long private chain.
C2B_DEV void wsolve(int lane, int J, int NR, double *Mx, double *R) {
  const int NC = J + NR;  // columns of the augmented matrix [Mx | R]
  for (int k = 0; k < J; ++k) {
    // pivot search over rows k..J-1 (J <= 32)
    double v = (lane >= k && lane < J) ? fabs(Mx[lane * J + k]) : -1.0;
    int piv  = lane;
    for (int o = 16; o > 0; o >>= 1) {
      const double v2 = __shfl_xor_sync(0xffffffffu, v, o);
      const int p2    = __shfl_xor_sync(0xffffffffu, piv, o);
      if (v2 > v || (v2 == v && p2 < piv)) {
        v   = v2;
        piv = p2;
      }
    }
    // swap rows k and piv, scale the pivot row
    const double inv = 1.0 / Mx[piv * J + k];
    __syncwarp();
    for (int cidx = lane; cidx < NC; cidx += 32) {
      double *col       = cidx < J ? Mx + cidx : R + (cidx - J);
      const int ld      = cidx < J ? J : NR;
      const double top  = col[k * ld];
      const double pivv = col[piv * ld];
      col[piv * ld]     = top;
      col[k * ld]       = pivv * inv;
    }
    __syncwarp();
    // eliminate column k from every other row (columns > k of Mx and all of R)
    const int ncols = (J - k - 1) + NR;
    for (int e = lane; e < (J - 1) * ncols; e += 32) {
      int i       = e / ncols;
      const int q = e - i * ncols;
      if (i >= k) ++i;  // skip the pivot row
      const double f = Mx[i * J + k];
      if (q < J - k - 1) {
        const int cidx = k + 1 + q;
        Mx[i * J + cidx] = fma(-f, Mx[k * J + cidx], Mx[i * J + cidx]);
      } else {
        const int cidx = q - (J - k - 1);
        R[i * NR + cidx] = fma(-f, R[k * NR + cidx], R[i * NR + cidx]);
      }
    }
    __syncwarp();
  }
}

// Compile-time-width version of wsolve with the augmented matrix [Mx | R] held in registers, lane c owning column c
// (columns c, c + 32, ... when J + NR > 32).  A pivot step is then: broadcast column k from its owner (J shuffles),
// every lane finds the pivot row itself (the same J - k comparisons on the same values), swaps two of its own
// entries, scales and eliminates -- no shared-memory phases, no barriers, no index arithmetic.  About 4x fewer
// instructions than wsolve for J = 6 (r01f: the J x J solve was ~3/4 of a 2,000-instruction combine).
template <int J, int NR>
C2B_DEV void wsolve_reg(int lane, const double *Mx, double *R) {
  constexpr int NC   = J + NR;
  constexpr int NCOL = (NC + 31) / 32;
  double col[NCOL][J];
#pragma unroll
  for (int q = 0; q < NCOL; ++q) {
    const int c = lane + 32 * q;
#pragma unroll
    for (int i = 0; i < J; ++i) col[q][i] = (c < J) ? Mx[i * J + c] : (c < NC ? R[i * NR + (c - J)] : 0.0);
  }
#pragma unroll
  for (int k = 0; k < J; ++k) {
    // column k lives in slot k / 32 of lane k % 32
    double f[J];
#pragma unroll
    for (int i = 0; i < J; ++i) f[i] = __shfl_sync(0xffffffffu, col[k / 32][i], k % 32);
    int piv     = k;
    double best = fabs(f[k]);
#pragma unroll
    for (int i = k + 1; i < J; ++i) {
      const double v = fabs(f[i]);
      if (v > best) {
        best = v;
        piv  = i;
      }
    }
    // swap rows k and piv (of the multipliers and of my columns)
    double pv = f[k];
#pragma unroll
    for (int i = k + 1; i < J; ++i)
      if (i == piv) {
        pv   = f[i];
        f[i] = f[k];
      }
    const double inv = 1.0 / pv;
#pragma unroll
    for (int q = 0; q < NCOL; ++q) {
      double top = col[q][k];
#pragma unroll
      for (int i = k + 1; i < J; ++i)
        if (i == piv) {
          const double x = col[q][i];
          col[q][i]      = top;
          top            = x;
        }
      top *= inv;
      col[q][k] = top;
#pragma unroll
      for (int i = 0; i < J; ++i)
        if (i != k) col[q][i] = fma(-f[i], top, col[q][i]);
    }
  }
#pragma unroll
  for (int q = 0; q < NCOL; ++q) {
    const int c = lane + 32 * q;
    if (c >= J && c < NC) {
#pragma unroll
      for (int i = 0; i < J; ++i) R[i * NR + (c - J)] = col[q][i];
    }
  }
  __syncwarp();
}

template <class Mon>
struct Coop;  // wsmem(J,K) doubles per warp; wcombine / wapply mirror Mon::combine / Mon::apply

// The operands of wcombine / wapply live in shared memory (block-resident trees of scan.cuh) and are read in place;
// `sm` is the warp's private scratch for intermediate products only.
template <int JT>
struct Coop<AffV<JT>> {
  __host__ __device__ static int wsmem(int, int) { return 0; }
  C2B_DEV static void wcombine(int lane, int J, int K, const double *e1, const double *e2, double *out, double *) {
    const int JJ = J * J;
    wmm_ct<JT, false, false>(lane, e2, J, e1, J, out, J);  // M = M2 M1
    // b = M2 b1 + b2 : b is J x K with element (j,k) at [j + J*k]
    const double *b1 = e1 + JJ, *b2 = e2 + JJ;
    double *b = out + JJ;
    for (int e = lane; e < J * K; e += 32) {
      const int k = e / J, i = e - k * J;
      double s    = b2[e];
      for (int j = 0; j < J; ++j) s = fma(e2[i * J + j], b1[j + J * k], s);
      b[e] = s;
    }
    __syncwarp();
  }
  C2B_DEV static void wapply(int lane, int J, int K, const double *e, const double *sin, double *sout, double *) {
    const double *b = e + J * J;
    for (int x = lane; x < J * K; x += 32) {
      const int k = x / J, i = x - k * J;
      double s    = b[x];
      for (int j = 0; j < J; ++j) s = fma(e[i * J + j], sin[j + J * k], s);
      sout[x] = s;
    }
    __syncwarp();
  }
};

template <int JT>
struct Coop<DiagV<JT>> {
  __host__ __device__ static int wsmem(int, int) { return 0; }
  C2B_DEV static void wcombine(int lane, int J, int K, const double *e1, const double *e2, double *out, double *) {
    for (int j = lane; j < J; j += 32) out[j] = e2[j] * e1[j];
    for (int x = lane; x < J * K; x += 32) {
      const int j = x % J;
      out[J + x]  = fma(e2[j], e1[J + x], e2[J + x]);
    }
    __syncwarp();
  }
  C2B_DEV static void wapply(int lane, int J, int K, const double *e, const double *sin, double *sout, double *) {
    for (int x = lane; x < J * K; x += 32) {
      const int j = x % J;
      sout[x]     = fma(e[j], sin[x], e[J + x]);
    }
    __syncwarp();
  }
};

template <int JT>
struct Coop<Cong<JT>> {
  __host__ __device__ static int wsmem(int J, int) { return 2 * J * J; }
  C2B_DEV static void wcombine(int lane, int J, int, const double *e1, const double *e2, double *out, double *sm) {
    const int JJ = J * J;
    double *T = sm, *R = sm + JJ;
    wmm_ct<JT, false, false>(lane, e1, J, e2, J, out, J);     // M = M1 M2
    wmm_ct<JT, false, false>(lane, e1 + JJ, J, e2, J, T, J);  // T = B1 M2
    __syncwarp();
    wmm_ct<JT, true, false>(lane, e2, J, T, J, R, J);  // R = M2^T T
    __syncwarp();
    wsym_add_ct<JT>(lane, R, e2 + JJ, out + JJ);
    __syncwarp();
  }
  C2B_DEV static void wapply(int lane, int J, int, const double *e, const double *sin, double *sout, double *sm) {
    const int JJ = J * J;
    double *T = sm, *R = sm + JJ;
    wmm_ct<JT, false, false>(lane, sin, J, e, J, T, J);  // T = G M
    __syncwarp();
    wmm_ct<JT, true, false>(lane, e, J, T, J, R, J);  // R = M^T T
    __syncwarp();
    wsym_add_ct<JT>(lane, R, e + JJ, sout);
    __syncwarp();
  }
};

template <int JT>
struct Coop<Lft<JT>> {
  __host__ __device__ static int wsmem(int J, int) { return 7 * J * J; }
  C2B_DEV static void wcombine(int lane, int J, int, const double *e1, const double *e2, double *out, double *sm) {
    const int JJ = J * J;
    const double *A1 = e1, *C1 = e1 + JJ, *A2 = e2, *Q2 = e2 + 2 * JJ;
    double *MX = sm;
    double *R  = sm + JJ;  // J x 2J : [A1 | C1] -> [X1 | X2]
    double *T2 = sm + 3 * JJ, *T3 = sm + 4 * JJ, *RC = sm + 5 * JJ, *RQ = sm + 6 * JJ;
    // MX = I + C1 Q2 ; R = [A1 | C1]
    for (int e = lane; e < JJ; e += 32) {
      const int i = e / J, j = e - i * J;
      double s    = (i == j) ? 1.0 : 0.0;
      for (int k = 0; k < J; ++k) s = fma(C1[i * J + k], Q2[k * J + j], s);
      MX[e]                = s;
      R[i * 2 * J + j]     = A1[e];
      R[i * 2 * J + J + j] = C1[e];
    }
    __syncwarp();
    wsolve_reg<JT, 2 * JT>(lane, MX, R);
    const double *X1 = R, *X2 = R + J;  // leading dimension 2J
    wmm_ct<JT, false, false>(lane, A2, J, X1, 2 * J, out, J);  // A = A2 X1
    wmm_ct<JT, false, false>(lane, A2, J, X2, 2 * J, T2, J);   // T2 = A2 X2
    wmm_ct<JT, false, false>(lane, Q2, J, X1, 2 * J, T3, J);   // T3 = Q2 X1
    __syncwarp();
    wmm_ct<JT, false, true>(lane, T2, J, A2, J, RC, J);  // RC = T2 A2^T
    wmm_ct<JT, true, false>(lane, A1, J, T3, J, RQ, J);  // RQ = A1^T T3
    __syncwarp();
    wsym_add_ct<JT>(lane, RC, e2 + JJ, out + JJ);
    wsym_add_ct<JT>(lane, RQ, e1 + 2 * JJ, out + 2 * JJ);
    __syncwarp();
  }
  C2B_DEV static void wapply(int lane, int J, int, const double *e, const double *sin, double *sout, double *sm) {
    const int JJ = J * J;
    const double *A = e, *Q = e + 2 * JJ, *X = sin;
    double *MX = sm, *Z = sm + JJ, *T = sm + 2 * JJ, *R = sm + 3 * JJ;
    for (int x = lane; x < JJ; x += 32) {
      const int i = x / J, j = x - i * J;
      double s    = (i == j) ? 1.0 : 0.0;
      for (int k = 0; k < J; ++k) s = fma(X[i * J + k], Q[k * J + j], s);
      MX[x] = s;
      Z[x]  = X[x];
    }
    __syncwarp();
    wsolve_reg<JT, JT>(lane, MX, Z);  // Z = (I + X Q)^-1 X
    wmm_ct<JT, false, false>(lane, A, J, Z, J, T, J);
    __syncwarp();
    wmm_ct<JT, false, true>(lane, T, J, A, J, R, J);
    __syncwarp();
    wsym_add_ct<JT>(lane, R, e + JJ, sout);
    __syncwarp();
  }
};

}  // namespace c2b
