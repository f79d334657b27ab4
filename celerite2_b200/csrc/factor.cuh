// factor (LDL^T of the semiseparable matrix) and factor_rev.
//
// Reference: celerite2::core::factor, c++/include/celerite2/forward.hpp:69-135, and factor_rev,
// c++/include/celerite2/reverse.hpp:10-85.
//
// Chunk-local form of factor.  Carry entering row n: Sigma_n = S_{n-1} + d_{n-1} W_{n-1}^T W_{n-1} (symmetric,
// zero for n = 0), i.e. the reference's Sn just after forward.hpp:115.  Row n then does exactly
// forward.hpp:116-131:
//     St = diag(p) Sigma         -> workspace row n (column-major)
//     S  = St diag(p)
//     tmp = U[n] S ; d[n] = a[n] - tmp.U[n] ; W[n] = (V[n] - tmp)/d[n] ; Sigma' = S + d[n] W[n]^T W[n]
// (p = 1 for n = 0, where Sigma = 0 anyway).
//
// The map Sigma_{n0} -> Sigma_{n1} of a chunk is linear-fractional in X = -Sigma:
//     X' = A X (I + Q X)^-1 A^T + C.
// Folding one row (u, v, a, p) into an accumulated (A, C, Q) needs no matrix inverse (Sherman-Morrison):
//     A <- P A ; C <- P C P ; q = C u ; gam = a + u.q ; r = A^T u ; g = (q + v)/gam
//     A <- A - g r^T ; C <- C - gam g g^T ; Q <- Q + r r^T / gam
// gam is the pivot the chunk would see if nothing preceded it; it bounds the true pivot from above, so it is
// positive whenever the true factorisation succeeds.
//
// Mapping: one thread per (series, chunk); the J x J state lives in registers (JT <= 8) or local memory.
#pragma once
#include "args.cuh"
#include "scan.cuh"

namespace c2b {


// Row fields staged through RowPipe (factor apply and summary): 0 t[n], 1 t[n-1], 2 a[n], 3.. U row, 3+J.. V row.
constexpr int kFactorSlots = 4;
template <int JT>
struct FactorRow {
  static constexpr int MJ   = Width<JT>::MAXJ;
  static constexpr int ROWD = 3 + 2 * MJ;
};

template <int JT, class RP>
C2B_DEV void factor_fetch_row(RP &rp, int J, int64_t n, const double *t, const double *a, const double *U, const double *V) {
  constexpr int MJ = Width<JT>::MAXJ;
  rp.put(0, t + n);
  if (n > 0) rp.put(1, t + n - 1);
  rp.put(2, a + n);
  C2B_UNROLL_J
  for (int j = 0; j < J; ++j) {
    rp.put(3 + j, U + n * J + j);
    rp.put(3 + MJ + j, V + n * J + j);
  }
}

template <int JT>
__global__ void factor_apply_kernel(FactorArgs g) {
  pdl_prologue();
  extern __shared__ double c2b_ring[];
  constexpr int MJ = Width<JT>::MAXJ;
  const int J      = JT ? JT : g.J;
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= g.batch * g.C) return;
  const int64_t chunk = gid % g.C, b = gid / g.C;
  int64_t n0 = chunk * g.L, n1 = n0 + g.L < g.N ? n0 + g.L : g.N;
  const double *t = g.t + b * g.t_bs;
  const double *c = g.c + b * g.c_bs;
  const int64_t NJ = g.N * J;
  const double *a = g.a + b * g.N, *U = g.U + b * NJ, *V = g.V + b * NJ;
  double *d = g.d + b * g.N, *W = g.W + b * NJ;
  double *S = g.S ? g.S + b * NJ * J : nullptr;
  double cj[MJ], p[MJ], u[MJ], tmp[MJ], w[MJ];
  unsigned dup[MJ];
  double Sg[MJ * MJ];  // Sg[j + J*i] = element (j, i); symmetric
  C2B_UNROLL_J
  for (int j = 0; j < J; ++j) cj[j] = c[j];
  dup_masks<JT>(J, cj, dup);
  if (g.carry) {
    // scan state is X = -Sigma (row-major, symmetric)
    const double *st = g.carry + (b * g.node_stride + chunk) * (int64_t)(J * J);
    for (int i = 0; i < J * J; ++i) Sg[i] = -st[i];
  } else {
    for (int i = 0; i < J * J; ++i) Sg[i] = 0.0;
  }
  RowPipe<(JT > 0), FactorRow<JT>::ROWD, kFactorSlots> rp;
  rp.init(c2b_ring);
  const int64_t cnt = n1 - n0;
  for (int pq = 0; pq < rp.D; ++pq) {
    const bool f = pq < cnt;
    if (f) factor_fetch_row<JT>(rp, J, n0 + pq, t, a, U, V);
    rp.commit(f);
  }
  for (int64_t n = n0; n < n1; ++n) {
    const bool f = n + rp.D < n1;
    if (f) factor_fetch_row<JT>(rp, J, n + rp.D, t, a, U, V);
    rp.commit(f);
    rp.acquire();
    if (n == 0) {
      C2B_UNROLL_J
      for (int j = 0; j < J; ++j) p[j] = 1.0;
    } else {
      compute_p<JT>(J, cj, dup, rp.get(1) - rp.get(0), p);
    }
    C2B_UNROLL_J
    for (int j = 0; j < J; ++j) u[j] = rp.get(3 + j);
    double *Sw = S ? S + n * (int64_t)J * J : nullptr;
    C2B_UNROLL_J
    for (int i = 0; i < J; ++i) {
      C2B_UNROLL_J
      for (int j = 0; j < J; ++j) {
        const double st = p[j] * Sg[j + J * i];
        if (Sw) Sw[j + J * i] = st;
        Sg[j + J * i] = st * p[i];
      }
    }
    double acc = 0.0;
    C2B_UNROLL_J
    for (int i = 0; i < J; ++i) {
      double s = 0.0;
      C2B_UNROLL_J
      for (int j = 0; j < J; ++j) s = fma(u[j], Sg[j + J * i], s);
      tmp[i] = s;
      acc    = fma(s, u[i], acc);
    }
    const double dn = rp.get(2) - acc;
    d[n]            = dn;
    if (n >= 1 && dn <= 0.0) {
      record_failure(g.status ? g.status + b : nullptr, n);
      cp_async_wait<0>();
      return;  // forward.hpp:128: later rows of this chunk stay untouched
    }
    C2B_UNROLL_J
    for (int j = 0; j < J; ++j) {
      w[j]         = (rp.get(3 + MJ + j) - tmp[j]) / dn;
      W[n * J + j] = w[j];
    }
    C2B_UNROLL_J
    for (int i = 0; i < J; ++i) {
      C2B_UNROLL_J
      for (int j = 0; j < J; ++j) Sg[j + J * i] = fma(dn * w[j], w[i], Sg[j + J * i]);
    }
    rp.release();
  }
}

template <int JT>
__global__ void factor_summary_kernel(FactorArgs g) {
  pdl_prologue();
  extern __shared__ double c2b_ring[];
  constexpr int MJ = Width<JT>::MAXJ;
  const int J      = JT ? JT : g.J;
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= g.batch * g.C) return;
  const int64_t chunk = gid % g.C, b = gid / g.C;
  int64_t n0 = chunk * g.L, n1 = n0 + g.L < g.N ? n0 + g.L : g.N;
  const double *t = g.t + b * g.t_bs;
  const double *c = g.c + b * g.c_bs;
  const int64_t NJ = g.N * J;
  const double *a = g.a + b * g.N, *U = g.U + b * NJ, *V = g.V + b * NJ;
  double cj[MJ], p[MJ], u[MJ], q[MJ], r[MJ], gv[MJ];
  unsigned dup[MJ];
  double A[MJ * MJ], Cm[MJ * MJ], Q[MJ * MJ];  // row-major
  C2B_UNROLL_J
  for (int j = 0; j < J; ++j) cj[j] = c[j];
  dup_masks<JT>(J, cj, dup);
  C2B_UNROLL_J
  for (int i = 0; i < J; ++i) {
    C2B_UNROLL_J
    for (int j = 0; j < J; ++j) {
      A[i * J + j]  = (i == j) ? 1.0 : 0.0;
      Cm[i * J + j] = 0.0;
      Q[i * J + j]  = 0.0;
    }
  }
  RowPipe<(JT > 0), FactorRow<JT>::ROWD, kFactorSlots> rp;
  rp.init(c2b_ring);
  const int64_t cnt = n1 - n0;
  for (int pq = 0; pq < rp.D; ++pq) {
    const bool f = pq < cnt;
    if (f) factor_fetch_row<JT>(rp, J, n0 + pq, t, a, U, V);
    rp.commit(f);
  }
  for (int64_t n = n0; n < n1; ++n) {
    const bool f = n + rp.D < n1;
    if (f) factor_fetch_row<JT>(rp, J, n + rp.D, t, a, U, V);
    rp.commit(f);
    rp.acquire();
    if (n == 0) {
      C2B_UNROLL_J
      for (int j = 0; j < J; ++j) p[j] = 1.0;
    } else {
      compute_p<JT>(J, cj, dup, rp.get(1) - rp.get(0), p);
    }
    C2B_UNROLL_J
    for (int j = 0; j < J; ++j) u[j] = rp.get(3 + j);
    // A <- P A ; C <- P C P ; q = C u ; r = A^T u
    C2B_UNROLL_J
    for (int j = 0; j < J; ++j) r[j] = 0.0;
    double uq = 0.0;
    C2B_UNROLL_J
    for (int i = 0; i < J; ++i) {
      double s = 0.0;
      C2B_UNROLL_J
      for (int j = 0; j < J; ++j) {
        A[i * J + j] *= p[i];
        Cm[i * J + j] *= p[i] * p[j];
        s    = fma(Cm[i * J + j], u[j], s);
        r[j] = fma(A[i * J + j], u[i], r[j]);
      }
      q[i] = s;
      uq   = fma(u[i], s, uq);
    }
    const double gam  = rp.get(2) + uq;
    const double igam = 1.0 / gam;
    C2B_UNROLL_J
    for (int j = 0; j < J; ++j) gv[j] = (q[j] + rp.get(3 + MJ + j)) * igam;
    C2B_UNROLL_J
    for (int i = 0; i < J; ++i) {
      const double ggi = gam * gv[i], rgi = r[i] * igam;
      C2B_UNROLL_J
      for (int j = 0; j < J; ++j) {
        A[i * J + j]  = fma(-gv[i], r[j], A[i * J + j]);
        Cm[i * J + j] = fma(-ggi, gv[j], Cm[i * J + j]);
        Q[i * J + j]  = fma(rgi, r[j], Q[i * J + j]);
      }
    }
    rp.release();
  }
  double *e = g.leaves + (b * g.node_stride + chunk) * (int64_t)(3 * J * J);
  for (int i = 0; i < J * J; ++i) {
    e[i]             = A[i];
    e[J * J + i]     = Cm[i];
    e[2 * J * J + i] = Q[i];
  }
}

// =================================================================================================================
// factor_rev.  Chunk-local form with G = sym(bS) as the carry entering row n (bS after the reference has
// finished row n+1, reverse.hpp:77; zero for n = N-1).  Only the symmetric part of bS reaches the outputs.
//   b0 = bW[n]/d[n] ; wG = W[n] G ; bV[n] = b0 + 2 wG ; ba[n] = bd[n] + wG.W[n] - W[n].bV[n]
//   n >= 1:  Sn = workspace row n ; x = bV[n] + 2 ba[n] U[n] ; bU[n] = -(x Sn) o p
//            y = bV[n] + ba[n] U[n] ; G <- G - sym(U[n]^T y)
//            bp = 2 p o colsum(Sn o G) ; bc += dt bp ; fac = c.bp ; bt[n] -= fac ; bt[n-1] += fac
//            G <- P G P
//   n == 0:  bU[0] = 0
// which is reverse.hpp:58-84 with the "Step 3" additions to row n-1 moved to the top of row n-1's own turn.
// The chunk map is G -> M^T G M + B with M the same transition as solve_lower's (M = prod (I - W^T U) P).
// =================================================================================================================

// Row fields staged through RowPipe: 0 t[n], 1 t[n-1], 2 d[n], 3 bd[n], 4.. U row, 4+J.. W row, 4+2J.. bW row,
// 4+3J.. S row (J*J, only for n >= 1).
constexpr int kFactorRevSlots = 3;
template <int JT>
struct FactorRevRow {
  static constexpr int MJ   = Width<JT>::MAXJ;
  static constexpr int ROWD = 4 + 3 * MJ + MJ * MJ;
};

// WITHM: additionally right-multiply the J rows of Mr (J x J row-major, initially the identity) by every row's
// transition T_n = (I - W[n]^T U[n]) P_n, n >= 1: the chunk's congruence matrix, built in the same pass as B.
template <int JT, bool WRITE, bool WITHM = false>
C2B_DEV void factor_rev_rows(int J, int64_t N, int64_t n0, int64_t n1, const double *t, const double *cj, const unsigned *dup,
                             const double *U, const double *d, const double *W, const double *S, const double *bd, const double *bW,
                             double *bt, double *ba, double *bU, double *bV, double *G, double *bc_acc, double *edge_out, double *smem,
                             double *Mr = nullptr, bool acc = false) {
  constexpr int MJ = Width<JT>::MAXJ;
  RowPipe<(JT > 0), FactorRevRow<JT>::ROWD, kFactorRevSlots> rp;
  rp.init(smem);
  double p[MJ], u[MJ], w[MJ], bv[MJ], x[MJ], y[MJ];
  double pend = 0.0;
  const int64_t cnt = n1 - n0;
  auto fetch = [&](int64_t i) {
    const int64_t n = n1 - 1 - i;
    rp.put(0, t + n);
    if (n >= 1) rp.put(1, t + n - 1);
    rp.put(2, d + n);
    rp.put(3, bd + n);
    C2B_UNROLL_J
    for (int j = 0; j < J; ++j) {
      rp.put(4 + j, U + n * J + j);
      rp.put(4 + MJ + j, W + n * J + j);
      rp.put(4 + 2 * MJ + j, bW + n * J + j);
    }
    if (n >= 1) {
      const double *Sn = S + n * (int64_t)J * J;
      for (int e = 0; e < J * J; ++e) rp.put(4 + 3 * MJ + e, Sn + e);
    }
  };
  for (int pq = 0; pq < rp.D; ++pq) {
    const bool f = pq < cnt;
    if (f) fetch(pq);
    rp.commit(f);
  }
  for (int64_t i = 0; i < cnt; ++i) {
    const bool f = i + rp.D < cnt;
    if (f) fetch(i + rp.D);
    rp.commit(f);
    rp.acquire();
    const int64_t n    = n1 - 1 - i;
    const double inv_d = 1.0 / rp.get(2);
    C2B_UNROLL_J
    for (int j = 0; j < J; ++j) {
      w[j] = rp.get(4 + MJ + j);
      u[j] = rp.get(4 + j);
    }
    double wGw = 0.0, wbv = 0.0;
    C2B_UNROLL_J
    for (int k = 0; k < J; ++k) {
      double s = 0.0;
      C2B_UNROLL_J
      for (int j = 0; j < J; ++j) s = fma(w[j], G[j * J + k], s);
      wGw   = fma(s, w[k], wGw);
      bv[k] = fma(2.0, s, rp.get(4 + 2 * MJ + k) * inv_d);
      wbv   = fma(w[k], bv[k], wbv);
    }
    const double ban = rp.get(3) + wGw - wbv;
    if (WRITE) {
      ba[n] = ban;
      C2B_UNROLL_J
      for (int j = 0; j < J; ++j) bV[n * J + j] = bv[j];
    }
    if (n >= 1) {
      const double dt = rp.get(1) - rp.get(0);
      compute_p<JT>(J, cj, dup, dt, p);
      if (WITHM) {
        C2B_UNROLL_J
        for (int k = 0; k < J; ++k) {
          double s = 0.0;
          C2B_UNROLL_J
          for (int j = 0; j < J; ++j) s = fma(Mr[k * J + j], w[j], s);
          C2B_UNROLL_J
          for (int j = 0; j < J; ++j) Mr[k * J + j] = fma(-s, u[j], Mr[k * J + j]) * p[j];
        }
      }
      C2B_UNROLL_J
      for (int j = 0; j < J; ++j) {
        x[j] = fma(2.0 * ban, u[j], bv[j]);
        y[j] = fma(ban, u[j], bv[j]);
      }
      C2B_UNROLL_J
      for (int k = 0; k < J; ++k) {
        C2B_UNROLL_J
        for (int j = 0; j < J; ++j) G[k * J + j] -= 0.5 * (u[k] * y[j] + u[j] * y[k]);
      }
      double fac = 0.0;
      C2B_UNROLL_J
      for (int k = 0; k < J; ++k) {
        double xs = 0.0, gs = 0.0;
        C2B_UNROLL_J
        for (int j = 0; j < J; ++j) {
          const double sjk = rp.get(4 + 3 * MJ + j + J * k);  // Sn(j,k) at [j + J*k]
          xs               = fma(x[j], sjk, xs);
          gs               = fma(sjk, G[j * J + k], gs);
        }
        if (WRITE) {
          out_store(bU + n * J + k, -(xs * p[k]), acc);
          const double bp = 2.0 * p[k] * gs;
          bc_acc[k]       = fma(dt, bp, bc_acc[k]);
          fac             = fma(cj[k], bp, fac);
        }
      }
      C2B_UNROLL_J
      for (int k = 0; k < J; ++k) {
        C2B_UNROLL_J
        for (int j = 0; j < J; ++j) G[k * J + j] *= p[k] * p[j];
      }
      if (WRITE) {
        out_store(bt + n, pend - fac, acc);
        pend = fac;
      }
    } else if (WRITE) {
      if (!acc) {
        C2B_UNROLL_J
        for (int j = 0; j < J; ++j) bU[j] = 0.0;
      }
      out_store(bt, pend, acc);
      pend = 0.0;
    }
    rp.release();
  }
  if (WRITE) *edge_out = pend;
}

template <int JT>
__global__ void factor_rev_apply_kernel(FactorRevArgs g) {
  pdl_prologue();
  extern __shared__ double c2b_ring[];
  constexpr int MJ = Width<JT>::MAXJ;
  const int J      = JT ? JT : g.J;
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= g.batch * g.C) return;
  const int64_t chunk = gid % g.C, b = gid / g.C;
  int64_t n0 = chunk * g.L, n1 = n0 + g.L < g.N ? n0 + g.L : g.N;
  const double *t = g.t + b * g.t_bs;
  const double *c = g.c + b * g.c_bs;
  const int64_t NJ = g.N * J;
  double cj[MJ], bc_acc[MJ], G[MJ * MJ];
  unsigned dup[MJ];
  C2B_UNROLL_J
  for (int j = 0; j < J; ++j) {
    cj[j]     = c[j];
    bc_acc[j] = 0.0;
  }
  dup_masks<JT>(J, cj, dup);
  if (g.carry) {
    const int64_t leaf = g.C - 1 - chunk;
    const double *st   = g.carry + (b * g.node_stride + leaf) * (int64_t)(J * J);
    for (int i = 0; i < J * J; ++i) G[i] = st[i];
  } else {
    for (int i = 0; i < J * J; ++i) G[i] = 0.0;
  }
  double edge = 0.0;
  factor_rev_rows<JT, true>(J, g.N, n0, n1, t, cj, dup, g.U + b * NJ, g.d + b * g.N, g.W + b * NJ, g.S + b * NJ * J, g.bd + b * g.N,
                            g.bW + b * NJ, g.bt + b * g.N, g.ba + b * g.N, g.bU + b * NJ, g.bV + b * NJ, G, bc_acc, &edge, c2b_ring, nullptr,
                            g.acc != 0);
  double *bcp = g.bc_part + (b * g.C + chunk) * J;
  C2B_UNROLL_J
  for (int j = 0; j < J; ++j) bcp[j] = bc_acc[j];
  g.bt_edge[b * g.C + chunk] = edge;
}

// summary: thread per (series, chunk, col), col in [0, J]:
//   col < J : row `col` of M = T_{n1-1} ... T_{n0'} (n0' = max(n0, 1)), T_n = (I - W[n]^T U[n]) P_n, built by
//             right-multiplying the unit row vector e_col in descending n:  r <- ((r - (r.W^T) U) o p)
//   col == J: B = the chunk run from G = 0 without outputs
template <int JT>
__global__ void factor_rev_summary_kernel(FactorRevArgs g) {
  pdl_prologue();
  extern __shared__ double c2b_ring[];
  constexpr int MJ = Width<JT>::MAXJ;
  const int J      = JT ? JT : g.J;
  const int ncol   = J + 1;
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= g.batch * g.C * ncol) return;
  const int col       = (int)(gid % ncol);
  const int64_t chunk = (gid / ncol) % g.C;
  const int64_t b     = gid / ((int64_t)ncol * g.C);
  int64_t n0 = chunk * g.L, n1 = n0 + g.L < g.N ? n0 + g.L : g.N;
  const double *t = g.t + b * g.t_bs;
  const double *c = g.c + b * g.c_bs;
  const int64_t NJ = g.N * J;
  double cj[MJ];
  unsigned dup[MJ];
  C2B_UNROLL_J
  for (int j = 0; j < J; ++j) cj[j] = c[j];
  dup_masks<JT>(J, cj, dup);
  const int64_t leaf = g.C - 1 - chunk;
  double *e          = g.leaves + (b * g.node_stride + leaf) * (int64_t)(2 * J * J);
  if (col < J) {
    double r[MJ], p[MJ];
    const double *U = g.U + b * NJ, *W = g.W + b * NJ;
    C2B_UNROLL_J
    for (int j = 0; j < J; ++j) r[j] = (j == col) ? 1.0 : 0.0;
    RowPipe<(JT > 0), FactorRevRow<JT>::ROWD, kFactorRevSlots> rp;
    rp.init(c2b_ring);
    const int64_t lo  = n0 >= 1 ? n0 : 1;  // row 0 applies no transition
    const int64_t cnt = n1 - lo;
    auto fetch = [&](int64_t i) {
      const int64_t n = n1 - 1 - i;
      rp.put(0, t + n);
      rp.put(1, t + n - 1);
      C2B_UNROLL_J
      for (int j = 0; j < J; ++j) {
        rp.put(4 + j, U + n * J + j);
        rp.put(4 + MJ + j, W + n * J + j);
      }
    };
    for (int pq = 0; pq < rp.D; ++pq) {
      const bool f = pq < cnt;
      if (f) fetch(pq);
      rp.commit(f);
    }
    for (int64_t i = 0; i < cnt; ++i) {
      const bool f = i + rp.D < cnt;
      if (f) fetch(i + rp.D);
      rp.commit(f);
      rp.acquire();
      compute_p<JT>(J, cj, dup, rp.get(1) - rp.get(0), p);
      double s = 0.0;
      C2B_UNROLL_J
      for (int j = 0; j < J; ++j) s = fma(r[j], rp.get(4 + MJ + j), s);
      C2B_UNROLL_J
      for (int j = 0; j < J; ++j) r[j] = fma(-s, rp.get(4 + j), r[j]) * p[j];
      rp.release();
    }
    C2B_UNROLL_J
    for (int j = 0; j < J; ++j) e[col * J + j] = r[j];
  } else {
    double G[MJ * MJ];
    for (int i = 0; i < J * J; ++i) G[i] = 0.0;
    factor_rev_rows<JT, false>(J, g.N, n0, n1, t, cj, dup, g.U + b * NJ, g.d + b * g.N, g.W + b * NJ, g.S + b * NJ * J, g.bd + b * g.N,
                               g.bW + b * NJ, nullptr, nullptr, nullptr, nullptr, G, nullptr, nullptr, c2b_ring);
    for (int i = 0; i < J * J; ++i) e[J * J + i] = G[i];
  }
}

// summary, one thread per (series, chunk): M and B in a single pass over the rows
template <int JT>
__global__ void factor_rev_summary1_kernel(FactorRevArgs g) {
  pdl_prologue();
  extern __shared__ double c2b_ring[];
  constexpr int MJ = Width<JT>::MAXJ;
  const int J      = JT ? JT : g.J;
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= g.batch * g.C) return;
  const int64_t chunk = gid % g.C, b = gid / g.C;
  int64_t n0 = chunk * g.L, n1 = n0 + g.L < g.N ? n0 + g.L : g.N;
  const double *t = g.t + b * g.t_bs;
  const double *c = g.c + b * g.c_bs;
  const int64_t NJ = g.N * J;
  double cj[MJ], G[MJ * MJ], Mr[MJ * MJ];
  unsigned dup[MJ];
  C2B_UNROLL_J
  for (int j = 0; j < J; ++j) cj[j] = c[j];
  dup_masks<JT>(J, cj, dup);
  C2B_UNROLL_J
  for (int i = 0; i < J; ++i) {
    C2B_UNROLL_J
    for (int j = 0; j < J; ++j) {
      G[i * J + j]  = 0.0;
      Mr[i * J + j] = (i == j) ? 1.0 : 0.0;
    }
  }
  factor_rev_rows<JT, false, true>(J, g.N, n0, n1, t, cj, dup, g.U + b * NJ, g.d + b * g.N, g.W + b * NJ, g.S + b * NJ * J, g.bd + b * g.N,
                                   g.bW + b * NJ, nullptr, nullptr, nullptr, nullptr, G, nullptr, nullptr, c2b_ring, Mr);
  const int64_t leaf = g.C - 1 - chunk;
  double *e          = g.leaves + (b * g.node_stride + leaf) * (int64_t)(2 * J * J);
  for (int i = 0; i < J * J; ++i) {
    e[i]         = Mr[i];
    e[J * J + i] = G[i];
  }
}

}  // namespace c2b
