// Linear sweeps: solve_lower / solve_upper / matmul_lower / matmul_upper and their reverse-mode counterparts.
//
// Reference: internal::forward / backward (c++/include/celerite2/internal.hpp:105-189) and
// internal::forward_rev / backward_rev (internal.hpp:191-303).  The reference keeps the running state F
// *between* rows; here every row's update is written "chunk-local": the carry entering row n is the value the
// reference stores in the workspace row n (the state before the multiplication by P_n), so a chunk [n0, n1)
// can start from any carried-in state and rows n0..n1-1 are then computed exactly as the reference does.
//
//   forward (lower), per right-hand-side column:     backward (upper): same with n descending,
//     p  = exp(c (t[n-1]-t[n]))   (n = 0: p = 1)       p = exp(c (t[n]-t[n+1])), roles of U and W/V swapped
//     [F row n = G]
//     Fp = p o G
//     solve : z = y[n] - A[n].Fp ;  G = Fp + B[n] z
//     matmul: z = z[n] + A[n].Fp ;  G = Fp + B[n] y[n]
//   with (A, B) = (U, W or V) for lower and (W or V, U) for upper.
//
// Mapping: one thread per (series, chunk, right-hand-side column); the J-vector state lives in registers.
#pragma once
#include "args.cuh"
#include "scan.cuh"

namespace c2b {


// Runs rows of one chunk for one column.  WRITE = false is the summary mode (no Z / F stores).
template <int JT, bool SOLVE, bool LOWER, bool WRITE>
C2B_DEV void sweep_rows(int J, int K, int64_t N, int64_t n0, int64_t n1, const double *t, const double *cj, const unsigned *dup,
                        const double *A, const double *B, const double *Ycol, double *Zcol, double *Fcol, double *G) {
  constexpr int MJ = Width<JT>::MAXJ;
  double p[MJ];
  const int64_t cnt = n1 - n0;
  for (int64_t i = 0; i < cnt; ++i) {
    const int64_t n = LOWER ? n0 + i : n1 - 1 - i;
    bool edge;
    double dt;
    if (LOWER) {
      edge = (n == 0);
      dt   = edge ? 0.0 : t[n - 1] - t[n];
    } else {
      edge = (n == N - 1);
      dt   = edge ? 0.0 : t[n] - t[n + 1];
    }
    if (edge) {
      C2B_UNROLL_J
      for (int j = 0; j < J; ++j) p[j] = 1.0;
    } else {
      compute_p<JT>(J, cj, dup, dt, p);
    }
    if (WRITE && Fcol) {
      C2B_UNROLL_J
      for (int j = 0; j < J; ++j) Fcol[n * (int64_t)J * K + j] = G[j];
    }
    const double *a = A + n * J, *b = B + n * J;
    double s = 0.0;
    C2B_UNROLL_J
    for (int j = 0; j < J; ++j) {
      G[j] = p[j] * G[j];
      s    = fma(a[j], G[j], s);
    }
    const double y = Ycol ? Ycol[n * (int64_t)K] : 0.0;
    double z, r;
    if (SOLVE) {
      z = y - s;
      r = z;
    } else {
      z = (WRITE ? Zcol[n * (int64_t)K] : 0.0) + s;
      r = y;
    }
    if (WRITE) Zcol[n * (int64_t)K] = z;
    C2B_UNROLL_J
    for (int j = 0; j < J; ++j) G[j] = fma(b[j], r, G[j]);
  }
}

C2B_DEV void chunk_range(int64_t N, int64_t L, int64_t chunk, int64_t *n0, int64_t *n1) {
  *n0 = chunk * L;
  int64_t e = *n0 + L;
  *n1 = e < N ? e : N;
}

// ---- apply: thread per (series, chunk, column) ---------------------------------------------------------------
template <int JT, bool SOLVE, bool LOWER>
__global__ void sweep_apply_kernel(SweepArgs a) {
  constexpr int MJ = Width<JT>::MAXJ;
  const int J      = JT ? JT : a.J;
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= a.batch * a.C * a.K) return;
  const int col       = (int)(gid % a.K);
  const int64_t chunk = (gid / a.K) % a.C;
  const int64_t b     = gid / (a.K * a.C);
  int64_t n0, n1;
  chunk_range(a.N, a.L, chunk, &n0, &n1);
  const double *t = a.t + b * a.t_bs;
  const double *c = a.c + b * a.c_bs;
  double cj[MJ], G[MJ];
  unsigned dup[MJ];
  C2B_UNROLL_J
  for (int j = 0; j < J; ++j) cj[j] = c[j];
  dup_masks<JT>(J, cj, dup);
  if (a.carry) {
    const int64_t leaf = LOWER ? chunk : a.C - 1 - chunk;
    const double *st   = a.carry + (b * a.node_stride + leaf) * (int64_t)(J * a.K) + (int64_t)J * col;
    C2B_UNROLL_J
    for (int j = 0; j < J; ++j) G[j] = st[j];
  } else {
    C2B_UNROLL_J
    for (int j = 0; j < J; ++j) G[j] = 0.0;
  }
  const int64_t NJ = a.N * J, NK = a.N * a.K;
  sweep_rows<JT, SOLVE, LOWER, true>(J, a.K, a.N, n0, n1, t, cj, dup, a.A + b * NJ, a.B + b * NJ, a.Y + b * a.y_bs + col,
                                     a.Z + b * NK + col, a.F ? a.F + b * NK * J + (int64_t)J * col : nullptr, G);
}

// ---- summary: thread per (series, chunk, column'), column' in [0, J+K) for solves, [0, K) for matmuls ---------
//   solve  : columns 0..J-1 propagate the unit vectors (-> the chunk's J x J transition M),
//            columns J..J+K-1 propagate zero with the real right-hand sides (-> b)
//   matmul : the transition is diag(prod p) = exp(c * (t at chunk entry - t at chunk exit)); only b is swept
template <int JT, bool SOLVE, bool LOWER>
__global__ void sweep_summary_kernel(SweepArgs a) {
  constexpr int MJ = Width<JT>::MAXJ;
  const int J      = JT ? JT : a.J;
  const int ncol   = SOLVE ? J + a.K : a.K;
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= a.batch * a.C * ncol) return;
  const int col       = (int)(gid % ncol);
  const int64_t chunk = (gid / ncol) % a.C;
  const int64_t b     = gid / ((int64_t)ncol * a.C);
  int64_t n0, n1;
  chunk_range(a.N, a.L, chunk, &n0, &n1);
  const double *t = a.t + b * a.t_bs;
  const double *c = a.c + b * a.c_bs;
  double cj[MJ], G[MJ];
  unsigned dup[MJ];
  C2B_UNROLL_J
  for (int j = 0; j < J; ++j) cj[j] = c[j];
  dup_masks<JT>(J, cj, dup);
  const int64_t leaf = LOWER ? chunk : a.C - 1 - chunk;
  const int es       = SOLVE ? J * J + J * a.K : J + J * a.K;
  double *e          = a.leaves + (b * a.node_stride + leaf) * (int64_t)es;
  const int64_t NJ   = a.N * J;
  const bool basis   = SOLVE && col < J;
  const int k        = SOLVE ? col - J : col;
  C2B_UNROLL_J
  for (int j = 0; j < J; ++j) G[j] = (basis && j == col) ? 1.0 : 0.0;
  sweep_rows<JT, SOLVE, LOWER, false>(J, a.K, a.N, n0, n1, t, cj, dup, a.A + b * NJ, a.B + b * NJ,
                                      basis ? nullptr : a.Y + b * a.y_bs + k, nullptr, nullptr, G);
  if (basis) {
    C2B_UNROLL_J
    for (int j = 0; j < J; ++j) e[j * J + col] = G[j];  // column `col` of M (row-major)
  } else {
    double *bb = e + (SOLVE ? J * J : J);
    C2B_UNROLL_J
    for (int j = 0; j < J; ++j) bb[j + J * k] = G[j];
    if (!SOLVE && k == 0) {
      // diagonal transition of the chunk
      double dt;
      if (LOWER)
        dt = t[n0 > 0 ? n0 - 1 : 0] - t[n1 - 1];
      else
        dt = t[n0] - t[n1 < a.N ? n1 : a.N - 1];
      C2B_UNROLL_J
      for (int j = 0; j < J; ++j) e[j] = exp(cj[j] * dt);
    }
  }
}

// =================================================================================================================
// Reverse mode.  Chunk-local form (internal.hpp:226-245 / 283-302 rearranged so that the carry entering row n
// is bF after the reference has finished row n+1 (lower) / n-1 (upper)):
//   solve : bz = bZ[n] + B[n].bF ;  bB[n] = z[n] bF ;  bY[n] = bz
//   matmul: bz = bZ[n]           ;  bB[n] = y[n] bF ;  bY[n] = B[n].bF
//   if row n has a neighbour o (n-1 for lower, n+1 for upper):
//     dt = s (t[o]-t[n]) (s = +1 lower, -1 upper), p = exp(c dt), Fn = workspace row n
//     bA[n] = -/+ bz (p o Fn) ;  bF' = bF -/+ A[n] bz ;  bp = Fn o bF' o p
//     bc += dt bp ; fac = c.bp ; bt[n] -= s fac ; bt[o] += s fac ;  bF = p o bF'
//   else bA[n] = 0.
// Cotangents that sum over right-hand-side columns (bt, bc, bA, bB) are accumulated column by column.
// =================================================================================================================

// MODE 0: full (all outputs).  MODE 1: summary with real cotangents, no outputs.  MODE 2: homogeneous (bZ = 0).
template <int JT, bool SOLVE, bool LOWER, int MODE>
C2B_DEV void sweep_rev_rows(int J, int K, int col, bool first_col, int64_t N, int64_t n0, int64_t n1, const double *t, const double *cj,
                            const unsigned *dup, const double *A, const double *B, const double *Y, const double *Z, const double *F,
                            const double *bZ, double *bt, double *bA, double *bB, double *bY, double *bF, double *bc_acc,
                            double *edge_out) {
  constexpr int MJ = Width<JT>::MAXJ;
  double p[MJ];
  const double sgn = LOWER ? 1.0 : -1.0;
  double pend      = 0.0;  // bt contribution owed to the current row by the previously processed one
  const int64_t cnt = n1 - n0;
  for (int64_t i = 0; i < cnt; ++i) {
    const int64_t n = LOWER ? n1 - 1 - i : n0 + i;
    const int64_t o = LOWER ? n - 1 : n + 1;
    const double *a = A + n * J, *b = B + n * J;
    double bz = (MODE == 2) ? 0.0 : bZ[n * (int64_t)K + col];
    if (SOLVE) {
      double s = 0.0;
      C2B_UNROLL_J
      for (int j = 0; j < J; ++j) s = fma(b[j], bF[j], s);
      bz += s;
      if (MODE == 0) {
        const double z = Z[n * (int64_t)K + col];
        C2B_UNROLL_J
        for (int j = 0; j < J; ++j) {
          const double v = z * bF[j];
          bB[n * J + j]  = first_col ? v : bB[n * J + j] + v;
        }
        bY[n * (int64_t)K + col] = bz;
      }
    } else {
      if (MODE == 0) {
        const double y = Y[n * (int64_t)K + col];
        double s       = 0.0;
        C2B_UNROLL_J
        for (int j = 0; j < J; ++j) {
          const double v = y * bF[j];
          bB[n * J + j]  = first_col ? v : bB[n * J + j] + v;
          s              = fma(b[j], bF[j], s);
        }
        bY[n * (int64_t)K + col] = s;
      }
    }
    if (o >= 0 && o < N) {
      const double dt = sgn * (t[o] - t[n]);
      compute_p<JT>(J, cj, dup, dt, p);
      double fac = 0.0;
      C2B_UNROLL_J
      for (int j = 0; j < J; ++j) {
        const double au = a[j] * bz;
        const double nf = SOLVE ? bF[j] - au : bF[j] + au;
        if (MODE == 0) {
          const double Fn = F[n * (int64_t)J * K + (int64_t)J * col + j];
          const double pf = p[j] * Fn;
          const double v  = SOLVE ? -(bz * pf) : bz * pf;
          bA[n * J + j]   = first_col ? v : bA[n * J + j] + v;
          const double bp = Fn * nf * p[j];
          bc_acc[j]       = fma(dt, bp, bc_acc[j]);
          fac             = fma(cj[j], bp, fac);
        }
        bF[j] = p[j] * nf;
      }
      if (MODE == 0) {
        const double v = pend - sgn * fac;
        bt[n]          = first_col ? v : bt[n] + v;
        pend           = sgn * fac;
      }
    } else if (MODE == 0) {
      C2B_UNROLL_J
      for (int j = 0; j < J; ++j)
        if (first_col) bA[n * J + j] = 0.0;
      bt[n] = first_col ? pend : bt[n] + pend;
      pend  = 0.0;
    }
  }
  if (MODE == 0) *edge_out += pend;
}

// apply: thread per (series, chunk); loops over the K columns
template <int JT, bool SOLVE, bool LOWER>
__global__ void sweep_rev_apply_kernel(SweepRevArgs a) {
  constexpr int MJ = Width<JT>::MAXJ;
  const int J      = JT ? JT : a.J;
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= a.batch * a.C) return;
  const int64_t chunk = gid % a.C, b = gid / a.C;
  int64_t n0, n1;
  chunk_range(a.N, a.L, chunk, &n0, &n1);
  const double *t = a.t + b * a.t_bs;
  const double *c = a.c + b * a.c_bs;
  double cj[MJ], bF[MJ], bc_acc[MJ];
  unsigned dup[MJ];
  C2B_UNROLL_J
  for (int j = 0; j < J; ++j) {
    cj[j]     = c[j];
    bc_acc[j] = 0.0;
  }
  dup_masks<JT>(J, cj, dup);
  const int64_t NJ = a.N * J, NK = a.N * a.K;
  const int64_t leaf = LOWER ? a.C - 1 - chunk : chunk;  // processing order: descending for lower
  double edge        = 0.0;
  for (int col = 0; col < a.K; ++col) {
    if (a.carry) {
      const double *st = a.carry + (b * a.node_stride + leaf) * (int64_t)(J * a.K) + (int64_t)J * col;
      C2B_UNROLL_J
      for (int j = 0; j < J; ++j) bF[j] = st[j];
    } else {
      C2B_UNROLL_J
      for (int j = 0; j < J; ++j) bF[j] = 0.0;
    }
    sweep_rev_rows<JT, SOLVE, LOWER, 0>(J, a.K, col, col == 0, a.N, n0, n1, t, cj, dup, a.A + b * NJ, a.B + b * NJ, a.Y + b * a.y_bs,
                                        a.Z + b * NK, a.F + b * NK * J, a.bZ + b * NK, a.bt + b * a.N, a.bA + b * NJ, a.bB + b * NJ,
                                        a.bY + b * NK, bF, bc_acc, &edge);
  }
  double *bcp = a.bc_part + (b * a.C + chunk) * J;
  C2B_UNROLL_J
  for (int j = 0; j < J; ++j) bcp[j] = bc_acc[j];
  a.bt_edge[b * a.C + chunk] = edge;
}

// summary: thread per (series, chunk, column') -- same column convention as the forward summary
template <int JT, bool SOLVE, bool LOWER>
__global__ void sweep_rev_summary_kernel(SweepRevArgs a) {
  constexpr int MJ = Width<JT>::MAXJ;
  const int J      = JT ? JT : a.J;
  const int ncol   = SOLVE ? J + a.K : a.K;
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= a.batch * a.C * ncol) return;
  const int col       = (int)(gid % ncol);
  const int64_t chunk = (gid / ncol) % a.C;
  const int64_t b     = gid / ((int64_t)ncol * a.C);
  int64_t n0, n1;
  chunk_range(a.N, a.L, chunk, &n0, &n1);
  const double *t = a.t + b * a.t_bs;
  const double *c = a.c + b * a.c_bs;
  double cj[MJ], bF[MJ];
  unsigned dup[MJ];
  C2B_UNROLL_J
  for (int j = 0; j < J; ++j) cj[j] = c[j];
  dup_masks<JT>(J, cj, dup);
  const int64_t leaf = LOWER ? a.C - 1 - chunk : chunk;
  const int es       = SOLVE ? J * J + J * a.K : J + J * a.K;
  double *e          = a.leaves + (b * a.node_stride + leaf) * (int64_t)es;
  const int64_t NJ = a.N * J, NK = a.N * a.K;
  const bool basis = SOLVE && col < J;
  const int k      = SOLVE ? col - J : col;
  C2B_UNROLL_J
  for (int j = 0; j < J; ++j) bF[j] = (basis && j == col) ? 1.0 : 0.0;
  if (basis)
    sweep_rev_rows<JT, SOLVE, LOWER, 2>(J, a.K, 0, false, a.N, n0, n1, t, cj, dup, a.A + b * NJ, a.B + b * NJ, nullptr, nullptr, nullptr,
                                        nullptr, nullptr, nullptr, nullptr, nullptr, bF, nullptr, nullptr);
  else
    sweep_rev_rows<JT, SOLVE, LOWER, 1>(J, a.K, k, false, a.N, n0, n1, t, cj, dup, a.A + b * NJ, a.B + b * NJ, nullptr, nullptr, nullptr,
                                        a.bZ + b * NK, nullptr, nullptr, nullptr, nullptr, bF, nullptr, nullptr);
  if (basis) {
    C2B_UNROLL_J
    for (int j = 0; j < J; ++j) e[j * J + col] = bF[j];
  } else {
    double *bb = e + (SOLVE ? J * J : J);
    C2B_UNROLL_J
    for (int j = 0; j < J; ++j) bb[j + J * k] = bF[j];
    if (!SOLVE && k == 0) {
      // product of the p's the chunk applies: rows with a neighbour only
      double dt;
      if (LOWER)
        dt = t[n0 > 0 ? n0 - 1 : 0] - t[n1 - 1];
      else
        dt = t[n0] - t[n1 < a.N ? n1 : a.N - 1];
      C2B_UNROLL_J
      for (int j = 0; j < J; ++j) e[j] = exp(cj[j] * dt);
    }
  }
}

}  // namespace c2b
