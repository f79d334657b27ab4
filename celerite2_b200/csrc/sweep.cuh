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
// Row fields staged through RowPipe: 0 t[n], 1 t[neighbour], 2 y, 3 old z (matmul), 4.. A row, 4+J.. B row.
constexpr int kSweepSlots = 4;
template <int JT>
struct SweepRow {
  static constexpr int MJ   = Width<JT>::MAXJ;
  static constexpr int ROWD = 4 + 2 * MJ;
};

template <int JT, bool SOLVE, bool LOWER, bool WRITE>
C2B_DEV void sweep_rows(int J, int K, int64_t N, int64_t n0, int64_t n1, const double *t, const double *cj, const unsigned *dup,
                        const double *A, const double *B, const double *Ycol, double *Zcol, double *Fcol, double *G, double *smem,
                        const double *zdiv = nullptr) {
  constexpr int MJ = Width<JT>::MAXJ;
  RowPipe<(JT > 0), SweepRow<JT>::ROWD, kSweepSlots> rp;
  rp.init(smem);
  double p[MJ];
  const int64_t cnt = n1 - n0;
  auto fetch = [&](int64_t i) {
    const int64_t n = LOWER ? n0 + i : n1 - 1 - i;
    const int64_t o = LOWER ? n - 1 : n + 1;
    rp.put(0, t + n);
    if (o >= 0 && o < N) rp.put(1, t + o);
    if (Ycol) rp.put(2, Ycol + n * (int64_t)K);
    if (!SOLVE && WRITE) rp.put(3, Zcol + n * (int64_t)K);
    C2B_UNROLL_J
    for (int j = 0; j < J; ++j) {
      rp.put(4 + j, A + n * J + j);
      rp.put(4 + MJ + j, B + n * J + j);
    }
  };
  for (int q = 0; q < rp.D; ++q) {
    const bool f = q < cnt;
    if (f) fetch(q);
    rp.commit(f);
  }
  for (int64_t i = 0; i < cnt; ++i) {
    const bool f = i + rp.D < cnt;
    if (f) fetch(i + rp.D);
    rp.commit(f);
    rp.acquire();
    const int64_t n = LOWER ? n0 + i : n1 - 1 - i;
    const bool edge = LOWER ? (n == 0) : (n == N - 1);
    if (edge) {
      C2B_UNROLL_J
      for (int j = 0; j < J; ++j) p[j] = 1.0;
    } else {
      const double dt = LOWER ? rp.get(1) - rp.get(0) : rp.get(0) - rp.get(1);
      compute_p<JT>(J, cj, dup, dt, p);
    }
    if (WRITE && Fcol) {
      C2B_UNROLL_J
      for (int j = 0; j < J; ++j) Fcol[n * (int64_t)J * K + j] = G[j];
    }
    double s = 0.0;
    C2B_UNROLL_J
    for (int j = 0; j < J; ++j) {
      G[j] = p[j] * G[j];
      s    = fma(rp.get(4 + j), G[j], s);
    }
    const double y = Ycol ? rp.get(2) : 0.0;
    double z, r;
    if (SOLVE) {
      z = y - s;
      r = z;
    } else {
      z = (WRITE ? rp.get(3) : 0.0) + s;
      r = y;
    }
    if (WRITE) Zcol[n * (int64_t)K] = (SOLVE && zdiv) ? z / zdiv[n] : z;
    C2B_UNROLL_J
    for (int j = 0; j < J; ++j) G[j] = fma(rp.get(4 + MJ + j), r, G[j]);
    rp.release();
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
  pdl_prologue();
  extern __shared__ double c2b_ring[];
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
                                     a.Z + b * NK + col, a.F ? a.F + b * NK * J + (int64_t)J * col : nullptr, G, c2b_ring,
                                     a.zdiv ? a.zdiv + b * a.N : nullptr);
}

// ---- summary: thread per (series, chunk, column'), column' in [0, J+K) for solves, [0, K) for matmuls ---------
//   solve  : columns 0..J-1 propagate the unit vectors (-> the chunk's J x J transition M),
//            columns J..J+K-1 propagate zero with the real right-hand sides (-> b)
//   matmul : the transition is diag(prod p) = exp(c * (t at chunk entry - t at chunk exit)); only b is swept
template <int JT, bool SOLVE, bool LOWER>
__global__ void sweep_summary_kernel(SweepArgs a) {
  pdl_prologue();
  extern __shared__ double c2b_ring[];
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
                                      basis ? nullptr : a.Y + b * a.y_bs + k, nullptr, nullptr, G, c2b_ring);
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

// K == 1 solves: one thread per (series, chunk) propagates the J unit vectors AND the real right-hand side in
// one pass over the rows (one row fetch, J+1 independent dependency chains).
template <int JT, bool LOWER>
__global__ void sweep_summary1_kernel(SweepArgs a) {
  pdl_prologue();
  extern __shared__ double c2b_ring[];
  constexpr int MJ = Width<JT>::MAXJ;
  const int J      = JT ? JT : a.J;
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= a.batch * a.C) return;
  const int64_t chunk = gid % a.C, b = gid / a.C;
  int64_t n0, n1;
  chunk_range(a.N, a.L, chunk, &n0, &n1);
  const double *t = a.t + b * a.t_bs;
  const double *c = a.c + b * a.c_bs;
  const double *A = a.A + b * a.N * J, *B = a.B + b * a.N * J, *Y = a.Y + b * a.y_bs;
  double cj[MJ], p[MJ], G[(MJ + 1) * MJ];
  unsigned dup[MJ];
  C2B_UNROLL_J
  for (int j = 0; j < J; ++j) cj[j] = c[j];
  dup_masks<JT>(J, cj, dup);
  C2B_PRAGMA(unroll Width<JT>::UNR + (JT > 0 && JT <= 8 ? 1 : 0))
  for (int q = 0; q <= J; ++q) {
    C2B_UNROLL_J
    for (int j = 0; j < J; ++j) G[q * MJ + j] = (q == j) ? 1.0 : 0.0;
  }
  RowPipe<(JT > 0), SweepRow<JT>::ROWD, kSweepSlots> rp;
  rp.init(c2b_ring);
  const int64_t cnt = n1 - n0;
  auto fetch = [&](int64_t i) {
    const int64_t n = LOWER ? n0 + i : n1 - 1 - i;
    const int64_t o = LOWER ? n - 1 : n + 1;
    rp.put(0, t + n);
    if (o >= 0 && o < a.N) rp.put(1, t + o);
    rp.put(2, Y + n);
    C2B_UNROLL_J
    for (int j = 0; j < J; ++j) {
      rp.put(4 + j, A + n * J + j);
      rp.put(4 + MJ + j, B + n * J + j);
    }
  };
  for (int q = 0; q < rp.D; ++q) {
    const bool f = q < cnt;
    if (f) fetch(q);
    rp.commit(f);
  }
  for (int64_t i = 0; i < cnt; ++i) {
    const bool f = i + rp.D < cnt;
    if (f) fetch(i + rp.D);
    rp.commit(f);
    rp.acquire();
    const int64_t n = LOWER ? n0 + i : n1 - 1 - i;
    const bool edge = LOWER ? (n == 0) : (n == a.N - 1);
    if (edge) {
      C2B_UNROLL_J
      for (int j = 0; j < J; ++j) p[j] = 1.0;
    } else {
      const double dt = LOWER ? rp.get(1) - rp.get(0) : rp.get(0) - rp.get(1);
      compute_p<JT>(J, cj, dup, dt, p);
    }
    const double y = rp.get(2);
    C2B_PRAGMA(unroll Width<JT>::UNR + (JT > 0 && JT <= 8 ? 1 : 0))
    for (int q = 0; q <= J; ++q) {
      double s = 0.0;
      C2B_UNROLL_J
      for (int j = 0; j < J; ++j) {
        G[q * MJ + j] = p[j] * G[q * MJ + j];
        s             = fma(rp.get(4 + j), G[q * MJ + j], s);
      }
      const double r = (q == J ? y : 0.0) - s;
      C2B_UNROLL_J
      for (int j = 0; j < J; ++j) G[q * MJ + j] = fma(rp.get(4 + MJ + j), r, G[q * MJ + j]);
    }
    rp.release();
  }
  const int64_t leaf = LOWER ? chunk : a.C - 1 - chunk;
  double *e          = a.leaves + (b * a.node_stride + leaf) * (int64_t)(J * J + J);
  C2B_UNROLL_J
  for (int q = 0; q < J; ++q) {
    C2B_UNROLL_J
    for (int j = 0; j < J; ++j) e[j * J + q] = G[q * MJ + j];
  }
  C2B_UNROLL_J
  for (int j = 0; j < J; ++j) e[J * J + j] = G[J * MJ + j];
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
// Row fields staged through RowPipe: 0 t[n], 1 t[neighbour], 2 bZ, 3 z or y, 4.. A row, 4+J.. B row, 4+2J.. F row.
constexpr int kSweepRevSlots = 4;
template <int JT>
struct SweepRevRow {
  static constexpr int MJ   = Width<JT>::MAXJ;
  static constexpr int ROWD = 4 + 3 * MJ;
};

template <int JT, bool SOLVE, bool LOWER, int MODE>
C2B_DEV void sweep_rev_rows(int J, int K, int col, bool first_col, int64_t N, int64_t n0, int64_t n1, const double *t, const double *cj,
                            const unsigned *dup, const double *A, const double *B, const double *Y, const double *Z, const double *F,
                            const double *bZ, double *bt, double *bA, double *bB, double *bY, double *bF, double *bc_acc,
                            double *edge_out, double *smem) {
  constexpr int MJ = Width<JT>::MAXJ;
  RowPipe<(JT > 0), SweepRevRow<JT>::ROWD, kSweepRevSlots> rp;
  rp.init(smem);
  double p[MJ];
  const double sgn = LOWER ? 1.0 : -1.0;
  double pend      = 0.0;  // bt contribution owed to the current row by the previously processed one
  const int64_t cnt = n1 - n0;
  auto fetch = [&](int64_t i) {
    const int64_t n = LOWER ? n1 - 1 - i : n0 + i;
    const int64_t o = LOWER ? n - 1 : n + 1;
    rp.put(0, t + n);
    if (o >= 0 && o < N) rp.put(1, t + o);
    if (MODE != 2) rp.put(2, bZ + n * (int64_t)K + col);
    if (MODE == 0) rp.put(3, (SOLVE ? Z : Y) + n * (int64_t)K + col);
    C2B_UNROLL_J
    for (int j = 0; j < J; ++j) {
      rp.put(4 + j, A + n * J + j);
      if (SOLVE || MODE == 0) rp.put(4 + MJ + j, B + n * J + j);
      if (MODE == 0 && o >= 0 && o < N) rp.put(4 + 2 * MJ + j, F + n * (int64_t)J * K + (int64_t)J * col + j);
    }
  };
  for (int q = 0; q < rp.D; ++q) {
    const bool f = q < cnt;
    if (f) fetch(q);
    rp.commit(f);
  }
  for (int64_t i = 0; i < cnt; ++i) {
    const bool f = i + rp.D < cnt;
    if (f) fetch(i + rp.D);
    rp.commit(f);
    rp.acquire();
    const int64_t n = LOWER ? n1 - 1 - i : n0 + i;
    const int64_t o = LOWER ? n - 1 : n + 1;
    double bz = (MODE == 2) ? 0.0 : rp.get(2);
    if (SOLVE) {
      double s = 0.0;
      C2B_UNROLL_J
      for (int j = 0; j < J; ++j) s = fma(rp.get(4 + MJ + j), bF[j], s);
      bz += s;
      if (MODE == 0) {
        const double z = rp.get(3);
        C2B_UNROLL_J
        for (int j = 0; j < J; ++j) {
          const double v = z * bF[j];
          bB[n * J + j]  = first_col ? v : bB[n * J + j] + v;
        }
        bY[n * (int64_t)K + col] = bz;
      }
    } else {
      if (MODE == 0) {
        const double y = rp.get(3);
        double s       = 0.0;
        C2B_UNROLL_J
        for (int j = 0; j < J; ++j) {
          const double v = y * bF[j];
          bB[n * J + j]  = first_col ? v : bB[n * J + j] + v;
          s              = fma(rp.get(4 + MJ + j), bF[j], s);
        }
        bY[n * (int64_t)K + col] = s;
      }
    }
    if (o >= 0 && o < N) {
      const double dt = sgn * (rp.get(1) - rp.get(0));
      compute_p<JT>(J, cj, dup, dt, p);
      double fac = 0.0;
      C2B_UNROLL_J
      for (int j = 0; j < J; ++j) {
        const double au = rp.get(4 + j) * bz;
        const double nf = SOLVE ? bF[j] - au : bF[j] + au;
        if (MODE == 0) {
          const double Fn = rp.get(4 + 2 * MJ + j);
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
    rp.release();
  }
  if (MODE == 0) *edge_out += pend;
}

// apply: thread per (series, chunk); loops over the K columns
template <int JT, bool SOLVE, bool LOWER>
__global__ void sweep_rev_apply_kernel(SweepRevArgs a) {
  pdl_prologue();
  extern __shared__ double c2b_ring[];
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
                                        a.bY + b * NK, bF, bc_acc, &edge, c2b_ring);
  }
  double *bcp = a.bc_part + (b * a.C + chunk) * J;
  C2B_UNROLL_J
  for (int j = 0; j < J; ++j) bcp[j] = bc_acc[j];
  a.bt_edge[b * a.C + chunk] = edge;
}

// summary: thread per (series, chunk, column') -- same column convention as the forward summary
template <int JT, bool SOLVE, bool LOWER>
__global__ void sweep_rev_summary_kernel(SweepRevArgs a) {
  pdl_prologue();
  extern __shared__ double c2b_ring[];
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
                                        nullptr, nullptr, nullptr, nullptr, nullptr, bF, nullptr, nullptr, c2b_ring);
  else
    sweep_rev_rows<JT, SOLVE, LOWER, 1>(J, a.K, k, false, a.N, n0, n1, t, cj, dup, a.A + b * NJ, a.B + b * NJ, nullptr, nullptr, nullptr,
                                        a.bZ + b * NK, nullptr, nullptr, nullptr, nullptr, bF, nullptr, nullptr, c2b_ring);
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

// K == 1 solves, reverse mode: one thread per (series, chunk) carries the J unit columns and the real column.
template <int JT, bool LOWER>
__global__ void sweep_rev_summary1_kernel(SweepRevArgs a) {
  pdl_prologue();
  extern __shared__ double c2b_ring[];
  constexpr int MJ = Width<JT>::MAXJ;
  const int J      = JT ? JT : a.J;
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= a.batch * a.C) return;
  const int64_t chunk = gid % a.C, b = gid / a.C;
  int64_t n0, n1;
  chunk_range(a.N, a.L, chunk, &n0, &n1);
  const double *t = a.t + b * a.t_bs;
  const double *c = a.c + b * a.c_bs;
  const double *A = a.A + b * a.N * J, *B = a.B + b * a.N * J, *bZ = a.bZ + b * a.N;
  double cj[MJ], p[MJ], G[(MJ + 1) * MJ];
  unsigned dup[MJ];
  C2B_UNROLL_J
  for (int j = 0; j < J; ++j) cj[j] = c[j];
  dup_masks<JT>(J, cj, dup);
  C2B_PRAGMA(unroll Width<JT>::UNR + (JT > 0 && JT <= 8 ? 1 : 0))
  for (int q = 0; q <= J; ++q) {
    C2B_UNROLL_J
    for (int j = 0; j < J; ++j) G[q * MJ + j] = (q == j) ? 1.0 : 0.0;
  }
  RowPipe<(JT > 0), SweepRevRow<JT>::ROWD, kSweepRevSlots> rp;
  rp.init(c2b_ring);
  const int64_t cnt = n1 - n0;
  const double sgn  = LOWER ? 1.0 : -1.0;
  auto fetch = [&](int64_t i) {
    const int64_t n = LOWER ? n1 - 1 - i : n0 + i;
    const int64_t o = LOWER ? n - 1 : n + 1;
    rp.put(0, t + n);
    if (o >= 0 && o < a.N) rp.put(1, t + o);
    rp.put(2, bZ + n);
    C2B_UNROLL_J
    for (int j = 0; j < J; ++j) {
      rp.put(4 + j, A + n * J + j);
      rp.put(4 + MJ + j, B + n * J + j);
    }
  };
  for (int q = 0; q < rp.D; ++q) {
    const bool f = q < cnt;
    if (f) fetch(q);
    rp.commit(f);
  }
  for (int64_t i = 0; i < cnt; ++i) {
    const bool f = i + rp.D < cnt;
    if (f) fetch(i + rp.D);
    rp.commit(f);
    rp.acquire();
    const int64_t n = LOWER ? n1 - 1 - i : n0 + i;
    const int64_t o = LOWER ? n - 1 : n + 1;
    if (o >= 0 && o < a.N) {
      compute_p<JT>(J, cj, dup, sgn * (rp.get(1) - rp.get(0)), p);
      const double bz0 = rp.get(2);
      C2B_PRAGMA(unroll Width<JT>::UNR + (JT > 0 && JT <= 8 ? 1 : 0))
      for (int q = 0; q <= J; ++q) {
        double bz = (q == J) ? bz0 : 0.0;
        C2B_UNROLL_J
        for (int j = 0; j < J; ++j) bz = fma(rp.get(4 + MJ + j), G[q * MJ + j], bz);
        C2B_UNROLL_J
        for (int j = 0; j < J; ++j) G[q * MJ + j] = p[j] * fma(-rp.get(4 + j), bz, G[q * MJ + j]);
      }
    }
    rp.release();
  }
  const int64_t leaf = LOWER ? a.C - 1 - chunk : chunk;
  double *e          = a.leaves + (b * a.node_stride + leaf) * (int64_t)(J * J + J);
  C2B_UNROLL_J
  for (int q = 0; q < J; ++q) {
    C2B_UNROLL_J
    for (int j = 0; j < J; ++j) e[j * J + q] = G[q * MJ + j];
  }
  C2B_UNROLL_J
  for (int j = 0; j < J; ++j) e[J * J + j] = G[J * MJ + j];
}

}  // namespace c2b
