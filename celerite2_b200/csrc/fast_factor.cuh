// Fast paths (compile-time J, TMA tile pipeline) of factor: apply and summary.  Arithmetic identical to the generic
// kernels in factor.cuh; only the data movement differs (tile.cuh).
#pragma once
#include "factor.cuh"
#include "tile.cuh"

namespace c2b {

constexpr int kFT = 4;  // rows per tile

template <int J>
struct FFactorLayout {
  // input slot: t (T+1 rows, one halo row below), a, U, V
  static constexpr int oT = 0;
  static constexpr int oA = oT + ((kFT + 1 + 1) & ~1);
  static constexpr int oU = oA + kFT;
  static constexpr int oV = oU + kFT * J;
  static constexpr int nIn = oV + kFT * J;
  static constexpr int SIN = (nIn % 4 == 2) ? nIn : ((nIn + 1) & ~1) + (((((nIn + 1) & ~1)) % 4 == 2) ? 0 : 2);
  // output slot: d, W, S
  static constexpr int oD = 0;
  static constexpr int oW = oD + kFT;
  static constexpr int oS = oW + ((kFT * J + 1) & ~1);
  static constexpr int nOut = oS + kFT * J * J;
  static constexpr int SOUT = (((nOut + 1) & ~1) % 4 == 2) ? ((nOut + 1) & ~1) : ((nOut + 1) & ~1) + 2;
  static constexpr int SOUT_SUMMARY = 2;  // the summary writes no per-row outputs
};

template <int J, class TP>
C2B_DEV void ffactor_fetch(TP &tp, int stage, int tile, const double *t, const double *a, const double *U, const double *V) {
  using Lay = FFactorLayout<J>;
  tp.begin_fetch();
  tp.template fetch<1, 1, 0>(stage, tile, t, Lay::oT);
  tp.template fetch<1, 0, 0>(stage, tile, a, Lay::oA);
  tp.template fetch<J, 0, 0>(stage, tile, U, Lay::oU);
  tp.template fetch<J, 0, 0>(stage, tile, V, Lay::oV);
  tp.end_fetch(stage);
}

template <int J>
__global__ void __launch_bounds__(32) ffactor_apply_kernel(FactorArgs g) {
  extern __shared__ __align__(16) double c2b_tiles[];
  using Lay = FFactorLayout<J>;
  using TP  = TilePipe<kFT, Lay::SIN, Lay::SOUT>;
  const int64_t gid   = (int64_t)blockIdx.x * 32 + threadIdx.x;
  const bool active   = gid < g.batch * g.C;
  const int64_t chunk = active ? gid % g.C : 0, b = active ? gid / g.C : 0;
  ChunkGeom cg;
  cg.n0         = chunk * g.L;
  cg.cnt        = active ? (int)((cg.n0 + g.L < g.N ? cg.n0 + g.L : g.N) - cg.n0) : 0;
  cg.N          = g.N;
  cg.descending = false;
  TP tp;
  tp.init(c2b_tiles, cg);
  if (!active) return;
  const double *t = g.t + b * g.t_bs;
  const double *c = g.c + b * g.c_bs;
  const int64_t NJ = g.N * J;
  const double *a = g.a + b * g.N, *U = g.U + b * NJ, *V = g.V + b * NJ;
  double *d = g.d + b * g.N, *W = g.W + b * NJ;
  double *S = g.S ? g.S + b * NJ * J : nullptr;
  double cj[J], p[J], u[J], tmp[J], w[J];
  unsigned dup[J];
  double Sg[J * J];
#pragma unroll
  for (int j = 0; j < J; ++j) cj[j] = c[j];
  dup_masks<J>(J, cj, dup);
  if (g.carry) {
    const double *st = g.carry + (b * g.node_stride + chunk) * (int64_t)(J * J);
#pragma unroll
    for (int i = 0; i < J * J; ++i) Sg[i] = -st[i];
  } else {
#pragma unroll
    for (int i = 0; i < J * J; ++i) Sg[i] = 0.0;
  }
  ffactor_fetch<J>(tp, 0, 0, t, a, U, V);
  bool failed = false;
  for (int tile = 0; tile < tp.ntiles && !failed; ++tile) {
    const int stage = tile & 1;
    if (tile + 1 < tp.ntiles) ffactor_fetch<J>(tp, stage ^ 1, tile + 1, t, a, U, V);
    tp.wait_tile(stage);
    tp.begin_tile_outputs();
    const int rows = tp.rows_in_tile(tile);
    int done       = rows;
    for (int rp = 0; rp < rows; ++rp) {
      const int64_t n = tp.row_of(tile, rp);
      if (n == 0) {
#pragma unroll
        for (int j = 0; j < J; ++j) p[j] = 1.0;
      } else {
        const double dt = tp.template in<1, 1>(stage, Lay::oT, rp, 0, -1) - tp.template in<1, 1>(stage, Lay::oT, rp, 0);
        compute_p<J>(J, cj, dup, dt, p);
      }
#pragma unroll
      for (int j = 0; j < J; ++j) u[j] = tp.template in<J, 0>(stage, Lay::oU, rp, j);
#pragma unroll
      for (int i = 0; i < J; ++i) {
#pragma unroll
        for (int j = 0; j < J; ++j) {
          const double st = p[j] * Sg[j + J * i];
          if (S) tp.template out<J * J>(Lay::oS, rp, j + J * i, st);
          Sg[j + J * i] = st * p[i];
        }
      }
      double acc = 0.0;
#pragma unroll
      for (int i = 0; i < J; ++i) {
        double s = 0.0;
#pragma unroll
        for (int j = 0; j < J; ++j) s = fma(u[j], Sg[j + J * i], s);
        tmp[i] = s;
        acc    = fma(s, u[i], acc);
      }
      const double dn = tp.template in<1, 0>(stage, Lay::oA, rp, 0) - acc;
      tp.template out<1>(Lay::oD, rp, 0, dn);
      if (n >= 1 && dn <= 0.0) {
        // forward.hpp:128: d[n] is written, W[n] and everything after it is not
        record_failure(g.status ? g.status + b : nullptr, n);
        d[n]   = dn;
        done   = rp;
        failed = true;
        break;
      }
#pragma unroll
      for (int j = 0; j < J; ++j) {
        w[j] = (tp.template in<J, 0>(stage, Lay::oV, rp, j) - tmp[j]) / dn;
        tp.template out<J>(Lay::oW, rp, j, w[j]);
      }
#pragma unroll
      for (int i = 0; i < J; ++i) {
#pragma unroll
        for (int j = 0; j < J; ++j) Sg[j + J * i] = fma(dn * w[j], w[i], Sg[j + J * i]);
      }
    }
    tp.begin_stores();
    tp.template store<1>(tile, d, Lay::oD, done);
    tp.template store<J>(tile, W, Lay::oW, done);
    if (S) tp.template store<J * J>(tile, S, Lay::oS, done);
    tp.end_tile_outputs();
    if (failed && tile + 1 < tp.ntiles) tp.wait_tile(stage ^ 1);  // never leave with a bulk load in flight into our slot
  }
  tp.finish();
}

template <int J>
__global__ void __launch_bounds__(32) ffactor_summary_kernel(FactorArgs g) {
  extern __shared__ __align__(16) double c2b_tiles[];
  using Lay = FFactorLayout<J>;
  using TP  = TilePipe<kFT, Lay::SIN, Lay::SOUT_SUMMARY>;
  const int64_t gid   = (int64_t)blockIdx.x * 32 + threadIdx.x;
  const bool active   = gid < g.batch * g.C;
  const int64_t chunk = active ? gid % g.C : 0, b = active ? gid / g.C : 0;
  ChunkGeom cg;
  cg.n0         = chunk * g.L;
  cg.cnt        = active ? (int)((cg.n0 + g.L < g.N ? cg.n0 + g.L : g.N) - cg.n0) : 0;
  cg.N          = g.N;
  cg.descending = false;
  TP tp;
  tp.init(c2b_tiles, cg);
  if (!active) return;
  const double *t = g.t + b * g.t_bs;
  const double *c = g.c + b * g.c_bs;
  const int64_t NJ = g.N * J;
  const double *a = g.a + b * g.N, *U = g.U + b * NJ, *V = g.V + b * NJ;
  double cj[J], p[J], u[J], q[J], r[J], gv[J];
  unsigned dup[J];
  double A[J * J], Cm[J * J], Q[J * J];
#pragma unroll
  for (int j = 0; j < J; ++j) cj[j] = c[j];
  dup_masks<J>(J, cj, dup);
#pragma unroll
  for (int i = 0; i < J; ++i) {
#pragma unroll
    for (int j = 0; j < J; ++j) {
      A[i * J + j]  = (i == j) ? 1.0 : 0.0;
      Cm[i * J + j] = 0.0;
      Q[i * J + j]  = 0.0;
    }
  }
  ffactor_fetch<J>(tp, 0, 0, t, a, U, V);
  for (int tile = 0; tile < tp.ntiles; ++tile) {
    const int stage = tile & 1;
    if (tile + 1 < tp.ntiles) ffactor_fetch<J>(tp, stage ^ 1, tile + 1, t, a, U, V);
    tp.wait_tile(stage);
    const int rows = tp.rows_in_tile(tile);
    for (int rp = 0; rp < rows; ++rp) {
      const int64_t n = tp.row_of(tile, rp);
      if (n == 0) {
#pragma unroll
        for (int j = 0; j < J; ++j) p[j] = 1.0;
      } else {
        const double dt = tp.template in<1, 1>(stage, Lay::oT, rp, 0, -1) - tp.template in<1, 1>(stage, Lay::oT, rp, 0);
        compute_p<J>(J, cj, dup, dt, p);
      }
#pragma unroll
      for (int j = 0; j < J; ++j) {
        u[j] = tp.template in<J, 0>(stage, Lay::oU, rp, j);
        r[j] = 0.0;
      }
      double uq = 0.0;
#pragma unroll
      for (int i = 0; i < J; ++i) {
        double s = 0.0;
#pragma unroll
        for (int j = 0; j < J; ++j) {
          A[i * J + j] *= p[i];
          Cm[i * J + j] *= p[i] * p[j];
          s    = fma(Cm[i * J + j], u[j], s);
          r[j] = fma(A[i * J + j], u[i], r[j]);
        }
        q[i] = s;
        uq   = fma(u[i], s, uq);
      }
      const double gam  = tp.template in<1, 0>(stage, Lay::oA, rp, 0) + uq;
      const double igam = 1.0 / gam;
#pragma unroll
      for (int j = 0; j < J; ++j) gv[j] = (q[j] + tp.template in<J, 0>(stage, Lay::oV, rp, j)) * igam;
#pragma unroll
      for (int i = 0; i < J; ++i) {
        const double ggi = gam * gv[i], rgi = r[i] * igam;
#pragma unroll
        for (int j = 0; j < J; ++j) {
          A[i * J + j]  = fma(-gv[i], r[j], A[i * J + j]);
          Cm[i * J + j] = fma(-ggi, gv[j], Cm[i * J + j]);
          Q[i * J + j]  = fma(rgi, r[j], Q[i * J + j]);
        }
      }
    }
  }
  double *e = g.leaves + (b * g.node_stride + chunk) * (int64_t)(3 * J * J);
#pragma unroll
  for (int i = 0; i < J * J; ++i) {
    e[i]             = A[i];
    e[J * J + i]     = Cm[i];
    e[2 * J * J + i] = Q[i];
  }
}

}  // namespace c2b
