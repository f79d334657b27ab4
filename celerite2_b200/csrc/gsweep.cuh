// Lane-group kernels of the K == 1 sweeps (solve / matmul, lower / upper) and of their reverse mode.
// apply: lane c of a group holds element c of the J-vector state; summary of the solves: lane q holds the whole
// J-vector of column q (q < J: unit vectors -> transition M, q == J: the real right-hand side -> b).
#pragma once
#include "group.cuh"
#include "sweep.cuh"

namespace c2b {

constexpr int kGroupThreadsS = 64;

constexpr int kSweepDepth = 8;  // rows of look-ahead in the per-group input ring (group.cuh)

// ---- forward apply -----------------------------------------------------------------------------------------------
// ring row: A[J] B[J] | y, t_n, t_other, z_old
template <int J, bool SOLVE, bool LOWER>
__global__ void __launch_bounds__(kGroupThreadsS) gsweep_apply_kernel(SweepArgs a) {
  constexpr int G    = GroupOf<J>::G;
  constexpr int GP   = kGroupThreadsS / G;
  constexpr int oX   = 2 * J;
  constexpr int ROWD = (oX + 4 + 1) & ~1;
  __shared__ __align__(16) double sring[GP * GRing<G, ROWD, kSweepDepth>::kStride];
  const int c         = threadIdx.x & (G - 1);
  const int gin       = threadIdx.x / G;
  const int64_t unit  = (int64_t)blockIdx.x * GP + gin;
  const bool has      = unit < a.batch * a.C;
  const bool act      = has && c < J;
  const int cj        = c < J ? c : 0;
  const int64_t chunk = has ? unit % a.C : 0, b = has ? unit / a.C : 0;
  const int64_t n0    = chunk * a.L;
  const int64_t n1    = has ? (n0 + a.L < a.N ? n0 + a.L : a.N) : n0;
  const double *t     = a.t + b * a.t_bs;
  const double *A = a.A + b * a.N * J, *B = a.B + b * a.N * J, *Y = a.Y + b * a.y_bs;
  double *Z = a.Z + b * a.N, *F = a.F ? a.F + b * a.N * J : nullptr;
  GRing<G, ROWD, kSweepDepth> ring{sring + gin * GRing<G, ROWD, kSweepDepth>::kStride, c};
  const double cc = act ? (a.c + b * a.c_bs)[c] : 0.0;
  double Fc = 0.0;
  if (a.carry && act) {
    const int64_t leaf = LOWER ? chunk : a.C - 1 - chunk;
    Fc                 = a.carry[(b * a.node_stride + leaf) * (int64_t)J + c];
  }
  const int cnt   = (int)(n1 - n0);
  const int trips = warp_max(cnt);
  // rows are visited in the order nf, nf + dir, ...
  constexpr int dir = LOWER ? 1 : -1;
  const int64_t nf  = LOWER ? n0 : n1 - 1;
  const double *pA = A + nf * J + cj, *pB = B + nf * J + cj;
  double *pF = F ? F + nf * J + cj : nullptr, *pZ = Z + nf;
  constexpr int NS = SOLVE ? 3 : 4;
  ScalarStreams<G, NS> sc;
  {
    const double *src[4] = {Y, t, t, Z};
    const int delta[4]   = {0, 0, -dir, 0};
    sc.init(c, has, src, delta);
  }
  auto issue = [&](int i) {
    if (i < cnt) {
      if (c < J) {
        ring.copy1(i, c, pA + (int64_t)(dir * J) * i);
        ring.copy1(i, J + c, pB + (int64_t)(dir * J) * i);
      }
      sc.issue(ring, i, oX, c, nf + dir * i, a.N);
    }
    ring.commit();
  };
#pragma unroll
  for (int s = 0; s < kSweepDepth; ++s) issue(s);
  const int64_t nedge = LOWER ? 0 : a.N - 1;
  // exp(c dt) of row i does not depend on the state: it is evaluated one trip ahead
  auto prop = [&](int i) {
    const double *rb = ring.stage(i);
    const double dt  = LOWER ? rb[oX + 2] - rb[oX + 1] : rb[oX + 1] - rb[oX + 2];
    return (i >= cnt || nf + dir * i == nedge) ? 1.0 : exp(cc * dt);
  };
  ring.acquire2();
  double p_next = prop(0);
  for (int i = 0; i < trips; ++i) {
    const bool live = i < cnt;
    ring.acquire2();
    const double *rb  = ring.stage(i);
    const double a_c  = (live && act) ? rb[cj] : 0.0;
    const double b_c  = (live && act) ? rb[J + cj] : 0.0;
    const double y    = rb[oX];
    const double zold = SOLVE ? 0.0 : rb[oX + 3];
    const double p_c  = p_next;
    p_next            = prop(i + 1);
    ring.release();
    issue(i + kSweepDepth);
    if (pF && live && act) pF[(int64_t)(dir * J) * i] = Fc;
    Fc             = p_c * Fc;
    const double s = gsum<G>(a_c * Fc);
    double z, r;
    if (SOLVE) {
      z = y - s;
      r = z;
    } else {
      z = zold + s;
      r = y;
    }
    if (live && c == G - 1) pZ[(int64_t)dir * i] = z;
    Fc = fma(b_c, r, Fc);
  }
}

// ---- forward summary, solves: lane q carries column q ----------------------------------------------------------------
// ring row: A[J] B[J] | y, t_n, t_other
template <int J, bool LOWER>
__global__ void __launch_bounds__(kGroupThreadsS) gsweep_summary_kernel(SweepArgs a) {
  constexpr int G    = GroupOf<J + 1>::G;  // J + 1 columns
  constexpr int GP   = kGroupThreadsS / G;
  constexpr int oX   = 2 * J;
  constexpr int ROWD = (oX + 3 + 1) & ~1;
  constexpr int JE   = (J + 1) & ~1;
  __shared__ __align__(16) double sring[GP * GRing<G, ROWD, kSweepDepth>::kStride];
  __shared__ __align__(16) double sxch[GP * JE];
  const int q         = threadIdx.x & (G - 1);
  const int gin       = threadIdx.x / G;
  const int64_t unit  = (int64_t)blockIdx.x * GP + gin;
  const bool has      = unit < a.batch * a.C;
  const bool act      = has && q < J;  // lanes that also own an element of the rows
  const int64_t chunk = has ? unit % a.C : 0, b = has ? unit / a.C : 0;
  const int64_t n0    = chunk * a.L;
  const int64_t n1    = has ? (n0 + a.L < a.N ? n0 + a.L : a.N) : n0;
  const double *t     = a.t + b * a.t_bs;
  const double *A = a.A + b * a.N * J, *B = a.B + b * a.N * J, *Y = a.Y + b * a.y_bs;
  GRing<G, ROWD, kSweepDepth> ring{sring + gin * GRing<G, ROWD, kSweepDepth>::kStride, q};
  double *xch     = sxch + gin * JE;
  const double cc = act ? (a.c + b * a.c_bs)[q] : 0.0;
  double Gq[J];
#pragma unroll
  for (int j = 0; j < J; ++j) Gq[j] = (j == q) ? 1.0 : 0.0;
  const int cnt   = (int)(n1 - n0);
  const int trips = warp_max(cnt);
  auto row = [&](int i) { return LOWER ? n0 + i : n1 - 1 - i; };
  auto issue = [&](int i) {
    if (i < cnt) {
      const int64_t n = row(i), o = LOWER ? n - 1 : n + 1;
      if (q < J) {
        ring.copy1(i, q, A + n * J + q);
        ring.copy1(i, J + q, B + n * J + q);
      }
      for (int e = G - 1 - q; e < 3; e += G) {
        const double *src = e == 0 ? Y + n : (e == 1 ? t + n : t + ((o >= 0 && o < a.N) ? o : n));
        ring.copy1(i, oX + e, src);
      }
    }
    ring.commit();
  };
#pragma unroll
  for (int s = 0; s < kSweepDepth; ++s) issue(s);
  for (int i = 0; i < trips; ++i) {
    const bool live = i < cnt;
    const int64_t n = live ? row(i) : 0;
    ring.acquire();
    const double *rb = ring.stage(i);
    double av[J], bv[J];
    lds_row<J>(rb, av);
    lds_row<J>(rb + J, bv);
    const double y     = rb[oX];
    const double t_cur = rb[oX + 1], t_oth = rb[oX + 2];
    const bool edge    = LOWER ? (n == 0) : (n == a.N - 1);
    const double dt    = LOWER ? t_oth - t_cur : t_cur - t_oth;
    if (act) xch[q] = (edge || !live) ? 1.0 : exp(cc * dt);
    ring.release();  // also publishes xch
    issue(i + kSweepDepth);
    double p[J];
    lds_row<J>(xch, p);
    if (live) {
      double s = 0.0;
#pragma unroll
      for (int j = 0; j < J; ++j) {
        Gq[j] = p[j] * Gq[j];
        s     = fma(av[j], Gq[j], s);
      }
      const double r = (q == J ? y : 0.0) - s;
#pragma unroll
      for (int j = 0; j < J; ++j) Gq[j] = fma(bv[j], r, Gq[j]);
    }
  }
  if (has && q <= J) {
    const int64_t leaf = LOWER ? chunk : a.C - 1 - chunk;
    double *e          = a.leaves + (b * a.node_stride + leaf) * (int64_t)(J * J + J);
    if (q < J) {
#pragma unroll
      for (int j = 0; j < J; ++j) e[j * J + q] = Gq[j];
    } else {
#pragma unroll
      for (int j = 0; j < J; ++j) e[J * J + j] = Gq[j];
    }
  }
}

// ---- reverse apply -----------------------------------------------------------------------------------------------------
// ring row: A[J] B[J] F[J] | bZ, Z or Y, t_n, t_other
template <int J, bool SOLVE, bool LOWER>
__global__ void __launch_bounds__(kGroupThreadsS) gsweep_rev_apply_kernel(SweepRevArgs a) {
  constexpr int G    = GroupOf<J>::G;
  constexpr int GP   = kGroupThreadsS / G;
  constexpr int oX   = 3 * J;
  constexpr int ROWD = (oX + 4 + 1) & ~1;
  __shared__ __align__(16) double sring[GP * GRing<G, ROWD, kSweepDepth>::kStride];
  const int c         = threadIdx.x & (G - 1);
  const int gin       = threadIdx.x / G;
  const int64_t unit  = (int64_t)blockIdx.x * GP + gin;
  const bool has      = unit < a.batch * a.C;
  const bool act      = has && c < J;
  const int cj        = c < J ? c : 0;
  const int64_t chunk = has ? unit % a.C : 0, b = has ? unit / a.C : 0;
  const int64_t n0    = chunk * a.L;
  const int64_t n1    = has ? (n0 + a.L < a.N ? n0 + a.L : a.N) : n0;
  const double *t     = a.t + b * a.t_bs;
  const int64_t NJ    = a.N * J;
  const double *A = a.A + b * NJ, *B = a.B + b * NJ, *F = a.F + b * NJ, *bZ = a.bZ + b * a.N;
  const double *ZY = SOLVE ? a.Z + b * a.N : a.Y + b * a.y_bs;
  double *bt = a.bt + b * a.N, *bA = a.bA + b * NJ, *bB = a.bB + b * NJ, *bY = a.bY + b * a.N;
  GRing<G, ROWD, kSweepDepth> ring{sring + gin * GRing<G, ROWD, kSweepDepth>::kStride, c};
  const double cc = act ? (a.c + b * a.c_bs)[c] : 0.0;
  double bF = 0.0, bc_acc = 0.0, pend = 0.0;
  if (a.carry && act) {
    const int64_t leaf = LOWER ? a.C - 1 - chunk : chunk;
    bF                 = a.carry[(b * a.node_stride + leaf) * (int64_t)J + c];
  }
  const double sgn = LOWER ? 1.0 : -1.0;
  const int cnt    = (int)(n1 - n0);
  const int trips  = warp_max(cnt);
  auto row = [&](int i) { return LOWER ? n1 - 1 - i : n0 + i; };
  auto issue = [&](int i) {
    if (i < cnt) {
      const int64_t n = row(i), o = LOWER ? n - 1 : n + 1;
      if (c < J) {
        ring.copy1(i, c, A + n * J + c);
        ring.copy1(i, J + c, B + n * J + c);
        ring.copy1(i, 2 * J + c, F + n * J + c);
      }
      for (int e = G - 1 - c; e < 4; e += G) {
        const double *src = e == 0 ? bZ + n : (e == 1 ? ZY + n : (e == 2 ? t + n : t + ((o >= 0 && o < a.N) ? o : n)));
        ring.copy1(i, oX + e, src);
      }
    }
    ring.commit();
  };
#pragma unroll
  for (int s = 0; s < kSweepDepth; ++s) issue(s);
  for (int i = 0; i < trips; ++i) {
    const bool live = i < cnt;
    const int64_t n = live ? row(i) : 0;
    const int64_t o = LOWER ? n - 1 : n + 1;
    const bool nb   = live && o >= 0 && o < a.N;
    ring.acquire();
    const double *rb   = ring.stage(i);
    const double a_c   = (live && act) ? rb[cj] : 0.0;
    const double b_c   = (live && act) ? rb[J + cj] : 0.0;
    const double Fn    = (nb && act) ? rb[2 * J + cj] : 0.0;
    const double bz0   = rb[oX], zy = rb[oX + 1];
    const double t_cur = rb[oX + 2], t_oth = rb[oX + 3];
    ring.release();
    issue(i + kSweepDepth);
    const double dot = gsum<G>(b_c * bF);
    double bz;
    if (SOLVE) {
      bz = bz0 + dot;
      if (live && c == G - 1) bY[n] = bz;
    } else {
      bz = bz0;
      if (live && c == G - 1) bY[n] = dot;
    }
    if (live && act) bB[n * J + c] = zy * bF;
    const double dt  = sgn * (t_oth - t_cur);
    const double p_c = nb ? exp(cc * dt) : 1.0;
    const double au  = a_c * bz;
    const double nf  = SOLVE ? bF - au : bF + au;
    const double pf  = p_c * Fn;
    const double bp  = nb ? Fn * nf * p_c : 0.0;
    const double fac = gsum<G>(cc * bp);
    if (nb) {
      if (act) bA[n * J + c] = SOLVE ? -(bz * pf) : bz * pf;
      bc_acc = fma(dt, bp, bc_acc);
      if (c == 0) bt[n] = pend - sgn * fac;
      pend = sgn * fac;
      bF   = p_c * nf;
    } else if (live) {
      if (act) bA[n * J + c] = 0.0;
      if (c == 0) bt[n] = pend;
      pend = 0.0;
    }
  }
  if (act) a.bc_part[(b * a.C + chunk) * J + c] = bc_acc;
  if (has && c == 0) a.bt_edge[b * a.C + chunk] = pend;
}

// ---- reverse summary, solves -----------------------------------------------------------------------------------------
// ring row: A[J] B[J] | bZ, t_n, t_other
template <int J, bool LOWER>
__global__ void __launch_bounds__(kGroupThreadsS) gsweep_rev_summary_kernel(SweepRevArgs a) {
  constexpr int G    = GroupOf<J + 1>::G;
  constexpr int GP   = kGroupThreadsS / G;
  constexpr int oX   = 2 * J;
  constexpr int ROWD = (oX + 3 + 1) & ~1;
  constexpr int JE   = (J + 1) & ~1;
  __shared__ __align__(16) double sring[GP * GRing<G, ROWD, kSweepDepth>::kStride];
  __shared__ __align__(16) double sxch[GP * JE];
  const int q         = threadIdx.x & (G - 1);
  const int gin       = threadIdx.x / G;
  const int64_t unit  = (int64_t)blockIdx.x * GP + gin;
  const bool has      = unit < a.batch * a.C;
  const bool act      = has && q < J;
  const int64_t chunk = has ? unit % a.C : 0, b = has ? unit / a.C : 0;
  const int64_t n0    = chunk * a.L;
  const int64_t n1    = has ? (n0 + a.L < a.N ? n0 + a.L : a.N) : n0;
  const double *t     = a.t + b * a.t_bs;
  const double *A = a.A + b * a.N * J, *B = a.B + b * a.N * J, *bZ = a.bZ + b * a.N;
  GRing<G, ROWD, kSweepDepth> ring{sring + gin * GRing<G, ROWD, kSweepDepth>::kStride, q};
  double *xch     = sxch + gin * JE;
  const double cc = act ? (a.c + b * a.c_bs)[q] : 0.0;
  double Gq[J];
#pragma unroll
  for (int j = 0; j < J; ++j) Gq[j] = (j == q) ? 1.0 : 0.0;
  const double sgn = LOWER ? 1.0 : -1.0;
  const int cnt    = (int)(n1 - n0);
  const int trips  = warp_max(cnt);
  auto row = [&](int i) { return LOWER ? n1 - 1 - i : n0 + i; };
  auto issue = [&](int i) {
    if (i < cnt) {
      const int64_t n = row(i), o = LOWER ? n - 1 : n + 1;
      if (q < J) {
        ring.copy1(i, q, A + n * J + q);
        ring.copy1(i, J + q, B + n * J + q);
      }
      for (int e = G - 1 - q; e < 3; e += G) {
        const double *src = e == 0 ? bZ + n : (e == 1 ? t + n : t + ((o >= 0 && o < a.N) ? o : n));
        ring.copy1(i, oX + e, src);
      }
    }
    ring.commit();
  };
#pragma unroll
  for (int s = 0; s < kSweepDepth; ++s) issue(s);
  for (int i = 0; i < trips; ++i) {
    const bool live = i < cnt;
    const int64_t n = live ? row(i) : 0;
    const int64_t o = LOWER ? n - 1 : n + 1;
    const bool nb   = live && o >= 0 && o < a.N;
    ring.acquire();
    const double *rb = ring.stage(i);
    double av[J], bv[J];
    lds_row<J>(rb, av);
    lds_row<J>(rb + J, bv);
    const double bz0   = rb[oX];
    const double t_cur = rb[oX + 1], t_oth = rb[oX + 2];
    if (act) xch[q] = nb ? exp(cc * sgn * (t_oth - t_cur)) : 1.0;
    ring.release();  // also publishes xch
    issue(i + kSweepDepth);
    double p[J];
    lds_row<J>(xch, p);
    if (nb) {
      double bz = (q == J) ? bz0 : 0.0;
#pragma unroll
      for (int j = 0; j < J; ++j) bz = fma(bv[j], Gq[j], bz);
#pragma unroll
      for (int j = 0; j < J; ++j) Gq[j] = p[j] * fma(-av[j], bz, Gq[j]);
    }
  }
  if (has && q <= J) {
    const int64_t leaf = LOWER ? a.C - 1 - chunk : chunk;
    double *e          = a.leaves + (b * a.node_stride + leaf) * (int64_t)(J * J + J);
    if (q < J) {
#pragma unroll
      for (int j = 0; j < J; ++j) e[j * J + q] = Gq[j];
    } else {
#pragma unroll
      for (int j = 0; j < J; ++j) e[J * J + j] = Gq[j];
    }
  }
}

// =================================================================================================================
// Multi-right-hand-side solves (K > 1): one warp per (series, chunk, tile of 32 columns); lane k of the warp
// carries the J-vector state of ITS column, so the columns never talk to each other.  The row's A / B entries and
// exp(c dt) are fetched / evaluated once per row by lanes 0..J-1 and broadcast with shuffles.  Y and Z accesses
// are coalesced across the lanes.  Summary: the same kernel over J + K columns (J unit vectors, then the real ones).
// =================================================================================================================
template <int J, bool LOWER, bool SUMMARY>
__global__ void __launch_bounds__(128) csweep_kernel(SweepArgs a) {
  static_assert(J <= 32, "one lane per row element");
  const int lane       = threadIdx.x & 31;
  const int ncol       = SUMMARY ? J + a.K : a.K;
  const int tiles      = (ncol + 31) / 32;
  const int64_t warp   = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (warp >= a.batch * a.C * tiles) return;  // whole warps only
  const int tile       = (int)(warp % tiles);
  const int64_t chunk  = (warp / tiles) % a.C;
  const int64_t b      = warp / ((int64_t)tiles * a.C);
  const int64_t n0     = chunk * a.L;
  const int64_t n1     = n0 + a.L < a.N ? n0 + a.L : a.N;
  const int q          = tile * 32 + lane;  // my column
  const bool have      = q < ncol;
  const bool basis     = SUMMARY && q < J;
  const int k          = SUMMARY ? q - J : q;  // real column index
  const double *t      = a.t + b * a.t_bs;
  const double *A = a.A + b * a.N * J, *B = a.B + b * a.N * J;
  const double *Y = a.Y + b * a.y_bs;
  double *Z = a.Z + b * a.N * (int64_t)a.K;
  double *F = (!SUMMARY && a.F) ? a.F + b * a.N * (int64_t)J * a.K : nullptr;
  const double cc = lane < J ? (a.c + b * a.c_bs)[lane] : 0.0;
  const int64_t leaf = LOWER ? chunk : a.C - 1 - chunk;
  double Fk[J];
#pragma unroll
  for (int j = 0; j < J; ++j) Fk[j] = (basis && j == q) ? 1.0 : 0.0;
  if (!SUMMARY && a.carry && have) {
    const double *st = a.carry + (b * a.node_stride + leaf) * (int64_t)(J * a.K) + (int64_t)J * k;
#pragma unroll
    for (int j = 0; j < J; ++j) Fk[j] = st[j];
  }
  const int cnt = (int)(n1 - n0);
  auto row = [&](int i) { return LOWER ? n0 + i : n1 - 1 - i; };
  double tn = 0.0, to = 0.0, an = 0.0, bn = 0.0, yn = 0.0;
  auto load = [&](int i) {
    const int64_t n = row(i), o = LOWER ? n - 1 : n + 1;
    tn = t[n];
    to = (o >= 0 && o < a.N) ? t[o] : tn;
    an = lane < J ? A[n * J + lane] : 0.0;
    bn = lane < J ? B[n * J + lane] : 0.0;
    yn = (have && !basis) ? Y[n * (int64_t)a.K + k] : 0.0;
  };
  if (cnt > 0) load(0);
  for (int i = 0; i < cnt; ++i) {
    const int64_t n    = row(i);
    const double t_cur = tn, t_oth = to, a_l = an, b_l = bn, y = yn;
    if (i + 1 < cnt) load(i + 1);
    const bool edge  = LOWER ? (n == 0) : (n == a.N - 1);
    const double dt  = LOWER ? t_oth - t_cur : t_cur - t_oth;
    const double p_l = edge ? 1.0 : exp(cc * dt);
    if (F && have) {
#pragma unroll
      for (int j = 0; j < J; ++j) F[n * (int64_t)J * a.K + (int64_t)J * k + j] = Fk[j];
    }
    double s = 0.0;
#pragma unroll
    for (int j = 0; j < J; ++j) {
      const double pj = __shfl_sync(0xffffffffu, p_l, j);
      const double aj = __shfl_sync(0xffffffffu, a_l, j);
      Fk[j]           = pj * Fk[j];
      s               = fma(aj, Fk[j], s);
    }
    const double z = y - s;
    if (!SUMMARY && have) Z[n * (int64_t)a.K + k] = z;
#pragma unroll
    for (int j = 0; j < J; ++j) {
      const double bj = __shfl_sync(0xffffffffu, b_l, j);
      Fk[j]           = fma(bj, z, Fk[j]);
    }
  }
  if (SUMMARY && have) {
    double *e = a.leaves + (b * a.node_stride + leaf) * (int64_t)(J * J + J * a.K);
    if (basis) {
#pragma unroll
      for (int j = 0; j < J; ++j) e[j * J + q] = Fk[j];
    } else {
#pragma unroll
      for (int j = 0; j < J; ++j) e[J * J + j + J * k] = Fk[j];
    }
  }
}

}  // namespace c2b
