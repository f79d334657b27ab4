// factor_rev: lane-group apply and summary kernels (compile-time J <= 8).  Lane c owns column c of G = sym(bS);
// the J x J workspace row S[n] is fetched with contiguous accesses through a per-group shared-memory row buffer.
// Recurrences: see factor.cuh.
#pragma once
#include "factor.cuh"
#include "group.cuh"

namespace c2b {

constexpr int kGroupThreadsR = 128;

// WRITE: produce the outputs (apply); otherwise only carry G (and, WITHM, the J rows of the congruence matrix M,
// lane k holding row k) through the chunk (summary).
template <int J, bool WRITE>
__global__ void __launch_bounds__(kGroupThreadsR) gfactor_rev_kernel(FactorRevArgs g) {
  constexpr int G  = GroupOf<J>::G;
  constexpr int GP = kGroupThreadsR / G;
  __shared__ double srow[GP * J * J];
  const int c         = threadIdx.x & (G - 1);
  const int gin       = threadIdx.x / G;
  const int64_t unit  = (int64_t)blockIdx.x * GP + gin;
  const bool has      = unit < g.batch * g.C;
  const bool act      = has && c < J;
  const int64_t chunk = has ? unit % g.C : 0, b = has ? unit / g.C : 0;
  const int64_t n0    = chunk * g.L;
  const int64_t n1    = has ? (n0 + g.L < g.N ? n0 + g.L : g.N) : n0;
  const double *t     = g.t + b * g.t_bs;
  const int64_t NJ    = g.N * J;
  const double *U = g.U + b * NJ, *W = g.W + b * NJ, *S = g.S + b * NJ * J, *bW = g.bW + b * NJ;
  const double *d = g.d + b * g.N, *bd = g.bd + b * g.N;
  double *bt = WRITE ? g.bt + b * g.N : nullptr, *ba = WRITE ? g.ba + b * g.N : nullptr;
  double *bU = WRITE ? g.bU + b * NJ : nullptr, *bV = WRITE ? g.bV + b * NJ : nullptr;
  double *gbuf    = srow + gin * J * J;
  const double cc = act ? (g.c + b * g.c_bs)[c] : 0.0;
  double Gc[J];   // column c of G: Gc[j] = G(j, c)
  double Mr[J];   // summary: row c of M
  double Sn[J];   // column c of the workspace row: Sn[j] = Sn(j, c)
#pragma unroll
  for (int j = 0; j < J; ++j) {
    Gc[j] = 0.0;
    Mr[j] = (j == c) ? 1.0 : 0.0;
    Sn[j] = 0.0;
  }
  const int64_t leaf = g.C - 1 - chunk;
  if (WRITE && g.carry && act) {
    const double *st = g.carry + (b * g.node_stride + leaf) * (int64_t)(J * J);  // row-major, symmetric
#pragma unroll
    for (int j = 0; j < J; ++j) Gc[j] = st[j * J + c];
  }
  double bc_acc = 0.0, pend = 0.0;
  const int cnt   = (int)(n1 - n0);
  const int trips = warp_max(cnt);
  double tn = 0.0, to = 0.0, dn_ = 1.0, bdn = 0.0, un = 0.0, wn = 0.0, bwn = 0.0;
  auto load = [&](int i) {
    const int64_t n = n1 - 1 - i;
    tn  = t[n];
    to  = n >= 1 ? t[n - 1] : tn;
    dn_ = d[n];
    bdn = bd[n];
    un  = act ? U[n * J + c] : 0.0;
    wn  = act ? W[n * J + c] : 0.0;
    bwn = act ? bW[n * J + c] : 0.0;
  };
  if (cnt > 0) load(0);
  for (int i = 0; i < trips; ++i) {
    const bool live    = i < cnt;
    const int64_t n    = live ? n1 - 1 - i : 1;
    const bool nb      = live && n >= 1;
    const double t_cur = tn, t_prev = to, d_n = live ? dn_ : 1.0, bd_n = bdn;
    const double u_c = live ? un : 0.0, w_c = live ? wn : 0.0, bw_c = live ? bwn : 0.0;
    // the workspace row of this step (contiguous fetch, then every lane takes its column)
    gload_matrix_row<G, J>(gbuf, S + n * (int64_t)J * J, Sn, c, act, nb);
    if (i + 1 < cnt) load(i + 1);
    double w[J], u[J];
    gall<G, J>(w_c, w);
    gall<G, J>(u_c, u);
    double wG = 0.0;
#pragma unroll
    for (int j = 0; j < J; ++j) wG = fma(w[j], Gc[j], wG);
    const double bv_c = fma(2.0, wG, bw_c / d_n);
    const double wGw  = gsum<G>(wG * w_c);
    const double wbv  = gsum<G>(w_c * bv_c);
    const double ban  = bd_n + wGw - wbv;
    if (WRITE && live) {
      if (c == 0) ba[n] = ban;
      if (act) bV[n * J + c] = bv_c;
    }
    const double dt  = t_prev - t_cur;
    const double p_c = nb ? exp(cc * dt) : 1.0;
    const double y_c = fma(ban, u_c, bv_c);
    double y[J], p[J];
    gall<G, J>(y_c, y);
    gall<G, J>(p_c, p);
    if (nb) {
      if (!WRITE) {
        // M <- M T_n, lane c holding row c: r <- (r - (r.w) u) o p
        double s = 0.0;
#pragma unroll
        for (int j = 0; j < J; ++j) s = fma(Mr[j], w[j], s);
#pragma unroll
        for (int j = 0; j < J; ++j) Mr[j] = fma(-s, u[j], Mr[j]) * p[j];
      }
      double xs = 0.0, gs = 0.0;
#pragma unroll
      for (int j = 0; j < J; ++j) {
        Gc[j] -= 0.5 * (u[j] * y_c + u_c * y[j]);
        const double xj = fma(ban, u[j], y[j]);  // x = bV + 2 ba u = y + ba u
        xs              = fma(xj, Sn[j], xs);
        gs              = fma(Sn[j], Gc[j], gs);
      }
      const double bp = 2.0 * p_c * gs;
      if (WRITE && act) bU[n * J + c] = -(xs * p_c);
      bc_acc = fma(dt, bp, bc_acc);
#pragma unroll
      for (int j = 0; j < J; ++j) Gc[j] *= p[j] * p_c;
      const double fac = gsum<G>(cc * bp);
      if (WRITE && c == 0) bt[n] = pend - fac;
      pend = fac;
    } else {
      (void)gsum<G>(0.0);  // keep the warp's shuffle sequence uniform
      if (WRITE && live) {
        if (act) bU[n * J + c] = 0.0;
        if (c == 0) bt[n] = pend;
        pend = 0.0;
      }
    }
  }
  if (WRITE) {
    if (act) g.bc_part[(b * g.C + chunk) * J + c] = bc_acc;
    if (has && c == 0) g.bt_edge[b * g.C + chunk] = pend;
  } else if (act) {
    double *e = g.leaves + (b * g.node_stride + leaf) * (int64_t)(2 * J * J);
#pragma unroll
    for (int j = 0; j < J; ++j) {
      e[c * J + j]         = Mr[j];      // row c of M
      e[J * J + j * J + c] = Gc[j];      // column c of B
    }
  }
}

}  // namespace c2b
