// factor_rev: lane-group apply and summary kernels (compile-time J <= 16).  Lane c owns column c of G = sym(bS);
// every row's inputs (for the apply pass including the J x J workspace row S[n]) arrive through the per-group
// cp.async ring of group.cuh, DEPTH rows ahead of their use.
// Recurrences: see factor.cuh.
#pragma once
#include "factor.cuh"
#include "group.cuh"

namespace c2b {

constexpr int kGroupThreadsR = 64;

// WRITE: produce the outputs (apply); otherwise only carry G (and the J rows of the congruence matrix M, lane k
// holding row k) through the chunk (summary).  The summary never needs the workspace S: S enters only bU and the
// bc / bt terms (reverse.hpp:47-60), not the recursion of G.
template <int J, bool WRITE>
__global__ void __launch_bounds__(kGroupThreadsR) gfactor_rev_kernel(FactorRevArgs g) {
  constexpr int G     = GroupOf<J>::G;
  constexpr int GP    = kGroupThreadsR / G;
  constexpr int JJ    = WRITE ? J * J : 0;
  constexpr int oU    = JJ, oW = JJ + J, oB = JJ + 2 * J, oX = JJ + 3 * J;  // then d, bd, t_n, t_{n-1}
  constexpr int ROWD  = (oX + 4 + 1) & ~1;
  constexpr int DEPTH = (J <= 8) ? 4 : 2;
  constexpr int JE    = (J + 1) & ~1;
  __shared__ __align__(16) double sring[GP * GRing<G, ROWD, DEPTH>::kStride];
  __shared__ __align__(16) double sxch[GP * 2 * JE];
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
  GRing<G, ROWD, DEPTH> ring{sring + gin * GRing<G, ROWD, DEPTH>::kStride, c};
  double *xch         = sxch + gin * 2 * JE;
  const bool s16      = WRITE && (J % 2 == 0) && ((reinterpret_cast<uintptr_t>(S) & 15) == 0);
  const double cc     = act ? (g.c + b * g.c_bs)[c] : 0.0;
  double Gc[J];  // column c of G: Gc[j] = G(j, c)
  double Mr[J];  // summary: row c of M
#pragma unroll
  for (int j = 0; j < J; ++j) {
    Gc[j] = 0.0;
    Mr[j] = (j == c) ? 1.0 : 0.0;
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
  // rows are visited from n1-1 down to n0; trip i handles row n1-1-i
  auto issue = [&](int i) {
    if (i < cnt) {
      const int64_t n = n1 - 1 - i;
      if (WRITE) {
        if (s16)
          ring.copy16(i, 0, S + n * (int64_t)(J * J), J * J);
        else
          ring.copy(i, 0, S + n * (int64_t)(J * J), J * J);
      }
      if (c < J) {
        ring.copy1(i, oU + c, U + n * J + c);
        ring.copy1(i, oW + c, W + n * J + c);
        ring.copy1(i, oB + c, bW + n * J + c);
      }
      for (int e = G - 1 - c; e < 4; e += G) {  // the idle lanes of the group first
        const double *src = e == 0 ? d + n : (e == 1 ? bd + n : (e == 2 ? t + n : t + (n >= 1 ? n - 1 : n)));
        ring.copy1(i, oX + e, src);
      }
    }
    ring.commit();
  };
#pragma unroll
  for (int s = 0; s < DEPTH; ++s) issue(s);
  for (int i = 0; i < trips; ++i) {
    const bool live = i < cnt;
    const int64_t n = live ? n1 - 1 - i : 1;
    const bool nb   = live && n >= 1;
    ring.acquire();
    const double *rb = ring.stage(i);
    double w[J], u[J], Sn[J];  // Sn[j] = S_n(j, c)
    lds_row<J>(rb + oU, u);
    lds_row<J>(rb + oW, w);
    if (WRITE) lds_row<J>(rb + (c < J ? J * c : 0), Sn);
    const int cj       = c < J ? c : 0;
    const double u_c   = act ? rb[oU + cj] : 0.0;
    const double w_c   = act ? rb[oW + cj] : 0.0;
    const double bw_c  = act ? rb[oB + cj] : 0.0;
    const double d_n   = live ? rb[oX] : 1.0;
    const double bd_n  = rb[oX + 1];
    const double t_cur = rb[oX + 2], t_prev = rb[oX + 3];
    ring.release();
    issue(i + DEPTH);
    double wG = 0.0;
#pragma unroll
    for (int j = 0; j < J; ++j) wG = fma(w[j], Gc[j], wG);
    const double bv_c = fma(2.0, wG, bw_c / d_n);
    // ba_n = bd_n + w G w^T - w . bV_n   (reverse.hpp:30-33), one reduction for both
    const double ban = bd_n + gsum<G>(w_c * (wG - bv_c));
    const double dt  = t_prev - t_cur;
    const double p_c = nb ? exp(cc * dt) : 1.0;
    const double y_c = fma(ban, u_c, bv_c);
    if (act) {
      xch[c]      = y_c;
      xch[JE + c] = p_c;
    }
    if (WRITE && live) {
      if (c == G - 1) ba[n] = ban;
      if (act) bV[n * J + c] = bv_c;
    }
    __syncwarp();
    double y[J], p[J];
    lds_row<J>(xch, y);
    lds_row<J>(xch + JE, p);
    double bp = 0.0, xs = 0.0;
    if (nb) {
      if (!WRITE) {
        // M <- M T_n, lane c holding row c: r <- (r - (r.w) u) o p
        double s = 0.0;
#pragma unroll
        for (int j = 0; j < J; ++j) s = fma(Mr[j], w[j], s);
#pragma unroll
        for (int j = 0; j < J; ++j) Mr[j] = fma(-s, u[j], Mr[j]) * p[j];
      }
      double gs = 0.0;
#pragma unroll
      for (int j = 0; j < J; ++j) {
        Gc[j] -= 0.5 * (u[j] * y_c + u_c * y[j]);
        if (WRITE) {
          const double xj = fma(ban, u[j], y[j]);  // x = bV + 2 ba u = y + ba u
          xs              = fma(xj, Sn[j], xs);
          gs              = fma(Sn[j], Gc[j], gs);
        }
      }
      bp = 2.0 * p_c * gs;
      if (WRITE) bc_acc = fma(dt, bp, bc_acc);
#pragma unroll
      for (int j = 0; j < J; ++j) Gc[j] *= p[j] * p_c;
    }
    if (WRITE) {
      const double fac = gsum<G>(cc * bp);  // bp = 0 on rows without a predecessor
      if (live) {
        if (act) bU[n * J + c] = nb ? -(xs * p_c) : 0.0;
        if (c == 0) bt[n] = pend - fac;
        pend = fac;
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
      e[c * J + j]         = Mr[j];  // row c of M
      e[J * J + j * J + c] = Gc[j];  // column c of B
    }
  }
}

}  // namespace c2b
