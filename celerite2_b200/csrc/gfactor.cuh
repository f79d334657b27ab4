// factor: lane-group apply and summary kernels (compile-time J).  Same arithmetic as factor.cuh, lane c of a
// group owning column c of the J x J state.
#pragma once
#include "factor.cuh"
#include "group.cuh"

namespace c2b {

constexpr int kGroupThreads = 64;   // threads per CTA of the lane-group kernels

template <int J>
__global__ void __launch_bounds__(kGroupThreads) gfactor_apply_kernel(FactorArgs g) {
  constexpr int G  = GroupOf<J>::G;
  constexpr int GP = kGroupThreads / G;  // groups per CTA
  __shared__ double srow[GP * J * J];
  const int c          = threadIdx.x & (G - 1);
  const int gin        = threadIdx.x / G;
  const int64_t unit   = (int64_t)blockIdx.x * GP + gin;
  const bool has       = unit < g.batch * g.C;
  const bool act       = has && c < J;
  const int64_t chunk  = has ? unit % g.C : 0, b = has ? unit / g.C : 0;
  const int64_t n0     = chunk * g.L;
  const int64_t n1     = has ? (n0 + g.L < g.N ? n0 + g.L : g.N) : n0;
  const double *t      = g.t + b * g.t_bs;
  const int64_t NJ     = g.N * J;
  const double *a = g.a + b * g.N, *U = g.U + b * NJ, *V = g.V + b * NJ;
  double *d = g.d + b * g.N, *W = g.W + b * NJ;
  double *S = g.S ? g.S + b * NJ * J : nullptr;
  double *gbuf       = srow + gin * J * J;
  const double cc    = act ? (g.c + b * g.c_bs)[c] : 0.0;
  double Sg[J];  // column c of Sigma: Sg[j] = Sigma(j, c)
#pragma unroll
  for (int j = 0; j < J; ++j) Sg[j] = 0.0;
  if (g.carry && act) {
    const double *st = g.carry + (b * g.node_stride + chunk) * (int64_t)(J * J);  // X = -Sigma, row-major, symmetric
#pragma unroll
    for (int j = 0; j < J; ++j) Sg[j] = -st[j * J + c];
  }
  // software pipeline: the next row's inputs are loaded while the current row is computed
  double tn = 0.0, tp = 0.0, an = 0.0, un = 0.0, vn = 0.0;
  if (n0 < n1) {
    tn = t[n0];
    tp = n0 > 0 ? t[n0 - 1] : tn;
    an = a[n0];
    un = act ? U[n0 * J + c] : 0.0;
    vn = act ? V[n0 * J + c] : 0.0;
  }
  bool alive       = has;
  const int trips  = warp_max((int)(n1 - n0));
  for (int i = 0; i < trips; ++i) {
    const int64_t n = n0 + i;
    if (n >= n1) alive = false;  // this group's chunk is done; keep the warp converged
    const double t_cur = tn, t_prev = tp, a_cur = an, u_c = un, v_c = vn;
    if (n + 1 < n1) {
      tn = t[n + 1];
      an = a[n + 1];
      un = act ? U[(n + 1) * J + c] : 0.0;
      vn = act ? V[(n + 1) * J + c] : 0.0;
    }
    tp = t_cur;
    const double p_c = (n == 0) ? 1.0 : exp(cc * (t_prev - t_cur));
    double p[J], u[J];
    gall<G, J>(p_c, p);
    gall<G, J>(u_c, u);
    double St[J];
#pragma unroll
    for (int j = 0; j < J; ++j) {
      St[j] = p[j] * Sg[j];
      Sg[j] = St[j] * p_c;
    }
    if (S) gstore_matrix_row<G, J>(gbuf, S + n * (int64_t)J * J, St, c, act, alive);
    double tmp = 0.0;
#pragma unroll
    for (int j = 0; j < J; ++j) tmp = fma(u[j], Sg[j], tmp);
    const double dn = a_cur - gsum<G>(tmp * u_c);
    if (alive && c == 0) d[n] = dn;
    if (alive && n >= 1 && dn <= 0.0) {
      if (c == 0) record_failure(g.status ? g.status + b : nullptr, n);
      alive = false;  // forward.hpp:128: later rows stay untouched (lanes keep participating in the shuffles)
    }
    const double w_c = (v_c - tmp) / dn;
    if (alive && act) W[n * J + c] = w_c;
    double w[J];
    gall<G, J>(w_c, w);
    const double dw = dn * w_c;
#pragma unroll
    for (int j = 0; j < J; ++j) Sg[j] = fma(w[j], dw, Sg[j]);
  }
}

// summary: lane c owns column c of A, C, Q (row-major element (j,c)); see factor.cuh for the recurrences
template <int J>
__global__ void __launch_bounds__(kGroupThreads) gfactor_summary_kernel(FactorArgs g) {
  constexpr int G  = GroupOf<J>::G;
  constexpr int GP = kGroupThreads / G;
  const int c          = threadIdx.x & (G - 1);
  const int gin        = threadIdx.x / G;
  const int64_t unit   = (int64_t)blockIdx.x * GP + gin;
  const bool has       = unit < g.batch * g.C;
  const bool act       = has && c < J;
  const int64_t chunk  = has ? unit % g.C : 0, b = has ? unit / g.C : 0;
  const int64_t n0     = chunk * g.L;
  const int64_t n1     = has ? (n0 + g.L < g.N ? n0 + g.L : g.N) : n0;
  const double *t      = g.t + b * g.t_bs;
  const int64_t NJ     = g.N * J;
  const double *a = g.a + b * g.N, *U = g.U + b * NJ, *V = g.V + b * NJ;
  const double cc = act ? (g.c + b * g.c_bs)[c] : 0.0;
  double A[J], Cm[J], Q[J];  // column c
#pragma unroll
  for (int j = 0; j < J; ++j) {
    A[j]  = (j == c) ? 1.0 : 0.0;
    Cm[j] = 0.0;
    Q[j]  = 0.0;
  }
  double tn = 0.0, tp = 0.0, an = 0.0, un = 0.0, vn = 0.0;
  if (n0 < n1) {
    tn = t[n0];
    tp = n0 > 0 ? t[n0 - 1] : tn;
    an = a[n0];
    un = act ? U[n0 * J + c] : 0.0;
    vn = act ? V[n0 * J + c] : 0.0;
  }
  const int trips = warp_max((int)(n1 - n0));
  for (int i = 0; i < trips; ++i) {
    const int64_t n  = n0 + i;
    const bool live  = n < n1;
    const double t_cur = tn, t_prev = tp, a_cur = live ? an : 1.0, u_c = live ? un : 0.0, v_c = live ? vn : 0.0;
    if (n + 1 < n1) {
      tn = t[n + 1];
      an = a[n + 1];
      un = act ? U[(n + 1) * J + c] : 0.0;
      vn = act ? V[(n + 1) * J + c] : 0.0;
    }
    tp = t_cur;
    // past the end of the chunk the row is the identity: p = 1, u = v = 0, a = 1
    const double p_c = (n == 0 || !live) ? 1.0 : exp(cc * (t_prev - t_cur));
    double p[J], u[J];
    gall<G, J>(p_c, p);
    gall<G, J>(u_c, u);
    double q_c = 0.0, r_c = 0.0;
#pragma unroll
    for (int j = 0; j < J; ++j) {
      A[j] *= p[j];
      Cm[j] *= p[j] * p_c;
      q_c = fma(Cm[j], u[j], q_c);  // C symmetric: column c dotted with u = (C u)_c
      r_c = fma(A[j], u[j], r_c);   // (A^T u)_c
    }
    const double gam  = a_cur + gsum<G>(u_c * q_c);
    const double igam = 1.0 / gam;
    const double g_c  = (q_c + v_c) * igam;
    double gv[J], r[J];
    gall<G, J>(g_c, gv);
    gall<G, J>(r_c, r);
    const double ggc = gam * g_c, rgc = r_c * igam;
#pragma unroll
    for (int j = 0; j < J; ++j) {
      A[j]  = fma(-gv[j], r_c, A[j]);
      Cm[j] = fma(-gv[j], ggc, Cm[j]);
      Q[j]  = fma(r[j], rgc, Q[j]);
    }
  }
  if (act) {
    double *e = g.leaves + (b * g.node_stride + chunk) * (int64_t)(3 * J * J);
#pragma unroll
    for (int j = 0; j < J; ++j) {
      e[j * J + c]             = A[j];
      e[J * J + j * J + c]     = Cm[j];
      e[2 * J * J + j * J + c] = Q[j];
    }
  }
}

}  // namespace c2b
