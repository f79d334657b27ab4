// factor: lane-group apply and summary kernels (compile-time J).  Same arithmetic as factor.cuh, lane c of a
// group owning column c of the J x J state.
#pragma once
#include "factor.cuh"
#include "group.cuh"

namespace c2b {

constexpr int kGroupThreads = 64;   // threads per CTA of the lane-group kernels

// rows per staged tile (group.cuh); the apply pass collects SR workspace rows (SR*J*J doubles) in shared memory
// before writing them out with wide stores (shared memory per CTA decides how many CTAs fit on an SM)
template <int J>
struct FactorTile {
  static constexpr int R  = J <= 8 ? 8 : 2;
  static constexpr int SR = J <= 8 ? 4 : 2;
};

// tile buffer: U[R*J] V[R*J] | t[R+2] a[R]
// DIRECT: lane c writes its column of a workspace row (J contiguous doubles at offset J*c of the row) straight from
// registers with 16-byte stores instead of collecting SR rows in shared memory first (A/B switch C2B_S_DIRECT).
// BULK: the lanes collect BR rows in one of two shared-memory buffers (16-byte stores of their columns) and lane 0 of the
// group hands the finished buffer to the TMA engine (one bulk store of BR * J * J doubles); needs J even and S 16-byte
// aligned (the launcher checks), otherwise the DIRECT form runs.
template <int J, bool DIRECT = false, bool BULK = false>
__global__ void __launch_bounds__(kGroupThreads, C2B_MINB_FACTOR) gfactor_apply_kernel(FactorArgs g) {
  pdl_prologue();
  static_assert(!(DIRECT && BULK), "one way of writing the workspace rows");
  constexpr int BR = 2;  // rows per bulk store
  constexpr int G  = GroupOf<J>::G;
  constexpr int GP = kGroupThreads / G;  // groups per CTA
  constexpr int R  = FactorTile<J>::R;
  constexpr int JE = (J + 1) & ~1;
  constexpr int oV = R * J, oT = 2 * R * J, oA = oT + R + 2;
  constexpr int TB = (oA + R + 1) & ~1;
  constexpr int GS = GTileStride<G, TB, 2>::value;
  constexpr int SR = FactorTile<J>::SR;
  constexpr int SS = GTileStride<G, SR * J * J, 1>::value;
  __shared__ __align__(16) double stile[GP * GS];
  __shared__ __align__(16) double sprop[GP * GPad<G, R * JE>::value];
  __shared__ __align__(16) double sxch[GP * GPad<G, 2 * JE>::value];
  __shared__ double scc[GP * JE];
  constexpr int BS = GTileStride<G, 2 * BR * J * J, 1>::value;
  __shared__ __align__(16) double srows[DIRECT ? 2 : (BULK ? GP * BS : GP * SS)];  // S rows on their way out
  const int c          = threadIdx.x & (G - 1);
  const int gin        = threadIdx.x / G;
  const int64_t unit   = (int64_t)blockIdx.x * GP + gin;
  const bool has       = unit < g.batch * g.C;
  const bool act       = has && c < J;
  const int cj         = c < J ? c : 0;
  const int64_t chunk  = has ? unit % g.C : 0, b = has ? unit / g.C : 0;
  const int64_t n0     = chunk * g.L;
  const int64_t n1     = has ? (n0 + g.L < g.N ? n0 + g.L : g.N) : n0;
  const double *t      = g.t + b * g.t_bs;
  const int64_t NJ     = g.N * J;
  const double *a = g.a + b * g.N, *U = g.U + b * NJ, *V = g.V + b * NJ;
  double *d = g.d + b * g.N, *W = g.W + b * NJ;
  double *S = g.S ? g.S + b * NJ * J : nullptr;
  double *gbuf = stile + gin * GS, *pb = sprop + gin * GPad<G, R * JE>::value, *xch = sxch + gin * GPad<G, 2 * JE>::value, *sbuf = srows + (DIRECT ? 0 : gin * (BULK ? BS : SS));
  const GCopy<G> cp{c};
  const bool wide  = (J % 2 == 0) && (((reinterpret_cast<uintptr_t>(U) | reinterpret_cast<uintptr_t>(V)) & 15) == 0);
  const bool swide = (J % 2 == 0) && ((reinterpret_cast<uintptr_t>(S) & 15) == 0);
  const double cc  = act ? (g.c + b * g.c_bs)[c] : 0.0;
  PropPairs<J> pairs;
  if (c < J) scc[gin * JE + c] = cc;
  __syncwarp();
  pairs.init(scc + gin * JE, c);
  double Sg[J];  // column c of Sigma: Sg[j] = Sigma(j, c)
#pragma unroll
  for (int j = 0; j < J; ++j) Sg[j] = 0.0;
  if (g.carry && act) {
    const double *st = g.carry + (b * g.node_stride + chunk) * (int64_t)(J * J);  // X = -Sigma, row-major, symmetric
#pragma unroll
    for (int j = 0; j < J; ++j) Sg[j] = -st[j * J + c];
  }
  const int cnt    = (int)(n1 - n0);
  const int ntile  = (cnt + R - 1) / R;
  const int ttrips = warp_max(ntile);
  auto issue = [&](int k) {
    if (k < ntile) {
      const int rows   = cnt - k * R < R ? cnt - k * R : R;
      const int64_t lo = n0 + (int64_t)k * R;
      double *tb       = gbuf + (k & 1) * TB;
      const bool fast   = GCopy<G>::template full_tile<R>(rows, lo, g.tmin, g.tmax);
      cp.template vec<R * J>(tb, U + lo * J, rows * J, wide, fast);
      cp.template vec<R * J>(tb + oV, V + lo * J, rows * J, wide, fast);
      cp.template tim<R + 2>(tb + oT, t, lo - 1, rows + 2, g.tmin, g.tmax, fast);
      cp.template sca<R>(tb + oA, a + lo, rows, fast);
    }
    cp_async_commit();
  };
  issue(0);
  issue(1);
  bool alive = has;
  int bsel   = 0;  // BULK: the buffer being filled
  for (int k = 0; k < ttrips; ++k) {
    cp_async_wait<1>();
    __syncwarp();
    const double *tb = gbuf + (k & 1) * TB;
    const int rows   = cnt - k * R < R ? (cnt - k * R > 0 ? cnt - k * R : 0) : R;
    const int64_t nt = n0 + (int64_t)k * R;
    pairs.template eval<R, JE>(pb, c, cc, [&](int r) { return tb[oT + r] - tb[oT + r + 1]; });  // dt = 0 at n = 0
    __syncwarp();
    int stored = 0;  // S rows of this tile that are to be written
#pragma unroll
    for (int r = 0; r < R; ++r) {
      if (r >= rows) alive = false;  // this group's chunk is done; keep the warp converged
      const int64_t n = nt + r;
      double p[J], u[J];
      lds_row<J>(pb + r * JE, p);
      lds_row<J>(tb + r * J, u);
      const double p_c = act ? pb[r * JE + cj] : 1.0;
      const double u_c = act ? tb[r * J + cj] : 0.0;
      const double v_c = act ? tb[oV + r * J + cj] : 0.0;
      const double a_n = tb[oA + r];
      double St[J];
#pragma unroll
      for (int j = 0; j < J; ++j) {
        St[j] = p[j] * Sg[j];
        Sg[j] = St[j] * p_c;
      }
      if (DIRECT) {
        if (S && act && alive) {
          double *dst = S + n * (int64_t)(J * J) + J * c;  // element (j, c) at j + J*c
          if (swide) {
#pragma unroll
            for (int j = 0; j + 1 < J; j += 2) *reinterpret_cast<double2 *>(dst + j) = make_double2(St[j], St[j + 1]);
          } else {
#pragma unroll
            for (int j = 0; j < J; ++j) dst[j] = St[j];
          }
        }
      } else if (BULK) {
        if (S && act) {
          double *dst = sbuf + bsel * (BR * J * J) + (r % BR) * J * J + J * c;
#pragma unroll
          for (int j = 0; j + 1 < J; j += 2) *reinterpret_cast<double2 *>(dst + j) = make_double2(St[j], St[j + 1]);
        }
        if ((r + 1) % BR == 0) fence_proxy_async_smem();  // my rows of the buffer, before the barrier below
      } else if (S && act) {
#pragma unroll
        for (int j = 0; j < J; ++j) sbuf[(r % SR) * J * J + J * c + j] = St[j];  // element (j, c) at j + J*c
      }
      if (alive) stored = r % (BULK ? BR : SR) + 1;
      double tmp = 0.0;
#pragma unroll
      for (int j = 0; j < J; ++j) tmp = fma(u[j], Sg[j], tmp);
      const double dn = a_n - gsum<G>(tmp * u_c);
      if (alive && c == G - 1) d[n] = dn;
      if (alive && n > g.tmin && dn <= 0.0) {
        if (c == 0) record_failure(g.status ? g.status + b : nullptr, n);
        alive = false;  // forward.hpp:128: later rows stay untouched (lanes keep participating in the barriers)
      }
      // lanes without a column divide 1 by d: a zero numerator sends the whole warp through the slow path of the division
      // on EVERY row (r02j: 27 % of the kernel's instructions at J = 6)
      const double w_c = (act ? v_c - tmp : 1.0) / dn;
      if (alive && act) W[n * J + c] = w_c;
      double *xw = xch + (r & 1) * JE;
      if (c < J) xw[c] = w_c;
      __syncwarp();
      double w[J];
      lds_row<J>(xw, w);
      const double dw = dn * w_c;
#pragma unroll
      for (int j = 0; j < J; ++j) Sg[j] = fma(w[j], dw, Sg[j]);
      if (BULK && (r + 1) % BR == 0) {
        // the barrier after the exchange of w made the buffer's rows visible; rows past a failure stay unwritten
        if (c == 0) {
          if (S && stored > 0) {
            bulk_store(S + (nt + r + 1 - BR) * (int64_t)(J * J), sbuf + bsel * (BR * J * J), (uint32_t)(stored * J * J * sizeof(double)));
          }
          bulk_commit();
          bulk_wait_read<1>();  // the other buffer (handed over BR rows ago) may be written again
        }
        stored = 0;
        bsel ^= 1;  // (tiles of two rows, J > 8, hand over one buffer per tile: the choice cannot hang on r)
        __syncwarp();
      }
      if (!DIRECT && !BULK && (r + 1) % SR == 0) {
        // flush the collected workspace rows (the barrier above made them visible)
        if (S && stored > 0) {
          double *dst     = S + (nt + r + 1 - SR) * (int64_t)(J * J);
          const int count = stored * J * J;
          if (swide) {
            for (int e = 2 * c; e < count; e += 2 * G) *reinterpret_cast<double2 *>(dst + e) = *reinterpret_cast<const double2 *>(sbuf + e);
          } else {
            for (int e = c; e < count; e += G) dst[e] = sbuf[e];
          }
        }
        stored = 0;
        __syncwarp();
      }
    }
    __syncwarp();
    issue(k + 2);
  }
  if (BULK && c == 0) bulk_wait_read<0>();  // shared memory must outlive the engine's reads
}

// summary: lane c owns column c of A, C, Q (row-major element (j,c)); see factor.cuh for the recurrences
// tile buffer: U[R*J] V[R*J] | t[R+2] a[R]
template <int J>
__global__ void __launch_bounds__(kGroupThreads, C2B_MINB_FACTOR) gfactor_summary_kernel(FactorArgs g) {
  pdl_prologue();
  constexpr int G  = GroupOf<J>::G;
  constexpr int GP = kGroupThreads / G;
  constexpr int R  = FactorTile<J>::R;
  constexpr int JE = (J + 1) & ~1;
  constexpr int oV = R * J, oT = 2 * R * J, oA = oT + R + 2;
  constexpr int TB = (oA + R + 1) & ~1;
  constexpr int GS = GTileStride<G, TB, 2>::value;
  __shared__ __align__(16) double stile[GP * GS];
  __shared__ __align__(16) double sprop[GP * GPad<G, R * JE>::value];
  __shared__ __align__(16) double sxch[GP * GPad<G, 4 * JE>::value];
  __shared__ double scc[GP * JE];
  const int c          = threadIdx.x & (G - 1);
  const int gin        = threadIdx.x / G;
  const int64_t unit   = (int64_t)blockIdx.x * GP + gin;
  const bool has       = unit < g.batch * g.C;
  const bool act       = has && c < J;
  const int cj         = c < J ? c : 0;
  const int64_t chunk  = has ? unit % g.C : 0, b = has ? unit / g.C : 0;
  const int64_t n0     = chunk * g.L;
  const int64_t n1     = has ? (n0 + g.L < g.N ? n0 + g.L : g.N) : n0;
  const double *t      = g.t + b * g.t_bs;
  const int64_t NJ     = g.N * J;
  const double *a = g.a + b * g.N, *U = g.U + b * NJ, *V = g.V + b * NJ;
  double *gbuf = stile + gin * GS, *pb = sprop + gin * GPad<G, R * JE>::value, *xch = sxch + gin * GPad<G, 4 * JE>::value;
  const GCopy<G> cp{c};
  const bool wide = (J % 2 == 0) && (((reinterpret_cast<uintptr_t>(U) | reinterpret_cast<uintptr_t>(V)) & 15) == 0);
  const double cc = act ? (g.c + b * g.c_bs)[c] : 0.0;
  PropPairs<J> pairs;
  if (c < J) scc[gin * JE + c] = cc;
  __syncwarp();
  pairs.init(scc + gin * JE, c);
  double A[J], Cm[J], Q[J];  // column c
#pragma unroll
  for (int j = 0; j < J; ++j) {
    A[j]  = (j == c) ? 1.0 : 0.0;
    Cm[j] = 0.0;
    Q[j]  = 0.0;
  }
  const int cnt    = (int)(n1 - n0);
  const int ntile  = (cnt + R - 1) / R;
  const int ttrips = warp_max(ntile);
  auto issue = [&](int k) {
    if (k < ntile) {
      const int rows   = cnt - k * R < R ? cnt - k * R : R;
      const int64_t lo = n0 + (int64_t)k * R;
      double *tb       = gbuf + (k & 1) * TB;
      const bool fast   = GCopy<G>::template full_tile<R>(rows, lo, g.tmin, g.tmax);
      cp.template vec<R * J>(tb, U + lo * J, rows * J, wide, fast);
      cp.template vec<R * J>(tb + oV, V + lo * J, rows * J, wide, fast);
      cp.template tim<R + 2>(tb + oT, t, lo - 1, rows + 2, g.tmin, g.tmax, fast);
      cp.template sca<R>(tb + oA, a + lo, rows, fast);
    }
    cp_async_commit();
  };
  issue(0);
  issue(1);
  for (int k = 0; k < ttrips; ++k) {
    cp_async_wait<1>();
    __syncwarp();
    const double *tb = gbuf + (k & 1) * TB;
    const int rows   = cnt - k * R < R ? (cnt - k * R > 0 ? cnt - k * R : 0) : R;
    pairs.template eval<R, JE>(pb, c, cc, [&](int r) { return tb[oT + r] - tb[oT + r + 1]; });  // dt = 0 at n = 0
    __syncwarp();
#pragma unroll
    for (int r = 0; r < R; ++r) {
      // past the end of the chunk the row is the identity: p = 1, u = v = 0, a = 1
      const bool live = r < rows;
      double p[J], u[J];
      lds_row<J>(pb + r * JE, p);
      lds_row<J>(tb + r * J, u);
      const double p_c = (live && act) ? pb[r * JE + cj] : 1.0;
      const double u_c = (live && act) ? tb[r * J + cj] : 0.0;
      const double v_c = (live && act) ? tb[oV + r * J + cj] : 0.0;
      const double a_n = live ? tb[oA + r] : 1.0;
      double q_c = 0.0, r_c = 0.0;
      if (live) {
#pragma unroll
        for (int j = 0; j < J; ++j) {
          A[j] *= p[j];
          Cm[j] *= p[j] * p_c;
          q_c = fma(Cm[j], u[j], q_c);  // C symmetric: column c dotted with u = (C u)_c
          r_c = fma(A[j], u[j], r_c);   // (A^T u)_c
        }
      }
      const double gam  = a_n + gsum<G>(u_c * q_c);
      const double igam = 1.0 / gam;
      const double g_c  = (q_c + v_c) * igam;
      double *xw = xch + (r & 1) * 2 * JE;
      if (c < J) {
        xw[c]      = g_c;
        xw[JE + c] = r_c;
      }
      __syncwarp();
      double gv[J], rv[J];
      lds_row<J>(xw, gv);
      lds_row<J>(xw + JE, rv);
      const double ggc = gam * g_c, rgc = r_c * igam;
      if (live) {
#pragma unroll
        for (int j = 0; j < J; ++j) {
          A[j]  = fma(-gv[j], r_c, A[j]);
          Cm[j] = fma(-gv[j], ggc, Cm[j]);
          Q[j]  = fma(rv[j], rgc, Q[j]);
        }
      }
    }
    __syncwarp();
    issue(k + 2);
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
