// factor_rev: lane-group apply and summary kernels (compile-time J <= 16).  Lane c owns column c of G = sym(bS);
// the inputs (for the apply pass including the J x J workspace rows) are staged in per-group tiles of rows
// (group.cuh), the next tile in flight while the current one is consumed.
// Recurrences: see factor.cuh.
#pragma once
#include "factor.cuh"
#include "group.cuh"

namespace c2b {

constexpr int kGroupThreadsR = 64;

// WRITE: produce the outputs (apply); otherwise only carry G (and the J rows of the congruence matrix M, lane k
// holding row k) through the chunk (summary).  The summary never needs the workspace S: S enters only bU and the
// bc / bt terms (reverse.hpp:47-60), not the recursion of G.
// tile buffer (rows walked downwards): [S[R*J*J]] U[R*J] W[R*J] bW[R*J] | t[R+2] d[R] bd[R]
// BULK (apply pass, J even, S 16-byte aligned): the workspace rows of a tile, one contiguous run of R * J * J doubles, come
// in by one bulk copy of the TMA engine per group and tile (mbarrier-tracked) instead of R * J * J / 2 16-byte cp.async.
template <int J, bool WRITE, bool BULK = false>
__global__ void __launch_bounds__(kGroupThreadsR, C2B_MINB_FREV) gfactor_rev_kernel(FactorRevArgs g) {
  pdl_prologue();
  static_assert(!BULK || (WRITE && J % 2 == 0), "bulk loads are for the workspace rows of the apply pass");
  constexpr int G  = GroupOf<J>::G;
  constexpr int GP = kGroupThreadsR / G;
  constexpr int R  = J <= 7 ? (WRITE ? C2B_FREV_TILE : 8) : (J <= 8 ? (WRITE ? 2 : 8) : 2);  // static shared memory stays under 48 KB
  constexpr int JE = (J + 1) & ~1;
  constexpr int JJ = WRITE ? J * J : 0;
  constexpr int oU = R * JJ, oW = oU + R * J, oB = oW + R * J, oT = oB + R * J, oD = oT + R + 2, oE = oD + R;
  constexpr int TB = (oE + R + 1) & ~1;
  constexpr int GS = GTileStride<G, TB, 2>::value;
  __shared__ __align__(16) double stile[GP * GS];
  __shared__ __align__(16) double sprop[GP * R * JE];  // (dense: padded, the J = 4 kernels drop from 6 to 5 CTAs per SM, r02f)
  __shared__ __align__(16) double sxch[GP * 2 * JE];
  __shared__ double scc[GP * JE];
  __shared__ uint64_t sbar[BULK ? GP * 2 : 1];
  const int c         = threadIdx.x & (G - 1);
  const int gin       = threadIdx.x / G;
  if (BULK) {
    if (c == 0) {
      mbar_init(sbar + gin * 2, 1);
      mbar_init(sbar + gin * 2 + 1, 1);
      fence_proxy_async_smem();
    }
    __syncwarp();
  }
  const int64_t unit  = (int64_t)blockIdx.x * GP + gin;
  const bool has      = unit < g.batch * g.C;
  const bool act      = has && c < J;
  const int cj        = c < J ? c : 0;
  const int64_t chunk = has ? unit % g.C : 0, b = has ? unit / g.C : 0;
  const int64_t n0    = chunk * g.L;
  const int64_t n1    = has ? (n0 + g.L < g.N ? n0 + g.L : g.N) : n0;
  const double *t     = g.t + b * g.t_bs;
  const int64_t NJ    = g.N * J;
  const double *U = g.U + b * NJ, *W = g.W + b * NJ, *S = g.S + b * NJ * J, *bW = g.bW + b * NJ;
  const double *d = g.d + b * g.N, *bd = g.bd + b * g.N;
  double *bt = WRITE ? g.bt + b * g.N : nullptr, *ba = WRITE ? g.ba + b * g.N : nullptr;
  double *bU = WRITE ? g.bU + b * NJ : nullptr, *bV = WRITE ? g.bV + b * NJ : nullptr;
  double *gbuf = stile + gin * GS, *pb = sprop + gin * R * JE, *xch = sxch + gin * 2 * JE;
  const GCopy<G> cp{c};
  const bool wide  = (J % 2 == 0) && (((reinterpret_cast<uintptr_t>(U) | reinterpret_cast<uintptr_t>(W) | reinterpret_cast<uintptr_t>(bW)) & 15) == 0);
  const bool swide = (J % 2 == 0) && ((reinterpret_cast<uintptr_t>(S) & 15) == 0);
  const double cc  = act ? (g.c + b * g.c_bs)[c] : 0.0;
  PropPairs<J> pairs;
  if (c < J) scc[gin * JE + c] = cc;
  __syncwarp();
  pairs.init(scc + gin * JE, c);
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
  const bool acc = g.acc != 0;
  const int cnt    = (int)(n1 - n0);
  const int ntile  = (cnt + R - 1) / R;
  const int ttrips = warp_max(ntile);
  // tile k = rows [lo, lo + rows), the short last tile aligned to the end of the buffer
  auto issue = [&](int k) {
    if (k < ntile) {
      const int rows   = cnt - k * R < R ? cnt - k * R : R;
      const int64_t lo = n1 - (int64_t)k * R - rows;
      const int pad    = R - rows;
      double *tb       = gbuf + (k & 1) * TB;
      const bool fast   = GCopy<G>::template full_tile<R>(rows, lo, g.tmin, g.tmax);
      if (BULK) {
        if (c == 0) {
          const uint32_t bytes = (uint32_t)(rows * J * J * sizeof(double));
          mbar_expect_tx(sbar + gin * 2 + (k & 1), bytes);
          bulk_load(tb + pad * J * J, S + lo * (int64_t)(J * J), bytes, sbar + gin * 2 + (k & 1));
        }
      } else if (WRITE) {
        cp.template vec<R * J * J>(tb + pad * J * J, S + lo * (int64_t)(J * J), rows * J * J, swide, fast);
      }
      cp.template vec<R * J>(tb + oU + pad * J, U + lo * J, rows * J, wide, fast);
      cp.template vec<R * J>(tb + oW + pad * J, W + lo * J, rows * J, wide, fast);
      cp.template vec<R * J>(tb + oB + pad * J, bW + lo * J, rows * J, wide, fast);
      cp.template tim<R + 2>(tb + oT + pad, t, lo - 1, rows + 2, g.tmin, g.tmax, fast);
      cp.template sca<R>(tb + oD + pad, d + lo, rows, fast);
      cp.template sca<R>(tb + oE + pad, bd + lo, rows, fast);
    }
    cp_async_commit();
  };
  issue(0);
  issue(1);
  for (int k = 0; k < ttrips; ++k) {
    cp_async_wait<1>();
    if (BULK && k < ntile) mbar_wait(sbar + gin * 2 + (k & 1), (uint32_t)((k >> 1) & 1));
    __syncwarp();
    const double *tb = gbuf + (k & 1) * TB;
    const int rows   = cnt - k * R < R ? (cnt - k * R > 0 ? cnt - k * R : 0) : R;
    const int64_t nt = n1 - 1 - (int64_t)k * R;  // first row of the tile in visiting order
    double *bat = WRITE ? ba + nt : nullptr, *btt = WRITE ? bt + nt : nullptr;
    double *bUt = WRITE ? bU + nt * J + cj : nullptr, *bVt = WRITE ? bV + nt * J + cj : nullptr;
    // dt = t_{n-1} - t_n, 0 at n = 0
    pairs.template eval<R, JE>(pb, c, cc, [&](int r) { return tb[oT + R - 1 - r] - tb[oT + R - r]; });
    __syncwarp();
    // also independent of the state: bW_n / d_n of the tile's rows (a division each, kept off the dependent chain)
    double bwd[R], cbp[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int s = R - 1 - r;
      bwd[r]      = (act && r < rows) ? tb[oB + s * J + cj] / tb[oD + s] : 0.0;
      cbp[r]      = 0.0;
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int s     = R - 1 - r;
      const bool live = r < rows;
      const int64_t n = nt - r;
      const bool nb   = live && n > g.tmin;
      double w[J], u[J], p[J], Sn[J];  // Sn[j] = S_n(j, c)
      lds_row<J>(tb + oU + s * J, u);
      lds_row<J>(tb + oW + s * J, w);
      lds_row<J>(pb + r * JE, p);
      if (WRITE) lds_row<J>(tb + s * J * J + J * cj, Sn);
      const double u_c  = act ? tb[oU + s * J + cj] : 0.0;
      const double w_c  = act ? tb[oW + s * J + cj] : 0.0;
      const double p_c  = act ? pb[r * JE + cj] : 1.0;
      const double bd_n = tb[oE + s];
      double wG = 0.0;
#pragma unroll
      for (int j = 0; j < J; ++j) wG = fma(w[j], Gc[j], wG);
      const double bv_c = fma(2.0, wG, bwd[r]);
      // ba_n = bd_n + w G w^T - w . bV_n   (reverse.hpp:30-33), one reduction for both
      const double ban = bd_n + gsum<G>(w_c * (wG - bv_c));
      const double y_c = fma(ban, u_c, bv_c);
      double *xy = xch + (r & 1) * JE;
      if (c < J) xy[c] = y_c;
      if (WRITE && live) {
        if (c == G - 1) bat[-r] = ban;
        if (act) bVt[-r * J] = bv_c;
      }
      __syncwarp();
      double y[J];
      lds_row<J>(xy, y);
      double bp = 0.0, xs = 0.0;
      if (nb) {
        if (!WRITE) {
          // M <- M T_n, lane c holding row c: r <- (r - (r.w) u) o p
          double sm = 0.0;
#pragma unroll
          for (int j = 0; j < J; ++j) sm = fma(Mr[j], w[j], sm);
#pragma unroll
          for (int j = 0; j < J; ++j) Mr[j] = fma(-sm, u[j], Mr[j]) * p[j];
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
        if (WRITE) bc_acc = fma(tb[oT + s] - tb[oT + s + 1], bp, bc_acc);
#pragma unroll
        for (int j = 0; j < J; ++j) Gc[j] *= p[j] * p_c;
      }
      if (WRITE) {
        cbp[r] = cc * bp;  // bp = 0 on rows without a predecessor
        if (live && act) out_store(bUt - r * J, nb ? -(xs * p_c) : 0.0, acc);
      }
    }
    if (WRITE) {
      // bt_n = (term handed down by row n+1) - c . bp_n: the tile's R reductions at once, off the dependent chain
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const double fac = gsum<G>(cbp[r]);
        if (r < rows) {
          if (c == 0) out_store(btt - r, pend - fac, acc);
          pend = fac;
        }
      }
    }
    __syncwarp();
    issue(k + 2);
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
