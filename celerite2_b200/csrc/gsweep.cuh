// Lane-group kernels of the K == 1 sweeps (solve / matmul, lower / upper) and of their reverse mode.
// apply: lane c of a group holds element c of the J-vector state; summary of the solves: lane q holds the whole
// J-vector of column q (q < J: unit vectors -> transition M, q == J: the real right-hand side -> b).
#pragma once
#include "group.cuh"
#include "sweep.cuh"

namespace c2b {

constexpr int kGroupThreadsS = 64;

constexpr int kTileRows = 8;  // rows per staged tile (group.cuh)

// ---- forward apply -----------------------------------------------------------------------------------------------
// tile buffer: A[R*J] B[R*J] | t[R+2] y[R] z_old[R]
template <int J, bool SOLVE, bool LOWER>
__global__ void __launch_bounds__(kGroupThreadsS, C2B_MINB_SWEEP) gsweep_apply_kernel(SweepArgs a) {
  pdl_prologue();
  constexpr int G   = GroupOf<J>::G;
  constexpr int GP  = kGroupThreadsS / G;
  constexpr int R   = kTileRows;
  constexpr int oB  = R * J, oT = 2 * R * J, oY = oT + R + 2, oZ = oY + R;
  constexpr int TB  = (oZ + (SOLVE ? 0 : R) + 1) & ~1;
  constexpr int GS  = GTileStride<G, TB, 2>::value;
  constexpr int dir = LOWER ? 1 : -1;
  constexpr int JE  = (J + 1) & ~1;
  __shared__ __align__(16) double stile[GP * GS];
  __shared__ double scc[GP * JE];
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
  const double *zdiv = (SOLVE && a.zdiv) ? a.zdiv + b * a.N : nullptr;
  double *gbuf = stile + gin * GS;
  const GCopy<G> cp{c};
  const bool wide = (J % 2 == 0) && (((reinterpret_cast<uintptr_t>(A) | reinterpret_cast<uintptr_t>(B)) & 15) == 0);
  const double cc = act ? (a.c + b * a.c_bs)[c] : 0.0;
  PropPairs<J> pairs;
  if (c < J) scc[gin * JE + c] = cc;
  __syncwarp();
  pairs.init(scc + gin * JE, c);
  double Fc = 0.0;
  if (a.carry && act) {
    const int64_t leaf = LOWER ? chunk : a.C - 1 - chunk;
    Fc                 = a.carry[(b * a.node_stride + leaf) * (int64_t)J + c];
  }
  const int cnt    = (int)(n1 - n0);
  const int ntile  = (cnt + R - 1) / R;
  const int ttrips = warp_max(ntile);
  // tile k (processing order) = rows [lo, lo + rows)
  auto issue = [&](int k) {
    if (k < ntile) {
      const int rows   = cnt - k * R < R ? cnt - k * R : R;
      const int64_t lo = LOWER ? n0 + (int64_t)k * R : n1 - (int64_t)k * R - rows;
      const int pad    = LOWER ? 0 : R - rows;
      double *tb       = gbuf + (k & 1) * TB;
      const bool fast   = GCopy<G>::template full_tile<R>(rows, lo, a.tmin, a.tmax);
      cp.template vec<R * J>(tb + pad * J, A + lo * J, rows * J, wide, fast);
      cp.template vec<R * J>(tb + oB + pad * J, B + lo * J, rows * J, wide, fast);
      cp.template tim<R + 2>(tb + oT + pad, t, lo - 1, rows + 2, a.tmin, a.tmax, fast);
      cp.template sca<R>(tb + oY + pad, Y + lo, rows, fast);
      if (!SOLVE) cp.template sca<R>(tb + oZ + pad, Z + lo, rows, fast);
    }
    cp_async_commit();
  };
  issue(0);
  issue(1);
  const int64_t nf = LOWER ? n0 : n1 - 1;  // first row visited
  for (int k = 0; k < ttrips; ++k) {
    cp_async_wait<1>();
    __syncwarp();
    const double *tb = gbuf + (k & 1) * TB;
    const int rows   = cnt - k * R < R ? (cnt - k * R > 0 ? cnt - k * R : 0) : R;
    // state-independent part of the whole tile first: independent exp evaluations the compiler can interleave
    double p[R];
    eval_props_reg<J, G, R>(pairs, c, cc, [&](int r) {
      const int s = LOWER ? r : R - 1 - r;
      return LOWER ? tb[oT + s] - tb[oT + s + 1] : tb[oT + s + 1] - tb[oT + s + 2];
    }, p);
    double *Ft = F ? F + (nf + (int64_t)dir * k * R) * J + cj : nullptr;
    double *Zt = Z + (nf + (int64_t)dir * k * R);
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int s       = LOWER ? r : R - 1 - r;
      const bool live   = r < rows;
      const double a_c  = (live && act) ? tb[s * J + cj] : 0.0;
      const double b_c  = (live && act) ? tb[oB + s * J + cj] : 0.0;
      const double y    = tb[oY + s];
      const double zold = SOLVE ? 0.0 : tb[oZ + s];
      if (Ft && live && act) Ft[dir * J * r] = Fc;
      Fc             = p[r] * Fc;
      const double sm = gsum<G>(a_c * Fc);
      double z, rr;
      if (SOLVE) {
        z  = y - sm;
        rr = z;
      } else {
        z  = zold + sm;
        rr = y;
      }
      if (live && c == G - 1) Zt[dir * r] = zdiv ? z / zdiv[nf + (int64_t)dir * (k * R + r)] : z;
      Fc = fma(b_c, rr, Fc);
    }
    __syncwarp();
    issue(k + 2);
  }
}

// ---- forward summary, solves: lane q carries column q ----------------------------------------------------------------
// tile buffer: A[R*J] B[R*J] | t[R+2] y[R]; plus one R x J block per group for the propagators of the tile
template <int J, bool LOWER>
__global__ void __launch_bounds__(kGroupThreadsS, C2B_MINB_SWEEP) gsweep_summary_kernel(SweepArgs a) {
  pdl_prologue();
  constexpr int G  = GroupOf<J + 1>::G;  // J + 1 columns
  constexpr int GP = kGroupThreadsS / G;
  constexpr int R  = kTileRows;
  constexpr int JE = (J + 1) & ~1;
  constexpr int oB = R * J, oT = 2 * R * J, oY = oT + R + 2;
  constexpr int TB = (oY + R + 1) & ~1;
  constexpr int GS = GTileStride<G, TB, 2>::value;
  __shared__ __align__(16) double stile[GP * GS];
  __shared__ __align__(16) double sprop[GP * GPad<G, R * JE>::value];
  __shared__ double scc[GP * JE];
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
  double *gbuf = stile + gin * GS, *pb = sprop + gin * GPad<G, R * JE>::value;
  const GCopy<G> cp{q};
  const bool wide = (J % 2 == 0) && (((reinterpret_cast<uintptr_t>(A) | reinterpret_cast<uintptr_t>(B)) & 15) == 0);
  const double cc = act ? (a.c + b * a.c_bs)[q] : 0.0;
  PropPairs<J> pairs;
  if (q < J) scc[gin * JE + q] = cc;
  __syncwarp();
  pairs.init(scc + gin * JE, q);
  double Gq[J];
#pragma unroll
  for (int j = 0; j < J; ++j) Gq[j] = (j == q) ? 1.0 : 0.0;
  const int cnt    = (int)(n1 - n0);
  const int ntile  = (cnt + R - 1) / R;
  const int ttrips = warp_max(ntile);
  auto issue = [&](int k) {
    if (k < ntile) {
      const int rows   = cnt - k * R < R ? cnt - k * R : R;
      const int64_t lo = LOWER ? n0 + (int64_t)k * R : n1 - (int64_t)k * R - rows;
      const int pad    = LOWER ? 0 : R - rows;
      double *tb       = gbuf + (k & 1) * TB;
      const bool fast   = GCopy<G>::template full_tile<R>(rows, lo, a.tmin, a.tmax);
      cp.template vec<R * J>(tb + pad * J, A + lo * J, rows * J, wide, fast);
      cp.template vec<R * J>(tb + oB + pad * J, B + lo * J, rows * J, wide, fast);
      cp.template tim<R + 2>(tb + oT + pad, t, lo - 1, rows + 2, a.tmin, a.tmax, fast);
      cp.template sca<R>(tb + oY + pad, Y + lo, rows, fast);
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
    pairs.template eval<R, JE>(pb, q, cc, [&](int r) {
      const int s = LOWER ? r : R - 1 - r;
      return LOWER ? tb[oT + s] - tb[oT + s + 1] : tb[oT + s + 1] - tb[oT + s + 2];
    });
    __syncwarp();
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int s = LOWER ? r : R - 1 - r;
      double av[J], bv[J], p[J];
      lds_row<J>(tb + s * J, av);
      lds_row<J>(tb + oB + s * J, bv);
      lds_row<J>(pb + r * JE, p);
      const double y = tb[oY + s];
      if (r < rows) {
        double sm = 0.0;
#pragma unroll
        for (int j = 0; j < J; ++j) {
          Gq[j] = p[j] * Gq[j];
          sm    = fma(av[j], Gq[j], sm);
        }
        const double rr = (q == J ? y : 0.0) - sm;
#pragma unroll
        for (int j = 0; j < J; ++j) Gq[j] = fma(bv[j], rr, Gq[j]);
      }
    }
    __syncwarp();
    issue(k + 2);
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

// ---- forward summary, matmuls: diagonal-affine chunk map G -> m o G + b, lane c holding element c ----------------------
// STORE = false: the chunk's summary (m, b) goes to the leaves.  STORE = true: the walk starts from the carried state and
// the state AFTER every row is written to a.F (N x J): the running sums H_m of the rectangular products
// (general_matmul_lower / upper, forward.hpp:285-392).
// tile buffer: B[R*J] | t[R+2] y[R]
template <int J, bool LOWER, bool STORE = false>
__global__ void __launch_bounds__(kGroupThreadsS, C2B_MINB_SWEEP) gsweep_msummary_kernel(SweepArgs a) {
  pdl_prologue();
  constexpr int G  = GroupOf<J>::G;
  constexpr int GP = kGroupThreadsS / G;
  constexpr int R  = kTileRows;
  constexpr int JE = (J + 1) & ~1;
  constexpr int oT = R * J, oY = oT + R + 2;
  constexpr int TB = (oY + R + 1) & ~1;
  constexpr int GS = GTileStride<G, TB, 2>::value;
  __shared__ __align__(16) double stile[GP * GS];
  __shared__ double scc[GP * JE];
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
  const double *B = a.B + b * a.N * J, *Y = a.Y + b * a.y_bs;
  double *gbuf = stile + gin * GS;
  const GCopy<G> cp{c};
  const bool wide = (J % 2 == 0) && ((reinterpret_cast<uintptr_t>(B) & 15) == 0);
  const double cc = act ? (a.c + b * a.c_bs)[c] : 0.0;
  PropPairs<J> pairs;
  if (c < J) scc[gin * JE + c] = cc;
  __syncwarp();
  pairs.init(scc + gin * JE, c);
  double Fc = 0.0, mc = 1.0;
  if (STORE && a.carry && act) {
    const int64_t leaf = LOWER ? chunk : a.C - 1 - chunk;
    Fc                 = a.carry[(b * a.node_stride + leaf) * (int64_t)J + c];
  }
  constexpr int dir = LOWER ? 1 : -1;
  const int64_t nf  = LOWER ? n0 : n1 - 1;
  double *H         = STORE ? a.F + b * a.N * J : nullptr;
  const int cnt    = (int)(n1 - n0);
  const int ntile  = (cnt + R - 1) / R;
  const int ttrips = warp_max(ntile);
  auto issue = [&](int k) {
    if (k < ntile) {
      const int rows   = cnt - k * R < R ? cnt - k * R : R;
      const int64_t lo = LOWER ? n0 + (int64_t)k * R : n1 - (int64_t)k * R - rows;
      const int pad    = LOWER ? 0 : R - rows;
      double *tb       = gbuf + (k & 1) * TB;
      const bool fast   = GCopy<G>::template full_tile<R>(rows, lo, a.tmin, a.tmax);
      cp.template vec<R * J>(tb + pad * J, B + lo * J, rows * J, wide, fast);
      cp.template tim<R + 2>(tb + oT + pad, t, lo - 1, rows + 2, a.tmin, a.tmax, fast);
      cp.template sca<R>(tb + oY + pad, Y + lo, rows, fast);
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
    double p[R];
    eval_props_reg<J, G, R>(pairs, c, cc, [&](int r) {
      const int s = LOWER ? r : R - 1 - r;
      return LOWER ? tb[oT + s] - tb[oT + s + 1] : tb[oT + s + 1] - tb[oT + s + 2];
    }, p);
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int s = LOWER ? r : R - 1 - r;
      if (r < rows) {
        mc *= p[r];
        Fc = fma(act ? tb[s * J + cj] : 0.0, tb[oY + s], p[r] * Fc);
        if (STORE && act) H[(nf + (int64_t)dir * (k * R + r)) * J + c] = Fc;
      }
    }
    __syncwarp();
    issue(k + 2);
  }
  if (!STORE && act) {
    const int64_t leaf = LOWER ? chunk : a.C - 1 - chunk;
    double *e          = a.leaves + (b * a.node_stride + leaf) * (int64_t)(2 * J);
    e[c]               = mc;
    e[J + c]           = Fc;
  }
}

// ---- reverse summary, matmuls ------------------------------------------------------------------------------------------
// tile buffer: A[R*J] | t[R+2] bZ[R]
template <int J, bool LOWER>
__global__ void __launch_bounds__(kGroupThreadsS, C2B_MINB_SWEEP) gsweep_rev_msummary_kernel(SweepRevArgs a) {
  pdl_prologue();
  constexpr int G   = GroupOf<J>::G;
  constexpr int GP  = kGroupThreadsS / G;
  constexpr int R   = kTileRows;
  constexpr int JE  = (J + 1) & ~1;
  constexpr int oT  = R * J, oZ = oT + R + 2;
  constexpr int TB  = (oZ + R + 1) & ~1;
  constexpr int GS  = GTileStride<G, TB, 2>::value;
  constexpr bool UP = !LOWER;
  constexpr int dir = UP ? 1 : -1;
  __shared__ __align__(16) double stile[GP * GS];
  __shared__ double scc[GP * JE];
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
  const double *A = a.A + b * a.N * J, *bZ = a.bZ + b * a.N;
  double *gbuf = stile + gin * GS;
  const GCopy<G> cp{c};
  const bool wide = (J % 2 == 0) && ((reinterpret_cast<uintptr_t>(A) & 15) == 0);
  const double cc = act ? (a.c + b * a.c_bs)[c] : 0.0;
  PropPairs<J> pairs;
  if (c < J) scc[gin * JE + c] = cc;
  __syncwarp();
  pairs.init(scc + gin * JE, c);
  double bF = 0.0, mc = 1.0;
  const int cnt    = (int)(n1 - n0);
  const int ntile  = (cnt + R - 1) / R;
  const int ttrips = warp_max(ntile);
  auto issue = [&](int k) {
    if (k < ntile) {
      const int rows   = cnt - k * R < R ? cnt - k * R : R;
      const int64_t lo = UP ? n0 + (int64_t)k * R : n1 - (int64_t)k * R - rows;
      const int pad    = UP ? 0 : R - rows;
      double *tb       = gbuf + (k & 1) * TB;
      const bool fast   = GCopy<G>::template full_tile<R>(rows, lo, a.tmin, a.tmax);
      cp.template vec<R * J>(tb + pad * J, A + lo * J, rows * J, wide, fast);
      cp.template tim<R + 2>(tb + oT + pad, t, lo - 1, rows + 2, a.tmin, a.tmax, fast);
      cp.template sca<R>(tb + oZ + pad, bZ + lo, rows, fast);
    }
    cp_async_commit();
  };
  issue(0);
  issue(1);
  const int64_t nf    = UP ? n0 : n1 - 1;
  const int64_t nlast = LOWER ? a.tmin : a.tmax;
  for (int k = 0; k < ttrips; ++k) {
    cp_async_wait<1>();
    __syncwarp();
    const double *tb = gbuf + (k & 1) * TB;
    const int rows   = cnt - k * R < R ? (cnt - k * R > 0 ? cnt - k * R : 0) : R;
    double p[R];
    eval_props_reg<J, G, R>(pairs, c, cc, [&](int r) {
      const int s = UP ? r : R - 1 - r;
      return LOWER ? tb[oT + s] - tb[oT + s + 1] : tb[oT + s + 1] - tb[oT + s + 2];
    }, p);
    const int64_t nt = nf + (int64_t)dir * k * R;
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int s = UP ? r : R - 1 - r;
      if (r < rows && nt + dir * r != nlast) {
        mc *= p[r];
        bF = p[r] * fma(act ? tb[s * J + cj] : 0.0, tb[oZ + s], bF);  // bF <- p (bF + a bZ)
      }
    }
    __syncwarp();
    issue(k + 2);
  }
  if (act) {
    const int64_t leaf = LOWER ? a.C - 1 - chunk : chunk;
    double *e          = a.leaves + (b * a.node_stride + leaf) * (int64_t)(2 * J);
    e[c]               = mc;
    e[J + c]           = bF;
  }
}

// ---- reverse apply -----------------------------------------------------------------------------------------------------
// tile buffer: A[R*J] B[R*J] F[R*J] | t[R+2] bZ[R] (Z or Y)[R]
template <int J, bool SOLVE, bool LOWER>
__global__ void __launch_bounds__(kGroupThreadsS, C2B_MINB_SWEEP) gsweep_rev_apply_kernel(SweepRevArgs a) {
  pdl_prologue();
  constexpr int G   = GroupOf<J>::G;
  constexpr int GP  = kGroupThreadsS / G;
  constexpr int R   = kTileRows;
  constexpr int oB  = R * J, oF = 2 * R * J, oT = 3 * R * J, oZ = oT + R + 2, oY = oZ + R;
  constexpr int TB  = (oY + R + 1) & ~1;
  constexpr int GS  = GTileStride<G, TB, 2>::value;
  constexpr bool UP = !LOWER;  // the reverse of a lower sweep walks downwards
  constexpr int dir = UP ? 1 : -1;
  constexpr int JE  = (J + 1) & ~1;
  __shared__ __align__(16) double stile[GP * GS];
  __shared__ double scc[GP * JE];
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
  double *gbuf = stile + gin * GS;
  const GCopy<G> cp{c};
  const bool wide = (J % 2 == 0) && (((reinterpret_cast<uintptr_t>(A) | reinterpret_cast<uintptr_t>(B) | reinterpret_cast<uintptr_t>(F)) & 15) == 0);
  const double cc = act ? (a.c + b * a.c_bs)[c] : 0.0;
  PropPairs<J> pairs;
  if (c < J) scc[gin * JE + c] = cc;
  __syncwarp();
  pairs.init(scc + gin * JE, c);
  double bF = 0.0, bc_acc = 0.0, pend = 0.0;
  if (a.carry && act) {
    const int64_t leaf = LOWER ? a.C - 1 - chunk : chunk;
    bF                 = a.carry[(b * a.node_stride + leaf) * (int64_t)J + c];
  }
  const double sgn = LOWER ? 1.0 : -1.0;
  const int cnt    = (int)(n1 - n0);
  const int ntile  = (cnt + R - 1) / R;
  const int ttrips = warp_max(ntile);
  auto issue = [&](int k) {
    if (k < ntile) {
      const int rows   = cnt - k * R < R ? cnt - k * R : R;
      const int64_t lo = UP ? n0 + (int64_t)k * R : n1 - (int64_t)k * R - rows;
      const int pad    = UP ? 0 : R - rows;
      double *tb       = gbuf + (k & 1) * TB;
      const bool fast   = GCopy<G>::template full_tile<R>(rows, lo, a.tmin, a.tmax);
      cp.template vec<R * J>(tb + pad * J, A + lo * J, rows * J, wide, fast);
      cp.template vec<R * J>(tb + oB + pad * J, B + lo * J, rows * J, wide, fast);
      cp.template vec<R * J>(tb + oF + pad * J, F + lo * J, rows * J, wide, fast);
      cp.template tim<R + 2>(tb + oT + pad, t, lo - 1, rows + 2, a.tmin, a.tmax, fast);
      cp.template sca<R>(tb + oZ + pad, bZ + lo, rows, fast);
      cp.template sca<R>(tb + oY + pad, ZY + lo, rows, fast);
    }
    cp_async_commit();
  };
  issue(0);
  issue(1);
  const int64_t nf    = UP ? n0 : n1 - 1;  // first row visited
  const int64_t nlast = LOWER ? a.tmin : a.tmax;  // the row without a predecessor in this direction
  for (int k = 0; k < ttrips; ++k) {
    cp_async_wait<1>();
    __syncwarp();
    const double *tb = gbuf + (k & 1) * TB;
    const int rows   = cnt - k * R < R ? (cnt - k * R > 0 ? cnt - k * R : 0) : R;
    double p[R];
    eval_props_reg<J, G, R>(pairs, c, cc, [&](int r) {
      const int s = UP ? r : R - 1 - r;
      return LOWER ? tb[oT + s] - tb[oT + s + 1] : tb[oT + s + 1] - tb[oT + s + 2];
    }, p);
    const int64_t nt = nf + (int64_t)dir * k * R;  // first row of this tile in visiting order
    double cbp[R];
    double *bAt = bA + nt * J + cj, *bBt = bB + nt * J + cj, *btt = bt + nt, *bYt = bY + nt;
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int s      = UP ? r : R - 1 - r;
      const bool live  = r < rows;
      const bool nb    = live && (nt + dir * r != nlast);
      const double a_c = (live && act) ? tb[s * J + cj] : 0.0;
      const double b_c = (live && act) ? tb[oB + s * J + cj] : 0.0;
      const double Fn  = (nb && act) ? tb[oF + s * J + cj] : 0.0;
      const double bz0 = tb[oZ + s], zy = tb[oY + s];
      const double dot = gsum<G>(b_c * bF);
      double bz;
      if (SOLVE) {
        bz = bz0 + dot;
        if (live && c == G - 1) bYt[dir * r] = bz;
      } else {
        bz = bz0;
        if (live && c == G - 1) bYt[dir * r] = dot;
      }
      if (live && act) bBt[dir * J * r] = zy * bF;
      const double p_c = p[r];
      const double au  = a_c * bz;
      const double nf_ = SOLVE ? bF - au : bF + au;
      const double pf  = p_c * Fn;
      const double bp  = nb ? Fn * nf_ * p_c : 0.0;
      cbp[r]           = cc * bp;
      if (nb) {
        if (act) bAt[dir * J * r] = SOLVE ? -(bz * pf) : bz * pf;
        bc_acc = fma(LOWER ? tb[oT + s] - tb[oT + s + 1] : tb[oT + s + 1] - tb[oT + s + 2], bp, bc_acc);
        bF     = p_c * nf_;
      } else if (live) {
        if (act) bAt[dir * J * r] = 0.0;
      }
    }
    // bt_n = (term handed over by the previous row) -/+ c . bp_n: the tile's R reductions at once, off the dependent chain
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const double fac = gsum<G>(cbp[r]);  // 0 on rows without a predecessor
      if (r < rows) {
        if (c == 0) btt[dir * r] = pend - sgn * fac;
        pend = sgn * fac;
      }
    }
    __syncwarp();
    issue(k + 2);
  }
  if (act) a.bc_part[(b * a.C + chunk) * J + c] = bc_acc;
  if (has && c == 0) a.bt_edge[b * a.C + chunk] = pend;
}

// ---- reverse summary, solves -----------------------------------------------------------------------------------------
// tile buffer: A[R*J] B[R*J] | t[R+2] bZ[R]; plus the R x J propagator block
template <int J, bool LOWER>
__global__ void __launch_bounds__(kGroupThreadsS, C2B_MINB_SWEEP) gsweep_rev_summary_kernel(SweepRevArgs a) {
  pdl_prologue();
  constexpr int G   = GroupOf<J + 1>::G;
  constexpr int GP  = kGroupThreadsS / G;
  constexpr int R   = kTileRows;
  constexpr int JE  = (J + 1) & ~1;
  constexpr int oB  = R * J, oT = 2 * R * J, oZ = oT + R + 2;
  constexpr int TB  = (oZ + R + 1) & ~1;
  constexpr int GS  = GTileStride<G, TB, 2>::value;
  constexpr bool UP = !LOWER;
  constexpr int dir = UP ? 1 : -1;
  __shared__ __align__(16) double stile[GP * GS];
  __shared__ __align__(16) double sprop[GP * GPad<G, R * JE>::value];
  __shared__ double scc[GP * JE];
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
  double *gbuf = stile + gin * GS, *pb = sprop + gin * GPad<G, R * JE>::value;
  const GCopy<G> cp{q};
  const bool wide = (J % 2 == 0) && (((reinterpret_cast<uintptr_t>(A) | reinterpret_cast<uintptr_t>(B)) & 15) == 0);
  const double cc = act ? (a.c + b * a.c_bs)[q] : 0.0;
  PropPairs<J> pairs;
  if (q < J) scc[gin * JE + q] = cc;
  __syncwarp();
  pairs.init(scc + gin * JE, q);
  double Gq[J];
#pragma unroll
  for (int j = 0; j < J; ++j) Gq[j] = (j == q) ? 1.0 : 0.0;
  const int cnt    = (int)(n1 - n0);
  const int ntile  = (cnt + R - 1) / R;
  const int ttrips = warp_max(ntile);
  auto issue = [&](int k) {
    if (k < ntile) {
      const int rows   = cnt - k * R < R ? cnt - k * R : R;
      const int64_t lo = UP ? n0 + (int64_t)k * R : n1 - (int64_t)k * R - rows;
      const int pad    = UP ? 0 : R - rows;
      double *tb       = gbuf + (k & 1) * TB;
      const bool fast   = GCopy<G>::template full_tile<R>(rows, lo, a.tmin, a.tmax);
      cp.template vec<R * J>(tb + pad * J, A + lo * J, rows * J, wide, fast);
      cp.template vec<R * J>(tb + oB + pad * J, B + lo * J, rows * J, wide, fast);
      cp.template tim<R + 2>(tb + oT + pad, t, lo - 1, rows + 2, a.tmin, a.tmax, fast);
      cp.template sca<R>(tb + oZ + pad, bZ + lo, rows, fast);
    }
    cp_async_commit();
  };
  issue(0);
  issue(1);
  const int64_t nf    = UP ? n0 : n1 - 1;
  const int64_t nlast = LOWER ? a.tmin : a.tmax;
  for (int k = 0; k < ttrips; ++k) {
    cp_async_wait<1>();
    __syncwarp();
    const double *tb = gbuf + (k & 1) * TB;
    const int rows   = cnt - k * R < R ? (cnt - k * R > 0 ? cnt - k * R : 0) : R;
    pairs.template eval<R, JE>(pb, q, cc, [&](int r) {
      const int s = UP ? r : R - 1 - r;
      return LOWER ? tb[oT + s] - tb[oT + s + 1] : tb[oT + s + 1] - tb[oT + s + 2];
    });
    __syncwarp();
    const int64_t nt = nf + (int64_t)dir * k * R;
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int s = UP ? r : R - 1 - r;
      double av[J], bv[J], p[J];
      lds_row<J>(tb + s * J, av);
      lds_row<J>(tb + oB + s * J, bv);
      lds_row<J>(pb + r * JE, p);
      const double bz0 = tb[oZ + s];
      if (r < rows && nt + dir * r != nlast) {
        double bz = (q == J) ? bz0 : 0.0;
#pragma unroll
        for (int j = 0; j < J; ++j) bz = fma(bv[j], Gq[j], bz);
#pragma unroll
        for (int j = 0; j < J; ++j) Gq[j] = p[j] * fma(-av[j], bz, Gq[j]);
      }
    }
    __syncwarp();
    issue(k + 2);
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
// Multi-right-hand-side solves (K > 1) on staged tiles: one CTA per (series, chunk); thread q owns column q and
// carries its J-vector state in registers, so the columns never talk to each other.  The summary pass runs J + K
// columns (J unit vectors -> transition M, then the K right-hand sides -> b), the apply pass K.  The chunk's rows
// are staged in tiles of R rows shared by the whole CTA (A, B rows, time stamps, the tile of Y), the tile's
// exp(c dt) are evaluated once per CTA, and every thread reads a row's A / B / propagator vectors as 16-byte
// shared-memory broadcasts (the first version broadcast them with 6 J shuffles per row: c4 at 3.9 % of the roofline).
// More columns than threads: the CTA walks the chunk once per group of kColThreads columns.
// =================================================================================================================
constexpr int kColThreads = 256;

struct CtaCopy {
  int c, n;  // my thread, threads in the CTA
  C2B_DEV void run(double *dst, const double *src, int cnt, bool wide) const {
    if (wide) {
      for (int e = 2 * c; e < cnt; e += 2 * n) cp_async16(dst + e, src + e);
    } else {
      for (int e = c; e < cnt; e += n) cp_async8(dst + e, src + e);
    }
  }
  C2B_DEV void run_t(double *dst, const double *t, int64_t first, int cnt, int64_t lo, int64_t hi) const {
    for (int e = c; e < cnt; e += n) {
      int64_t i = first + e;
      i         = i < lo ? lo : (i > hi ? hi : i);
      cp_async8(dst + e, t + i);
    }
  }
};

// dynamic shared memory: 2 x [A R*J | B R*J | t R+2 (+pad) | Y R*YW] then P R*J, c J
template <int J>
__host__ __device__ inline size_t csweep_smem_doubles(int yw) {
  constexpr int R = kTileRows;
  const int tb    = 2 * R * J + ((R + 2 + 1) & ~1) + R * yw;
  return (size_t)2 * tb + R * J + J + R;
}

// SOLVE = false: the products (matmul_lower / upper: z_n += a_n . F_n, F fed by the INPUT y_n, forward.hpp:228-271).  Their
// chunk summary is diagonal-affine (m = the chunk's product of propagators, b = the state reached from zero): K columns,
// no unit vectors; the thread of column 0 also carries m.
template <int J, bool LOWER, bool SUMMARY, bool SOLVE = true>
__global__ void __launch_bounds__(kColThreads, 2) csweep_tile_kernel(SweepArgs a) {
  pdl_prologue();
  constexpr int R  = kTileRows;
  constexpr int oB = R * J, oT = 2 * R * J, oY = oT + ((R + 2 + 1) & ~1);
  extern __shared__ __align__(16) double cs_smem[];
  const int NT         = blockDim.x;
  const int tid        = threadIdx.x;
  const int64_t chunk  = blockIdx.x % a.C, b = blockIdx.x / a.C;
  const int64_t n0     = chunk * a.L;
  const int64_t n1     = n0 + a.L < a.N ? n0 + a.L : a.N;
  constexpr bool BASIS = SUMMARY && SOLVE;  // J unit-vector columns in front of the real ones
  const int ncol       = BASIS ? J + a.K : a.K;
  const int yw         = NT < a.K ? NT : a.K;  // widest slice of Y a column group can need
  const int TB         = oY + R * yw;
  double *pb           = cs_smem + 2 * TB;
  double *csm          = pb + R * J;
  double *dinv         = csm + J;  // 1 / zdiv of the tile's rows (apply pass of apply_inverse)
  const double *t      = a.t + b * a.t_bs;
  const double *A = a.A + b * a.N * J, *B = a.B + b * a.N * J;
  const double *Y = a.Y + b * a.y_bs;
  double *Z = a.Z + b * a.N * (int64_t)a.K;
  double *F = (!SUMMARY && a.F) ? a.F + b * a.N * (int64_t)J * a.K : nullptr;
  const double *zdiv = (!SUMMARY && a.zdiv) ? a.zdiv + b * a.N : nullptr;
  const int64_t leaf = LOWER ? chunk : a.C - 1 - chunk;
  const CtaCopy cp{tid, NT};
  const bool wide = (J % 2 == 0) && (((reinterpret_cast<uintptr_t>(A) | reinterpret_cast<uintptr_t>(B)) & 15) == 0);
  if (tid < J) csm[tid] = (a.c + b * a.c_bs)[tid];
  const int cnt   = (int)(n1 - n0);
  const int ntile = (cnt + R - 1) / R;
  constexpr int dir = LOWER ? 1 : -1;
  const int64_t nf  = LOWER ? n0 : n1 - 1;
  for (int q0 = 0; q0 < ncol; q0 += NT) {
    const int q      = q0 + tid;  // my column
    const bool have  = q < ncol;
    const bool basis = BASIS && q < J;
    const int k      = BASIS ? q - J : q;  // real column index
    // real columns of this group: [kc0, kc0 + kw)
    const int kc0 = BASIS ? (q0 > J ? q0 - J : 0) : q0;
    const int kc1 = BASIS ? (q0 + NT - J < a.K ? q0 + NT - J : a.K) : (q0 + NT < a.K ? q0 + NT : a.K);
    const int kw  = kc1 > kc0 ? kc1 - kc0 : 0;
    double Fk[J];
    double mk[(SUMMARY && !SOLVE) ? J : 1];  // products: the chunk's propagator product (thread of column 0)
#pragma unroll
    for (int j = 0; j < ((SUMMARY && !SOLVE) ? J : 1); ++j) mk[j] = 1.0;
#pragma unroll
    for (int j = 0; j < J; ++j) Fk[j] = (basis && j == q) ? 1.0 : 0.0;
    if (!SUMMARY && a.carry && have) {
      const double *st = a.carry + (b * a.node_stride + leaf) * (int64_t)(J * a.K) + (int64_t)J * k;
#pragma unroll
      for (int j = 0; j < J; ++j) Fk[j] = st[j];
    }
    auto issue = [&](int kt) {
      if (kt < ntile) {
        const int rows   = cnt - kt * R < R ? cnt - kt * R : R;
        const int64_t lo = LOWER ? n0 + (int64_t)kt * R : n1 - (int64_t)kt * R - rows;
        const int pad    = LOWER ? 0 : R - rows;
        double *tb       = cs_smem + (kt & 1) * TB;
        cp.run(tb + pad * J, A + lo * J, rows * J, wide);
        cp.run(tb + oB + pad * J, B + lo * J, rows * J, wide);
        cp.run_t(tb + oT + pad, t, lo - 1, rows + 2, a.tmin, a.tmax);
        for (int r = 0; r < rows; ++r) cp.run(tb + oY + (pad + r) * yw, Y + (lo + r) * (int64_t)a.K + kc0, kw, false);
      }
      cp_async_commit();
    };
    __syncthreads();  // the previous column group is done with the buffers
    issue(0);
    issue(1);
    for (int kt = 0; kt < ntile; ++kt) {
      cp_async_wait<1>();
      __syncthreads();
      const double *tb = cs_smem + (kt & 1) * TB;
      const int rows   = cnt - kt * R < R ? cnt - kt * R : R;
      for (int i = tid; i < R * J; i += NT) {
        const int r = i / J, j = i - r * J;
        const int s = LOWER ? r : R - 1 - r;
        const double dt = LOWER ? tb[oT + s] - tb[oT + s + 1] : tb[oT + s + 1] - tb[oT + s + 2];
        pb[i] = exp(csm[j] * dt);  // dt = 0 (propagator 1) at the end of the series
      }
      const int64_t nt = nf + (int64_t)dir * kt * R;  // first row of the tile in visiting order
      if (zdiv && tid < rows) dinv[tid] = 1.0 / zdiv[nt + dir * tid];
      __syncthreads();
      if (have) {
#pragma unroll 2
        for (int r = 0; r < R; ++r) {
          if (r < rows) {
            const int s     = LOWER ? r : R - 1 - r;
            const int64_t n = nt + dir * r;
            if (F) {
#pragma unroll
              for (int j = 0; j < J; ++j) F[n * (int64_t)J * a.K + (int64_t)J * k + j] = Fk[j];  // (j, k) at j + J*k
            }
            double p[J], av[J], bv[J];
            lds_row<J>(pb + r * J, p);
            lds_row<J>(tb + s * J, av);
            double sm = 0.0;
#pragma unroll
            for (int j = 0; j < J; ++j) {
              Fk[j] = p[j] * Fk[j];
              sm    = fma(av[j], Fk[j], sm);
            }
            const double y = basis ? 0.0 : tb[oY + s * yw + (k - kc0)];
            double rr;
            if (SOLVE) {
              rr = y - sm;
              if (!SUMMARY) Z[n * (int64_t)a.K + k] = zdiv ? rr * dinv[r] : rr;
            } else {
              rr = y;
              if (!SUMMARY) Z[n * (int64_t)a.K + k] += sm;
              if (SUMMARY && q == 0) {
#pragma unroll
                for (int j = 0; j < J; ++j) mk[j] *= p[j];
              }
            }
            lds_row<J>(tb + oB + s * J, bv);
#pragma unroll
            for (int j = 0; j < J; ++j) Fk[j] = fma(bv[j], rr, Fk[j]);
          }
        }
      }
      __syncthreads();
      issue(kt + 2);
    }
    if (SUMMARY && !SOLVE && have) {
      double *e = a.leaves + (b * a.node_stride + leaf) * (int64_t)(J + J * a.K);
      if (q == 0) {
#pragma unroll
        for (int j = 0; j < J; ++j) e[j] = mk[j];
      }
#pragma unroll
      for (int j = 0; j < J; ++j) e[J + j + J * k] = Fk[j];
    }
    if (SUMMARY && SOLVE && have) {
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
}

// =================================================================================================================
// Multi-right-hand-side solves (K > 1): one warp per (series, chunk, tile of 32 columns); lane k of the warp
// carries the J-vector state of ITS column, so the columns never talk to each other.  The row's A / B entries and
// exp(c dt) are fetched / evaluated once per row by lanes 0..J-1 and broadcast with shuffles.  Y and Z accesses
// are coalesced across the lanes.  Summary: the same kernel over J + K columns (J unit vectors, then the real ones).
// =================================================================================================================
template <int J, bool LOWER, bool SUMMARY>
__global__ void __launch_bounds__(128) csweep_kernel(SweepArgs a) {
  pdl_prologue();
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
    if (!SUMMARY && have) Z[n * (int64_t)a.K + k] = a.zdiv ? z / a.zdiv[b * a.N + n] : z;
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
