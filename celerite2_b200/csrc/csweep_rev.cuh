// Reverse mode of the multi-right-hand-side sweeps (K > 1), apply pass, on staged tiles: one CTA per (series, chunk), thread k
// carrying the J-vector adjoint state bF(:, k) of ITS column (reverse.hpp:87-217 -> internal.hpp:191-303, column by column).
// Unlike the forward pass the columns meet again: bA_n, bB_n (the cotangents of the two row vectors) and the propagator
// cotangent bp_n are SUMS over the K columns.  Each thread writes its 3 J contributions of a row into a shared-memory
// panel; after RB rows the CTA's threads add the panel up (one thread per (row, vector entry), a column-strided read made
// conflict-free by an odd row pitch) and store bA / bB, while bp goes on to bc (per-chunk partial sums) and bt.
// The generic kernel this replaces walks the K columns one after the other in ONE thread per chunk with read-modify-write
// accumulation in global memory: 417 ms for solve_lower_rev at J = 16, N = 2e5, K = 64 (r03b).
#pragma once
#include "gsweep.cuh"

namespace c2b {

// rows per reduction panel: the largest of 4, 2, 1 whose panel fits the budget
__host__ __device__ inline int csweep_rev_rb(int J, int nt) {
  const size_t budget = 96 * 1024;
  for (int rb = 4; rb > 1; rb >>= 1)
    if ((size_t)rb * 3 * J * (nt + 1) * sizeof(double) <= budget) return rb;
  return 1;
}
// dynamic shared memory (doubles): 2 x [A R*J | B R*J | t R+2 (+pad) | bZ R*yw | ZY R*yw], P R*J, c J, dt R, bp RB*J, panel
template <int J>
__host__ __device__ inline size_t csweep_rev_smem_doubles(int yw, int nt) {
  constexpr int R = kTileRows;
  const int tb    = 2 * R * J + ((R + 2 + 1) & ~1) + 2 * R * yw;
  const int rb    = csweep_rev_rb(J, nt);
  return (size_t)2 * tb + R * J + J + R + (size_t)rb * J + (size_t)rb * 3 * J * (nt + 1);
}

template <int J, bool SOLVE, bool LOWER>
__global__ void __launch_bounds__(kColThreads, 1) csweep_rev_tile_kernel(SweepRevArgs a) {
  pdl_prologue();
  constexpr int R   = kTileRows;
  constexpr bool UP = !LOWER;  // the reverse of a lower sweep walks downwards
  constexpr int dir = UP ? 1 : -1;
  constexpr int oB = R * J, oT = 2 * R * J, oZ = oT + ((R + 2 + 1) & ~1);
  extern __shared__ __align__(16) double cr_smem[];
  const int NT        = blockDim.x;
  const int tid       = threadIdx.x;
  const int64_t chunk = blockIdx.x % a.C, b = blockIdx.x / a.C;
  const int64_t n0    = chunk * a.L;
  const int64_t n1    = n0 + a.L < a.N ? n0 + a.L : a.N;
  const int K         = a.K;
  const int yw        = NT < K ? NT : K;
  const int oY        = oZ + R * yw;
  const int TB        = oY + R * yw;
  const int RB        = csweep_rev_rb(J, NT);
  const int pitch     = NT + 1;
  double *pb    = cr_smem + 2 * TB;
  double *csm   = pb + R * J;
  double *dts   = csm + J;
  double *bps   = dts + R;
  double *panel = bps + RB * J;
  const double *t = a.t + b * a.t_bs;
  const int64_t NJ = a.N * J, NK = a.N * (int64_t)K;
  const double *A = a.A + b * NJ, *B = a.B + b * NJ, *F = a.F + b * NK * J, *bZ = a.bZ + b * NK;
  const double *ZY = SOLVE ? a.Z + b * NK : a.Y + b * a.y_bs;
  double *bt = a.bt + b * a.N, *bA = a.bA + b * NJ, *bB = a.bB + b * NJ, *bY = a.bY + b * NK;
  const int64_t leaf = LOWER ? a.C - 1 - chunk : chunk;
  const CtaCopy cp{tid, NT};
  const bool wide = (J % 2 == 0) && (((reinterpret_cast<uintptr_t>(A) | reinterpret_cast<uintptr_t>(B)) & 15) == 0);
  if (tid < J) csm[tid] = (a.c + b * a.c_bs)[tid];
  const int cnt       = (int)(n1 - n0);
  const int ntile     = (cnt + R - 1) / R;
  const int64_t nf    = UP ? n0 : n1 - 1;              // first row visited
  const int64_t nlast = LOWER ? a.tmin : a.tmax;       // the row without a predecessor in this direction
  const double sgn    = LOWER ? 1.0 : -1.0;
  double bc_acc = 0.0;  // thread j < J: this chunk's share of bc_j (all column groups)
  double edge   = 0.0;  // thread 0: the bt term owed to the row just outside the chunk
  for (int q0 = 0; q0 < K; q0 += NT) {
    const int k     = q0 + tid;
    const bool have = k < K;
    const int kw    = q0 + NT < K ? NT : K - q0;
    const bool first_group = q0 == 0;
    double bF[J];
#pragma unroll
    for (int j = 0; j < J; ++j) bF[j] = 0.0;
    if (a.carry && have) {
      const double *st = a.carry + (b * a.node_stride + leaf) * (int64_t)(J * K) + (int64_t)J * k;
#pragma unroll
      for (int j = 0; j < J; ++j) bF[j] = st[j];
    }
    double pend = 0.0;  // thread 0
    auto issue = [&](int kt) {
      if (kt < ntile) {
        const int rows   = cnt - kt * R < R ? cnt - kt * R : R;
        const int64_t lo = UP ? n0 + (int64_t)kt * R : n1 - (int64_t)kt * R - rows;
        const int pad    = UP ? 0 : R - rows;
        double *tb       = cr_smem + (kt & 1) * TB;
        cp.run(tb + pad * J, A + lo * J, rows * J, wide);
        cp.run(tb + oB + pad * J, B + lo * J, rows * J, wide);
        cp.run_t(tb + oT + pad, t, lo - 1, rows + 2, a.tmin, a.tmax);
        for (int r = 0; r < rows; ++r) {
          cp.run(tb + oZ + (pad + r) * yw, bZ + (lo + r) * (int64_t)K + q0, kw, false);
          cp.run(tb + oY + (pad + r) * yw, ZY + (lo + r) * (int64_t)K + q0, kw, false);
        }
      }
      cp_async_commit();
    };
    __syncthreads();  // the previous column group is done with the buffers
    issue(0);
    issue(1);
    for (int kt = 0; kt < ntile; ++kt) {
      cp_async_wait<1>();
      __syncthreads();
      const double *tb = cr_smem + (kt & 1) * TB;
      const int rows   = cnt - kt * R < R ? cnt - kt * R : R;
      for (int i = tid; i < R * J; i += NT) {
        const int r = i / J, j = i - r * J;
        const int s = UP ? r : R - 1 - r;
        const double dt = LOWER ? tb[oT + s] - tb[oT + s + 1] : tb[oT + s + 1] - tb[oT + s + 2];
        pb[i] = exp(csm[j] * dt);
        if (j == 0) dts[r] = dt;
      }
      __syncthreads();
      const int64_t nt = nf + (int64_t)dir * kt * R;  // first row of the tile in visiting order
      for (int r0 = 0; r0 < rows; r0 += RB) {
        const int nr = rows - r0 < RB ? rows - r0 : RB;
        // ---- every column walks nr rows and leaves its contributions in the panel ------------------------------------
        for (int rr = 0; rr < nr; ++rr) {
          const int r     = r0 + rr;
          const int s     = UP ? r : R - 1 - r;
          const int64_t n = nt + dir * r;
          const bool nb   = n != nlast;
          double *pan     = panel + (size_t)rr * 3 * J * pitch + tid;
          if (have) {
            double p[J], av[J], bv[J];
            lds_row<J>(pb + r * J, p);
            lds_row<J>(tb + s * J, av);
            lds_row<J>(tb + oB + s * J, bv);
            const double bz0 = tb[oZ + s * yw + tid], zy = tb[oY + s * yw + tid];
            double dot = 0.0;
#pragma unroll
            for (int j = 0; j < J; ++j) dot = fma(bv[j], bF[j], dot);
            const double bz = SOLVE ? bz0 + dot : bz0;
            bY[n * (int64_t)K + k] = SOLVE ? bz : dot;
            const double *Fn = F + n * (int64_t)J * K + (int64_t)J * k;
#pragma unroll
            for (int j = 0; j < J; ++j) {
              pan[(size_t)j * pitch] = zy * bF[j];  // -> bB_n[j]
              double cA = 0.0, bp = 0.0;
              if (nb) {
                const double au = av[j] * bz;
                const double nw = SOLVE ? bF[j] - au : bF[j] + au;
                const double fn = Fn[j];
                const double pf = p[j] * fn;
                cA              = SOLVE ? -(bz * pf) : bz * pf;
                bp              = fn * nw * p[j];
                bF[j]           = p[j] * nw;
              }
              pan[(size_t)(J + j) * pitch]     = cA;  // -> bA_n[j]
              pan[(size_t)(2 * J + j) * pitch] = bp;  // -> bp_n[j]
            }
          } else {
#pragma unroll
            for (int j = 0; j < 3 * J; ++j) pan[(size_t)j * pitch] = 0.0;
          }
        }
        __syncthreads();
        // ---- panel sums: one thread per (row, entry) ---------------------------------------------------------------
        for (int w = tid; w < nr * 3 * J; w += NT) {
          const int rr = w / (3 * J), v = w - rr * 3 * J;
          const double *src = panel + (size_t)w * pitch;
          double s0 = 0.0, s1 = 0.0;
          int x = 0;
          for (; x + 1 < NT; x += 2) {
            s0 += src[x];
            s1 += src[x + 1];
          }
          if (x < NT) s0 += src[x];
          const double sum = s0 + s1;
          const int64_t n  = nt + dir * (r0 + rr);
          if (v < J) {
            double *dst = bB + n * J + v;
            *dst        = first_group ? sum : *dst + sum;
          } else if (v < 2 * J) {
            double *dst = bA + n * J + (v - J);
            *dst        = first_group ? sum : *dst + sum;
          } else {
            bps[rr * J + (v - 2 * J)] = sum;
          }
        }
        __syncthreads();
        // ---- bp -> bc (thread j) and bt (thread 0), rows in visiting order -----------------------------------------------
        if (tid < J) {
          for (int rr = 0; rr < nr; ++rr) bc_acc = fma(dts[r0 + rr], bps[rr * J + tid], bc_acc);
        }
        if (tid == 0) {
          for (int rr = 0; rr < nr; ++rr) {
            double fac = 0.0;
#pragma unroll
            for (int j = 0; j < J; ++j) fac = fma(csm[j], bps[rr * J + j], fac);  // 0 on the row without a predecessor
            const int64_t n = nt + dir * (r0 + rr);
            const double v  = pend - sgn * fac;
            bt[n]           = first_group ? v : bt[n] + v;
            pend            = sgn * fac;
          }
        }
        // (the next panel is written only after the next barrier-separated phase: the writers above run before the
        // barrier that follows the walk, and the readers of bps finish before the walk's writes to the panel matter)
        __syncthreads();
      }
      issue(kt + 2);
    }
    if (tid == 0) edge += pend;
  }
  if (tid < J) a.bc_part[(b * a.C + chunk) * J + tid] = bc_acc;
  if (tid == 0) a.bt_edge[b * a.C + chunk] = edge;
}

// ---- summary pass: the chunk's map of the adjoint state, column by column ----------------------------------------------
// solves: bF -> P (I -/+ a b^T) bF -/+ P a bz0 per row: J unit-vector columns give the J x J head, the K real columns (from
// zero, fed by bZ) the offsets: the AffV element of the forward pass with the roles of the row vectors exchanged.
// products: bF -> p o (bF + a bz0): diagonal-affine; the head is the chunk's product of propagators, exp(c * (time spanned)).
// dynamic shared memory (doubles): 2 x [A R*J | B R*J | t R+2 (+pad) | bZ R*yw], P R*J, c J
template <int J>
__host__ __device__ inline size_t csweep_rev_summary_smem_doubles(int yw) {
  constexpr int R = kTileRows;
  const int tb    = 2 * R * J + ((R + 2 + 1) & ~1) + R * yw;
  return (size_t)2 * tb + R * J + J;
}

template <int J, bool SOLVE, bool LOWER>
__global__ void __launch_bounds__(kColThreads, 2) csweep_rev_summary_kernel(SweepRevArgs a) {
  pdl_prologue();
  constexpr int R   = kTileRows;
  constexpr bool UP = !LOWER;
  constexpr int dir = UP ? 1 : -1;
  constexpr int oB = R * J, oT = 2 * R * J, oZ = oT + ((R + 2 + 1) & ~1);
  extern __shared__ __align__(16) double cr_smem[];
  const int NT        = blockDim.x;
  const int tid       = threadIdx.x;
  const int64_t chunk = blockIdx.x % a.C, b = blockIdx.x / a.C;
  const int64_t n0    = chunk * a.L;
  const int64_t n1    = n0 + a.L < a.N ? n0 + a.L : a.N;
  const int K         = a.K;
  const int ncol      = SOLVE ? J + K : K;
  const int yw        = NT < K ? NT : K;
  const int TB        = oZ + R * yw;
  double *pb  = cr_smem + 2 * TB;
  double *csm = pb + R * J;
  const double *t = a.t + b * a.t_bs;
  const int64_t NJ = a.N * J, NK = a.N * (int64_t)K;
  const double *A = a.A + b * NJ, *B = a.B + b * NJ, *bZ = a.bZ + b * NK;
  const int64_t leaf = LOWER ? a.C - 1 - chunk : chunk;
  const int es       = SOLVE ? J * J + J * K : J + J * K;
  double *e          = a.leaves + (b * a.node_stride + leaf) * (int64_t)es;
  const CtaCopy cp{tid, NT};
  const bool wide = (J % 2 == 0) && (((reinterpret_cast<uintptr_t>(A) | reinterpret_cast<uintptr_t>(B)) & 15) == 0);
  if (tid < J) csm[tid] = (a.c + b * a.c_bs)[tid];
  const int cnt       = (int)(n1 - n0);
  const int ntile     = (cnt + R - 1) / R;
  const int64_t nf    = UP ? n0 : n1 - 1;
  const int64_t nlast = LOWER ? a.tmin : a.tmax;
  for (int q0 = 0; q0 < ncol; q0 += NT) {
    const int q      = q0 + tid;
    const bool have  = q < ncol;
    const bool basis = SOLVE && q < J;
    const int k      = SOLVE ? q - J : q;
    const int kc0 = SOLVE ? (q0 > J ? q0 - J : 0) : q0;
    const int kc1 = SOLVE ? (q0 + NT - J < K ? q0 + NT - J : K) : (q0 + NT < K ? q0 + NT : K);
    const int kw  = kc1 > kc0 ? kc1 - kc0 : 0;
    double G[J];
#pragma unroll
    for (int j = 0; j < J; ++j) G[j] = (basis && j == q) ? 1.0 : 0.0;
    auto issue = [&](int kt) {
      if (kt < ntile) {
        const int rows   = cnt - kt * R < R ? cnt - kt * R : R;
        const int64_t lo = UP ? n0 + (int64_t)kt * R : n1 - (int64_t)kt * R - rows;
        const int pad    = UP ? 0 : R - rows;
        double *tb       = cr_smem + (kt & 1) * TB;
        cp.run(tb + pad * J, A + lo * J, rows * J, wide);
        if (SOLVE) cp.run(tb + oB + pad * J, B + lo * J, rows * J, wide);
        cp.run_t(tb + oT + pad, t, lo - 1, rows + 2, a.tmin, a.tmax);
        for (int r = 0; r < rows; ++r) cp.run(tb + oZ + (pad + r) * yw, bZ + (lo + r) * (int64_t)K + kc0, kw, false);
      }
      cp_async_commit();
    };
    __syncthreads();
    issue(0);
    issue(1);
    for (int kt = 0; kt < ntile; ++kt) {
      cp_async_wait<1>();
      __syncthreads();
      const double *tb = cr_smem + (kt & 1) * TB;
      const int rows   = cnt - kt * R < R ? cnt - kt * R : R;
      for (int i = tid; i < R * J; i += NT) {
        const int r = i / J, j = i - r * J;
        const int s = UP ? r : R - 1 - r;
        const double dt = LOWER ? tb[oT + s] - tb[oT + s + 1] : tb[oT + s + 1] - tb[oT + s + 2];
        pb[i] = exp(csm[j] * dt);
      }
      __syncthreads();
      const int64_t nt = nf + (int64_t)dir * kt * R;
      if (have) {
#pragma unroll 2
        for (int r = 0; r < R; ++r) {
          if (r < rows && nt + dir * r != nlast) {
            const int s = UP ? r : R - 1 - r;
            double p[J], av[J];
            lds_row<J>(pb + r * J, p);
            lds_row<J>(tb + s * J, av);
            double bz = basis ? 0.0 : tb[oZ + s * yw + (k - kc0)];
            if (SOLVE) {
              double bv[J];
              lds_row<J>(tb + oB + s * J, bv);
#pragma unroll
              for (int j = 0; j < J; ++j) bz = fma(bv[j], G[j], bz);
#pragma unroll
              for (int j = 0; j < J; ++j) G[j] = p[j] * fma(-av[j], bz, G[j]);
            } else {
#pragma unroll
              for (int j = 0; j < J; ++j) G[j] = p[j] * fma(av[j], bz, G[j]);
            }
          }
        }
      }
      __syncthreads();
      issue(kt + 2);
    }
    if (have) {
      if (basis) {
#pragma unroll
        for (int j = 0; j < J; ++j) e[j * J + q] = G[j];
      } else {
        double *bb = e + (SOLVE ? J * J : J);
#pragma unroll
        for (int j = 0; j < J; ++j) bb[j + J * k] = G[j];
      }
    }
  }
  if (!SOLVE && tid < J) {
    // product of the propagators the chunk applies (rows with a predecessor only): exp(c * time spanned)
    double dt;
    if (LOWER)
      dt = t[n0 > 0 ? n0 - 1 : 0] - t[n1 - 1];
    else
      dt = t[n0] - t[n1 < a.N ? n1 : a.N - 1];
    e[tid] = exp(csm[tid] * dt);
  }
}

}  // namespace c2b
