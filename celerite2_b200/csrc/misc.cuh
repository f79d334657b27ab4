// a/U/V assembly and its adjoint, the log-det / chi^2 reductions, and the rectangular (predict) products.
#pragma once
#include "scan.cuh"

namespace c2b {

// ---- assembly --------------------------------------------------------------------------------------------------
// Reference: Term::get_celerite_matrices (c++/include/celerite2/terms.hpp:112-139, blocked columns) and the
// driver's get_celerite_matrices (python/celerite2/driver.cpp:422-477, interleaved columns).
// One thread per (series, row).
struct AsmArgs {
  int64_t N, batch;
  int Jr, Jc, layout;
  const double *ar, *cr, *ac, *bc, *cc, *dc, *x, *diag;
  double *c_out, *a_out, *U_out, *V_out;
  int64_t x_bs, co_r_bs, co_c_bs;  // batch strides of x, of the real and of the complex coefficient vectors
};

C2B_DEV int col_cos(int layout, int Jr, int Jc, int j) { return layout ? Jr + 2 * j : Jr + j; }
C2B_DEV int col_sin(int layout, int Jr, int Jc, int j) { return layout ? Jr + 2 * j + 1 : Jr + Jc + j; }

// One thread per (series, row) computes the row; the CTA's 128 rows of U and V (contiguous in global memory, rows
// being 8 J bytes) go through shared memory so that the stores are coalesced: thread-per-row stores have a stride
// of 8 J bytes and write every 32-byte sector in pieces (r01e: 0.056 ms for 120 MB).
constexpr int kAsmThreads = 128;

__global__ void __launch_bounds__(kAsmThreads) assemble_kernel(AsmArgs g) {
  pdl_prologue();
  extern __shared__ double asm_tile[];  // U rows [kAsmThreads * J], then V rows
  const int64_t gid0 = (int64_t)blockIdx.x * kAsmThreads;
  const int64_t gid  = gid0 + threadIdx.x;
  const int J        = g.Jr + 2 * g.Jc;
  const int64_t tot  = g.batch * g.N;
  double *sU = asm_tile + (size_t)threadIdx.x * J, *sV = asm_tile + (size_t)kAsmThreads * J + (size_t)threadIdx.x * J;
  if (gid < tot) {
    const int64_t b = gid / g.N, n = gid % g.N;
    const double *ar = g.ar + b * g.co_r_bs, *cr = g.cr + b * g.co_r_bs;
    const double *ac = g.ac + b * g.co_c_bs, *bc = g.bc + b * g.co_c_bs, *cc = g.cc + b * g.co_c_bs, *dc = g.dc + b * g.co_c_bs;
    double sum = 0.0;
    for (int j = 0; j < g.Jr; ++j) sum += ar[j];
    for (int j = 0; j < g.Jc; ++j) sum += ac[j];
    const double xn = g.x[b * g.x_bs + n];
    g.a_out[gid]    = g.diag[gid] + sum;
    for (int j = 0; j < g.Jr; ++j) {
      sV[j] = 1.0;
      sU[j] = ar[j];
    }
    for (int j = 0; j < g.Jc; ++j) {
      const int ic = col_cos(g.layout, g.Jr, g.Jc, j), is = col_sin(g.layout, g.Jr, g.Jc, j);
      double sn, cs;
      sincos(dc[j] * xn, &sn, &cs);
      sV[ic] = cs;
      sV[is] = sn;
      sU[ic] = ac[j] * cs + bc[j] * sn;
      sU[is] = ac[j] * sn - bc[j] * cs;
    }
    if (n == 0 && g.c_out) {
      double *c = g.c_out + b * J;
      for (int j = 0; j < g.Jr; ++j) c[j] = cr[j];
      for (int j = 0; j < g.Jc; ++j) {
        c[col_cos(g.layout, g.Jr, g.Jc, j)] = cc[j];
        c[col_sin(g.layout, g.Jr, g.Jc, j)] = cc[j];
      }
    }
  }
  __syncthreads();
  const int64_t rows = tot - gid0 < kAsmThreads ? tot - gid0 : kAsmThreads;
  const int cnt      = (int)rows * J;
  double *U = g.U_out + gid0 * J, *V = g.V_out + gid0 * J;
  const double *tU = asm_tile, *tV = asm_tile + (size_t)kAsmThreads * J;
  for (int e = threadIdx.x; e < cnt; e += kAsmThreads) {
    U[e] = tU[e];
    V[e] = tV[e];
  }
}

// Adjoint of the assembly.  Partial sums per CTA go to scratch, a second kernel reduces them in a fixed order.
//   part layout per (series, cta): [b_ar (Jr) | b_ac (Jc) | b_bc (Jc) | b_dc (Jc)]
struct AsmRevArgs {
  int64_t N, batch;
  int Jr, Jc, layout;
  const double *ar, *ac, *bc, *dc, *x;
  const double *b_c, *b_a, *b_U, *b_V;
  double *b_ar, *b_cr, *b_ac, *b_bc, *b_cc, *b_dc, *b_x, *b_diag;
  int64_t x_bs, co_r_bs, co_c_bs;
  double *part;
  int ctas_per_series;
};

constexpr int kAsmRevThreads = 256;

__global__ void assemble_rev_kernel(AsmRevArgs g) {
  pdl_prologue();
  extern __shared__ double sh[];  // [nred][kAsmRevThreads]
  const int J     = g.Jr + 2 * g.Jc;
  const int nred  = g.Jr + 3 * g.Jc;
  const int64_t b = blockIdx.y;
  const double *ac = g.ac + b * g.co_c_bs, *bc = g.bc + b * g.co_c_bs, *dc = g.dc + b * g.co_c_bs;
  for (int r = 0; r < nred; ++r) sh[r * kAsmRevThreads + threadIdx.x] = 0.0;
  for (int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; n < g.N; n += (int64_t)gridDim.x * blockDim.x) {
    const int64_t row = b * g.N + n;
    const double ba   = g.b_a[row];
    const double *bU = g.b_U + row * J, *bV = g.b_V + row * J;
    const double xn = g.x[b * g.x_bs + n];
    if (g.b_diag) g.b_diag[row] = ba;
    for (int j = 0; j < g.Jr; ++j) sh[j * kAsmRevThreads + threadIdx.x] += ba + bU[j];
    double bx = 0.0;
    for (int j = 0; j < g.Jc; ++j) {
      const int ic = col_cos(g.layout, g.Jr, g.Jc, j), is = col_sin(g.layout, g.Jr, g.Jc, j);
      double sn, cs;
      sincos(dc[j] * xn, &sn, &cs);
      const double buc = bU[ic], bus = bU[is];
      sh[(g.Jr + j) * kAsmRevThreads + threadIdx.x] += ba + buc * cs + bus * sn;
      sh[(g.Jr + g.Jc + j) * kAsmRevThreads + threadIdx.x] += buc * sn - bus * cs;
      const double barg = buc * (bc[j] * cs - ac[j] * sn) + bus * (ac[j] * cs + bc[j] * sn) - bV[ic] * sn + bV[is] * cs;
      sh[(g.Jr + 2 * g.Jc + j) * kAsmRevThreads + threadIdx.x] += barg * xn;
      bx = fma(barg, dc[j], bx);
    }
    if (g.b_x) g.b_x[row] += bx;
  }
  __syncthreads();
  for (int w = kAsmRevThreads / 2; w > 0; w >>= 1) {
    if ((int)threadIdx.x < w)
      for (int r = 0; r < nred; ++r) sh[r * kAsmRevThreads + threadIdx.x] += sh[r * kAsmRevThreads + threadIdx.x + w];
    __syncthreads();
  }
  if ((int)threadIdx.x < nred) g.part[(b * g.ctas_per_series + blockIdx.x) * nred + threadIdx.x] = sh[threadIdx.x * kAsmRevThreads];
}

__global__ void assemble_rev_final_kernel(AsmRevArgs g) {
  pdl_prologue();
  // one warp per reduced quantity: lanes stride over the per-CTA partials, fixed-shape shuffle tree (deterministic)
  const int J     = g.Jr + 2 * g.Jc;
  const int nred  = g.Jr + 3 * g.Jc;
  const int64_t b = blockIdx.x;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
  for (int r = wid; r < nred; r += nw) {
    double s = 0.0;
    for (int k = lane; k < g.ctas_per_series; k += 32) s += g.part[(b * g.ctas_per_series + k) * nred + r];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) {
      if (r < g.Jr)
        g.b_ar[b * g.Jr + r] = s;
      else if (r < g.Jr + g.Jc)
        g.b_ac[b * g.Jc + (r - g.Jr)] = s;
      else if (r < g.Jr + 2 * g.Jc)
        g.b_bc[b * g.Jc + (r - g.Jr - g.Jc)] = s;
      else
        g.b_dc[b * g.Jc + (r - g.Jr - 2 * g.Jc)] = s;
    }
  }
  const double *b_c = g.b_c ? g.b_c + b * J : nullptr;
  for (int j = threadIdx.x; j < g.Jr; j += blockDim.x) g.b_cr[b * g.Jr + j] = b_c ? b_c[j] : 0.0;
  for (int j = threadIdx.x; j < g.Jc; j += blockDim.x)
    g.b_cc[b * g.Jc + j] = b_c ? b_c[col_cos(g.layout, g.Jr, g.Jc, j)] + b_c[col_sin(g.layout, g.Jr, g.Jc, j)] : 0.0;
}

// ---- log-det / chi^2 (python/celerite2/numpy.py:84-87,104-109) --------------------------------------------------
constexpr int kRedThreads = 256;
__global__ void norm_logdet_kernel(int64_t N, const double *d, const double *z, double *bz, double *bd, double *part, int ctas) {
  pdl_prologue();
  __shared__ double s1[kRedThreads], s2[kRedThreads];
  const int64_t b = blockIdx.y;
  double ld = 0.0, ch = 0.0;
  for (int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; n < N; n += (int64_t)gridDim.x * blockDim.x) {
    const double dn = d[b * N + n], zn = z[b * N + n];
    const double inv = 1.0 / dn;
    ld += log(dn);
    ch = fma(zn * zn, inv, ch);
    if (bz) bz[b * N + n] = -zn * inv;
    if (bd) bd[b * N + n] = 0.5 * zn * zn * inv * inv - 0.5 * inv;
  }
  s1[threadIdx.x] = ld;
  s2[threadIdx.x] = ch;
  __syncthreads();
  for (int w = kRedThreads / 2; w > 0; w >>= 1) {
    if ((int)threadIdx.x < w) {
      s1[threadIdx.x] += s1[threadIdx.x + w];
      s2[threadIdx.x] += s2[threadIdx.x + w];
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    part[(b * ctas + blockIdx.x) * 2]     = s1[0];
    part[(b * ctas + blockIdx.x) * 2 + 1] = s2[0];
  }
}
// ll (nullable): loglike[b] = -0.5 (chi2 + logdet) - half_n_log_2pi, written by the same launch (python/celerite2/core.py:407-428)
__global__ void norm_logdet_final_kernel(const double *part, int ctas, double *out, double *ll, double half_n_log_2pi) {
  pdl_prologue();
  // warp 0: log-det, warp 1: chi^2
  __shared__ double both[2];
  const int64_t b = blockIdx.x;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (wid < 2) {
    double s = 0.0;
    for (int k = lane; k < ctas; k += 32) s += part[(b * ctas + k) * 2 + wid];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) {
      out[b * 2 + wid] = s;
      both[wid]        = s;
    }
  }
  __syncthreads();
  if (ll && threadIdx.x == 0) ll[b] = -0.5 * (both[0] + both[1]) - half_n_log_2pi;
}

__global__ void add_inplace_kernel(int64_t n, const double *x, double *y) {
  pdl_prologue();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) y[i] += x[i];
}

// ---- general_matmul (forward.hpp:285-392) ------------------------------------------------------------------------
// Step 1: running state over t2, H_m = P_m H_{m-1} + V[m]^T Y[m] (lower, ascending m; upper is the mirror).  This
// is a diagonal-transition scan; chunked exactly like matmul_lower.  States are stored ROW-major (j*K + k) as
// the reference's workspace is (forward.hpp:299).  Step 2: every output row finds its source row m* by binary
// search (lower: last t2 <= t1[n]; upper: first t2 > t1[n]) and gathers
//   Z[n] += (U[n] o exp(-c |t1[n] - t2[m*]|)) H_{m*}.
struct GmArgs {
  int64_t N, M, C, L;
  int J, K;
  const double *t1, *t2, *c, *U, *V, *Y;
  double *Z, *H;      // H: (M, J*K) all rows (scratch)
  double *Fout;       // user workspace or null
  double *leaves;
  const double *carry;
  int64_t node_stride;
};

template <bool UPPER, bool WRITE>
C2B_DEV void gm_rows(const GmArgs &g, int64_t m0, int64_t m1, int k, double *Hs /* J values */) {
  const int J = g.J;
  for (int64_t i = 0; i < m1 - m0; ++i) {
    const int64_t m = UPPER ? m1 - 1 - i : m0 + i;
    const int64_t o = UPPER ? m + 1 : m - 1;
    const bool has  = (o >= 0 && o < g.M);
    const double dt = has ? (UPPER ? g.t2[m] - g.t2[o] : g.t2[o] - g.t2[m]) : 0.0;
    const double y  = g.Y[m * g.K + k];
    for (int j = 0; j < J; ++j) {
      const double pj = has ? exp(g.c[j] * dt) : 1.0;
      Hs[j]           = fma(g.V[m * J + j], y, pj * Hs[j]);
      if (WRITE) g.H[m * (int64_t)J * g.K + (int64_t)j * g.K + k] = Hs[j];
    }
  }
}

// thread per (chunk, column)
template <bool UPPER, bool SUMMARY>
__global__ void gm_scan_kernel(GmArgs g) {
  pdl_prologue();
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= g.C * g.K) return;
  const int k         = (int)(gid % g.K);
  const int64_t chunk = gid / g.K;
  const int64_t m0 = chunk * g.L, m1 = (m0 + g.L < g.M) ? m0 + g.L : g.M;
  const int64_t leaf = UPPER ? g.C - 1 - chunk : chunk;
  const int J        = g.J;
  double Hs[C2B_MAX_J];
  if (SUMMARY) {
    for (int j = 0; j < J; ++j) Hs[j] = 0.0;
    gm_rows<UPPER, false>(g, m0, m1, k, Hs);
    double *e = g.leaves + leaf * (int64_t)(J + J * g.K);
    for (int j = 0; j < J; ++j) e[J + j + J * k] = Hs[j];
    if (k == 0) {
      double dt;
      if (UPPER)
        dt = g.t2[m0] - g.t2[m1 < g.M ? m1 : g.M - 1];
      else
        dt = g.t2[m0 > 0 ? m0 - 1 : 0] - g.t2[m1 - 1];
      for (int j = 0; j < J; ++j) e[j] = exp(g.c[j] * dt);
    }
  } else {
    if (g.carry) {
      const double *st = g.carry + leaf * (int64_t)(J * g.K) + (int64_t)J * k;
      for (int j = 0; j < J; ++j) Hs[j] = st[j];
    } else {
      for (int j = 0; j < J; ++j) Hs[j] = 0.0;
    }
    gm_rows<UPPER, true>(g, m0, m1, k, Hs);
  }
}

// thread per (output row, column)
template <bool UPPER>
__global__ void gm_gather_kernel(GmArgs g) {
  pdl_prologue();
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= g.N * g.K) return;
  const int k     = (int)(gid % g.K);
  const int64_t n = gid / g.K;
  const double tn = g.t1[n];
  // lo = number of t2 <= tn
  int64_t lo = 0, hi = g.M;
  while (lo < hi) {
    const int64_t mid = (lo + hi) >> 1;
    if (g.t2[mid] <= tn)
      lo = mid + 1;
    else
      hi = mid;
  }
  const int64_t ms = UPPER ? lo : lo - 1;
  if (ms < 0 || ms >= g.M) return;  // forward.hpp:316-317 / 378-379: rows outside the range are untouched
  const double dt = UPPER ? tn - g.t2[ms] : g.t2[ms] - tn;
  double s        = 0.0;
  for (int j = 0; j < g.J; ++j) s = fma(g.U[n * g.J + j] * exp(g.c[j] * dt), g.H[ms * (int64_t)g.J * g.K + (int64_t)j * g.K + k], s);
  g.Z[gid] += s;
}

// Copy into the user's workspace exactly the rows the reference writes (forward.hpp:307,313,327 / 369,387).
template <bool UPPER>
__global__ void gm_workspace_kernel(GmArgs g) {
  pdl_prologue();
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t JK  = (int64_t)g.J * g.K;
  if (gid >= g.M * JK) return;
  const int64_t m = gid / JK;
  bool visited;
  if (!UPPER) {
    // rows 1.. are visited while t2[m] <= max t1 (and some t1 >= t2[0]); row 0 is always written
    visited = (m == 0) || (g.t1[g.N - 1] >= g.t2[0] && g.t2[m] <= g.t1[g.N - 1]);
  } else {
    // rows M-2.. down are visited while t2[m] > min t1 (and some t1 < t2[M-1]); row M-1 is never written
    visited = (m <= g.M - 2) && (g.t1[0] < g.t2[g.M - 1]) && (g.t2[m] > g.t1[0]);
  }
  if (visited)
    g.Fout[gid] = g.H[gid];
  else if (m == 0)
    g.Fout[gid] = 0.0;
}

}  // namespace c2b
