// Lane-group kernels: G = 2, 4 or 8 lanes (G >= J, compile-time) share one chunk; lane c of the group owns column c
// (or element c) of the recursion's state, so a row's U/V/W entries are one coalesced access per group, 32/G chunks
// share a warp, and the per-row dependency chain is spread over the lanes (shuffles for the few all-to-all steps).
// This replaces the thread-per-chunk mapping on the hot path: the r01b profile showed that mapping bound by L1/LSU
// wavefronts (32 divergent streams per warp) and by single-thread instruction latency.
#pragma once
#include "common.cuh"

namespace c2b {

template <int J>
struct GroupOf {
  static constexpr int G = J <= 2 ? 2 : (J <= 4 ? 4 : (J <= 8 ? 8 : (J <= 16 ? 16 : 32)));
};

// All lanes of a warp must run the same number of loop trips (the shuffles and warp barriers need the full warp),
// but the groups of a warp may own chunks of different length: loop to the warp-wide maximum.
C2B_DEV int warp_max(int v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const int w = __shfl_xor_sync(0xffffffffu, v, o);
    v           = w > v ? w : v;
  }
  return v;
}

// sum over the lanes of my group (all 32 lanes must call)
template <int G>
C2B_DEV double gsum(double x) {
#pragma unroll
  for (int o = G / 2; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
  return x;
}
// value held by lane j of my group
template <int G>
C2B_DEV double gget(double x, int j) {
  return __shfl_sync(0xffffffffu, x, j, G);
}
// all-gather: out[j] = value of lane j of my group, j < J
template <int G, int J>
C2B_DEV void gall(double x, double *out) {
#pragma unroll
  for (int j = 0; j < J; ++j) out[j] = __shfl_sync(0xffffffffu, x, j, G);
}

// Per-group shared-memory row buffer used to turn "lane c owns column c of a J x J matrix" into contiguous
// global accesses: the row (J*J doubles, element (j,c) at j + J*c) is written/read by its owners in shared memory
// and moved to/from global memory with lanes striding over consecutive elements.
// Every lane of the warp must call (warp-level barriers inside); `active` = lane owns a column, `doit` = this group
// really stores.
template <int G, int J>
C2B_DEV void gstore_matrix_row(double *gbuf /* smem, J*J doubles of my group */, double *grow /* global row */, const double *col, int c,
                               bool active, bool doit) {
  if (active) {
#pragma unroll
    for (int j = 0; j < J; ++j) gbuf[j + J * c] = col[j];
  }
  __syncwarp();
  if (doit) {
#pragma unroll
    for (int e = c; e < J * J; e += G) grow[e] = gbuf[e];
  }
  __syncwarp();
}
// Mirror image: fetch a J*J row from global memory with contiguous accesses, hand every lane its column.
template <int G, int J>
C2B_DEV void gload_matrix_row(double *gbuf, const double *grow, double *col, int c, bool active, bool doit) {
  if (doit) {
#pragma unroll
    for (int e = c; e < J * J; e += G) gbuf[e] = grow[e];
  }
  __syncwarp();
  if (active) {
#pragma unroll
    for (int j = 0; j < J; ++j) col[j] = gbuf[j + J * c];
  }
  __syncwarp();
}

// ---- per-group ring of input rows ------------------------------------------------------------------------------
// The lane-group kernels walk a chunk row by row; a row's inputs (a few J-vectors, a few scalars, for factor_rev
// also the J x J workspace row) must be in flight several rows ahead or every row pays the DRAM latency (r01d
// profile: 45-48 % of the stall samples are long-scoreboard with one row of look-ahead).  The lanes of a group
// copy row i+DEPTH into stage (i mod DEPTH) of a shared-memory ring with cp.async (no registers held) while row i
// is consumed; every lane then reads whole J-vectors of the row from shared memory (broadcast within the group),
// which also replaces the shuffle all-gathers of the inputs.
// Use: prologue issue(0..DEPTH-1) each followed by commit(); per trip: acquire(); read; release(); issue(i+DEPTH);
// commit().  All 32 lanes run every call (warp barriers inside acquire/release).
template <int G, int ROWD, int DEPTH>
struct GRing {
  static_assert((DEPTH & (DEPTH - 1)) == 0, "DEPTH must be a power of two");
  static_assert(ROWD % 2 == 0, "stages must stay 16-byte aligned");
  // Doubles between the rings of neighbouring groups.  The groups of a warp access the same offsets of their rings
  // in one instruction; a stride that is a multiple of 128 bytes puts them all on the same banks (r01e profile:
  // twice the ideal shared-memory wavefronts, L1 data pipe 68 % busy), so the stride is padded to G doubles
  // modulo 16: the 32/G groups of a warp then tile the 32 banks.
  static constexpr int kStride = DEPTH * ROWD + ((G % 16) - (DEPTH * ROWD) % 16 + 16) % 16;
  double *base;  // shared memory of my group: kStride doubles, 16-byte aligned
  int c;         // my lane within the group
  C2B_DEV double *stage(int i) const { return base + (i & (DEPTH - 1)) * ROWD; }
  // cnt doubles at src -> offset off of stage i, the lanes of the group striding over them
  C2B_DEV void copy(int i, int off, const double *src, int cnt) const {
    double *dst = stage(i) + off;
    for (int e = c; e < cnt; e += G) cp_async8(dst + e, src + e);
  }
  // same with 16-byte copies (src, off and cnt even / 16-byte aligned)
  C2B_DEV void copy16(int i, int off, const double *src, int cnt) const {
    double *dst = stage(i) + off;
    for (int e = 2 * c; e < cnt; e += 2 * G) cp_async16(dst + e, src + e);
  }
  C2B_DEV void copy1(int i, int off, const double *src) const { cp_async8(stage(i) + off, src); }
  C2B_DEV void commit() const { cp_async_commit(); }
  C2B_DEV void acquire() const {
    cp_async_wait<DEPTH - 1>();
    __syncwarp();
  }
  // rows i and i+1 are readable (one row less of look-ahead): lets a kernel evaluate the state-independent part of
  // row i+1 (the exp) in the shadow of row i's dependent chain
  C2B_DEV void acquire2() const {
    cp_async_wait<DEPTH - 2>();
    __syncwarp();
  }
  C2B_DEV void release() const { __syncwarp(); }
};

// One lane's share of the scalar inputs of a row (t_n, y_n, ...): scalar e of the row is copied by lane G-1-e (the
// lanes that own no column first), scalars e >= G wrap around.  Everything that depends only on the lane is
// resolved before the loop, so a row costs one clamp and one copy per scalar.
template <int G, int NS>
struct ScalarStreams {
  static constexpr int NSL = (NS + G - 1) / G;
  const double *base[NSL];
  int64_t del[NSL];
  bool on[NSL];
  // src[e] = array of scalar e, delta[e] = row offset (0, or -1/+1 for the neighbouring time stamp, clamped to [0, N-1])
  C2B_DEV void init(int c, bool has, const double *const *src, const int *delta) {
#pragma unroll
    for (int k = 0; k < NSL; ++k) {
      const int e = G - 1 - c + k * G;
      on[k]       = has && e < NS;
      base[k]     = src[0];
      del[k]      = 0;
#pragma unroll
      for (int x = 0; x < NS; ++x)
        if (e == x) {
          base[k] = src[x];
          del[k]  = delta[x];
        }
    }
  }
  template <class Ring>
  C2B_DEV void issue(const Ring &ring, int i, int off, int c, int64_t n, int64_t N) const {
#pragma unroll
    for (int k = 0; k < NSL; ++k)
      if (on[k]) {
        int64_t idx = n + del[k];
        idx         = idx < 0 ? 0 : (idx > N - 1 ? N - 1 : idx);
        ring.copy1(i, off + (G - 1 - c + k * G), base[k] + idx);
      }
  }
};

// ---- per-group tiles of input rows -------------------------------------------------------------------------------
// Second generation of the input staging (r01e profile: with one row per trip the loop is one long serial
// instruction stream of ~150 instructions at ~6 cycles each; a third of it is the per-row copy bookkeeping and
// another third the exp, which does not depend on the state).  A group now stages TILES of R consecutive rows:
// the rows of a tile are contiguous in global memory, so a field of the tile is one run of R*J doubles that the
// lanes copy with 16-byte cp.async; the kernel's row loop is fully unrolled over the tile, which gives constant
// shared-memory offsets, one wait + barrier per tile instead of per row, and lets the compiler interleave the R
// independent exp evaluations of the tile.  Two tile buffers per group: tile k+1 lands while tile k is consumed.
//
// Buffer slots: slot s of a tile holds global row lo + s - pad, where pad = 0 for kernels walking upwards and
// R - rows for kernels walking downwards (a short last tile is aligned to the END of the buffer), so that step r of
// the unrolled loop always reads slot r (upwards) or R-1-r (downwards).
template <int G>
struct GCopy {
  int c;  // my lane within the group
  // cnt contiguous doubles; wide: cnt even, src and dst 16-byte aligned
  C2B_DEV void run(double *dst, const double *src, int cnt, bool wide) const {
    if (wide) {
      for (int e = 2 * c; e < cnt; e += 2 * G) cp_async16(dst + e, src + e);
    } else {
      for (int e = c; e < cnt; e += G) cp_async8(dst + e, src + e);
    }
  }
  // ---- full tiles (r02j: the copy loops with run-time counts and the clamped time stamps below were 10 - 25 % of the
  // instructions of every lane-group kernel; all but the first and last tile of a chunk take this path) -----------------
  // a full tile of R rows starting at row lo whose time stamps lo-1 .. lo+R exist
  template <int R>
  C2B_DEV static bool full_tile(int rows, int64_t lo, int64_t tmin, int64_t tmax) {
    return rows == R && lo - 1 >= tmin && lo + R <= tmax;
  }
  // CNT doubles (even), dst and src 16-byte aligned
  template <int CNT>
  C2B_DEV void full16(double *dst, const double *src) const {
    static_assert(CNT % 2 == 0, "16-byte copies move pairs");
#pragma unroll
    for (int e0 = 0; e0 < CNT; e0 += 2 * G) {
      const int e = e0 + 2 * c;
      if (e0 + 2 * G <= CNT || e < CNT) cp_async16(dst + e, src + e);
    }
  }
  template <int CNT>
  C2B_DEV void full8(double *dst, const double *src) const {
#pragma unroll
    for (int e0 = 0; e0 < CNT; e0 += G) {
      const int e = e0 + c;
      if (e0 + G <= CNT || e < CNT) cp_async8(dst + e, src + e);
    }
  }
  // rows of J-vectors (CNT = doubles of a full tile), scalars per row, time stamps: `fast` = full_tile
  template <int CNT>
  C2B_DEV void vec(double *dst, const double *src, int cnt, bool wide, bool fast) const {
    if (fast && wide) {
      if constexpr (CNT % 2 == 0) full16<CNT>(dst, src); else run(dst, src, cnt, wide);
    } else {
      run(dst, src, cnt, wide);
    }
  }
  template <int CNT>
  C2B_DEV void sca(double *dst, const double *src, int cnt, bool fast) const {
    if (fast) full8<CNT>(dst, src); else run(dst, src, cnt, false);
  }
  template <int CNT>
  C2B_DEV void tim(double *dst, const double *t, int64_t first, int cnt, int64_t lo, int64_t hi, bool fast) const {
    if (fast) full8<CNT>(dst, t + first); else run_t(dst, t, first, cnt, lo, hi);
  }
  // time stamps with one neighbour on either side: dst[e] = t[clamp(first + e, lo, hi)], e < cnt, [lo, hi] = [0, N-1]
  // for a whole series.  The clamp makes dt = 0 (propagator exactly 1) at the two ends of the series, which is what
  // the recursions want there; a segment of a longer series has a real neighbour at -1 and / or N instead.
  C2B_DEV void run_t(double *dst, const double *t, int64_t first, int cnt, int64_t lo, int64_t hi) const {
    for (int e = c; e < cnt; e += G) {
      int64_t i = first + e;
      i         = i < lo ? lo : (i > hi ? hi : i);
      cp_async8(dst + e, t + i);
    }
  }
};

// Per-group scratch arrays (propagators of a tile, exchanged row vectors) need the same treatment as the tiles: the groups
// of a half-warp (8-byte accesses) or quarter-warp (16-byte accesses) read the same offsets of their own arrays in one
// instruction, so a per-group stride that is a multiple of 16 doubles puts them on the same banks (r02p: 52 % of the
// shared-memory wavefronts of the J = 4 factor kernel were conflicts).  Stride = N padded to G doubles modulo 16.
template <int G, int N>
struct GPad {
  // (G = 2 packs 32 groups into a CTA: its arrays stay dense, or the padding alone would exceed the static limit)
  static constexpr int value = G < 4 ? N : N + ((G % 16) - N % 16 + 16) % 16;
};

// doubles per group for NBUF tile buffers of TB doubles each, padded like GRing::kStride
template <int G, int TB, int NBUF>
struct GTileStride {
  static constexpr int value = NBUF * TB + ((G % 16) - (NBUF * TB) % 16 + 16) % 16;
};

// Propagators exp(c_j dt) of a tile: pb[r * JE + j] for the R steps r of the tile and the J decay rates.  The exp is
// half of a summary pass (r01g profile).  A complex term contributes two columns with the same decay rate
// (terms.hpp:125; adjacent in the interleaved layout, Jc apart in the blocked one): two lanes that share their rate with
// nobody else split the rows of the tile between them and each stores its results for both, which halves the exp
// work of all-complex kernels (SHO, rotation, Matern terms).  Unpaired lanes evaluate all R rows.
template <int J>
struct PropPairs {
  int mate;     // the lane that shares my decay rate, or -1
  bool second;  // of a pair, I take the second half of the rows
  // csm: the J decay rates of my group in shared memory, visible to all its lanes
  C2B_DEV void init(const double *csm, int c) {
    mate    = -1;
    int cnt = 0;
    if (c < J) {
#pragma unroll
      for (int j = 0; j < J; ++j)
        if (j != c && csm[j] == csm[c]) {
          mate = j;
          ++cnt;
        }
    }
    if (cnt != 1) mate = -1;
    if (c >= J) mate = c;  // lanes without a column: nothing to evaluate, must not pull the warp into the unpaired path
    second = mate >= 0 && c > mate;
  }
  // dt(r): time difference of step r (any lane may ask for any step).  No barriers inside.
  template <int R, int JE, class DT>
  C2B_DEV void eval(double *pb, int c, double cc, DT dt) const {
    static_assert(R % 2 == 0, "tiles have an even number of rows");
    if (c >= J) return;
    const int r0 = second ? R / 2 : 0;
#pragma unroll
    for (int q = 0; q < R / 2; ++q) {
      const int r    = r0 + q;
      const double v = exp(cc * dt(r));
      pb[r * JE + c] = v;
      if (mate >= 0) pb[r * JE + mate] = v;
    }
    if (mate < 0) {
#pragma unroll
      for (int q = R / 2; q < R; ++q) pb[q * JE + c] = exp(cc * dt(q));
    }
  }
};

// Register version for the kernels where lane c only needs its own propagators: p[r], r < R.  A pair computes half
// of the tile each and swaps halves with shuffles (all 32 lanes must call).
template <int J, int G, int R, class DT>
C2B_DEV void eval_props_reg(const PropPairs<J> &pr, int c, double cc, DT dt, double *p) {
  static_assert(R % 2 == 0, "tiles have an even number of rows");
  const int r0 = pr.second ? R / 2 : 0;
  double mine[R / 2], other[R / 2];
#pragma unroll
  for (int q = 0; q < R / 2; ++q) mine[q] = exp(cc * dt(r0 + q));
  const int src = pr.mate >= 0 ? pr.mate : c;  // lane within the group
#pragma unroll
  for (int q = 0; q < R / 2; ++q) other[q] = __shfl_sync(0xffffffffu, mine[q], src, G);
  if (pr.mate < 0) {
#pragma unroll
    for (int q = 0; q < R / 2; ++q) other[q] = exp(cc * dt(R / 2 + q));
  }
#pragma unroll
  for (int q = 0; q < R / 2; ++q) {
    p[q]         = pr.second ? other[q] : mine[q];
    p[R / 2 + q] = pr.second ? mine[q] : other[q];
  }
}

// J consecutive doubles from 16-byte aligned shared memory
template <int J>
C2B_DEV void lds_row(const double *p, double *out) {
  if constexpr (J % 2 == 0) {
    const double2 *q = reinterpret_cast<const double2 *>(p);
#pragma unroll
    for (int j = 0; j < J / 2; ++j) {
      const double2 v = q[j];
      out[2 * j]      = v.x;
      out[2 * j + 1]  = v.y;
    }
  } else {
#pragma unroll
    for (int j = 0; j < J; ++j) out[j] = p[j];
  }
}

}  // namespace c2b
