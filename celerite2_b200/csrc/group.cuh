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

}  // namespace c2b
