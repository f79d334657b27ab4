// Generic exclusive scan of chunk summaries ("elements") into per-chunk carry-in states.
//
// The chunks of a series, taken in PROCESSING order (ascending n for forward sweeps, descending for backward
// ones), are the leaves of a binary tree stored level by level:
//   up-sweep    node[l][i] = combine(node[l-1][2i], node[l-1][2i+1])
//   down-sweep  state[l-1][2i] = state[l][i];  state[l-1][2i+1] = apply(node[l-1][2i], state[l][i])
// Levels 1..5 are handled by one warp per block of 32 leaves (work spread over the whole GPU), the levels
// above by one CTA per series, the way back down mirrors that.  The state entering the first chunk is zero or
// a caller-supplied carry (the hand-off point for sharding one series across GPUs).
#pragma once
#include "coop.cuh"

namespace c2b {

constexpr int kMaxLevels = 24;
constexpr int kWarpLevels = 5;  // 2^5 = 32 leaves per warp block

struct Levels {
  int nlev;                      // number of levels, count[nlev-1] == 1
  int64_t count[kMaxLevels];     // nodes per level (level 0 = leaves = chunks per series)
  int64_t off[kMaxLevels];       // node offset of each level inside one series
  int64_t total;                 // nodes per series
};

static inline Levels make_levels(int64_t leaves) {
  Levels lv;
  memset(&lv, 0, sizeof(lv));
  int l       = 0;
  int64_t cnt = leaves, off = 0;
  for (;;) {
    lv.count[l] = cnt;
    lv.off[l]   = off;
    off += cnt;
    ++l;
    if (cnt == 1) break;
    cnt = (cnt + 1) / 2;
  }
  lv.nlev  = l;
  lv.total = off;
  return lv;
}

template <class Mon>
C2B_DEV void up_node(int J, int K, const Levels &lv, double *nodes, int es, int l, int64_t idx) {
  const double *left = nodes + (lv.off[l - 1] + 2 * idx) * es;
  double *out        = nodes + (lv.off[l] + idx) * es;
  if (2 * idx + 1 < lv.count[l - 1]) {
    Mon::combine(J, K, left, left + es, out);
  } else {
    for (int i = 0; i < es; ++i) out[i] = left[i];
  }
}

template <class Mon>
C2B_DEV void down_node(int J, int K, const Levels &lv, const double *nodes, double *states, int es, int ss, int l, int64_t idx) {
  const double *sin = states + (lv.off[l] + idx) * ss;
  double *s0        = states + (lv.off[l - 1] + 2 * idx) * ss;
  for (int i = 0; i < ss; ++i) s0[i] = sin[i];
  if (2 * idx + 1 < lv.count[l - 1]) Mon::apply(J, K, nodes + (lv.off[l - 1] + 2 * idx) * es, sin, s0 + ss);
}

// one warp per (series, block of 32 leaves): levels 1..min(5, top)
template <class Mon>
__global__ void scan_up_warp(int J_, int K, int64_t batch, Levels lv, double *nodes_all) {
  const int J = Mon::kJT ? Mon::kJT : J_;  // compile-time width => unrolled, register-resident
  const int64_t nblk = (lv.count[0] + 31) / 32;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane     = threadIdx.x & 31;
  if (warp >= batch * nblk) return;
  const int64_t b = warp / nblk, blk = warp % nblk;
  const int es  = Mon::esize(J, K);
  double *nodes = nodes_all + b * lv.total * es;
  const int top = lv.nlev - 1 < kWarpLevels ? lv.nlev - 1 : kWarpLevels;
  for (int l = 1; l <= top; ++l) {
    const int64_t idx = ((blk * 32) >> l) + lane;
    if (lane < (32 >> l) && idx < lv.count[l]) up_node<Mon>(J, K, lv, nodes, es, l, idx);
    __syncwarp();
  }
}

// one CTA per series: levels 6..top up, then top..6 down (leaves states at level 5)
template <class Mon>
__global__ void scan_top(int J_, int K, Levels lv, double *nodes_all, double *states_all, const double *init) {
  const int J = Mon::kJT ? Mon::kJT : J_;  // compile-time width => unrolled, register-resident
  const int64_t b = blockIdx.x;
  const int es = Mon::esize(J, K), ss = Mon::ssize(J, K);
  double *nodes  = nodes_all + b * lv.total * es;
  double *states = states_all + b * lv.total * ss;
  const int top  = lv.nlev - 1;
  for (int l = kWarpLevels + 1; l <= top; ++l) {
    for (int64_t idx = threadIdx.x; idx < lv.count[l]; idx += blockDim.x) up_node<Mon>(J, K, lv, nodes, es, l, idx);
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    double *st = states + lv.off[top] * ss;
    for (int i = 0; i < ss; ++i) st[i] = init ? init[b * ss + i] : 0.0;
  }
  __syncthreads();
  for (int l = top; l > kWarpLevels; --l) {
    for (int64_t idx = threadIdx.x; idx < lv.count[l]; idx += blockDim.x) down_node<Mon>(J, K, lv, nodes, states, es, ss, l, idx);
    __syncthreads();
  }
}

// one warp per (series, block of 32 leaves): levels min(5, top)..1 down
template <class Mon>
__global__ void scan_down_warp(int J_, int K, int64_t batch, Levels lv, const double *nodes_all, double *states_all, const double *init) {
  const int J = Mon::kJT ? Mon::kJT : J_;  // compile-time width => unrolled, register-resident
  const int64_t nblk = (lv.count[0] + 31) / 32;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane     = threadIdx.x & 31;
  if (warp >= batch * nblk) return;
  const int64_t b = warp / nblk, blk = warp % nblk;
  const int es = Mon::esize(J, K), ss = Mon::ssize(J, K);
  const double *nodes = nodes_all + b * lv.total * es;
  double *states      = states_all + b * lv.total * ss;
  const int top       = lv.nlev - 1 < kWarpLevels ? lv.nlev - 1 : kWarpLevels;
  if (lv.nlev - 1 <= kWarpLevels) {  // the whole tree lives in this warp: seed the top state here
    if (lane == 0) {
      double *st = states + lv.off[top] * ss;
      for (int i = 0; i < ss; ++i) st[i] = init ? init[b * ss + i] : 0.0;
    }
    __syncwarp();
  }
  for (int l = top; l >= 1; --l) {
    const int64_t idx = ((blk * 32) >> l) + lane;
    if (lane < (32 >> l) && idx < lv.count[l]) down_node<Mon>(J, K, lv, nodes, states, es, ss, l, idx);
    __syncwarp();
  }
}

// ---- warp-cooperative variants (compile-time widths) -------------------------------------------------------------
template <class Mon>
C2B_DEV void wup_node(int lane, int J, int K, const Levels &lv, double *nodes, int es, int l, int64_t idx, double *sm) {
  const double *left = nodes + (lv.off[l - 1] + 2 * idx) * es;
  double *out        = nodes + (lv.off[l] + idx) * es;
  if (2 * idx + 1 < lv.count[l - 1]) {
    Coop<Mon>::wcombine(lane, J, K, left, left + es, out, sm);
  } else {
    wcopy(lane, es, left, out);
    __syncwarp();
  }
}

template <class Mon>
C2B_DEV void wdown_node(int lane, int J, int K, const Levels &lv, const double *nodes, double *states, int es, int ss, int l,
                        int64_t idx, double *sm) {
  const double *sin = states + (lv.off[l] + idx) * ss;
  double *s0        = states + (lv.off[l - 1] + 2 * idx) * ss;
  wcopy(lane, ss, sin, s0);
  if (2 * idx + 1 < lv.count[l - 1]) Coop<Mon>::wapply(lane, J, K, nodes + (lv.off[l - 1] + 2 * idx) * es, sin, s0 + ss, sm);
  __syncwarp();
}

constexpr int kCoopLeaves = 32;  // nodes per block
constexpr int kCoopLevels = 5;
constexpr int kCoopWarps  = 16;  // warps per CTA (fewer when the per-warp shared memory is large): one combine each per round

// One CTA per (series, block of 32 nodes at level `lb`): builds levels lb+1 .. min(lb+5, top).  Every level is one
// round of independent warp-cooperative combines, separated by CTA barriers.
template <class Mon>
__global__ void __launch_bounds__(kCoopWarps * 32) scan_up_block(int J_, int K, int64_t batch, Levels lv, int lb, double *nodes_all) {
  const int J = Mon::kJT ? Mon::kJT : J_;
  extern __shared__ double sm_all[];
  const int wsm      = Coop<Mon>::wsmem(J, K);
  const int wib      = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double *sm         = sm_all + (size_t)wib * wsm;
  const int64_t nblk = (lv.count[lb] + kCoopLeaves - 1) / kCoopLeaves;
  const int64_t b = blockIdx.x / nblk, blk = blockIdx.x % nblk;
  const int es  = Mon::esize(J, K);
  double *nodes = nodes_all + b * lv.total * es;
  const int top = lv.nlev - 1 < lb + kCoopLevels ? lv.nlev - 1 : lb + kCoopLevels;
  for (int l = lb + 1; l <= top; ++l) {
    const int sh        = l - lb;
    const int64_t first = (blk * kCoopLeaves) >> sh;
    for (int r = wib; r < (kCoopLeaves >> sh); r += (int)(blockDim.x >> 5))
      if (first + r < lv.count[l]) wup_node<Mon>(lane, J, K, lv, nodes, es, l, first + r, sm);
    __syncthreads();
  }
}

// Mirror image: states of levels min(lb+5, top) .. lb+1 are pushed down to level lb.  The CTA that owns the top of the
// tree seeds it with the initial state.
template <class Mon>
__global__ void __launch_bounds__(kCoopWarps * 32) scan_down_block(int J_, int K, int64_t batch, Levels lv, int lb,
                                                                    const double *nodes_all, double *states_all, const double *init) {
  const int J = Mon::kJT ? Mon::kJT : J_;
  extern __shared__ double sm_all[];
  const int wsm      = Coop<Mon>::wsmem(J, K);
  const int wib      = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double *sm         = sm_all + (size_t)wib * wsm;
  const int64_t nblk = (lv.count[lb] + kCoopLeaves - 1) / kCoopLeaves;
  const int64_t b = blockIdx.x / nblk, blk = blockIdx.x % nblk;
  const int es = Mon::esize(J, K), ss = Mon::ssize(J, K);
  const double *nodes = nodes_all + b * lv.total * es;
  double *states      = states_all + b * lv.total * ss;
  const int top       = lv.nlev - 1 < lb + kCoopLevels ? lv.nlev - 1 : lb + kCoopLevels;
  if (top == lv.nlev - 1) {
    double *st = states + lv.off[top] * ss;
    for (int i = threadIdx.x; i < ss; i += blockDim.x) st[i] = init ? init[b * ss + i] : 0.0;
    __syncthreads();
  }
  for (int l = top; l > lb; --l) {
    const int sh        = l - lb;
    const int64_t first = (blk * kCoopLeaves) >> sh;
    for (int r = wib; r < (kCoopLeaves >> sh); r += (int)(blockDim.x >> 5))
      if (first + r < lv.count[l]) wdown_node<Mon>(lane, J, K, lv, nodes, states, es, ss, l, first + r, sm);
    __syncthreads();
  }
}

// Scratch needed by a scan over `leaves` chunks per series.
template <class Mon>
static inline void scan_sizes(int J, int K, int64_t batch, int64_t leaves, size_t *node_doubles, size_t *state_doubles) {
  Levels lv      = make_levels(leaves);
  *node_doubles  = (size_t)batch * lv.total * Mon::esize(J, K);
  *state_doubles = (size_t)batch * lv.total * Mon::ssize(J, K);
}

// Leaves must already be in nodes[b][0..leaves).  On return states[b][i] (i < leaves) is the carry into chunk i.
template <class Mon>
static inline int run_scan(int J, int K, int64_t batch, int64_t leaves, double *nodes, double *states, const double *init,
                           cudaStream_t stream) {
  Levels lv          = make_levels(leaves);
  const int64_t nblk = (leaves + 31) / 32;
  if constexpr (Mon::kCoop) {
    const int wsm     = Coop<Mon>::wsmem(J, K);
    int nwarps = kCoopWarps;
    while (nwarps > 1 && (size_t)nwarps * wsm * sizeof(double) > 200 * 1024) nwarps >>= 1;
    const size_t smem = (size_t)nwarps * wsm * sizeof(double);
    if (smem > 48 * 1024) {
      C2B_CUDA(cudaFuncSetAttribute(scan_up_block<Mon>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      C2B_CUDA(cudaFuncSetAttribute(scan_down_block<Mon>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    }
    const int top = lv.nlev - 1;
    int bases[kMaxLevels], nb = 0;
    for (int lb = 0; lb < top || nb == 0; lb += kCoopLevels) bases[nb++] = lb;
    for (int q = 0; q < nb; ++q) {
      const int lb        = bases[q];
      const unsigned grid = (unsigned)(batch * ((lv.count[lb] + kCoopLeaves - 1) / kCoopLeaves));
      if (lb < top)
        C2B_LAUNCH("scan_up_block", stream, scan_up_block<Mon><<<grid, nwarps * 32, smem, stream>>>(J, K, batch, lv, lb, nodes));
    }
    for (int q = nb - 1; q >= 0; --q) {
      const int lb        = bases[q];
      const unsigned grid = (unsigned)(batch * ((lv.count[lb] + kCoopLeaves - 1) / kCoopLeaves));
      C2B_LAUNCH("scan_down_block", stream,
                 scan_down_block<Mon><<<grid, nwarps * 32, smem, stream>>>(J, K, batch, lv, lb, nodes, states, init));
    }
    return C2B_OK;
  } else {
  const int threads  = 128;  // 4 warps
  const int64_t warps = batch * nblk;
  const unsigned grid = (unsigned)((warps * 32 + threads - 1) / threads);
  if (lv.nlev > 1) {
    C2B_LAUNCH("scan_up_warp", stream, scan_up_warp<Mon><<<grid, threads, 0, stream>>>(J, K, batch, lv, nodes));
  }
  if (lv.nlev - 1 > kWarpLevels) {
    C2B_LAUNCH("scan_top", stream, scan_top<Mon><<<(unsigned)batch, 256, 0, stream>>>(J, K, lv, nodes, states, init));
  }
  C2B_LAUNCH("scan_down_warp", stream, scan_down_warp<Mon><<<grid, threads, 0, stream>>>(J, K, batch, lv, nodes, states, init));
  return C2B_OK;
  }
}

// Finalise a reverse op: add the bt term each chunk owes its neighbour (one thread per chunk) and, in the first CTA
// of every series, reduce the per-chunk bc partials in a fixed order (deterministic).  grid = (ceil(C/1024), batch).
//   descending == true  (lower / factor_rev): chunk k owes bt[n0_k - 1]
//   descending == false (upper)             : chunk k owes bt[n1_k]
static __global__ void __launch_bounds__(1024) rev_finalize_kernel(int64_t N, int64_t C, int64_t L, int J, bool descending,
                                                                   const double *bc_part, const double *bt_edge, double *bt, double *bc) {
  const int64_t b = blockIdx.y;
  const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k < C) {
    const int64_t n0 = k * L, n1 = (n0 + L < N) ? n0 + L : N;
    const int64_t tgt = descending ? n0 - 1 : n1;
    if (tgt >= 0 && tgt < N) bt[b * N + tgt] += bt_edge[b * C + k];
  }
  if (blockIdx.x != 0) return;
  // bc[j] = sum over chunks: strided partial sums per thread, then a fixed-shape tree
  __shared__ double sh[32];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
  for (int j = 0; j < J; ++j) {
    double s = 0.0;
    for (int64_t q = threadIdx.x; q < C; q += blockDim.x) s += bc_part[(b * C + q) * J + j];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) sh[wid] = s;
    __syncthreads();
    if (wid == 0) {
      double v = lane < nw ? sh[lane] : 0.0;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0) bc[b * J + j] = v;
    }
    __syncthreads();
  }
}

}  // namespace c2b
