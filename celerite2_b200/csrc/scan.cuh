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
#include <atomic>

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
  lv.total = off + 2;  // stride of one series in the scratch arrays: two spare slots behind the levels (fold.cuh's aggregates)
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
  pdl_prologue();
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
  pdl_prologue();
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
  pdl_prologue();
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

// ---- block-resident tree kernels ----------------------------------------------------------------------------------
// One CTA owns a block of LB = 2^nlev nodes of level `lb` and the nlev tree levels above them.  The block's nodes are
// loaded into shared memory once (coalesced), every level is one round of independent warp-cooperative combines
// from shared memory to shared memory separated by CTA barriers, and the new nodes go back to global memory in one
// coalesced pass at the end: one L2 round trip per launch instead of three per level (r01f: 2.5 us per level for a
// 6 x 6 product).  The down-sweep mirrors that with a tree of states; the launch that holds the top of the tree
// does both directions without leaving shared memory.
//
// Local tree layout: local level q (global level lb + q) has LB >> q slots starting at slot 2 LB - (2 LB >> q).
struct BlockTree {
  int nlev, LB, nl;     // levels per launch, 2^nlev, levels actually present above lb in this launch
  int64_t blk, b;       // block index within the series, series
  int cnt[8];           // valid nodes per local level
  C2B_DEV int off(int q) const { return 2 * LB - ((2 * LB) >> q); }
};

C2B_DEV BlockTree make_block_tree(const Levels &lv, int lb, int nlev) {
  BlockTree bt;
  bt.nlev            = nlev;
  bt.LB              = 1 << nlev;
  const int64_t nblk = (lv.count[lb] + bt.LB - 1) / bt.LB;
  bt.b               = blockIdx.x / nblk;
  bt.blk             = blockIdx.x % nblk;
  const int top      = lv.nlev - 1;
  bt.nl              = top - lb < nlev ? top - lb : nlev;
  const int64_t rem  = lv.count[lb] - bt.blk * bt.LB;
  bt.cnt[0]          = rem < bt.LB ? (int)rem : bt.LB;
  for (int q = 1; q <= bt.nl; ++q) bt.cnt[q] = (bt.cnt[q - 1] + 1) >> 1;
  return bt;
}

template <class Mon>
C2B_DEV void block_up(const BlockTree &bt, int J, int K, int es, double *tree, double *sm, int wib, int lane, int nwarps) {
  for (int q = 1; q <= bt.nl; ++q) {
    for (int r = wib; r < bt.cnt[q]; r += nwarps) {
      const double *left = tree + (size_t)(bt.off(q - 1) + 2 * r) * es;
      double *out        = tree + (size_t)(bt.off(q) + r) * es;
      if (2 * r + 1 < bt.cnt[q - 1]) {
        Coop<Mon>::wcombine(lane, J, K, left, left + es, out, sm);
      } else {
        wcopy(lane, es, left, out);
        __syncwarp();
      }
    }
    __syncthreads();
  }
}

// states of local level bt.nl must be in `st`; fills the levels below
template <class Mon>
C2B_DEV void block_down(const BlockTree &bt, int J, int K, int es, int ss, const double *tree, double *st, double *sm, int wib, int lane,
                        int nwarps) {
  for (int q = bt.nl; q >= 1; --q) {
    for (int r = wib; r < bt.cnt[q]; r += nwarps) {
      const double *sin = st + (size_t)(bt.off(q) + r) * ss;
      double *s0        = st + (size_t)(bt.off(q - 1) + 2 * r) * ss;
      wcopy(lane, ss, sin, s0);
      if (2 * r + 1 < bt.cnt[q - 1]) Coop<Mon>::wapply(lane, J, K, tree + (size_t)(bt.off(q - 1) + 2 * r) * es, sin, s0 + ss, sm);
      __syncwarp();
    }
    __syncthreads();
  }
}

constexpr int kCoopThreadsMax = 512;

// Elements of the multi-right-hand-side sweeps are a J x J head plus J doubles per column, and with many columns a
// whole element no longer fits a block-resident tree.  The columns do not interact, so the scan is cut into slices
// of kb columns (blockIdx.y): every slice keeps [head | its columns] in shared memory, recomputes the (cheap) head
// products, and only slice 0 writes the heads back.
// n local elements of `loc` doubles  <->  global elements of `glob` doubles: [0, head) maps to [0, head), the rest
// to goff + (r - head)
C2B_DEV void slice_in(double *dst, const double *src, int n, int loc, int glob, int head, int goff) {
  for (int e = threadIdx.x; e < n * loc; e += blockDim.x) {
    const int i = e / loc, r = e - i * loc;
    dst[e]      = src[(size_t)i * glob + (r < head ? r : r - head + goff)];
  }
}
C2B_DEV void slice_out(double *dst, const double *src, int n, int loc, int glob, int head, int goff, bool with_head) {
  for (int e = threadIdx.x; e < n * loc; e += blockDim.x) {
    const int i = e / loc, r = e - i * loc;
    if (r >= head || with_head) dst[(size_t)i * glob + (r < head ? r : r - head + goff)] = src[e];
  }
}


}  // namespace c2b
#include "fold.cuh"
namespace c2b {

// ---- prefix blocks (round 2) ---------------------------------------------------------------------------------------
// The tree above spends one combine latency per level on the way up AND one apply latency per level on the way down
// (26 dependent rounds for 7,000 chunks; r01i profile: the scans are 27 % of the c2 step at ~0 bytes moved -- pure
// depth x latency).  Here a CTA owns a block of BS = 2 nwarps consecutive elements of one "level group" and turns them
// into their INCLUSIVE PREFIXES P_i = e_0 o ... o e_i inside shared memory with the divide-and-conquer (Sklansky)
// network: log2(BS) rounds, BS / 2 independent warp-cooperative combines per round, one per warp.  The block's last
// prefix is the block aggregate and becomes one element of the next level group.  On the way down a block receives
// the state s entering it and produces the states entering its elements in ONE round of independent applies:
// state_0 = s, state_i = apply(P_{i-1}, s).  Depth for C = 7,104: 5 + 5 + 3 combine rounds up, 1 + 2 + 2 apply rounds
// down (BS = 32) instead of 13 + 13.
//
// Level groups: group 0 = the C leaves, group g + 1 = the ceil(cnt_g / BS) block aggregates of group g; the top group
// has at most BS elements.  nodes[off_g + i] holds element i of group g, overwritten in place by its prefix within its
// block; states[off_g + i] the state entering it.
struct PLevels {
  int ng;            // number of level groups; cnt[ng-1] <= BS
  int64_t cnt[12];   // elements per group
  int64_t off[12];   // offset of each group inside one series (in elements)
};

static inline PLevels make_plevels(int64_t leaves, int BS) {
  PLevels pl;
  memset(&pl, 0, sizeof(pl));
  int g       = 0;
  int64_t cnt = leaves, off = 0;
  for (;;) {
    pl.cnt[g] = cnt;
    pl.off[g] = off;
    off += cnt;
    ++g;
    if (cnt <= BS) break;
    cnt = (cnt + BS - 1) / BS;
  }
  pl.ng = g;
  return pl;
}

// where element i of a block lives after the rounds of the network: it moved to the other buffer every time it was
// updated, i.e. once per set bit of its index within the block
C2B_DEV int pbuf_of(int i) { return __popc((unsigned)i) & 1; }

// one CTA per (series, block of BS elements of group g) and column slice (blockIdx.y).
//   do_up   : scan the block (elements -> prefixes, written back in place); not the top group: aggregate -> group g + 1
//   do_down : states of the block's elements from the state entering the block: the top group is seeded from `init`
//             (NULL = zero state), the others read states[off_{g+1} + blk]
template <class Mon>
__global__ void __launch_bounds__(kCoopThreadsMax) pscan_block(int J_, int K, int kb, PLevels pl, int g, int BS, int64_t stride,
                                                              bool do_up, bool do_down, double *nodes_all, double *states_all,
                                                              const double *init) {
  pdl_prologue();
  const int J = Mon::kJT ? Mon::kJT : J_;
  extern __shared__ __align__(16) double sm_all[];
  const int k0  = blockIdx.y * kb;
  const int kl  = K - k0 < kb ? K - k0 : kb;
  const int esg = Mon::esize(J, K), ssg = Mon::ssize(J, K);
  const int es = Mon::esize(J, kl), ss = Mon::ssize(J, kl);
  const int eh = Mon::ehead(J), sh = Mon::shead(J);
  const int ego = eh + Mon::ecol(J) * k0, sgo = sh + Mon::scol(J) * k0;
  const int wsm    = Coop<Mon>::wsmem(J, kl);
  const int nwarps = blockDim.x >> 5, wib = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool top     = g == pl.ng - 1;
  const int64_t nblk = (pl.cnt[g] + BS - 1) / BS;
  const int64_t b = blockIdx.x / nblk, blk = blockIdx.x % nblk;
  const int64_t rem = pl.cnt[g] - blk * BS;
  const int cnt     = rem < BS ? (int)rem : BS;
  double *const buf0 = sm_all, *const buf1 = buf0 + (size_t)BS * es;
  auto bsel          = [&](int p) { return p ? buf1 : buf0; };
  double *sin_ = buf1 + (size_t)BS * es;            // ss: the state entering the block
  double *sout = sin_ + ss;                         // BS * ss
  double *sm   = sout + (size_t)BS * ss + (size_t)wib * wsm;
  double *elems = nodes_all + (b * stride + pl.off[g] + blk * BS) * esg;
  slice_in(buf0, elems, cnt, es, esg, eh, ego);
  __syncthreads();
  if (do_up) {
    for (int s = 0; (1 << s) < cnt; ++s) {
      // the BS / 2 elements with bit s set combine with the last element of the lower half of their 2^(s+1) group
      for (int r = wib; r < BS / 2; r += nwarps) {
        const int i = ((r >> s) << (s + 1)) | (1 << s) | (r & ((1 << s) - 1));
        if (i < cnt) {
          const int j  = ((i >> s) << s) - 1;
          const int lo = i & ((1 << s) - 1);  // bits of i below s: the updates i has seen so far
          const double *ej = bsel(pbuf_of(j & ((1 << s) - 1))) + (size_t)j * es;
          const double *ei = bsel(pbuf_of(lo)) + (size_t)i * es;
          double *eo       = bsel(pbuf_of(lo) ^ 1) + (size_t)i * es;
          Coop<Mon>::wcombine(lane, J, kl, ej, ei, eo, sm);
        }
      }
      __syncthreads();
    }
    // prefixes back in place (element i sits in buffer popc(i) & 1)
    for (int e = threadIdx.x; e < cnt * es; e += blockDim.x) {
      const int i = e / es, r = e - i * es;
      if (r >= eh || blockIdx.y == 0) elems[(size_t)i * esg + (r < eh ? r : r - eh + ego)] = bsel(pbuf_of(i))[e];
    }
    if (!top) {
      double *agg       = nodes_all + (b * stride + pl.off[g + 1] + blk) * esg;
      const double *src = bsel(pbuf_of(cnt - 1)) + (size_t)(cnt - 1) * es;
      for (int r = threadIdx.x; r < es; r += blockDim.x)
        if (r >= eh || blockIdx.y == 0) agg[r < eh ? r : r - eh + ego] = src[r];
    }
  }
  if (do_down) {
    double *states = states_all + b * stride * ssg;
    if (top) {
      for (int e = threadIdx.x; e < ss; e += blockDim.x) sin_[e] = init ? init[b * ssg + (e < sh ? e : e - sh + sgo)] : 0.0;
    } else {
      slice_in(sin_, states + (pl.off[g + 1] + blk) * ssg, 1, ss, ssg, sh, sgo);
    }
    __syncthreads();
    for (int i = wib; i < cnt; i += nwarps) {
      if (i == 0) {
        wcopy(lane, ss, sin_, sout);
      } else {
        const double *e = (do_up ? bsel(pbuf_of(i - 1)) : buf0) + (size_t)(i - 1) * es;
        Coop<Mon>::wapply(lane, J, kl, e, sin_, sout + (size_t)i * ss, sm);
      }
    }
    __syncthreads();
    slice_out(states + (pl.off[g] + blk * BS) * ssg, sout, cnt, ss, ssg, sh, sgo, true);
  }
}

// DIR: 0 = up only, 1 = down only, 2 = up then down (the launch that contains the top of the tree)
template <class Mon, int DIR>
__global__ void __launch_bounds__(kCoopThreadsMax) scan_block(int J_, int K, int kb, int64_t batch, Levels lv, int lb, int nlev,
                                                             double *nodes_all, double *states_all, const double *init) {
  pdl_prologue();
  const int J = Mon::kJT ? Mon::kJT : J_;
  extern __shared__ __align__(16) double sm_all[];
  const int k0  = blockIdx.y * kb;                       // my columns: [k0, k0 + kl)
  const int kl  = K - k0 < kb ? K - k0 : kb;
  const int esg = Mon::esize(J, K), ssg = Mon::ssize(J, K);   // global element / state
  const int es = Mon::esize(J, kl), ss = Mon::ssize(J, kl);   // local (my slice)
  const int eh = Mon::ehead(J), sh = Mon::shead(J);
  const int ego = eh + Mon::ecol(J) * k0, sgo = sh + Mon::scol(J) * k0;
  const int wsm    = Coop<Mon>::wsmem(J, kl);
  const int nwarps = blockDim.x >> 5, wib = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const BlockTree bt = make_block_tree(lv, lb, nlev);
  const int slots    = 2 * bt.LB - 1;
  double *tree = sm_all;                                         // slots * es
  double *st   = tree + (size_t)slots * es;                      // slots * ss (DIR >= 1)
  double *sm   = st + (DIR >= 1 ? (size_t)slots * ss : 0) + (size_t)wib * wsm;
  double *nodes  = nodes_all + bt.b * lv.total * esg;
  double *states = DIR >= 1 ? states_all + bt.b * lv.total * ssg : nullptr;
  // load: level lb nodes (up) / all present local levels below the top one (down)
  const int qload = DIR == 1 ? bt.nl - 1 : 0;
  for (int q = 0; q <= qload; ++q)
    slice_in(tree + (size_t)bt.off(q) * es, nodes + (lv.off[lb + q] + ((bt.blk * bt.LB) >> q)) * esg, bt.cnt[q], es, esg, eh, ego);
  if (DIR == 1)
    slice_in(st + (size_t)bt.off(bt.nl) * ss, states + (lv.off[lb + bt.nl] + ((bt.blk * bt.LB) >> bt.nl)) * ssg, bt.cnt[bt.nl], ss, ssg,
             sh, sgo);
  __syncthreads();
  if (DIR != 1) {
    block_up<Mon>(bt, J, kl, es, tree, sm, wib, lane, nwarps);
    for (int q = 1; q <= bt.nl; ++q)
      slice_out(nodes + (lv.off[lb + q] + ((bt.blk * bt.LB) >> q)) * esg, tree + (size_t)bt.off(q) * es, bt.cnt[q], es, esg, eh, ego,
                blockIdx.y == 0);
  }
  if (DIR == 2) {
    // this CTA holds the root: seed it
    double *dst = st + (size_t)bt.off(bt.nl) * ss;
    for (int e = threadIdx.x; e < ss; e += blockDim.x) dst[e] = init ? init[bt.b * ssg + (e < sh ? e : e - sh + sgo)] : 0.0;
    __syncthreads();
  }
  if (DIR >= 1) {
    block_down<Mon>(bt, J, kl, es, ss, tree, st, sm, wib, lane, nwarps);
    slice_out(states + (lv.off[lb] + bt.blk * bt.LB) * ssg, st, bt.cnt[0], ss, ssg, sh, sgo, true);
  }
}


// Which scan runs (measured on the B200, profiles/r02b_ab_runs.json, r02c_*):
//   trees         (scan.cuh, round 1)  64 us (AffV), 72 us (Cong), 128 us (Lft) at C = 7,104, J = 6; run when K > 1 and the
//                 algebra is not affine (there is none today) or on request;
//   blocked fold  (fold.cuh)           the default for the affine algebras with K > 1: the J x K state makes a tree
//                 combine a warp-serial J x J x K product (c4: 255 us per sweep), a fold step is one CTA-parallel apply
//                 (134 us per sweep).  At K = 1 its ~170 steps of ~0.6 us lose to the tree (102 us vs 64 us);
//   prefix blocks (below)              the default for one right-hand side: 2.5 % faster than the trees on the c2 step
//                 (r02g, r02k: 0.971 -> 0.946 ms; full GPU suite green with them).  Their in-place prefixes race between
//                 column slices, so they never run for K > 1.
// C2B_SCAN (environment, read once) overrides: "tree", "prefix", "fold" (every affine scan), "fold+prefix".
static inline const char *scan_mode() {
  static const char *v = [] {
    const char *e = getenv("C2B_SCAN");
    return (e && e[0]) ? e : "auto";
  }();
  return v;
}
// fold: every affine scan with K > 1.  With one right-hand side both fold variants lost to the prefix blocks on the c2
// step: CTA-wide walks 102 us per scan (r02c), one-warp walks without barriers 71 us (r02n, commit 33fd0f1: a lone warp
// issues one instruction per ~6 cycles, 139 instructions per step), prefix blocks 52 us.
template <class Mon>
static inline bool scan_use_fold(int K, int64_t) {
  if (strstr(scan_mode(), "fold")) return true;
  if (strstr(scan_mode(), "auto")) return K > 1;
  return false;
}
static inline bool scan_use_tree(int K) {
  if (K > 1) return true;
  return strstr(scan_mode(), "prefix") == nullptr && strstr(scan_mode(), "auto") == nullptr;
}

// Block size / warps / columns per slice of the prefix-block scan: the largest block whose shared memory (two element
// buffers, the states, the warps' scratch) lets two CTAs share an SM, else one CTA per SM, narrower slices last.
template <class Mon>
struct PScanPlan {
  int BS, nwarps, kb;
  size_t smem;
  bool ok;
};
template <class Mon>
static inline size_t pscan_smem(int J, int kb, int BS, int nwarps) {
  return ((size_t)2 * BS * Mon::esize(J, kb) + (size_t)(BS + 1) * Mon::ssize(J, kb) + (size_t)nwarps * Coop<Mon>::wsmem(J, kb)) * sizeof(double);
}
template <class Mon>
static inline PScanPlan<Mon> pscan_plan(int J, int K) {
  static const int cand[] = {32, 16, 8, 4, 2};
  PScanPlan<Mon> pp{2, 1, K, 0, false};
  const int ec = Mon::ecol(J) + Mon::scol(J);
  for (int pass = 0; pass < 2 && !pp.ok; ++pass) {
    const size_t budget = pass == 0 ? 110 * 1024 : 200 * 1024;
    for (int BS : cand) {
      if (pass == 0 && BS < 16) break;
      const int nwarps = BS / 2;
      int kb           = K;
      if (ec > 0) {
        while (kb > 1 && pscan_smem<Mon>(J, kb, BS, nwarps) > budget) kb = (kb + 1) / 2;
        if (kb < (K < 4 ? K : 4) && BS > 2) continue;  // slices that narrow: prefer a smaller block
        const int nsl = (K + kb - 1) / kb;
        kb            = (K + nsl - 1) / nsl;  // balanced slices
      }
      const size_t smem = pscan_smem<Mon>(J, kb, BS, nwarps);
      if (smem <= budget) {
        pp = PScanPlan<Mon>{BS, nwarps, kb, smem, true};
        break;
      }
    }
  }
  return pp;
}

// mode 0: everything.  mode 1: up only (the aggregate of the whole series is the last prefix of the top group).
// mode 2: down only, the state entering the series comes from `init` (NULL = zero).
template <class Mon>
static inline int run_pscan(int J, int K, int64_t batch, int64_t leaves, double *nodes, double *states, const double *init,
                            cudaStream_t stream, int mode) {
  const PScanPlan<Mon> pp = pscan_plan<Mon>(J, K);
  if (!pp.ok) {
    set_error("scan: summary element too large for a shared-memory block (J = %d, K = %d)", J, K);
    return C2B_ERR_INVALID_ARGUMENT;
  }
  const Levels lv  = make_levels(leaves);  // lv.total is the per-series stride of the scratch arrays (host_common.cuh: take_scan)
  const PLevels pl = make_plevels(leaves, pp.BS);
  static std::atomic<bool> attr_done[64];  // per instantiation and device; a race only repeats the (idempotent) call
  int dev = 0;
  C2B_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !attr_done[dev].load(std::memory_order_acquire)) {
    C2B_CUDA(cudaFuncSetAttribute(pscan_block<Mon>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
    if (dev >= 0 && dev < 64) attr_done[dev].store(true, std::memory_order_release);
  }
  const unsigned nslices = (unsigned)((K + pp.kb - 1) / pp.kb);
  const int top          = pl.ng - 1;
  auto launch = [&](const char *name, int g, bool up, bool down) -> int {
    const dim3 grid((unsigned)(batch * ((pl.cnt[g] + pp.BS - 1) / pp.BS)), nslices);
    C2B_LAUNCH(name, stream,
               (launch_k(pscan_block<Mon>, grid, pp.nwarps * 32, pp.smem, stream, J, K, pp.kb, pl, g, pp.BS, lv.total, up, down, nodes, states, init)));
    return C2B_OK;
  };
  int rc;
  if (mode != 2)
    for (int g = 0; g < top; ++g)
      if ((rc = launch("scan_up_block", g, true, false))) return rc;
  if (mode == 0) {
    if ((rc = launch("scan_top_block", top, true, true))) return rc;
  } else if (mode == 1) {
    if (pl.cnt[top] > 1 && (rc = launch("scan_up_block", top, true, false))) return rc;
  } else {
    if ((rc = launch("scan_down_block", top, false, true))) return rc;
  }
  if (mode != 1)
    for (int g = top - 1; g >= 0; --g)
      if ((rc = launch("scan_down_block", g, false, true))) return rc;
  return C2B_OK;
}

// Scratch needed by a scan over `leaves` chunks per series.
template <class Mon>
static inline void scan_sizes(int J, int K, int64_t batch, int64_t leaves, size_t *node_doubles, size_t *state_doubles) {
  Levels lv      = make_levels(leaves);
  *node_doubles  = (size_t)batch * lv.total * Mon::esize(J, K);
  *state_doubles = (size_t)batch * lv.total * Mon::ssize(J, K);
}

// Leaves must already be in nodes[b][0..leaves).  On return states[b][i] (i < leaves) is the carry into chunk i.
// mode 0: everything.  mode 1: up-sweep only (the root summary ends up in nodes[b][lv.off[top]]).  mode 2: down-sweep
// only, from a root state the caller has put into states[b][lv.off[top]] (segments of a longer series).
template <class Mon>
static inline int run_scan(int J, int K, int64_t batch, int64_t leaves, double *nodes, double *states, const double *init,
                           cudaStream_t stream, int mode = 0) {
  Levels lv          = make_levels(leaves);
  const int64_t nblk = (leaves + 31) / 32;
  if constexpr (Mon::kCoop && Fold<Mon>::kAffine) {
    if (scan_use_fold<Mon>(K, leaves)) {
      const Levels lvf = make_levels(leaves);
      double *heads    = nodes + (size_t)batch * lvf.total * Mon::esize(J, K);  // behind the nodes of all series (take_scan)
      return run_fold<Mon>(J, K, batch, leaves, lvf.total, nodes, states, heads, init, stream, mode);
    }
  }
  if constexpr (Mon::kCoop) {
    if (!scan_use_tree(K)) return run_pscan<Mon>(J, K, batch, leaves, nodes, states, init, stream, mode);
    // levels per launch / warps per CTA / columns per slice: the largest configuration whose shared memory (sized
    // for the launch that holds nodes and states) lets two CTAs share an SM
    static const int cand[][2] = {{5, 16}, {5, 8}, {4, 8}, {4, 4}, {3, 4}, {2, 2}, {1, 1}};
    const int eh = Mon::ehead(J) + Mon::shead(J), ec = Mon::ecol(J) + Mon::scol(J);
    int nlev = 1, nwarps = 1, kb = K;
    size_t smem = 0;
    bool found = false;
    // first choice: at least four levels per launch with two CTAs per SM; otherwise one CTA per SM
    for (int pass = 0; pass < 2 && !found; ++pass) {
      const long budget_bytes = pass == 0 ? 110 * 1024 : 200 * 1024;
      for (auto &cd : cand) {
        if (pass == 0 && cd[0] < 4) break;
        const long slots  = (2L << cd[0]) - 1;
        nlev              = cd[0];
        nwarps            = cd[1];
        const long budget = budget_bytes / (long)sizeof(double);
        kb                = K;
        if (ec > 0) {
          // widest slice that fits: per-warp scratch is taken at the full width (an upper bound)
          const long per_slot = (budget - (long)nwarps * Coop<Mon>::wsmem(J, K)) / slots;
          long fit            = (per_slot - eh) / ec;
          if (fit > K) fit = K;
          kb = (int)fit;
        }
        const bool ok = ec > 0 ? kb >= (K < 4 ? K : 4) : true;
        if (kb < 1) kb = 1;
        const int nsl = (K + kb - 1) / kb;
        kb            = (K + nsl - 1) / nsl;  // balanced slices
        smem = ((size_t)slots * (Mon::esize(J, kb) + Mon::ssize(J, kb)) + (size_t)nwarps * Coop<Mon>::wsmem(J, kb)) * sizeof(double);
        if (ok && smem <= (size_t)budget_bytes) {
          found = true;
          break;
        }
      }
    }
    if (smem > 220 * 1024) {
      set_error("scan: summary element too large for the block-resident tree (%zu bytes of shared memory)", smem);
      return C2B_ERR_INVALID_ARGUMENT;
    }
    const unsigned nslices = (unsigned)((K + kb - 1) / kb);
    static std::atomic<bool> attr_done[64];  // per instantiation and device; a race only repeats the (idempotent) calls
    int dev = 0;
    C2B_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_done[dev].load(std::memory_order_acquire)) {
      C2B_CUDA(cudaFuncSetAttribute(scan_block<Mon, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
      C2B_CUDA(cudaFuncSetAttribute(scan_block<Mon, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
      C2B_CUDA(cudaFuncSetAttribute(scan_block<Mon, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
      if (dev >= 0 && dev < 64) attr_done[dev].store(true, std::memory_order_release);
    }
    const int top = lv.nlev - 1;
    int bases[kMaxLevels], nb = 0;
    for (int lb = 0; lb < top || nb == 0; lb += nlev) bases[nb++] = lb;
    auto grid_of = [&](int lb) { return dim3((unsigned)(batch * ((lv.count[lb] + (1 << nlev) - 1) >> nlev)), nslices); };
    if (mode == 0) {
      for (int q = 0; q + 1 < nb; ++q)
        C2B_LAUNCH("scan_up_block", stream,
                   (launch_k(scan_block<Mon, 0>, grid_of(bases[q]), nwarps * 32, smem, stream, J, K, kb, batch, lv, bases[q], nlev, nodes, states, init)));
      C2B_LAUNCH("scan_top_block", stream,
                 (launch_k(scan_block<Mon, 2>, grid_of(bases[nb - 1]), nwarps * 32, smem, stream, J, K, kb, batch, lv, bases[nb - 1], nlev, nodes, states, init)));
      for (int q = nb - 2; q >= 0; --q)
        C2B_LAUNCH("scan_down_block", stream,
                   (launch_k(scan_block<Mon, 1>, grid_of(bases[q]), nwarps * 32, smem, stream, J, K, kb, batch, lv, bases[q], nlev, nodes, states, init)));
    } else if (mode == 1) {
      if (top > 0)
        for (int q = 0; q < nb; ++q)
          C2B_LAUNCH("scan_up_block", stream,
                     (launch_k(scan_block<Mon, 0>, grid_of(bases[q]), nwarps * 32, smem, stream, J, K, kb, batch, lv, bases[q], nlev, nodes, states, init)));
    } else {
      if (top > 0)
        for (int q = nb - 1; q >= 0; --q)
          C2B_LAUNCH("scan_down_block", stream,
                     (launch_k(scan_block<Mon, 1>, grid_of(bases[q]), nwarps * 32, smem, stream, J, K, kb, batch, lv, bases[q], nlev, nodes, states, init)));
    }
    return C2B_OK;
  } else {
  if (mode != 0) {
    set_error("segments of a longer series need a compiled width (J in 1..16)");
    return C2B_ERR_INVALID_ARGUMENT;
  }
  const int threads  = 128;  // 4 warps
  const int64_t warps = batch * nblk;
  const unsigned grid = (unsigned)((warps * 32 + threads - 1) / threads);
  if (lv.nlev > 1) {
    C2B_LAUNCH("scan_up_warp", stream, launch_k(scan_up_warp<Mon>, grid, threads, 0, stream, J, K, batch, lv, nodes));
  }
  if (lv.nlev - 1 > kWarpLevels) {
    C2B_LAUNCH("scan_top", stream, launch_k(scan_top<Mon>, (unsigned)batch, 256, 0, stream, J, K, lv, nodes, states, init));
  }
  C2B_LAUNCH("scan_down_warp", stream, launch_k(scan_down_warp<Mon>, grid, threads, 0, stream, J, K, batch, lv, nodes, states, init));
  return C2B_OK;
  }
}

// Segment plumbing shared by the launchers (batch == 1): where the root of the tree lives
template <class Mon>
static inline int seg_export_root(int J, int K, int64_t leaves, const double *nodes, double *summary_out, cudaStream_t stream) {
  const Levels lv = make_levels(leaves);
  const int es    = Mon::esize(J, K);
  int64_t root    = lv.off[lv.nlev - 1];
  if constexpr (Mon::kCoop && Fold<Mon>::kAffine) {
    if (scan_use_fold<Mon>(K, leaves)) {  // fold: the aggregate of the series sits behind the group aggregates
      const FoldPlan fp = fold_plan<Mon>(J, K, leaves, lv.total);
      C2B_CUDA(cudaMemcpyAsync(summary_out, nodes + (fp.C + fp.ng) * es, sizeof(double) * es, cudaMemcpyDeviceToDevice, stream));
      return C2B_OK;
    }
  }
  if constexpr (Mon::kCoop) {
    if (!scan_use_tree(K)) {  // prefix blocks: the aggregate of the series is the last prefix of the top group
      const PLevels pl = make_plevels(leaves, pscan_plan<Mon>(J, K).BS);
      root             = pl.off[pl.ng - 1] + pl.cnt[pl.ng - 1] - 1;
    }
  }
  C2B_CUDA(cudaMemcpyAsync(summary_out, nodes + root * es, sizeof(double) * es, cudaMemcpyDeviceToDevice, stream));
  return C2B_OK;
}
template <class Mon>
static inline int seg_import_state(int J, int K, int64_t leaves, double *states, const double *state_in, cudaStream_t stream) {
  if constexpr (Mon::kCoop && Fold<Mon>::kAffine) {
    if (scan_use_fold<Mon>(K, leaves)) return C2B_OK;  // the fold takes the entering state through run_scan's `init`
  }
  if constexpr (Mon::kCoop) {
    if (!scan_use_tree(K)) return C2B_OK;  // prefix blocks take the entering state through run_scan's `init`
  }
  const Levels lv = make_levels(leaves);
  const int ss    = Mon::ssize(J, K);
  double *dst     = states + lv.off[lv.nlev - 1] * ss;
  if (state_in)
    C2B_CUDA(cudaMemcpyAsync(dst, state_in, sizeof(double) * ss, cudaMemcpyDeviceToDevice, stream));
  else
    C2B_CUDA(cudaMemsetAsync(dst, 0, sizeof(double) * ss, stream));
  return C2B_OK;
}

// State entering segment `which`: the summaries of the segments that come earlier in processing order, applied one
// after the other to the zero state by one warp (at most a handful of segments: one per GPU).
template <class Mon>
__global__ void __launch_bounds__(32) seg_carry_kernel(int J_, int K, int nseg, int which, bool descending, const double *summaries,
                                                       double *out) {
  pdl_prologue();
  const int J = Mon::kJT ? Mon::kJT : J_;
  extern __shared__ __align__(16) double sc_sm[];
  const int es = Mon::esize(J, K), ss = Mon::ssize(J, K);
  double *e = sc_sm, *s0 = e + es, *s1 = s0 + ss, *scratch = s1 + ss;
  const int lane = threadIdx.x;
  for (int i = lane; i < ss; i += 32) s0[i] = 0.0;
  __syncwarp();
  const int step = descending ? -1 : 1;
  for (int s = descending ? nseg - 1 : 0; s != which; s += step) {
    wcopy(lane, es, summaries + (size_t)s * es, e);
    __syncwarp();
    Coop<Mon>::wapply(lane, J, K, e, s0, s1, scratch);
    double *tmp = s0;
    s0          = s1;
    s1          = tmp;
  }
  wcopy(lane, ss, s0, out);
}
template <class Mon>
static inline int run_seg_carry(int J, int K, int nseg, int which, bool descending, const double *summaries, double *out,
                                cudaStream_t stream) {
  if constexpr (Mon::kCoop) {
    const size_t smem = (size_t)(Mon::esize(J, K) + 2 * Mon::ssize(J, K) + Coop<Mon>::wsmem(J, K)) * sizeof(double);
    if (smem > 48 * 1024) C2B_CUDA(cudaFuncSetAttribute(seg_carry_kernel<Mon>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    C2B_LAUNCH("seg_carry_kernel", stream, (launch_k(seg_carry_kernel<Mon>, 1, 32, smem, stream, J, K, nseg, which, descending, summaries, out)));
    return C2B_OK;
  } else {
    set_error("segments of a longer series need a compiled width (J in 1..16)");
    return C2B_ERR_INVALID_ARGUMENT;
  }
}

// Finalise a reverse op: add the bt term each chunk owes its neighbour (one thread per chunk) and, in the first CTA
// of every series, reduce the per-chunk bc partials in a fixed order (deterministic).  grid = (ceil(C/1024), batch).
//   descending == true  (lower / factor_rev): chunk k owes bt[n0_k - 1]
//   descending == false (upper)             : chunk k owes bt[n1_k]
static __global__ void __launch_bounds__(1024) rev_finalize_kernel(int64_t N, int64_t C, int64_t L, int J, bool descending,
                                                                   const double *bc_part, const double *bt_edge, double *bt, double *bc,
                                                                   bool acc_bc) {
  pdl_prologue();
  const int64_t b = blockIdx.y;
  const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k < C) {
    const int64_t n0 = k * L, n1 = (n0 + L < N) ? n0 + L : N;
    const int64_t tgt = descending ? n0 - 1 : n1;
    if (tgt >= 0 && tgt < N) bt[b * N + tgt] += bt_edge[b * C + k];
  }
  if (blockIdx.x != 0) return;
  // bc[j] = sum over chunks, all J at once: thread (w, j) of warp-sized groups... every thread owns one j and a strided
  // set of chunks (consecutive threads read consecutive doubles of bc_part), then a fixed-shape tree per j
  // (deterministic).  One pass over the partials and two barriers, instead of J passes.
  __shared__ double sh[32 * 32];  // [slot][j], slot = thread / J within the first J * (1024 / J) threads
  const int per = (int)blockDim.x / J;  // threads per j
  const int j = threadIdx.x % J, slot = threadIdx.x / J;
  double s = 0.0;
  if (slot < per) {
    const double *p = bc_part + b * C * J;
    for (int64_t q = slot; q < C; q += per) s += p[q * J + j];
  }
  // reduce over slots: first within groups of 32 slots through shared memory
  const int nslot = per;
  for (int base = 0; base < nslot; base += 32) {
    __syncthreads();
    if (slot >= base && slot < base + 32 && slot < per) sh[(slot - base) * 32 + j] = s;
    __syncthreads();
    if (base + 32 < nslot || base > 0) {
      // more than 32 slots: fold this group's 32 values into the first slot's accumulator of the group
      if (slot == base && slot < per) {
        double v        = 0.0;
        const int count = nslot - base < 32 ? nslot - base : 32;
        for (int x = 0; x < count; ++x) v += sh[x * 32 + j];
        s = v;
      }
    }
  }
  __syncthreads();
  if (nslot <= 32) {
    if (slot == 0) {
      double v = 0.0;
      for (int x = 0; x < nslot; ++x) v += sh[x * 32 + j];
      bc[b * J + j] = acc_bc ? bc[b * J + j] + v : v;
    }
  } else {
    // group leaders (slot = 0, 32, 64, ...) hold the group sums: gather them through shared memory
    if (slot % 32 == 0 && slot < per) sh[(slot / 32) * 32 + j] = s;
    __syncthreads();
    if (slot == 0) {
      double v       = 0.0;
      const int ngrp = (nslot + 31) / 32;
      for (int x = 0; x < ngrp; ++x) v += sh[x * 32 + j];
      bc[b * J + j] = acc_bc ? bc[b * J + j] + v : v;
    }
  }
}

}  // namespace c2b
