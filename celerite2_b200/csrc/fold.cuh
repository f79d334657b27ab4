// Blocked sequential fold: the scan of the AFFINE summary algebras (AffV, DiagV, Cong) in three short launches.
//
// Why not the tree (scan.cuh): a tree spends one element-times-element COMBINE per level (J^3 work plus a J x K block,
// executed by one warp: 2 - 10 us each, r01i / r02a profiles) on 2 log2(C) dependent levels.  For an affine map
// s -> L(s) + b the scan only needs states, and pushing a state through an element (APPLY: a J x J by J x K product) is
// both much cheaper than composing two elements and trivially parallel over the state's entries -- a whole CTA finishes
// one step in the latency of a J-term dot product.  So:
//
//   A  one CTA per group of B ~ sqrt(C) consecutive chunks walks its group sequentially from the ZERO state:
//        s_loc[i+1] = apply(e_i, s_loc[i]),  H[i+1] = head(e_i) (x) H[i]      (H = product of the heads so far, H[0] = 1)
//      stores s_loc[i] and H[i] for every chunk and the group aggregate (H[B], s_loc[B]): "from zero" gives the b part
//      of the composition for free, the head product rides along on other threads.
//   B  one CTA per series walks the C / B group aggregates sequentially from the state entering the series:
//        G[g+1] = apply(agg_g, G[g])
//   C  every chunk, in parallel:  carry_i = s_loc[i] + lift(H[i], G[g])   (the part of the entering state that the
//      group's earlier chunks pass on)
//
// Depth: 2 sqrt(C) apply latencies (~0.1 us each for J = 6) instead of 26 combine latencies (2 - 5 us each); the arithmetic
// order is that of a sequential recursion.  The linear-fractional algebra of `factor` is not affine in this sense (its
// aggregate needs the full combine) and keeps the tree / prefix blocks.
#pragma once
#include "monoids.cuh"

namespace c2b {

constexpr int kFoldMaxThreads = 512;

// per-algebra pieces; every function is called by ALL threads of the CTA (barriers inside where a step has two stages)
template <class Mon>
struct Fold {
  static constexpr bool kAffine = false;
  static constexpr bool kSplit  = false;
};

// ---- AffV: s -> M s + b, head M (J x J row-major), state J x K (index j + J k) ------------------------------------------
template <int JT>
struct Fold<AffV<JT>> {
  static constexpr bool kAffine = true;
  static constexpr bool kSplit  = true;  // apply and head_step have no barriers inside: two halves of the CTA run them side by side
  __host__ __device__ static int hsize(int J) { return J * J; }
  __host__ __device__ static int scratch(int, int) { return 0; }
  C2B_DEV static void head_identity(int tid, int nt, int J, double *H) {
    for (int o = tid; o < J * J; o += nt) H[o] = (o / J == o % J) ? 1.0 : 0.0;
  }
  // Consecutive threads own consecutive ROWS i of a state column (the state is stored j + J k), so they read M down a
  // column: the kernels keep M TRANSPOSED in shared memory (element (i, j) at j * J + i), or the reads of a warp would sit
  // J doubles apart (16-way bank conflicts at J = 16: r02b, 5 us per step).  head_slot says where entry r of the head goes.
  C2B_DEV static int head_slot(int J, int r) { return (r % J) * J + r / J; }
  // (compiled widths: the J-term dot products are fully unrolled on two accumulators; a step of the walk is a dependent
  // instruction stream per thread, r02y: ~200 instructions at ~6 cycles each with the rolled loops)
  C2B_DEV static double dot_col(int J, const double *mT, int i, const double *x, double init) {
    if constexpr (JT > 0) {
      double s0 = init, s1 = 0.0;
#pragma unroll
      for (int j = 0; j + 1 < JT; j += 2) {
        s0 = fma(mT[j * JT + i], x[j], s0);
        s1 = fma(mT[(j + 1) * JT + i], x[j + 1], s1);
      }
      if (JT & 1) s0 = fma(mT[(JT - 1) * JT + i], x[JT - 1], s0);
      return s0 + s1;
    } else {
      double s = init;
#pragma unroll 4
      for (int j = 0; j < J; ++j) s = fma(mT[j * J + i], x[j], s);
      return s;
    }
  }
  // sout = M sin + b            (eT: transposed head, then b)
  C2B_DEV static void apply(int tid, int nt, int J, int K, const double *eT, const double *sin, double *sout, double *) {
    const double *b = eT + J * J;
    for (int o = tid; o < J * K; o += nt) {
      const int k = o / J, i = o - k * J;
      sout[o]     = dot_col(J, eT, i, sin + J * k, b[o]);
    }
  }
  // Hout = M Hin   (composition "Hin first, then this element"); H row-major
  C2B_DEV static void head_step(int tid, int nt, int J, const double *eT, const double *Hin, double *Hout) {
    for (int o = tid; o < J * J; o += nt) {
      const int i = o / J, j = o - i * J;
      double s    = 0.0;
      if constexpr (JT > 0) {
        double s1 = 0.0;
#pragma unroll
        for (int k = 0; k + 1 < JT; k += 2) {
          s  = fma(eT[k * JT + i], Hin[k * JT + j], s);
          s1 = fma(eT[(k + 1) * JT + i], Hin[(k + 1) * JT + j], s1);
        }
        if (JT & 1) s = fma(eT[(JT - 1) * JT + i], Hin[(JT - 1) * JT + j], s);
        s += s1;
      } else {
#pragma unroll 4
        for (int k = 0; k < J; ++k) s = fma(eT[k * J + i], Hin[k * J + j], s);
      }
      Hout[o] = s;
    }
  }
  // out = sloc + H G             (HT: transposed, see head_slot)
  C2B_DEV static void lift(int tid, int nt, int J, int K, const double *HT, const double *sloc, const double *G, double *out, double *) {
    for (int o = tid; o < J * K; o += nt) {
      const int k = o / J, i = o - k * J;
      out[o]      = dot_col(J, HT, i, G + J * k, sloc[o]);
    }
  }
};

// ---- DiagV: s -> m o s + b, head m (J), state J x K ------------------------------------------------------------------------
template <int JT>
struct Fold<DiagV<JT>> {
  static constexpr bool kAffine = true;
  static constexpr bool kSplit  = false;
  __host__ __device__ static int hsize(int J) { return J; }
  __host__ __device__ static int scratch(int, int) { return 0; }
  C2B_DEV static void head_identity(int tid, int nt, int J, double *H) {
    for (int o = tid; o < J; o += nt) H[o] = 1.0;
  }
  C2B_DEV static int head_slot(int, int r) { return r; }
  C2B_DEV static void apply(int tid, int nt, int J, int K, const double *e, const double *sin, double *sout, double *) {
    for (int o = tid; o < J * K; o += nt) sout[o] = fma(e[o % J], sin[o], e[J + o]);
  }
  C2B_DEV static void head_step(int tid, int nt, int J, const double *e, const double *Hin, double *Hout) {
    for (int o = tid; o < J; o += nt) Hout[o] = e[o] * Hin[o];
  }
  C2B_DEV static void lift(int tid, int nt, int J, int K, const double *H, const double *sloc, const double *G, double *out, double *) {
    for (int o = tid; o < J * K; o += nt) out[o] = fma(H[o % J], G[o], sloc[o]);
  }
};

// ---- Cong: G -> M^T G M + B, head M (J x J), state J x J symmetric ------------------------------------------------------------
template <int JT>
struct Fold<Cong<JT>> {
  static constexpr bool kAffine = true;
  static constexpr bool kSplit  = false;
  __host__ __device__ static int hsize(int J) { return J * J; }
  __host__ __device__ static int scratch(int J, int) { return 2 * J * J; }
  C2B_DEV static void head_identity(int tid, int nt, int J, double *H) {
    for (int o = tid; o < J * J; o += nt) H[o] = (o / J == o % J) ? 1.0 : 0.0;
  }
  C2B_DEV static int head_slot(int, int r) { return r; }  // threads run along rows of M: no conflicts
  // out = 0.5 (R + R^T) + add with R = M^T (G M); two stages through the scratch (T, R)
  C2B_DEV static void sandwich(int tid, int nt, int J, const double *M, const double *G, const double *add, double *out, double *sc) {
    double *T = sc, *R = sc + J * J;
    for (int o = tid; o < J * J; o += nt) {
      const int i = o / J, j = o - i * J;
      double s    = 0.0;
#pragma unroll 4
      for (int k = 0; k < J; ++k) s = fma(G[i * J + k], M[k * J + j], s);
      T[o] = s;
    }
    __syncthreads();
    for (int o = tid; o < J * J; o += nt) {
      const int i = o / J, j = o - i * J;
      double s    = 0.0;
#pragma unroll 4
      for (int k = 0; k < J; ++k) s = fma(M[k * J + i], T[k * J + j], s);
      R[o] = s;
    }
    __syncthreads();
    for (int o = tid; o < J * J; o += nt) {
      const int i = o / J, j = o - i * J;
      out[o]      = 0.5 * (R[i * J + j] + R[j * J + i]) + add[o];
    }
  }
  C2B_DEV static void apply(int tid, int nt, int J, int, const double *e, const double *sin, double *sout, double *sc) {
    sandwich(tid, nt, J, e, sin, e + J * J, sout, sc);
  }
  // Hout = Hin M   (Cong composes heads as "first times second")
  C2B_DEV static void head_step(int tid, int nt, int J, const double *e, const double *Hin, double *Hout) {
    for (int o = tid; o < J * J; o += nt) {
      const int i = o / J, j = o - i * J;
      double s    = 0.0;
#pragma unroll 4
      for (int k = 0; k < J; ++k) s = fma(Hin[i * J + k], e[k * J + j], s);
      Hout[o] = s;
    }
  }
  C2B_DEV static void lift(int tid, int nt, int J, int, const double *H, const double *sloc, const double *G, double *out, double *sc) {
    sandwich(tid, nt, J, H, G, sloc, out, sc);
  }
};

}  // namespace c2b


// =================================================================================================================
// kernels (this header is included by scan.cuh after slice_in / slice_out, Levels and kCoopThreadsMax)
// =================================================================================================================
namespace c2b {

constexpr int kFoldDepth = 8;  // elements in flight ahead of the walk (a step is shorter than one trip to L2 / HBM)

struct FoldPlan {
  int64_t C;      // chunks per series
  int B;          // chunks per group
  int64_t ng;     // groups per series
  int64_t stride; // per-series stride of nodes / states (elements), = Levels::total
  int kb;         // columns per slice
  int nslices;
  int threads;    // phase A
  int tthreads;   // phase B: one thread per state entry (r02zb: 64 / 128 threads with more entries each are slower)
  int fthreads;   // phase C: threads per chunk
  size_t smemA, smemC;
};

// layout inside one series: nodes[0, C) leaves | nodes[C, C + ng) group aggregates | nodes[C + ng] aggregate of the series;
// states[0, C) carry of every chunk | states[C, C + ng) state entering every group.
// heads: H[i] of every chunk, (batch, C, hsize), behind the node arrays of all series (host_common.cuh: take_scan).

// one local element <- global element, 8-byte cp.async per double; the algebra says where a head entry goes (head_slot)
template <class Mon>
C2B_DEV void fold_fetch(int J, double *dst, const double *src, int loc, int head, int hs, int goff) {
  for (int r = threadIdx.x; r < loc; r += blockDim.x)
    cp_async8(dst + (r < hs ? Fold<Mon>::head_slot(J, r) : r), src + (r < head ? r : r - head + goff));
}

// Ring of kFoldDepth element slots.  issue(i) starts the copy of element i of `n` into slot i mod depth (an empty group
// past the end keeps the group count uniform); ready() makes the oldest outstanding element visible to the CTA.
struct FoldRing {
  C2B_DEV static void ready() {
    cp_async_wait<kFoldDepth - 1>();
    __syncthreads();
  }
};

// phase A: grid.x = batch * ng, grid.y = slices
template <class Mon>
__global__ void __launch_bounds__(kFoldMaxThreads) fold_groups_kernel(int J_, int K, FoldPlan fp, double *nodes_all, double *states_all,
                                                                     double *heads_all) {
  pdl_prologue();
  using F = Fold<Mon>;
  const int J = Mon::kJT ? Mon::kJT : J_;
  extern __shared__ __align__(16) double fsm[];
  const int k0  = blockIdx.y * fp.kb;
  const int kl  = K - k0 < fp.kb ? K - k0 : fp.kb;
  const int esg = Mon::esize(J, K), ssg = Mon::ssize(J, K);
  const int es = Mon::esize(J, kl), ss = Mon::ssize(J, kl), hs = F::hsize(J);
  const int eh = Mon::ehead(J), sh = Mon::shead(J);
  const int ego = eh + Mon::ecol(J) * k0, sgo = sh + Mon::scol(J) * k0;
  const int tid = threadIdx.x, nt = blockDim.x;
  const int64_t b = blockIdx.x / fp.ng, g = blockIdx.x % fp.ng;
  const int64_t first = g * fp.B;
  const int cnt       = (int)(fp.C - first < fp.B ? fp.C - first : fp.B);
  double *ebuf = fsm;                        // kFoldDepth * es
  double *sbuf = ebuf + kFoldDepth * es;     // 2 * ss
  double *hbuf = sbuf + 2 * ss;              // 2 * hs
  double *sc   = hbuf + 2 * hs;              // F::scratch
  double *nodes = nodes_all + b * fp.stride * esg, *states = states_all + b * fp.stride * ssg;
  double *heads = heads_all + (b * fp.C + first) * (int64_t)hs;
  auto issue = [&](int i) {
    if (i < cnt) fold_fetch<Mon>(J, ebuf + (i % kFoldDepth) * es, nodes + (first + i) * esg, es, eh, hs, ego);
    cp_async_commit();
  };
  for (int i = 0; i < kFoldDepth; ++i) issue(i);
  for (int o = tid; o < ss; o += nt) sbuf[o] = 0.0;
  F::head_identity(tid, nt, J, hbuf);
  for (int i = 0; i < cnt; ++i) {
    FoldRing::ready();
    const double *e = ebuf + (i % kFoldDepth) * es;
    const double *s = sbuf + (i & 1) * ss, *H = hbuf + (i & 1) * hs;
    double *sn = sbuf + ((i + 1) & 1) * ss, *Hn = hbuf + ((i + 1) & 1) * hs;
    // what enters chunk i when the group starts from zero, and the head product of the chunks before it
    double *sl = states + (first + i) * ssg;
    for (int o = tid; o < ss; o += nt) sl[o < sh ? o : o - sh + sgo] = s[o];
    if (blockIdx.y == 0)
      for (int o = tid; o < hs; o += nt) heads[(int64_t)i * hs + o] = H[o];
    if (F::kSplit && nt >= 64) {
      // the head product on the first half of the CTA, the state on the second (whole warps each)
      const int half = (nt / 64) * 32;
      if (tid < half)
        F::head_step(tid, half, J, e, H, Hn);
      else
        F::apply(tid - half, nt - half, J, kl, e, s, sn, sc);
    } else {
      F::head_step(tid, nt, J, e, H, Hn);
      F::apply(tid, nt, J, kl, e, s, sn, sc);
    }
    __syncthreads();  // the slot is free, the new state and head are complete
    issue(i + kFoldDepth);
  }
  // the group's aggregate: element = [head product | state reached from zero]
  const double *s = sbuf + (cnt & 1) * ss, *H = hbuf + (cnt & 1) * hs;
  double *agg = nodes + (fp.C + g) * esg;
  if (blockIdx.y == 0) {
    for (int o = tid; o < hs; o += nt) agg[o] = H[o];
    for (int o = tid; o < sh; o += nt) agg[hs + o] = s[o];  // state head (Cong: B)
  }
  for (int o = tid + sh; o < ss; o += nt) agg[ego + (o - sh)] = s[o];
}

// phase B: grid.x = batch, grid.y = slices.  States entering the groups from `init` (NULL = zero) when want_states;
// the aggregate of the whole series (from zero, with the head product) when want_root.
template <class Mon>
__global__ void __launch_bounds__(kFoldMaxThreads) fold_top_kernel(int J_, int K, FoldPlan fp, bool want_states, bool want_root,
                                                                  double *nodes_all, double *states_all, const double *init) {
  pdl_prologue();
  using F = Fold<Mon>;
  const int J = Mon::kJT ? Mon::kJT : J_;
  extern __shared__ __align__(16) double fsm[];
  const int k0  = blockIdx.y * fp.kb;
  const int kl  = K - k0 < fp.kb ? K - k0 : fp.kb;
  const int esg = Mon::esize(J, K), ssg = Mon::ssize(J, K);
  const int es = Mon::esize(J, kl), ss = Mon::ssize(J, kl), hs = F::hsize(J);
  const int eh = Mon::ehead(J), sh = Mon::shead(J);
  const int ego = eh + Mon::ecol(J) * k0, sgo = sh + Mon::scol(J) * k0;
  const int tid = threadIdx.x, nt = blockDim.x;
  const int64_t b = blockIdx.x;
  double *ebuf = fsm;                        // kFoldDepth * es
  double *sbuf = ebuf + kFoldDepth * es;     // 2 * ss: chain from init
  double *zbuf = sbuf + 2 * ss;              // 2 * ss: chain from zero (root)
  double *hbuf = zbuf + 2 * ss;              // 2 * hs
  double *sc   = hbuf + 2 * hs;
  double *nodes = nodes_all + b * fp.stride * esg, *states = states_all + b * fp.stride * ssg;
  const int n = (int)fp.ng;
  auto issue = [&](int g) {
    if (g < n) fold_fetch<Mon>(J, ebuf + (g % kFoldDepth) * es, nodes + (fp.C + g) * esg, es, eh, hs, ego);
    cp_async_commit();
  };
  for (int g = 0; g < kFoldDepth; ++g) issue(g);
  for (int o = tid; o < ss; o += nt) {
    sbuf[o] = init ? init[b * ssg + (o < sh ? o : o - sh + sgo)] : 0.0;
    zbuf[o] = 0.0;
  }
  F::head_identity(tid, nt, J, hbuf);
  for (int g = 0; g < n; ++g) {
    FoldRing::ready();
    const double *e = ebuf + (g % kFoldDepth) * es;
    const int cur = g & 1, nxt = cur ^ 1;
    if (want_states) {
      double *st = states + (fp.C + g) * ssg;
      for (int o = tid; o < ss; o += nt) st[o < sh ? o : o - sh + sgo] = sbuf[cur * ss + o];
      F::apply(tid, nt, J, kl, e, sbuf + cur * ss, sbuf + nxt * ss, sc);
    }
    if (want_root) {
      F::head_step(tid, nt, J, e, hbuf + cur * hs, hbuf + nxt * hs);
      __syncthreads();  // (the scratch of apply is reused)
      F::apply(tid, nt, J, kl, e, zbuf + cur * ss, zbuf + nxt * ss, sc);
    }
    __syncthreads();
    issue(g + kFoldDepth);
  }
  if (want_root) {
    const double *s = zbuf + (n & 1) * ss, *H = hbuf + (n & 1) * hs;
    double *root = nodes + (fp.C + fp.ng) * esg;
    if (blockIdx.y == 0) {
      for (int o = tid; o < hs; o += nt) root[o] = H[o];
      for (int o = tid; o < sh; o += nt) root[hs + o] = s[o];
    }
    for (int o = tid + sh; o < ss; o += nt) root[ego + (o - sh)] = s[o];
  }
}

// phase C: carry_i = s_loc[i] + lift(H[i], G[group of i]), in place.  A CTA of 256 threads handles 256 / fthreads chunks
// (grid.x = batch * ceil(C / chunks per CTA), grid.y = slices); every sub-group runs the same sequence, so the CTA-wide
// barriers inside `lift` line up.
template <class Mon>
__global__ void __launch_bounds__(256) fold_finish_kernel(int J_, int K, FoldPlan fp, double *states_all, const double *heads_all) {
  pdl_prologue();
  using F = Fold<Mon>;
  const int J = Mon::kJT ? Mon::kJT : J_;
  extern __shared__ __align__(16) double fsm[];
  const int k0  = blockIdx.y * fp.kb;
  const int kl  = K - k0 < fp.kb ? K - k0 : fp.kb;
  const int ssg = Mon::ssize(J, K);
  const int ss = Mon::ssize(J, kl), hs = F::hsize(J);
  const int sh = Mon::shead(J);
  const int sgo = sh + Mon::scol(J) * k0;
  const int nt  = fp.fthreads, cpb = blockDim.x / nt;
  const int sub = threadIdx.x / nt, tid = threadIdx.x - sub * nt;
  const int64_t per = (fp.C + cpb - 1) / cpb;  // CTAs per series
  const int64_t b = blockIdx.x / per;
  const int64_t i = (blockIdx.x % per) * cpb + sub;
  const bool have = i < fp.C;
  const int64_t ic = have ? i : fp.C - 1, g = ic / fp.B;
  const int per_sub = hs + 3 * ss + F::scratch(J, kl);
  double *H   = fsm + (size_t)sub * per_sub;  // hs
  double *sl  = H + hs;                       // ss
  double *G   = sl + ss;                      // ss
  double *out = G + ss;                       // ss
  double *sc  = out + ss;
  double *states = states_all + b * fp.stride * ssg;
  const double *hg = heads_all + (b * fp.C + ic) * (int64_t)hs;
  for (int o = tid; o < hs; o += nt) H[F::head_slot(J, o)] = hg[o];
  for (int o = tid; o < ss; o += nt) {
    const int go = o < sh ? o : o - sh + sgo;
    sl[o]        = states[ic * ssg + go];
    G[o]         = states[(fp.C + g) * ssg + go];
  }
  __syncthreads();
  F::lift(tid, nt, J, kl, H, sl, G, out, sc);
  __syncthreads();
  if (have)
    for (int o = tid; o < ss; o += nt) states[i * ssg + (o < sh ? o : o - sh + sgo)] = out[o];
}

template <class Mon>
static inline FoldPlan fold_plan(int J, int K, int64_t leaves, int64_t stride) {
  using F = Fold<Mon>;
  FoldPlan fp;
  fp.C = leaves;
  int B = 1;
  while ((int64_t)B * B < leaves) ++B;  // ceil(sqrt(C)): the two sequential walks are equally long
  if (B > 256) B = 256;
  fp.B      = B;
  fp.ng     = (leaves + B - 1) / B;
  fp.stride = stride;
  const int hs = F::hsize(J);
  auto doubles = [&](int kb) {
    return (size_t)kFoldDepth * Mon::esize(J, kb) + (size_t)4 * Mon::ssize(J, kb) + 2 * hs + F::scratch(J, kb);
  };
  int kb = K;
  while (kb > 1 && doubles(kb) * sizeof(double) > 160 * 1024) kb = (kb + 1) / 2;
  // Slices of about J columns (at least 8): the walks are latency, not work -- narrower slices walk on more SMs at once,
  // each with a shorter step, and the head product every slice repeats costs no wall time (r02y, c4: J = 16, K = 64:
  // one slice 0.267 ms of fold kernels per apply_inverse, slices of 32 / 16 / 8 / 4 columns 0.208 / 0.185 / 0.192 / 0.253).
  // C2B_FOLD_KB overrides.
  static const int env_kb = getenv("C2B_FOLD_KB") ? atoi(getenv("C2B_FOLD_KB")) : 0;
  const int want_kb       = env_kb > 0 ? env_kb : (J > 8 ? J : 8);
  if (want_kb < kb) kb = want_kb;
  const int nsl = (K + kb - 1) / kb;
  kb            = (K + nsl - 1) / nsl;
  fp.kb      = kb;
  fp.nslices = nsl;
  const int work = Mon::ssize(J, kb) > hs ? Mon::ssize(J, kb) : hs;
  int th         = (work + 31) / 32 * 32;
  if (th < 64) th = 64;
  fp.tthreads = th > kFoldMaxThreads ? kFoldMaxThreads : th;
  if (F::kSplit) th *= 2;  // (fold_groups_kernel: head product and state on separate halves)
  if (th > kFoldMaxThreads) th = kFoldMaxThreads;
  fp.threads = th;
  int ft     = 32;
  while (ft < Mon::ssize(J, kb) && ft < 256) ft *= 2;
  fp.fthreads = ft;
  fp.smemA    = doubles(kb) * sizeof(double);
  fp.smemC    = (size_t)(256 / ft) * ((size_t)hs + 3 * Mon::ssize(J, kb) + F::scratch(J, kb)) * sizeof(double);
  return fp;
}

// mode 0: all three phases.  mode 1: A + the aggregate of the whole series (nodes[C + ng]).  mode 2: B from `init` + C
// (A ran in an earlier mode-1 call on the same scratch).  `heads`: (batch, C, hsize) doubles.
template <class Mon>
static inline int run_fold(int J, int K, int64_t batch, int64_t leaves, int64_t stride, double *nodes, double *states, double *heads,
                           const double *init, cudaStream_t stream, int mode) {
  const FoldPlan fp = fold_plan<Mon>(J, K, leaves, stride);
  if (fp.smemA > 48 * 1024) {
    C2B_CUDA(cudaFuncSetAttribute(fold_groups_kernel<Mon>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fp.smemA));
    C2B_CUDA(cudaFuncSetAttribute(fold_top_kernel<Mon>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fp.smemA));
  }
  if (fp.smemC > 48 * 1024)
    C2B_CUDA(cudaFuncSetAttribute(fold_finish_kernel<Mon>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fp.smemC));
  const unsigned ns = (unsigned)fp.nslices;
  if (mode != 2)
    C2B_LAUNCH("fold_groups_kernel", stream,
               (launch_k(fold_groups_kernel<Mon>, dim3((unsigned)(batch * fp.ng), ns), fp.threads, fp.smemA, stream, J, K, fp, nodes, states, heads)));
  C2B_LAUNCH("fold_top_kernel", stream,
             (launch_k(fold_top_kernel<Mon>, dim3((unsigned)batch, ns), fp.tthreads, fp.smemA, stream, J, K, fp, mode != 1, mode == 1, nodes, states, init)));
  if (mode != 1) {
    const int cpb     = 256 / fp.fthreads;
    const int64_t per = (fp.C + cpb - 1) / cpb;
    C2B_LAUNCH("fold_finish_kernel", stream,
               (launch_k(fold_finish_kernel<Mon>, dim3((unsigned)(batch * per), ns), 256, fp.smemC, stream, J, K, fp, states, heads)));
  }
  return C2B_OK;
}

}  // namespace c2b
