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
};

// ---- AffV: s -> M s + b, head M (J x J row-major), state J x K (index j + J k) ------------------------------------------
template <int JT>
struct Fold<AffV<JT>> {
  static constexpr bool kAffine = true;
  __host__ __device__ static int hsize(int J) { return J * J; }
  __host__ __device__ static int scratch(int, int) { return 0; }
  C2B_DEV static void head_identity(int tid, int nt, int J, double *H) {
    for (int o = tid; o < J * J; o += nt) H[o] = (o / J == o % J) ? 1.0 : 0.0;
  }
  // sout = M sin + b
  C2B_DEV static void apply(int tid, int nt, int J, int K, const double *e, const double *sin, double *sout, double *) {
    const double *b = e + J * J;
    for (int o = tid; o < J * K; o += nt) {
      const int k = o / J, i = o - k * J;
      double s    = b[o];
#pragma unroll 4
      for (int j = 0; j < J; ++j) s = fma(e[i * J + j], sin[j + J * k], s);
      sout[o] = s;
    }
  }
  // Hout = M Hin   (composition "Hin first, then this element")
  C2B_DEV static void head_step(int tid, int nt, int J, const double *e, const double *Hin, double *Hout) {
    for (int o = tid; o < J * J; o += nt) {
      const int i = o / J, j = o - i * J;
      double s    = 0.0;
#pragma unroll 4
      for (int k = 0; k < J; ++k) s = fma(e[i * J + k], Hin[k * J + j], s);
      Hout[o] = s;
    }
  }
  // out = sloc + H G
  C2B_DEV static void lift(int tid, int nt, int J, int K, const double *H, const double *sloc, const double *G, double *out, double *) {
    for (int o = tid; o < J * K; o += nt) {
      const int k = o / J, i = o - k * J;
      double s    = sloc[o];
#pragma unroll 4
      for (int j = 0; j < J; ++j) s = fma(H[i * J + j], G[j + J * k], s);
      out[o] = s;
    }
  }
};

// ---- DiagV: s -> m o s + b, head m (J), state J x K ------------------------------------------------------------------------
template <int JT>
struct Fold<DiagV<JT>> {
  static constexpr bool kAffine = true;
  __host__ __device__ static int hsize(int J) { return J; }
  __host__ __device__ static int scratch(int, int) { return 0; }
  C2B_DEV static void head_identity(int tid, int nt, int J, double *H) {
    for (int o = tid; o < J; o += nt) H[o] = 1.0;
  }
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
  __host__ __device__ static int hsize(int J) { return J * J; }
  __host__ __device__ static int scratch(int J, int) { return 2 * J * J; }
  C2B_DEV static void head_identity(int tid, int nt, int J, double *H) {
    for (int o = tid; o < J * J; o += nt) H[o] = (o / J == o % J) ? 1.0 : 0.0;
  }
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

struct FoldPlan {
  int64_t C;      // chunks per series
  int B;          // chunks per group
  int64_t ng;     // groups per series
  int64_t stride; // per-series stride of nodes / states (elements), = Levels::total
  int kb;         // columns per slice
  int nslices;
  int threads;
  size_t smemA, smemC;
};

// layout inside one series: nodes[0, C) leaves | nodes[C, C + ng) group aggregates | nodes[C + ng] aggregate of the series;
// states[0, C) carry of every chunk | states[C, C + ng) state entering every group.
// heads: H[i] of every chunk, (batch, C, hsize), behind the node arrays of all series (host_common.cuh: take_scan).

C2B_DEV void fold_fetch(double *dst, const double *src, int loc, int head, int goff) {
  // one local element <- global element, 8-byte cp.async per double (the next step's element lands while this one is used)
  for (int r = threadIdx.x; r < loc; r += blockDim.x) cp_async8(dst + r, src + (r < head ? r : r - head + goff));
}

// phase A: grid.x = batch * ng, grid.y = slices
template <class Mon>
__global__ void __launch_bounds__(kFoldMaxThreads) fold_groups_kernel(int J_, int K, FoldPlan fp, double *nodes_all, double *states_all,
                                                                     double *heads_all) {
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
  double *ebuf = fsm;                  // 2 * es
  double *sbuf = ebuf + 2 * es;        // 2 * ss
  double *hbuf = sbuf + 2 * ss;        // 2 * hs
  double *sc   = hbuf + 2 * hs;        // F::scratch
  double *nodes = nodes_all + b * fp.stride * esg, *states = states_all + b * fp.stride * ssg;
  double *heads = heads_all + (b * fp.C + first) * (int64_t)hs;
  for (int o = tid; o < ss; o += nt) sbuf[o] = 0.0;
  F::head_identity(tid, nt, J, hbuf);
  fold_fetch(ebuf, nodes + first * esg, es, eh, ego);
  cp_async_commit();
  for (int i = 0; i < cnt; ++i) {
    if (i + 1 < cnt) fold_fetch(ebuf + ((i + 1) & 1) * es, nodes + (first + i + 1) * esg, es, eh, ego);
    cp_async_commit();
    cp_async_wait<1>();
    __syncthreads();
    const double *e = ebuf + (i & 1) * es;
    const double *s = sbuf + (i & 1) * ss, *H = hbuf + (i & 1) * hs;
    double *sn = sbuf + ((i + 1) & 1) * ss, *Hn = hbuf + ((i + 1) & 1) * hs;
    // what enters chunk i when the group starts from zero, and the head product of the chunks before it
    double *sl = states + (first + i) * ssg;
    for (int o = tid; o < ss; o += nt) sl[o < sh ? o : o - sh + sgo] = s[o];
    if (blockIdx.y == 0)
      for (int o = tid; o < hs; o += nt) heads[(int64_t)i * hs + o] = H[o];
    F::head_step(tid, nt, J, e, H, Hn);
    F::apply(tid, nt, J, kl, e, s, sn, sc);
    __syncthreads();
  }
  // the group's aggregate: element = [head product | state reached from zero]
  const double *s = sbuf + (cnt & 1) * ss, *H = hbuf + (cnt & 1) * hs;
  double *agg = nodes + (fp.C + g) * esg;
  if (blockIdx.y == 0) {
    for (int o = tid; o < hs; o += nt) agg[o] = H[o];
    for (int o = tid; o < sh; o += nt) agg[hs + o] = s[o];  // state head (Cong: B)
  }
  for (int o = tid + sh; o < ss; o += nt) agg[eh + (o - sh) + (ego - eh)] = s[o];
}

// phase B: grid.x = batch, grid.y = slices.  States entering the groups from `init` (NULL = zero) when want_states;
// the aggregate of the whole series (from zero, with the head product) when want_root.
template <class Mon>
__global__ void __launch_bounds__(kFoldMaxThreads) fold_top_kernel(int J_, int K, FoldPlan fp, bool want_states, bool want_root,
                                                                  double *nodes_all, double *states_all, const double *init) {
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
  double *ebuf = fsm;                  // 2 * es
  double *sbuf = ebuf + 2 * es;        // 2 * ss: chain from init
  double *zbuf = sbuf + 2 * ss;        // 2 * ss: chain from zero (root)
  double *hbuf = zbuf + 2 * ss;        // 2 * hs
  double *sc   = hbuf + 2 * hs;
  double *nodes = nodes_all + b * fp.stride * esg, *states = states_all + b * fp.stride * ssg;
  for (int o = tid; o < ss; o += nt) {
    sbuf[o] = init ? init[b * ssg + (o < sh ? o : o - sh + sgo)] : 0.0;
    zbuf[o] = 0.0;
  }
  F::head_identity(tid, nt, J, hbuf);
  fold_fetch(ebuf, nodes + fp.C * esg, es, eh, ego);
  cp_async_commit();
  const int n = (int)fp.ng;
  for (int g = 0; g < n; ++g) {
    if (g + 1 < n) fold_fetch(ebuf + ((g + 1) & 1) * es, nodes + (fp.C + g + 1) * esg, es, eh, ego);
    cp_async_commit();
    cp_async_wait<1>();
    __syncthreads();
    const double *e = ebuf + (g & 1) * es;
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
  }
  if (want_root) {
    const double *s = zbuf + (n & 1) * ss, *H = hbuf + (n & 1) * hs;
    double *root = nodes + (fp.C + fp.ng) * esg;
    if (blockIdx.y == 0) {
      for (int o = tid; o < hs; o += nt) root[o] = H[o];
      for (int o = tid; o < sh; o += nt) root[hs + o] = s[o];
    }
    for (int o = tid + sh; o < ss; o += nt) root[eh + (o - sh) + (ego - eh)] = s[o];
  }
}

// phase C: one CTA per chunk (grid.x = batch * C, grid.y = slices): carry_i = s_loc[i] + lift(H[i], G[group of i]), in place
template <class Mon>
__global__ void __launch_bounds__(kFoldMaxThreads) fold_finish_kernel(int J_, int K, FoldPlan fp, double *states_all,
                                                                     const double *heads_all) {
  using F = Fold<Mon>;
  const int J = Mon::kJT ? Mon::kJT : J_;
  extern __shared__ __align__(16) double fsm[];
  const int k0  = blockIdx.y * fp.kb;
  const int kl  = K - k0 < fp.kb ? K - k0 : fp.kb;
  const int ssg = Mon::ssize(J, K);
  const int ss = Mon::ssize(J, kl), hs = F::hsize(J);
  const int sh = Mon::shead(J);
  const int sgo = sh + Mon::scol(J) * k0;
  const int tid = threadIdx.x, nt = blockDim.x;
  const int64_t b = blockIdx.x / fp.C, i = blockIdx.x % fp.C, g = i / fp.B;
  double *H  = fsm;           // hs
  double *sl = H + hs;        // ss
  double *G  = sl + ss;       // ss
  double *out = G + ss;       // ss
  double *sc  = out + ss;
  double *states = states_all + b * fp.stride * ssg;
  const double *hg = heads_all + (b * fp.C + i) * (int64_t)hs;
  for (int o = tid; o < hs; o += nt) H[o] = hg[o];
  for (int o = tid; o < ss; o += nt) {
    const int go = o < sh ? o : o - sh + sgo;
    sl[o]        = states[i * ssg + go];
    G[o]         = states[(fp.C + g) * ssg + go];
  }
  __syncthreads();
  F::lift(tid, nt, J, kl, H, sl, G, out, sc);
  __syncthreads();
  for (int o = tid; o < ss; o += nt) states[i * ssg + (o < sh ? o : o - sh + sgo)] = out[o];
}

}  // namespace c2b

namespace c2b {

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
  auto doubles = [&](int kb, bool top) {
    return (size_t)2 * Mon::esize(J, kb) + (size_t)(top ? 4 : 2) * Mon::ssize(J, kb) + 2 * hs + F::scratch(J, kb);
  };
  int kb = K;
  while (kb > 1 && doubles(kb, true) * sizeof(double) > 96 * 1024) kb = (kb + 1) / 2;
  const int nsl = (K + kb - 1) / kb;
  kb            = (K + nsl - 1) / nsl;
  fp.kb      = kb;
  fp.nslices = nsl;
  int work   = Mon::ssize(J, kb) > hs ? Mon::ssize(J, kb) : hs;
  int th     = (work + 31) / 32 * 32;
  if (th < 64) th = 64;
  if (th > kFoldMaxThreads) th = kFoldMaxThreads;
  fp.threads = th;
  fp.smemA   = doubles(kb, true) * sizeof(double);
  fp.smemC   = ((size_t)hs + 3 * Mon::ssize(J, kb) + F::scratch(J, kb)) * sizeof(double);
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
               (fold_groups_kernel<Mon><<<dim3((unsigned)(batch * fp.ng), ns), fp.threads, fp.smemA, stream>>>(J, K, fp, nodes, states, heads)));
  C2B_LAUNCH("fold_top_kernel", stream,
             (fold_top_kernel<Mon><<<dim3((unsigned)batch, ns), fp.threads, fp.smemA, stream>>>(J, K, fp, mode != 1, mode == 1, nodes, states, init)));
  if (mode != 1) {
    int th = (Mon::ssize(J, fp.kb) + 31) / 32 * 32;
    if (th < 32) th = 32;
    if (th > 256) th = 256;
    C2B_LAUNCH("fold_finish_kernel", stream,
               (fold_finish_kernel<Mon><<<dim3((unsigned)(batch * fp.C), ns), th, fp.smemC, stream>>>(J, K, fp, states, heads)));
  }
  return C2B_OK;
}

}  // namespace c2b
