// C ABI: factor and factor_rev.
#include "host_common.cuh"

namespace c2b {

static void sizes_factor(int64_t N, int J, int64_t batch, Plan *pl, Arena &ar, ScanBuffers *sb) {
  static const int64_t units = env_units("C2B_UNITS_FACTOR");
  *pl = make_plan(N, batch, 1, J, C2B_MINB_FACTOR, units);
  // a single series gets its (one-node) tree even un-chunked: it may be run as a segment of a longer series
  if (pl->C > 1 || batch == 1) take_scan(ar, batch, pl->C, 3 * J * J, J * J, sb);
}
size_t scratch_factor(int64_t N, int J, int64_t batch) {
  Plan pl;
  Arena ar = sizing_arena();
  ScanBuffers sb;
  sizes_factor(N, J, batch, &pl, ar, &sb);
  return ar.used;
}

static void sizes_factor_rev(int64_t N, int J, int64_t batch, Plan *pl, Arena &ar, ScanBuffers *sb, double **bc_part, double **bt_edge) {
  static const int64_t units = env_units("C2B_UNITS_FREV");
  *pl = make_plan(N, batch, 1, J, C2B_MINB_FREV, units);
  if (pl->C > 1 || batch == 1) take_scan(ar, batch, pl->C, 2 * J * J, J * J, sb);
  *bc_part = ar.take<double>((size_t)batch * pl->C * J);
  *bt_edge = ar.take<double>((size_t)batch * pl->C);
}
size_t scratch_factor_rev(int64_t N, int J, int64_t batch) {
  Plan pl;
  Arena ar = sizing_arena();
  ScanBuffers sb;
  double *a, *b;
  sizes_factor_rev(N, J, batch, &pl, ar, &sb, &a, &b);
  return ar.used;
}



}  // namespace c2b

namespace c2b {

// Common checks of the *_segment entry points; fills the kernel-side descriptor and the time-stamp index bounds.
int segment_setup(const c2b_segment_t *seg, int64_t N, int64_t batch, const void *scratch, bool reverse_op, Seg *out, int64_t *tmin,
                  int64_t *tmax) {
  memset(out, 0, sizeof(*out));
  *tmin = 0;
  *tmax = N - 1;
  if (!seg) return C2B_OK;
  C2B_REQUIRE(seg->phase == C2B_SEG_SUMMARISE || seg->phase == C2B_SEG_APPLY, "segment: phase must be C2B_SEG_SUMMARISE or C2B_SEG_APPLY");
  C2B_REQUIRE(batch == 1, "segment: batch must be 1");
  C2B_REQUIRE(scratch != nullptr, "segment: the caller must provide the scratch buffer (it carries the tree between the two phases)");
  C2B_REQUIRE(seg->phase != C2B_SEG_SUMMARISE || seg->summary_out, "segment: summary_out is NULL");
  (void)reverse_op;
  out->phase       = seg->phase;
  out->state_in    = seg->state_in;
  out->summary_out = seg->summary_out;
  out->edge_out    = seg->edge_out;
  if (seg->has_left) *tmin = -1;
  if (seg->has_right) *tmax = N;
  return C2B_OK;
}

static int factor_impl(int64_t N, int32_t J, int64_t batch, const double *t, const double *c, const double *a, const double *U,
                       const double *V, double *d, double *W, double *S, int64_t *status, int32_t flags, void *scratch,
                       size_t scratch_bytes, void *stream_, const c2b_segment_t *seg) {
  int rc = check_common(N, J, batch);
  if (rc) return rc;
  C2B_REQUIRE(t && c && a && U && V && d && W, "null data pointer");
  Seg sg;
  int64_t tmin, tmax;
  rc = segment_setup(seg, N, batch, scratch, false, &sg, &tmin, &tmax);
  if (rc) return rc;
  cudaStream_t stream = (cudaStream_t)stream_;
  Plan pl;
  ScanBuffers sb;
  Arena sz = sizing_arena();
  sizes_factor(N, J, batch, &pl, sz, &sb);
  Scratch sc;
  rc = sc.init(scratch, scratch_bytes, sz.used, stream);
  if (rc) return rc;
  Arena ar(sc.ptr, sc.bytes);
  sizes_factor(N, J, batch, &pl, ar, &sb);
  if (status && sg.phase != 1) C2B_CUDA(cudaMemsetAsync(status, 0, sizeof(int64_t) * (size_t)batch, stream));
  FactorArgs g;
  memset(&g, 0, sizeof(g));
  g.N = N; g.batch = batch; g.C = pl.C; g.L = pl.L; g.J = J; g.tmin = tmin; g.tmax = tmax; g.seg = sg;
  g.t = t; g.c = c; g.a = a; g.U = U; g.V = V; g.d = d; g.W = W; g.S = S; g.status = status;
  g.t_bs = (flags & C2B_SHARED_T) ? 0 : N;
  g.c_bs = (flags & C2B_SHARED_C) ? 0 : J;
#define CALL(JT) rc = launch_factor<JT>(g, sb, stream)
  C2B_DISPATCH_J_FACTOR(J, CALL)
#undef CALL
  return rc;
}

static int factor_rev_impl(int64_t N, int32_t J, int64_t batch, const double *t, const double *c, const double *a, const double *U,
                           const double *V, const double *d, const double *W, const double *S, const double *bd, const double *bW,
                           double *bt, double *bc, double *ba, double *bU, double *bV, int32_t flags, void *scratch,
                           size_t scratch_bytes, void *stream_, const c2b_segment_t *seg) {
  (void)a;
  (void)V;
  int rc = check_common(N, J, batch);
  if (rc) return rc;
  C2B_REQUIRE(t && c && U && d && W && S && bd && bW && bt && bc && ba && bU && bV, "null data pointer");
  Seg sg;
  int64_t tmin, tmax;
  rc = segment_setup(seg, N, batch, scratch, true, &sg, &tmin, &tmax);
  if (rc) return rc;
  cudaStream_t stream = (cudaStream_t)stream_;
  Plan pl;
  ScanBuffers sb;
  double *bc_part, *bt_edge;
  Arena sz = sizing_arena();
  sizes_factor_rev(N, J, batch, &pl, sz, &sb, &bc_part, &bt_edge);
  Scratch sc;
  rc = sc.init(scratch, scratch_bytes, sz.used, stream);
  if (rc) return rc;
  Arena ar(sc.ptr, sc.bytes);
  sizes_factor_rev(N, J, batch, &pl, ar, &sb, &bc_part, &bt_edge);
  FactorRevArgs g;
  memset(&g, 0, sizeof(g));
  g.N = N; g.batch = batch; g.C = pl.C; g.L = pl.L; g.J = J; g.tmin = tmin; g.tmax = tmax; g.seg = sg;
  g.t = t; g.c = c; g.U = U; g.d = d; g.W = W; g.S = S; g.bd = bd; g.bW = bW;
  g.bt = bt; g.ba = ba; g.bU = bU; g.bV = bV;
  g.t_bs = (flags & C2B_SHARED_T) ? 0 : N;
  g.c_bs = (flags & C2B_SHARED_C) ? 0 : J;
  g.bc_part = bc_part;
  g.bt_edge = bt_edge;
  g.acc     = (flags & C2B_ACCUMULATE) ? 1 : 0;
#define CALL(JT) rc = launch_factor_rev<JT>(g, sb, bc, stream)
  C2B_DISPATCH_J_FACTOR(J, CALL)
#undef CALL
  return rc;
}

}  // namespace c2b

extern "C" {

int c2b_factor(int64_t N, int32_t J, int64_t batch, const double *t, const double *c, const double *a, const double *U, const double *V,
               double *d, double *W, double *S, int64_t *status, int32_t flags, void *scratch, size_t scratch_bytes, void *stream) {
  c2b::OpScope _op("c2b_factor");
  return c2b::factor_impl(N, J, batch, t, c, a, U, V, d, W, S, status, flags, scratch, scratch_bytes, stream, nullptr);
}
int c2b_factor_segment(int64_t N, int32_t J, int64_t batch, const double *t, const double *c, const double *a, const double *U,
                       const double *V, double *d, double *W, double *S, int64_t *status, int32_t flags, void *scratch,
                       size_t scratch_bytes, void *stream, const c2b_segment_t *seg) {
  c2b::OpScope _op("c2b_factor");
  if (!seg) {
    c2b::set_error("c2b_factor_segment: seg is NULL");
    return C2B_ERR_INVALID_ARGUMENT;
  }
  return c2b::factor_impl(N, J, batch, t, c, a, U, V, d, W, S, status, flags, scratch, scratch_bytes, stream, seg);
}
int c2b_factor_rev(int64_t N, int32_t J, int64_t batch, const double *t, const double *c, const double *a, const double *U,
                   const double *V, const double *d, const double *W, const double *S, const double *bd, const double *bW, double *bt,
                   double *bc, double *ba, double *bU, double *bV, int32_t flags, void *scratch, size_t scratch_bytes, void *stream) {
  c2b::OpScope _op("c2b_factor_rev");
  return c2b::factor_rev_impl(N, J, batch, t, c, a, U, V, d, W, S, bd, bW, bt, bc, ba, bU, bV, flags, scratch, scratch_bytes, stream,
                              nullptr);
}
int c2b_factor_rev_segment(int64_t N, int32_t J, int64_t batch, const double *t, const double *c, const double *a, const double *U,
                           const double *V, const double *d, const double *W, const double *S, const double *bd, const double *bW,
                           double *bt, double *bc, double *ba, double *bU, double *bV, int32_t flags, void *scratch,
                           size_t scratch_bytes, void *stream, const c2b_segment_t *seg) {
  c2b::OpScope _op("c2b_factor_rev");
  if (!seg) {
    c2b::set_error("c2b_factor_rev_segment: seg is NULL");
    return C2B_ERR_INVALID_ARGUMENT;
  }
  return c2b::factor_rev_impl(N, J, batch, t, c, a, U, V, d, W, S, bd, bW, bt, bc, ba, bU, bV, flags, scratch, scratch_bytes, stream, seg);
}

int c2b_segment_sizes(int32_t op, int32_t J, int32_t K, int64_t *summary_doubles, int64_t *state_doubles) {
  using namespace c2b;
  int64_t e = 0, s = 0;
  switch (op) {
    case C2B_OP_FACTOR: e = 3 * J * J; s = J * J; break;
    case C2B_OP_FACTOR_REV: e = 2 * J * J; s = J * J; break;
    case C2B_OP_SOLVE_LOWER:
    case C2B_OP_SOLVE_UPPER:
    case C2B_OP_SOLVE_LOWER_REV:
    case C2B_OP_SOLVE_UPPER_REV: e = J * J + J * K; s = J * K; break;
    default: set_error("c2b_segment_sizes: operator %d cannot run as a segment", (int)op); return C2B_ERR_INVALID_ARGUMENT;
  }
  if (summary_doubles) *summary_doubles = e;
  if (state_doubles) *state_doubles = s;
  return C2B_OK;
}

int c2b_segment_carry(int32_t op, int32_t J, int32_t K, int32_t nseg, int32_t which, const double *summaries, double *state_out,
                      void *stream_) {
  using namespace c2b;
  C2B_REQUIRE(summaries && state_out, "null data pointer");
  C2B_REQUIRE(nseg >= 1 && which >= 0 && which < nseg, "segment index out of range");
  C2B_REQUIRE(K == 1, "segments: one right-hand side only");
  cudaStream_t stream = (cudaStream_t)stream_;
  int rc              = C2B_ERR_INVALID_ARGUMENT;
  if (op == C2B_OP_FACTOR || op == C2B_OP_FACTOR_REV) {
    const int kind = op == C2B_OP_FACTOR ? 0 : 1;
#define CALL(JT) rc = launch_seg_carry_factor<JT>(kind, nseg, which, summaries, state_out, stream)
    C2B_DISPATCH_J_FACTOR(J, CALL)
#undef CALL
  } else if (op == C2B_OP_SOLVE_LOWER || op == C2B_OP_SOLVE_LOWER_REV || op == C2B_OP_SOLVE_UPPER || op == C2B_OP_SOLVE_UPPER_REV) {
    const int kind = op == C2B_OP_SOLVE_LOWER ? 2 : (op == C2B_OP_SOLVE_LOWER_REV ? 3 : (op == C2B_OP_SOLVE_UPPER ? 4 : 5));
#define CALL(JT) rc = launch_seg_carry_sweep<JT>(kind, nseg, which, summaries, state_out, stream)
    C2B_DISPATCH_J_SWEEP(J, CALL)
#undef CALL
  } else {
    set_error("c2b_segment_carry: operator %d cannot run as a segment", (int)op);
  }
  return rc;
}

}  // extern "C"
