// C ABI: factor and factor_rev.
#include "host_common.cuh"

namespace c2b {

static void sizes_factor(int64_t N, int J, int64_t batch, Plan *pl, Arena &ar, ScanBuffers *sb) {
  *pl = make_plan(N, batch, 1, J);
  if (pl->C > 1) take_scan(ar, batch, pl->C, 3 * J * J, J * J, sb);
}
size_t scratch_factor(int64_t N, int J, int64_t batch) {
  Plan pl;
  Arena ar = sizing_arena();
  ScanBuffers sb;
  sizes_factor(N, J, batch, &pl, ar, &sb);
  return ar.used;
}

static void sizes_factor_rev(int64_t N, int J, int64_t batch, Plan *pl, Arena &ar, ScanBuffers *sb, double **bc_part, double **bt_edge) {
  *pl = make_plan(N, batch, 1, J);
  if (pl->C > 1) take_scan(ar, batch, pl->C, 2 * J * J, J * J, sb);
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

extern "C" {

int c2b_factor(int64_t N, int32_t J, int64_t batch, const double *t, const double *c, const double *a, const double *U, const double *V,
               double *d, double *W, double *S, int64_t *status, int32_t flags, void *scratch, size_t scratch_bytes, void *stream_) {
  c2b::OpScope _op("c2b_factor");
  using namespace c2b;
  int rc = check_common(N, J, batch);
  if (rc) return rc;
  C2B_REQUIRE(t && c && a && U && V && d && W, "null data pointer");
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
  if (status) C2B_CUDA(cudaMemsetAsync(status, 0, sizeof(int64_t) * (size_t)batch, stream));
  FactorArgs g;
  memset(&g, 0, sizeof(g));
  g.N = N; g.batch = batch; g.C = pl.C; g.L = pl.L; g.J = J;
  g.t = t; g.c = c; g.a = a; g.U = U; g.V = V; g.d = d; g.W = W; g.S = S; g.status = status;
  g.t_bs = (flags & C2B_SHARED_T) ? 0 : N;
  g.c_bs = (flags & C2B_SHARED_C) ? 0 : J;
#define CALL(JT) rc = launch_factor<JT>(g, sb, stream)
  C2B_DISPATCH_J_FACTOR(J, CALL)
#undef CALL
  return rc;
}

int c2b_factor_rev(int64_t N, int32_t J, int64_t batch, const double *t, const double *c, const double *a, const double *U,
                   const double *V, const double *d, const double *W, const double *S, const double *bd, const double *bW, double *bt,
                   double *bc, double *ba, double *bU, double *bV, int32_t flags, void *scratch, size_t scratch_bytes, void *stream_) {
  c2b::OpScope _op("c2b_factor_rev");
  using namespace c2b;
  (void)a;
  (void)V;
  int rc = check_common(N, J, batch);
  if (rc) return rc;
  C2B_REQUIRE(t && c && U && d && W && S && bd && bW && bt && bc && ba && bU && bV, "null data pointer");
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
  g.N = N; g.batch = batch; g.C = pl.C; g.L = pl.L; g.J = J;
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

}  // extern "C"
