// C ABI: solve_lower / solve_upper / matmul_lower / matmul_upper and their reverse-mode counterparts.
#include "host_common.cuh"

namespace c2b {

// CTAs per SM the chunk plan of a K = 1 sweep is sized for: what the kernels of the width can really keep resident (the
// launch bounds only promise C2B_MINB_SWEEP; at J = 6 the registers and tiles allow 8, r02e: summary and apply passes
// -10 % with 8 resident CTAs), never more than 8 (more chunks cost the scans more than the passes gain)
static int sweep_ctas(int J, int K) {
  if (K != 1) return C2B_MINB_SWEEP;
  int occ = 0;
#define CALL(JT) occ = sweep_resident_ctas<JT>()
  C2B_DISPATCH_J_SWEEP(J, CALL)
#undef CALL
  if (occ < C2B_MINB_SWEEP) return C2B_MINB_SWEEP;
  return occ > 8 ? 8 : occ;
}

static int sizes_sweep(bool solve, int64_t N, int J, int K, int64_t batch, Plan *pl, Arena &ar, ScanBuffers *sb) {
  static const int64_t units = env_units("C2B_UNITS_SWEEP");
  *pl = make_plan(N, batch, (K + 7) / 8, J, sweep_ctas(J, K), K == 1 ? units : 0);  // K columns already supply parallelism: fewer chunks per series
  if (pl->C > 1 || batch == 1) {
    const int es = solve ? J * J + J * K : J + J * K;
    take_scan(ar, batch, pl->C, es, J * K, sb);
  }
  return C2B_OK;
}

size_t scratch_sweep(bool solve, int64_t N, int J, int K, int64_t batch) {
  Plan pl;
  Arena ar = sizing_arena();
  ScanBuffers sb;
  sizes_sweep(solve, N, J, K, batch, &pl, ar, &sb);
  return ar.used;
}

static int sizes_sweep_rev(bool solve, int64_t N, int J, int K, int64_t batch, Plan *pl, Arena &ar, ScanBuffers *sb, double **bc_part,
                           double **bt_edge) {
  static const int64_t units = env_units("C2B_UNITS_SWEEP");
  *pl = make_plan(N, batch, (K + 7) / 8, J, sweep_ctas(J, K), K == 1 ? units : 0);  // K columns already supply parallelism: fewer chunks per series
  if (pl->C > 1 || batch == 1) {
    const int es = solve ? J * J + J * K : J + J * K;
    take_scan(ar, batch, pl->C, es, J * K, sb);
  }
  *bc_part = ar.take<double>((size_t)batch * pl->C * J);
  *bt_edge = ar.take<double>((size_t)batch * pl->C);
  return C2B_OK;
}

size_t scratch_sweep_rev(bool solve, int64_t N, int J, int K, int64_t batch) {
  Plan pl;
  Arena ar = sizing_arena();
  ScanBuffers sb;
  double *a, *b;
  sizes_sweep_rev(solve, N, J, K, batch, &pl, ar, &sb, &a, &b);
  return ar.used;
}


int segment_setup(const c2b_segment_t *seg, int64_t N, int64_t batch, const void *scratch, bool reverse_op, Seg *out, int64_t *tmin,
                  int64_t *tmax);  // api_factor.cu

template <bool SOLVE, bool LOWER>
static int sweep_entry(int64_t N, int J, int K, int64_t batch, const double *t, const double *c, const double *U, const double *W,
                       const double *Y, double *Z, double *F, int flags, void *scratch, size_t scratch_bytes, void *stream_,
                       const c2b_segment_t *seg = nullptr, const double *zdiv = nullptr) {
  int rc = check_common(N, J, batch);
  if (rc) return rc;
  C2B_REQUIRE(K >= 1, "K (number of right-hand sides) must be >= 1");
  C2B_REQUIRE(t && c && U && W && Y && Z, "null data pointer");
  Seg sg;
  int64_t tmin, tmax;
  rc = segment_setup(seg, N, batch, scratch, false, &sg, &tmin, &tmax);
  if (rc) return rc;
  cudaStream_t stream = (cudaStream_t)stream_;
  Plan pl;
  ScanBuffers sb;
  {
    Arena sz = sizing_arena();
    sizes_sweep(SOLVE, N, J, K, batch, &pl, sz, &sb);
    Scratch sc;
    rc = sc.init(scratch, scratch_bytes, sz.used, stream);
    if (rc) return rc;
    Arena ar(sc.ptr, sc.bytes);
    sizes_sweep(SOLVE, N, J, K, batch, &pl, ar, &sb);
    SweepArgs a;
    memset(&a, 0, sizeof(a));
    a.N = N; a.batch = batch; a.C = pl.C; a.L = pl.L; a.J = J; a.tmin = tmin; a.tmax = tmax; a.seg = sg; a.K = K;
    a.t = t; a.c = c; a.Y = Y; a.Z = Z; a.F = F;
    a.zdiv = SOLVE ? zdiv : nullptr;
    a.A = LOWER ? U : W;
    a.B = LOWER ? W : U;
    a.t_bs = (flags & C2B_SHARED_T) ? 0 : N;
    a.c_bs = (flags & C2B_SHARED_C) ? 0 : J;
    a.y_bs = (flags & C2B_SHARED_Y) ? 0 : N * (int64_t)K;
    if (SOLVE && (flags & C2B_SHARED_Y) && batch > 1) C2B_REQUIRE(Z != Y, "in-place solve needs a per-series Y");
#define CALL(JT) rc = launch_sweep<JT, SOLVE, LOWER>(a, sb, stream)
    C2B_DISPATCH_J_SWEEP(J, CALL)
#undef CALL
    return rc;
  }
}


template <bool SOLVE, bool LOWER>
static int sweep_rev_entry(int64_t N, int J, int K, int64_t batch, const double *t, const double *c, const double *U, const double *W,
                           const double *Y, const double *Z, const double *F, const double *bZ, double *bt, double *bc, double *bU,
                           double *bW, double *bY, int flags, void *scratch, size_t scratch_bytes, void *stream_,
                           const c2b_segment_t *seg = nullptr) {
  int rc = check_common(N, J, batch);
  if (rc) return rc;
  C2B_REQUIRE(K >= 1, "K (number of right-hand sides) must be >= 1");
  C2B_REQUIRE(t && c && U && W && Y && Z && F && bZ && bt && bc && bU && bW && bY, "null data pointer");
  Seg sg;
  int64_t tmin, tmax;
  rc = segment_setup(seg, N, batch, scratch, true, &sg, &tmin, &tmax);
  if (rc) return rc;
  cudaStream_t stream = (cudaStream_t)stream_;
  Plan pl;
  ScanBuffers sb;
  double *bc_part, *bt_edge;
  Arena sz = sizing_arena();
  sizes_sweep_rev(SOLVE, N, J, K, batch, &pl, sz, &sb, &bc_part, &bt_edge);
  Scratch sc;
  rc = sc.init(scratch, scratch_bytes, sz.used, stream);
  if (rc) return rc;
  Arena ar(sc.ptr, sc.bytes);
  sizes_sweep_rev(SOLVE, N, J, K, batch, &pl, ar, &sb, &bc_part, &bt_edge);
  SweepRevArgs a;
  memset(&a, 0, sizeof(a));
  a.N = N; a.batch = batch; a.C = pl.C; a.L = pl.L; a.J = J; a.tmin = tmin; a.tmax = tmax; a.seg = sg; a.K = K;
  a.t = t; a.c = c; a.Y = Y; a.Z = Z; a.F = F; a.bZ = bZ;
  a.A = LOWER ? U : W;
  a.B = LOWER ? W : U;
  a.bt = bt; a.bY = bY;
  a.bA = LOWER ? bU : bW;
  a.bB = LOWER ? bW : bU;
  a.t_bs = (flags & C2B_SHARED_T) ? 0 : N;
  a.c_bs = (flags & C2B_SHARED_C) ? 0 : J;
  a.y_bs = (flags & C2B_SHARED_Y) ? 0 : N * (int64_t)K;
  a.bc_part = bc_part;
  a.bt_edge = bt_edge;
#define CALL(JT) rc = launch_sweep_rev<JT, SOLVE, LOWER>(a, sb, bc, stream)
  C2B_DISPATCH_J_SWEEP(J, CALL)
#undef CALL
  return rc;
}

// Z[b][n][k] = Y[b][n][k] * sqrt(d[b][n]): the first half of dot_tril (python/celerite2/numpy.py:100-102)
static __global__ void __launch_bounds__(256) scale_sqrt_rows_kernel(int64_t total, int64_t NK, int K, int64_t y_bs, const double *d,
                                                                     const double *Y, double *Z) {
  pdl_prologue();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t b = i / NK, r = i - b * NK;
    Z[i]            = Y[b * y_bs + r] * sqrt(d[i / K]);
  }
}

}  // namespace c2b

extern "C" {

int c2b_apply_inverse(int64_t N, int32_t J, int32_t K, int64_t batch, const double *t, const double *c, const double *U, const double *W,
                      const double *d, const double *Y, double *Z, int32_t flags, void *scratch, size_t scratch_bytes, void *stream) {
  c2b::OpScope _op("c2b_apply_inverse");
  C2B_REQUIRE(d != nullptr, "null data pointer");
  // Z = L^-1 Y with the division by d folded into the store of every row, then Z = L^-T Z in place
  int rc = c2b::sweep_entry<true, true>(N, J, K, batch, t, c, U, W, Y, Z, nullptr, flags, scratch, scratch_bytes, stream, nullptr, d);
  if (rc) return rc;
  return c2b::sweep_entry<true, false>(N, J, K, batch, t, c, U, W, Z, Z, nullptr, flags & ~C2B_SHARED_Y, scratch, scratch_bytes, stream);
}

int c2b_dot_tril(int64_t N, int32_t J, int32_t K, int64_t batch, const double *t, const double *c, const double *U, const double *W,
                 const double *d, const double *Y, double *Z, int32_t flags, void *scratch, size_t scratch_bytes, void *stream) {
  c2b::OpScope _op("c2b_dot_tril");
  using namespace c2b;
  int rc = check_common(N, J, batch);
  if (rc) return rc;
  C2B_REQUIRE(K >= 1 && d && Y && Z, "bad arguments");
  const int64_t NK = N * (int64_t)K, total = batch * NK;
  int64_t blocks = (total + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  C2B_LAUNCH("scale_sqrt_rows_kernel", (cudaStream_t)stream,
             (launch_k(scale_sqrt_rows_kernel, (unsigned)blocks, 256, 0, (cudaStream_t)stream, total, NK, K, (flags & C2B_SHARED_Y) ? 0 : NK, d, Y, Z)));
  // z <- L z in place: the sweep keeps the previous INPUT row (internal.hpp:134-141), so aliasing Y and Z is exact
  return sweep_entry<false, true>(N, J, K, batch, t, c, U, W, Z, Z, nullptr, flags & ~C2B_SHARED_Y, scratch, scratch_bytes, stream);
}

int c2b_solve_lower(int64_t N, int32_t J, int32_t K, int64_t batch, const double *t, const double *c, const double *U, const double *W,
                    const double *Y, double *Z, double *F, int32_t flags, void *scratch, size_t scratch_bytes, void *stream) {
  c2b::OpScope _op("c2b_solve_lower");
  return c2b::sweep_entry<true, true>(N, J, K, batch, t, c, U, W, Y, Z, F, flags, scratch, scratch_bytes, stream);
}
int c2b_solve_upper(int64_t N, int32_t J, int32_t K, int64_t batch, const double *t, const double *c, const double *U, const double *W,
                    const double *Y, double *Z, double *F, int32_t flags, void *scratch, size_t scratch_bytes, void *stream) {
  c2b::OpScope _op("c2b_solve_upper");
  return c2b::sweep_entry<true, false>(N, J, K, batch, t, c, U, W, Y, Z, F, flags, scratch, scratch_bytes, stream);
}
int c2b_matmul_lower(int64_t N, int32_t J, int32_t K, int64_t batch, const double *t, const double *c, const double *U, const double *V,
                     const double *Y, double *Z, double *F, int32_t flags, void *scratch, size_t scratch_bytes, void *stream) {
  c2b::OpScope _op("c2b_matmul_lower");
  return c2b::sweep_entry<false, true>(N, J, K, batch, t, c, U, V, Y, Z, F, flags, scratch, scratch_bytes, stream);
}
int c2b_matmul_upper(int64_t N, int32_t J, int32_t K, int64_t batch, const double *t, const double *c, const double *U, const double *V,
                     const double *Y, double *Z, double *F, int32_t flags, void *scratch, size_t scratch_bytes, void *stream) {
  c2b::OpScope _op("c2b_matmul_upper");
  return c2b::sweep_entry<false, false>(N, J, K, batch, t, c, U, V, Y, Z, F, flags, scratch, scratch_bytes, stream);
}

int c2b_solve_lower_rev(int64_t N, int32_t J, int32_t K, int64_t batch, const double *t, const double *c, const double *U,
                        const double *W, const double *Y, const double *Z, const double *F, const double *bZ, double *bt, double *bc,
                        double *bU, double *bW, double *bY, int32_t flags, void *scratch, size_t scratch_bytes, void *stream) {
  c2b::OpScope _op("c2b_solve_lower_rev");
  return c2b::sweep_rev_entry<true, true>(N, J, K, batch, t, c, U, W, Y, Z, F, bZ, bt, bc, bU, bW, bY, flags, scratch, scratch_bytes,
                                          stream);
}
int c2b_solve_upper_rev(int64_t N, int32_t J, int32_t K, int64_t batch, const double *t, const double *c, const double *U,
                        const double *W, const double *Y, const double *Z, const double *F, const double *bZ, double *bt, double *bc,
                        double *bU, double *bW, double *bY, int32_t flags, void *scratch, size_t scratch_bytes, void *stream) {
  c2b::OpScope _op("c2b_solve_upper_rev");
  return c2b::sweep_rev_entry<true, false>(N, J, K, batch, t, c, U, W, Y, Z, F, bZ, bt, bc, bU, bW, bY, flags, scratch, scratch_bytes,
                                           stream);
}
int c2b_matmul_lower_rev(int64_t N, int32_t J, int32_t K, int64_t batch, const double *t, const double *c, const double *U,
                         const double *V, const double *Y, const double *Z, const double *F, const double *bZ, double *bt, double *bc,
                         double *bU, double *bV, double *bY, int32_t flags, void *scratch, size_t scratch_bytes, void *stream) {
  c2b::OpScope _op("c2b_matmul_lower_rev");
  return c2b::sweep_rev_entry<false, true>(N, J, K, batch, t, c, U, V, Y, Z, F, bZ, bt, bc, bU, bV, bY, flags, scratch, scratch_bytes,
                                           stream);
}
int c2b_matmul_upper_rev(int64_t N, int32_t J, int32_t K, int64_t batch, const double *t, const double *c, const double *U,
                         const double *V, const double *Y, const double *Z, const double *F, const double *bZ, double *bt, double *bc,
                         double *bU, double *bV, double *bY, int32_t flags, void *scratch, size_t scratch_bytes, void *stream) {
  c2b::OpScope _op("c2b_matmul_upper_rev");
  return c2b::sweep_rev_entry<false, false>(N, J, K, batch, t, c, U, V, Y, Z, F, bZ, bt, bc, bU, bV, bY, flags, scratch, scratch_bytes,
                                            stream);
}

int c2b_solve_lower_segment(int64_t N, int32_t J, int32_t K, int64_t batch, const double *t, const double *c, const double *U,
                            const double *W, const double *Y, double *Z, double *F, int32_t flags, void *scratch, size_t scratch_bytes,
                            void *stream, const c2b_segment_t *seg) {
  c2b::OpScope _op("c2b_solve_lower");
  if (!seg) {
    c2b::set_error("c2b_solve_lower_segment: seg is NULL");
    return C2B_ERR_INVALID_ARGUMENT;
  }
  return c2b::sweep_entry<true, true>(N, J, K, batch, t, c, U, W, Y, Z, F, flags, scratch, scratch_bytes, stream, seg);
}
int c2b_solve_lower_rev_segment(int64_t N, int32_t J, int32_t K, int64_t batch, const double *t, const double *c, const double *U,
                                const double *W, const double *Y, const double *Z, const double *F, const double *bZ, double *bt,
                                double *bc, double *bU, double *bW, double *bY, int32_t flags, void *scratch, size_t scratch_bytes,
                                void *stream, const c2b_segment_t *seg) {
  c2b::OpScope _op("c2b_solve_lower_rev");
  if (!seg) {
    c2b::set_error("c2b_solve_lower_rev_segment: seg is NULL");
    return C2B_ERR_INVALID_ARGUMENT;
  }
  return c2b::sweep_rev_entry<true, true>(N, J, K, batch, t, c, U, W, Y, Z, F, bZ, bt, bc, bU, bW, bY, flags, scratch, scratch_bytes,
                                          stream, seg);
}

int c2b_solve_upper_segment(int64_t N, int32_t J, int32_t K, int64_t batch, const double *t, const double *c, const double *U,
                            const double *W, const double *Y, double *Z, double *F, int32_t flags, void *scratch, size_t scratch_bytes,
                            void *stream, const c2b_segment_t *seg) {
  c2b::OpScope _op("c2b_solve_upper");
  if (!seg) {
    c2b::set_error("c2b_solve_upper_segment: seg is NULL");
    return C2B_ERR_INVALID_ARGUMENT;
  }
  return c2b::sweep_entry<true, false>(N, J, K, batch, t, c, U, W, Y, Z, F, flags, scratch, scratch_bytes, stream, seg);
}
int c2b_solve_upper_rev_segment(int64_t N, int32_t J, int32_t K, int64_t batch, const double *t, const double *c, const double *U,
                                const double *W, const double *Y, const double *Z, const double *F, const double *bZ, double *bt,
                                double *bc, double *bU, double *bW, double *bY, int32_t flags, void *scratch, size_t scratch_bytes,
                                void *stream, const c2b_segment_t *seg) {
  c2b::OpScope _op("c2b_solve_upper_rev");
  if (!seg) {
    c2b::set_error("c2b_solve_upper_rev_segment: seg is NULL");
    return C2B_ERR_INVALID_ARGUMENT;
  }
  return c2b::sweep_rev_entry<true, false>(N, J, K, batch, t, c, U, W, Y, Z, F, bZ, bt, bc, bU, bW, bY, flags, scratch, scratch_bytes,
                                           stream, seg);
}

}  // extern "C"
