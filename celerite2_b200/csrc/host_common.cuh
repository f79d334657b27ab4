// Host-side helpers shared by the API translation units.
#pragma once
#include "args.cuh"
#include "scan.cuh"

namespace c2b {

// Compile-time widths.  Must match WIDTHS_SWEEP / WIDTHS_FACTOR in celerite2_b200/build.py: every width listed
// here needs its inst_*.cu object.  Other widths (up to C2B_MAX_J) take the run-time-width path (JT = 0).
#define C2B_DISPATCH_J_SWEEP(J, CALL)                                                                                                     \
  switch (J) {                                                                                                                            \
    case 1: CALL(1); break;                                                                                                               \
    case 2: CALL(2); break;                                                                                                               \
    case 3: CALL(3); break;                                                                                                               \
    case 4: CALL(4); break;                                                                                                               \
    case 5: CALL(5); break;                                                                                                               \
    case 6: CALL(6); break;                                                                                                               \
    case 7: CALL(7); break;                                                                                                               \
    case 8: CALL(8); break;                                                                                                               \
    case 9: CALL(9); break;                                                                                                               \
    case 10: CALL(10); break;                                                                                                             \
    case 11: CALL(11); break;                                                                                                             \
    case 12: CALL(12); break;                                                                                                             \
    case 13: CALL(13); break;                                                                                                             \
    case 14: CALL(14); break;                                                                                                             \
    case 15: CALL(15); break;                                                                                                             \
    case 16: CALL(16); break;                                                                                                             \
    default: CALL(0); break;                                                                                                              \
  }
#define C2B_DISPATCH_J_FACTOR(J, CALL) C2B_DISPATCH_J_SWEEP(J, CALL)

// Thread-per-unit kernels: few units => small CTAs so that every SM gets some.
static inline void launch_shape(int64_t units, unsigned *grid, unsigned *block) {
  const unsigned b = units < 148 * 128 ? 32u : 128u;
  *block           = b;
  *grid            = (unsigned)((units + b - 1) / b);
  if (*grid == 0) *grid = 1;
}

// Staged kernels (compile-time widths) keep a per-thread ring of input rows in dynamic shared memory:
// slots * row_doubles * 8 bytes per thread.  One warp per CTA keeps the footprint small and lets many CTAs share an SM.
template <typename Kernel>
static inline int staged_shape(Kernel kernel, int64_t units, bool staged, int slots, int row_doubles, unsigned *grid, unsigned *block,
                               size_t *smem) {
  if (!staged) {
    launch_shape(units, grid, block);
    *smem = 0;
    return C2B_OK;
  }
  *block = 32;
  *grid  = (unsigned)((units + 31) / 32);
  if (*grid == 0) *grid = 1;
  *smem = (size_t)slots * row_doubles * 32 * sizeof(double);
  if (*smem > 48 * 1024) C2B_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)*smem));
  return C2B_OK;
}

static inline int check_common(int64_t N, int J, int64_t batch) {
  if (N < 1) {
    set_error("N must be >= 1 (got %lld)", (long long)N);
    return C2B_ERR_INVALID_ARGUMENT;
  }
  if (J < 1 || J > C2B_MAX_J) {
    set_error("J must be in 1..%d (got %d)", C2B_MAX_J, J);
    return C2B_ERR_INVALID_ARGUMENT;
  }
  if (batch < 1) {
    set_error("batch must be >= 1 (got %lld)", (long long)batch);
    return C2B_ERR_INVALID_ARGUMENT;
  }
  return C2B_OK;
}

// Sizing arena: counts bytes without touching memory.
static inline Arena sizing_arena() { return Arena((void *)nullptr, (size_t)-1); }

static inline void take_scan(Arena &ar, int64_t batch, int64_t C, int esize, int ssize, ScanBuffers *sb) {
  Levels lv       = make_levels(C);
  sb->node_stride = lv.total;
  // nodes of all series, then one head product per chunk for the blocked fold (fold.cuh; a head is never larger than an element)
  sb->nodes       = ar.take<double>((size_t)batch * lv.total * esize + (size_t)batch * C * esize);
  sb->states      = ar.take<double>((size_t)batch * lv.total * ssize);
}

size_t scratch_sweep(bool solve, int64_t N, int J, int K, int64_t batch);
size_t scratch_sweep_rev(bool solve, int64_t N, int J, int K, int64_t batch);
size_t scratch_factor(int64_t N, int J, int64_t batch);
size_t scratch_factor_rev(int64_t N, int J, int64_t batch);
size_t scratch_general_matmul(int64_t M, int J, int K);
size_t scratch_assemble_rev(int64_t N, int Jr, int Jc, int64_t batch);
size_t scratch_norm_logdet(int64_t N, int64_t batch);
size_t scratch_predict_variance(int64_t N, int64_t M);

}  // namespace c2b
