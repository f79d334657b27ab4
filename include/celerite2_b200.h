/*
 * celerite2_b200 -- C ABI of the B200-native semiseparable Gaussian-process core.
 *
 * One entry point per operation of the reference's hot path.  Each comment names
 * the reference interface the entry point replaces (paths relative to the reference checkout).
 *
 * Conventions
 *   - every data pointer is a DEVICE pointer to FP64, C-contiguous, batch-major:
 *       t (batch,N) | c (batch,J) | a,d (batch,N) | U,V,W (batch,N,J) | Y,Z (batch,N,K)
 *       S (batch,N,J*J)  row n = column-major J x J   (c++/include/celerite2/forward.hpp:97-103,116-120)
 *       F (batch,N,J*K)  row n = column-major J x K   (c++/include/celerite2/internal.hpp:127-129)
 *     `flags` may drop the batch dimension of t, c or Y (C2B_SHARED_*): one array shared by all series.
 *   - batch == 1 is the reference's single-series call.
 *   - calls are asynchronous on `stream` (a cudaStream_t passed as void*; NULL = default stream).
 *   - `scratch`: chunked-scan workspace in device memory.  Ask c2b_scratch_bytes() for the size; pass a
 *     buffer of at least that many bytes (256-byte aligned), or NULL to let the library allocate and free
 *     one around the call with cudaMallocAsync/cudaFreeAsync on `stream`.
 *   - return value: 0 on success, a negative C2B_ERR_* code otherwise (text via c2b_last_error()).
 *     Nothing here synchronises the device; numerical failure of `factor` is reported through the
 *     device-side `status` array, exactly one int64 per series (see c2b_factor).
 *   - supported widths: 1 <= J <= 32 (compiled fast paths for J = 1..16), any K >= 1, any N >= 1.
 */
#ifndef CELERITE2_B200_H_
#define CELERITE2_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define C2B_OK 0
#define C2B_ERR_INVALID_ARGUMENT (-1) /* bad size / null pointer / unsupported width */
#define C2B_ERR_SCRATCH_TOO_SMALL (-2)
#define C2B_ERR_CUDA (-3) /* a CUDA runtime call failed; see c2b_last_error() */
#define C2B_ERR_NO_DEVICE (-4)

#define C2B_SHARED_T 1 /* t has shape (N) for all series          */
#define C2B_SHARED_C 2 /* c (and the coefficient vectors of the assembly) are shared */
#define C2B_SHARED_Y 4 /* Y (or y) has no batch dimension          */
#define C2B_ACCUMULATE 8 /* c2b_factor_rev only: add to bt, bc and bU instead of overwriting them */

#define C2B_MAX_J 32

/* Operation ids for c2b_scratch_bytes */
enum c2b_op {
  C2B_OP_FACTOR           = 0,
  C2B_OP_FACTOR_REV       = 1,
  C2B_OP_SOLVE_LOWER      = 2,
  C2B_OP_SOLVE_UPPER      = 3,
  C2B_OP_MATMUL_LOWER     = 4,
  C2B_OP_MATMUL_UPPER     = 5,
  C2B_OP_SOLVE_LOWER_REV  = 6,
  C2B_OP_SOLVE_UPPER_REV  = 7,
  C2B_OP_MATMUL_LOWER_REV = 8,
  C2B_OP_MATMUL_UPPER_REV = 9,
  C2B_OP_GENERAL_MATMUL   = 10,
  C2B_OP_ASSEMBLE_REV     = 11,
  C2B_OP_PREDICT_VARIANCE = 12
};

const char *c2b_version(void);
/* Text of the last error on the calling thread ("" if none). */
const char *c2b_last_error(void);
/* Number of CUDA devices visible to the library (0 => every compute call fails with C2B_ERR_NO_DEVICE). */
int32_t c2b_device_count(void);
/* Bytes of scratch the given operation needs for these sizes (M only matters for GENERAL_MATMUL). */
size_t c2b_scratch_bytes(int32_t op, int64_t N, int64_t M, int32_t J, int32_t K, int64_t batch);
/*
 * Built-in kernel timing for benchmarks.  While enabled, every kernel the library launches is bracketed by a
 * pair of CUDA events on its stream.  c2b_profile_report() waits for the recorded events and returns one line
 * per (entry point, kernel): "<entry>/<kernel> <launches> <total ms>".  Enabling (or disabling) clears the
 * record.  The returned string stays valid until the next call.
 */
void c2b_profile_enable(int32_t on);
const char *c2b_profile_report(void);
/* Override the chunk planner (testing / tuning): chunks per series; 0 restores the heuristic. */
void c2b_testing_set_chunks_per_series(int64_t chunks);

/*
 * Dense N x N matrix of the semiseparable representation: replaces celerite2::core::to_dense,
 * c++/include/celerite2/forward.hpp:19-42 (a test utility there as here: O(N^2 J) work, N^2 doubles of output).
 * t (N)  c (J)  a (N)  U,V (N,J)  ->  K_out (N,N) row-major, symmetric.
 */
int c2b_to_dense(int64_t N, int32_t J, const double *t, const double *c, const double *a, const double *U, const double *V, double *K_out,
                 void *stream);

/*
 * Kernel-term algebra on the device: physical hyper-parameters -> coefficient vectors, for a batch of
 * parameter sets resident on the GPU, with the adjoint maps.  Replaces Term.get_coefficients() of
 * python/celerite2/terms.py: RealTerm 485-521 (kind 0; params a, c), ComplexTerm 524-569 (kind 1; a, b, c, d),
 * SHOTerm 604-691 (kind 2; S0, w0, Q; `underdamped` selects the one-complex-term (Q >= 1/2) or two-real-terms form for the
 * whole batch, `eps` as in the reference), Matern32Term 694-745 (kind 3; sigma, rho; eps), RotationTerm 748-814 (kind 4;
 * sigma, period, Q0, dQ, f).  params is (batch, P) row-major.  TermSum (terms.py:210-235) is concatenation: every term
 * writes its own slots [r_off, ..) / [c_off, ..) of coefficient arrays with rows of Jr_total / Jc_total entries.
 * `bad` (nullable, one int32, caller zeroes it) is set to 1 if a parameter set falls on the other side of Q = 1/2.
 * The adjoint reads the cotangents of the coefficient arrays (same slots) and overwrites b_params (batch, P).
 * c2b_term_product: coefficients of the product of two kernels (TermProduct, terms.py:238-300), Jr = Jr1 Jr2,
 * Jc = Jr1 Jc2 + Jr2 Jc1 + 2 Jc1 Jc2 in the reference's order; _rev overwrites the cotangents of both factors.
 * TermDiff and TermConvolution (terms.py:304-410) are not provided.
 */
int c2b_term_coefficients(int32_t kind, int64_t batch, const double *params, double eps, int32_t underdamped, int32_t Jr_total,
                          int32_t Jc_total, int32_t r_off, int32_t c_off, double *ar, double *cr, double *ac, double *bc, double *cc,
                          double *dc, int32_t *bad, void *stream);
int c2b_term_coefficients_rev(int32_t kind, int64_t batch, const double *params, double eps, int32_t underdamped, int32_t Jr_total,
                              int32_t Jc_total, int32_t r_off, int32_t c_off, const double *b_ar, const double *b_cr, const double *b_ac,
                              const double *b_bc, const double *b_cc, const double *b_dc, double *b_params, void *stream);
int c2b_term_product(int64_t batch, int32_t Jr1, int32_t Jc1, const double *ar1, const double *cr1, const double *ac1, const double *bc1,
                     const double *cc1, const double *dc1, int32_t Jr2, int32_t Jc2, const double *ar2, const double *cr2,
                     const double *ac2, const double *bc2, const double *cc2, const double *dc2, double *ar, double *cr, double *ac,
                     double *bc, double *cc, double *dc, void *stream);
int c2b_term_product_rev(int64_t batch, int32_t Jr1, int32_t Jc1, const double *ar1, const double *cr1, const double *ac1,
                         const double *bc1, const double *cc1, const double *dc1, int32_t Jr2, int32_t Jc2, const double *ar2,
                         const double *cr2, const double *ac2, const double *bc2, const double *cc2, const double *dc2,
                         const double *b_ar, const double *b_cr, const double *b_ac, const double *b_bc, const double *b_cc,
                         const double *b_dc, double *g1_ar, double *g1_cr, double *g1_ac, double *g1_bc, double *g1_cc, double *g1_dc,
                         double *g2_ar, double *g2_cr, double *g2_ac, double *g2_bc, double *g2_cc, double *g2_dc, void *stream);

/*
 * Dense pieces of the predictive distribution, python/celerite2/core.py:52-66,134-150.  The reference
 * builds K* = kernel.get_value(x[:, None] - xs[None, :]) (N x M) in numpy, applies K^-1 to it and takes
 * k(0) - einsum("ij,ij->j", K*, K^-1 K*) (numpy.py:24-25).
 *   c2b_kernel_matrix      K_out (N,M) row-major = k(x1[n] - x2[m]) from the six coefficient vectors (terms.py:64-85):
 *                          K* with x1 = x, x2 = xs; the prior covariance of the predictions with x1 = x2 = xs.
 *   c2b_predict_variance   var_out (M) = k(0) - sum_n k(x1[n] - x2[m]) Z[n][m] with Z (N,M) = K^-1 K* (c2b_apply_inverse on
 *                          the output of c2b_kernel_matrix); K* is not read back, its entries are re-evaluated in the
 *                          reduction.  Scratch: c2b_scratch_bytes(C2B_OP_PREDICT_VARIANCE, N, M, ...).
 * Coefficients in device memory: ar, cr (Jr)  ac, bc, cc, dc (Jc).
 */
int c2b_kernel_matrix(int64_t N, int64_t M, int32_t Jr, int32_t Jc, const double *ar, const double *cr, const double *ac, const double *bc,
                      const double *cc, const double *dc, const double *x1, const double *x2, double *K_out, void *stream);
int c2b_predict_variance(int64_t N, int64_t M, int32_t Jr, int32_t Jc, const double *ar, const double *cr, const double *ac,
                         const double *bc, const double *cc, const double *dc, const double *x1, const double *x2, const double *Z,
                         double *var_out, void *scratch, size_t scratch_bytes, void *stream);

/*
 * a/U/V (and c) assembly from the kernel coefficients.
 *   layout 0 "blocked"     [real.. | cos.. | sin..]       replaces celerite2::Term::get_celerite_matrices,
 *                                                         c++/include/celerite2/terms.hpp:112-139
 *   layout 1 "interleaved" [real.. | cos0 sin0 cos1 ..]   replaces driver.get_celerite_matrices,
 *                                                         python/celerite2/driver.cpp:422-477 (+ c fill, terms.py:171-173)
 * ar,cr (batch,Jr)  ac,bc,cc,dc (batch,Jc)  x,diag (batch,N)  ->  c (batch,J)  a (batch,N)  U,V (batch,N,J),
 * J = Jr + 2 Jc.  c_out may be NULL.  C2B_SHARED_T: x shared; C2B_SHARED_C: coefficient vectors shared.
 */
int c2b_get_celerite_matrices(int32_t layout, int64_t N, int32_t Jr, int32_t Jc, int64_t batch, const double *ar, const double *cr,
                              const double *ac, const double *bc, const double *cc, const double *dc, const double *x,
                              const double *diag, double *c_out, double *a_out, double *U_out, double *V_out, int32_t flags,
                              void *stream);

/*
 * Adjoint of the assembly: cotangents of (c, a, U, V) -> cotangents of the six coefficient vectors, of x and
 * of diag.  The reference has no C++ counterpart: JAX differentiates python/celerite2/jax/terms.py:119-135
 * itself; this kernel is that derivative written out.  b_c may be NULL (treated as zero).  b_x / b_diag may be
 * NULL.  b_x is ACCUMULATED into (so the caller can pre-load the `bt` of the sweeps); everything else is overwritten.
 * Coefficient cotangents have shape (batch,Jr) / (batch,Jc) even when the coefficients are shared.
 */
int c2b_get_celerite_matrices_rev(int32_t layout, int64_t N, int32_t Jr, int32_t Jc, int64_t batch, const double *ar, const double *cr,
                                  const double *ac, const double *bc, const double *cc, const double *dc, const double *x,
                                  const double *b_c, const double *b_a, const double *b_U, const double *b_V, double *b_ar,
                                  double *b_cr, double *b_ac, double *b_bc, double *b_cc, double *b_dc, double *b_x, double *b_diag,
                                  int32_t flags, void *scratch, size_t scratch_bytes, void *stream);

/*
 * ---- One series split into contiguous segments (one per GPU) ------------------------------------------------------------
 * No reference counterpart: the reference is sequential over N ("one long series").  A segment is a
 * run of rows [r0, r1) of the series; the caller passes pointers to the segment's first row (t + r0, U + r0*J, ...) and
 * N = r1 - r0.  The time stamps of the neighbouring rows must be addressable through the same pointer when has_left /
 * has_right say so (t[-1], t[N]): replicate t on every device or keep a one-element halo.  batch must be 1 and the
 * width one of the compiled ones (J in 1..16; sweeps: K = 1).  Each of the four operators of the
 * log-likelihood + gradient pipeline runs in two calls with the SAME scratch buffer (which holds the tree in between):
 *   phase C2B_SEG_SUMMARISE: summary pass + up-sweep; the segment's summary element (c2b_segment_sizes doubles) is
 *                            written to summary_out.  Exchange the summaries of all segments (all-gather),
 *   c2b_segment_carry:       folds the summaries of the segments that come earlier in the operator's processing order
 *                            (lower rows for factor / solve_lower, higher rows for the two reverse operators) into
 *                            the state entering this segment,
 *   phase C2B_SEG_APPLY:     state_in -> down-sweep -> apply pass; outputs as for the plain entry point, for this
 *                            segment's rows.  Reverse operators also return in edge_out the one bt term owed to the row
 *                            just below the segment (add it to the last bt of the lower neighbour); bc is this
 *                            segment's partial sum (add over segments).
 */
typedef struct c2b_segment {
  int32_t phase;     /* C2B_SEG_SUMMARISE or C2B_SEG_APPLY */
  int32_t has_left;  /* the series continues below row 0 of the segment */
  int32_t has_right; /* the series continues above row N-1 */
  int32_t reserved;
  const double *state_in; /* APPLY: device, c2b_segment_sizes state doubles; NULL = zero state */
  double *summary_out;    /* SUMMARISE: device, c2b_segment_sizes summary doubles */
  double *edge_out;       /* APPLY, reverse operators: device, 1 double; may be NULL */
} c2b_segment_t;
#define C2B_SEG_SUMMARISE 1
#define C2B_SEG_APPLY 2

int c2b_segment_sizes(int32_t op, int32_t J, int32_t K, int64_t *summary_doubles, int64_t *state_doubles);
/* summaries: device, nseg elements in series order.  state_out: device.  op as for c2b_scratch_bytes. */
int c2b_segment_carry(int32_t op, int32_t J, int32_t K, int32_t nseg, int32_t which, const double *summaries, double *state_out,
                      void *stream);
int c2b_factor_segment(int64_t N, int32_t J, int64_t batch, const double *t, const double *c, const double *a, const double *U,
                       const double *V, double *d, double *W, double *S, int64_t *status, int32_t flags, void *scratch,
                       size_t scratch_bytes, void *stream, const c2b_segment_t *seg);
int c2b_factor_rev_segment(int64_t N, int32_t J, int64_t batch, const double *t, const double *c, const double *a, const double *U,
                           const double *V, const double *d, const double *W, const double *S, const double *bd, const double *bW,
                           double *bt, double *bc, double *ba, double *bU, double *bV, int32_t flags, void *scratch,
                           size_t scratch_bytes, void *stream, const c2b_segment_t *seg);
int c2b_solve_lower_segment(int64_t N, int32_t J, int32_t K, int64_t batch, const double *t, const double *c, const double *U,
                            const double *W, const double *Y, double *Z, double *F, int32_t flags, void *scratch, size_t scratch_bytes,
                            void *stream, const c2b_segment_t *seg);
int c2b_solve_lower_rev_segment(int64_t N, int32_t J, int32_t K, int64_t batch, const double *t, const double *c, const double *U,
                                const double *W, const double *Y, const double *Z, const double *F, const double *bZ, double *bt,
                                double *bc, double *bU, double *bW, double *bY, int32_t flags, void *scratch, size_t scratch_bytes,
                                void *stream, const c2b_segment_t *seg);
/* solve_upper walks downwards: the segments that come earlier are the higher rows; its reverse operator walks upwards and
 * owes its edge term to the row just ABOVE the segment (add it to the first bt of the upper neighbour). */
int c2b_solve_upper_segment(int64_t N, int32_t J, int32_t K, int64_t batch, const double *t, const double *c, const double *U,
                            const double *W, const double *Y, double *Z, double *F, int32_t flags, void *scratch, size_t scratch_bytes,
                            void *stream, const c2b_segment_t *seg);
int c2b_solve_upper_rev_segment(int64_t N, int32_t J, int32_t K, int64_t batch, const double *t, const double *c, const double *U,
                                const double *W, const double *Y, const double *Z, const double *F, const double *bZ, double *bt,
                                double *bc, double *bU, double *bW, double *bY, int32_t flags, void *scratch, size_t scratch_bytes,
                                void *stream, const c2b_segment_t *seg);

/*
 * LDL^T factorisation.  Replaces celerite2::core::factor, c++/include/celerite2/forward.hpp:69-135
 * (S != NULL) and the no-workspace overload c++/include/celerite2/interface.hpp:37-48 (S == NULL).
 * d may alias a and W may alias V (forward.hpp:55-58).
 * status (batch int64, may be NULL): 0, or the first n >= 1 with d[n] <= 0 (forward.hpp:128).  When a
 * series runs as ONE chunk (always the case for N below the chunking threshold) rows after that index are
 * left untouched, as in the reference; when it runs chunked, rows after it hold unspecified values.
 */
int c2b_factor(int64_t N, int32_t J, int64_t batch, const double *t, const double *c, const double *a, const double *U,
               const double *V, double *d, double *W, double *S, int64_t *status, int32_t flags, void *scratch,
               size_t scratch_bytes, void *stream);

/*
 * Reverse mode of factor.  Replaces celerite2::core::factor_rev, c++/include/celerite2/reverse.hpp:10-85.
 * a and V are accepted for signature parity and never read (reverse.hpp:29-30).  All outputs are overwritten,
 * unless flags has C2B_ACCUMULATE: then bt, bc and bU -- the cotangents of inputs that other operators of a
 * pipeline read as well (t, c, U) -- are added to, so that the caller can pass the arrays solve_*_rev filled and
 * skip three separate additions (what an autodiff framework does after the reference's factor_rev; no reference
 * counterpart).  ba and bV are always overwritten.  bc has shape (batch,J) even when c is shared.
 */
int c2b_factor_rev(int64_t N, int32_t J, int64_t batch, const double *t, const double *c, const double *a, const double *U,
                   const double *V, const double *d, const double *W, const double *S, const double *bd, const double *bW,
                   double *bt, double *bc, double *ba, double *bU, double *bV, int32_t flags, void *scratch,
                   size_t scratch_bytes, void *stream);

/*
 * The four linear sweeps.  F may be NULL (the no-workspace overloads, interface.hpp:70-152).  Z may alias Y.
 *   c2b_solve_lower   Z = L^-1 Y          celerite2::core::solve_lower,  forward.hpp:156-170 -> internal.hpp:105-146
 *   c2b_solve_upper   Z = L^-T Y          celerite2::core::solve_upper,  forward.hpp:193-207 -> internal.hpp:148-189
 *   c2b_matmul_lower  Z += tril(U V^T) Y  celerite2::core::matmul_lower, forward.hpp:228-239 (Z is accumulated into)
 *   c2b_matmul_upper  Z += triu(V U^T) Y  celerite2::core::matmul_upper, forward.hpp:260-271 (Z is accumulated into)
 */
int c2b_solve_lower(int64_t N, int32_t J, int32_t K, int64_t batch, const double *t, const double *c, const double *U,
                    const double *W, const double *Y, double *Z, double *F, int32_t flags, void *scratch, size_t scratch_bytes,
                    void *stream);
int c2b_solve_upper(int64_t N, int32_t J, int32_t K, int64_t batch, const double *t, const double *c, const double *U,
                    const double *W, const double *Y, double *Z, double *F, int32_t flags, void *scratch, size_t scratch_bytes,
                    void *stream);
int c2b_matmul_lower(int64_t N, int32_t J, int32_t K, int64_t batch, const double *t, const double *c, const double *U,
                     const double *V, const double *Y, double *Z, double *F, int32_t flags, void *scratch, size_t scratch_bytes,
                     void *stream);
int c2b_matmul_upper(int64_t N, int32_t J, int32_t K, int64_t batch, const double *t, const double *c, const double *U,
                     const double *V, const double *Y, double *Z, double *F, int32_t flags, void *scratch, size_t scratch_bytes,
                     void *stream);

/*
 * The two compositions the reference's GP object makes of these operators, python/celerite2/numpy.py:94-102
 * (SURVEY 8a15), as single calls so that the element-wise step between / before the sweeps costs no extra pass:
 *   c2b_apply_inverse  Z = K^-1 Y = L^-T D^-1 L^-1 Y   (_do_solve: solve_lower, z /= d[:, None], solve_upper); the division
 *                      is folded into the store of the first sweep.  Z may alias Y.
 *   c2b_dot_tril       Z = L (sqrt(d) o Y)              (_do_dot_tril: z = y * sqrt(d)[:, None], matmul_lower in place).
 *                      Z may alias Y.
 * d (batch,N) and W (batch,N,J) are the outputs of c2b_factor.  Scratch: the size c2b_scratch_bytes reports for
 * C2B_OP_SOLVE_LOWER (apply_inverse) / C2B_OP_MATMUL_LOWER (dot_tril) with the same N, J, K, batch.
 */
int c2b_apply_inverse(int64_t N, int32_t J, int32_t K, int64_t batch, const double *t, const double *c, const double *U, const double *W,
                      const double *d, const double *Y, double *Z, int32_t flags, void *scratch, size_t scratch_bytes, void *stream);
int c2b_dot_tril(int64_t N, int32_t J, int32_t K, int64_t batch, const double *t, const double *c, const double *U, const double *W,
                 const double *d, const double *Y, double *Z, int32_t flags, void *scratch, size_t scratch_bytes, void *stream);

/*
 * Reverse mode of the sweeps.  Replace celerite2::core::{solve_lower,solve_upper,matmul_lower,matmul_upper}_rev,
 * c++/include/celerite2/reverse.hpp:87-217 -> internal.hpp:191-303.  All outputs are overwritten.
 * bc has shape (batch,J); bY has shape (batch,N,K) even when Y is shared.
 */
int c2b_solve_lower_rev(int64_t N, int32_t J, int32_t K, int64_t batch, const double *t, const double *c, const double *U,
                        const double *W, const double *Y, const double *Z, const double *F, const double *bZ, double *bt,
                        double *bc, double *bU, double *bW, double *bY, int32_t flags, void *scratch, size_t scratch_bytes,
                        void *stream);
int c2b_solve_upper_rev(int64_t N, int32_t J, int32_t K, int64_t batch, const double *t, const double *c, const double *U,
                        const double *W, const double *Y, const double *Z, const double *F, const double *bZ, double *bt,
                        double *bc, double *bU, double *bW, double *bY, int32_t flags, void *scratch, size_t scratch_bytes,
                        void *stream);
int c2b_matmul_lower_rev(int64_t N, int32_t J, int32_t K, int64_t batch, const double *t, const double *c, const double *U,
                         const double *V, const double *Y, const double *Z, const double *F, const double *bZ, double *bt,
                         double *bc, double *bU, double *bV, double *bY, int32_t flags, void *scratch, size_t scratch_bytes,
                         void *stream);
int c2b_matmul_upper_rev(int64_t N, int32_t J, int32_t K, int64_t batch, const double *t, const double *c, const double *U,
                         const double *V, const double *Y, const double *Z, const double *F, const double *bZ, double *bt,
                         double *bc, double *bU, double *bV, double *bY, int32_t flags, void *scratch, size_t scratch_bytes,
                         void *stream);

/*
 * Rectangular products used by predict / conditional means.  Replace celerite2::core::general_matmul_lower,
 * c++/include/celerite2/forward.hpp:285-332 and general_matmul_upper, forward.hpp:346-392.
 * t1 (N)  t2 (M)  c (J)  U (N,J)  V (M,J)  Y (M,K)  ->  Z (N,K) accumulated into.  Single series (no batch).
 * F (M,J*K, ROW-major rows, forward.hpp:299) may be NULL; when given, the rows the reference writes hold the
 * reference's values and row 0 follows forward.hpp:307,369; other rows are left untouched.
 */
int c2b_general_matmul_lower(int64_t N, int64_t M, int32_t J, int32_t K, const double *t1, const double *t2, const double *c,
                             const double *U, const double *V, const double *Y, double *Z, double *F, void *scratch,
                             size_t scratch_bytes, void *stream);
int c2b_general_matmul_upper(int64_t N, int64_t M, int32_t J, int32_t K, const double *t1, const double *t2, const double *c,
                             const double *U, const double *V, const double *Y, double *Z, double *F, void *scratch,
                             size_t scratch_bytes, void *stream);

/*
 * The elementwise/reduction glue of the reference's Python layer, fused:
 *   logdet = sum log d                      python/celerite2/numpy.py:84
 *   chi2   = sum z^2 / d                    python/celerite2/numpy.py:104-109
 * out (batch,2) = {logdet, chi2}.  If bz / bd are non-NULL they receive the cotangents of
 * loglike = -0.5 (chi2 + logdet + N log 2 pi):  bz = -z/d,  bd = 0.5 z^2/d^2 - 0.5/d.
 * d (batch,N), z (batch,N).
 */
int c2b_norm_logdet(int64_t N, int64_t batch, const double *d, const double *z, double *out, double *bz, double *bd,
                    void *scratch, size_t scratch_bytes, void *stream);
/* The same, and loglike (batch) = -0.5 (chi2 + logdet + N log 2 pi) from the same launch: what
 * GaussianProcess.log_likelihood returns for a zero-mean process (python/celerite2/core.py:407-428). */
int c2b_norm_logdet_loglike(int64_t N, int64_t batch, const double *d, const double *z, double *out, double *loglike, double *bz,
                            double *bd, void *scratch, size_t scratch_bytes, void *stream);

/*
 * solve_lower for callers that cannot allocate device memory.  (No counterpart in the reference.)
 *
 * Some hosts of this library own a fixed set of device buffers and cannot get more while a call is in
 * flight: an XLA/JAX custom-call handler only sees the buffers the compiler planned for it, a service that
 * gave the whole GPU to its own memory pool has nothing left to give, and a callback running in the middle of
 * someone else's step cannot risk the device-wide synchronisation an allocation or a free may trigger.  They
 * can neither pass the scan scratch c2b_solve_lower wants nor let it call cudaMallocAsync.  This variant
 * computes exactly what c2b_solve_lower computes, with the scratch in an array built into the library.
 *
 *   - limit: c2b_noalloc_scratch_limit() bytes (8 MiB).  Calls whose c2b_scratch_bytes(C2B_OP_SOLVE_LOWER, ...)
 *     exceeds it fail with C2B_ERR_SCRATCH_TOO_SMALL and an explanatory c2b_last_error(); 8 MiB covers one series
 *     of N = 1,000,000 at J = 6, K = 1.  The array occupies that much memory on every GPU that loads the library,
 *     until the process exits.
 *   - the call itself performs no cudaMalloc / cudaFree / cudaMallocAsync and calls no allocating library
 *     routine (with c2b_profile_enable off).  Inputs and outputs are still the caller's buffers.
 *   - first call on a device: looks up the device address of the built-in array (cudaGetSymbolAddress, with
 *     the calling thread's current device, which must be the device of `stream`) and caches it.  Make this
 *     first call before starting any stream capture on that device.
 *   - every call on a device shares the one array, and it keeps the previous call's contents.  Nothing in it
 *     needs to start at zero (each word is written before it is read within a call), so nothing is cleared.
 *     Calls that overlap on the GPU -- two streams, two host threads -- race on the array: keep calls to this
 *     entry point on one stream or order them yourself.
 */
size_t c2b_noalloc_scratch_limit(void);
int c2b_solve_lower_noalloc(int64_t N, int32_t J, int32_t K, int64_t batch, const double *t, const double *c, const double *U,
                            const double *W, const double *Y, double *Z, double *F, int32_t flags, void *stream);

/*
 * y[i] += x[i], i < n.  The sum of two cotangents of the same array (e.g. the bU of solve_lower_rev and of
 * factor_rev), which the reference leaves to the autodiff framework (python/celerite2/jax/ops.py:140-151).
 */
int c2b_add_inplace(int64_t n, const double *x, double *y, void *stream);

/*
 * Host-staging helpers for callers that hold HOST data and do not want to include CUDA headers (the drop-in C++
 * header include/celerite2_b200/core.hpp uses them; the reference's callers pass host Eigen objects,
 * c++/include/celerite2/forward.hpp).  Thin wrappers over cudaMalloc / cudaFree / cudaMemcpy / cudaStreamSynchronize
 * on the current device.
 */
void *c2b_device_malloc(size_t bytes);
void c2b_device_free(void *ptr);
int c2b_copy_to_device(void *dst_device, const void *src_host, size_t bytes);
int c2b_copy_to_host(void *dst_host, const void *src_device, size_t bytes);
int c2b_stream_synchronize(void *stream);

#ifdef __cplusplus
}
#endif
#endif /* CELERITE2_B200_H_ */
