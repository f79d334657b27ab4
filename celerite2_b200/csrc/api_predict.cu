// C ABI: the dense pieces of the predictive distribution (python/celerite2/core.py:52-66,134-150): the kernel function
// evaluated on a grid of time differences, and the predictive variance k(0) - diag(K*^T K^-1 K*) reduced without reading
// K* back (its entries are re-evaluated from the coefficients inside the reduction).
#include "host_common.cuh"

namespace c2b {

constexpr int kMaxTerms = C2B_MAX_J;  // Jr + Jc <= J <= 32

struct KernelCoeffs {
  int Jr, Jc;
  double ar[kMaxTerms], cr[kMaxTerms], ac[kMaxTerms], bc[kMaxTerms], cc[kMaxTerms], dc[kMaxTerms];
};

// k(tau) = sum_r ar exp(-cr |tau|) + sum_c exp(-cc |tau|) (ac cos(dc |tau|) + bc sin(dc |tau|))   (python/celerite2/terms.py:64-85)
C2B_DEV double kernel_value(const KernelCoeffs &k, double tau) {
  tau      = fabs(tau);
  double v = 0.0;
  for (int i = 0; i < k.Jr; ++i) v = fma(k.ar[i], exp(-k.cr[i] * tau), v);
  for (int i = 0; i < k.Jc; ++i) {
    double sn, cs;
    sincos(k.dc[i] * tau, &sn, &cs);
    v = fma(exp(-k.cc[i] * tau), fma(k.ac[i], cs, k.bc[i] * sn), v);
  }
  return v;
}

// coefficients live in device memory (they are outputs of device-side term algebra as often as host constants): every
// CTA copies them into shared memory once
C2B_DEV void load_coeffs(KernelCoeffs *dst, int Jr, int Jc, const double *ar, const double *cr, const double *ac, const double *bc,
                         const double *cc, const double *dc) {
  if (threadIdx.x == 0) {
    dst->Jr = Jr;
    dst->Jc = Jc;
  }
  for (int i = threadIdx.x; i < Jr; i += blockDim.x) {
    dst->ar[i] = ar[i];
    dst->cr[i] = cr[i];
  }
  for (int i = threadIdx.x; i < Jc; i += blockDim.x) {
    dst->ac[i] = ac[i];
    dst->bc[i] = bc[i];
    dst->cc[i] = cc[i];
    dst->dc[i] = dc[i];
  }
  __syncthreads();
}

// out[n][m] = k(x1[n] - x2[m]); consecutive threads take consecutive m (coalesced stores)
static __global__ void __launch_bounds__(256) kernel_matrix_kernel(int64_t N, int64_t M, int Jr, int Jc, const double *ar, const double *cr,
                                                                   const double *ac, const double *bc, const double *cc, const double *dc,
                                                                   const double *x1, const double *x2, double *out) {
  pdl_prologue();
  __shared__ KernelCoeffs k;
  load_coeffs(&k, Jr, Jc, ar, cr, ac, bc, cc, dc);
  const int64_t total = N * M;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t n = i / M, m = i - n * M;
    out[i]          = kernel_value(k, x1[n] - x2[m]);
  }
}

// part[chunk][m] = sum over the rows n of the chunk of k(x1[n] - x2[m]) * Z[n][m].  CTA = 32 columns x 8 row lanes.
constexpr int kVarRows = 8;
static __global__ void __launch_bounds__(32 * kVarRows) variance_partial_kernel(int64_t N, int64_t M, int64_t rows_per_chunk, int Jr, int Jc,
                                                                              const double *ar, const double *cr, const double *ac,
                                                                              const double *bc, const double *cc, const double *dc,
                                                                              const double *x1, const double *x2, const double *Z,
                                                                              double *part) {
  pdl_prologue();
  __shared__ KernelCoeffs k;
  __shared__ double red[kVarRows][33];
  load_coeffs(&k, Jr, Jc, ar, cr, ac, bc, cc, dc);
  const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const int64_t m     = (int64_t)blockIdx.x * 32 + cx;
  const int64_t chunk = blockIdx.y;
  const int64_t n0 = chunk * rows_per_chunk, n1 = n0 + rows_per_chunk < N ? n0 + rows_per_chunk : N;
  double s = 0.0;
  if (m < M) {
    const double xm = x2[m];
    for (int64_t n = n0 + ry; n < n1; n += kVarRows) s = fma(kernel_value(k, x1[n] - xm), Z[n * M + m], s);
  }
  red[ry][cx] = s;
  __syncthreads();
  if (ry == 0 && m < M) {
    double v = 0.0;
#pragma unroll
    for (int r = 0; r < kVarRows; ++r) v += red[r][cx];  // fixed order: deterministic
    part[chunk * M + m] = v;
  }
}

static __global__ void __launch_bounds__(256) variance_final_kernel(int64_t M, int64_t nchunk, int Jr, int Jc, const double *ar,
                                                                    const double *ac, const double *part, double *var) {
  pdl_prologue();
  const int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (m >= M) return;
  double k0 = 0.0;  // k(0) = sum ar + sum ac
  for (int i = 0; i < Jr; ++i) k0 += ar[i];
  for (int i = 0; i < Jc; ++i) k0 += ac[i];
  double s = 0.0;
  for (int64_t c = 0; c < nchunk; ++c) s += part[c * M + m];
  var[m] = k0 - s;
}

static int64_t variance_chunks(int64_t N, int64_t M) {
  // enough CTAs to fill the GPU: (M / 32) column tiles x chunks ~ 4 CTAs per SM
  const int64_t tiles = (M + 31) / 32;
  int64_t want        = (148 * 4 + tiles - 1) / tiles;
  const int64_t maxc  = (N + 255) / 256;
  if (want > maxc) want = maxc;
  if (want < 1) want = 1;
  return want;
}

size_t scratch_predict_variance(int64_t N, int64_t M) { return align256((size_t)variance_chunks(N, M) * M * sizeof(double)); }

static int check_terms(int Jr, int Jc, const double *ar, const double *cr, const double *ac, const double *bc, const double *cc,
                       const double *dc) {
  C2B_REQUIRE(Jr >= 0 && Jc >= 0 && Jr + 2 * Jc >= 1 && Jr + 2 * Jc <= C2B_MAX_J, "term counts out of range");
  C2B_REQUIRE((Jr == 0 || (ar && cr)) && (Jc == 0 || (ac && bc && cc && dc)), "null coefficient pointer");
  return C2B_OK;
}

}  // namespace c2b

extern "C" {

int c2b_kernel_matrix(int64_t N, int64_t M, int32_t Jr, int32_t Jc, const double *ar, const double *cr, const double *ac, const double *bc,
                      const double *cc, const double *dc, const double *x1, const double *x2, double *K_out, void *stream) {
  c2b::OpScope _op("c2b_kernel_matrix");
  using namespace c2b;
  int rc = check_terms(Jr, Jc, ar, cr, ac, bc, cc, dc);
  if (rc) return rc;
  C2B_REQUIRE(N >= 1 && M >= 1 && x1 && x2 && K_out, "bad arguments");
  int64_t blocks = (N * M + 255) / 256;
  if (blocks > 148 * 32) blocks = 148 * 32;
  C2B_LAUNCH("kernel_matrix_kernel", (cudaStream_t)stream,
             (launch_k(kernel_matrix_kernel, (unsigned)blocks, 256, 0, (cudaStream_t)stream, N, M, Jr, Jc, ar, cr, ac, bc, cc, dc, x1, x2, K_out)));
  return C2B_OK;
}

int c2b_predict_variance(int64_t N, int64_t M, int32_t Jr, int32_t Jc, const double *ar, const double *cr, const double *ac,
                         const double *bc, const double *cc, const double *dc, const double *x1, const double *x2, const double *Z,
                         double *var_out, void *scratch, size_t scratch_bytes, void *stream_) {
  c2b::OpScope _op("c2b_predict_variance");
  using namespace c2b;
  int rc = check_terms(Jr, Jc, ar, cr, ac, bc, cc, dc);
  if (rc) return rc;
  C2B_REQUIRE(N >= 1 && M >= 1 && x1 && x2 && Z && var_out, "bad arguments");
  cudaStream_t stream = (cudaStream_t)stream_;
  Scratch sc;
  rc = sc.init(scratch, scratch_bytes, scratch_predict_variance(N, M), stream);
  if (rc) return rc;
  const int64_t nchunk = variance_chunks(N, M);
  const int64_t rows   = (N + nchunk - 1) / nchunk;
  dim3 grid((unsigned)((M + 31) / 32), (unsigned)nchunk);
  C2B_LAUNCH("variance_partial_kernel", stream,
             (launch_k(variance_partial_kernel, grid, 32 * kVarRows, 0, stream, N, M, rows, Jr, Jc, ar, cr, ac, bc, cc, dc, x1, x2, Z, (double *)sc.ptr)));
  C2B_LAUNCH("variance_final_kernel", stream,
             (launch_k(variance_final_kernel, (unsigned)((M + 255) / 256), 256, 0, stream, M, nchunk, Jr, Jc, ar, ac, (const double *)sc.ptr, var_out)));
  return C2B_OK;
}

}  // extern "C"
