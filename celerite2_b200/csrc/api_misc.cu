// C ABI: library plumbing, assembly (+ adjoint), log-det / chi^2 reductions, general_matmul.
#include <stdarg.h>
#include <stdlib.h>

#include <atomic>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include <nvtx3/nvToolsExt.h>

#include "host_common.cuh"
#include "misc.cuh"

namespace c2b {

static thread_local char g_err[512] = "";
static int64_t g_chunk_override    = -1;  // -1: not yet read from the environment

bool g_profile_on               = false;
thread_local const char *g_op   = "";
namespace {
struct ProfRec {
  const char *op, *kernel;
  cudaEvent_t e0, e1;
};
std::vector<ProfRec> g_recs;
std::vector<cudaEvent_t> g_event_pool;
std::mutex g_prof_mutex;
std::string g_report;
cudaEvent_t get_event() {
  if (!g_event_pool.empty()) {
    cudaEvent_t e = g_event_pool.back();
    g_event_pool.pop_back();
    return e;
  }
  cudaEvent_t e;
  cudaEventCreate(&e);
  return e;
}
}  // namespace
void prof_begin(const char *kernel, cudaStream_t s) {
  std::lock_guard<std::mutex> lock(g_prof_mutex);
  ProfRec r{g_op, kernel, get_event(), get_event()};
  cudaEventRecord(r.e0, s);
  g_recs.push_back(r);
}
void prof_end(cudaStream_t s) {
  std::lock_guard<std::mutex> lock(g_prof_mutex);
  if (!g_recs.empty()) cudaEventRecord(g_recs.back().e1, s);
}

static bool nvtx_on() {
  static const bool v = getenv("C2B_NVTX") != nullptr;
  return v;
}
void nvtx_push(const char *name) {
  if (nvtx_on()) nvtxRangePushA(name);
}
void nvtx_pop() {
  if (nvtx_on()) nvtxRangePop();
}

void set_error(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char *what) {
  set_error("CUDA error %d (%s) in %s", (int)e, cudaGetErrorString(e), what);
  return (e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver) ? C2B_ERR_NO_DEVICE : C2B_ERR_CUDA;
}

int64_t get_chunk_override() {
  if (g_chunk_override < 0) {
    const char *s    = getenv("C2B_CHUNKS");
    g_chunk_override = s ? atoll(s) : 0;
    if (g_chunk_override < 0) g_chunk_override = 0;
  }
  return g_chunk_override;
}

// Chunks along N exist only to expose parallelism: batch * cols * C independent units should fill the GPU
// (148 SMs x a few warps).  Short chunks would be dominated by the summary pass, so L >= 64.
//
// For the lane-group kernels (K == 1, J > 0 given) the unit is a group of G(J) lanes and a CTA holds 64 / G groups.
// Every CTA of an op must be resident at once or the last partial wave runs the whole chunk length again
// (r01f profile: 1.14 waves cost factor_rev 25 %), and the heaviest of those kernels fits 6 CTAs per SM
// (registers, shared memory).  So the default is 6 CTAs per SM exactly -- a grid that is a multiple of the SM count --
// and never more than 8192 chunks (scan cost).
static int sm_count() {
  // per device (a process may drive several), lock-free: a race only repeats the query
  static std::atomic<int> cache[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  int n = cache[dev].load(std::memory_order_relaxed);
  if (n == 0) {
    int v = 0;
    n = (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && v > 0) ? v : 148;
    cache[dev].store(n, std::memory_order_relaxed);
  }
  return n;
}

// Developer switch, read from the environment once per process: C2B_NO_GROUP=1 routes the compiled widths to the
// generic thread-per-chunk kernels (A/B comparisons on the GPU box).  Not part of the ABI.
// C2B_PDL: 0 no programmatic dependent launches, 1 (default) the ops of a chunked batch, 2 every launch (common.cuh)
int pdl_mode() {
  static const int v = getenv("C2B_PDL") ? atoi(getenv("C2B_PDL")) : 1;
  return v;
}
thread_local bool g_pdl_wave = false;
bool no_group_kernels() {
  static const bool v = getenv("C2B_NO_GROUP") != nullptr;
  return v;
}

int64_t env_units(const char *name) {
  const char *s = getenv(name);
  const int64_t v = s ? atoll(s) : 0;
  return v > 0 ? v : 0;
}

Plan make_plan(int64_t N, int64_t batch, int64_t cols, int J, int ctas_per_sm, int64_t units) {
  Plan pl;
  pl.N     = N;
  pl.batch = batch;
  int64_t C;
  const int64_t ov = get_chunk_override();
  if (ov > 0) {
    C = ov;
  } else {
    static int64_t env_target = -1;
    if (env_target < 0) {
      const char *s = getenv("C2B_TARGET_UNITS");
      env_target    = s ? atoll(s) : 0;
      if (env_target < 0) env_target = 0;
    }
    int64_t target = 8192;
    if (J > 0 && cols <= 1) {
      // one wave of lane-group CTAs, shared by the series of a batch; the 8192 limit is on the chunks of ONE series
      // (the depth of its tree), not on the product
      const int lanes = J <= 2 ? 2 : (J <= 4 ? 4 : (J <= 8 ? 8 : (J <= 16 ? 16 : 32)));
      target          = (int64_t)sm_count() * ctas_per_sm * (64 / lanes);
    }
    if (env_target > 0) target = env_target;
    if (units > 0) target = units;
    const int64_t have = batch * (cols > 0 ? cols : 1);
    C                  = have >= target ? 1 : target / have;
    // chunking doubles the passes over the rows (summary + apply): two chunks per series buy nothing in the latency-bound
    // regime (half the rows per lane group, twice the passes; r02r: c3 30.4 -> 35.6 ms when its 8,192 series were cut in two)
    if (C == 2) C = 1;
    if (C > 16384 && units <= 0) C = 16384;
    // shortest chunk worth a summary pass; short series are latency-bound (one warp walks a chunk row by row), so they
    // are cut finer (c1: N = 1000)
    static int64_t env_min_rows = -1;
    if (env_min_rows < 0) {
      const char *s = getenv("C2B_MIN_CHUNK_ROWS");
      env_min_rows  = s ? atoll(s) : 0;
      if (env_min_rows < 0) env_min_rows = 0;
    }
    // r02b: c1 (N = 1000) 137 us with 64-row chunks, 88 us with 16-row chunks (one warp per 16 chunks, latency-bound rows)
    // r02g: a rank's share of a split series (N = 125k / 250k): 0.517 / 0.549 ms per evaluation with 64-row chunks, 0.420 /
    // 0.499 ms with 32-row chunks (16 rows: no faster, and grad_dc leaves the 1e-10 band); from 500k on the one-wave
    // target decides and the floor is not reached
    const int64_t pts      = N * batch;
    // (r03g / r03h: c1, N = 1000: 69.9 us with 16-row chunks, 53.5 us with 8-row chunks, no further gain below; at
    // N = 4000 - 5000 the extra scan levels of 8-row chunks push W of the ill-conditioned J = 6 kernel to 1.4e-10, so the
    // 8-row floor is for the really short series only)
    const int64_t min_rows = env_min_rows > 0 ? env_min_rows : (pts <= 2048 ? 8 : (pts <= 16384 ? 16 : (pts <= 400000 ? 32 : 64)));
    const int64_t maxC     = N / min_rows;
    if (C > maxC) C = maxC;
  }
  if (C > N) C = N;
  if (C < 1) C = 1;
  if (C > 32768) C = 32768;
  pl.L = (N + C - 1) / C;
  if (C > 1 && (pl.L & 1)) pl.L += 1;  // even chunk starts keep the 8-byte-per-row arrays 16-byte aligned for bulk copies
  pl.C = (N + pl.L - 1) / pl.L;
  g_pdl_wave = batch > 1 && pl.C > 1;  // (the plan of the op about to be launched on this thread)
  return pl;
}

int Scratch::init(void *user, size_t user_bytes, size_t need, cudaStream_t s) {
  stream = s;
  if (need == 0) return C2B_OK;
  if (user) {
    if (user_bytes < need) {
      set_error("scratch buffer too small: need %zu bytes, got %zu", need, user_bytes);
      return C2B_ERR_SCRATCH_TOO_SMALL;
    }
    ptr   = user;
    bytes = user_bytes;
    return C2B_OK;
  }
  C2B_CUDA(cudaMallocAsync(&ptr, need, s));
  bytes = need;
  owned = true;
  return C2B_OK;
}
Scratch::~Scratch() {
  if (owned && ptr) cudaFreeAsync(ptr, stream);
}

size_t scratch_general_matmul(int64_t M, int J, int K) {
  Arena ar = sizing_arena();
  Plan pl  = make_plan(M, 1, K, 0);
  ScanBuffers sb;
  if (pl.C > 1) take_scan(ar, 1, pl.C, J + J * K, J * K, &sb);
  ar.take<double>((size_t)M * J * K);
  return ar.used;
}
static int asm_rev_ctas(int64_t N) {
  int64_t c = (N + kAsmRevThreads * 4 - 1) / (kAsmRevThreads * 4);
  if (c > 592) c = 592;
  if (c < 1) c = 1;
  return (int)c;
}
size_t scratch_assemble_rev(int64_t N, int Jr, int Jc, int64_t batch) {
  return align256((size_t)batch * asm_rev_ctas(N) * (Jr + 3 * Jc) * sizeof(double));
}
static int red_ctas(int64_t N) {
  int64_t c = (N + kRedThreads * 8 - 1) / (kRedThreads * 8);
  if (c > 592) c = 592;
  if (c < 1) c = 1;
  return (int)c;
}
size_t scratch_norm_logdet(int64_t N, int64_t batch) { return align256((size_t)batch * red_ctas(N) * 2 * sizeof(double)); }

template <bool UPPER>
static int general_matmul_entry(int64_t N, int64_t M, int J, int K, const double *t1, const double *t2, const double *c, const double *U,
                                const double *V, const double *Y, double *Z, double *F, void *scratch, size_t scratch_bytes,
                                void *stream_) {
  int rc = check_common(N, J, 1);
  if (rc) return rc;
  C2B_REQUIRE(M >= 1 && K >= 1, "M and K must be >= 1");
  C2B_REQUIRE(t1 && t2 && c && U && V && Y && Z, "null data pointer");
  cudaStream_t stream = (cudaStream_t)stream_;
  Scratch sc;
  rc = sc.init(scratch, scratch_bytes, scratch_general_matmul(M, J, K), stream);
  if (rc) return rc;
  Arena ar(sc.ptr, sc.bytes);
  Plan pl = make_plan(M, 1, K, 0);
  ScanBuffers sb;
  if (pl.C > 1) take_scan(ar, 1, pl.C, J + J * K, J * K, &sb);
  GmArgs g;
  memset(&g, 0, sizeof(g));
  g.N = N; g.M = M; g.C = pl.C; g.L = pl.L; g.J = J; g.K = K;
  g.t1 = t1; g.t2 = t2; g.c = c; g.U = U; g.V = V; g.Y = Y; g.Z = Z;
  g.H    = ar.take<double>((size_t)M * J * K);
  g.Fout = F;
  unsigned grid, block;
  launch_shape(pl.C * K, &grid, &block);
  // one right-hand side and a compiled width: the running sums come from the lane-group tile kernels (gsweep.cuh)
  bool states_done = false;
  if (K == 1 && !no_group_kernels()) {
    SweepArgs a;
    memset(&a, 0, sizeof(a));
    a.N = M; a.batch = 1; a.C = pl.C; a.L = pl.L; a.J = J; a.K = 1; a.tmin = 0; a.tmax = M - 1;
    a.t = t2; a.c = c; a.B = V; a.A = V; a.Y = Y; a.F = g.H;
    int rg = C2B_ERR_INVALID_ARGUMENT;
#define CALL(JT) rg = launch_gm_states<JT, UPPER>(a, sb, stream)
    C2B_DISPATCH_J_SWEEP(J, CALL)
#undef CALL
    if (rg != C2B_OK && rg != C2B_ERR_INVALID_ARGUMENT) return rg;
    states_done = rg == C2B_OK;
  }
  if (states_done) {
    // fall through to the gather
  } else if (pl.C > 1) {
    g.leaves      = sb.nodes;
    g.node_stride = sb.node_stride;
    C2B_LAUNCH("gm_scan_kernel", stream, launch_k(gm_scan_kernel<UPPER, true>, grid, block, 0, stream, g));
    rc = run_scan<DiagV<0>>(J, K, 1, pl.C, sb.nodes, sb.states, nullptr, stream);
    if (rc) return rc;
    g.carry = sb.states;
  }
  if (!states_done) C2B_LAUNCH("gm_scan_kernel", stream, launch_k(gm_scan_kernel<UPPER, false>, grid, block, 0, stream, g));
  C2B_LAUNCH("gm_gather_kernel", stream, launch_k(gm_gather_kernel<UPPER>, (unsigned)((N * K + 127) / 128), 128, 0, stream, g));
  if (F) {
    C2B_LAUNCH("gm_workspace_kernel", stream, launch_k(gm_workspace_kernel<UPPER>, (unsigned)((M * (int64_t)J * K + 127) / 128), 128, 0, stream, g));
  }
  return C2B_OK;
}

// dense N x N matrix of the semiseparable representation (forward.hpp:19-42): K(n,n) = a_n,
// K(n,m) = K(m,n) = sum_j U(n,j) V(m,j) exp(-c_j (t_n - t_m)) for n > m.  One thread per element of the lower triangle.
static __global__ void __launch_bounds__(256) to_dense_kernel(int64_t N, int J, const double *t, const double *c, const double *a,
                                                              const double *U, const double *V, double *K) {
  pdl_prologue();
  const int64_t total = N * N;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t n = i / N, m = i - n * N;
    if (m > n) continue;
    if (m == n) {
      K[i] = a[n];
      continue;
    }
    const double dt = t[n] - t[m];
    double s        = 0.0;
    for (int j = 0; j < J; ++j) s = fma(U[n * J + j] * V[m * J + j], exp(-c[j] * dt), s);
    K[n * N + m] = s;
    K[m * N + n] = s;
  }
}

}  // namespace c2b

extern "C" {

int c2b_to_dense(int64_t N, int32_t J, const double *t, const double *c, const double *a, const double *U, const double *V, double *K_out,
                 void *stream) {
  c2b::OpScope _op("c2b_to_dense");
  using namespace c2b;
  int rc = check_common(N, J, 1);
  if (rc) return rc;
  C2B_REQUIRE(t && c && a && U && V && K_out, "null data pointer");
  int64_t blocks = (N * N + 255) / 256;
  if (blocks > 148 * 32) blocks = 148 * 32;
  C2B_LAUNCH("to_dense_kernel", (cudaStream_t)stream, (launch_k(to_dense_kernel, (unsigned)blocks, 256, 0, (cudaStream_t)stream, N, J, t, c, a, U, V, K_out)));
  return C2B_OK;
}

const char *c2b_version(void) { return "celerite2_b200 0.1.0 (sm_100a)"; }
const char *c2b_last_error(void) { return c2b::g_err; }

int32_t c2b_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

void c2b_profile_enable(int32_t on) {
  std::lock_guard<std::mutex> lock(c2b::g_prof_mutex);
  for (auto &r : c2b::g_recs) {
    c2b::g_event_pool.push_back(r.e0);
    c2b::g_event_pool.push_back(r.e1);
  }
  c2b::g_recs.clear();
  c2b::g_profile_on = on != 0;
}

const char *c2b_profile_report(void) {
  std::lock_guard<std::mutex> lock(c2b::g_prof_mutex);
  std::map<std::string, std::pair<long, double>> agg;
  std::vector<std::string> order;
  for (auto &r : c2b::g_recs) {
    if (cudaEventSynchronize(r.e1) != cudaSuccess) continue;
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, r.e0, r.e1) != cudaSuccess) continue;
    std::string key = std::string(r.op) + "/" + r.kernel;
    if (!agg.count(key)) order.push_back(key);
    agg[key].first += 1;
    agg[key].second += ms;
  }
  c2b::g_report.clear();
  char buf[256];
  for (auto &k : order) {
    snprintf(buf, sizeof(buf), "%s %ld %.6f\n", k.c_str(), agg[k].first, agg[k].second);
    c2b::g_report += buf;
  }
  return c2b::g_report.c_str();
}

void c2b_testing_set_chunks_per_series(int64_t chunks) { c2b::g_chunk_override = chunks > 0 ? chunks : 0; }

size_t c2b_scratch_bytes(int32_t op, int64_t N, int64_t M, int32_t J, int32_t K, int64_t batch) {
  using namespace c2b;
  if (N < 1 || J < 1 || batch < 1) return 0;
  if (K < 1) K = 1;
  switch (op) {
    case C2B_OP_FACTOR: return scratch_factor(N, J, batch);
    case C2B_OP_FACTOR_REV: return scratch_factor_rev(N, J, batch);
    case C2B_OP_SOLVE_LOWER:
    case C2B_OP_SOLVE_UPPER: return scratch_sweep(true, N, J, K, batch);
    case C2B_OP_MATMUL_LOWER:
    case C2B_OP_MATMUL_UPPER: return scratch_sweep(false, N, J, K, batch);
    case C2B_OP_SOLVE_LOWER_REV:
    case C2B_OP_SOLVE_UPPER_REV: return scratch_sweep_rev(true, N, J, K, batch);
    case C2B_OP_MATMUL_LOWER_REV:
    case C2B_OP_MATMUL_UPPER_REV: return scratch_sweep_rev(false, N, J, K, batch);
    case C2B_OP_GENERAL_MATMUL: return scratch_general_matmul(M > 0 ? M : N, J, K);
    case C2B_OP_ASSEMBLE_REV: return scratch_assemble_rev(N, J, J, batch);  // upper bound: Jr, Jc <= J
    case C2B_OP_PREDICT_VARIANCE: return scratch_predict_variance(N, M > 0 ? M : N);
    default: return 0;
  }
}

int c2b_get_celerite_matrices(int32_t layout, int64_t N, int32_t Jr, int32_t Jc, int64_t batch, const double *ar, const double *cr,
                              const double *ac, const double *bc, const double *cc, const double *dc, const double *x,
                              const double *diag, double *c_out, double *a_out, double *U_out, double *V_out, int32_t flags,
                              void *stream_) {
  c2b::OpScope _op("c2b_get_celerite_matrices");
  using namespace c2b;
  C2B_REQUIRE(Jr >= 0 && Jc >= 0, "negative term count");
  int rc = check_common(N, Jr + 2 * Jc, batch);
  if (rc) return rc;
  C2B_REQUIRE(layout == 0 || layout == 1, "layout must be 0 (blocked) or 1 (interleaved)");
  C2B_REQUIRE(x && diag && a_out && U_out && V_out, "null data pointer");
  C2B_REQUIRE((Jr == 0 || (ar && cr)) && (Jc == 0 || (ac && bc && cc && dc)), "null coefficient pointer");
  AsmArgs g;
  memset(&g, 0, sizeof(g));
  g.N = N; g.batch = batch; g.Jr = Jr; g.Jc = Jc; g.layout = layout;
  g.ar = ar; g.cr = cr; g.ac = ac; g.bc = bc; g.cc = cc; g.dc = dc; g.x = x; g.diag = diag;
  g.c_out = c_out; g.a_out = a_out; g.U_out = U_out; g.V_out = V_out;
  g.x_bs    = (flags & C2B_SHARED_T) ? 0 : N;
  g.co_r_bs = (flags & C2B_SHARED_C) ? 0 : Jr;
  g.co_c_bs = (flags & C2B_SHARED_C) ? 0 : Jc;
  const int64_t total = batch * N;
  const size_t smem = (size_t)2 * kAsmThreads * (Jr + 2 * Jc) * sizeof(double);
  if (smem > 48 * 1024) C2B_CUDA(cudaFuncSetAttribute(assemble_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
  C2B_LAUNCH("assemble_kernel", (cudaStream_t)stream_,
             launch_k(assemble_kernel, (unsigned)((total + kAsmThreads - 1) / kAsmThreads), kAsmThreads, smem, (cudaStream_t)stream_, g));
  return C2B_OK;
}

int c2b_get_celerite_matrices_rev(int32_t layout, int64_t N, int32_t Jr, int32_t Jc, int64_t batch, const double *ar, const double *cr,
                                  const double *ac, const double *bc, const double *cc, const double *dc, const double *x,
                                  const double *b_c, const double *b_a, const double *b_U, const double *b_V, double *b_ar,
                                  double *b_cr, double *b_ac, double *b_bc, double *b_cc, double *b_dc, double *b_x, double *b_diag,
                                  int32_t flags, void *scratch, size_t scratch_bytes, void *stream_) {
  c2b::OpScope _op("c2b_get_celerite_matrices_rev");
  using namespace c2b;
  (void)cr;
  (void)cc;
  C2B_REQUIRE(Jr >= 0 && Jc >= 0, "negative term count");
  int rc = check_common(N, Jr + 2 * Jc, batch);
  if (rc) return rc;
  C2B_REQUIRE(layout == 0 || layout == 1, "layout must be 0 (blocked) or 1 (interleaved)");
  C2B_REQUIRE(x && b_a && b_U && b_V, "null data pointer");
  C2B_REQUIRE((Jr == 0 || (ar && b_ar && b_cr)) && (Jc == 0 || (ac && bc && dc && b_ac && b_bc && b_cc && b_dc)),
              "null coefficient pointer");
  cudaStream_t stream = (cudaStream_t)stream_;
  Scratch sc;
  rc = sc.init(scratch, scratch_bytes, scratch_assemble_rev(N, Jr, Jc, batch), stream);
  if (rc) return rc;
  AsmRevArgs g;
  memset(&g, 0, sizeof(g));
  g.N = N; g.batch = batch; g.Jr = Jr; g.Jc = Jc; g.layout = layout;
  g.ar = ar; g.ac = ac; g.bc = bc; g.dc = dc; g.x = x;
  g.b_c = b_c; g.b_a = b_a; g.b_U = b_U; g.b_V = b_V;
  g.b_ar = b_ar; g.b_cr = b_cr; g.b_ac = b_ac; g.b_bc = b_bc; g.b_cc = b_cc; g.b_dc = b_dc; g.b_x = b_x; g.b_diag = b_diag;
  g.x_bs    = (flags & C2B_SHARED_T) ? 0 : N;
  g.co_r_bs = (flags & C2B_SHARED_C) ? 0 : Jr;
  g.co_c_bs = (flags & C2B_SHARED_C) ? 0 : Jc;
  g.part    = (double *)sc.ptr;
  g.ctas_per_series = asm_rev_ctas(N);
  const int nred    = Jr + 3 * Jc;
  dim3 grid((unsigned)g.ctas_per_series, (unsigned)batch);
  C2B_LAUNCH("assemble_rev_kernel", stream, launch_k(assemble_rev_kernel, grid, kAsmRevThreads, (size_t)nred * kAsmRevThreads * sizeof(double), stream, g));
  C2B_LAUNCH("assemble_rev_final_kernel", stream, launch_k(assemble_rev_final_kernel, (unsigned)batch, 256, 0, stream, g));
  return C2B_OK;
}

static int norm_logdet_impl(int64_t N, int64_t batch, const double *d, const double *z, double *out, double *loglike, double *bz,
                            double *bd, void *scratch, size_t scratch_bytes, void *stream_) {
  using namespace c2b;
  int rc = check_common(N, 1, batch);
  if (rc) return rc;
  C2B_REQUIRE(d && z && out, "null data pointer");
  cudaStream_t stream = (cudaStream_t)stream_;
  Scratch sc;
  rc = sc.init(scratch, scratch_bytes, scratch_norm_logdet(N, batch), stream);
  if (rc) return rc;
  const int ctas = red_ctas(N);
  dim3 grid((unsigned)ctas, (unsigned)batch);
  C2B_LAUNCH("norm_logdet_kernel", stream, launch_k(norm_logdet_kernel, grid, kRedThreads, 0, stream, N, d, z, bz, bd, (double *)sc.ptr, ctas));
  C2B_LAUNCH("norm_logdet_final_kernel", stream,
             launch_k(norm_logdet_final_kernel, (unsigned)batch, 64, 0, stream, (const double *)sc.ptr, ctas, out, loglike,
                      0.5 * (double)N * 1.8378770664093453 /* log(2 pi) */));
  return C2B_OK;
}
int c2b_norm_logdet(int64_t N, int64_t batch, const double *d, const double *z, double *out, double *bz, double *bd, void *scratch,
                    size_t scratch_bytes, void *stream_) {
  c2b::OpScope _op("c2b_norm_logdet");
  return norm_logdet_impl(N, batch, d, z, out, nullptr, bz, bd, scratch, scratch_bytes, stream_);
}
int c2b_norm_logdet_loglike(int64_t N, int64_t batch, const double *d, const double *z, double *out, double *loglike, double *bz,
                            double *bd, void *scratch, size_t scratch_bytes, void *stream_) {
  c2b::OpScope _op("c2b_norm_logdet");
  C2B_REQUIRE(loglike, "null data pointer");
  return norm_logdet_impl(N, batch, d, z, out, loglike, bz, bd, scratch, scratch_bytes, stream_);
}

int c2b_add_inplace(int64_t n, const double *x, double *y, void *stream_) {
  c2b::OpScope _op("c2b_add_inplace");
  using namespace c2b;
  C2B_REQUIRE(n >= 0 && (n == 0 || (x && y)), "bad arguments");
  if (n == 0) return C2B_OK;
  int64_t blocks = (n + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  C2B_LAUNCH("add_inplace_kernel", (cudaStream_t)stream_, launch_k(add_inplace_kernel, (unsigned)blocks, 256, 0, (cudaStream_t)stream_, n, x, y));
  return C2B_OK;
}

void *c2b_device_malloc(size_t bytes) {
  void *p = nullptr;
  if (cudaMalloc(&p, bytes ? bytes : 1) != cudaSuccess) {
    c2b::set_error("c2b_device_malloc: cudaMalloc of %zu bytes failed: %s", bytes, cudaGetErrorString(cudaGetLastError()));
    return nullptr;
  }
  return p;
}
void c2b_device_free(void *ptr) {
  if (ptr) cudaFree(ptr);
}
int c2b_copy_to_device(void *dst_device, const void *src_host, size_t bytes) {
  if (bytes == 0) return C2B_OK;
  C2B_CUDA(cudaMemcpy(dst_device, src_host, bytes, cudaMemcpyHostToDevice));
  return C2B_OK;
}
int c2b_copy_to_host(void *dst_host, const void *src_device, size_t bytes) {
  if (bytes == 0) return C2B_OK;
  C2B_CUDA(cudaMemcpy(dst_host, src_device, bytes, cudaMemcpyDeviceToHost));
  return C2B_OK;
}
int c2b_stream_synchronize(void *stream) {
  C2B_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  return C2B_OK;
}

int c2b_general_matmul_lower(int64_t N, int64_t M, int32_t J, int32_t K, const double *t1, const double *t2, const double *c,
                             const double *U, const double *V, const double *Y, double *Z, double *F, void *scratch,
                             size_t scratch_bytes, void *stream) {
  c2b::OpScope _op("c2b_general_matmul_lower");
  return c2b::general_matmul_entry<false>(N, M, J, K, t1, t2, c, U, V, Y, Z, F, scratch, scratch_bytes, stream);
}
int c2b_general_matmul_upper(int64_t N, int64_t M, int32_t J, int32_t K, const double *t1, const double *t2, const double *c,
                             const double *U, const double *V, const double *Y, double *Z, double *F, void *scratch,
                             size_t scratch_bytes, void *stream) {
  c2b::OpScope _op("c2b_general_matmul_upper");
  return c2b::general_matmul_entry<true>(N, M, J, K, t1, t2, c, U, V, Y, Z, F, scratch, scratch_bytes, stream);
}

}  // extern "C"
