// Shared device/host helpers for the celerite2_b200 kernels (sm_100a, FP64 throughout).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <math.h>

#include "../../include/celerite2_b200.h"

#define C2B_DEV __device__ __forceinline__

// Resident CTAs per SM the lane-group kernels (64 threads each) are compiled for, per kernel family: caps their
// registers at 65536 / (64 * n) and sizes the chunk plan of the op (make_plan: exactly one wave of CTAs; every op has its
// own plan, so the families may differ).  r02p profile: at 6 CTAs (3 warps per scheduler) the kernels issue on 40 % of
// the cycles, one instruction per warp every ~7 cycles: latency-bound, more resident warps is the lever.
#ifndef C2B_MINB_SWEEP
#define C2B_MINB_SWEEP 6
#endif
#ifndef C2B_MINB_FACTOR
#define C2B_MINB_FACTOR 6
#endif
#ifndef C2B_MINB_FREV
#define C2B_MINB_FREV 6
#endif
#ifndef C2B_FREV_TILE
#define C2B_FREV_TILE 4  // rows per tile of the factor_rev apply pass (its tiles carry the J x J workspace rows)
#endif

// Compile-time width JT in 1..16 gives register arrays and fully unrolled loops; JT == 0 is the
// run-time-width fallback (J <= C2B_MAX_J, arrays live in local memory).
template <int JT>
struct Width {
  static constexpr int MAXJ = JT > 0 ? JT : C2B_MAX_J;
  static constexpr int UNR  = (JT > 0 && JT <= 8) ? JT : 1;  // unroll factor for loops over J
};
#define C2B_PRAGMA(x) _Pragma(#x)
#define C2B_UNROLL_J C2B_PRAGMA(unroll Width<JT>::UNR)

namespace c2b {

// ---- error reporting (host) ---------------------------------------------------------------------------------
void set_error(const char *fmt, ...);
int cuda_fail(cudaError_t e, const char *what);
#define C2B_CUDA(call)                                                                                                                    \
  do {                                                                                                                                    \
    cudaError_t _e = (call);                                                                                                              \
    if (_e != cudaSuccess) return c2b::cuda_fail(_e, #call);                                                                              \
  } while (0)
#define C2B_CHECK_LAUNCH() C2B_CUDA(cudaGetLastError())
#define C2B_REQUIRE(cond, msg)                                                                                                            \
  do {                                                                                                                                    \
    if (!(cond)) {                                                                                                                        \
      c2b::set_error("%s: %s", __func__, msg);                                                                                            \
      return C2B_ERR_INVALID_ARGUMENT;                                                                                                    \
    }                                                                                                                                     \
  } while (0)

// ---- built-in kernel timing (bench support) -------------------------------------------------------------------
// When enabled (c2b_profile_enable) every kernel launch is bracketed by a pair of CUDA events on its stream;
// c2b_profile_report() turns them into per-(entry point, kernel) totals.
extern bool g_profile_on;
extern thread_local const char *g_op;
void prof_begin(const char *kernel, cudaStream_t s);
void prof_end(cudaStream_t s);
// Every C-ABI entry point opens one of these: it names the op for the built-in profile and, when the environment has
// C2B_NVTX=1, brackets the call with an NVTX range (host-side; shows the entry points above their kernels in a timeline).
void nvtx_push(const char *name);
void nvtx_pop();
struct OpScope {
  const char *prev;
  explicit OpScope(const char *name) : prev(g_op) {
    g_op = name;
    nvtx_push(name);
  }
  ~OpScope() {
    nvtx_pop();
    g_op = prev;
  }
};
// ---- programmatic dependent launch ---------------------------------------------------------------------------
// An evaluation is a chain of ~35 dependent kernels.  Every kernel opens with pdl_prologue(): it lets the NEXT kernel of
// the stream be scheduled right away and then waits until the PREVIOUS kernel has completed and flushed, before it
// touches any memory.  A launch that carries the matching attribute may therefore be set up while its predecessor is
// still running; without the attribute both instructions are no-ops and the launch is an ordinary stream-ordered one.
// Measured (r02t / r02u; attribute on every launch against none): c5 (512 series x 27 chunks) 11.03 -> 10.08 ms, c1 no
// change, c2 (one series, 7,104 chunks) 0.937 -> 0.972 ms, c3 (8,192 whole series, 512 CTAs) 30.4 -> 34.5 ms -- CTAs of
// a wave-filling kernel that become resident early, beside the scan's CTAs or in the slots a 512-CTA grid leaves free,
// end up unevenly spread over the SMs; on the small kernels alone c2 0.948 ms.  Hence the default: the attribute is set
// for the ops of a CHUNKED BATCH only (g_pdl_wave, set where the op's chunk plan is made).  C2B_PDL=0 never, 2 always.
C2B_DEV void pdl_prologue() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");
}
int pdl_mode();
extern thread_local bool g_pdl_wave;
template <typename... KArgs, typename... Args>
static inline cudaError_t launch_k(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, Args &&...args) {
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim          = grid;
  cfg.blockDim         = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream           = stream;
  cudaLaunchAttribute at[1];
  at[0].id                                         = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs                                        = at;
  const int mode                                   = pdl_mode();
  cfg.numAttrs                                     = (mode >= 2 || (mode == 1 && g_pdl_wave)) ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

#define C2B_LAUNCH(name, stream, ...)                                                                                                     \
  do {                                                                                                                                    \
    if (c2b::g_profile_on) c2b::prof_begin(name, stream);                                                                                 \
    __VA_ARGS__;                                                                                                                          \
    if (c2b::g_profile_on) c2b::prof_end(stream);                                                                                         \
    C2B_CHECK_LAUNCH();                                                                                                                   \
  } while (0)

// ---- chunk plan ---------------------------------------------------------------------------------------------
// A series of N points is cut into C chunks of L points (the last one shorter).  Chunks are the unit of
// parallelism along N; carries between chunks come from an associative scan over per-chunk summaries.
struct Plan {
  int64_t N, batch;
  int64_t C;  // chunks per series
  int64_t L;  // points per chunk
};
Plan make_plan(int64_t N, int64_t batch, int64_t cols, int J, int ctas_per_sm = 6, int64_t units = 0);
int64_t env_units(const char *name);  // per-family chunk targets from the environment (A/B runs), 0 when unset
int64_t get_chunk_override();
bool no_group_kernels();

static inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

// Bump allocator over the caller's scratch buffer.
struct Arena {
  char *base;
  size_t size, used;
  bool overflow;
  Arena(void *p, size_t n) : base((char *)p), size(n), used(0), overflow(false) {}
  template <typename T>
  T *take(size_t count) {
    size_t bytes = align256(count * sizeof(T));
    if (used + bytes > size) {
      overflow = true;
      return nullptr;
    }
    T *r = (T *)(base + used);
    used += bytes;
    return r;
  }
};

// RAII scratch: uses the caller's buffer or allocates one on the stream.
struct Scratch {
  void *ptr    = nullptr;
  size_t bytes = 0;
  bool owned   = false;
  cudaStream_t stream;
  int init(void *user, size_t user_bytes, size_t need, cudaStream_t s);
  ~Scratch();
};

}  // namespace c2b

// ---- device helpers -----------------------------------------------------------------------------------------
// p[j] = exp(c[j] * dt), computing each distinct value of c once: complex terms contribute pairs of equal
// c (c++/include/celerite2/terms.hpp:125, python/celerite2/terms.py:172-173).  dup[j] is a bitmask of earlier
// indices i < j with c[i] == c[j] (0 => compute).
template <int JT>
C2B_DEV void dup_masks(int J, const double *c, unsigned *dup) {
  C2B_UNROLL_J
  for (int j = 0; j < J; ++j) {
    unsigned m = 0;
    C2B_UNROLL_J
    for (int i = 0; i < J; ++i)
      if (i < j && c[i] == c[j] && m == 0) m = 1u << i;
    dup[j] = m;
  }
}

template <int JT>
C2B_DEV void compute_p(int J, const double *c, const unsigned *dup, double dt, double *p) {
  C2B_UNROLL_J
  for (int j = 0; j < J; ++j) {
    if (dup[j] == 0) {
      p[j] = exp(c[j] * dt);
    } else {
      double v = 0.0;
      C2B_UNROLL_J
      for (int i = 0; i < J; ++i)
        if (i < j && ((dup[j] >> i) & 1u)) v = p[i];
      p[j] = v;
    }
  }
}

// ---- per-thread software pipeline of input rows -----------------------------------------------------------------
// The recursions walk their chunk row by row with a dependent chain per row, and every row needs a handful of
// small, thread-private loads (t, a, one row of U, V, ...).  Issued at the point of use they expose the full
// DRAM latency three or four times per row (r01 profile: >70 % of the stall samples).  RowPipe issues the loads
// of row i+D with cp.async (LDGSTS, 8 bytes each, no registers held) into a ring of D+1 shared-memory slots
// while row i is being computed.  Layout [slot][field][thread]: lanes read consecutive words, no bank conflicts.
// STAGE = false (run-time widths) degenerates to plain loads into a local array at the top of each row.
C2B_DEV void cp_async8(double *smem_dst, const double *gmem_src) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(s), "l"(gmem_src) : "memory");
}
C2B_DEV void cp_async16(double *smem_dst, const double *gmem_src) {  // both 16-byte aligned
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem_src) : "memory");
}
C2B_DEV void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }

// ---- bulk copies (TMA engine, no tensor map): one instruction of one lane moves a contiguous run between shared and
// global memory, outside the load/store pipe.  Sizes are multiples of 16 bytes, both addresses 16-byte aligned.
// r02h: the lane-group kernels are bound by L1 line-wavefronts (every per-lane 16-byte store of a J x J workspace row
// touches its own 128-byte line: ~20 cycles per row at J = 6); a bulk store costs the pipe nothing.
C2B_DEV void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
C2B_DEV void bulk_store(double *gmem_dst, const double *smem_src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem_dst),
               "r"((uint32_t)__cvta_generic_to_shared(smem_src)), "r"(bytes)
               : "memory");
}
C2B_DEV void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// all but the N most recent bulk groups of this thread have finished READING their shared-memory source
template <int N>
C2B_DEV void bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
// mbarrier-tracked bulk load into shared memory
C2B_DEV void mbar_init(uint64_t *bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(count) : "memory");
}
C2B_DEV void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(bytes) : "memory");
}
C2B_DEV void mbar_wait(uint64_t *bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"((uint32_t)__cvta_generic_to_shared(bar)),
      "r"(parity)
      : "memory");
}
C2B_DEV void bulk_load(double *smem_dst, const double *gmem_src, uint32_t bytes, uint64_t *bar) {
  asm volatile("cp.async.bulk.shared::cta.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   (uint32_t)__cvta_generic_to_shared(smem_dst)),
               "l"(gmem_src), "r"(bytes), "r"((uint32_t)__cvta_generic_to_shared(bar))
               : "memory");
}
template <int N>
C2B_DEV void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

template <bool STAGE, int ROWD, int NSLOT>
struct RowPipe {
  static constexpr int D = STAGE ? NSLOT - 1 : 0;  // rows of look-ahead
  double *base;                                     // shared memory (STAGE)
  double row[STAGE ? 1 : ROWD];                     // direct mode: the current row
  int nt, tid, slot_put, slot_get;
  C2B_DEV void init(double *smem) {
    base     = smem;
    nt       = blockDim.x;
    tid      = threadIdx.x;
    slot_put = 0;
    slot_get = 0;
  }
  // store field f of the row being fetched
  C2B_DEV void put(int f, const double *g) {
    if (STAGE)
      cp_async8(base + ((size_t)(slot_put * ROWD + f) * nt + tid), g);
    else
      row[f] = *g;
  }
  C2B_DEV void put_value(int f, double v) {
    if (STAGE)
      base[(size_t)(slot_put * ROWD + f) * nt + tid] = v;
    else
      row[f] = v;
  }
  // end of one row's fetch (called once per loop trip even when nothing was fetched, to keep group counts uniform)
  C2B_DEV void commit(bool fetched) {
    if (STAGE) {
      cp_async_commit();
      if (fetched) slot_put = (slot_put + 1 == NSLOT) ? 0 : slot_put + 1;
    }
  }
  // make the oldest outstanding row readable
  C2B_DEV void acquire() {
    if (STAGE) cp_async_wait<D>();
  }
  C2B_DEV double get(int f) const {
    if (STAGE) return base[(size_t)(slot_get * ROWD + f) * nt + tid];
    return row[f];
  }
  C2B_DEV void release() {
    if (STAGE) slot_get = (slot_get + 1 == NSLOT) ? 0 : slot_get + 1;
  }
};

// Store or accumulate an output element.  Every element is touched by exactly one thread per launch, so the
// accumulate form is a fire-and-forget reduction (no value returned, nothing waits for it).
C2B_DEV void out_store(double *p, double v, bool acc) {
  if (acc)
    atomicAdd(p, v);
  else
    *p = v;
}

// First failing row of factor(): keep the smallest positive index (0 = success).
C2B_DEV void record_failure(int64_t *status, int64_t n) {
  if (!status) return;
  unsigned long long *s   = (unsigned long long *)status;
  unsigned long long old = atomicCAS(s, 0ull, (unsigned long long)n);
  while (old != 0ull && (unsigned long long)n < old) {
    unsigned long long prev = atomicCAS(s, old, (unsigned long long)n);
    if (prev == old) break;
    old = prev;
  }
}
