// C ABI: c2b_solve_lower_noalloc -- solve_lower for callers that must not allocate device memory.
//
// Who this is for.  Some hosts of this library are handed every byte of device memory they may touch and have
// no way to ask for more while a call is in flight: an XLA/JAX custom-call handler only sees the buffers the
// compiler planned for it; an inference or sampling loop with a fixed memory budget has given the whole device
// to its own pool; a callback running while the rest of the process is mid-step cannot afford the implicit
// device-wide synchronisation a cudaMalloc/cudaFree can cause.  The ordinary entry point needs a scratch
// buffer for its chunked scan (from the caller, or from cudaMallocAsync when the caller passes NULL).  This
// variant needs neither: the scratch is a fixed array that is part of the library image itself.
//
// Cost and limit.  The array below is C2B_STATIC_SCRATCH_BYTES (8 MiB) of device memory on EVERY GPU that
// loads the library, from load until the process exits, whether or not this entry point is ever called.
// Problems whose scan needs more than that are refused (C2B_ERR_SCRATCH_TOO_SMALL) -- use c2b_solve_lower with a
// caller-owned buffer for those.  8 MiB covers one series of N = 1,000,000 at J = 6, K = 1 (6.3 MB).
//
// Sharing.  All calls on a device use the same array and it keeps whatever the previous call left in it.
// Nothing in it has to start at zero: every word a call reads has been written earlier by that same call
// (chunk summaries, then tree nodes, then carried states), so no clearing pass is needed.  Because the array is
// shared, two calls that overlap on the GPU (different streams, or different host threads) race on it; run
// calls to this entry point on one stream, or otherwise order them yourself.
//
// Set-up.  In host code the name of a __device__ array is only a stand-in; passing it to a kernel would hand
// the kernel a host address.  The first call on each device therefore asks the runtime for the array's real
// device address (cudaGetSymbolAddress, with that device current) and remembers it.  Make that first call
// before starting any stream capture on the device; later calls do no set-up at all.
#include <mutex>

#include "host_common.cuh"

#define C2B_STATIC_SCRATCH_BYTES (8u << 20)

static __device__ __align__(256) unsigned char c2b_static_scratch[C2B_STATIC_SCRATCH_BYTES];

namespace {
constexpr int kMaxDevices = 64;
void *g_static_scratch_addr[kMaxDevices] = {nullptr};  // per device: device address of c2b_static_scratch
std::mutex g_static_scratch_mutex;
}  // namespace

extern "C" {

size_t c2b_noalloc_scratch_limit(void) { return (size_t)C2B_STATIC_SCRATCH_BYTES; }

int c2b_solve_lower_noalloc(int64_t N, int32_t J, int32_t K, int64_t batch, const double *t, const double *c, const double *U,
                            const double *W, const double *Y, double *Z, double *F, int32_t flags, void *stream) {
  using namespace c2b;
  int rc = check_common(N, J, batch);
  if (rc) return rc;
  C2B_REQUIRE(K >= 1, "K (number of right-hand sides) must be >= 1");
  const size_t need = c2b_scratch_bytes(C2B_OP_SOLVE_LOWER, N, 0, J, K, batch);
  if (need > (size_t)C2B_STATIC_SCRATCH_BYTES) {
    set_error("c2b_solve_lower_noalloc: this problem (N=%lld, J=%d, K=%d, batch=%lld) needs %zu bytes of scan scratch but the "
              "built-in array holds %zu; use c2b_solve_lower with a caller-owned scratch buffer",
              (long long)N, (int)J, (int)K, (long long)batch, need, (size_t)C2B_STATIC_SCRATCH_BYTES);
    return C2B_ERR_SCRATCH_TOO_SMALL;
  }
  int dev = -1;
  C2B_CUDA(cudaGetDevice(&dev));
  C2B_REQUIRE(dev >= 0 && dev < kMaxDevices, "device ordinal out of range");
  void *scratch = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_static_scratch_mutex);
    scratch = g_static_scratch_addr[dev];
    if (!scratch) {
      // once per device, with that device current: resolve the device-side address of the array
      C2B_CUDA(cudaGetSymbolAddress((void **)&scratch, c2b_static_scratch));
      g_static_scratch_addr[dev] = scratch;
    }
  }
  // The regular entry point never allocates when it is given a buffer.
  return c2b_solve_lower(N, J, K, batch, t, c, U, W, Y, Z, F, flags, scratch, (size_t)C2B_STATIC_SCRATCH_BYTES, stream);
}

}  // extern "C"
