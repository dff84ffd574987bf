// la_b200 — la_b200_dgetrf_dev_noalloc: the device-resident FP64 LU factorisation for callers
// that must not allocate.
//
// Who needs this.  la_b200_dgetrf_dev sizes a device workspace with cudaMalloc the first time
// it runs on a thread (pivot-exchange area of the panel kernel, index lists of the row
// interchanges).  Some hosts cannot tolerate that: a solver embedded in a service whose GPU
// memory was carved up in advance by another allocator and where any driver allocation may
// fail or stall; code that runs between a framework's "no allocation from here on" fence
// (for example while its caching allocator is frozen for a measurement, or inside a
// latency-critical loop where cudaMalloc's implicit device synchronisation is not acceptable);
// and processes that run under a memory cgroup/MIG slice sized exactly for their operands.
// For them this entry point does the same factorisation with ALL scratch taken from one
// fixed array that is part of the library image.
//
// What it costs.  The array below (LA_B200_NOALLOC_SCRATCH_BYTES = 384 KiB) is resident on
// every GPU on which the library is loaded, from load until the process exits, whether or not
// this entry point is ever called.
//
// Rules for callers.
//  * m and n must not exceed LA_B200_NOALLOC_MAX_DIM (32768); larger problems are refused with
//    return code -100 and a message on stderr — use la_b200_dgetrf_dev for those.
//  * The first call on each device performs a one-time set-up (it looks up the device address
//    of the array and sets kernel attributes).  Make that first call before starting any
//    stream capture.
//  * All calls share the one array and it keeps what the previous call left in it.  Calls are
//    safe back to back on one stream; two calls that can overlap on the GPU (different streams
//    or host threads without ordering between them) race on the scratch and give wrong results.
//  * The caller has nothing to clear between calls: each call first wipes the pivot-exchange
//    part of the array on its stream (packets left by the previous call would otherwise carry
//    tags the new call uses again), and the interchange lists are fully rewritten before they
//    are read.
//  * The array is all the scratch there is.  Should a code path ever ask for more than fits, the
//    call returns -102 with a message instead of falling back to cudaMalloc.
#include "shim_common.h"
#include <mutex>

namespace la {
// lu_panel.cu
size_t lu_panel_exchange_bytes();
size_t lu_panel_state_bytes();
void lu_panel_adopt_exchange(Context& ctx, void* device_mem, void* state_storage);
}  // namespace la

#define LA_B200_NOALLOC_SCRATCH_BYTES (384 * 1024)
#define LA_B200_NOALLOC_MAX_DIM 32768

// The scratch itself.  In host code this name is only a handle for cudaGetSymbolAddress; the
// device address obtained from it is what the kernels are given.
static __device__ __align__(256) unsigned char g_noalloc_scratch[LA_B200_NOALLOC_SCRATCH_BYTES];

namespace {

constexpr int kMaxDevices = 64;

struct NoAllocDevice {
  bool ready = false;
  la::Context ctx;                       // never passed to init_context(): owns no CUDA objects
  alignas(16) unsigned char panel_state[64];
};

NoAllocDevice g_dev[kMaxDevices];
std::mutex g_mu;

// One-time set-up for the current device: resolve the array's device address and lay it out as
// [ pivot-exchange area | arena for the interchange lists ].
NoAllocDevice* setup_current_device() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return nullptr;
  std::lock_guard<std::mutex> lk(g_mu);
  NoAllocDevice& d = g_dev[dev];
  if (d.ready) return &d;
  unsigned char* base = nullptr;
  cudaError_t e = cudaGetSymbolAddress((void**)&base, g_noalloc_scratch);
  if (e != cudaSuccess || base == nullptr) {
    fprintf(stderr, "la_b200: dgetrf_dev_noalloc: cudaGetSymbolAddress failed: %s\n",
            cudaGetErrorString(e));
    return nullptr;
  }
  int sms = 0;
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return nullptr;
  const size_t xch = (size_t)la::round_up((int64_t)la::lu_panel_exchange_bytes(), 256);
  if (xch + 32 * 1024 > LA_B200_NOALLOC_SCRATCH_BYTES ||
      la::lu_panel_state_bytes() > sizeof(d.panel_state)) {
    fprintf(stderr, "la_b200: dgetrf_dev_noalloc: internal scratch layout does not fit\n");
    return nullptr;
  }
  d.ctx.device = dev;
  d.ctx.allow_side_stream = false;   // look-ahead would create a stream and an arena
  d.ctx.sm_count = sms;
  la::lu_panel_adopt_exchange(d.ctx, base, d.panel_state);
  // the rest of the array becomes the (only) block of this context's bump arena
  d.ctx.arena.fixed = true;          // never grows: a request that does not fit is an error
  d.ctx.arena.blocks.reserve(1);
  d.ctx.arena.blocks.push_back(
      la::Arena::Block{(char*)base + xch, LA_B200_NOALLOC_SCRATCH_BYTES - xch, 0});
  d.ready = true;
  return &d;
}

}  // namespace

extern "C" int la_b200_dgetrf_dev_noalloc(int64_t m, int64_t n, double* a, int64_t lda, int* ipiv,
                                          int* info, void* stream) {
  if (m < 0) return -1;
  if (n < 0) return -2;
  if (lda < (m > 1 ? m : 1)) return -4;
  if (m > LA_B200_NOALLOC_MAX_DIM || n > LA_B200_NOALLOC_MAX_DIM) {
    fprintf(stderr,
            "la_b200: dgetrf_dev_noalloc: %lld x %lld exceeds the no-allocation limit of %d rows/"
            "columns; use la_b200_dgetrf_dev\n",
            (long long)m, (long long)n, LA_B200_NOALLOC_MAX_DIM);
    return -100;
  }
  NoAllocDevice* d = setup_current_device();
  if (!d) return -101;
  d->ctx.stream = (cudaStream_t)stream;
  la::Arena::Mark mk = d->ctx.arena.mark();
  int rc = 0;
  try {
    la::getrf<double>(d->ctx, m, n, a, lda, ipiv, info);
  } catch (const la::ArenaExhausted& e) {
    fprintf(stderr,
            "la_b200: dgetrf_dev_noalloc: a scratch request of %zu bytes does not fit the fixed "
            "%d-byte array; nothing was allocated, the factorisation is incomplete\n",
            e.bytes, LA_B200_NOALLOC_SCRATCH_BYTES);
    rc = -102;
  }
  d->ctx.arena.release(mk);
  return rc;
}
