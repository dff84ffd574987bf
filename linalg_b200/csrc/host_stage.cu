// la_b200 — host <-> device transfers of the drop-in symbols when the caller's buffers are
// PAGEABLE (the normal case for the reference: Crystal allocates matrices on its GC heap,
// src/matrix/general_matrix.cr:26, and hands raw slices to the library).
//
// A cudaMemcpy from pageable memory is staged by the driver through one thread and one small
// buffer; it moves a fraction of what the PCIe link can and it blocks the calling thread, so the
// three-stream overlap of the host-pointer GEMM / factorisations (shim_cblas.cu, shim_lapack.cu)
// collapses.  Here large pageable transfers go through a ring of pinned slots instead:
//   * host -> device: a few worker threads copy column ranges of the caller's matrix into a free
//     slot in parallel, the slot is then sent with one cudaMemcpy2DAsync on the caller's stream;
//     the caller blocks only while the slots are being filled, never for the DMA;
//   * device -> host: a drain thread per context receives (event, source, destination) requests,
//     issues the device -> slot copies on its own stream once the event has fired and lets the
//     workers copy landed slots into the caller's matrix; `drain()` waits for the queue to empty.
// Pinned or registered host memory, device pointers and small transfers take the direct path.
#include "shim_common.h"
#include <atomic>
#include <condition_variable>
#include <deque>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

namespace la {

namespace {

constexpr size_t kSlotBytes = size_t(16) << 20;
constexpr int kSlots = 6;                           // per direction and context: 96 MiB pinned
constexpr size_t kMinStaged = size_t(4) << 20;      // smaller transfers go direct

// ---- a small fork/join pool for the memcpy work -----------------------------------------------
class CopyPool {
 public:
  explicit CopyPool(int nthreads) {
    for (int i = 0; i < nthreads; i++) threads_.emplace_back([this] { loop(); });
  }
  // run fn(0..n-1) on the workers and the caller; returns when all are done
  void parallel(int n, const std::function<void(int)>& fn) {
    if (n <= 0) return;
    std::atomic<int> next{0}, done{0};
    auto body = [&] {
      for (;;) {
        const int i = next.fetch_add(1);
        if (i >= n) break;
        fn(i);
        done.fetch_add(1);
      }
    };
    const int helpers = n - 1 < (int)threads_.size() ? n - 1 : (int)threads_.size();
    std::atomic<int> left{helpers};
    {
      std::lock_guard<std::mutex> lk(mu_);
      for (int h = 0; h < helpers; h++)
        jobs_.push_back([&] { body(); left.fetch_sub(1); });
    }
    cv_.notify_all();
    body();
    while (done.load() < n || left.load() > 0) std::this_thread::yield();
  }

 private:
  void loop() {
    for (;;) {
      std::function<void()> job;
      {
        std::unique_lock<std::mutex> lk(mu_);
        cv_.wait(lk, [this] { return !jobs_.empty(); });
        job = std::move(jobs_.front());
        jobs_.pop_front();
      }
      job();
    }
  }
  std::vector<std::thread> threads_;
  std::mutex mu_;
  std::condition_variable cv_;
  std::deque<std::function<void()>> jobs_;
};

int pool_threads() {
  static const int n = [] {
    if (const char* e = getenv("LA_B200_STAGE_THREADS")) return atoi(e) > 0 ? atoi(e) : 1;
    const unsigned hw = std::thread::hardware_concurrency();
    int t = hw >= 16 ? 6 : (hw >= 8 ? 4 : 2);
    return t;
  }();
  return n;
}
// pools are process-wide and never destroyed (their threads sleep on a condition variable)
CopyPool& up_pool() { static CopyPool* p = new CopyPool(pool_threads() - 1); return *p; }
CopyPool& down_pool() { static CopyPool* p = new CopyPool(pool_threads() - 1); return *p; }

// columns [c0, c1) of a pitched matrix: dst/src pitches differ, `width` bytes per column
void copy_cols(char* dst, size_t dpitch, const char* src, size_t spitch, size_t width, size_t c0,
               size_t c1) {
  if (dpitch == width && spitch == width) { memcpy(dst + c0 * width, src + c0 * width, (c1 - c0) * width); return; }
  for (size_t c = c0; c < c1; c++) memcpy(dst + c * dpitch, src + c * spitch, width);
}
void parallel_copy(CopyPool& pool, char* dst, size_t dpitch, const char* src, size_t spitch,
                   size_t width, size_t ncols) {
  const size_t bytes = width * ncols;
  int parts = (int)(bytes / (size_t(1) << 20));
  if (parts > pool_threads() * 2) parts = pool_threads() * 2;
  if (parts < 1) parts = 1;
  if ((size_t)parts > ncols) parts = (int)ncols;
  if (ncols == 1 && width >= (size_t(2) << 20)) {   // one long column: split it by bytes
    parts = pool_threads();
    pool.parallel(parts, [&](int i) {
      const size_t b0 = width * i / parts, b1 = width * (i + 1) / parts;
      memcpy(dst + b0, src + b0, b1 - b0);
    });
    return;
  }
  pool.parallel(parts, [&](int i) {
    copy_cols(dst, dpitch, src, spitch, width, ncols * i / parts, ncols * (i + 1) / parts);
  });
}

struct Slot {
  char* buf = nullptr;
  cudaEvent_t ev = nullptr;
  bool busy = false;
};

bool staging_enabled() {
  static const bool off = getenv("LA_B200_NO_STAGING") != nullptr;
  return !off;
}

bool is_pageable(const void* p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return true; }
  return a.type == cudaMemoryTypeUnregistered;
}

}  // namespace

struct HostStager {
  int device = 0;
  Slot up[kSlots], down[kSlots];
  int up_next = 0;
  // device -> host requests, served by the drain thread
  struct Request {
    cudaEvent_t ready;
    char* dst; size_t dpitch; const char* src; size_t spitch; size_t width, ncols;
  };
  std::mutex mu;
  std::condition_variable cv, cv_idle;
  std::deque<Request> queue;
  int in_flight = 0;
  std::vector<cudaEvent_t> free_events;
  cudaStream_t down_stream = nullptr;
  std::thread drain_thread;

  explicit HostStager(int dev) : device(dev) {
    for (auto& s : up) {
      LA_CUDA_CHECK(cudaMallocHost(&s.buf, kSlotBytes));
      LA_CUDA_CHECK(cudaEventCreateWithFlags(&s.ev, cudaEventDisableTiming));
    }
    for (auto& s : down) {
      LA_CUDA_CHECK(cudaMallocHost(&s.buf, kSlotBytes));
      LA_CUDA_CHECK(cudaEventCreateWithFlags(&s.ev, cudaEventDisableTiming));
    }
    LA_CUDA_CHECK(cudaStreamCreateWithFlags(&down_stream, cudaStreamNonBlocking));
    drain_thread = std::thread([this] { drain_loop(); });
    drain_thread.detach();
  }

  // columns per slot for a transfer of `width` bytes per column
  static size_t cols_per_slot(size_t width) { return width >= kSlotBytes ? 1 : kSlotBytes / width; }

  void upload(void* dst, size_t dpitch, const void* src, size_t spitch, size_t width, size_t ncols,
              cudaStream_t s) {
    const size_t cps = cols_per_slot(width);
    for (size_t c0 = 0; c0 < ncols; c0 += cps) {
      const size_t nc = cps < ncols - c0 ? cps : ncols - c0;
      // a column wider than a slot travels in pieces
      for (size_t b0 = 0; b0 < width; b0 += kSlotBytes) {
        const size_t w = kSlotBytes < width - b0 ? kSlotBytes : width - b0;
        Slot& sl = up[up_next];
        up_next = (up_next + 1) % kSlots;
        if (sl.busy) LA_CUDA_CHECK(cudaEventSynchronize(sl.ev));
        parallel_copy(up_pool(), sl.buf, w, (const char*)src + c0 * spitch + b0, spitch, w, nc);
        LA_CUDA_CHECK(cudaMemcpy2DAsync((char*)dst + c0 * dpitch + b0, dpitch, sl.buf, w, w, nc,
                                        cudaMemcpyHostToDevice, s));
        LA_CUDA_CHECK(cudaEventRecord(sl.ev, s));
        sl.busy = true;
      }
    }
  }

  void download(void* dst, size_t dpitch, const void* src, size_t spitch, size_t width,
                size_t ncols, cudaStream_t s) {
    Request r;
    {
      std::lock_guard<std::mutex> lk(mu);
      if (free_events.empty()) {
        cudaEvent_t e;
        LA_CUDA_CHECK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        free_events.push_back(e);
      }
      r.ready = free_events.back();
      free_events.pop_back();
    }
    LA_CUDA_CHECK(cudaEventRecord(r.ready, s));
    r.dst = (char*)dst; r.dpitch = dpitch; r.src = (const char*)src; r.spitch = spitch;
    r.width = width; r.ncols = ncols;
    {
      std::lock_guard<std::mutex> lk(mu);
      queue.push_back(r);
      in_flight++;
    }
    cv.notify_one();
  }

  void drain() {
    std::unique_lock<std::mutex> lk(mu);
    cv_idle.wait(lk, [this] { return in_flight == 0; });
  }

  void drain_loop() {
    LA_CUDA_CHECK(cudaSetDevice(device));
    for (;;) {
      Request r;
      {
        std::unique_lock<std::mutex> lk(mu);
        cv.wait(lk, [this] { return !queue.empty(); });
        r = queue.front();
        queue.pop_front();
      }
      LA_CUDA_CHECK(cudaStreamWaitEvent(down_stream, r.ready, 0));
      // pieces of the transfer, each at most one slot; the DMA of piece i+1.. runs while piece i
      // is copied out to the caller's memory
      struct Piece { size_t c0, nc, b0, w; };
      std::vector<Piece> pieces;
      const size_t cps = cols_per_slot(r.width);
      for (size_t c0 = 0; c0 < r.ncols; c0 += cps)
        for (size_t b0 = 0; b0 < r.width; b0 += kSlotBytes)
          pieces.push_back(Piece{c0, cps < r.ncols - c0 ? cps : r.ncols - c0, b0,
                                 kSlotBytes < r.width - b0 ? kSlotBytes : r.width - b0});
      size_t issued = 0;
      for (size_t done = 0; done < pieces.size(); done++) {
        while (issued < pieces.size() && issued < done + kSlots) {
          const Piece& p = pieces[issued];
          Slot& sl = down[issued % kSlots];
          LA_CUDA_CHECK(cudaMemcpy2DAsync(sl.buf, p.w, r.src + p.c0 * r.spitch + p.b0, r.spitch,
                                          p.w, p.nc, cudaMemcpyDeviceToHost, down_stream));
          LA_CUDA_CHECK(cudaEventRecord(sl.ev, down_stream));
          issued++;
        }
        const Piece& p = pieces[done];
        Slot& sl = down[done % kSlots];
        LA_CUDA_CHECK(cudaEventSynchronize(sl.ev));
        parallel_copy(down_pool(), r.dst + p.c0 * r.dpitch + p.b0, r.dpitch, sl.buf, p.w, p.w, p.nc);
      }
      {
        std::lock_guard<std::mutex> lk(mu);
        free_events.push_back(r.ready);
        in_flight--;
        if (in_flight == 0) cv_idle.notify_all();
      }
    }
  }
};

static HostStager* stager(Context& ctx) {
  if (!ctx.stager) ctx.stager = new HostStager(ctx.device);
  return (HostStager*)ctx.stager;
}

void h2d_2d(Context& ctx, void* dst, size_t dpitch, const void* src, size_t spitch, size_t width,
            size_t ncols, cudaStream_t s) {
  if (width == 0 || ncols == 0) return;
  if (staging_enabled() && width * ncols >= kMinStaged && is_pageable(src)) {
    stager(ctx)->upload(dst, dpitch, src, spitch, width, ncols, s);
    return;
  }
  LA_CUDA_CHECK(cudaMemcpy2DAsync(dst, dpitch, src, spitch, width, ncols, cudaMemcpyHostToDevice, s));
}

void d2h_2d(Context& ctx, void* dst, size_t dpitch, const void* src, size_t spitch, size_t width,
            size_t ncols, cudaStream_t s) {
  if (width == 0 || ncols == 0) return;
  if (staging_enabled() && width * ncols >= kMinStaged && is_pageable(dst)) {
    stager(ctx)->download(dst, dpitch, src, spitch, width, ncols, s);
    return;
  }
  LA_CUDA_CHECK(cudaMemcpy2DAsync(dst, dpitch, src, spitch, width, ncols, cudaMemcpyDeviceToHost, s));
}

void drain_host_transfers(Context& ctx) {
  if (ctx.stager) ((HostStager*)ctx.stager)->drain();
}

}  // namespace la
