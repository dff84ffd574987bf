// la_b200 — context pool and device workspace arena (see context.h).
#include "common.cuh"
#include "context.h"
#include <mutex>
#include <vector>

namespace la {

static constexpr size_t kAlign = 256;
static constexpr size_t kMinBlock = size_t(64) << 20;

void* Arena::alloc(size_t bytes) {
  bytes = (size_t)round_up((int64_t)(bytes ? bytes : 1), (int64_t)kAlign);
  if (!blocks.empty()) {
    Block& b = blocks.back();
    if (b.used + bytes <= b.size) {
      void* p = b.base + b.used;
      b.used += bytes;
      return p;
    }
  }
  if (fixed) throw ArenaExhausted{bytes};
  size_t sz = bytes > kMinBlock ? bytes : kMinBlock;
  if (!blocks.empty() && blocks.back().size * 2 > sz && bytes <= blocks.back().size * 2)
    sz = blocks.back().size * 2;
  if (sz > (size_t(8) << 30) && bytes <= (size_t(8) << 30)) sz = size_t(8) << 30;
  if (sz < bytes) sz = bytes;
  char* p = nullptr;
  cudaError_t e = cudaMalloc(&p, sz);
  if (e != cudaSuccess && sz > bytes) { sz = bytes; e = cudaMalloc(&p, sz); }
  if (e != cudaSuccess) {
    fprintf(stderr, "la_b200: device workspace allocation of %zu bytes failed: %s\n", sz,
            cudaGetErrorString(e));
    abort();
  }
  total += sz;
  blocks.push_back(Block{p, sz, bytes});
  return p;
}

Arena::Mark Arena::mark() const {
  return Mark{blocks.size(), blocks.empty() ? 0 : blocks.back().used};
}

void Arena::release(Mark m) {
  // Blocks added after the mark stay allocated (they are reused by later calls) but
  // are emptied; the block that was last at mark time is rewound.
  for (size_t i = m.nblocks; i < blocks.size(); i++) blocks[i].used = 0;
  if (m.nblocks > 0) blocks[m.nblocks - 1].used = m.used_last;
  // If several blocks exist and everything is released, coalesce into one large
  // block next time by dropping the small ones.
  if (m.nblocks == 0 && blocks.size() > 1 && !fixed) {
    size_t want = 0;
    for (auto& b : blocks) want += b.size;
    for (auto& b : blocks) cudaFree(b.base);
    blocks.clear();
    total = 0;
    char* p = nullptr;
    if (cudaMalloc(&p, want) == cudaSuccess) {
      blocks.push_back(Block{p, want, 0});
      total = want;
    } else {
      cudaGetLastError();
    }
  }
}

void Arena::free_all() {
  if (fixed) return;               // caller-provided blocks
  for (auto& b : blocks) cudaFree(b.base);
  blocks.clear();
  total = 0;
}

static void init_context(Context* c, bool make_stream) {
  LA_CUDA_CHECK(cudaGetDevice(&c->device));
  cudaDeviceProp prop;
  LA_CUDA_CHECK(cudaGetDeviceProperties(&prop, c->device));
  if (prop.major != 10) {
    fprintf(stderr, "la_b200: this library is built for sm_100a (B200) only; found sm_%d%d\n",
            prop.major, prop.minor);
    abort();
  }
  c->sm_count = prop.multiProcessorCount;
  if (make_stream) {
    LA_CUDA_CHECK(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    c->owns_stream = true;
  }
  LA_CUDA_CHECK(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
  LA_CUDA_CHECK(cudaStreamCreateWithFlags(&c->d2h_stream, cudaStreamNonBlocking));
  for (auto& e : c->ev) LA_CUDA_CHECK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  LA_CUDA_CHECK(cudaMalloc(&c->d_info, 256));
  LA_CUDA_CHECK(cudaMemset(c->d_info, 0, 256));
}

static std::mutex g_mu;
static std::vector<Context*> g_free;

Context* acquire_context() {
  {
    std::lock_guard<std::mutex> lk(g_mu);
    int dev = 0;
    LA_CUDA_CHECK(cudaGetDevice(&dev));
    for (size_t i = 0; i < g_free.size(); i++) {
      if (g_free[i]->device == dev) {
        Context* c = g_free[i];
        g_free.erase(g_free.begin() + i);
        return c;
      }
    }
  }
  Context* c = new Context();
  init_context(c, true);
  return c;
}

void release_context(Context* c) {
  std::lock_guard<std::mutex> lk(g_mu);
  g_free.push_back(c);
}

static thread_local std::vector<Context*> per_device;

Context* thread_context(cudaStream_t stream) {
  int dev = 0;
  LA_CUDA_CHECK(cudaGetDevice(&dev));
  if ((int)per_device.size() <= dev) per_device.resize(dev + 1, nullptr);
  if (!per_device[dev]) {
    per_device[dev] = new Context();
    init_context(per_device[dev], false);
    LA_CUDA_CHECK(cudaEventCreateWithFlags(&per_device[dev]->tail_ev, cudaEventDisableTiming));
  }
  per_device[dev]->stream = stream;
  return per_device[dev];
}

size_t release_workspaces() {
  size_t freed = 0;
  auto drop = [&](Context* c) {
    if (!c) return;
    int cur = 0;
    LA_CUDA_CHECK(cudaGetDevice(&cur));
    if (cur != c->device) LA_CUDA_CHECK(cudaSetDevice(c->device));
    LA_CUDA_CHECK(cudaDeviceSynchronize());
    freed += c->arena.total;
    c->arena.free_all();
    if (c->side) { freed += c->side->arena.total; c->side->arena.free_all(); }
    if (cur != c->device) LA_CUDA_CHECK(cudaSetDevice(cur));
  };
  for (Context* c : per_device) drop(c);
  std::lock_guard<std::mutex> lk(g_mu);
  for (Context* c : g_free) drop(c);
  return freed;
}

Context* side_context(Context& c) {
  if (!c.side) {
    Context* s = new Context();
    s->device = c.device;
    s->sm_count = c.sm_count;
    s->allow_side_stream = false;
    int lo = 0, hi = 0;
    LA_CUDA_CHECK(cudaDeviceGetStreamPriorityRange(&lo, &hi));
    LA_CUDA_CHECK(cudaStreamCreateWithPriority(&s->stream, cudaStreamNonBlocking, hi));
    s->owns_stream = true;
    for (auto& e : s->ev) LA_CUDA_CHECK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    c.side = s;
  }
  if (!c.ev[0])
    for (auto& e : c.ev) LA_CUDA_CHECK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  return c.side;
}

void* pinned_scratch(Context& c, size_t bytes) {
  if (bytes > c.pinned_size) {
    if (c.pinned) cudaFreeHost(c.pinned);
    size_t sz = bytes < 4096 ? 4096 : bytes;
    LA_CUDA_CHECK(cudaMallocHost(&c.pinned, sz));
    c.pinned_size = sz;
  }
  return c.pinned;
}

}  // namespace la
