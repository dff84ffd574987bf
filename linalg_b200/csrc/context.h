// la_b200 — execution context: one CUDA stream plus a device workspace arena.
//
// The reference keeps LAPACK workspaces in a pooled host arena so that any number
// of fibers can call LAPACK concurrently (src/linalg/workpool.cr:6-73,
// spec/linalg_spec.cr:519-539).  The GPU backend has the same need on the device
// side: every exported symbol checks a Context out of a mutex-protected pool,
// runs on that context's stream and bump-allocates scratch from its arena.
#pragma once
#include <functional>
#include <cuda_runtime.h>
#include <cstddef>
#include <cstdint>
#include <vector>

namespace la {

struct ArenaExhausted { size_t bytes; };

struct Arena {
  struct Block { char* base; size_t size; size_t used; };
  std::vector<Block> blocks;
  struct Mark { size_t nblocks; size_t used_last; };
  void* alloc(size_t bytes);       // 256-byte aligned; grows with cudaMalloc on demand
  // fixed = true: the arena consists of caller-provided blocks only and never calls cudaMalloc;
  // a request that does not fit throws ArenaExhausted (caught by the entry point that set it).
  bool fixed = false;
  Mark mark() const;
  void release(Mark m);            // stack discipline
  void free_all();
  size_t total = 0;
};

struct Context {
  int device = 0;
  cudaStream_t stream = nullptr;   // compute stream of this context
  bool owns_stream = false;
  cudaStream_t copy_stream = nullptr;  // H2D overlap in the host-pointer shims
  cudaStream_t d2h_stream = nullptr;   // D2H overlap (the other PCIe direction)
  cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
  Arena arena;
  int sm_count = 0;
  int* d_info = nullptr;           // device-side LAPACK info word (in arena block 0)
  void* pinned = nullptr;          // small pinned bounce buffer (info, ipiv)
  size_t pinned_size = 0;
  void* lu_state = nullptr;        // LU panel exchange area + launch sequence (lu_panel.cu)
  void* stager = nullptr;          // pinned staging ring for pageable host operands (host_stage.cu)
  int sm_limit = 0;                // > 0: persistent kernels use at most this many CTAs
  bool allow_side_stream = true;   // look-ahead may create `side` (a stream, events, an arena)
  Context* side = nullptr;         // second context for work overlapped with the main stream
  // Device-pointer API: the stream the previous call ran on and an event recorded at its end.
  // A call that arrives on another stream first waits for that event, because the scratch of
  // this context (arena, LU exchange area) is reused from call to call.
  cudaStream_t last_stream = nullptr;
  bool has_last = false;
  cudaEvent_t tail_ev = nullptr;
  // Host-pointer shims only: called by a factorisation, in stream order, each time a block of
  // the result has reached its final value (Cholesky: block column / block row [c0, c1)), so
  // that the shim can send it home while the factorisation goes on.
  std::function<void(int64_t c0, int64_t c1)> on_block_final;
  template <typename T> T* alloc(size_t n) { return (T*)arena.alloc(n * sizeof(T)); }
};

// Pool of contexts for the host-pointer (drop-in) symbols.
Context* acquire_context();
void release_context(Context* c);
struct ContextLease {
  Context* c;
  ContextLease() : c(acquire_context()) {}
  ~ContextLease() { release_context(c); }
  Context* operator->() { return c; }
  Context& operator*() { return *c; }
};

// Per-thread context for the device-pointer API, re-targeted at a caller stream.
Context* thread_context(cudaStream_t stream);

void* pinned_scratch(Context& c, size_t bytes);

// Free the device workspace of the calling thread's device-API contexts and of every pooled
// (idle) context; returns the number of bytes given back.  Contexts in use are not touched.
size_t release_workspaces();

// Lazily created companion context (own high-priority stream and arena) for look-ahead.
Context* side_context(Context& c);

}  // namespace la
