// la_b200 — helpers shared by the C-ABI translation units (host-pointer drop-in symbols
// and the device-pointer API).
#pragma once
#include "la_internal.h"
#include <cstring>

namespace la {

// host_stage.cu: 2-D copies between the caller's host matrix and the device (`width` bytes per
// column, `ncols` columns).  Large transfers from / to PAGEABLE memory go through the context's
// ring of pinned slots (parallel memcpy + DMA); everything else is a plain cudaMemcpy2DAsync.
// Staged device -> host copies complete asynchronously: drain_host_transfers() (called by sync())
// returns once they have all landed in the caller's memory.
void h2d_2d(Context& ctx, void* dst, size_t dpitch, const void* src, size_t spitch, size_t width,
            size_t ncols, cudaStream_t s);
void d2h_2d(Context& ctx, void* dst, size_t dpitch, const void* src, size_t spitch, size_t width,
            size_t ncols, cudaStream_t s);
void drain_host_transfers(Context& ctx);

// A matrix staged on the device for one host-pointer call.
template <typename T>
struct DevMat {
  T* p = nullptr;
  int64_t ld = 0;
  int64_t rows = 0, cols = 0;
};

template <typename T>
inline DevMat<T> dev_alloc(Context& ctx, int64_t rows, int64_t cols) {
  DevMat<T> d;
  d.rows = rows;
  d.cols = cols;
  d.ld = round_up(rows > 0 ? rows : 1, 16);  // keeps every column 128-byte aligned (TMA)
  d.p = ctx.alloc<T>((size_t)d.ld * (size_t)(cols > 0 ? cols : 1));
  return d;
}

template <typename T>
inline void upload(Context& ctx, DevMat<T>& d, const T* h, int64_t ldh, cudaStream_t s = nullptr) {
  if (d.rows <= 0 || d.cols <= 0) return;
  h2d_2d(ctx, d.p, (size_t)d.ld * sizeof(T), h, (size_t)ldh * sizeof(T),
         (size_t)d.rows * sizeof(T), (size_t)d.cols, s ? s : ctx.stream);
}

template <typename T>
inline void download(Context& ctx, const DevMat<T>& d, T* h, int64_t ldh, cudaStream_t s = nullptr) {
  if (d.rows <= 0 || d.cols <= 0) return;
  d2h_2d(ctx, h, (size_t)ldh * sizeof(T), d.p, (size_t)d.ld * sizeof(T),
         (size_t)d.rows * sizeof(T), (size_t)d.cols, s ? s : ctx.stream);
}

template <typename T>
inline DevMat<T> dev_from_host(Context& ctx, int64_t rows, int64_t cols, const T* h, int64_t ldh) {
  DevMat<T> d = dev_alloc<T>(ctx, rows, cols);
  upload<T>(ctx, d, h, ldh);
  return d;
}

// Transfer only the named triangle of an n x n matrix (block columns of `bw` columns, each as
// one 2D copy of its trapezoid): ?potrf / ?posv neither read nor write the other triangle
// (cholesky.cr:50-64 relies on that), so half of the PCIe traffic can stay at home.
template <typename T>
inline void copy_triangle_hd(Context& ctx, bool to_device, bool lower, int64_t n, T* dev,
                             int64_t ldd, T* host, int64_t ldh, int64_t bw,
                             cudaStream_t s = nullptr) {
  // (diagonal blocks of width bw travel whole, so the other triangle inside them makes a
  // round trip; bw must therefore match the block width the factorisation reports)
  if (bw <= 0 || bw > n) bw = n;
  for (int64_t c0 = 0; c0 < n; c0 += bw) {
    const int64_t w = bw < n - c0 ? bw : n - c0;
    const int64_t r0 = lower ? c0 : 0;
    const int64_t rows = lower ? n - c0 : c0 + w;
    T* d = dev + r0 + c0 * ldd;
    T* h = host + r0 + c0 * ldh;
    if (to_device)
      h2d_2d(ctx, d, (size_t)ldd * sizeof(T), h, (size_t)ldh * sizeof(T), (size_t)rows * sizeof(T),
             (size_t)w, s ? s : ctx.stream);
    else
      d2h_2d(ctx, h, (size_t)ldh * sizeof(T), d, (size_t)ldd * sizeof(T), (size_t)rows * sizeof(T),
             (size_t)w, s ? s : ctx.stream);
  }
}

inline void sync(Context& ctx) {
  drain_host_transfers(ctx);
  LA_CUDA_CHECK(cudaStreamSynchronize(ctx.stream));
}

// Read the device info word after the stream has drained.
inline int read_info(Context& ctx, const int* d_info) {
  int* h = (int*)pinned_scratch(ctx, sizeof(int));
  LA_CUDA_CHECK(cudaMemcpyAsync(h, d_info, sizeof(int), cudaMemcpyDeviceToHost, ctx.stream));
  sync(ctx);
  return *h;
}

inline bool valid_op(int t) { return t == OpN || t == OpT || t == OpC; }
inline bool valid_uplo(int u) { return u == Upper || u == Lower; }
inline bool valid_side(int s) { return s == Left || s == Right; }
inline bool valid_diag(int d) { return d == Unit || d == NonUnit; }

inline int op_from_char(char c) {
  switch (c) {
    case 'N': case 'n': return OpN;
    case 'T': case 't': return OpT;
    case 'C': case 'c': return OpC;
    default: return 0;
  }
}
inline int uplo_from_char(char c) {
  switch (c) {
    case 'U': case 'u': return Upper;
    case 'L': case 'l': return Lower;
    default: return 0;
  }
}
inline int diag_from_char(char c) {
  switch (c) {
    case 'N': case 'n': return NonUnit;
    case 'U': case 'u': return Unit;
    default: return 0;
  }
}
inline int side_from_char(char c) {
  switch (c) {
    case 'L': case 'l': return Left;
    case 'R': case 'r': return Right;
    default: return 0;
  }
}

constexpr int LA_ROW_MAJOR_ = 101, LA_COL_MAJOR_ = 102;

}  // namespace la
