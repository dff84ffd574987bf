// la_b200 — helpers shared by the C-ABI translation units (host-pointer drop-in symbols
// and the device-pointer API).
#pragma once
#include "la_internal.h"
#include <cstring>

namespace la {

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
  LA_CUDA_CHECK(cudaMemcpy2DAsync(d.p, (size_t)d.ld * sizeof(T), h, (size_t)ldh * sizeof(T),
                                  (size_t)d.rows * sizeof(T), (size_t)d.cols,
                                  cudaMemcpyHostToDevice, s ? s : ctx.stream));
}

template <typename T>
inline void download(Context& ctx, const DevMat<T>& d, T* h, int64_t ldh, cudaStream_t s = nullptr) {
  if (d.rows <= 0 || d.cols <= 0) return;
  LA_CUDA_CHECK(cudaMemcpy2DAsync(h, (size_t)ldh * sizeof(T), d.p, (size_t)d.ld * sizeof(T),
                                  (size_t)d.rows * sizeof(T), (size_t)d.cols,
                                  cudaMemcpyDeviceToHost, s ? s : ctx.stream));
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
      LA_CUDA_CHECK(cudaMemcpy2DAsync(d, (size_t)ldd * sizeof(T), h, (size_t)ldh * sizeof(T),
                                      (size_t)rows * sizeof(T), (size_t)w, cudaMemcpyHostToDevice,
                                      s ? s : ctx.stream));
    else
      LA_CUDA_CHECK(cudaMemcpy2DAsync(h, (size_t)ldh * sizeof(T), d, (size_t)ldd * sizeof(T),
                                      (size_t)rows * sizeof(T), (size_t)w, cudaMemcpyDeviceToHost,
                                      s ? s : ctx.stream));
  }
}

inline void sync(Context& ctx) { LA_CUDA_CHECK(cudaStreamSynchronize(ctx.stream)); }

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
