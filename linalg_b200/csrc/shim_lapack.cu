// la_b200 — LAPACK drop-in symbols (host pointers, Fortran-77 ABI).  See include/la_b200.h,
// Part 2.  Replaces the `fun`s of src/linalg/libLAPACK_linux.cr cited there; the call
// protocol (all-pointer arguments, lwork = -1 query, info) is the one the reference's
// `lapack` macro drives (src/linalg/lapack_helper.cr:215-372).
//
// Error convention: info = -i for an illegal i-th argument (nothing is
// touched), info > 0 with LAPACK's meaning (zero pivot / non-PD minor / zero R diagonal).
// CUDA failures abort the process with a message; they never return success.
#include "shim_common.h"
#include <vector>

namespace la {

static inline int64_t imax(int64_t a, int64_t b) { return a > b ? a : b; }
static inline int64_t imin(int64_t a, int64_t b) { return a < b ? a : b; }

template <typename T>
static void set_work(T* work, double v) {
  if (work) work[0] = from_real<T>(v);
}

// ------------------------------------------------------------------ getrf / getrs / gesv
// getrf on the staged matrix with the factors sent home block row by block row: the rows of a
// panel's diagonal block are final for the whole matrix width as soon as U12 has been solved
// (later panels only interchange rows below), so their device->host copy runs on the second copy
// stream while the factorisation continues; only the last block row is exposed.
template <typename T>
static void getrf_streamed_home(Context& ctx, int64_t m, int64_t n, DevMat<T>& dA, int* d_ipiv,
                                T* A, int64_t lda) {
  int64_t done = 0;
  std::vector<cudaEvent_t> evs;
  if (m >= 4096 && n >= 4096) {
    ctx.on_block_final = [&](int64_t r0, int64_t r1) {
      if (r0 != done) return;  // only contiguous progress is streamed; the rest goes at the end
      cudaEvent_t e;
      LA_CUDA_CHECK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
      LA_CUDA_CHECK(cudaEventRecord(e, ctx.stream));
      LA_CUDA_CHECK(cudaStreamWaitEvent(ctx.d2h_stream, e, 0));
      evs.push_back(e);
      d2h_2d(ctx, A + r0, (size_t)lda * sizeof(T), dA.p + r0, (size_t)dA.ld * sizeof(T),
             (size_t)(r1 - r0) * sizeof(T), (size_t)n, ctx.d2h_stream);
      done = r1;
    };
  }
  getrf<T>(ctx, m, n, dA.p, dA.ld, d_ipiv, ctx.d_info);
  ctx.on_block_final = nullptr;
  if (m - done > 0)
    d2h_2d(ctx, A + done, (size_t)lda * sizeof(T), dA.p + done, (size_t)dA.ld * sizeof(T),
           (size_t)(m - done) * sizeof(T), (size_t)n, ctx.stream);
  LA_CUDA_CHECK(cudaStreamSynchronize(ctx.d2h_stream));
  for (auto e : evs) cudaEventDestroy(e);
}

template <typename T>
static void h_getrf(const int* pm, const int* pn, T* A, const int* plda, int* ipiv, int* info) {
  const int64_t m = *pm, n = *pn, lda = *plda;
  *info = 0;
  if (m < 0) { *info = -1; return; }
  if (n < 0) { *info = -2; return; }
  if (lda < imax(1, m)) { *info = -4; return; }
  if (m == 0 || n == 0) return;
  ContextLease ctx;
  Arena::Mark mk = ctx->arena.mark();
  const int64_t k = imin(m, n);
  DevMat<T> dA = dev_from_host<T>(*ctx, m, n, A, lda);
  int* d_ipiv = ctx->alloc<int>(k);
  getrf_streamed_home<T>(*ctx, m, n, dA, d_ipiv, A, lda);
  LA_CUDA_CHECK(cudaMemcpyAsync(ipiv, d_ipiv, k * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  *info = read_info(*ctx, ctx->d_info);
  ctx->arena.release(mk);
}

template <typename T>
static void h_getrs(const char* ptrans, const int* pn, const int* pnrhs, const T* A,
                    const int* plda, const int* ipiv, T* B, const int* pldb, int* info) {
  const int64_t n = *pn, nrhs = *pnrhs, lda = *plda, ldb = *pldb;
  const int trans = op_from_char(*ptrans);
  *info = 0;
  if (!trans) { *info = -1; return; }
  if (n < 0) { *info = -2; return; }
  if (nrhs < 0) { *info = -3; return; }
  if (lda < imax(1, n)) { *info = -5; return; }
  if (ldb < imax(1, n)) { *info = -8; return; }
  if (n == 0 || nrhs == 0) return;
  ContextLease ctx;
  Arena::Mark mk = ctx->arena.mark();
  DevMat<T> dA = dev_from_host<T>(*ctx, n, n, A, lda);
  DevMat<T> dB = dev_from_host<T>(*ctx, n, nrhs, B, ldb);
  int* d_ipiv = ctx->alloc<int>(n);
  LA_CUDA_CHECK(cudaMemcpyAsync(d_ipiv, ipiv, n * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
  getrs<T>(*ctx, (Op)trans, n, nrhs, dA.p, dA.ld, d_ipiv, dB.p, dB.ld);
  download<T>(*ctx, dB, B, ldb);
  sync(*ctx);
  ctx->arena.release(mk);
}

template <typename T>
static void h_gesv(const int* pn, const int* pnrhs, T* A, const int* plda, int* ipiv, T* B,
                   const int* pldb, int* info) {
  const int64_t n = *pn, nrhs = *pnrhs, lda = *plda, ldb = *pldb;
  *info = 0;
  if (n < 0) { *info = -1; return; }
  if (nrhs < 0) { *info = -2; return; }
  if (lda < imax(1, n)) { *info = -4; return; }
  if (ldb < imax(1, n)) { *info = -7; return; }
  if (n == 0) return;
  ContextLease ctx;
  Arena::Mark mk = ctx->arena.mark();
  DevMat<T> dA = dev_from_host<T>(*ctx, n, n, A, lda);
  DevMat<T> dB = dev_from_host<T>(*ctx, n, nrhs, B, ldb);
  int* d_ipiv = ctx->alloc<int>(n);
  getrf_streamed_home<T>(*ctx, n, n, dA, d_ipiv, A, lda);
  LA_CUDA_CHECK(cudaMemcpyAsync(ipiv, d_ipiv, n * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  const int inf = read_info(*ctx, ctx->d_info);
  if (inf == 0 && nrhs > 0) {  // LAPACK solves only when U is non-singular
    getrs<T>(*ctx, OpN, n, nrhs, dA.p, dA.ld, d_ipiv, dB.p, dB.ld);
    download<T>(*ctx, dB, B, ldb);
    sync(*ctx);
  }
  *info = inf;
  ctx->arena.release(mk);
}

// potrf on the staged matrix with the factor sent home block by block: every block column (L)
// or block row (U) is final as soon as its panel has been solved, so its device->host copy runs
// on the second copy stream while the factorisation continues; only the last block is exposed.
template <typename T>
static void potrf_streamed_home(Context& ctx, int uplo, int64_t n, DevMat<T>& dA, T* A, int64_t lda) {
  const bool lower = uplo == Lower;
  int64_t done = 0;
  std::vector<cudaEvent_t> evs;
  ctx.on_block_final = [&](int64_t c0, int64_t c1) {
    if (c0 != done) return;  // only contiguous progress is streamed; the rest goes at the end
    cudaEvent_t e;
    LA_CUDA_CHECK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    LA_CUDA_CHECK(cudaEventRecord(e, ctx.stream));
    LA_CUDA_CHECK(cudaStreamWaitEvent(ctx.d2h_stream, e, 0));
    evs.push_back(e);
    const T* d = dA.p + c0 + c0 * dA.ld;
    T* h = A + c0 + c0 * lda;
    // lower: columns [c0, c1), rows c0..n ; upper: rows [c0, c1), columns c0..n
    const size_t width = (size_t)(lower ? n - c0 : c1 - c0) * sizeof(T);
    const size_t height = (size_t)(lower ? c1 - c0 : n - c0);
    d2h_2d(ctx, h, (size_t)lda * sizeof(T), d, (size_t)dA.ld * sizeof(T), width, height,
           ctx.d2h_stream);
    done = c1;
  };
  potrf<T>(ctx, (Uplo)uplo, n, dA.p, dA.ld, ctx.d_info);
  ctx.on_block_final = nullptr;
  // (streamed blocks and uploaded blocks have the same width, so the part of a diagonal block
  // that lies in the other triangle only makes a round trip)
  if (n - done > 0)
    copy_triangle_hd<T>(ctx, false, lower, n - done, dA.p + done + done * dA.ld, dA.ld,
                        A + done + done * lda, lda, potrf_block_width(n));
  LA_CUDA_CHECK(cudaStreamSynchronize(ctx.d2h_stream));
  for (auto e : evs) cudaEventDestroy(e);
}

// ------------------------------------------------------------------ potrf / potrs / posv
template <typename T>
static void h_potrf(const char* puplo, const int* pn, T* A, const int* plda, int* info) {
  const int64_t n = *pn, lda = *plda;
  const int uplo = uplo_from_char(*puplo);
  *info = 0;
  if (!uplo) { *info = -1; return; }
  if (n < 0) { *info = -2; return; }
  if (lda < imax(1, n)) { *info = -4; return; }
  if (n == 0) return;
  ContextLease ctx;
  Arena::Mark mk = ctx->arena.mark();
  // the kernels touch only the named triangle: only that triangle travels, and the other one
  // stays bit-identical on the host ("other triangle untouched", cholesky.cr:50-64 relies on it)
  DevMat<T> dA = dev_alloc<T>(*ctx, n, n);
  copy_triangle_hd<T>(*ctx, true, uplo == Lower, n, dA.p, dA.ld, A, lda, potrf_block_width(n));
  potrf_streamed_home<T>(*ctx, uplo, n, dA, A, lda);
  *info = read_info(*ctx, ctx->d_info);
  ctx->arena.release(mk);
}

template <typename T>
static void h_potrs(const char* puplo, const int* pn, const int* pnrhs, const T* A,
                    const int* plda, T* B, const int* pldb, int* info) {
  const int64_t n = *pn, nrhs = *pnrhs, lda = *plda, ldb = *pldb;
  const int uplo = uplo_from_char(*puplo);
  *info = 0;
  if (!uplo) { *info = -1; return; }
  if (n < 0) { *info = -2; return; }
  if (nrhs < 0) { *info = -3; return; }
  if (lda < imax(1, n)) { *info = -5; return; }
  if (ldb < imax(1, n)) { *info = -7; return; }
  if (n == 0 || nrhs == 0) return;
  ContextLease ctx;
  Arena::Mark mk = ctx->arena.mark();
  DevMat<T> dA = dev_from_host<T>(*ctx, n, n, A, lda);
  DevMat<T> dB = dev_from_host<T>(*ctx, n, nrhs, B, ldb);
  potrs<T>(*ctx, (Uplo)uplo, n, nrhs, dA.p, dA.ld, dB.p, dB.ld);
  download<T>(*ctx, dB, B, ldb);
  sync(*ctx);
  ctx->arena.release(mk);
}

template <typename T>
static void h_posv(const char* puplo, const int* pn, const int* pnrhs, T* A, const int* plda,
                   T* B, const int* pldb, int* info) {
  const int64_t n = *pn, nrhs = *pnrhs, lda = *plda, ldb = *pldb;
  const int uplo = uplo_from_char(*puplo);
  *info = 0;
  if (!uplo) { *info = -1; return; }
  if (n < 0) { *info = -2; return; }
  if (nrhs < 0) { *info = -3; return; }
  if (lda < imax(1, n)) { *info = -5; return; }
  if (ldb < imax(1, n)) { *info = -7; return; }
  if (n == 0) return;
  ContextLease ctx;
  Arena::Mark mk = ctx->arena.mark();
  DevMat<T> dA = dev_alloc<T>(*ctx, n, n);
  copy_triangle_hd<T>(*ctx, true, uplo == Lower, n, dA.p, dA.ld, A, lda, potrf_block_width(n));
  DevMat<T> dB = dev_from_host<T>(*ctx, n, nrhs, B, ldb);
  potrf_streamed_home<T>(*ctx, uplo, n, dA, A, lda);
  const int inf = read_info(*ctx, ctx->d_info);
  if (inf == 0 && nrhs > 0) {
    potrs<T>(*ctx, (Uplo)uplo, n, nrhs, dA.p, dA.ld, dB.p, dB.ld);
    download<T>(*ctx, dB, B, ldb);
    sync(*ctx);
  }
  *info = inf;
  ctx->arena.release(mk);
}

// ------------------------------------------------------------------ trtrs
template <typename T>
static void h_trtrs(const char* puplo, const char* ptrans, const char* pdiag, const int* pn,
                    const int* pnrhs, const T* A, const int* plda, T* B, const int* pldb,
                    int* info) {
  const int64_t n = *pn, nrhs = *pnrhs, lda = *plda, ldb = *pldb;
  const int uplo = uplo_from_char(*puplo), trans = op_from_char(*ptrans),
            diag = diag_from_char(*pdiag);
  *info = 0;
  if (!uplo) { *info = -1; return; }
  if (!trans) { *info = -2; return; }
  if (!diag) { *info = -3; return; }
  if (n < 0) { *info = -4; return; }
  if (nrhs < 0) { *info = -5; return; }
  if (lda < imax(1, n)) { *info = -7; return; }
  if (ldb < imax(1, n)) { *info = -9; return; }
  if (n == 0) return;
  ContextLease ctx;
  Arena::Mark mk = ctx->arena.mark();
  DevMat<T> dA = dev_from_host<T>(*ctx, n, n, A, lda);
  DevMat<T> dB = dev_from_host<T>(*ctx, n, nrhs, B, ldb);
  trtrs<T>(*ctx, (Uplo)uplo, (Op)trans, (Diag)diag, n, nrhs, dA.p, dA.ld, dB.p, dB.ld,
           ctx->d_info);
  const int inf = read_info(*ctx, ctx->d_info);
  if (inf == 0) {  // singular: LAPACK returns before solving, B stays as it was
    download<T>(*ctx, dB, B, ldb);
    sync(*ctx);
  }
  *info = inf;
  ctx->arena.release(mk);
}

// ------------------------------------------------------------------ inverses
template <typename T>
static void h_getri(const int* pn, T* A, const int* plda, const int* ipiv, T* work,
                    const int* plwork, int* info) {
  const int64_t n = *pn, lda = *plda, lwork = *plwork;
  *info = 0;
  if (n < 0) { *info = -1; return; }
  if (lda < imax(1, n)) { *info = -3; return; }
  if (lwork != -1 && lwork < imax(1, n)) { *info = -6; return; }
  if (lwork == -1) { set_work<T>(work, (double)imax(1, n)); return; }
  set_work<T>(work, (double)imax(1, n));
  if (n == 0) return;
  ContextLease ctx;
  Arena::Mark mk = ctx->arena.mark();
  DevMat<T> dA = dev_from_host<T>(*ctx, n, n, A, lda);
  int* d_ipiv = ctx->alloc<int>(n);
  LA_CUDA_CHECK(cudaMemcpyAsync(d_ipiv, ipiv, n * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
  getri<T>(*ctx, n, dA.p, dA.ld, d_ipiv, ctx->d_info);
  download<T>(*ctx, dA, A, lda);
  *info = read_info(*ctx, ctx->d_info);
  ctx->arena.release(mk);
}

template <typename T>
static void h_potri(const char* puplo, const int* pn, T* A, const int* plda, int* info) {
  const int64_t n = *pn, lda = *plda;
  const int uplo = uplo_from_char(*puplo);
  *info = 0;
  if (!uplo) { *info = -1; return; }
  if (n < 0) { *info = -2; return; }
  if (lda < imax(1, n)) { *info = -4; return; }
  if (n == 0) return;
  ContextLease ctx;
  Arena::Mark mk = ctx->arena.mark();
  DevMat<T> dA = dev_from_host<T>(*ctx, n, n, A, lda);
  potri<T>(*ctx, (Uplo)uplo, n, dA.p, dA.ld, ctx->d_info);
  download<T>(*ctx, dA, A, lda);
  *info = read_info(*ctx, ctx->d_info);
  ctx->arena.release(mk);
}

template <typename T>
static void h_trtri(const char* puplo, const char* pdiag, const int* pn, T* A, const int* plda,
                    int* info) {
  const int64_t n = *pn, lda = *plda;
  const int uplo = uplo_from_char(*puplo), diag = diag_from_char(*pdiag);
  *info = 0;
  if (!uplo) { *info = -1; return; }
  if (!diag) { *info = -2; return; }
  if (n < 0) { *info = -3; return; }
  if (lda < imax(1, n)) { *info = -5; return; }
  if (n == 0) return;
  ContextLease ctx;
  Arena::Mark mk = ctx->arena.mark();
  DevMat<T> dA = dev_from_host<T>(*ctx, n, n, A, lda);
  trtri<T>(*ctx, (Uplo)uplo, (Diag)diag, n, dA.p, dA.ld, ctx->d_info);
  download<T>(*ctx, dA, A, lda);
  *info = read_info(*ctx, ctx->d_info);
  ctx->arena.release(mk);
}

}  // namespace la

using namespace la;

#define LA_LAPACK_FACTOR_SYMBOLS(X, T, VT)                                                        \
  void X##getrf_(const int* m, const int* n, VT* a, const int* lda, int* ipiv, int* info) {       \
    h_getrf<T>(m, n, (T*)a, lda, ipiv, info);                                                     \
  }                                                                                               \
  void X##getrs_(const char* trans, const int* n, const int* nrhs, const VT* a, const int* lda,   \
                 const int* ipiv, VT* b, const int* ldb, int* info) {                             \
    h_getrs<T>(trans, n, nrhs, (const T*)a, lda, ipiv, (T*)b, ldb, info);                         \
  }                                                                                               \
  void X##gesv_(const int* n, const int* nrhs, VT* a, const int* lda, int* ipiv, VT* b,           \
                const int* ldb, int* info) {                                                      \
    h_gesv<T>(n, nrhs, (T*)a, lda, ipiv, (T*)b, ldb, info);                                       \
  }                                                                                               \
  void X##potrf_(const char* uplo, const int* n, VT* a, const int* lda, int* info) {              \
    h_potrf<T>(uplo, n, (T*)a, lda, info);                                                        \
  }                                                                                               \
  void X##potrs_(const char* uplo, const int* n, const int* nrhs, const VT* a, const int* lda,    \
                 VT* b, const int* ldb, int* info) {                                              \
    h_potrs<T>(uplo, n, nrhs, (const T*)a, lda, (T*)b, ldb, info);                                \
  }                                                                                               \
  void X##posv_(const char* uplo, const int* n, const int* nrhs, VT* a, const int* lda, VT* b,    \
                const int* ldb, int* info) {                                                      \
    h_posv<T>(uplo, n, nrhs, (T*)a, lda, (T*)b, ldb, info);                                       \
  }                                                                                               \
  void X##trtrs_(const char* uplo, const char* trans, const char* diag, const int* n,             \
                 const int* nrhs, const VT* a, const int* lda, VT* b, const int* ldb,             \
                 int* info) {                                                                     \
    h_trtrs<T>(uplo, trans, diag, n, nrhs, (const T*)a, lda, (T*)b, ldb, info);                   \
  }                                                                                               \
  void X##getri_(const int* n, VT* a, const int* lda, const int* ipiv, VT* work,                  \
                 const int* lwork, int* info) {                                                   \
    h_getri<T>(n, (T*)a, lda, ipiv, (T*)work, lwork, info);                                       \
  }                                                                                               \
  void X##potri_(const char* uplo, const int* n, VT* a, const int* lda, int* info) {              \
    h_potri<T>(uplo, n, (T*)a, lda, info);                                                        \
  }                                                                                               \
  void X##trtri_(const char* uplo, const char* diag, const int* n, VT* a, const int* lda,         \
                 int* info) {                                                                     \
    h_trtri<T>(uplo, diag, n, (T*)a, lda, info);                                                  \
  }

extern "C" {

void ilaver_(int* major, int* minor, int* patch) {
  *major = 3; *minor = 12; *patch = 0;
}

LA_LAPACK_FACTOR_SYMBOLS(s, float, float)
LA_LAPACK_FACTOR_SYMBOLS(d, double, double)
LA_LAPACK_FACTOR_SYMBOLS(z, zcomplex, void)

}  // extern "C"
