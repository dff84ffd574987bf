// la_b200 — device-resident C API (include/la_b200.h, Part 3): same routines as the drop-in
// symbols but on DEVICE pointers, 64-bit sizes and the caller's stream; nothing is copied
// and nothing synchronises.  Used by bench.py (kernel-only timing), by the multi-GPU
// drivers in linalg_b200/dist.py and by hosts that keep operands resident across a chain of
// products.
#include "shim_common.h"

namespace la {
namespace {

struct DevCall {
  Context* ctx;
  Arena::Mark mk;
  // The per-thread context's scratch is shared by every call of this thread on this device:
  // a call on a different stream than the previous one is ordered after it with an event.
  explicit DevCall(void* stream) : ctx(thread_context((cudaStream_t)stream)), mk(ctx->arena.mark()) {
    if (ctx->has_last && ctx->last_stream != ctx->stream)
      LA_CUDA_CHECK(cudaStreamWaitEvent(ctx->stream, ctx->tail_ev, 0));
  }
  ~DevCall() {
    ctx->arena.release(mk);
    LA_CUDA_CHECK(cudaEventRecord(ctx->tail_ev, ctx->stream));
    ctx->last_stream = ctx->stream;
    ctx->has_last = true;
  }
};

template <typename T>
int d_gemm(int ta, int tb, int64_t m, int64_t n, int64_t k, T alpha, const T* A, int64_t lda,
           const T* B, int64_t ldb, T beta, T* C, int64_t ldc, void* stream) {
  if (!valid_op(ta)) return -1;
  if (!valid_op(tb)) return -2;
  if (m < 0) return -3;
  if (n < 0) return -4;
  if (k < 0) return -5;
  const int64_t ar = ta == OpN ? m : k, br = tb == OpN ? k : n;
  if (lda < (ar > 1 ? ar : 1)) return -8;
  if (ldb < (br > 1 ? br : 1)) return -10;
  if (ldc < (m > 1 ? m : 1)) return -13;
  DevCall c(stream);
  gemm<T>(*c.ctx, (Op)ta, (Op)tb, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc);
  return 0;
}

template <typename T>
int d_gemm_stair(int ta, int tb, int64_t m, int64_t n, int64_t k, T alpha, const T* A, int64_t lda,
                 const T* B, int64_t ldb, T beta, T* C, int64_t ldc, int64_t nb, int64_t row_first,
                 int64_t row_step, int64_t col_first, int64_t col_step, void* stream) {
  if (!valid_op(ta)) return -1;
  if (!valid_op(tb)) return -2;
  if (m < 0) return -3;
  if (n < 0) return -4;
  if (k < 0) return -5;
  const int64_t ar = ta == OpN ? m : k, br = tb == OpN ? k : n;
  if (lda < (ar > 1 ? ar : 1)) return -8;
  if (ldb < (br > 1 ? br : 1)) return -10;
  if (ldc < (m > 1 ? m : 1)) return -13;
  if (nb < 1) return -14;
  if (row_first < 0) return -15;
  if (row_step < 1) return -16;
  if (col_first < 0) return -17;
  if (col_step < 1) return -18;
  DevCall c(stream);
  StairMask mk;
  mk.nb = nb; mk.row_first = row_first; mk.row_step = row_step;
  mk.col_first = col_first; mk.col_step = col_step;
  gemm_stair<T>(*c.ctx, (Op)ta, (Op)tb, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, mk);
  return 0;
}

template <typename T>
int d_lange(int norm, int64_t m, int64_t n, const T* A, int64_t lda, double* result, void* stream) {
  const int kind = norm_kind_from_char((char)norm);
  if (kind < 0) return -1;
  if (m < 0) return -2;
  if (n < 0) return -3;
  if (lda < (m > 1 ? m : 1)) return -5;
  DevCall c(stream);
  matrix_norm<T>(*c.ctx, kind, ShapeGeneral, Upper, false, m, n, A, lda, result);
  return 0;
}

template <typename T>
int d_trsm(int side, int uplo, int ta, int diag, int64_t m, int64_t n, const void* alpha,
           const T* A, int64_t lda, T* B, int64_t ldb, void* stream) {
  if (!valid_side(side)) return -1;
  if (!valid_uplo(uplo)) return -2;
  if (!valid_op(ta)) return -3;
  if (!valid_diag(diag)) return -4;
  if (m < 0) return -5;
  if (n < 0) return -6;
  const int64_t na = side == Left ? m : n;
  if (lda < (na > 1 ? na : 1)) return -9;
  if (ldb < (m > 1 ? m : 1)) return -11;
  DevCall c(stream);
  trsm<T>(*c.ctx, (Side)side, (Uplo)uplo, (Op)ta, (Diag)diag, m, n, *(const T*)alpha, A, lda, B,
          ldb);
  return 0;
}

template <typename T>
int d_herk(int uplo, int trans, int64_t n, int64_t k, double alpha, const T* A, int64_t lda,
           double beta, T* C, int64_t ldc, void* stream) {
  if (!valid_uplo(uplo)) return -1;
  if (!valid_op(trans)) return -2;
  if (n < 0) return -3;
  if (k < 0) return -4;
  const int64_t ar = trans == OpN ? n : k;
  if (lda < (ar > 1 ? ar : 1)) return -7;
  if (ldc < (n > 1 ? n : 1)) return -10;
  DevCall c(stream);
  herk<T>(*c.ctx, (Uplo)uplo, trans == OpN ? OpN : OpC, n, k, alpha, A, lda, beta, C, ldc);
  return 0;
}

template <typename T>
int d_getrf(int64_t m, int64_t n, T* A, int64_t lda, int* ipiv, int* info, void* stream) {
  if (m < 0) return -1;
  if (n < 0) return -2;
  if (lda < (m > 1 ? m : 1)) return -4;
  DevCall c(stream);
  getrf<T>(*c.ctx, m, n, A, lda, ipiv, info);
  return 0;
}
template <typename T>
int d_laswp(int64_t m, int64_t n, T* A, int64_t lda, int64_t k1, int64_t k2, const int* ipiv,
            void* stream) {
  if (m < 0) return -1;
  if (n < 0) return -2;
  if (lda < (m > 1 ? m : 1)) return -4;
  if (k1 < 0 || k1 > m) return -5;
  if (k2 < k1 || k2 > m) return -6;
  DevCall c(stream);
  laswp_rows<T>(*c.ctx, m, n, A, lda, k1, k2, ipiv, false);
  return 0;
}
template <typename T>
int d_getrs(int trans, int64_t n, int64_t nrhs, const T* A, int64_t lda, const int* ipiv, T* B,
            int64_t ldb, void* stream) {
  if (!valid_op(trans)) return -1;
  if (n < 0) return -2;
  if (nrhs < 0) return -3;
  if (lda < (n > 1 ? n : 1)) return -5;
  if (ldb < (n > 1 ? n : 1)) return -8;
  DevCall c(stream);
  getrs<T>(*c.ctx, (Op)trans, n, nrhs, A, lda, ipiv, B, ldb);
  return 0;
}
template <typename T>
int d_potrf(int uplo, int64_t n, T* A, int64_t lda, int* info, void* stream) {
  if (!valid_uplo(uplo)) return -1;
  if (n < 0) return -2;
  if (lda < (n > 1 ? n : 1)) return -4;
  DevCall c(stream);
  potrf<T>(*c.ctx, (Uplo)uplo, n, A, lda, info);
  return 0;
}
template <typename T>
int d_potrs(int uplo, int64_t n, int64_t nrhs, const T* A, int64_t lda, T* B, int64_t ldb,
            void* stream) {
  if (!valid_uplo(uplo)) return -1;
  if (n < 0) return -2;
  if (nrhs < 0) return -3;
  if (lda < (n > 1 ? n : 1)) return -5;
  if (ldb < (n > 1 ? n : 1)) return -7;
  DevCall c(stream);
  potrs<T>(*c.ctx, (Uplo)uplo, n, nrhs, A, lda, B, ldb);
  return 0;
}
template <typename T>
int d_geqrf(int64_t m, int64_t n, T* A, int64_t lda, T* tau, void* stream) {
  if (m < 0) return -1;
  if (n < 0) return -2;
  if (lda < (m > 1 ? m : 1)) return -4;
  DevCall c(stream);
  geqrf<T>(*c.ctx, m, n, A, lda, tau);
  return 0;
}
template <typename T>
int d_ormqr(int side, int trans, int64_t m, int64_t n, int64_t k, const T* A, int64_t lda,
            const T* tau, T* C, int64_t ldc, void* stream) {
  if (!valid_side(side)) return -1;
  if (!valid_op(trans)) return -2;
  if (m < 0) return -3;
  if (n < 0) return -4;
  const int64_t nq = side == Left ? m : n;
  if (k < 0 || k > nq) return -5;
  if (lda < (nq > 1 ? nq : 1)) return -7;
  if (ldc < (m > 1 ? m : 1)) return -10;
  DevCall c(stream);
  ormqr<T>(*c.ctx, (Side)side, trans == OpN ? OpN : OpC, m, n, k, A, lda, tau, C, ldc);
  return 0;
}
template <typename T>
int d_gels(int trans, int64_t m, int64_t n, int64_t nrhs, T* A, int64_t lda, T* B, int64_t ldb,
           int* info, void* stream) {
  if (!valid_op(trans)) return -1;
  if (m < 0) return -2;
  if (n < 0) return -3;
  if (nrhs < 0) return -4;
  if (lda < (m > 1 ? m : 1)) return -6;
  const int64_t mx = m > n ? m : n;
  if (ldb < (mx > 1 ? mx : 1)) return -8;
  DevCall c(stream);
  gels<T>(*c.ctx, trans == OpN ? OpN : OpC, m, n, nrhs, A, lda, B, ldb, info);
  return 0;
}
template <typename T>
int d_trtrs(int uplo, int trans, int diag, int64_t n, int64_t nrhs, const T* A, int64_t lda, T* B,
            int64_t ldb, int* info, void* stream) {
  if (!valid_uplo(uplo)) return -1;
  if (!valid_op(trans)) return -2;
  if (!valid_diag(diag)) return -3;
  if (n < 0) return -4;
  if (nrhs < 0) return -5;
  if (lda < (n > 1 ? n : 1)) return -7;
  if (ldb < (n > 1 ? n : 1)) return -9;
  DevCall c(stream);
  trtrs<T>(*c.ctx, (Uplo)uplo, (Op)trans, (Diag)diag, n, nrhs, A, lda, B, ldb, info);
  return 0;
}

}  // namespace
}  // namespace la

using namespace la;

#define LA_DEV_SYMBOLS(X, T, VT)                                                                  \
  int la_b200_##X##trsm_dev(int side, int uplo, int ta, int diag, int64_t m, int64_t n,           \
                            const void* alpha, const VT* a, int64_t lda, VT* b, int64_t ldb,      \
                            void* stream) {                                                       \
    return d_trsm<T>(side, uplo, ta, diag, m, n, alpha, (const T*)a, lda, (T*)b, ldb, stream);    \
  }                                                                                               \
  int la_b200_##X##herk_dev(int uplo, int trans, int64_t n, int64_t k, double alpha,              \
                            const VT* a, int64_t lda, double beta, VT* c, int64_t ldc,            \
                            void* stream) {                                                       \
    return d_herk<T>(uplo, trans, n, k, alpha, (const T*)a, lda, beta, (T*)c, ldc, stream);       \
  }                                                                                               \
  int la_b200_##X##getrf_dev(int64_t m, int64_t n, VT* a, int64_t lda, int* ipiv, int* info,      \
                             void* stream) {                                                      \
    return d_getrf<T>(m, n, (T*)a, lda, ipiv, info, stream);                                      \
  }                                                                                               \
  int la_b200_##X##laswp_dev(int64_t m, int64_t n, VT* a, int64_t lda, int64_t k1, int64_t k2,    \
                             const int* ipiv, void* stream) {                                     \
    return d_laswp<T>(m, n, (T*)a, lda, k1, k2, ipiv, stream);                                    \
  }                                                                                               \
  int la_b200_##X##lange_dev(int norm, int64_t m, int64_t n, const VT* a, int64_t lda,            \
                             double* result, void* stream) {                                      \
    return d_lange<T>(norm, m, n, (const T*)a, lda, result, stream);                              \
  }                                                                                               \
  int la_b200_##X##getrs_dev(int trans, int64_t n, int64_t nrhs, const VT* a, int64_t lda,        \
                             const int* ipiv, VT* b, int64_t ldb, void* stream) {                 \
    return d_getrs<T>(trans, n, nrhs, (const T*)a, lda, ipiv, (T*)b, ldb, stream);                \
  }                                                                                               \
  int la_b200_##X##potrf_dev(int uplo, int64_t n, VT* a, int64_t lda, int* info, void* stream) {  \
    return d_potrf<T>(uplo, n, (T*)a, lda, info, stream);                                         \
  }                                                                                               \
  int la_b200_##X##potrs_dev(int uplo, int64_t n, int64_t nrhs, const VT* a, int64_t lda, VT* b,  \
                             int64_t ldb, void* stream) {                                         \
    return d_potrs<T>(uplo, n, nrhs, (const T*)a, lda, (T*)b, ldb, stream);                       \
  }                                                                                               \
  int la_b200_##X##geqrf_dev(int64_t m, int64_t n, VT* a, int64_t lda, VT* tau, void* stream) {   \
    return d_geqrf<T>(m, n, (T*)a, lda, (T*)tau, stream);                                         \
  }                                                                                               \
  int la_b200_##X##ormqr_dev(int side, int trans, int64_t m, int64_t n, int64_t k, const VT* a,   \
                             int64_t lda, const VT* tau, VT* c, int64_t ldc, void* stream) {      \
    return d_ormqr<T>(side, trans, m, n, k, (const T*)a, lda, (const T*)tau, (T*)c, ldc, stream); \
  }                                                                                               \
  int la_b200_##X##gels_dev(int trans, int64_t m, int64_t n, int64_t nrhs, VT* a, int64_t lda,    \
                            VT* b, int64_t ldb, int* info, void* stream) {                        \
    return d_gels<T>(trans, m, n, nrhs, (T*)a, lda, (T*)b, ldb, info, stream);                    \
  }                                                                                               \
  int la_b200_##X##trtrs_dev(int uplo, int trans, int diag, int64_t n, int64_t nrhs,              \
                             const VT* a, int64_t lda, VT* b, int64_t ldb, int* info,             \
                             void* stream) {                                                      \
    return d_trtrs<T>(uplo, trans, diag, n, nrhs, (const T*)a, lda, (T*)b, ldb, info, stream);    \
  }

extern "C" {

int la_b200_sgemm_dev(int ta, int tb, int64_t m, int64_t n, int64_t k, float alpha, const float* a,
                      int64_t lda, const float* b, int64_t ldb, float beta, float* c, int64_t ldc,
                      void* stream) {
  return d_gemm<float>(ta, tb, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc, stream);
}
int la_b200_dgemm_dev(int ta, int tb, int64_t m, int64_t n, int64_t k, double alpha,
                      const double* a, int64_t lda, const double* b, int64_t ldb, double beta,
                      double* c, int64_t ldc, void* stream) {
  return d_gemm<double>(ta, tb, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc, stream);
}
int la_b200_zgemm_dev(int ta, int tb, int64_t m, int64_t n, int64_t k, const void* alpha,
                      const void* a, int64_t lda, const void* b, int64_t ldb, const void* beta,
                      void* c, int64_t ldc, void* stream) {
  return d_gemm<zcomplex>(ta, tb, m, n, k, *(const zcomplex*)alpha, (const zcomplex*)a, lda,
                          (const zcomplex*)b, ldb, *(const zcomplex*)beta, (zcomplex*)c, ldc,
                          stream);
}

int la_b200_sgemm_stair_dev(int ta, int tb, int64_t m, int64_t n, int64_t k, float alpha,
                            const float* a, int64_t lda, const float* b, int64_t ldb, float beta,
                            float* c, int64_t ldc, int64_t nb, int64_t row_first, int64_t row_step,
                            int64_t col_first, int64_t col_step, void* stream) {
  return d_gemm_stair<float>(ta, tb, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc, nb, row_first,
                             row_step, col_first, col_step, stream);
}
int la_b200_dgemm_stair_dev(int ta, int tb, int64_t m, int64_t n, int64_t k, double alpha,
                            const double* a, int64_t lda, const double* b, int64_t ldb,
                            double beta, double* c, int64_t ldc, int64_t nb, int64_t row_first,
                            int64_t row_step, int64_t col_first, int64_t col_step, void* stream) {
  return d_gemm_stair<double>(ta, tb, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc, nb, row_first,
                              row_step, col_first, col_step, stream);
}
int la_b200_zgemm_stair_dev(int ta, int tb, int64_t m, int64_t n, int64_t k, const void* alpha,
                            const void* a, int64_t lda, const void* b, int64_t ldb,
                            const void* beta, void* c, int64_t ldc, int64_t nb, int64_t row_first,
                            int64_t row_step, int64_t col_first, int64_t col_step, void* stream) {
  return d_gemm_stair<zcomplex>(ta, tb, m, n, k, *(const zcomplex*)alpha, (const zcomplex*)a, lda,
                                (const zcomplex*)b, ldb, *(const zcomplex*)beta, (zcomplex*)c, ldc,
                                nb, row_first, row_step, col_first, col_step, stream);
}

// Persistent kernels of the calling thread's device-API context use at most `sms` CTAs
// (0 = all): leaves SMs free for communication kernels that must run beside a long product.
int la_b200_set_sm_limit(int sms) {
  Context* c = thread_context(nullptr);
  const int old = c->sm_limit;
  c->sm_limit = sms > 0 && sms < c->sm_count ? sms : 0;
  return old;
}

LA_DEV_SYMBOLS(s, float, float)
LA_DEV_SYMBOLS(d, double, double)
LA_DEV_SYMBOLS(z, zcomplex, void)

const char* la_b200_version(void) { return "la_b200 0.1.0 (sm_100a)"; }

int64_t la_b200_workspace_bytes(void) {
  Context* c = thread_context(nullptr);
  return (int64_t)c->arena.total;
}

int la_b200_last_gemm_path(void) { return last_gemm_path(); }

int64_t la_b200_release_workspace(void) { return (int64_t)release_workspaces(); }

}  // extern "C"
