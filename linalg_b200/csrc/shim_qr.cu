// la_b200 — LAPACK drop-in symbols for the QR / least-squares family (host pointers):
// ?geqrf_, ?ormqr_/zunmqr_, ?orgqr_/zungqr_, ?gels_.  Workspace protocol as driven by the
// reference's `lapack` macro for WORK_DETECT routines (src/linalg/lapack_helper.cr:143-183,
// :215-280): lwork = -1 returns the size in work[0]; the real call never writes past work[0]
// (the debug guard of src/linalg/workpool.cr:37-49 would trip otherwise).
#include "shim_common.h"

namespace la {

static inline int64_t imax(int64_t a, int64_t b) { return a > b ? a : b; }
static inline int64_t imin(int64_t a, int64_t b) { return a < b ? a : b; }
template <typename T>
static void set_work(T* work, double v) {
  if (work) work[0] = from_real<T>(v);
}
constexpr int kNB = 32;  // the block size the sizes below are quoted for, as LAPACK does

template <typename T>
static void h_geqrf(const int* pm, const int* pn, T* A, const int* plda, T* tau, T* work,
                    const int* plwork, int* info) {
  const int64_t m = *pm, n = *pn, lda = *plda, lwork = *plwork;
  *info = 0;
  if (m < 0) { *info = -1; return; }
  if (n < 0) { *info = -2; return; }
  if (lda < imax(1, m)) { *info = -4; return; }
  if (lwork != -1 && lwork < imax(1, n)) { *info = -7; return; }
  const double opt = (double)imax(1, n * kNB);
  if (lwork == -1) { set_work<T>(work, opt); return; }
  set_work<T>(work, opt);
  const int64_t k = imin(m, n);
  if (k == 0) return;
  ContextLease ctx;
  Arena::Mark mk = ctx->arena.mark();
  DevMat<T> dA = dev_from_host<T>(*ctx, m, n, A, lda);
  T* d_tau = ctx->alloc<T>(k);
  geqrf<T>(*ctx, m, n, dA.p, dA.ld, d_tau);
  download<T>(*ctx, dA, A, lda);
  LA_CUDA_CHECK(cudaMemcpyAsync(tau, d_tau, k * sizeof(T), cudaMemcpyDeviceToHost, ctx->stream));
  sync(*ctx);
  ctx->arena.release(mk);
}

template <typename T>
static void h_ormqr(const char* pside, const char* ptrans, const int* pm, const int* pn,
                    const int* pk, const T* A, const int* plda, const T* tau, T* C,
                    const int* pldc, T* work, const int* plwork, int* info) {
  const int64_t m = *pm, n = *pn, k = *pk, lda = *plda, ldc = *pldc, lwork = *plwork;
  const int side = side_from_char(*pside);
  int trans = op_from_char(*ptrans);
  *info = 0;
  // real routines accept N/T, the complex one N/C
  if (Traits<T>::is_complex ? trans == OpT : trans == OpC) trans = 0;
  const int64_t nq = side == Left ? m : n, nw = side == Left ? n : m;
  if (!side) { *info = -1; return; }
  if (!trans) { *info = -2; return; }
  if (m < 0) { *info = -3; return; }
  if (n < 0) { *info = -4; return; }
  if (k < 0 || k > nq) { *info = -5; return; }
  if (lda < imax(1, nq)) { *info = -7; return; }
  if (ldc < imax(1, m)) { *info = -10; return; }
  if (lwork != -1 && lwork < imax(1, nw)) { *info = -12; return; }
  const double opt = (double)imax(1, nw * kNB);
  if (lwork == -1) { set_work<T>(work, opt); return; }
  set_work<T>(work, opt);
  if (m == 0 || n == 0 || k == 0) return;
  ContextLease ctx;
  Arena::Mark mk = ctx->arena.mark();
  DevMat<T> dA = dev_from_host<T>(*ctx, nq, k, A, lda);
  DevMat<T> dC = dev_from_host<T>(*ctx, m, n, C, ldc);
  T* d_tau = ctx->alloc<T>(k);
  LA_CUDA_CHECK(cudaMemcpyAsync(d_tau, tau, k * sizeof(T), cudaMemcpyHostToDevice, ctx->stream));
  ormqr<T>(*ctx, (Side)side, trans == OpN ? OpN : OpC, m, n, k, dA.p, dA.ld, d_tau, dC.p, dC.ld);
  download<T>(*ctx, dC, C, ldc);
  sync(*ctx);
  ctx->arena.release(mk);
}

template <typename T>
static void h_orgqr(const int* pm, const int* pn, const int* pk, T* A, const int* plda,
                    const T* tau, T* work, const int* plwork, int* info) {
  const int64_t m = *pm, n = *pn, k = *pk, lda = *plda, lwork = *plwork;
  *info = 0;
  if (m < 0) { *info = -1; return; }
  if (n < 0 || n > m) { *info = -2; return; }
  if (k < 0 || k > n) { *info = -3; return; }
  if (lda < imax(1, m)) { *info = -5; return; }
  if (lwork != -1 && lwork < imax(1, n)) { *info = -8; return; }
  const double opt = (double)imax(1, n * kNB);
  if (lwork == -1) { set_work<T>(work, opt); return; }
  set_work<T>(work, opt);
  if (n == 0) return;
  ContextLease ctx;
  Arena::Mark mk = ctx->arena.mark();
  DevMat<T> dA = dev_from_host<T>(*ctx, m, n, A, lda);
  T* d_tau = ctx->alloc<T>(imax(k, 1));
  if (k > 0)
    LA_CUDA_CHECK(cudaMemcpyAsync(d_tau, tau, k * sizeof(T), cudaMemcpyHostToDevice, ctx->stream));
  orgqr<T>(*ctx, m, n, k, dA.p, dA.ld, d_tau);
  download<T>(*ctx, dA, A, lda);
  sync(*ctx);
  ctx->arena.release(mk);
}

template <typename T>
static void h_gels(const char* ptrans, const int* pm, const int* pn, const int* pnrhs, T* A,
                   const int* plda, T* B, const int* pldb, T* work, const int* plwork,
                   int* info) {
  const int64_t m = *pm, n = *pn, nrhs = *pnrhs, lda = *plda, ldb = *pldb, lwork = *plwork;
  int trans = op_from_char(*ptrans);
  if (Traits<T>::is_complex ? trans == OpT : trans == OpC) trans = 0;
  const int64_t mn = imin(m, n);
  *info = 0;
  if (!trans) { *info = -1; return; }
  if (m < 0) { *info = -2; return; }
  if (n < 0) { *info = -3; return; }
  if (nrhs < 0) { *info = -4; return; }
  if (lda < imax(1, m)) { *info = -6; return; }
  if (ldb < imax(1, imax(m, n))) { *info = -8; return; }
  if (lwork != -1 && lwork < imax(1, mn + imax(mn, nrhs))) { *info = -10; return; }
  const double opt = (double)imax(1, mn + imax(mn, nrhs) * kNB);
  if (lwork == -1) { set_work<T>(work, opt); return; }
  set_work<T>(work, opt);
  const int64_t brows = imax(m, n);
  if (mn == 0 || nrhs == 0) {
    // LAPACK: quick return sets B (max(m,n) x nrhs) to zero
    for (int64_t j = 0; j < nrhs; j++)
      for (int64_t i = 0; i < brows; i++) B[i + j * ldb] = zero<T>();
    return;
  }
  ContextLease ctx;
  Arena::Mark mk = ctx->arena.mark();
  DevMat<T> dA = dev_from_host<T>(*ctx, m, n, A, lda);
  // B holds the right-hand sides in its first (trans==N ? m : n) rows and receives the
  // solution in its first (trans==N ? n : m) rows; stage all max(m,n) rows
  DevMat<T> dB = dev_alloc<T>(*ctx, brows, nrhs);
  {
    const int64_t in_rows = trans == OpN ? m : n;
    DevMat<T> view = dB;
    view.rows = in_rows;
    upload<T>(*ctx, view, B, ldb);
  }
  gels<T>(*ctx, trans == OpN ? OpN : OpC, m, n, nrhs, dA.p, dA.ld, dB.p, dB.ld, ctx->d_info);
  download<T>(*ctx, dA, A, lda);
  const int inf = read_info(*ctx, ctx->d_info);
  if (inf == 0) {
    download<T>(*ctx, dB, B, ldb);
    sync(*ctx);
  }
  *info = inf;
  ctx->arena.release(mk);
}

// ?gerqf_ / ?gelqf_ / ?geqlf_ (src/linalg/rq_lq_ql.cr:21-27): same argument list and workspace
// protocol as ?geqrf_
template <typename T>
static void h_geqrf_variant(char which, const int* pm, const int* pn, T* A, const int* plda, T* tau,
                            T* work, const int* plwork, int* info) {
  const int64_t m = *pm, n = *pn, lda = *plda, lwork = *plwork;
  *info = 0;
  if (m < 0) { *info = -1; return; }
  if (n < 0) { *info = -2; return; }
  if (lda < imax(1, m)) { *info = -4; return; }
  const int64_t wmin = which == 'Q' ? n : m;       // QL works on columns, RQ / LQ on rows
  if (lwork != -1 && lwork < imax(1, wmin)) { *info = -7; return; }
  const double opt = (double)imax(1, wmin * kNB);
  if (lwork == -1) { set_work<T>(work, opt); return; }
  set_work<T>(work, opt);
  const int64_t k = imin(m, n);
  if (k == 0) return;
  ContextLease ctx;
  Arena::Mark mk = ctx->arena.mark();
  DevMat<T> dA = dev_from_host<T>(*ctx, m, n, A, lda);
  T* d_tau = ctx->alloc<T>(k);
  geqrf_variant<T>(*ctx, which, m, n, dA.p, dA.ld, d_tau);
  download<T>(*ctx, dA, A, lda);
  LA_CUDA_CHECK(cudaMemcpyAsync(tau, d_tau, k * sizeof(T), cudaMemcpyDeviceToHost, ctx->stream));
  sync(*ctx);
  ctx->arena.release(mk);
}

// ?orgrq_ / ?orglq_ / ?orgql_ and zung* (src/linalg/rq_lq_ql.cr:12)
template <typename T>
static void h_orgqr_variant(char which, const int* pm, const int* pn, const int* pk, T* A,
                            const int* plda, const T* tau, T* work, const int* plwork, int* info) {
  const int64_t m = *pm, n = *pn, k = *pk, lda = *plda, lwork = *plwork;
  *info = 0;
  const bool rows = which != 'Q';                  // LQ / RQ generate orthonormal ROWS (m <= n)
  if (m < 0) { *info = -1; return; }
  if (rows ? n < m : (n < 0 || n > m)) { *info = -2; return; }
  if (k < 0 || k > (rows ? m : n)) { *info = -3; return; }
  if (lda < imax(1, m)) { *info = -5; return; }
  const int64_t wmin = rows ? m : n;
  if (lwork != -1 && lwork < imax(1, wmin)) { *info = -8; return; }
  const double opt = (double)imax(1, wmin * kNB);
  if (lwork == -1) { set_work<T>(work, opt); return; }
  set_work<T>(work, opt);
  if (m == 0 || n == 0) return;
  ContextLease ctx;
  Arena::Mark mk = ctx->arena.mark();
  DevMat<T> dA = dev_from_host<T>(*ctx, m, n, A, lda);
  T* d_tau = ctx->alloc<T>(imax(k, 1));
  if (k > 0)
    LA_CUDA_CHECK(cudaMemcpyAsync(d_tau, tau, k * sizeof(T), cudaMemcpyHostToDevice, ctx->stream));
  orgqr_variant<T>(*ctx, which, m, n, k, dA.p, dA.ld, d_tau);
  download<T>(*ctx, dA, A, lda);
  sync(*ctx);
  ctx->arena.release(mk);
}

// ?geqp3_ (src/linalg/qr.cr:38-41).  jpvt(j) != 0 on entry marks a column that must stay in
// front (LAPACK's "fixed" columns): they are swapped to the front here, as ?geqp3 does, and
// factored without pivot search.
template <typename T>
static void h_geqp3(const int* pm, const int* pn, T* A, const int* plda, int* jpvt, T* tau, T* work,
                    const int* plwork, int* info) {
  const int64_t m = *pm, n = *pn, lda = *plda, lwork = *plwork;
  *info = 0;
  if (m < 0) { *info = -1; return; }
  if (n < 0) { *info = -2; return; }
  if (lda < imax(1, m)) { *info = -4; return; }
  const int64_t k = imin(m, n);
  const int64_t wmin = k == 0 ? 1 : (Traits<T>::is_complex ? n + 1 : 3 * n + 1);
  if (lwork != -1 && lwork < wmin) { *info = -8; return; }
  const double opt = k == 0 ? 1.0 : (double)((Traits<T>::is_complex ? 1 : 2) * n + (n + 1) * kNB);
  if (lwork == -1) { set_work<T>(work, opt); return; }
  set_work<T>(work, opt);
  if (k == 0) return;
  ContextLease ctx;
  Arena::Mark mk = ctx->arena.mark();
  DevMat<T> dA = dev_from_host<T>(*ctx, m, n, A, lda);
  // fixed columns to the front (dgeqp3.f: "Move initial columns up front")
  int64_t nfixed = 0;
  for (int64_t j = 0; j < n; j++) {
    if (jpvt[j] != 0) {
      if (j != nfixed) {
        swap_columns<T>(*ctx, m, dA.p + j * dA.ld, dA.p + nfixed * dA.ld);
        jpvt[j] = jpvt[nfixed];
        jpvt[nfixed] = (int)j + 1;
      } else {
        jpvt[j] = (int)j + 1;
      }
      nfixed++;
    } else {
      jpvt[j] = (int)j + 1;
    }
  }
  if (nfixed > k) nfixed = k;
  int* d_jpvt = ctx->alloc<int>(n);
  T* d_tau = ctx->alloc<T>(k);
  LA_CUDA_CHECK(cudaMemcpyAsync(d_jpvt, jpvt, n * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
  geqp3<T>(*ctx, m, n, dA.p, dA.ld, d_jpvt, d_tau, nfixed);
  download<T>(*ctx, dA, A, lda);
  LA_CUDA_CHECK(cudaMemcpyAsync(jpvt, d_jpvt, n * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  LA_CUDA_CHECK(cudaMemcpyAsync(tau, d_tau, k * sizeof(T), cudaMemcpyDeviceToHost, ctx->stream));
  sync(*ctx);
  ctx->arena.release(mk);
}

}  // namespace la

using namespace la;

#define LA_LAPACK_QR_SYMBOLS(X, T, VT, ORM, ORG)                                                  \
  void X##geqrf_(const int* m, const int* n, VT* a, const int* lda, VT* tau, VT* work,            \
                 const int* lwork, int* info) {                                                   \
    h_geqrf<T>(m, n, (T*)a, lda, (T*)tau, (T*)work, lwork, info);                                 \
  }                                                                                               \
  void ORM(const char* side, const char* trans, const int* m, const int* n, const int* k,         \
           const VT* a, const int* lda, const VT* tau, VT* c, const int* ldc, VT* work,           \
           const int* lwork, int* info) {                                                         \
    h_ormqr<T>(side, trans, m, n, k, (const T*)a, lda, (const T*)tau, (T*)c, ldc, (T*)work,       \
               lwork, info);                                                                      \
  }                                                                                               \
  void ORG(const int* m, const int* n, const int* k, VT* a, const int* lda, const VT* tau,        \
           VT* work, const int* lwork, int* info) {                                               \
    h_orgqr<T>(m, n, k, (T*)a, lda, (const T*)tau, (T*)work, lwork, info);                        \
  }                                                                                               \
  void X##gels_(const char* trans, const int* m, const int* n, const int* nrhs, VT* a,            \
                const int* lda, VT* b, const int* ldb, VT* work, const int* lwork, int* info) {   \
    h_gels<T>(trans, m, n, nrhs, (T*)a, lda, (T*)b, ldb, (T*)work, lwork, info);                  \
  }

#define LA_LAPACK_QR_VARIANTS(X, T, VT, ORGRQ, ORGLQ, ORGQL)                                       \
  void X##gerqf_(const int* m, const int* n, VT* a, const int* lda, VT* tau, VT* work,            \
                 const int* lwork, int* info) {                                                   \
    h_geqrf_variant<T>('R', m, n, (T*)a, lda, (T*)tau, (T*)work, lwork, info);                    \
  }                                                                                               \
  void X##gelqf_(const int* m, const int* n, VT* a, const int* lda, VT* tau, VT* work,            \
                 const int* lwork, int* info) {                                                   \
    h_geqrf_variant<T>('L', m, n, (T*)a, lda, (T*)tau, (T*)work, lwork, info);                    \
  }                                                                                               \
  void X##geqlf_(const int* m, const int* n, VT* a, const int* lda, VT* tau, VT* work,            \
                 const int* lwork, int* info) {                                                   \
    h_geqrf_variant<T>('Q', m, n, (T*)a, lda, (T*)tau, (T*)work, lwork, info);                    \
  }                                                                                               \
  void ORGRQ(const int* m, const int* n, const int* k, VT* a, const int* lda, const VT* tau,      \
             VT* work, const int* lwork, int* info) {                                             \
    h_orgqr_variant<T>('R', m, n, k, (T*)a, lda, (const T*)tau, (T*)work, lwork, info);           \
  }                                                                                               \
  void ORGLQ(const int* m, const int* n, const int* k, VT* a, const int* lda, const VT* tau,      \
             VT* work, const int* lwork, int* info) {                                             \
    h_orgqr_variant<T>('L', m, n, k, (T*)a, lda, (const T*)tau, (T*)work, lwork, info);           \
  }                                                                                               \
  void ORGQL(const int* m, const int* n, const int* k, VT* a, const int* lda, const VT* tau,      \
             VT* work, const int* lwork, int* info) {                                             \
    h_orgqr_variant<T>('Q', m, n, k, (T*)a, lda, (const T*)tau, (T*)work, lwork, info);           \
  }

extern "C" {
LA_LAPACK_QR_VARIANTS(s, float, float, sorgrq_, sorglq_, sorgql_)
LA_LAPACK_QR_VARIANTS(d, double, double, dorgrq_, dorglq_, dorgql_)
LA_LAPACK_QR_VARIANTS(z, zcomplex, void, zungrq_, zunglq_, zungql_)
void sgeqp3_(const int* m, const int* n, float* a, const int* lda, int* jpvt, float* tau,
             float* work, const int* lwork, int* info) {
  h_geqp3<float>(m, n, a, lda, jpvt, tau, work, lwork, info);
}
void dgeqp3_(const int* m, const int* n, double* a, const int* lda, int* jpvt, double* tau,
             double* work, const int* lwork, int* info) {
  h_geqp3<double>(m, n, a, lda, jpvt, tau, work, lwork, info);
}
// (rwork, 2n doubles, is LAPACK's scratch for the partial norms: not used here)
void zgeqp3_(const int* m, const int* n, void* a, const int* lda, int* jpvt, void* tau, void* work,
             const int* lwork, double* rwork, int* info) {
  (void)rwork;
  h_geqp3<zcomplex>(m, n, (zcomplex*)a, lda, jpvt, (zcomplex*)tau, (zcomplex*)work, lwork, info);
}
LA_LAPACK_QR_SYMBOLS(s, float, float, sormqr_, sorgqr_)
LA_LAPACK_QR_SYMBOLS(d, double, double, dormqr_, dorgqr_)
LA_LAPACK_QR_SYMBOLS(z, zcomplex, void, zunmqr_, zungqr_)
}  // extern "C"
