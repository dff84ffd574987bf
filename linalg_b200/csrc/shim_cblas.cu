// la_b200 — CBLAS drop-in symbols (host pointers).  See include/la_b200.h, Part 1.
//
// Replaces the `fun`s the reference binds in src/linalg/libCBLAS.cr:174-203 and calls from
// src/linalg/mult.cr:6-15 (`blas` macro: always COL_MAJOR, NoTrans).  Row-major calls are
// mapped onto the column-major kernels by the usual transposition identities.
#include "shim_common.h"
#include <cstdlib>
#include <thread>
#include <utility>
#include <vector>

namespace la {

static void cblas_error(const char* name, int arg) {
  // CBLAS has no error return (cblas_xerbla is bound at libCBLAS.cr:206 and never called by
  // the reference); report and leave the output untouched, as the reference BLAS does.
  fprintf(stderr, "la_b200: %s: parameter %d had an illegal value\n", name, arg);
}

// GPUs one drop-in product may use: LA_B200_GPUS (default 1 — the reference's host calls one
// library, and a process normally owns one GPU; set it to let the big products of an unchanged
// host spread over the GPUs of the box).  Clamped to the number of visible devices.
static int gemm_gpu_count() {
  static const int n = [] {
    const char* e = getenv("LA_B200_GPUS");
    int want = e ? atoi(e) : 1, have = 0;
    if (want <= 1 || cudaGetDeviceCount(&have) != cudaSuccess) return 1;
    return want < have ? want : have;
  }();
  return n;
}
static thread_local bool tl_gemm_worker = false;   // set in the per-GPU threads of a split product

template <typename T>
static void h_gemm(const char* name, int order, int ta, int tb, int m, int n, int k, T alpha,
                   const T* A, int lda, const T* B, int ldb, T beta, T* C, int ldc) {
  if (order == LA_ROW_MAJOR_) {
    // C^T = op(B)^T op(A)^T in column-major terms
    std::swap(ta, tb); std::swap(m, n); std::swap(A, B); std::swap(lda, ldb);
  } else if (order != LA_COL_MAJOR_) return cblas_error(name, 1);
  if (!valid_op(ta)) return cblas_error(name, 2);
  if (!valid_op(tb)) return cblas_error(name, 3);
  if (m < 0) return cblas_error(name, 4);
  if (n < 0) return cblas_error(name, 5);
  if (k < 0) return cblas_error(name, 6);
  const int64_t ar = ta == OpN ? m : k, ac = ta == OpN ? k : m;
  const int64_t br = tb == OpN ? k : n, bc = tb == OpN ? n : k;
  if (lda < (ar > 1 ? ar : 1)) return cblas_error(name, 9);
  if (ldb < (br > 1 ? br : 1)) return cblas_error(name, 11);
  if (ldc < (m > 1 ? m : 1)) return cblas_error(name, 14);
  if (m == 0 || n == 0) return;
  if ((is_zero(alpha) || k == 0) && is_one(beta)) return;
  // Several GPUs (LA_B200_GPUS > 1): the columns of C and of op(B) are dealt in contiguous blocks
  // to one host thread per GPU, each of which runs the single-GPU path below on its block (its
  // own PCIe link for its share of B and C and for A; no data-path exchange between the GPUs).
  if (!tl_gemm_worker && gemm_gpu_count() > 1 && !is_zero(alpha) && k > 0 &&
      (double)m * (double)n * (double)k >= 6.4e10 && n >= 1024 * gemm_gpu_count()) {
    const int ng = gemm_gpu_count();
    int cur = 0;
    LA_CUDA_CHECK(cudaGetDevice(&cur));
    const int64_t nb = round_up(ceil_div((int64_t)n, (int64_t)ng), 128);
    std::vector<std::thread> workers;
    for (int d = 0; d < ng; d++) {
      const int64_t c0 = (int64_t)d * nb;
      if (c0 >= n) break;
      const int cn = (int)(nb < n - c0 ? nb : n - c0);
      const T* Bd = tb == OpN ? B + c0 * (int64_t)ldb : B + c0;
      T* Cd = C + c0 * (int64_t)ldc;
      workers.emplace_back([=] {
        LA_CUDA_CHECK(cudaSetDevice(d));
        tl_gemm_worker = true;
        h_gemm<T>(name, LA_COL_MAJOR_, ta, tb, m, cn, k, alpha, A, lda, Bd, ldb, beta, Cd, ldc);
      });
    }
    for (auto& w : workers) w.join();
    LA_CUDA_CHECK(cudaSetDevice(cur));
    return;
  }
  ContextLease ctx;
  Arena::Mark mk = ctx->arena.mark();
  const bool need_ab = !is_zero(alpha) && k > 0;
  DevMat<T> dA, dB;
  DevMat<T> dC = dev_alloc<T>(*ctx, m, n);
  // Large products are pipelined over column blocks of C: while block j is being multiplied,
  // block j+1 of B (and of C when beta != 0) travels host->device and block j-1 of C travels
  // back, so that only the upload of A and of the first block is exposed.
  const double work = (double)m * (double)n * (double)k;
  const int64_t nchunks = (need_ab && work >= 8e9 && n >= 2048) ? 8 : 1;
  if (nchunks == 1) {
    if (need_ab) {
      dA = dev_from_host<T>(*ctx, ar, ac, A, lda);
      dB = dev_from_host<T>(*ctx, br, bc, B, ldb);
    }
    if (!is_zero(beta)) upload<T>(*ctx, dC, C, ldc);
    gemm<T>(*ctx, (Op)ta, (Op)tb, m, n, need_ab ? k : 0, alpha, dA.p, dA.ld, dB.p, dB.ld, beta,
            dC.p, dC.ld);
    download<T>(*ctx, dC, C, ldc);
    sync(*ctx);
  } else {
    dA = dev_alloc<T>(*ctx, ar, ac);
    dB = dev_alloc<T>(*ctx, br, bc);
    const int64_t nb = round_up(ceil_div(n, nchunks), 128);
    cudaStream_t up = ctx->copy_stream, down = ctx->d2h_stream;
    cudaEvent_t ev_up[16], ev_done[16], ev_a[16];
    // A travels in k-chunks BEHIND the first block of B and C, and the first column block is
    // multiplied chunk by chunk as they land: only B0, C0 and the first k-chunk of A are exposed
    // instead of the whole of A.
    const int64_t kc = 4096;
    const int nka = (k >= 2 * kc) ? (int)(ceil_div(k, kc) < 16 ? ceil_div(k, kc) : 16) : 1;
    const int64_t kstep = nka == 1 ? k : round_up(ceil_div(k, (int64_t)nka), 128);
    int nc = 0, na_ev = 0;
    for (int64_t c0 = 0; c0 < n; c0 += nb, nc++) {
      const int64_t cn = nb < n - c0 ? nb : n - c0;
      LA_CUDA_CHECK(cudaEventCreateWithFlags(&ev_up[nc], cudaEventDisableTiming));
      LA_CUDA_CHECK(cudaEventCreateWithFlags(&ev_done[nc], cudaEventDisableTiming));
      // columns [c0, c0+cn) of op(B): a column block of B (tb == N) or a row block (otherwise)
      DevMat<T> vB = dB, vC = dC;
      const T* hB;
      if (tb == OpN) { vB.p = dB.p + c0 * dB.ld; vB.rows = k; vB.cols = cn; hB = B + c0 * (int64_t)ldb; }
      else           { vB.p = dB.p + c0;         vB.rows = cn; vB.cols = k; hB = B + c0; }
      upload<T>(*ctx, vB, hB, ldb, up);
      vC.p = dC.p + c0 * dC.ld; vC.cols = cn;
      if (!is_zero(beta)) upload<T>(*ctx, vC, C + c0 * (int64_t)ldc, ldc, up);
      LA_CUDA_CHECK(cudaEventRecord(ev_up[nc], up));
      LA_CUDA_CHECK(cudaStreamWaitEvent(ctx->stream, ev_up[nc], 0));
      if (nc == 0) {
        for (int l = 0; l < nka; l++) {
          const int64_t l0 = l * kstep;
          const int64_t kl = kstep < k - l0 ? kstep : k - l0;
          if (kl <= 0) break;
          // k-chunk l of op(A): columns of A (ta == N) or rows (otherwise), same for op(B)
          DevMat<T> vA = dA;
          const T* hA;
          if (ta == OpN) { vA.p = dA.p + l0 * dA.ld; vA.rows = m;  vA.cols = kl; hA = A + l0 * (int64_t)lda; }
          else           { vA.p = dA.p + l0;         vA.rows = kl; vA.cols = m;  hA = A + l0; }
          upload<T>(*ctx, vA, hA, lda, up);
          LA_CUDA_CHECK(cudaEventCreateWithFlags(&ev_a[l], cudaEventDisableTiming));
          na_ev = l + 1;
          LA_CUDA_CHECK(cudaEventRecord(ev_a[l], up));
          LA_CUDA_CHECK(cudaStreamWaitEvent(ctx->stream, ev_a[l], 0));
          const T* pB = tb == OpN ? vB.p + l0 : vB.p + l0 * dB.ld;
          gemm<T>(*ctx, (Op)ta, (Op)tb, m, cn, kl, alpha, vA.p, dA.ld, pB, dB.ld,
                  l == 0 ? beta : one<T>(), vC.p, dC.ld);
        }
      } else {
        gemm<T>(*ctx, (Op)ta, (Op)tb, m, cn, k, alpha, dA.p, dA.ld, vB.p, dB.ld, beta, vC.p, dC.ld);
      }
      LA_CUDA_CHECK(cudaEventRecord(ev_done[nc], ctx->stream));
      LA_CUDA_CHECK(cudaStreamWaitEvent(down, ev_done[nc], 0));
      download<T>(*ctx, vC, C + c0 * (int64_t)ldc, ldc, down);
    }
    LA_CUDA_CHECK(cudaStreamSynchronize(down));
    sync(*ctx);
    for (int i = 0; i < nc; i++) { cudaEventDestroy(ev_up[i]); cudaEventDestroy(ev_done[i]); }
    for (int l = 0; l < na_ev; l++) cudaEventDestroy(ev_a[l]);
  }
  ctx->arena.release(mk);
}

template <typename T>
static void h_symm(const char* name, bool herm, int order, int side, int uplo, int m, int n,
                   T alpha, const T* A, int lda, const T* B, int ldb, T beta, T* C, int ldc) {
  if (order == LA_ROW_MAJOR_) {
    // row-major C (m x n) is column-major C^T (n x m): A on the other side, other triangle
    side = side == Left ? Right : (side == Right ? Left : side);
    uplo = uplo == Upper ? Lower : (uplo == Lower ? Upper : uplo);
    std::swap(m, n);
  } else if (order != LA_COL_MAJOR_) return cblas_error(name, 1);
  if (!valid_side(side)) return cblas_error(name, 2);
  if (!valid_uplo(uplo)) return cblas_error(name, 3);
  if (m < 0) return cblas_error(name, 4);
  if (n < 0) return cblas_error(name, 5);
  const int64_t na = side == Left ? m : n;
  if (lda < (na > 1 ? na : 1)) return cblas_error(name, 8);
  if (ldb < (m > 1 ? m : 1)) return cblas_error(name, 10);
  if (ldc < (m > 1 ? m : 1)) return cblas_error(name, 13);
  if (m == 0 || n == 0) return;
  ContextLease ctx;
  Arena::Mark mk = ctx->arena.mark();
  // Row-major note: the column-major view of a row-major symmetric/Hermitian operand is its
  // transpose, which is exactly the factor the transposed product needs (C^T = B^T A^T), so
  // swapping side and triangle above is the whole mapping, Hermitian included.
  DevMat<T> dA = dev_from_host<T>(*ctx, na, na, A, lda);
  DevMat<T> dB = dev_from_host<T>(*ctx, m, n, B, ldb);
  DevMat<T> dC = dev_alloc<T>(*ctx, m, n);
  if (!is_zero(beta)) upload<T>(*ctx, dC, C, ldc);
  symm<T>(*ctx, herm, (Side)side, (Uplo)uplo, m, n, alpha, dA.p, dA.ld, dB.p, dB.ld, beta, dC.p,
          dC.ld);
  download<T>(*ctx, dC, C, ldc);
  sync(*ctx);
  ctx->arena.release(mk);
}

// trmm (solve == false) and trsm (solve == true) share validation and staging
template <typename T>
static void h_trxm(const char* name, bool solve, int order, int side, int uplo, int ta, int diag,
                   int m, int n, T alpha, const T* A, int lda, T* B, int ldb) {
  if (order == LA_ROW_MAJOR_) {
    side = side == Left ? Right : (side == Right ? Left : side);
    uplo = uplo == Upper ? Lower : (uplo == Lower ? Upper : uplo);
    std::swap(m, n);
  } else if (order != LA_COL_MAJOR_) return cblas_error(name, 1);
  if (!valid_side(side)) return cblas_error(name, 2);
  if (!valid_uplo(uplo)) return cblas_error(name, 3);
  if (!valid_op(ta)) return cblas_error(name, 4);
  if (!valid_diag(diag)) return cblas_error(name, 5);
  if (m < 0) return cblas_error(name, 6);
  if (n < 0) return cblas_error(name, 7);
  const int64_t na = side == Left ? m : n;
  if (lda < (na > 1 ? na : 1)) return cblas_error(name, 10);
  if (ldb < (m > 1 ? m : 1)) return cblas_error(name, 12);
  if (m == 0 || n == 0) return;
  ContextLease ctx;
  Arena::Mark mk = ctx->arena.mark();
  DevMat<T> dA = dev_from_host<T>(*ctx, na, na, A, lda);
  DevMat<T> dB = dev_from_host<T>(*ctx, m, n, B, ldb);
  if (solve)
    trsm<T>(*ctx, (Side)side, (Uplo)uplo, (Op)ta, (Diag)diag, m, n, alpha, dA.p, dA.ld, dB.p, dB.ld);
  else
    trmm<T>(*ctx, (Side)side, (Uplo)uplo, (Op)ta, (Diag)diag, m, n, alpha, dA.p, dA.ld, dB.p, dB.ld);
  download<T>(*ctx, dB, B, ldb);
  sync(*ctx);
  ctx->arena.release(mk);
}

// syrk (herm == false, real types) / herk (herm == true)
template <typename T>
static void h_herk(const char* name, int order, int uplo, int trans, int n, int k, double alpha,
                   const T* A, int lda, double beta, T* C, int ldc) {
  if (order == LA_ROW_MAJOR_) {
    uplo = uplo == Upper ? Lower : (uplo == Lower ? Upper : uplo);
    // row-major A (n x k, trans N) is column-major k x n: A A^H = (A_cm)^T conj(A_cm); for the
    // real routines that is trans flipped; for herk the result is the conjugate, which on the
    // transposed (swapped-triangle) storage is again the right Hermitian matrix.
    trans = trans == OpN ? OpC : OpN;
  } else if (order != LA_COL_MAJOR_) return cblas_error(name, 1);
  if (!valid_uplo(uplo)) return cblas_error(name, 2);
  if (!valid_op(trans)) return cblas_error(name, 3);
  if (n < 0) return cblas_error(name, 4);
  if (k < 0) return cblas_error(name, 5);
  const int64_t ar = trans == OpN ? n : k, ac = trans == OpN ? k : n;
  if (lda < (ar > 1 ? ar : 1)) return cblas_error(name, 8);
  if (ldc < (n > 1 ? n : 1)) return cblas_error(name, 11);
  if (n == 0) return;
  ContextLease ctx;
  Arena::Mark mk = ctx->arena.mark();
  DevMat<T> dA = dev_from_host<T>(*ctx, ar, ac, A, lda);
  DevMat<T> dC = dev_from_host<T>(*ctx, n, n, C, ldc);
  herk<T>(*ctx, (Uplo)uplo, trans == OpN ? OpN : OpC, n, k, alpha, dA.p, dA.ld, beta, dC.p, dC.ld);
  // the masked GEMM writes only the named triangle of dC; the rest still holds C's
  // original values, so a full download leaves the other triangle unchanged on the host
  download<T>(*ctx, dC, C, ldc);
  sync(*ctx);
  ctx->arena.release(mk);
}

}  // namespace la

using namespace la;

extern "C" {

void cblas_sgemm(int order, int ta, int tb, int m, int n, int k, float alpha, const float* a,
                 int lda, const float* b, int ldb, float beta, float* c, int ldc) {
  h_gemm<float>("cblas_sgemm", order, ta, tb, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc);
}
void cblas_dgemm(int order, int ta, int tb, int m, int n, int k, double alpha, const double* a,
                 int lda, const double* b, int ldb, double beta, double* c, int ldc) {
  h_gemm<double>("cblas_dgemm", order, ta, tb, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc);
}
void cblas_zgemm(int order, int ta, int tb, int m, int n, int k, const void* alpha, const void* a,
                 int lda, const void* b, int ldb, const void* beta, void* c, int ldc) {
  h_gemm<zcomplex>("cblas_zgemm", order, ta, tb, m, n, k, *(const zcomplex*)alpha,
                   (const zcomplex*)a, lda, (const zcomplex*)b, ldb, *(const zcomplex*)beta,
                   (zcomplex*)c, ldc);
}
void cblas_ssymm(int order, int side, int uplo, int m, int n, float alpha, const float* a, int lda,
                 const float* b, int ldb, float beta, float* c, int ldc) {
  h_symm<float>("cblas_ssymm", false, order, side, uplo, m, n, alpha, a, lda, b, ldb, beta, c, ldc);
}
void cblas_dsymm(int order, int side, int uplo, int m, int n, double alpha, const double* a,
                 int lda, const double* b, int ldb, double beta, double* c, int ldc) {
  h_symm<double>("cblas_dsymm", false, order, side, uplo, m, n, alpha, a, lda, b, ldb, beta, c, ldc);
}
void cblas_zsymm(int order, int side, int uplo, int m, int n, const void* alpha, const void* a,
                 int lda, const void* b, int ldb, const void* beta, void* c, int ldc) {
  h_symm<zcomplex>("cblas_zsymm", false, order, side, uplo, m, n, *(const zcomplex*)alpha,
                   (const zcomplex*)a, lda, (const zcomplex*)b, ldb, *(const zcomplex*)beta,
                   (zcomplex*)c, ldc);
}
void cblas_zhemm(int order, int side, int uplo, int m, int n, const void* alpha, const void* a,
                 int lda, const void* b, int ldb, const void* beta, void* c, int ldc) {
  h_symm<zcomplex>("cblas_zhemm", true, order, side, uplo, m, n, *(const zcomplex*)alpha,
                   (const zcomplex*)a, lda, (const zcomplex*)b, ldb, *(const zcomplex*)beta,
                   (zcomplex*)c, ldc);
}
void cblas_strmm(int order, int side, int uplo, int ta, int diag, int m, int n, float alpha,
                 const float* a, int lda, float* b, int ldb) {
  h_trxm<float>("cblas_strmm", false, order, side, uplo, ta, diag, m, n, alpha, a, lda, b, ldb);
}
void cblas_dtrmm(int order, int side, int uplo, int ta, int diag, int m, int n, double alpha,
                 const double* a, int lda, double* b, int ldb) {
  h_trxm<double>("cblas_dtrmm", false, order, side, uplo, ta, diag, m, n, alpha, a, lda, b, ldb);
}
void cblas_ztrmm(int order, int side, int uplo, int ta, int diag, int m, int n, const void* alpha,
                 const void* a, int lda, void* b, int ldb) {
  h_trxm<zcomplex>("cblas_ztrmm", false, order, side, uplo, ta, diag, m, n,
                   *(const zcomplex*)alpha, (const zcomplex*)a, lda, (zcomplex*)b, ldb);
}
void cblas_strsm(int order, int side, int uplo, int ta, int diag, int m, int n, float alpha,
                 const float* a, int lda, float* b, int ldb) {
  h_trxm<float>("cblas_strsm", true, order, side, uplo, ta, diag, m, n, alpha, a, lda, b, ldb);
}
void cblas_dtrsm(int order, int side, int uplo, int ta, int diag, int m, int n, double alpha,
                 const double* a, int lda, double* b, int ldb) {
  h_trxm<double>("cblas_dtrsm", true, order, side, uplo, ta, diag, m, n, alpha, a, lda, b, ldb);
}
void cblas_ztrsm(int order, int side, int uplo, int ta, int diag, int m, int n, const void* alpha,
                 const void* a, int lda, void* b, int ldb) {
  h_trxm<zcomplex>("cblas_ztrsm", true, order, side, uplo, ta, diag, m, n, *(const zcomplex*)alpha,
                   (const zcomplex*)a, lda, (zcomplex*)b, ldb);
}
void cblas_ssyrk(int order, int uplo, int trans, int n, int k, float alpha, const float* a, int lda,
                 float beta, float* c, int ldc) {
  h_herk<float>("cblas_ssyrk", order, uplo, trans, n, k, alpha, a, lda, beta, c, ldc);
}
void cblas_dsyrk(int order, int uplo, int trans, int n, int k, double alpha, const double* a,
                 int lda, double beta, double* c, int ldc) {
  h_herk<double>("cblas_dsyrk", order, uplo, trans, n, k, alpha, a, lda, beta, c, ldc);
}
void cblas_zherk(int order, int uplo, int trans, int n, int k, double alpha, const void* a, int lda,
                 double beta, void* c, int ldc) {
  h_herk<zcomplex>("cblas_zherk", order, uplo, trans, n, k, alpha, (const zcomplex*)a, lda, beta,
                   (zcomplex*)c, ldc);
}

void openblas_set_num_threads(int) {}
int openblas_get_num_threads(void) { return 1; }
int openblas_get_num_procs(void) { return 1; }
char* openblas_get_config(void) {
  static char cfg[] = "la_b200 0.1 CUDA sm_100a (B200) DMMA/tcgen05 backend";
  return cfg;
}
char* openblas_get_corename(void) {
  static char core[] = "B200";
  return core;
}
int openblas_get_parallel(void) { return 0; }

}  // extern "C"
