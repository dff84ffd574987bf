// la_b200 — matrix inverses on top of the factors:
// ?trtri_, ?getri_, ?potri_ as reached from GeneralMatrix#inv! (src/linalg/linalg.cr:366-397).
// Each is a triangular/LU/Cholesky solve against the identity held in workspace; only the
// part of A that LAPACK defines on exit is written back.
#include "la_internal.h"

namespace la {

namespace {
template <typename T>
__global__ void diag_zero_min_kernel(int64_t n, const T* __restrict__ A, int64_t lda, int* info) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && is_zero(A[i + i * lda])) atomicMin(info, (int)(i + 1));
}
__global__ void info_begin_kernel(int* info) { *info = 0x7fffffff; }
__global__ void info_end_kernel(int* info) { if (*info == 0x7fffffff) *info = 0; }

template <typename T>
void singular_check(Context& ctx, int64_t n, const T* A, int64_t lda, int* d_info) {
  info_begin_kernel<<<1, 1, 0, ctx.stream>>>(d_info);
  diag_zero_min_kernel<T><<<(unsigned)ceil_div(n, 256), 256, 0, ctx.stream>>>(n, A, lda, d_info);
  info_end_kernel<<<1, 1, 0, ctx.stream>>>(d_info);
  LA_CUDA_CHECK(cudaGetLastError());
}

// copy X to A where *info == 0 (LAPACK leaves A untouched on a singular factor);
// tri: 0 full, TriLower / TriUpper restrict the written part
template <typename T>
__global__ void guarded_copy_kernel(int tri, int64_t n, const T* __restrict__ X, int64_t ldx,
                                    T* __restrict__ A, int64_t lda, const int* info) {
  if (*info != 0) return;
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  for (int64_t j = blockIdx.y; j < n; j += gridDim.y) {
    if (tri == TriLower && i < j) continue;
    if (tri == TriUpper && i > j) continue;
    A[i + j * lda] = X[i + j * ldx];
  }
}
template <typename T>
void guarded_copy(Context& ctx, int tri, int64_t n, const T* X, int64_t ldx, T* A, int64_t lda,
                  const int* d_info) {
  dim3 g((unsigned)ceil_div(n, 256), (unsigned)(n < 1024 ? n : 1024));
  guarded_copy_kernel<T><<<g, 256, 0, ctx.stream>>>(tri, n, X, ldx, A, lda, d_info);
  LA_CUDA_CHECK(cudaGetLastError());
}
}  // namespace

template <typename T>
void trtri(Context& ctx, Uplo uplo, Diag diag, int64_t n, T* A, int64_t lda, int* d_info) {
  LA_CUDA_CHECK(cudaMemsetAsync(d_info, 0, sizeof(int), ctx.stream));
  if (n <= 0) return;
  if (diag == NonUnit) singular_check<T>(ctx, n, A, lda, d_info);
  Arena::Mark mk = ctx.arena.mark();
  const int64_t ldx = round_up(n, 16);
  T* X = ctx.alloc<T>(ldx * n);
  set_matrix<T>(ctx, n, n, zero<T>(), one<T>(), X, ldx);
  trsm<T>(ctx, Left, uplo, OpN, diag, n, n, one<T>(), A, lda, X, ldx);
  guarded_copy<T>(ctx, uplo == Lower ? TriLower : TriUpper, n, X, ldx, A, lda, d_info);
  ctx.arena.release(mk);
}

template <typename T>
void getri(Context& ctx, int64_t n, T* A, int64_t lda, const int* d_ipiv, int* d_info) {
  LA_CUDA_CHECK(cudaMemsetAsync(d_info, 0, sizeof(int), ctx.stream));
  if (n <= 0) return;
  singular_check<T>(ctx, n, A, lda, d_info);
  Arena::Mark mk = ctx.arena.mark();
  const int64_t ldx = round_up(n, 16);
  T* X = ctx.alloc<T>(ldx * n);
  set_matrix<T>(ctx, n, n, zero<T>(), one<T>(), X, ldx);
  getrs<T>(ctx, OpN, n, n, A, lda, d_ipiv, X, ldx);
  guarded_copy<T>(ctx, TriFull, n, X, ldx, A, lda, d_info);
  ctx.arena.release(mk);
}

template <typename T>
void potri(Context& ctx, Uplo uplo, int64_t n, T* A, int64_t lda, int* d_info) {
  LA_CUDA_CHECK(cudaMemsetAsync(d_info, 0, sizeof(int), ctx.stream));
  if (n <= 0) return;
  singular_check<T>(ctx, n, A, lda, d_info);
  Arena::Mark mk = ctx.arena.mark();
  const int64_t ldx = round_up(n, 16);
  T* X = ctx.alloc<T>(ldx * n);
  set_matrix<T>(ctx, n, n, zero<T>(), one<T>(), X, ldx);
  potrs<T>(ctx, uplo, n, n, A, lda, X, ldx);
  guarded_copy<T>(ctx, uplo == Lower ? TriLower : TriUpper, n, X, ldx, A, lda, d_info);
  ctx.arena.release(mk);
}

#define LA_INST_INV(T)                                                              \
  template void trtri<T>(Context&, Uplo, Diag, int64_t, T*, int64_t, int*);         \
  template void getri<T>(Context&, int64_t, T*, int64_t, const int*, int*);         \
  template void potri<T>(Context&, Uplo, int64_t, T*, int64_t, int*);
LA_INST_INV(float)
LA_INST_INV(double)
LA_INST_INV(zcomplex)

}  // namespace la
