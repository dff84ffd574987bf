// la_b200 — Cholesky factorisation and solve.
//
// Reference call sites replaced: ?potrf_ / ?potrs_ / ?posv_ (src/linalg/cholesky.cr:47-90,
// src/linalg/linalg.cr:433-434; PD detection src/matrix/flag_checks.cr:32-40).
//
// Two-level right-looking algorithm.  Inside an outer panel of NBO columns, a one-CTA leaf kernel
// factors a diagonal block of NB (128) columns, a tall TRSM solves the rows below it and a
// rank-NB masked GEMM updates the rest of the panel — every step is parallel over the rows
// below, only the leaf runs on a single SM.  The outer trailing matrix then takes one rank-NBO
// HERK on the tensor cores, which is where almost all of the n^3/3 flops are.
// Only the named triangle of A is read or written (cholesky.cr:50-64 relies on that).
#include "la_internal.h"
#include "tri_small.cuh"

namespace la {

namespace {

template <typename T> struct PotfCfg { static constexpr int NB = 128; };
template <> struct PotfCfg<zcomplex> { static constexpr int NB = 64; };
constexpr int POTF_IB = 32;
constexpr int POTF_NT = 256;

template <typename T> __device__ __forceinline__ T shfl_T(T v, int src) {
  return __shfl_sync(0xffffffffu, v, src);
}
template <> __device__ __forceinline__ zcomplex shfl_T<zcomplex>(zcomplex v, int src) {
  return zcomplex{__shfl_sync(0xffffffffu, v.x, src), __shfl_sync(0xffffffffu, v.y, src)};
}

template <typename T>
constexpr size_t potf2_smem_bytes() {
  return sizeof(T) * ((size_t)PotfCfg<T>::NB * (PotfCfg<T>::NB + 1) + POTF_IB * (POTF_IB + 1) +
                      (size_t)POTF_IB * PotfCfg<T>::NB + POTF_IB);
}

// Leaf: one CTA factors an n x n diagonal block (n <= NB) held in shared memory, 32 columns at a
// time:
//   (1) warp 0 factors the 32x32 diagonal block column by column (__syncwarp only);
//   (2) one thread per row below solves x * D^H = b with the row in registers and stores the
//       result in place and transposed (Pt) for conflict-free reads;
//   (3) all threads apply the rank-32 update to the trailing lower triangle in 4x4 tiles.
// The block is always factored as lower; for uplo = U the conjugate transpose is loaded and
// stored.  On a non-positive pivot, info = info_off + (1-based column) is set once; columns
// of earlier 32-blocks stay factored, as LAPACK leaves a partially factored matrix.
template <typename T>
__global__ void __launch_bounds__(POTF_NT)
potf2_kernel(int lower, int n, T* __restrict__ A, int64_t lda, int* info, int info_off) {
  using R = typename Traits<T>::real;
  constexpr int NB = PotfCfg<T>::NB, LDS = NB + 1, LDM = POTF_IB + 1;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T* S = reinterpret_cast<T*>(smem_raw);  // S[i * LDS + j], lower triangle
  T* Ms = S + NB * LDS;                   // conj(D) of the current diagonal block
  T* Pt = Ms + POTF_IB * LDM;             // Pt[c * NB + i]: solved panel, transposed
  T* rdiag = Pt + POTF_IB * NB;           // 1 / diag(D)
  __shared__ int fail;
  if (*info != 0) return;  // an earlier block already failed: LAPACK would have stopped there
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) fail = 0;
  for (int idx = tid; idx < n * n; idx += POTF_NT) {
    int i = idx % n, j = idx / n;
    if (lower) { if (i >= j) S[i * LDS + j] = A[i + (int64_t)j * lda]; }
    else       { if (i <= j) S[j * LDS + i] = conj_(A[i + (int64_t)j * lda]); }
  }
  __syncthreads();
  int done_cols = n;
  for (int j0 = 0; j0 < n; j0 += POTF_IB) {
    const int jb = min(POTF_IB, n - j0);
    if (warp == 0) {  // ---- (1) diagonal block: lane i keeps row i in registers
      T* D = S + j0 * LDS + j0;
      T a[POTF_IB];
#pragma unroll
      for (int c = 0; c < POTF_IB; c++)
        a[c] = (lane < jb && c <= lane) ? D[lane * LDS + c] : zero<T>();
      int f = 0;
#pragma unroll
      for (int j = 0; j < POTF_IB; j++) {
        if (j < jb && !f) {
          const double d = __shfl_sync(0xffffffffu, (double)real_(a[j]), j);
          if (!(d > 0.0)) {
            f = j0 + j + 1;  // also catches NaN; uniform over the warp
          } else {
            // sqrt and reciprocal sqrt are independent, so neither waits for the other
            const double sq = sqrt(d);
            const double rs = rsqrt(d);
            const R rinv = (R)rs;
            if (lane == j) { a[j] = from_real<T>(sq); rdiag[j] = from_real<T>(rs); }
            if (lane > j) a[j] = scale_real(a[j], rinv);
            // a(i,k) -= a(i,j) * conj(a(k,j)) for j < k <= i; a(k,j) lives in lane k
#pragma unroll
            for (int k = j + 1; k < POTF_IB; k++) {
              const T akj = shfl_T(a[j], k);
              if (lane >= k) {
                T v = fnma_(a[j], conj_(akj), a[k]);
                if (lane == k) v = from_real<T>((double)real_(v));
                a[k] = v;
              }
            }
          }
        }
      }
      if (f && lane == 0) fail = f;
      if (!f && lane < jb) {
#pragma unroll
        for (int c = 0; c < POTF_IB; c++)
          if (c < jb) {
            if (c <= lane) D[lane * LDS + c] = a[c];
            Ms[lane * LDM + c] = c <= lane ? conj_(a[c]) : zero<T>();
          }
      }
    }
    __syncthreads();
    if (fail) { done_cols = j0; break; }
    const int rem = n - j0 - jb;
    if (rem <= 0) break;
    if (tid < rem) {  // ---- (2) x * D^H = b  <=>  conj(D) x^T = b^T, one thread per row
      const int i = j0 + jb + tid;
      T x[POTF_IB];
#pragma unroll
      for (int c = 0; c < POTF_IB; c++) x[c] = c < jb ? S[i * LDS + j0 + c] : zero<T>();
      solve_vec_regs<T, POTF_IB, LDM>(Ms, rdiag, jb, true, false, one<T>(), x);
#pragma unroll
      for (int c = 0; c < POTF_IB; c++)
        if (c < jb) { S[i * LDS + j0 + c] = x[c]; Pt[c * NB + tid] = x[c]; }
    }
    __syncthreads();
    // ---- (3) trailing lower triangle -= P P^H in 4x4 tiles (tile row >= tile column)
    const int nt = (rem + 3) / 4;
    for (int t = tid; t < nt * nt; t += POTF_NT) {
      const int ti = t % nt, tk = t / nt;
      if (ti < tk) continue;
      T acc[4][4];
#pragma unroll
      for (int a = 0; a < 4; a++)
#pragma unroll
        for (int b = 0; b < 4; b++) acc[a][b] = zero<T>();
      for (int c = 0; c < jb; c++) {
        T pa[4], pb[4];
#pragma unroll
        for (int a = 0; a < 4; a++) {
          pa[a] = (4 * ti + a < rem) ? Pt[c * NB + 4 * ti + a] : zero<T>();
          pb[a] = (4 * tk + a < rem) ? conj_(Pt[c * NB + 4 * tk + a]) : zero<T>();
        }
#pragma unroll
        for (int a = 0; a < 4; a++)
#pragma unroll
          for (int b = 0; b < 4; b++) acc[a][b] = fma_(pa[a], pb[b], acc[a][b]);
      }
#pragma unroll
      for (int a = 0; a < 4; a++)
#pragma unroll
        for (int b = 0; b < 4; b++) {
          const int li = 4 * ti + a, lk = 4 * tk + b;
          if (li < rem && lk < rem && li >= lk) {
            const int i = j0 + jb + li, k = j0 + jb + lk;
            T v = sub(S[i * LDS + k], acc[a][b]);
            if (i == k) v = from_real<T>((double)real_(v));
            S[i * LDS + k] = v;
          }
        }
    }
    __syncthreads();
  }
  if (fail && tid == 0) atomicCAS(info, 0, info_off + fail);
  for (int idx = tid; idx < n * n; idx += POTF_NT) {
    int i = idx % n, j = idx / n;
    if (lower) { if (i >= j && j < done_cols) A[i + (int64_t)j * lda] = S[i * LDS + j]; }
    else       { if (i <= j && i < done_cols) A[i + (int64_t)j * lda] = conj_(S[j * LDS + i]); }
  }
}

}  // namespace

template <typename T>
void potrf(Context& ctx, Uplo uplo, int64_t n, T* A, int64_t lda, int* d_info) {
  LA_CUDA_CHECK(cudaMemsetAsync(d_info, 0, sizeof(int), ctx.stream));
  if (n <= 0) return;
  static bool attr_set = false;
  if (!attr_set) {
    LA_CUDA_CHECK(cudaFuncSetAttribute(potf2_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)potf2_smem_bytes<T>()));
    attr_set = true;
  }
  constexpr int NB = PotfCfg<T>::NB;
  const int64_t NBO = n >= 8192 ? 1024 : (n >= 2048 ? 512 : 256);
  const size_t smem = potf2_smem_bytes<T>();
  auto at = [&](int64_t i, int64_t j) { return A + i + j * lda; };
  const T mone = from_real<T>(-1.0);
  for (int64_t K0 = 0; K0 < n; K0 += NBO) {
    const int64_t kb = NBO < n - K0 ? NBO : n - K0;
    for (int64_t j0 = K0; j0 < K0 + kb; j0 += NB) {
      const int64_t jb = NB < K0 + kb - j0 ? NB : K0 + kb - j0;
      potf2_kernel<T><<<1, POTF_NT, smem, ctx.stream>>>(uplo == Lower, (int)jb, at(j0, j0), lda,
                                                       d_info, (int)j0);
      LA_CUDA_CHECK(cudaGetLastError());
      const int64_t r = n - j0 - jb;         // rows (L) / columns (U) beyond the diagonal block
      const int64_t pc = K0 + kb - j0 - jb;  // those of them that lie inside the outer panel
      if (r <= 0) continue;
      if (uplo == Lower) {
        trsm<T>(ctx, Right, Lower, OpC, NonUnit, r, jb, one<T>(), at(j0, j0), lda, at(j0 + jb, j0), lda);
        if (pc > 0)  // A[j0+jb:n, j0+jb:K0+kb] -= P P(0:pc)^H, lower part only
          gemm<T>(ctx, OpN, OpC, r, pc, jb, mone, at(j0 + jb, j0), lda, at(j0 + jb, j0), lda,
                  one<T>(), at(j0 + jb, j0 + jb), lda, TriLower, Traits<T>::is_complex);
      } else {
        trsm<T>(ctx, Left, Upper, OpC, NonUnit, jb, r, one<T>(), at(j0, j0), lda, at(j0, j0 + jb), lda);
        if (pc > 0)
          gemm<T>(ctx, OpC, OpN, pc, r, jb, mone, at(j0, j0 + jb), lda, at(j0, j0 + jb), lda,
                  one<T>(), at(j0 + jb, j0 + jb), lda, TriUpper, Traits<T>::is_complex);
      }
    }
    const int64_t r2 = n - K0 - kb;
    if (r2 > 0) {
      if (uplo == Lower)
        herk<T>(ctx, Lower, OpN, r2, kb, -1.0, at(K0 + kb, K0), lda, 1.0, at(K0 + kb, K0 + kb), lda);
      else
        herk<T>(ctx, Upper, OpC, r2, kb, -1.0, at(K0, K0 + kb), lda, 1.0, at(K0 + kb, K0 + kb), lda);
    }
  }
}

template <typename T>
void potrs(Context& ctx, Uplo uplo, int64_t n, int64_t nrhs, const T* A, int64_t lda, T* B,
           int64_t ldb) {
  if (n <= 0 || nrhs <= 0) return;
  if (uplo == Lower) {
    trsm<T>(ctx, Left, Lower, OpN, NonUnit, n, nrhs, one<T>(), A, lda, B, ldb);
    trsm<T>(ctx, Left, Lower, OpC, NonUnit, n, nrhs, one<T>(), A, lda, B, ldb);
  } else {
    trsm<T>(ctx, Left, Upper, OpC, NonUnit, n, nrhs, one<T>(), A, lda, B, ldb);
    trsm<T>(ctx, Left, Upper, OpN, NonUnit, n, nrhs, one<T>(), A, lda, B, ldb);
  }
}

#define LA_INST_CHOL(T)                                                   \
  template void potrf<T>(Context&, Uplo, int64_t, T*, int64_t, int*);     \
  template void potrs<T>(Context&, Uplo, int64_t, int64_t, const T*, int64_t, T*, int64_t);
LA_INST_CHOL(float)
LA_INST_CHOL(double)
LA_INST_CHOL(zcomplex)

}  // namespace la
