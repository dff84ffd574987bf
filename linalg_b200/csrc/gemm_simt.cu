// la_b200 — generic register-tiled SIMT GEMM for s/d/z, any transposition, any
// leading dimension or alignment.  This is the always-correct path behind
// cblas_?gemm (reference call site: src/linalg/mult.cr:64-69 -> libCBLAS.cr:174-178);
// the tensor-core kernels in gemm_dmma.cu / gemm_tf32x3.cu take over whenever the
// operands satisfy their TMA alignment rules.
#include "la_internal.h"

namespace la {

namespace {

constexpr int BM = 64, BN = 64, BK = 16, TM = 4, TN = 4, NT = 256;

template <typename T>
__global__ void __launch_bounds__(NT)
gemm_simt_kernel(int ta, int tb, int64_t m, int64_t n, int64_t k, T alpha,
                 const T* __restrict__ A, int64_t lda, const T* __restrict__ B, int64_t ldb,
                 T beta, T* __restrict__ C, int64_t ldc, int tri, int herm_diag) {
  __shared__ T As[BK][BM + 1];
  __shared__ T Bs[BK][BN + 1];
  const int64_t m0 = (int64_t)blockIdx.x * BM, n0 = (int64_t)blockIdx.y * BN;
  if (tri == TriLower && m0 + BM - 1 < n0) return;
  if (tri == TriUpper && m0 > n0 + BN - 1) return;
  const int tid = threadIdx.x;
  const int tx = tid % 16, ty = tid / 16;
  T acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; i++)
#pragma unroll
    for (int j = 0; j < TN; j++) acc[i][j] = zero<T>();

  const bool a_n = (ta == OpN), b_n = (tb == OpN);
  const bool a_c = (ta == OpC), b_c = (tb == OpC);

  for (int64_t k0 = 0; k0 < k; k0 += BK) {
    // ---- stage A tile: As[kk][i] = op(A)(m0+i, k0+kk)
    if (a_n) {
#pragma unroll
      for (int p = 0; p < BM * BK / NT; p++) {
        int idx = tid + p * NT;
        int i = idx % BM, kk = idx / BM;
        int64_t gi = m0 + i, gk = k0 + kk;
        As[kk][i] = (gi < m && gk < k) ? A[gi + gk * lda] : zero<T>();
      }
    } else {
#pragma unroll
      for (int p = 0; p < BM * BK / NT; p++) {
        int idx = tid + p * NT;
        int kk = idx % BK, i = idx / BK;
        int64_t gi = m0 + i, gk = k0 + kk;
        T v = (gi < m && gk < k) ? A[gk + gi * lda] : zero<T>();
        As[kk][i] = apply_conj(v, a_c);
      }
    }
    // ---- stage B tile: Bs[kk][j] = op(B)(k0+kk, n0+j)
    if (b_n) {
#pragma unroll
      for (int p = 0; p < BN * BK / NT; p++) {
        int idx = tid + p * NT;
        int kk = idx % BK, j = idx / BK;
        int64_t gj = n0 + j, gk = k0 + kk;
        Bs[kk][j] = (gj < n && gk < k) ? B[gk + gj * ldb] : zero<T>();
      }
    } else {
#pragma unroll
      for (int p = 0; p < BN * BK / NT; p++) {
        int idx = tid + p * NT;
        int j = idx % BN, kk = idx / BN;
        int64_t gj = n0 + j, gk = k0 + kk;
        T v = (gj < n && gk < k) ? B[gj + gk * ldb] : zero<T>();
        Bs[kk][j] = apply_conj(v, b_c);
      }
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < BK; kk++) {
      T a[TM], b[TN];
#pragma unroll
      for (int i = 0; i < TM; i++) a[i] = As[kk][tx + 16 * i];
#pragma unroll
      for (int j = 0; j < TN; j++) b[j] = Bs[kk][ty + 16 * j];
#pragma unroll
      for (int i = 0; i < TM; i++)
#pragma unroll
        for (int j = 0; j < TN; j++) acc[i][j] = fma_(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }

  const bool beta_zero = is_zero(beta);
#pragma unroll
  for (int j = 0; j < TN; j++) {
    int64_t c = n0 + ty + 16 * j;
    if (c >= n) continue;
#pragma unroll
    for (int i = 0; i < TM; i++) {
      int64_t r = m0 + tx + 16 * i;
      if (r >= m) continue;
      if (tri == TriLower && r < c) continue;
      if (tri == TriUpper && r > c) continue;
      T v = mul(alpha, acc[i][j]);
      if (!beta_zero) v = fma_(beta, C[r + c * ldc], v);
      if (herm_diag && r == c) v = from_real<T>((double)real_(v));
      C[r + c * ldc] = v;
    }
  }
}

}  // namespace

template <typename T>
void gemm_simt(Context& ctx, Op ta, Op tb, int64_t m, int64_t n, int64_t k, T alpha, const T* A,
               int64_t lda, const T* B, int64_t ldb, T beta, T* C, int64_t ldc, int tri,
               bool herm_diag) {
  if (m <= 0 || n <= 0) return;
  if (!Traits<T>::is_complex) {  // 'C' == 'T' for real types
    if (ta == OpC) ta = OpT;
    if (tb == OpC) tb = OpT;
  }
  // grid.y is limited to 65535: loop over column super-blocks if needed
  const int64_t max_cols = (int64_t)65535 * BN;
  for (int64_t c0 = 0; c0 < n; c0 += max_cols) {
    int64_t nc = n - c0 < max_cols ? n - c0 : max_cols;
    const T* Bp = (tb == OpN) ? B + c0 * ldb : B + c0;
    // the triangle mask is relative to C's origin: only valid for a single super-block
    dim3 grid((unsigned)ceil_div(m, BM), (unsigned)ceil_div(nc, BN));
    gemm_simt_kernel<T><<<grid, NT, 0, ctx.stream>>>((int)ta, (int)tb, m, nc, k, alpha, A, lda, Bp,
                                                     ldb, beta, C + c0 * ldc, ldc,
                                                     c0 == 0 ? tri : TriFull, herm_diag ? 1 : 0);
    LA_CUDA_CHECK(cudaGetLastError());
  }
}

template void gemm_simt<float>(Context&, Op, Op, int64_t, int64_t, int64_t, float, const float*,
                               int64_t, const float*, int64_t, float, float*, int64_t, int, bool);
template void gemm_simt<double>(Context&, Op, Op, int64_t, int64_t, int64_t, double,
                                const double*, int64_t, const double*, int64_t, double, double*,
                                int64_t, int, bool);
template void gemm_simt<zcomplex>(Context&, Op, Op, int64_t, int64_t, int64_t, zcomplex,
                                  const zcomplex*, int64_t, const zcomplex*, int64_t, zcomplex,
                                  zcomplex*, int64_t, int, bool);

}  // namespace la
