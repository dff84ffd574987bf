// la_b200 — level-3 BLAS on the device: GEMM dispatch, HERK/SYRK, TRSM, TRMM,
// SYMM/HEMM and the O(n^2) layout helpers they need.
//
// Reference call sites these replace: cblas_?gemm / ?symm / zhemm / ?trmm from
// src/linalg/mult.cr:26-91 (libCBLAS.cr:174-201); TRSM/HERK are the trailing-update
// kernels inside ?getrf/?potrf/?getrs/?potrs/?trtrs.
#include "la_internal.h"
#include "tri_small.cuh"
#include <type_traits>

namespace la {

// trsm_dmma.cu
bool trsm_base_dmma_f64(Context& ctx, Side side, Uplo uplo, Op op, Diag diag, int64_t m, int64_t n,
                        double alpha, const double* A, int64_t lda, double* B, int64_t ldb);

// ------------------------------------------------------------------ GEMM dispatch
int& last_gemm_path() {
  static thread_local int path = 0;
  return path;
}

// op(A)^H-times-B products with a tiny result and a very long reduction (block reflectors of a
// tall-skinny QR) have their own streaming kernel
template <typename T>
static bool try_tall_gram(Context& ctx, Op ta, Op tb, int64_t m, int64_t n, int64_t k, T alpha,
                          const T* A, int64_t lda, const T* B, int64_t ldb, T beta, T* C,
                          int64_t ldc, int tri) {
  const bool conj_a = ta == OpC || (ta == OpT && !Traits<T>::is_complex);
  if (!conj_a || tb != OpN || tri != TriFull || m > 64 || n > 64 || k < 8192) return false;
  if (!is_one(alpha) || !is_zero(beta)) return false;
  if (!tall_gram<T>(ctx, m, n, k, A, lda, B, ldb, C, ldc)) return false;
  last_gemm_path() = 4;
  return true;
}

template <>
void gemm<double>(Context& ctx, Op ta, Op tb, int64_t m, int64_t n, int64_t k, double alpha,
                  const double* A, int64_t lda, const double* B, int64_t ldb, double beta,
                  double* C, int64_t ldc, int tri, bool herm_diag) {
  if (m <= 0 || n <= 0) return;
  if (try_tall_gram<double>(ctx, ta, tb, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, tri)) return;
  if (gemm_dmma_f64(ctx, ta, tb, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, tri)) {
    last_gemm_path() = 1;
    return;
  }
  last_gemm_path() = 0;
  gemm_simt<double>(ctx, ta, tb, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, tri, false);
}
template <>
void gemm<zcomplex>(Context& ctx, Op ta, Op tb, int64_t m, int64_t n, int64_t k, zcomplex alpha,
                    const zcomplex* A, int64_t lda, const zcomplex* B, int64_t ldb,
                    zcomplex beta, zcomplex* C, int64_t ldc, int tri, bool herm_diag) {
  if (m <= 0 || n <= 0) return;
  if (try_tall_gram<zcomplex>(ctx, ta, tb, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, tri)) return;
  if (gemm_dmma_z(ctx, ta, tb, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, tri, herm_diag)) {
    last_gemm_path() = 2;
    return;
  }
  last_gemm_path() = 0;
  gemm_simt<zcomplex>(ctx, ta, tb, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, tri, herm_diag);
}
template <>
void gemm<float>(Context& ctx, Op ta, Op tb, int64_t m, int64_t n, int64_t k, float alpha,
                 const float* A, int64_t lda, const float* B, int64_t ldb, float beta, float* C,
                 int64_t ldc, int tri, bool herm_diag) {
  if (m <= 0 || n <= 0) return;
  if (try_tall_gram<float>(ctx, ta, tb, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, tri)) return;
  if (gemm_tf32x3(ctx, ta, tb, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, tri)) {
    last_gemm_path() = 3;
    return;
  }
  last_gemm_path() = 0;
  gemm_simt<float>(ctx, ta, tb, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, tri, false);
}

// Product restricted to the blocks on or below a block staircase (la_internal.h: StairMask).
// FP64 on TMA-legal operands: ONE launch of the tensor-core kernel that skips the tiles above the
// staircase; otherwise one product per block column, from its first active row block down.
template <typename T>
void gemm_stair(Context& ctx, Op ta, Op tb, int64_t m, int64_t n, int64_t k, T alpha, const T* A,
                int64_t lda, const T* B, int64_t ldb, T beta, T* C, int64_t ldc,
                const StairMask& mask) {
  if (m <= 0 || n <= 0) return;
  if constexpr (std::is_same<T, double>::value) {
    if (gemm_dmma_f64_ex(ctx, ta, tb, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, TriFull, 0,
                         false, &mask)) {
      last_gemm_path() = 1;
      return;
    }
  }
  const int64_t nb = mask.nb;
  for (int64_t j0 = 0; j0 < n; j0 += nb) {
    const int64_t jb = nb < n - j0 ? nb : n - j0;
    const int64_t cb = mask.col_first + (j0 / nb) * mask.col_step;
    // first local row block whose global index reaches cb
    int64_t li = 0;
    if (cb > mask.row_first) li = ceil_div(cb - mask.row_first, mask.row_step);
    const int64_t i0 = li * nb;
    if (i0 >= m) continue;
    const T* Ab = ta == OpN ? A + i0 : A + i0 * lda;
    const T* Bb = tb == OpN ? B + j0 * ldb : B + j0;
    gemm<T>(ctx, ta, tb, m - i0, jb, k, alpha, Ab, lda, Bb, ldb, beta, C + i0 + j0 * ldc, ldc);
  }
}
template void gemm_stair<float>(Context&, Op, Op, int64_t, int64_t, int64_t, float, const float*,
                                int64_t, const float*, int64_t, float, float*, int64_t,
                                const StairMask&);
template void gemm_stair<double>(Context&, Op, Op, int64_t, int64_t, int64_t, double,
                                 const double*, int64_t, const double*, int64_t, double, double*,
                                 int64_t, const StairMask&);
template void gemm_stair<zcomplex>(Context&, Op, Op, int64_t, int64_t, int64_t, zcomplex,
                                   const zcomplex*, int64_t, const zcomplex*, int64_t, zcomplex,
                                   zcomplex*, int64_t, const StairMask&);

template <typename T>
void herk(Context& ctx, Uplo uplo, Op trans, int64_t n, int64_t k, double alpha, const T* A,
          int64_t lda, double beta, T* C, int64_t ldc) {
  // C = alpha*A*A^H + beta*C (trans==N) or alpha*A^H*A + beta*C, one triangle only.
  int tri = uplo == Lower ? TriLower : TriUpper;
  if (trans == OpN)
    gemm<T>(ctx, OpN, OpC, n, n, k, from_real<T>(alpha), A, lda, A, lda, from_real<T>(beta), C,
            ldc, tri, Traits<T>::is_complex);
  else
    gemm<T>(ctx, OpC, OpN, n, n, k, from_real<T>(alpha), A, lda, A, lda, from_real<T>(beta), C,
            ldc, tri, Traits<T>::is_complex);
}
template void herk<float>(Context&, Uplo, Op, int64_t, int64_t, double, const float*, int64_t,
                          double, float*, int64_t);
template void herk<double>(Context&, Uplo, Op, int64_t, int64_t, double, const double*, int64_t,
                           double, double*, int64_t);
template void herk<zcomplex>(Context&, Uplo, Op, int64_t, int64_t, double, const zcomplex*,
                             int64_t, double, zcomplex*, int64_t);

// ------------------------------------------------------------------ elementwise helpers
namespace {

template <typename T>
__global__ void copy_kernel(int64_t m, int64_t n, const T* __restrict__ A, int64_t lda,
                            T* __restrict__ B, int64_t ldb) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  for (int64_t j = blockIdx.y; j < n; j += gridDim.y) B[i + j * ldb] = A[i + j * lda];
}

template <typename T>
__global__ void transpose_kernel(int conj, int64_t m, int64_t n, const T* __restrict__ A,
                                 int64_t lda, T* __restrict__ B, int64_t ldb) {
  __shared__ T tile[32][33];
  int64_t nbx = ceil_div(m, 32);
  for (int64_t blk = blockIdx.x; blk < nbx * ((n + 31) / 32); blk += gridDim.x) {
    int64_t i0 = (blk % nbx) * 32, j0 = (blk / nbx) * 32;
    for (int jj = threadIdx.y; jj < 32; jj += blockDim.y) {
      int64_t i = i0 + threadIdx.x, j = j0 + jj;
      if (i < m && j < n) tile[jj][threadIdx.x] = A[i + j * lda];
    }
    __syncthreads();
    for (int ii = threadIdx.y; ii < 32; ii += blockDim.y) {
      int64_t j = j0 + threadIdx.x, i = i0 + ii;
      if (i < m && j < n) B[j + i * ldb] = apply_conj(tile[threadIdx.x][ii], conj != 0);
    }
    __syncthreads();
  }
}

template <typename T>
__global__ void set_kernel(int64_t m, int64_t n, T offdiag, T diag, T* A, int64_t lda) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  for (int64_t j = blockIdx.y; j < n; j += gridDim.y) A[i + j * lda] = (i == j) ? diag : offdiag;
}

template <typename T>
__global__ void copy_triangle_kernel(int lower, int unit, int64_t m, int64_t n,
                                     const T* __restrict__ A, int64_t lda, T* __restrict__ B,
                                     int64_t ldb) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  for (int64_t j = blockIdx.y; j < n; j += gridDim.y) {
    T v;
    if (i == j) v = unit ? one<T>() : A[i + j * lda];
    else if ((lower && i > j) || (!lower && i < j)) v = A[i + j * lda];
    else v = zero<T>();
    B[i + j * ldb] = v;
  }
}

template <typename T>
__global__ void symmetrize_kernel(int herm, int lower, int64_t n, const T* __restrict__ A,
                                  int64_t lda, T* __restrict__ B, int64_t ldb) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  for (int64_t j = blockIdx.y; j < n; j += gridDim.y) {
    bool stored = lower ? (i >= j) : (i <= j);
    T v = stored ? A[i + j * lda] : apply_conj(A[j + i * lda], herm != 0);
    if (herm && i == j) v = from_real<T>((double)real_(v));
    B[i + j * ldb] = v;
  }
}

inline dim3 grid2d(int64_t m, int64_t n) {
  return dim3((unsigned)ceil_div(m, 256), (unsigned)(n < 1024 ? (n < 1 ? 1 : n) : 1024));
}

}  // namespace

template <typename T>
void copy_matrix(Context& ctx, int64_t m, int64_t n, const T* A, int64_t lda, T* B, int64_t ldb) {
  if (m <= 0 || n <= 0) return;
  copy_kernel<T><<<grid2d(m, n), 256, 0, ctx.stream>>>(m, n, A, lda, B, ldb);
  LA_CUDA_CHECK(cudaGetLastError());
}
template <typename T>
void transpose_matrix(Context& ctx, bool conj, int64_t m, int64_t n, const T* A, int64_t lda,
                      T* B, int64_t ldb) {
  if (m <= 0 || n <= 0) return;
  int64_t nblk = ceil_div(m, 32) * ceil_div(n, 32);
  unsigned g = (unsigned)(nblk < 65535 * 16 ? nblk : 65535 * 16);
  transpose_kernel<T><<<g, dim3(32, 8), 0, ctx.stream>>>(conj ? 1 : 0, m, n, A, lda, B, ldb);
  LA_CUDA_CHECK(cudaGetLastError());
}
template <typename T>
void set_matrix(Context& ctx, int64_t m, int64_t n, T offdiag, T diag, T* A, int64_t lda) {
  if (m <= 0 || n <= 0) return;
  set_kernel<T><<<grid2d(m, n), 256, 0, ctx.stream>>>(m, n, offdiag, diag, A, lda);
  LA_CUDA_CHECK(cudaGetLastError());
}
template <typename T>
void copy_triangle(Context& ctx, Uplo uplo, bool unit, int64_t m, int64_t n, const T* A,
                   int64_t lda, T* B, int64_t ldb) {
  if (m <= 0 || n <= 0) return;
  copy_triangle_kernel<T><<<grid2d(m, n), 256, 0, ctx.stream>>>(uplo == Lower, unit ? 1 : 0, m, n,
                                                                A, lda, B, ldb);
  LA_CUDA_CHECK(cudaGetLastError());
}
template <typename T>
void symmetrize(Context& ctx, bool herm, Uplo uplo, int64_t n, const T* A, int64_t lda, T* B,
                int64_t ldb) {
  if (n <= 0) return;
  symmetrize_kernel<T><<<grid2d(n, n), 256, 0, ctx.stream>>>(herm ? 1 : 0, uplo == Lower, n, A,
                                                             lda, B, ldb);
  LA_CUDA_CHECK(cudaGetLastError());
}

// ------------------------------------------------------------------ TRSM
namespace {

// Base case of the recursive TRSM: one thread per right-hand-side vector (a column of B for
// side = Left, a row for side = Right) solves against a triangular block of order <= NB held
// in shared memory, with the vector in registers (tri_small.cuh).
template <typename T> struct TrsmCfg { static constexpr int NB = 64, NT = 64; };
template <> struct TrsmCfg<zcomplex> { static constexpr int NB = 32, NT = 64; };

template <typename T>
constexpr size_t trsm_smem_bytes() {
  return sizeof(T) * (size_t)TrsmCfg<T>::NB * (TrsmCfg<T>::NB + 1) +
         sizeof(T) * (size_t)TrsmCfg<T>::NB * (TrsmCfg<T>::NT + 1) + sizeof(T) * TrsmCfg<T>::NB;
}

// NT vectors per CTA (one solver thread each) and TRSM_LT threads for the loads and stores.  The
// vectors live in shared memory (Xs[i][v]: conflict free across the solver threads) and are
// solved by blocked substitution: eight unknowns at a time in registers, first the updates from
// all finished unknowns (a rolled loop: one load of x_k and eight multiply-adds per step), then
// the 8 x 8 diagonal block.  The previous form kept the whole vector in registers and unrolled
// the full substitution — 2016 multiply-adds and as many shared-memory loads of straight-line
// code per instantiation, executed once per launch, i.e. always from a cold instruction cache:
// 124 us per launch for a 64 x 64 triangle and 16384 vectors, 56 % of sgetrf at N = 16384.
constexpr int TRSM_LT = 256;
constexpr int TRSM_BS = 8;

template <typename T>
__global__ void __launch_bounds__(TRSM_LT)
trsm_base_kernel(int right, int a_lower, int op, int unit, int na, int64_t nvec, T alpha,
                 const T* __restrict__ A, int64_t lda, T* __restrict__ B, int64_t ldb) {
  constexpr int NB = TrsmCfg<T>::NB, NT = TrsmCfg<T>::NT, LDM = NB + 1, LDX = NT + 1;
  constexpr int BS = TRSM_BS;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T* Ms = reinterpret_cast<T*>(smem_raw);   // effective matrix M (see below), row pitch LDM
  T* Xs = Ms + NB * LDM;                    // the vectors: Xs[i * LDX + v]
  T* rdiag = Xs + NB * LDX;                 // reciprocals of M's diagonal
  // M = op(A) for Left;  M = op(A)^T for Right (x*op(A) = b  <=>  M x^T = b^T).
#pragma unroll 4
  for (int idx = threadIdx.x; idx < na * na; idx += TRSM_LT) {
    int i = idx % na, j = idx / na;  // stored element A(i,j); only the stored triangle is read
    bool stored = a_lower ? (i >= j) : (i <= j);
    T v = stored ? A[i + (int64_t)j * lda] : zero<T>();
    int r = i, c = j;
    if (op != OpN) { r = j; c = i; }
    if (op == OpC) v = conj_(v);
    if (right) { int t = r; r = c; c = t; }
    Ms[r * LDM + c] = v;
  }
  bool eff_lower = (a_lower != 0) != (op != OpN);
  if (right) eff_lower = !eff_lower;
  const int64_t v0 = (int64_t)blockIdx.x * NT;
  const int nv = (int)((nvec - v0) < NT ? (nvec - v0) : NT);
  const int t = threadIdx.x;
  const bool scale = !is_one(alpha);
  // ---- stage alpha * b (coalesced along the contiguous direction of B)
#pragma unroll 4
  for (int idx = t; idx < na * nv; idx += TRSM_LT) {
    int i, v;
    T b;
    if (right) { v = idx % nv; i = idx / nv; b = B[v0 + v + (int64_t)i * ldb]; }
    else       { i = idx % na; v = idx / na; b = B[i + (v0 + v) * ldb]; }
    Xs[i * LDX + v] = scale ? mul(alpha, b) : b;
  }
  __syncthreads();
  if (t < na && !unit) rdiag[t] = recip(Ms[t * LDM + t]);
  __syncthreads();
  if (t < NT && t < nv) {
    T* xv = Xs + t;                        // this thread's vector: xv[i * LDX]
    if (eff_lower) {
      for (int jb = 0; jb < na; jb += BS) {
        T xr[BS];
#pragma unroll
        for (int r = 0; r < BS; r++) xr[r] = jb + r < na ? xv[(jb + r) * LDX] : zero<T>();
        for (int k = 0; k < jb; k++) {
          const T xk = xv[k * LDX];
#pragma unroll
          for (int r = 0; r < BS; r++) xr[r] = fnma_(Ms[(jb + r) * LDM + k], xk, xr[r]);
        }
#pragma unroll
        for (int c = 0; c < BS; c++) {
          if (jb + c < na) {
            if (!unit) xr[c] = mul(xr[c], rdiag[jb + c]);
#pragma unroll
            for (int r = c + 1; r < BS; r++)
              if (jb + r < na) xr[r] = fnma_(Ms[(jb + r) * LDM + jb + c], xr[c], xr[r]);
          }
        }
#pragma unroll
        for (int r = 0; r < BS; r++)
          if (jb + r < na) xv[(jb + r) * LDX] = xr[r];
      }
    } else {
      for (int jb = ((na - 1) / BS) * BS; jb >= 0; jb -= BS) {
        T xr[BS];
#pragma unroll
        for (int r = 0; r < BS; r++) xr[r] = jb + r < na ? xv[(jb + r) * LDX] : zero<T>();
        for (int k = jb + BS; k < na; k++) {
          const T xk = xv[k * LDX];
#pragma unroll
          for (int r = 0; r < BS; r++)
            if (jb + r < na) xr[r] = fnma_(Ms[(jb + r) * LDM + k], xk, xr[r]);
        }
#pragma unroll
        for (int cc = 0; cc < BS; cc++) {
          const int c = BS - 1 - cc;
          if (jb + c < na) {
            if (!unit) xr[c] = mul(xr[c], rdiag[jb + c]);
#pragma unroll
            for (int r = 0; r < c; r++) xr[r] = fnma_(Ms[(jb + r) * LDM + jb + c], xr[c], xr[r]);
          }
        }
#pragma unroll
        for (int r = 0; r < BS; r++)
          if (jb + r < na) xv[(jb + r) * LDX] = xr[r];
      }
    }
  }
  __syncthreads();
#pragma unroll 4
  for (int idx = t; idx < na * nv; idx += TRSM_LT) {
    if (right) { const int v = idx % nv, i = idx / nv; B[v0 + v + (int64_t)i * ldb] = Xs[i * LDX + v]; }
    else       { const int i = idx % na, v = idx / na; B[i + (v0 + v) * ldb] = Xs[i * LDX + v]; }
  }
}

template <typename T>
void trsm_rec(Context& ctx, Side side, Uplo uplo, Op op, Diag diag, int64_t m, int64_t n, T alpha,
              const T* A, int64_t lda, T* B, int64_t ldb) {
  const int64_t na = side == Left ? m : n;
  const int64_t nvec = side == Left ? n : m;
  constexpr int TRSM_NB = TrsmCfg<T>::NB;
  if constexpr (std::is_same<T, double>::value) {
    // triangles of 65..128 with many vectors: one launch of the 128-wide tensor-core base case
    if (na > TRSM_NB && na <= 128 &&
        trsm_base_dmma_f64(ctx, side, uplo, op, diag, m, n, alpha, A, lda, B, ldb))
      return;
  }
  if (na <= TRSM_NB) {
    if constexpr (std::is_same<T, double>::value) {
      if (trsm_base_dmma_f64(ctx, side, uplo, op, diag, m, n, alpha, A, lda, B, ldb)) return;
    }
    constexpr int TRSM_NT = TrsmCfg<T>::NT;
    static DeviceOnce attr;
    if (attr.pending(ctx.device)) {
      LA_CUDA_CHECK(cudaFuncSetAttribute(trsm_base_kernel<T>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)trsm_smem_bytes<T>()));
      attr.mark(ctx.device);
    }
    unsigned g = (unsigned)ceil_div(nvec, TRSM_NT);
    trsm_base_kernel<T><<<g, TRSM_LT, trsm_smem_bytes<T>(), ctx.stream>>>(side == Right, uplo == Lower, (int)op,
                                                       diag == Unit, (int)na, nvec, alpha, A, lda,
                                                       B, ldb);
    LA_CUDA_CHECK(cudaGetLastError());
    return;
  }
  // split na = n1 + n2, n1 a multiple of the base order
  int64_t n1 = round_up(na / 2, TRSM_NB);
  if (n1 >= na) n1 = na - TRSM_NB > 0 ? TRSM_NB : na / 2;
  int64_t n2 = na - n1;
  const T* A11 = A;
  const T* A22 = A + n1 + n1 * lda;
  const T* A21 = A + n1;        // n2 x n1 (stored when uplo == Lower)
  const T* A12 = A + n1 * lda;  // n1 x n2 (stored when uplo == Upper)
  const bool eff_lower = (uplo == Lower) != (op != OpN);
  const T mone = from_real<T>(-1.0);
  const T onev = one<T>();
  if (side == Left) {
    T* B1 = B;
    T* B2 = B + n1;
    if (eff_lower) {
      // [M11 0; M21 M22] [X1; X2] = alpha [B1; B2]
      trsm_rec<T>(ctx, side, uplo, op, diag, n1, n, alpha, A11, lda, B1, ldb);
      if (uplo == Lower)  // M21 = A21 (op == N)
        gemm<T>(ctx, OpN, OpN, n2, n, n1, mone, A21, lda, B1, ldb, alpha, B2, ldb);
      else  // M21 = op(A12)
        gemm<T>(ctx, op, OpN, n2, n, n1, mone, A12, lda, B1, ldb, alpha, B2, ldb);
      trsm_rec<T>(ctx, side, uplo, op, diag, n2, n, onev, A22, lda, B2, ldb);
    } else {
      // [M11 M12; 0 M22]
      trsm_rec<T>(ctx, side, uplo, op, diag, n2, n, alpha, A22, lda, B2, ldb);
      if (uplo == Upper)  // M12 = A12
        gemm<T>(ctx, OpN, OpN, n1, n, n2, mone, A12, lda, B2, ldb, alpha, B1, ldb);
      else  // M12 = op(A21)
        gemm<T>(ctx, op, OpN, n1, n, n2, mone, A21, lda, B2, ldb, alpha, B1, ldb);
      trsm_rec<T>(ctx, side, uplo, op, diag, n1, n, onev, A11, lda, B1, ldb);
    }
  } else {
    T* B1 = B;
    T* B2 = B + n1 * ldb;
    if (eff_lower) {
      // [X1 X2] [M11 0; M21 M22] = alpha [B1 B2]:  X2 first, then X1 M11 = alpha B1 - X2 M21
      trsm_rec<T>(ctx, side, uplo, op, diag, m, n2, alpha, A22, lda, B2, ldb);
      if (uplo == Lower)
        gemm<T>(ctx, OpN, OpN, m, n1, n2, mone, B2, ldb, A21, lda, alpha, B1, ldb);
      else
        gemm<T>(ctx, OpN, op, m, n1, n2, mone, B2, ldb, A12, lda, alpha, B1, ldb);
      trsm_rec<T>(ctx, side, uplo, op, diag, m, n1, onev, A11, lda, B1, ldb);
    } else {
      // [X1 X2] [M11 M12; 0 M22]:  X1 first, then X2 M22 = alpha B2 - X1 M12
      trsm_rec<T>(ctx, side, uplo, op, diag, m, n1, alpha, A11, lda, B1, ldb);
      if (uplo == Upper)
        gemm<T>(ctx, OpN, OpN, m, n2, n1, mone, B1, ldb, A12, lda, alpha, B2, ldb);
      else
        gemm<T>(ctx, OpN, op, m, n2, n1, mone, B1, ldb, A21, lda, alpha, B2, ldb);
      trsm_rec<T>(ctx, side, uplo, op, diag, m, n2, onev, A22, lda, B2, ldb);
    }
  }
}

}  // namespace

template <typename T>
void trsm(Context& ctx, Side side, Uplo uplo, Op op, Diag diag, int64_t m, int64_t n, T alpha,
          const T* A, int64_t lda, T* B, int64_t ldb) {
  if (m <= 0 || n <= 0) return;
  if (!Traits<T>::is_complex && op == OpC) op = OpT;
  // a vector or a handful of right-hand sides (solve / cho_solve / solvels with a column):
  // one pass over the triangle in a single launch instead of ~m/32 tiny ones
  if (side == Left && trsv_small<T>(ctx, uplo, op, diag, m, n, alpha, A, lda, B, ldb)) return;
  trsm_rec<T>(ctx, side, uplo, op, diag, m, n, alpha, A, lda, B, ldb);
}
template void trsm<float>(Context&, Side, Uplo, Op, Diag, int64_t, int64_t, float, const float*,
                          int64_t, float*, int64_t);
template void trsm<double>(Context&, Side, Uplo, Op, Diag, int64_t, int64_t, double,
                           const double*, int64_t, double*, int64_t);
template void trsm<zcomplex>(Context&, Side, Uplo, Op, Diag, int64_t, int64_t, zcomplex,
                             const zcomplex*, int64_t, zcomplex*, int64_t);

// ------------------------------------------------------------------ TRMM / SYMM / HEMM
// The flag-selected siblings of Matrix#* (mult.cr:35-62, :78-91).  Both are expressed as
// a dense GEMM on an explicitly completed operand held in workspace: the O(n^2) expansion
// is noise next to the O(n^3) product and lets them ride the tensor-core GEMM.
template <typename T>
void trmm(Context& ctx, Side side, Uplo uplo, Op op, Diag diag, int64_t m, int64_t n, T alpha,
          const T* A, int64_t lda, T* B, int64_t ldb) {
  if (m <= 0 || n <= 0) return;
  const int64_t na = side == Left ? m : n;
  Arena::Mark mk = ctx.arena.mark();
  const int64_t ldt = round_up(na, 16), ldw = round_up(m, 16);
  T* Tf = ctx.alloc<T>(ldt * na);
  T* W = ctx.alloc<T>(ldw * n);
  copy_triangle<T>(ctx, uplo, diag == Unit, na, na, A, lda, Tf, ldt);
  copy_matrix<T>(ctx, m, n, B, ldb, W, ldw);
  if (side == Left)
    gemm<T>(ctx, op, OpN, m, n, m, alpha, Tf, ldt, W, ldw, zero<T>(), B, ldb);
  else
    gemm<T>(ctx, OpN, op, m, n, n, alpha, W, ldw, Tf, ldt, zero<T>(), B, ldb);
  ctx.arena.release(mk);
}
template void trmm<float>(Context&, Side, Uplo, Op, Diag, int64_t, int64_t, float, const float*,
                          int64_t, float*, int64_t);
template void trmm<double>(Context&, Side, Uplo, Op, Diag, int64_t, int64_t, double,
                           const double*, int64_t, double*, int64_t);
template void trmm<zcomplex>(Context&, Side, Uplo, Op, Diag, int64_t, int64_t, zcomplex,
                             const zcomplex*, int64_t, zcomplex*, int64_t);

template <typename T>
void symm(Context& ctx, bool herm, Side side, Uplo uplo, int64_t m, int64_t n, T alpha,
          const T* A, int64_t lda, const T* B, int64_t ldb, T beta, T* C, int64_t ldc) {
  if (m <= 0 || n <= 0) return;
  const int64_t na = side == Left ? m : n;
  Arena::Mark mk = ctx.arena.mark();
  const int64_t ldf = round_up(na, 16);
  T* Af = ctx.alloc<T>(ldf * na);
  symmetrize<T>(ctx, herm, uplo, na, A, lda, Af, ldf);
  if (side == Left)
    gemm<T>(ctx, OpN, OpN, m, n, m, alpha, Af, ldf, B, ldb, beta, C, ldc);
  else
    gemm<T>(ctx, OpN, OpN, m, n, n, alpha, B, ldb, Af, ldf, beta, C, ldc);
  ctx.arena.release(mk);
}
template void symm<float>(Context&, bool, Side, Uplo, int64_t, int64_t, float, const float*,
                          int64_t, const float*, int64_t, float, float*, int64_t);
template void symm<double>(Context&, bool, Side, Uplo, int64_t, int64_t, double, const double*,
                           int64_t, const double*, int64_t, double, double*, int64_t);
template void symm<zcomplex>(Context&, bool, Side, Uplo, int64_t, int64_t, zcomplex,
                             const zcomplex*, int64_t, const zcomplex*, int64_t, zcomplex,
                             zcomplex*, int64_t);

#define LA_INST_HELPERS(T)                                                                      \
  template void copy_matrix<T>(Context&, int64_t, int64_t, const T*, int64_t, T*, int64_t);     \
  template void transpose_matrix<T>(Context&, bool, int64_t, int64_t, const T*, int64_t, T*,    \
                                    int64_t);                                                   \
  template void set_matrix<T>(Context&, int64_t, int64_t, T, T, T*, int64_t);                   \
  template void copy_triangle<T>(Context&, Uplo, bool, int64_t, int64_t, const T*, int64_t, T*, \
                                 int64_t);                                                      \
  template void symmetrize<T>(Context&, bool, Uplo, int64_t, const T*, int64_t, T*, int64_t);
LA_INST_HELPERS(float)
LA_INST_HELPERS(double)
LA_INST_HELPERS(zcomplex)

}  // namespace la
