// la_b200 — Cholesky and LU factorisations and the solves built on them.
//
// Reference call sites replaced: ?potrf_/?potrs_/?posv_ (src/linalg/cholesky.cr:47-90,
// src/linalg/linalg.cr:433-434), ?getrf_/?getrs_/?gesv_ (src/linalg/lu.cr:74-80,151-162,
// src/linalg/linalg.cr:443-445), ?trtrs_ (src/linalg/linalg.cr:431-432).
//
// Both factorisations are recursive (the shape LAPACK 3.12's ?getrf2/?potrf2 use), so that
// almost all flops land in large GEMM / TRSM calls, with hand-written panel kernels at the
// leaves:
//   * potf2_kernel   one CTA, block <= 64x64 resident in shared memory.
//   * lu_panel_kernel one cooperative launch per leaf panel (<= 32 columns, any height):
//       the panel is distributed by rows over the CTAs' shared memory and each column
//       costs ONE grid barrier (local arg-max + candidate row published together).
//   * laswp_kernel    row interchanges.
#include "la_internal.h"
#include <cooperative_groups.h>

namespace la {

// =========================================================================== Cholesky
namespace {

constexpr int POTF2_NB = 64;

// Factor one n x n (n <= 64) diagonal block in shared memory.  The block is always
// factored as lower (for uplo = U the conjugate transpose is loaded / stored), only the
// named triangle of A is read or written.
template <typename T>
__global__ void __launch_bounds__(256)
potf2_kernel(int lower, int n, T* __restrict__ A, int64_t lda, int* info, int info_off) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T(*S)[POTF2_NB + 1] = reinterpret_cast<T(*)[POTF2_NB + 1]>(smem_raw);
  __shared__ int fail;
  if (*info != 0) return;  // an earlier block already failed: LAPACK would have stopped
  const int tid = threadIdx.x;
  if (tid == 0) fail = 0;
  for (int idx = tid; idx < n * n; idx += blockDim.x) {
    int i = idx % n, j = idx / n;
    if (lower) { if (i >= j) S[i][j] = A[i + (int64_t)j * lda]; }
    else       { if (i <= j) S[j][i] = conj_(A[i + (int64_t)j * lda]); }
  }
  __syncthreads();
  for (int j = 0; j < n; j++) {
    if (tid == 0) {
      double d = (double)real_(S[j][j]);
      if (!(d > 0.0)) {  // also catches NaN, as LAPACK's disnan test does
        fail = j + 1;
      } else {
        S[j][j] = from_real<T>(sqrt(d));
      }
    }
    __syncthreads();
    if (fail) break;
    const double rinv = 1.0 / (double)real_(S[j][j]);
    for (int i = j + 1 + tid; i < n; i += blockDim.x) S[i][j] = scale_real(S[i][j], (typename Traits<T>::real)rinv);
    __syncthreads();
    // trailing update of the lower triangle: S(i,k) -= S(i,j)*conj(S(k,j)), k>j, i>=k
    const int rem = n - j - 1;
    for (int idx = tid; idx < rem * rem; idx += blockDim.x) {
      int i = j + 1 + idx % rem, k = j + 1 + idx / rem;
      if (i >= k) {
        T v = fnma_(S[i][j], conj_(S[k][j]), S[i][k]);
        if (i == k) v = from_real<T>((double)real_(v));
        S[i][k] = v;
      }
    }
    __syncthreads();
  }
  if (fail) {
    if (tid == 0) atomicCAS(info, 0, info_off + fail);
    // LAPACK leaves the leading (fail-1) columns factored; write those back
  }
  const int ncols_done = fail ? fail - 1 : n;
  for (int idx = tid; idx < n * n; idx += blockDim.x) {
    int i = idx % n, j = idx / n;
    if (lower) { if (i >= j && j < ncols_done) A[i + (int64_t)j * lda] = S[i][j]; }
    else       { if (i <= j && i < ncols_done) A[i + (int64_t)j * lda] = conj_(S[j][i]); }
  }
}

template <typename T>
void potrf_rec(Context& ctx, Uplo uplo, int64_t n, T* A, int64_t lda, int* d_info,
               int64_t off) {
  if (n <= POTF2_NB) {
    size_t smem = sizeof(T) * POTF2_NB * (POTF2_NB + 1);
    potf2_kernel<T><<<1, 256, smem, ctx.stream>>>(uplo == Lower, (int)n, A, lda, d_info, (int)off);
    LA_CUDA_CHECK(cudaGetLastError());
    return;
  }
  int64_t n1 = round_up(n / 2, POTF2_NB);
  if (n1 >= n) n1 = n / 2;
  int64_t n2 = n - n1;
  T* A11 = A;
  T* A22 = A + n1 + n1 * lda;
  potrf_rec<T>(ctx, uplo, n1, A11, lda, d_info, off);
  if (uplo == Lower) {
    T* A21 = A + n1;
    // A21 <- A21 * L11^-H ;  A22 <- A22 - A21*A21^H
    trsm<T>(ctx, Right, Lower, OpC, NonUnit, n2, n1, one<T>(), A11, lda, A21, lda);
    herk<T>(ctx, Lower, OpN, n2, n1, -1.0, A21, lda, 1.0, A22, lda);
  } else {
    T* A12 = A + n1 * lda;
    // A12 <- U11^-H * A12 ;  A22 <- A22 - A12^H*A12
    trsm<T>(ctx, Left, Upper, OpC, NonUnit, n1, n2, one<T>(), A11, lda, A12, lda);
    herk<T>(ctx, Upper, OpC, n2, n1, -1.0, A12, lda, 1.0, A22, lda);
  }
  potrf_rec<T>(ctx, uplo, n2, A22, lda, d_info, off + n1);
}

}  // namespace

template <typename T>
void potrf(Context& ctx, Uplo uplo, int64_t n, T* A, int64_t lda, int* d_info) {
  LA_CUDA_CHECK(cudaMemsetAsync(d_info, 0, sizeof(int), ctx.stream));
  if (n <= 0) return;
  static bool attr_set = false;
  if (!attr_set) {
    LA_CUDA_CHECK(cudaFuncSetAttribute(potf2_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)(sizeof(T) * POTF2_NB * (POTF2_NB + 1))));
  }
  potrf_rec<T>(ctx, uplo, n, A, lda, d_info, 0);
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

// =========================================================================== LU
namespace {

constexpr int LU_NT = 256;       // threads per CTA of the panel kernel
constexpr int LU_MAX_JB = 32;    // widest leaf panel
constexpr int LU_MAX_CTAS = 160; // slots reserved in the exchange buffer

// Exchange area in global memory for one panel launch (double-buffered by column parity).
template <typename T>
struct LuExchange {
  unsigned int barrier;              // monotonically increasing arrival counter
  unsigned int pad[63];
  struct Slot {
    typename Traits<T>::real val;    // best |a(i,j)| of this CTA (-1: no candidate)
    int idx;                         // its row (panel-relative)
    int pad;
    T row[LU_MAX_JB];                // the candidate row's current contents
  } slot[2][LU_MAX_CTAS];
  T rowj[2][LU_MAX_JB];              // current contents of row j (the row being swapped out)
};

__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

__device__ __forceinline__ void grid_barrier(unsigned* counter, unsigned target) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(counter, 1u);
    while (ld_acquire_u32(counter) < target) { }
    __threadfence();
  }
  __syncthreads();
}

template <typename R>
__device__ __forceinline__ void argmax_combine(R& v, int& i, R ov, int oi) {
  if (ov > v || (ov == v && oi < i)) { v = ov; i = oi; }
}

// Block-wide (value, index) arg-max with "first maximal element" tie-break (i?amax rule).
template <typename R>
__device__ __forceinline__ void block_argmax(R& v, int& i, R* sval, int* sidx) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    R ov = __shfl_xor_sync(0xffffffffu, v, o);
    int oi = __shfl_xor_sync(0xffffffffu, i, o);
    argmax_combine(v, i, ov, oi);
  }
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  __syncthreads();  // protect sval/sidx reuse
  if (l == 0) { sval[w] = v; sidx[w] = i; }
  __syncthreads();
  if (w == 0) {
    const int nw = blockDim.x >> 5;
    v = l < nw ? sval[l] : (R)-2;
    i = l < nw ? sidx[l] : 0x7fffffff;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      R ov = __shfl_xor_sync(0xffffffffu, v, o);
      int oi = __shfl_xor_sync(0xffffffffu, i, o);
      argmax_combine(v, i, ov, oi);
    }
    if (l == 0) { sval[0] = v; sidx[0] = i; }
  }
  __syncthreads();
  v = sval[0];
  i = sidx[0];
}

// LU with partial pivoting of an m x jb panel (m >= jb, jb <= 32).
//   A      : panel, column-major, leading dimension lda
//   ipiv   : pivot entries of the panel's columns; written as row0 + p + 1 (1-based, global)
//   rpc    : rows per CTA (contiguous ownership), pitch of the shared-memory panel
template <typename T>
__global__ void __launch_bounds__(LU_NT)
lu_panel_kernel(int m, int jb, T* __restrict__ A, int64_t lda, int* __restrict__ ipiv, int row0,
                int* info, int info_off, LuExchange<T>* xch, int rpc) {
  using R = typename Traits<T>::real;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T* S = reinterpret_cast<T*>(smem_raw);           // S[c * rpc + r]
  __shared__ T prow[LU_MAX_JB];
  __shared__ T rowj[LU_MAX_JB];
  __shared__ R sval[LU_NT / 32];
  __shared__ int sidx[LU_NT / 32];
  __shared__ int s_pcta;

  const int tid = threadIdx.x;
  const int G = gridDim.x;
  const int rs = blockIdx.x * rpc;                       // first owned row
  const int nr = max(0, min(m - rs, rpc));               // owned rows

  for (int c = 0; c < jb; c++)
    for (int r = tid; r < nr; r += LU_NT) S[c * rpc + r] = A[(int64_t)(rs + r) + (int64_t)c * lda];
  __syncthreads();

  unsigned target = 0;
  for (int j = 0; j < jb; j++) {
    const int par = j & 1;
    // ---- 1. local arg-max of |a(i,j)| over owned rows i >= j
    R bv = (R)-1;
    int bi = 0x7fffffff;
    for (int r = tid; r < nr; r += LU_NT) {
      int gr = rs + r;
      if (gr >= j) {
        R v = abs1(S[j * rpc + r]);
        if (v > bv) { bv = v; bi = gr; }
      }
    }
    block_argmax<R>(bv, bi, sval, sidx);
    // ---- 2. publish candidate (value, index, row contents) and row j
    if (tid == 0) {
      __stcg(&xch->slot[par][blockIdx.x].val, bv);
      __stcg(&xch->slot[par][blockIdx.x].idx, bi);
    }
    if (bv >= (R)0 && tid < jb) {
      T v = S[tid * rpc + (bi - rs)];
      T* dst = &xch->slot[par][blockIdx.x].row[tid];
      if constexpr (Traits<T>::is_complex) { __stcg(&dst->x, v.x); __stcg(&dst->y, v.y); }
      else __stcg(dst, v);
    }
    if (j >= rs && j < rs + nr && tid < jb) {
      T v = S[tid * rpc + (j - rs)];
      T* dst = &xch->rowj[par][tid];
      if constexpr (Traits<T>::is_complex) { __stcg(&dst->x, v.x); __stcg(&dst->y, v.y); }
      else __stcg(dst, v);
    }
    // ---- 3. one grid barrier per column
    target += (unsigned)G;
    grid_barrier(&xch->barrier, target);
    // ---- 4. every CTA picks the global winner
    R pv = (R)-1;
    int pi = 0x7fffffff;
    if (tid < G) {
      pv = __ldcg(&xch->slot[par][tid].val);
      pi = __ldcg(&xch->slot[par][tid].idx);
      if (!(pv >= (R)0)) { pv = (R)-1; pi = 0x7fffffff; }
    }
    block_argmax<R>(pv, pi, sval, sidx);
    if (tid < G) {
      R v = __ldcg(&xch->slot[par][tid].val);
      int i = __ldcg(&xch->slot[par][tid].idx);
      if (v == pv && i == pi) s_pcta = tid;
    }
    __syncthreads();
    if (tid < jb) {
      const T* src = &xch->slot[par][s_pcta].row[tid];
      const T* srj = &xch->rowj[par][tid];
      T a, b;
      if constexpr (Traits<T>::is_complex) {
        a.x = __ldcg(&src->x); a.y = __ldcg(&src->y);
        b.x = __ldcg(&srj->x); b.y = __ldcg(&srj->y);
      } else { a = __ldcg(src); b = __ldcg(srj); }
      prow[tid] = a;
      rowj[tid] = b;
    }
    __syncthreads();
    // ---- 5. row interchange j <-> pi
    if (pi != j && tid < jb) {
      if (pi >= rs && pi < rs + nr) S[tid * rpc + (pi - rs)] = rowj[tid];
      if (j >= rs && j < rs + nr)   S[tid * rpc + (j - rs)] = prow[tid];
    }
    if (blockIdx.x == 0 && tid == 0) ipiv[j] = row0 + pi + 1;
    __syncthreads();
    // ---- 6. scale the column and rank-1 update the rest of the panel
    const T pivot = prow[j];
    if (is_zero(pivot)) {
      if (blockIdx.x == 0 && tid == 0) atomicCAS(info, 0, info_off + j + 1);
      // as in getf2: no scaling, the (zero-multiplier) update is skipped
    } else {
      const bool use_recip = abs1(pivot) >= sfmin<R>();
      const T rinv = recip(pivot);
      for (int r = tid; r < nr; r += LU_NT) {
        int gr = rs + r;
        if (gr > j) {
          T a = S[j * rpc + r];
          T l = use_recip ? mul(a, rinv) : div_(a, pivot);
          S[j * rpc + r] = l;
          for (int c = j + 1; c < jb; c++) S[c * rpc + r] = fnma_(l, prow[c], S[c * rpc + r]);
        }
      }
    }
    __syncthreads();
  }
  for (int c = 0; c < jb; c++)
    for (int r = tid; r < nr; r += LU_NT) A[(int64_t)(rs + r) + (int64_t)c * lda] = S[c * rpc + r];
}

template <typename T>
__global__ void laswp_kernel(int64_t n, T* __restrict__ A, int64_t lda, int64_t k1, int64_t k2,
                             const int* __restrict__ ipiv, int reverse) {
  int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= n) return;
  T* col = A + c * lda;
  if (!reverse) {
    for (int64_t k = k1; k < k2; k++) {
      int64_t p = (int64_t)ipiv[k] - 1;
      if (p != k) { T t = col[k]; col[k] = col[p]; col[p] = t; }
    }
  } else {
    for (int64_t k = k2 - 1; k >= k1; k--) {
      int64_t p = (int64_t)ipiv[k] - 1;
      if (p != k) { T t = col[k]; col[k] = col[p]; col[p] = t; }
    }
  }
}

constexpr size_t LU_SMEM_BUDGET = 200 * 1024;

template <typename T>
struct LuPlan {
  int nb0;                 // leaf panel width for this matrix height
  LuExchange<T>* xch;
};

template <typename T>
void lu_panel(Context& ctx, int64_t m, int jb, T* A, int64_t lda, int* ipiv, int64_t row0,
              int* d_info, LuExchange<T>* xch) {
  // rows per CTA: as few CTAs as keep the panel resident and the per-column work short
  int64_t G = ceil_div(m, 512);
  if (G > ctx.sm_count) G = ctx.sm_count;
  if (G < 1) G = 1;
  int64_t rpc = ceil_div(m, G);
  // grow G until the slice fits shared memory
  while ((size_t)rpc * jb * sizeof(T) > LU_SMEM_BUDGET && G < ctx.sm_count) {
    G++;
    rpc = ceil_div(m, G);
  }
  rpc = round_up(rpc, 2);
  G = ceil_div(m, rpc);
  size_t smem = (size_t)rpc * jb * sizeof(T);
  if (smem > LU_SMEM_BUDGET) {
    fprintf(stderr, "la_b200: internal error: LU panel %lld x %d does not fit\n", (long long)m, jb);
    abort();
  }
  LA_CUDA_CHECK(cudaMemsetAsync(&xch->barrier, 0, sizeof(unsigned), ctx.stream));
  int mi = (int)m, row0i = (int)row0, info_off = (int)row0, rpci = (int)rpc;
  void* args[] = {&mi, &jb, &A, &lda, &ipiv, &row0i, &d_info, &info_off, &xch, &rpci};
  LA_CUDA_CHECK(cudaLaunchCooperativeKernel((void*)lu_panel_kernel<T>, dim3((unsigned)G),
                                            dim3(LU_NT), args, smem, ctx.stream));
}

// Factor columns [j0, j0+jn) of the m-row matrix A below row j0 (all earlier updates
// applied); pivots are global row indices.
template <typename T>
void getrf_rec(Context& ctx, int64_t m, T* A, int64_t lda, int* ipiv, int* d_info, int64_t j0,
               int64_t jn, const LuPlan<T>& plan) {
  if (jn <= plan.nb0) {
    lu_panel<T>(ctx, m - j0, (int)jn, A + j0 + j0 * lda, lda, ipiv + j0, j0, d_info, plan.xch);
    return;
  }
  int64_t n1 = jn / 2;
  if (n1 > 16) n1 = round_up(n1, 16);
  if (n1 >= jn) n1 = jn / 2;
  if (n1 < plan.nb0 && jn > plan.nb0) n1 = plan.nb0;
  int64_t n2 = jn - n1;
  getrf_rec<T>(ctx, m, A, lda, ipiv, d_info, j0, n1, plan);
  T* A12 = A + j0 + (j0 + n1) * lda;
  T* A22 = A + (j0 + n1) + (j0 + n1) * lda;
  const T* A11 = A + j0 + j0 * lda;
  const T* A21 = A + (j0 + n1) + j0 * lda;
  laswp<T>(ctx, n2, A + (j0 + n1) * lda, lda, j0, j0 + n1, ipiv, false);
  trsm<T>(ctx, Left, Lower, OpN, Unit, n1, n2, one<T>(), A11, lda, A12, lda);
  gemm<T>(ctx, OpN, OpN, m - j0 - n1, n2, n1, from_real<T>(-1.0), A21, lda, A12, lda, one<T>(),
          A22, lda);
  getrf_rec<T>(ctx, m, A, lda, ipiv, d_info, j0 + n1, n2, plan);
  laswp<T>(ctx, n1, A + j0 * lda, lda, j0 + n1, j0 + jn, ipiv, false);
}

}  // namespace

template <typename T>
void laswp(Context& ctx, int64_t n, T* A, int64_t lda, int64_t k1, int64_t k2, const int* d_ipiv,
           bool reverse) {
  if (n <= 0 || k2 <= k1) return;
  laswp_kernel<T><<<(unsigned)ceil_div(n, 128), 128, 0, ctx.stream>>>(n, A, lda, k1, k2, d_ipiv,
                                                                      reverse ? 1 : 0);
  LA_CUDA_CHECK(cudaGetLastError());
}

template <typename T>
void getrf(Context& ctx, int64_t m, int64_t n, T* A, int64_t lda, int* d_ipiv, int* d_info) {
  LA_CUDA_CHECK(cudaMemsetAsync(d_info, 0, sizeof(int), ctx.stream));
  const int64_t k = m < n ? m : n;
  if (k <= 0) return;
  static bool attr_set = false;
  if (!attr_set) {
    LA_CUDA_CHECK(cudaFuncSetAttribute(lu_panel_kernel<T>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)LU_SMEM_BUDGET));
    attr_set = true;
  }
  Arena::Mark mk = ctx.arena.mark();
  LuPlan<T> plan;
  plan.xch = (LuExchange<T>*)ctx.arena.alloc(sizeof(LuExchange<T>));
  // widest leaf panel that stays resident when the tallest panel (m rows) is spread
  // over all SMs
  int64_t rpc = round_up(ceil_div(m, ctx.sm_count), 2);
  int64_t nb0 = (int64_t)(LU_SMEM_BUDGET / (rpc * sizeof(T)));
  if (nb0 > LU_MAX_JB) nb0 = LU_MAX_JB;
  if (nb0 < 1) {
    fprintf(stderr, "la_b200: getrf: %lld rows exceed the panel kernel's capacity\n", (long long)m);
    abort();
  }
  plan.nb0 = (int)nb0;
  getrf_rec<T>(ctx, m, A, lda, d_ipiv, d_info, 0, k, plan);
  if (n > k) {  // wide matrix: U12 = L11^-1 * P * A12
    laswp<T>(ctx, n - k, A + k * lda, lda, 0, k, d_ipiv, false);
    trsm<T>(ctx, Left, Lower, OpN, Unit, k, n - k, one<T>(), A, lda, A + k * lda, lda);
  }
  ctx.arena.release(mk);
}

template <typename T>
void getrs(Context& ctx, Op trans, int64_t n, int64_t nrhs, const T* A, int64_t lda,
           const int* d_ipiv, T* B, int64_t ldb) {
  if (n <= 0 || nrhs <= 0) return;
  if (!Traits<T>::is_complex && trans == OpC) trans = OpT;
  if (trans == OpN) {
    laswp<T>(ctx, nrhs, B, ldb, 0, n, d_ipiv, false);
    trsm<T>(ctx, Left, Lower, OpN, Unit, n, nrhs, one<T>(), A, lda, B, ldb);
    trsm<T>(ctx, Left, Upper, OpN, NonUnit, n, nrhs, one<T>(), A, lda, B, ldb);
  } else {
    trsm<T>(ctx, Left, Upper, trans, NonUnit, n, nrhs, one<T>(), A, lda, B, ldb);
    trsm<T>(ctx, Left, Lower, trans, Unit, n, nrhs, one<T>(), A, lda, B, ldb);
    laswp<T>(ctx, nrhs, B, ldb, 0, n, d_ipiv, true);
  }
}

// =========================================================================== TRTRS
namespace {
template <typename T>
__global__ void diag_zero_check_kernel(int64_t n, const T* __restrict__ A, int64_t lda, int* info) {
  // info = smallest i (1-based) with A(i,i) == 0
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && is_zero(A[i + i * lda])) atomicMin(info, (int)(i + 1));
}
__global__ void info_fix_kernel(int* info) {
  if (*info == 0x7fffffff) *info = 0;
}
__global__ void info_init_kernel(int* info) { *info = 0x7fffffff; }
}  // namespace

template <typename T>
void trtrs(Context& ctx, Uplo uplo, Op trans, Diag diag, int64_t n, int64_t nrhs, const T* A,
           int64_t lda, T* B, int64_t ldb, int* d_info) {
  LA_CUDA_CHECK(cudaMemsetAsync(d_info, 0, sizeof(int), ctx.stream));
  if (n <= 0) return;
  if (diag == NonUnit) {
    info_init_kernel<<<1, 1, 0, ctx.stream>>>(d_info);
    diag_zero_check_kernel<T><<<(unsigned)ceil_div(n, 256), 256, 0, ctx.stream>>>(n, A, lda, d_info);
    info_fix_kernel<<<1, 1, 0, ctx.stream>>>(d_info);
    LA_CUDA_CHECK(cudaGetLastError());
    // LAPACK returns before solving when singular; the caller reads info and must treat B
    // as undefined in that case.
  }
  trsm<T>(ctx, Left, uplo, trans, diag, n, nrhs, one<T>(), A, lda, B, ldb);
}

#define LA_INST_FACTOR(T)                                                                         \
  template void potrf<T>(Context&, Uplo, int64_t, T*, int64_t, int*);                             \
  template void potrs<T>(Context&, Uplo, int64_t, int64_t, const T*, int64_t, T*, int64_t);       \
  template void laswp<T>(Context&, int64_t, T*, int64_t, int64_t, int64_t, const int*, bool);     \
  template void getrf<T>(Context&, int64_t, int64_t, T*, int64_t, int*, int*);                    \
  template void getrs<T>(Context&, Op, int64_t, int64_t, const T*, int64_t, const int*, T*,       \
                         int64_t);                                                                \
  template void trtrs<T>(Context&, Uplo, Op, Diag, int64_t, int64_t, const T*, int64_t, T*,       \
                         int64_t, int*);
LA_INST_FACTOR(float)
LA_INST_FACTOR(double)
LA_INST_FACTOR(zcomplex)

}  // namespace la
