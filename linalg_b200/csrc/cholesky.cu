// la_b200 — Cholesky factorisation and solve.
//
// Reference call sites replaced: ?potrf_ / ?potrs_ / ?posv_ (src/linalg/cholesky.cr:47-90,
// src/linalg/linalg.cr:433-434; PD detection src/matrix/flag_checks.cr:32-40).
//
// Large n: look-ahead driver (see potrf below).  Small n and the diagonal blocks of the driver:
// two-level right-looking algorithm.  Inside an outer panel of NBO columns, a one-CTA leaf kernel
// factors a diagonal block of NB (128) columns, a tall TRSM solves the rows below it and a
// rank-NB masked GEMM updates the rest of the panel — every step is parallel over the rows
// below, only the leaf runs on a single SM.  The outer trailing matrix then takes one rank-NBO
// HERK on the tensor cores, which is where almost all of the n^3/3 flops are.
// Only the named triangle of A is read or written (cholesky.cr:50-64 relies on that).
#include "la_internal.h"
#include "tri_small.cuh"
#include <type_traits>

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

// Compile-time loop: the body is instantiated once per index, so register arrays indexed by it
// stay in registers whatever the unroller's size heuristics would have decided.
template <int I, int N, typename F>
__device__ __forceinline__ void static_for(F&& f) {
  if constexpr (I < N) {
    f(std::integral_constant<int, I>{});
    static_for<I + 1, N>(f);
  }
}

// sqrt(d) and 1/sqrt(d) of a finite positive double without a single branch (the library
// routines carry special-case branches, and a potentially divergent branch between two warp
// shuffles makes the compiler re-converge the warp every time).  d = m * 4^h with m in [1, 4):
// a float reciprocal square root of m seeds two Newton steps (relative error ~1e-7 -> 1e-14 ->
// below 2^-53), one more correction gives sqrt(m) to within an ulp, and the scaling by 2^h is
// exact.  Subnormal and tiny d are pre-scaled by 4^256.
__device__ __forceinline__ void sqrt_rsqrt_pos(double d, double& sq, double& rs) {
  const bool tiny = d < 1.0e-280;
  d = tiny ? d * 1.3407807929942597e154 /* 2^512 */ : d;
  const int hi = __double2hiint(d), lo = __double2loint(d);
  const int e = ((hi >> 20) & 0x7ff) - 1023;
  const int ee = e & ~1;                                   // even part of the exponent
  const double m = __hiloint2double(hi - (ee << 20), lo);  // in [1, 4)
  double y = (double)rsqrtf((float)m);
  y = y * fma(-0.5 * m, y * y, 1.5);
  y = y * fma(-0.5 * m, y * y, 1.5);
  double sm = m * y;
  sm = fma(0.5 * y, fma(-sm, sm, m), sm);
  const int h = (ee >> 1) - (tiny ? 256 : 0);
  sq = sm * __hiloint2double((1023 + h) << 20, 0);
  rs = y * __hiloint2double((1023 - h) << 20, 0);
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
#ifdef LA_PANEL_TRACE
__device__ long long g_potf_trace[32];
#define PTRACE(slot) do { if (threadIdx.x == 0) g_potf_trace[slot] = clock64(); } while (0)
#else
#define PTRACE(slot) do { } while (0)
#endif

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
  const int tid = threadIdx.x, lane = tid & 31;
  // broadcast from lane 0: tells the compiler that the warp index is warp-uniform, so that the
  // shuffles under `if (warp == 0)` are plain SHFLs and not re-convergence sequences
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  if (tid == 0) fail = 0;
  PTRACE(0);
  for (int idx = tid; idx < n * n; idx += POTF_NT) {
    int i = idx % n, j = idx / n;
    if (lower) { if (i >= j) S[i * LDS + j] = A[i + (int64_t)j * lda]; }
    else       { if (i <= j) S[j * LDS + i] = conj_(A[i + (int64_t)j * lda]); }
  }
  __syncthreads();
  PTRACE(1);
  int done_cols = n;
  for (int j0 = 0; j0 < n; j0 += POTF_IB) {
    const int jb = min(POTF_IB, n - j0);
    if (warp == 0) {  // ---- (1) diagonal block: lane i keeps row i in registers
      // The column loop is branch free (a ragged block is padded with the identity, a failed
      // pivot is recorded and replaced by 1): shuffles inside data-dependent branches make the
      // compiler re-converge the warp around every one of them, which tripled the latency.
      T* D = S + j0 * LDS + j0;
      T a[POTF_IB];
#pragma unroll
      for (int c = 0; c < POTF_IB; c++)
        a[c] = (lane < jb && c <= lane) ? D[lane * LDS + c] : (c == lane ? one<T>() : zero<T>());
      int f = 0;
      double my_rs = 1.0;
      static_for<0, POTF_IB>([&](auto J) {
        constexpr int j = decltype(J)::value;
        const double d0 = __shfl_sync(0xffffffffu, (double)real_(a[j]), j);
        const bool bad = !(d0 > 0.0 && d0 < 1.7e308);   // also catches NaN; uniform over the warp
        f = (bad && f == 0) ? j0 + j + 1 : f;
        const double d = bad ? 1.0 : d0;
        double sq, rs;
        sqrt_rsqrt_pos(d, sq, rs);
        const R rinv = (R)rs;
        my_rs = lane == j ? rs : my_rs;
        a[j] = lane == j ? from_real<T>(sq) : (lane > j ? scale_real(a[j], rinv) : a[j]);
        // a(i,k) -= a(i,j) * conj(a(k,j)) for j < k <= i; a(k,j) lives in lane k
        static_for<j + 1, POTF_IB>([&](auto K) {
          constexpr int k = decltype(K)::value;
          const T akj = shfl_T(a[j], k);
          T v = fnma_(a[j], conj_(akj), a[k]);
          if (lane == k) v = from_real<T>((double)real_(v));
          a[k] = lane >= k ? v : a[k];
        });
      });
      rdiag[lane] = from_real<T>(my_rs);
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
    PTRACE(2 + 3 * (j0 / POTF_IB));
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
    PTRACE(3 + 3 * (j0 / POTF_IB));
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
    PTRACE(4 + 3 * (j0 / POTF_IB));
  }
  PTRACE(20);
  if (fail && tid == 0) atomicCAS(info, 0, info_off + fail);
  for (int idx = tid; idx < n * n; idx += POTF_NT) {
    int i = idx % n, j = idx / n;
    if (lower) { if (i >= j && j < done_cols) A[i + (int64_t)j * lda] = S[i * LDS + j]; }
    else       { if (i <= j && i < done_cols) A[i + (int64_t)j * lda] = conj_(S[j * LDS + i]); }
  }
  PTRACE(21);
}

}  // namespace

// Two-level right-looking factorisation on ONE stream (the whole job for small n, the diagonal
// blocks of the look-ahead driver below for large n).  info_off: global index of A(0,0).
template <typename T>
static void potrf_twolevel(Context& ctx, Uplo uplo, int64_t n, T* A, int64_t lda, int* d_info,
                           int64_t info_off) {
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
                                                       d_info, (int)(info_off + j0));
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

// Large n: blocked right-looking factorisation with look-ahead on the diagonal block.
//   step K:  P = A(K+kb:n, K:K+kb) * L_KK^-H            (one tall recursive TRSM, main stream)
//            next block column -= P P(0:kb2)^H           (masked GEMM, all SMs)
//            -> side stream: factor the NEXT diagonal block (the chain of one-CTA leaf kernels
//               and small updates, which needs a handful of SMs)
//            rest of the trailing matrix -= P' P'^H      (HERK on the remaining SMs, meanwhile)
// The serial part of the algorithm (n / 128 leaf launches) is thereby hidden behind the tensor
// core work as long as the trailing update of a step outlasts one diagonal block.
#ifdef LA_CHOL_TIMELINE
}  // namespace la
#include <vector>
namespace la {
static std::vector<cudaEvent_t> g_ctl;   // per step: start, after TRSM, after next-column GEMM, side done, HERK done
static void ctl_mark(cudaStream_t s) {
  cudaEvent_t e; cudaEventCreate(&e); cudaEventRecord(e, s); g_ctl.push_back(e);
}
extern "C" void la_b200_debug_chol_timeline() {
  cudaDeviceSynchronize();
  for (size_t i = 0; i + 4 < g_ctl.size(); i += 5) {
    float a, b, c, d;
    cudaEventElapsedTime(&a, g_ctl[i], g_ctl[i + 1]);
    cudaEventElapsedTime(&b, g_ctl[i + 1], g_ctl[i + 2]);
    cudaEventElapsedTime(&c, g_ctl[i + 2], g_ctl[i + 3]);
    cudaEventElapsedTime(&d, g_ctl[i + 2], g_ctl[i + 4]);
    printf("step %2zu: trsm %.2f | next-column gemm %.2f | after hand-over: diagonal block +%.2f, herk +%.2f ms\n", i / 5, a, b, c, d);
  }
  g_ctl.clear();
}
#define CTL(s) ctl_mark(s)
#else
#define CTL(s) (void)0
#endif

int64_t potrf_block_width(int64_t n) {
  static const bool la_disabled = getenv("LA_B200_NO_LOOKAHEAD") != nullptr;
  static const int nbo_env = getenv("LA_B200_CHOL_NBO") ? atoi(getenv("LA_B200_CHOL_NBO")) : 0;
  if (n < 4096 || la_disabled) return n;
  // (measured, tools/chol_sweep.sh: 512 beats 1024 by 17 % at N = 8192 and 4 % at 16384; from
  // N = 32768 on the two are equal and the wider block halves the number of steps)
  return nbo_env > 0 ? nbo_env : (n >= 32768 ? 1024 : 512);
}

template <typename T>
void potrf(Context& ctx, Uplo uplo, int64_t n, T* A, int64_t lda, int* d_info) {
  LA_CUDA_CHECK(cudaMemsetAsync(d_info, 0, sizeof(int), ctx.stream));
  if (n <= 0) return;
  static DeviceOnce attr_set;
  if (attr_set.pending(ctx.device)) {
    LA_CUDA_CHECK(cudaFuncSetAttribute(potf2_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)potf2_smem_bytes<T>()));
    attr_set.mark(ctx.device);
  }
  static const bool la_disabled = getenv("LA_B200_NO_LOOKAHEAD") != nullptr;
  if (n < 4096 || la_disabled || !ctx.allow_side_stream) {
    potrf_twolevel<T>(ctx, uplo, n, A, lda, d_info, 0);
    return;
  }
  static const int side_sms_env = getenv("LA_B200_CHOL_SIDE_SMS") ? atoi(getenv("LA_B200_CHOL_SIDE_SMS")) : 0;
  const int side_sms = side_sms_env > 0 ? side_sms_env : (n >= 16384 ? 4 : 8);
  Context* side = side_context(ctx);
  {
    // first-use allocations of the side context happen before anything is enqueued
    Arena::Mark smk = side->arena.mark();
    (void)side->arena.alloc(1 << 20);
    side->arena.release(smk);
    LA_CUDA_CHECK(cudaStreamSynchronize(side->stream));
  }
  const int64_t NBO = potrf_block_width(n);
  auto at = [&](int64_t i, int64_t j) { return A + i + j * lda; };
  const T mone = from_real<T>(-1.0);
  const bool L = uplo == Lower;
  potrf_twolevel<T>(ctx, uplo, NBO < n ? NBO : n, A, lda, d_info, 0);
  for (int64_t K0 = 0; K0 < n; K0 += NBO) {
    const int64_t kb = NBO < n - K0 ? NBO : n - K0;
    const int64_t r2 = n - K0 - kb;
    if (r2 <= 0) break;
    const int64_t K1 = K0 + kb;
    const int64_t kb2 = NBO < r2 ? NBO : r2;
    const int64_t r3 = r2 - kb2;  // order of the trailing matrix beyond the next block column
    CTL(ctx.stream);
    if (L) {
      trsm<T>(ctx, Right, Lower, OpC, NonUnit, r2, kb, one<T>(), at(K0, K0), lda, at(K1, K0), lda);
      if (ctx.on_block_final) ctx.on_block_final(K0, K1);
      CTL(ctx.stream);
      gemm<T>(ctx, OpN, OpC, r2, kb2, kb, mone, at(K1, K0), lda, at(K1, K0), lda, one<T>(),
              at(K1, K1), lda, TriLower, Traits<T>::is_complex);
    } else {
      trsm<T>(ctx, Left, Upper, OpC, NonUnit, kb, r2, one<T>(), at(K0, K0), lda, at(K0, K1), lda);
      if (ctx.on_block_final) ctx.on_block_final(K0, K1);
      gemm<T>(ctx, OpC, OpN, kb2, r2, kb, mone, at(K0, K1), lda, at(K0, K1), lda, one<T>(),
              at(K1, K1), lda, TriUpper, Traits<T>::is_complex);
    }
    CTL(ctx.stream);
    LA_CUDA_CHECK(cudaEventRecord(ctx.ev[0], ctx.stream));
    LA_CUDA_CHECK(cudaStreamWaitEvent(side->stream, ctx.ev[0], 0));
    side->sm_limit = side_sms;
    potrf_twolevel<T>(*side, uplo, kb2, at(K1, K1), lda, d_info, K1);
    CTL(side->stream);
    LA_CUDA_CHECK(cudaEventRecord(side->ev[0], side->stream));
    if (r3 > 0) {
      const int keep = ctx.sm_limit;
      ctx.sm_limit = ctx.sm_count - side_sms;
      if (L)
        herk<T>(ctx, Lower, OpN, r3, kb, -1.0, at(K1 + kb2, K0), lda, 1.0, at(K1 + kb2, K1 + kb2), lda);
      else
        herk<T>(ctx, Upper, OpC, r3, kb, -1.0, at(K0, K1 + kb2), lda, 1.0, at(K1 + kb2, K1 + kb2), lda);
      ctx.sm_limit = keep;
    }
    CTL(ctx.stream);
    LA_CUDA_CHECK(cudaStreamWaitEvent(ctx.stream, side->ev[0], 0));
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

#ifdef LA_PANEL_TRACE
extern "C" void la_b200_debug_potf_trace(long long* out) {
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(out, g_potf_trace, sizeof(long long) * 32);
}
#endif

#define LA_INST_CHOL(T)                                                   \
  template void potrf<T>(Context&, Uplo, int64_t, T*, int64_t, int*);     \
  template void potrs<T>(Context&, Uplo, int64_t, int64_t, const T*, int64_t, T*, int64_t);
LA_INST_CHOL(float)
LA_INST_CHOL(double)
LA_INST_CHOL(zcomplex)

}  // namespace la
