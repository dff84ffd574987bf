// la_b200 — LU factorisation and the solves built on it (Cholesky lives in cholesky.cu).
//
// Reference call sites replaced: ?getrf_/?getrs_/?gesv_ (src/linalg/lu.cr:74-80,151-162,
// src/linalg/linalg.cr:443-445), ?trtrs_ (src/linalg/linalg.cr:431-432).
//
// Blocked right-looking driver around a recursive panel factorisation; the leaf panel kernel
// is in lu_panel.cu, the row interchanges (net-permutation laswp) are here.
#include "la_internal.h"
#include <type_traits>
#include <cooperative_groups.h>
#include <vector>

namespace la {

// =========================================================================== LU
// lu_panel.cu
template <typename T>
void lu_panel(Context& ctx, int64_t m, int jb, T* A, int64_t lda, int* ipiv, int64_t row0,
              int* d_info);
int lu_panel_max_width(Context& ctx, int64_t m, size_t elem_bytes, bool wide);
int lu_panel_ctas(Context& ctx, int64_t m, size_t elem_bytes, bool wide);
bool lu_panel_wide_ok(int64_t m);
void lu_panel_prepare(Context& ctx);
void lu_panel_begin(Context& ctx, int64_t m);

namespace {

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

// Row interchanges inside a panel are queued and flushed together: in the recursion an
// interchange of a node's left columns is directly followed by interchanges of other columns
// (its parent's), with no other work between them, and a launch in this dependent chain costs
// more than the interchange itself.
struct SwapJob { int64_t col0; int ncols, k1, k2; };
struct LuPlan {
  int nb0;                       // leaf width for the panel being factored (depends on its height)
  std::vector<SwapJob> pending;  // queued interchanges of the panel being factored
};
template <typename T>
void flush_swaps(Context& ctx, int64_t m, T* A, int64_t lda, const int* ipiv, LuPlan& plan);

// Factor columns [j0, j0+jn) of the m-row matrix A below row j0 (all earlier updates
// applied); pivots are global row indices.  Interchanges may still be queued on return:
// the caller flushes them (flush_swaps) before it touches the panel's columns.
template <typename T>
void getrf_rec(Context& ctx, int64_t m, T* A, int64_t lda, int* ipiv, int* d_info, int64_t j0,
               int64_t jn, LuPlan& plan) {
  if (jn <= plan.nb0) {
    flush_swaps<T>(ctx, m, A, lda, ipiv, plan);
    lu_panel<T>(ctx, m - j0, (int)jn, A + j0 + j0 * lda, lda, ipiv + j0, j0, d_info);
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
  plan.pending.push_back(SwapJob{j0 + n1, (int)n2, (int)j0, (int)(j0 + n1)});
  flush_swaps<T>(ctx, m, A, lda, ipiv, plan);
  trsm<T>(ctx, Left, Lower, OpN, Unit, n1, n2, one<T>(), A11, lda, A12, lda);
  gemm<T>(ctx, OpN, OpN, m - j0 - n1, n2, n1, from_real<T>(-1.0), A21, lda, A12, lda, one<T>(),
          A22, lda);
  getrf_rec<T>(ctx, m, A, lda, ipiv, d_info, j0 + n1, n2, plan);
  plan.pending.push_back(SwapJob{j0, (int)n1, (int)(j0 + n1), (int)(j0 + jn)});
}


// Net permutation of the interchange sequence ipiv[k1..k2): one CTA simulates the swaps on an
// index map held in shared memory (O(1) per pivot) and emits, for every touched row, which
// original row ends up there: entries e < npiv are the top block rows k1+e, entries
// npiv..2*npiv-1 the pivot rows themselves (duplicates are harmless: same value, same place).
__global__ void __launch_bounds__(256)
laswp_prep_kernel(int k1, int npiv, int mapn, const int* __restrict__ ipiv, int* __restrict__ dst,
                  int* __restrict__ src) {
  extern __shared__ int content[];   // content[i] = original row now at row k1+i ; then pivots
  int* piv = content + mapn;
  for (int i = threadIdx.x; i < mapn; i += blockDim.x) content[i] = k1 + i;
  for (int e = threadIdx.x; e < npiv; e += blockDim.x) piv[e] = ipiv[k1 + e] - 1 - k1;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int e = 0; e < npiv; e++) {
      const int p = piv[e];
      const int a = content[e], b = content[p];
      content[e] = b;
      content[p] = a;
    }
  }
  __syncthreads();
  for (int e = threadIdx.x; e < npiv; e += blockDim.x) {
    dst[e] = k1 + e;
    src[e] = content[e];
    const int p = piv[e];
    dst[npiv + e] = k1 + p;
    src[npiv + e] = content[p];
  }
}

// Apply a net row permutation (dst[e] <- src[e], e < nent) to n columns: each CTA stages one
// column's touched rows in shared memory, then scatters them.
template <typename T>
__global__ void __launch_bounds__(256)
laswp_apply_kernel(int64_t n, T* __restrict__ A, int64_t lda, int nent, const int* __restrict__ dst,
                   const int* __restrict__ src) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T* buf = reinterpret_cast<T*>(smem_raw);
  for (int64_t c = blockIdx.x; c < n; c += gridDim.x) {
    T* col = A + c * lda;
    for (int e = threadIdx.x; e < nent; e += blockDim.x) buf[e] = col[src[e]];
    __syncthreads();
    for (int e = threadIdx.x; e < nent; e += blockDim.x) {
      const int d = dst[e];
      if (d != src[e]) col[d] = buf[e];
    }
    __syncthreads();
  }
}

constexpr int LASWP_CHUNK = 1024;        // pivots per prep/apply pair
constexpr int LASWP_MAX_MAP = 49152;     // rows the shared-memory index map can cover

}  // namespace

// `m` = number of rows of A (upper bound of the pivot indices); 0 = unknown (serial kernel).
template <typename T>
void laswp_rows(Context& ctx, int64_t m, int64_t n, T* A, int64_t lda, int64_t k1, int64_t k2,
                const int* d_ipiv, bool reverse) {
  if (n <= 0 || k2 <= k1) return;
  static DeviceOnce attr;
  if (attr.pending(ctx.device)) {
    LA_CUDA_CHECK(cudaFuncSetAttribute(laswp_prep_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (LASWP_MAX_MAP + LASWP_CHUNK) * (int)sizeof(int)));
    attr.mark(ctx.device);
  }
  Arena::Mark mk = ctx.arena.mark();
  int* lists = ctx.alloc<int>(4 * LASWP_CHUNK);
  // reverse = true undoes the interchanges (last first): the chunks are visited backwards and
  // each chunk's net permutation is applied inverted (source and destination lists swapped)
  const int64_t nchunk = ceil_div(k2 - k1, (int64_t)LASWP_CHUNK);
  for (int64_t ci = 0; ci < nchunk; ci++) {
    const int64_t c1 = k1 + (reverse ? nchunk - 1 - ci : ci) * LASWP_CHUNK;
    const int64_t c2 = c1 + LASWP_CHUNK < k2 ? c1 + LASWP_CHUNK : k2;
    const int npiv = (int)(c2 - c1);
    const int64_t mapn = m - c1;
    if (m <= 0 || mapn > LASWP_MAX_MAP) {
      laswp_kernel<T><<<(unsigned)ceil_div(n, 128), 128, 0, ctx.stream>>>(n, A, lda, c1, c2, d_ipiv,
                                                                          reverse ? 1 : 0);
      LA_CUDA_CHECK(cudaGetLastError());
      continue;
    }
    int* dst = lists;
    int* src = lists + 2 * LASWP_CHUNK;
    laswp_prep_kernel<<<1, 256, (size_t)(mapn + npiv) * sizeof(int), ctx.stream>>>(
        (int)c1, npiv, (int)mapn, d_ipiv, dst, src);
    const int nent = 2 * npiv;
    int64_t grid = n < (int64_t)ctx.sm_count * 8 ? n : (int64_t)ctx.sm_count * 8;
    laswp_apply_kernel<T><<<(unsigned)grid, 256, (size_t)nent * sizeof(T), ctx.stream>>>(
        n, A, lda, nent, reverse ? src : dst, reverse ? dst : src);
    LA_CUDA_CHECK(cudaGetLastError());
  }
  ctx.arena.release(mk);
}

namespace {

// The same on a capped number of CTAs (one per SM the caller may use, beside a kernel of another
// stream that needs its SMs whole): CB columns per pass keep enough loads in flight.
template <typename T, int CB>
__global__ void __launch_bounds__(1024)
laswp_apply_capped_kernel(int64_t n, T* __restrict__ A, int64_t lda, int nent,
                          const int* __restrict__ dst, const int* __restrict__ src) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T* buf = reinterpret_cast<T*>(smem_raw);
  for (int64_t c0 = (int64_t)blockIdx.x * CB; c0 < n; c0 += (int64_t)gridDim.x * CB) {
    for (int e = threadIdx.x; e < nent; e += blockDim.x) {
      const int64_t s = src[e];
#pragma unroll
      for (int i = 0; i < CB; i++)
        if (c0 + i < n) buf[i * nent + e] = A[(c0 + i) * lda + s];
    }
    __syncthreads();
    for (int e = threadIdx.x; e < nent; e += blockDim.x) {
      const int64_t d = dst[e];
      if (d != src[e]) {
#pragma unroll
        for (int i = 0; i < CB; i++)
          if (c0 + i < n) A[(c0 + i) * lda + d] = buf[i * nent + e];
      }
    }
    __syncthreads();
  }
}

// Net permutation of ipiv[k1, k2) into `lists` (4 * LASWP_CHUNK ints), for several applications
// to different column ranges.  False: not representable (too many pivots / rows): use laswp_rows.
bool laswp_lists(Context& ctx, int64_t m, int64_t k1, int64_t k2, const int* d_ipiv, int* lists) {
  const int64_t mapn = m - k1;
  if (k2 - k1 > LASWP_CHUNK || m <= 0 || mapn > LASWP_MAX_MAP) return false;
  static DeviceOnce attr;
  if (attr.pending(ctx.device)) {
    LA_CUDA_CHECK(cudaFuncSetAttribute(laswp_prep_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (LASWP_MAX_MAP + LASWP_CHUNK) * (int)sizeof(int)));
    attr.mark(ctx.device);
  }
  const int npiv = (int)(k2 - k1);
  laswp_prep_kernel<<<1, 256, (size_t)(mapn + npiv) * sizeof(int), ctx.stream>>>(
      (int)k1, npiv, (int)mapn, d_ipiv, lists, lists + 2 * LASWP_CHUNK);
  LA_CUDA_CHECK(cudaGetLastError());
  return true;
}

// Apply prepared lists to n columns; honours ctx.sm_limit (one CTA per permitted SM).
template <typename T>
void laswp_apply_lists(Context& ctx, int64_t n, T* A, int64_t lda, int npiv, const int* lists) {
  if (n <= 0) return;
  const int nent = 2 * npiv;
  const int* dst = lists;
  const int* src = lists + 2 * LASWP_CHUNK;
  if (ctx.sm_limit > 0 && ctx.sm_limit < ctx.sm_count) {
    constexpr int CB = 4;
    static DeviceOnce attr;
    if (attr.pending(ctx.device)) {
      LA_CUDA_CHECK(cudaFuncSetAttribute(laswp_apply_capped_kernel<T, CB>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         2 * LASWP_CHUNK * CB * (int)sizeof(T)));
      attr.mark(ctx.device);
    }
    int64_t grid = ceil_div(n, (int64_t)CB);
    if (grid > ctx.sm_limit) grid = ctx.sm_limit;
    laswp_apply_capped_kernel<T, CB><<<(unsigned)grid, 1024, (size_t)nent * CB * sizeof(T),
                                       ctx.stream>>>(n, A, lda, nent, dst, src);
  } else {
    int64_t grid = n < (int64_t)ctx.sm_count * 8 ? n : (int64_t)ctx.sm_count * 8;
    laswp_apply_kernel<T><<<(unsigned)grid, 256, (size_t)nent * sizeof(T), ctx.stream>>>(
        n, A, lda, nent, dst, src);
  }
  LA_CUDA_CHECK(cudaGetLastError());
}

constexpr int LASWP_MAX_JOBS = 6;
struct SwapJobs {
  int njobs;
  long long col0[LASWP_MAX_JOBS];
  int ncols[LASWP_MAX_JOBS], k1[LASWP_MAX_JOBS], k2[LASWP_MAX_JOBS];
  int cta0[LASWP_MAX_JOBS + 1];            // first CTA of each job
};

// Up to LASWP_MAX_JOBS interchange jobs (columns, pivot range) in one launch.  Every CTA builds
// the net permutation of its job in shared memory (as laswp_prep_kernel does) and applies it to
// its share of the job's columns.  For the narrow column ranges inside a panel.
template <typename T>
__global__ void __launch_bounds__(256)
laswp_jobs_kernel(int m, T* __restrict__ A, int64_t lda, const int* __restrict__ ipiv, SwapJobs jobs) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  int j = 0;
  while (j + 1 < jobs.njobs && (int)blockIdx.x >= jobs.cta0[j + 1]) j++;
  const int k1 = jobs.k1[j], npiv = jobs.k2[j] - k1, mapn = m - k1;
  int* content = reinterpret_cast<int*>(smem_raw);
  int* piv = content + mapn;
  T* buf = reinterpret_cast<T*>(smem_raw + (((size_t)(mapn + npiv) * sizeof(int) + 15) & ~(size_t)15));
  for (int i = threadIdx.x; i < mapn; i += blockDim.x) content[i] = k1 + i;
  for (int e = threadIdx.x; e < npiv; e += blockDim.x) piv[e] = ipiv[k1 + e] - 1 - k1;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int e = 0; e < npiv; e++) {
      const int p = piv[e];
      const int a = content[e], b = content[p];
      content[e] = b;
      content[p] = a;
    }
  }
  __syncthreads();
  const int nent = 2 * npiv;
  const int nc = jobs.ncols[j], first = (int)blockIdx.x - jobs.cta0[j], step = jobs.cta0[j + 1] - jobs.cta0[j];
  for (int c = first; c < nc; c += step) {
    T* col = A + (jobs.col0[j] + c) * lda;
    for (int e = threadIdx.x; e < nent; e += blockDim.x) {
      const int src = e < npiv ? content[e] : content[piv[e - npiv]];
      buf[e] = col[src];
    }
    __syncthreads();
    for (int e = threadIdx.x; e < nent; e += blockDim.x) {
      const int dst = e < npiv ? k1 + e : k1 + piv[e - npiv];
      const int src = e < npiv ? content[e] : content[piv[e - npiv]];
      if (dst != src) col[dst] = buf[e];
    }
    __syncthreads();
  }
}

constexpr size_t LASWP_JOBS_SMEM = 200 * 1024;

template <typename T>
void flush_swaps(Context& ctx, int64_t m, T* A, int64_t lda, const int* ipiv, LuPlan& plan) {
  if (plan.pending.empty()) return;
  static DeviceOnce attr;
  if (attr.pending(ctx.device)) {
    LA_CUDA_CHECK(cudaFuncSetAttribute(laswp_jobs_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)LASWP_JOBS_SMEM));
    attr.mark(ctx.device);
  }
  SwapJobs jobs;
  jobs.njobs = 0;
  jobs.cta0[0] = 0;
  size_t smem = 0;
  auto launch = [&] {
    if (jobs.njobs == 0) return;
    laswp_jobs_kernel<T><<<(unsigned)jobs.cta0[jobs.njobs], 256, smem, ctx.stream>>>((int)m, A, lda, ipiv, jobs);
    LA_CUDA_CHECK(cudaGetLastError());
    jobs.njobs = 0;
    smem = 0;
  };
  for (const SwapJob& jb : plan.pending) {
    if (jb.ncols <= 0 || jb.k2 <= jb.k1) continue;
    const size_t need = (((size_t)(m - jb.k1 + (jb.k2 - jb.k1)) * sizeof(int) + 15) & ~(size_t)15) +
                        (size_t)2 * (jb.k2 - jb.k1) * sizeof(T);
    if (need > LASWP_JOBS_SMEM || jb.ncols > 1024 || m > 0x7fffffff) {
      launch();                        // keep the order: everything queued before goes first
      laswp_rows<T>(ctx, m, jb.ncols, A + jb.col0 * lda, lda, jb.k1, jb.k2, ipiv);
      continue;
    }
    if (jobs.njobs == LASWP_MAX_JOBS) launch();
    const int q = jobs.njobs++;
    jobs.col0[q] = jb.col0; jobs.ncols[q] = jb.ncols; jobs.k1[q] = jb.k1; jobs.k2[q] = jb.k2;
    const int ctas = jb.ncols < 64 ? jb.ncols : 64;
    jobs.cta0[q + 1] = jobs.cta0[q] + ctas;
    if (need > smem) smem = need;
  }
  launch();
  plan.pending.clear();
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

#ifdef LA_LU_TIMELINE
#include <vector>
static std::vector<cudaEvent_t> g_tl;   // per step: start, ev0 (serial part done), side done, part1 done, part2 done
static cudaEvent_t tl_mark(cudaStream_t s) {
  cudaEvent_t e; cudaEventCreate(&e); cudaEventRecord(e, s); g_tl.push_back(e); return e;
}
extern "C" void la_b200_debug_lu_timeline() {
  cudaDeviceSynchronize();
  for (size_t i = 0; i + 4 < g_tl.size(); i += 5) {
    float a, b, c, d;
    cudaEventElapsedTime(&a, g_tl[i], g_tl[i + 1]);
    cudaEventElapsedTime(&b, g_tl[i + 1], g_tl[i + 2]);
    cudaEventElapsedTime(&c, g_tl[i + 1], g_tl[i + 3]);
    cudaEventElapsedTime(&d, g_tl[i + 1], g_tl[i + 4]);
    printf("step %2zu: serial %.2f | after ev0: panel done +%.2f, part1 done +%.2f, part2 done +%.2f ms\n", i / 5, a, b, c, d);
  }
  g_tl.clear();
}
#define TL(s) tl_mark(s)
#else
#define TL(s) (void)0
#endif

template <typename T>
void getrf(Context& ctx, int64_t m, int64_t n, T* A, int64_t lda, int* d_ipiv, int* d_info) {
  LA_CUDA_CHECK(cudaMemsetAsync(d_info, 0, sizeof(int), ctx.stream));
  const int64_t k = m < n ? m : n;
  if (k <= 0) return;
  LuPlan plan;
  const int64_t nb0 = lu_panel_max_width(ctx, m, sizeof(T), false);
  if (nb0 < 1) {
    fprintf(stderr, "la_b200: getrf: %lld rows exceed the panel kernel's capacity\n", (long long)m);
    abort();
  }
  plan.nb0 = (int)nb0;
  // Right-looking blocked LU: recursive factorisation of an m x NB panel, then row
  // interchanges of the columns left and right of it, U12 = L11^-1 A12 and the rank-NB
  // trailing update (where nearly all flops are).  Keeping the interchanges per panel bounds
  // the serial part of laswp's net-permutation set-up to NB steps.
  static const int nb_env = getenv("LA_B200_LU_NB") ? atoi(getenv("LA_B200_LU_NB")) : 0;
  const int64_t NB = nb_env > 0 ? nb_env : (k >= 8192 ? 512 : (k >= 2048 ? 256 : 128));
  // Look-ahead: once the columns of the NEXT panel have received their update, that panel is
  // factored on a side stream (a chain of latency-bound leaf launches that needs only a few
  // SMs) while the main stream finishes the trailing update on the remaining SMs.
  const int64_t leaf_ctas = ceil_div(m, 512);
  // Only when the overlapped trailing update is essentially ONE persistent kernel with a capped
  // grid (the FP64 DMMA GEMM on TMA-legal operands; FP32 below): the cooperative leaf needs its
  // SMs whole and all at once, and a trailing update made of many ordinary launches (SIMT GEMM)
  // keeps taking them away between two leaves.
  const bool dmma_ok = std::is_same<T, double>::value && (lda % 2 == 0) &&
                       (((uintptr_t)A & 15) == 0);
  // FP32: the update is the persistent 3xTF32 kernel behind two short packing launches, which
  // hold the leaf up for a few microseconds at most (LA_B200_LU_LOOKAHEAD_F32=0 turns it off)
  static const bool f32_la = !(getenv("LA_B200_LU_LOOKAHEAD_F32") && atoi(getenv("LA_B200_LU_LOOKAHEAD_F32")) == 0);
  const bool tf32_ok = f32_la && std::is_same<T, float>::value && (lda % 4 == 0) &&
                       (((uintptr_t)A & 15) == 0);
  // (complex128 was measured with the same arrangement and did not gain: 95.9 against 93.0 ms at
  // N = 8192, 470 against 456 ms at 16384; it stays on the main stream)
  const double gemm_flop_per_ms = std::is_same<T, float>::value ? 230.0e9 : 33.0e9;
  static const bool la_disabled = getenv("LA_B200_NO_LOOKAHEAD") != nullptr;
  const bool lookahead = !la_disabled && ctx.allow_side_stream && (dmma_ok || tf32_ok) && k >= 8192 &&
                         leaf_ctas <= 96 && k > NB;
  Context* side = lookahead ? side_context(ctx) : nullptr;
  if (side) {
    // every first-use allocation of the side context (stream, arena block, pivot-exchange
    // area) is made here, before anything is enqueued: creating them while the cooperative
    // leaf kernel of an earlier step is running has been seen to hang the device
    Arena::Mark smk = side->arena.mark();
    (void)side->arena.alloc(1 << 20);
    side->arena.release(smk);
    lu_panel_prepare(*side);
    lu_panel_begin(*side, m);
    LA_CUDA_CHECK(cudaStreamSynchronize(side->stream));
  }
  lu_panel_begin(ctx, m);
  bool panel_done = false;  // panel of the current j0 already factored (on the side stream)?
  for (int64_t j0 = 0; j0 < k; j0 += NB) {
    const int64_t jb = NB < k - j0 ? NB : k - j0;
    const bool tl_on = panel_done || j0 == 0;
    if (tl_on && lookahead) TL(ctx.stream);
    if (!panel_done) {
      plan.nb0 = lu_panel_max_width(ctx, m - j0, sizeof(T), true);   // nothing else runs beside it
      getrf_rec<T>(ctx, m, A, lda, d_ipiv, d_info, j0, jb, plan);
      flush_swaps<T>(ctx, m, A, lda, d_ipiv, plan);
    }
    panel_done = false;
    const int64_t nr = n - j0 - jb;  // columns to the right
    const int64_t mr = m - j0 - jb;
    const int64_t j1 = j0 + jb;
    const int64_t jb1 = j1 < k ? (NB < k - j1 ? NB : k - j1) : 0;
    const bool ahead = lookahead && jb1 > 0 && nr > jb1 && mr > 0;
    // Timing model of the step (ms; B200, FP64): the trailing update on the whole GPU, and the
    // next panel (m - j1 rows) when its own updates get plenty of SMs or just those of its
    // leaves.  The panel is a chain of latency-bound launches whose length hardly depends on
    // its height.
    static const double model_scale = getenv("LA_B200_LU_MODEL") ? atof(getenv("LA_B200_LU_MODEL")) : 1.1;
    static const bool split_env = !(getenv("LA_B200_LU_SPLIT") && atoi(getenv("LA_B200_LU_SPLIT")) == 0);
    const double h = (double)(m - j1) / 16384.0;
    const double upd_ms = 2.0 * (double)(m - j1) * (double)(nr - jb1) * (double)jb / gemm_flop_per_ms;
    const double pan_free_ms = 1.5 + 0.3 * h;
    const double pan_tight_ms = model_scale * (1.6 + 0.9 * h);
    // Steps whose update is shorter than the next panel are bound by the chain
    //   interchanges -> U12 -> update of the next panel's columns -> next panel,
    // so those columns go first and everything else (the other columns' interchanges and U12)
    // runs beside the panel, on the SMs it leaves free.
    Arena::Mark step_mk = ctx.arena.mark();
    int* lists = nullptr;
    static const double split_frac = getenv("LA_B200_LU_SPLIT_FRAC") ? atof(getenv("LA_B200_LU_SPLIT_FRAC")) : 0.6;
    bool split = ahead && split_env && upd_ms < split_frac * pan_free_ms;
    if (split) {
      lists = ctx.alloc<int>(4 * LASWP_CHUNK);
      split = laswp_lists(ctx, m, j0, j0 + jb, d_ipiv, lists);
    }
    T* A12 = A + j0 + (j0 + jb) * lda;
    if (!split) {
      if (j0 > 0) laswp_rows<T>(ctx, m, j0, A, lda, j0, j0 + jb, d_ipiv);
      if (nr <= 0) { ctx.arena.release(step_mk); continue; }
      laswp_rows<T>(ctx, m, nr, A + (j0 + jb) * lda, lda, j0, j0 + jb, d_ipiv);
      trsm<T>(ctx, Left, Lower, OpN, Unit, jb, nr, one<T>(), A + j0 + j0 * lda, lda, A12, lda);
      // rows [j0, j0 + jb) of the whole matrix are final now: later panels interchange rows below
      if (ctx.on_block_final) ctx.on_block_final(j0, j0 + jb);
    } else {
      laswp_apply_lists<T>(ctx, jb1, A + j1 * lda, lda, (int)jb, lists);
      trsm<T>(ctx, Left, Lower, OpN, Unit, jb, jb1, one<T>(), A + j0 + j0 * lda, lda, A12, lda);
    }
    if (mr <= 0) { ctx.arena.release(step_mk); continue; }
    const T* A21 = A + (j0 + jb) + j0 * lda;
    T* A22 = A + (j0 + jb) + (j0 + jb) * lda;
    if (ahead) {
      // the wide leaf needs twice the SMs per row: worth it once the update is small beside the panel
      const bool wide = lu_panel_wide_ok(m - j1) && upd_ms < 0.5 * pan_free_ms;
      const int64_t ctas1 = lu_panel_ctas(ctx, m - j1, sizeof(T), wide);
      int side_sms = (int)(ctas1 < 16 ? 16 : ctas1);
      {
        // late steps: the trailing update is shorter than the panel, so the panel's own
        // GEMMs / TRSMs get the SMs the update does not need to finish in the panel's time
        // (a split step also runs the deferred interchanges and U12 there: bandwidth-bound work
        // that wants a fair number of SMs)
        const double beside_ms = upd_ms + (split ? 0.15 + 0.25 * (double)(j0 + nr) / 16384.0 : 0.0);
        const int need = (int)(ctx.sm_count * beside_ms / pan_free_ms) + 1;
        const int floor_sms = split ? 64 : 32;
        int spare = ctx.sm_count - (need < floor_sms ? floor_sms : need);
        if (spare > 116) spare = 116;
        if (spare > side_sms) side_sms = spare;
      }
      gemm<T>(ctx, OpN, OpN, mr, jb1, jb, from_real<T>(-1.0), A21, lda, A12, lda, one<T>(), A22, lda);
      TL(ctx.stream);
      LA_CUDA_CHECK(cudaEventRecord(ctx.ev[0], ctx.stream));
      LA_CUDA_CHECK(cudaStreamWaitEvent(side->stream, ctx.ev[0], 0));
      side->sm_limit = side_sms;
      plan.nb0 = lu_panel_max_width(ctx, m - j1, sizeof(T), wide);
      getrf_rec<T>(*side, m, A, lda, d_ipiv, d_info, j1, jb1, plan);
      flush_swaps<T>(*side, m, A, lda, d_ipiv, plan);
      TL(side->stream);
      LA_CUDA_CHECK(cudaEventRecord(side->ev[0], side->stream));
      // The trailing update runs on the SMs the side stream leaves free — but only for as long
      // as the panel is expected to take (a model of the panel's duration, biased long: an
      // update that outlasts the panel merely keeps a few SMs idle, one that ends early stalls
      // the whole device); the rest of it is launched on all SMs once the panel is done.
      const int64_t rest = nr - jb1;
      const int keep = ctx.sm_limit;
      ctx.sm_limit = ctx.sm_count - side_sms;
      if (split) {
        if (j0 > 0) laswp_apply_lists<T>(ctx, j0, A, lda, (int)jb, lists);
        laswp_apply_lists<T>(ctx, rest, A + (j1 + jb1) * lda, lda, (int)jb, lists);
        trsm<T>(ctx, Left, Lower, OpN, Unit, jb, rest, one<T>(), A + j0 + j0 * lda, lda,
                A12 + jb1 * lda, lda);
        if (ctx.on_block_final) ctx.on_block_final(j0, j0 + jb);
      }
      const double panel_ms = side_sms > ctas1 + 16 ? pan_free_ms : pan_tight_ms;
      const double flop_per_ms = gemm_flop_per_ms * (double)(ctx.sm_count - side_sms) / 148.0;
      int64_t c1 = (int64_t)(panel_ms * flop_per_ms / (2.0 * (double)mr * (double)jb));
      c1 = round_up(c1 > 0 ? c1 : 1, (int64_t)128);
      if (c1 + 1024 >= rest) c1 = rest;
      gemm<T>(ctx, OpN, OpN, mr, c1, jb, from_real<T>(-1.0), A21, lda, A12 + jb1 * lda, lda,
              one<T>(), A22 + jb1 * lda, lda);
      ctx.sm_limit = keep;
      TL(ctx.stream);
      LA_CUDA_CHECK(cudaStreamWaitEvent(ctx.stream, side->ev[0], 0));
      if (c1 < rest)
        gemm<T>(ctx, OpN, OpN, mr, rest - c1, jb, from_real<T>(-1.0), A21, lda,
                A12 + (jb1 + c1) * lda, lda, one<T>(), A22 + (jb1 + c1) * lda, lda);
      TL(ctx.stream);
      panel_done = true;
    } else {
      gemm<T>(ctx, OpN, OpN, mr, nr, jb, from_real<T>(-1.0), A21, lda, A12, lda, one<T>(), A22, lda);
    }
    ctx.arena.release(step_mk);
  }
}

template <typename T>
void getrs(Context& ctx, Op trans, int64_t n, int64_t nrhs, const T* A, int64_t lda,
           const int* d_ipiv, T* B, int64_t ldb) {
  if (n <= 0 || nrhs <= 0) return;
  if (!Traits<T>::is_complex && trans == OpC) trans = OpT;
  if (trans == OpN) {
    laswp_rows<T>(ctx, n, nrhs, B, ldb, 0, n, d_ipiv);
    trsm<T>(ctx, Left, Lower, OpN, Unit, n, nrhs, one<T>(), A, lda, B, ldb);
    trsm<T>(ctx, Left, Upper, OpN, NonUnit, n, nrhs, one<T>(), A, lda, B, ldb);
  } else {
    trsm<T>(ctx, Left, Upper, trans, NonUnit, n, nrhs, one<T>(), A, lda, B, ldb);
    trsm<T>(ctx, Left, Lower, trans, Unit, n, nrhs, one<T>(), A, lda, B, ldb);
    laswp_rows<T>(ctx, n, nrhs, B, ldb, 0, n, d_ipiv, true);
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
  template void laswp<T>(Context&, int64_t, T*, int64_t, int64_t, int64_t, const int*, bool);     \
  template void laswp_rows<T>(Context&, int64_t, int64_t, T*, int64_t, int64_t, int64_t,          \
                              const int*, bool);                                                  \
  template void getrf<T>(Context&, int64_t, int64_t, T*, int64_t, int*, int*);                    \
  template void getrs<T>(Context&, Op, int64_t, int64_t, const T*, int64_t, const int*, T*,       \
                         int64_t);                                                                \
  template void trtrs<T>(Context&, Uplo, Op, Diag, int64_t, int64_t, const T*, int64_t, T*,       \
                         int64_t, int*);
LA_INST_FACTOR(float)
LA_INST_FACTOR(double)
LA_INST_FACTOR(zcomplex)

}  // namespace la
