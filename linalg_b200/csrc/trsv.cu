// la_b200 — triangular solve with a FEW right-hand sides (nrhs <= 8) in one cooperative launch.
//
// ?getrs / ?potrs / ?trtrs as the reference calls them from `solve`, `cho_solve` and `solvels`
// with a vector or a handful of columns (src/linalg/linalg.cr:425-450, cholesky.cr:75-90).  The
// recursive TRSM of level3.cu turns such a solve into ~n/64 * 2 tiny launches (30-50 ms at
// n = 16384); the work itself is one pass over the triangle (n^2/2 elements, HBM bound) plus a
// serial chain of n/64 diagonal-block solves.  Here:
//   * the rows are cut in blocks of 64, dealt round-robin to the CTAs in dependency order;
//   * the CTA that owns row block i accumulates  b_i - sum_j M(i,j) x_j  over the solved blocks j
//     (tile M(i,j) prefetched while it waits for x_j), solves the 64x64 diagonal block by warp
//     shuffles (one warp per right-hand side) and publishes x_i;
//   * x_j travels as 16-byte (payload, tag) packets polled by the readers (one L2 round trip, no
//     fence, no atomic), as in the LU panel kernel.
// Algorithmic bytes: n^2/2 * sizeof(T) read once; chain: n/64 steps of (poll + 64x64 update +
// diagonal solve + publish).
#include "la_internal.h"

namespace la {

namespace {

constexpr int TV_RB = 64;          // rows per block
constexpr int TV_NT = 256;         // threads per CTA: 64 rows x 4 column quarters
constexpr int TV_MAXRHS = 8;       // one warp per right-hand side in the diagonal solve

struct __align__(16) TvPacket {
  unsigned long long data;
  unsigned long long tag;
};

template <typename T> struct TvWords { static constexpr int N = 1; };
template <> struct TvWords<zcomplex> { static constexpr int N = 2; };

__device__ __forceinline__ void tv_store(TvPacket* p, unsigned long long d, unsigned long long t) {
  asm volatile("st.global.cg.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(d), "l"(t) : "memory");
}
__device__ __forceinline__ unsigned long long tv_wait(const TvPacket* p, unsigned long long tag) {
  unsigned long long d, t;
  do {
    asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(d), "=l"(t) : "l"(p) : "memory");
  } while (t != tag);
  return d;
}

__device__ __forceinline__ void tv_put(TvPacket* p, double v, unsigned long long tag) {
  tv_store(p, (unsigned long long)__double_as_longlong(v), tag);
}
__device__ __forceinline__ void tv_put(TvPacket* p, float v, unsigned long long tag) {
  tv_store(p, (unsigned long long)__float_as_uint(v), tag);
}
__device__ __forceinline__ void tv_put(TvPacket* p, zcomplex v, unsigned long long tag) {
  tv_store(p, (unsigned long long)__double_as_longlong(v.x), tag);
  tv_store(p + 1, (unsigned long long)__double_as_longlong(v.y), tag);
}
// raw 8-byte word -> element storage (xs is addressed in words for double / zcomplex, in
// elements for float)
template <typename T> __device__ __forceinline__ void tv_land(T* xs, int word, unsigned long long d);
template <> __device__ __forceinline__ void tv_land<double>(double* xs, int word, unsigned long long d) {
  xs[word] = __longlong_as_double((long long)d);
}
template <> __device__ __forceinline__ void tv_land<float>(float* xs, int word, unsigned long long d) {
  xs[word] = __uint_as_float((unsigned)d);
}
template <> __device__ __forceinline__ void tv_land<zcomplex>(zcomplex* xs, int word, unsigned long long d) {
  reinterpret_cast<double*>(xs)[word] = __longlong_as_double((long long)d);
}

template <typename T> __device__ __forceinline__ T tv_shfl(T v, int src) {
  return __shfl_sync(0xffffffffu, v, src);
}
template <> __device__ __forceinline__ zcomplex tv_shfl<zcomplex>(zcomplex v, int src) {
  return zcomplex{__shfl_sync(0xffffffffu, v.x, src), __shfl_sync(0xffffffffu, v.y, src)};
}

struct TvParams {
  int64_t n;          // order of the triangle
  int nrhs;
  int lower;          // the EFFECTIVE matrix M = op(A) is lower triangular
  int unit;
  int conj;           // M(i,j) = conj(stored element)
  int64_t ur, uc;     // M(i,j) = A[i * ur + j * uc]
  int64_t ldb;
  unsigned long long tag;
};

template <typename T>
__global__ void __launch_bounds__(TV_NT)
trsv_kernel(const TvParams p, const T alpha, const T* __restrict__ A, T* __restrict__ B,
            TvPacket* __restrict__ pk) {
  constexpr int W = TvWords<T>::N;
  constexpr int LDD = TV_RB + 1;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T* Ds = reinterpret_cast<T*>(smem_raw);            // Ds[r * LDD + c] = M(i0 + r, i0 + c)
  T* xs = Ds + TV_RB * LDD;                          // xs[2][64 * MAXRHS]: landed x_j
  T* red = xs + 2 * TV_RB * TV_MAXRHS;               // red[4][64][MAXRHS]: partial sums
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);   // warp-uniform for the compiler
  const int r = tid & (TV_RB - 1), q = tid >> 6;
  const int nb = (int)((p.n + TV_RB - 1) / TV_RB);
  const int nrhs = p.nrhs;
  const bool lower = p.lower != 0, cj = p.conj != 0;
  const size_t blk_words = (size_t)TV_RB * TV_MAXRHS * W;
  int buf = 0;
  const bool vec_ok = p.uc == 1 && !cj && (p.ur % 2 == 0) && (((uintptr_t)A & 15) == 0);

  for (int o = blockIdx.x; o < nb; o += gridDim.x) {   // o: position in dependency order
    const int i = lower ? o : nb - 1 - o;
    const int64_t i0 = (int64_t)i * TV_RB;
    const int rows = (int)((p.n - i0) < TV_RB ? (p.n - i0) : TV_RB);
    // ---- diagonal block (identity padding); fast thread index along the contiguous direction
    for (int idx = tid; idx < TV_RB * TV_RB; idx += TV_NT) {
      const int fa = idx & (TV_RB - 1), sl = idx >> 6;
      const int rr = p.uc == 1 ? sl : fa, cc = p.uc == 1 ? fa : sl;
      T v = rr == cc ? one<T>() : zero<T>();
      if (rr < rows && cc < rows) {
        const bool stored = lower ? rr >= cc : rr <= cc;
        v = zero<T>();
        if (stored)
          v = (rr == cc && p.unit) ? one<T>()
                                   : apply_conj(A[(i0 + rr) * p.ur + (i0 + cc) * p.uc], cj);
      }
      Ds[rr * LDD + cc] = v;
    }
    // the right-hand side values of the solving warps are fetched now, off the serial chain
    T b0 = zero<T>(), b1 = zero<T>();
    if (warp < nrhs) {
      if (lane < rows) b0 = B[i0 + lane + (int64_t)warp * p.ldb];
      if (lane + 32 < rows) b1 = B[i0 + lane + 32 + (int64_t)warp * p.ldb];
    }
    // ---- b_i - sum_j M(i, j) x_j over the blocks solved before this one
    T acc[TV_MAXRHS];
#pragma unroll
    for (int nn = 0; nn < TV_MAXRHS; nn++) acc[nn] = zero<T>();
    T tile[16], nxt[16];
    auto load_tile = [&](int jj, T (&t)[16]) {
      const int64_t j0 = (int64_t)(lower ? jj : nb - 1 - jj) * TV_RB + 16 * q;
      if constexpr (sizeof(T) == 8) {
        // transposed access: a thread's 16 elements are contiguous -> eight 16-byte loads
        if (vec_ok && i0 + r < p.n && j0 + 16 <= p.n) {
          const double2* src2 = reinterpret_cast<const double2*>(A + (i0 + r) * p.ur + j0);
#pragma unroll
          for (int cc = 0; cc < 8; cc++) {
            const double2 v = __ldg(src2 + cc);
            t[2 * cc] = v.x;
            t[2 * cc + 1] = v.y;
          }
          return;
        }
      }
#pragma unroll
      for (int cc = 0; cc < 16; cc++)
        t[cc] = (i0 + r < p.n && j0 + cc < p.n)
                    ? apply_conj(A[(i0 + r) * p.ur + (j0 + cc) * p.uc], cj) : zero<T>();
    };
    if (o > 0) load_tile(0, tile);
    for (int jj = 0; jj < o; jj++) {
      if (jj + 1 < o) load_tile(jj + 1, nxt);      // in flight while x_j is awaited
      const int jb = lower ? jj : nb - 1 - jj;
      const TvPacket* src = pk + (size_t)jb * blk_words;
      T* xb = xs + buf * (TV_RB * TV_MAXRHS);
      for (int w = tid; w < TV_RB * nrhs * W; w += TV_NT) tv_land<T>(xb, w, tv_wait(src + w, p.tag));
      __syncthreads();
#pragma unroll
      for (int cc = 0; cc < 16; cc++) {
#pragma unroll
        for (int nn = 0; nn < TV_MAXRHS; nn++)
          if (nn < nrhs) acc[nn] = fma_(tile[cc], xb[(16 * q + cc) * nrhs + nn], acc[nn]);
      }
#pragma unroll
      for (int cc = 0; cc < 16; cc++) tile[cc] = nxt[cc];
      buf ^= 1;
    }
#pragma unroll
    for (int nn = 0; nn < TV_MAXRHS; nn++)
      if (nn < nrhs) red[(q * TV_RB + r) * TV_MAXRHS + nn] = acc[nn];
    __syncthreads();
    // ---- diagonal solve: warp nn owns right-hand side nn, lane l holds rows l and l + 32
    if (warp < nrhs) {
      const int nn = warp;
      const int r0 = lane, r1 = lane + 32;
      T v0 = mul(alpha, b0), v1 = mul(alpha, b1);
#pragma unroll
      for (int qq = 0; qq < 4; qq++) {
        v0 = sub(v0, red[(qq * TV_RB + r0) * TV_MAXRHS + nn]);
        v1 = sub(v1, red[(qq * TV_RB + r1) * TV_MAXRHS + nn]);
      }
      const T d0 = recip(Ds[r0 * LDD + r0]), d1 = recip(Ds[r1 * LDD + r1]);
      if (lower) {
#pragma unroll
        for (int c = 0; c < 32; c++) {
          v0 = lane == c ? mul(v0, d0) : v0;
          const T xc = tv_shfl(v0, c);
          const T u0 = fnma_(Ds[r0 * LDD + c], xc, v0);
          v0 = lane > c ? u0 : v0;
          v1 = fnma_(Ds[r1 * LDD + c], xc, v1);
        }
#pragma unroll
        for (int c = 0; c < 32; c++) {
          v1 = lane == c ? mul(v1, d1) : v1;
          const T xc = tv_shfl(v1, c);
          const T u1 = fnma_(Ds[r1 * LDD + 32 + c], xc, v1);
          v1 = lane > c ? u1 : v1;
        }
      } else {
#pragma unroll
        for (int cr = 0; cr < 32; cr++) {
          const int c = 31 - cr;
          v1 = lane == c ? mul(v1, d1) : v1;
          const T xc = tv_shfl(v1, c);
          const T u1 = fnma_(Ds[r1 * LDD + 32 + c], xc, v1);
          v1 = lane < c ? u1 : v1;
          v0 = fnma_(Ds[r0 * LDD + 32 + c], xc, v0);
        }
#pragma unroll
        for (int cr = 0; cr < 32; cr++) {
          const int c = 31 - cr;
          v0 = lane == c ? mul(v0, d0) : v0;
          const T xc = tv_shfl(v0, c);
          const T u0 = fnma_(Ds[r0 * LDD + c], xc, v0);
          v0 = lane < c ? u0 : v0;
        }
      }
      // publish (padding rows carry zeros) and store the solution in place
      TvPacket* dst = pk + (size_t)i * blk_words;
      tv_put(dst + (size_t)(r0 * nrhs + nn) * W, v0, p.tag);
      tv_put(dst + (size_t)(r1 * nrhs + nn) * W, v1, p.tag);
      if (r0 < rows) B[i0 + r0 + (int64_t)nn * p.ldb] = v0;
      if (r1 < rows) B[i0 + r1 + (int64_t)nn * p.ldb] = v1;
    }
    __syncthreads();
  }
}

}  // namespace

// op(A) X = alpha B for nrhs <= 8 columns; returns false when the shape is not for this path.
template <typename T>
bool trsv_small(Context& ctx, Uplo uplo, Op op, Diag diag, int64_t n, int64_t nrhs, T alpha,
                const T* A, int64_t lda, T* B, int64_t ldb) {
  if (nrhs < 1 || nrhs > TV_MAXRHS || n < 256 || !ctx.allow_side_stream) return false;
  constexpr int W = TvWords<T>::N;
  const int64_t nb = ceil_div(n, (int64_t)TV_RB);
  const int sms = ctx.sm_limit > 0 && ctx.sm_limit < ctx.sm_count ? ctx.sm_limit : ctx.sm_count;
  const int64_t G = nb < sms ? nb : sms;
  static DeviceOnce attr;
  const size_t smem = sizeof(T) * ((size_t)TV_RB * (TV_RB + 1) + 2 * TV_RB * TV_MAXRHS +
                                   4 * TV_RB * TV_MAXRHS);
  if (attr.pending(ctx.device)) {
    LA_CUDA_CHECK(cudaFuncSetAttribute(trsv_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)smem));
    attr.mark(ctx.device);
  }
  Arena::Mark mk = ctx.arena.mark();
  const size_t npk = (size_t)nb * TV_RB * TV_MAXRHS * W;
  TvPacket* pk = ctx.alloc<TvPacket>(npk);
  LA_CUDA_CHECK(cudaMemsetAsync(pk, 0, npk * sizeof(TvPacket), ctx.stream));
  TvParams p;
  p.n = n;
  p.nrhs = (int)nrhs;
  const bool trans = op != OpN;
  p.lower = ((uplo == Lower) != trans) ? 1 : 0;
  p.unit = diag == Unit ? 1 : 0;
  p.conj = (op == OpC && Traits<T>::is_complex) ? 1 : 0;
  p.ur = trans ? lda : 1;
  p.uc = trans ? 1 : lda;
  p.ldb = ldb;
  p.tag = 1;
  void* args[] = {&p, &alpha, &A, &B, &pk};
  LA_CUDA_CHECK(cudaLaunchCooperativeKernel((void*)trsv_kernel<T>, dim3((unsigned)G), dim3(TV_NT),
                                            args, smem, ctx.stream));
  ctx.arena.release(mk);
  return true;
}

template bool trsv_small<float>(Context&, Uplo, Op, Diag, int64_t, int64_t, float, const float*,
                                int64_t, float*, int64_t);
template bool trsv_small<double>(Context&, Uplo, Op, Diag, int64_t, int64_t, double, const double*,
                                 int64_t, double*, int64_t);
template bool trsv_small<zcomplex>(Context&, Uplo, Op, Diag, int64_t, int64_t, zcomplex,
                                   const zcomplex*, int64_t, zcomplex*, int64_t);

}  // namespace la
