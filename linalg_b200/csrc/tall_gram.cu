// la_b200 — S = X^H Y for tall operands (X: m x p, Y: m x q, p, q <= 64, m in the hundred
// thousands): the products behind the block reflectors of a tall-skinny QR (?geqrf / ?gels,
// src/linalg/linalg.cr:493-506): Gram matrices V^H V, cross terms V1^H V2 of merged T factors
// and W = V^H C of the updates inside a block.
//
// On the general GEMM kernel these products put 8..64 useful rows into 128-row tiles; they are
// bound by the HBM stream of X and Y (p, q <= 32) or by the FP64 pipe (p = q = 64), not by
// anything a large tile helps with.  Here every CTA streams its row range of X and Y through a
// 3-stage cp.async ring in shared memory and keeps a P x Q accumulator spread over its threads:
//   * generic kernel (any element type): thread tile TP x TQ of FMAs, the threads that do not fit
//     the P x Q tile split the rows;
//   * FP64, both sides >= 32 columns: mma.sync.m8n8k4 (DMMA) fragments straight from the staged
//     chunks, a warp owns a (8 TPW) x (8 TQW) block of S.
// The per-CTA partial sums go to workspace and are added in a fixed order, so the result does not
// depend on scheduling.  Algorithmic bytes (p + q) * m * sizeof(T), flops 2 * p * q * m.
#include "la_internal.h"
#include <type_traits>

namespace la {

namespace {

constexpr int TG_NT = 256;
constexpr int TG_STAGES = 3;

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc, int src_bytes) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(d), "l"(gsrc), "r"(src_bytes)
               : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// Stage `chunk` of the CTA's rows [r_begin, r_end): P columns of X then Q columns of Y, R rows
// each at a pitch of RP elements; rows / columns beyond the operands are zero-filled.
template <typename T, int P, int Q, int R, int RP>
__device__ __forceinline__ void stage_chunk(T* st, int chunk, int64_t r_begin, int64_t r_end, int p,
                                            int q, const T* X, int64_t ldx, const T* Y, int64_t ldy) {
  constexpr int EPV = 16 / (int)sizeof(T);                   // elements per 16-byte piece
  constexpr int PIECES_PER_COL = R / EPV;
  const int64_t r0 = r_begin + (int64_t)chunk * R;
  for (int i = threadIdx.x; i < (P + Q) * PIECES_PER_COL; i += TG_NT) {
    const int col = i / PIECES_PER_COL, piece = i % PIECES_PER_COL;
    const int64_t r = r0 + (int64_t)piece * EPV;
    const bool isx = col < P;
    const int c = isx ? col : col - P;
    const bool col_ok = isx ? c < p : c < q;
    const int64_t left = r_end - r;                          // valid rows of this piece
    const int bytes = (!col_ok || left <= 0) ? 0 : (left >= EPV ? 16 : (int)left * (int)sizeof(T));
    const T* src = isx ? X + (int64_t)c * ldx : Y + (int64_t)c * ldy;
    cp_async16(st + col * RP + piece * EPV, bytes ? (const void*)(src + r) : (const void*)X, bytes);
  }
}

// P = TP * NTP columns of X, Q = TQ * NTQ columns of Y, KG = 256 / (NTP * NTQ) row groups, R rows
// per stage (R % 16 == 0), row pitch R + 16 bytes (keeps 16-byte alignment and spreads the
// columns of a warp's loads over the banks).
template <typename T, int TP, int TQ, int NTP, int NTQ, int R>
__global__ void __launch_bounds__(TG_NT)
tall_gram_kernel(int64_t m, int p, int q, const T* __restrict__ X, int64_t ldx,
                 const T* __restrict__ Y, int64_t ldy, T* __restrict__ part, int64_t rows_per_cta) {
  constexpr int P = TP * NTP, Q = TQ * NTQ;
  constexpr int KG = TG_NT / (NTP * NTQ);
  constexpr int RP = R + 16 / (int)sizeof(T);
  constexpr int STAGE_ELEMS = (P + Q) * RP;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T* smem = reinterpret_cast<T*>(smem_raw);

  const int tid = threadIdx.x;
  const int tp = tid % NTP, tq = (tid / NTP) % NTQ, kg = tid / (NTP * NTQ);
  const int64_t r_begin = (int64_t)blockIdx.x * rows_per_cta;
  int64_t r_end = r_begin + rows_per_cta;
  if (r_end > m) r_end = m;
  const int64_t nrows = r_end > r_begin ? r_end - r_begin : 0;
  const int nchunks = (int)((nrows + R - 1) / R);

  auto issue = [&](int chunk) {
    if (chunk < nchunks)
      stage_chunk<T, P, Q, R, RP>(smem + (chunk % TG_STAGES) * STAGE_ELEMS, chunk, r_begin, r_end, p,
                                  q, X, ldx, Y, ldy);
    cp_async_commit();
  };

  T acc[TP][TQ];
#pragma unroll
  for (int a = 0; a < TP; a++)
#pragma unroll
    for (int b = 0; b < TQ; b++) acc[a][b] = zero<T>();

  issue(0);
  issue(1);
  for (int ch = 0; ch < nchunks; ch++) {
    issue(ch + 2);
    cp_async_wait<2>();
    __syncthreads();
    const T* xs = smem + (ch % TG_STAGES) * STAGE_ELEMS;
    const T* ys = xs + P * RP;
#pragma unroll 4
    for (int kk = kg; kk < R; kk += KG) {
      T xv[TP], yv[TQ];
#pragma unroll
      for (int a = 0; a < TP; a++) xv[a] = conj_(xs[(tp + NTP * a) * RP + kk]);
#pragma unroll
      for (int b = 0; b < TQ; b++) yv[b] = ys[(tq + NTQ * b) * RP + kk];
#pragma unroll
      for (int a = 0; a < TP; a++)
#pragma unroll
        for (int b = 0; b < TQ; b++) acc[a][b] = fma_(xv[a], yv[b], acc[a][b]);
    }
    __syncthreads();
  }
  cp_async_wait<0>();
  __syncthreads();
  // reduction over the KG row groups in shared memory (reusing the stages), fixed order: group 0
  // keeps its own sums in registers and adds the others to them
  T* red = smem;                                               // [<= CAP][P * Q]
  constexpr int CAP = TG_STAGES * STAGE_ELEMS / (P * Q);       // row groups that fit at once
  static_assert(CAP >= 1, "reduction buffer");
  for (int g0 = 1; g0 < KG; g0 += CAP) {
    if (kg >= g0 && kg < g0 + CAP) {
#pragma unroll
      for (int a = 0; a < TP; a++)
#pragma unroll
        for (int b = 0; b < TQ; b++)
          red[(kg - g0) * P * Q + (tp + NTP * a) + (tq + NTQ * b) * P] = acc[a][b];
    }
    __syncthreads();
    if (kg == 0) {
      const int gn = KG - g0 < CAP ? KG - g0 : CAP;
      for (int g = 0; g < gn; g++)
#pragma unroll
        for (int a = 0; a < TP; a++)
#pragma unroll
          for (int b = 0; b < TQ; b++)
            acc[a][b] = add(acc[a][b], red[g * P * Q + (tp + NTP * a) + (tq + NTQ * b) * P]);
    }
    __syncthreads();
  }
  if (kg == 0) {
    T* out = part + (int64_t)blockIdx.x * P * Q;
#pragma unroll
    for (int a = 0; a < TP; a++)
#pragma unroll
      for (int b = 0; b < TQ; b++) out[(tp + NTP * a) + (tq + NTQ * b) * P] = acc[a][b];
  }
}

// FP64 on the tensor cores: A = X^T (8 columns of X x 4 rows), B = Y (4 rows x 8 columns of Y).
// Fragments of m8n8k4: a = A[lane / 4][lane % 4], b = B[lane % 4][lane / 4],
// c0, c1 = C[lane / 4][2 * (lane % 4) + {0, 1}].
template <int P, int Q, int TPW, int TQW, int R>
__global__ void __launch_bounds__(TG_NT)
tall_gram_dmma_kernel(int64_t m, int p, int q, const double* __restrict__ X, int64_t ldx,
                      const double* __restrict__ Y, int64_t ldy, double* __restrict__ part,
                      int64_t rows_per_cta) {
  constexpr int WPG = (P / (8 * TPW)) * (Q / (8 * TQW));   // warps per row group
  constexpr int KG = (TG_NT / 32) / WPG;
  static_assert(KG >= 1 && KG * WPG == TG_NT / 32, "warp layout");
  constexpr int RP = R + 4;                                 // pitch = 8 words mod 32
  constexpr int STAGE_ELEMS = (P + Q) * RP;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double* smem = reinterpret_cast<double*>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int gid = lane >> 2, tig = lane & 3;
  const int kg = warp / WPG, wi = warp % WPG;
  const int wp = wi % (P / (8 * TPW)), wq = wi / (P / (8 * TPW));
  const int64_t r_begin = (int64_t)blockIdx.x * rows_per_cta;
  int64_t r_end = r_begin + rows_per_cta;
  if (r_end > m) r_end = m;
  const int64_t nrows = r_end > r_begin ? r_end - r_begin : 0;
  const int nchunks = (int)((nrows + R - 1) / R);

  auto issue = [&](int chunk) {
    if (chunk < nchunks)
      stage_chunk<double, P, Q, R, RP>(smem + (chunk % TG_STAGES) * STAGE_ELEMS, chunk, r_begin, r_end,
                                       p, q, X, ldx, Y, ldy);
    cp_async_commit();
  };

  double acc[TPW][TQW][2];
#pragma unroll
  for (int a = 0; a < TPW; a++)
#pragma unroll
    for (int b = 0; b < TQW; b++) acc[a][b][0] = acc[a][b][1] = 0.0;

  issue(0);
  issue(1);
  for (int ch = 0; ch < nchunks; ch++) {
    issue(ch + 2);
    cp_async_wait<2>();
    __syncthreads();
    const double* st = smem + (ch % TG_STAGES) * STAGE_ELEMS;
    const double* xs = st + (wp * 8 * TPW + gid) * RP + tig;
    const double* ys = st + (P + wq * 8 * TQW + gid) * RP + tig;
#pragma unroll 2
    for (int kk = 4 * kg; kk < R; kk += 4 * KG) {
      double xa[TPW], yb[TQW];
#pragma unroll
      for (int a = 0; a < TPW; a++) xa[a] = xs[a * 8 * RP + kk];
#pragma unroll
      for (int b = 0; b < TQW; b++) yb[b] = ys[b * 8 * RP + kk];
#pragma unroll
      for (int a = 0; a < TPW; a++)
#pragma unroll
        for (int b = 0; b < TQW; b++)
          asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                       : "+d"(acc[a][b][0]), "+d"(acc[a][b][1])
                       : "d"(xa[a]), "d"(yb[b]));
    }
    __syncthreads();
  }
  cp_async_wait<0>();
  __syncthreads();
  double* red = smem;
  for (int g = 1; g < KG; g++) {
    if (kg == g) {
#pragma unroll
      for (int a = 0; a < TPW; a++)
#pragma unroll
        for (int b = 0; b < TQW; b++)
#pragma unroll
          for (int e = 0; e < 2; e++)
            red[(wp * 8 * TPW + a * 8 + gid) + (wq * 8 * TQW + b * 8 + 2 * tig + e) * P] = acc[a][b][e];
    }
    __syncthreads();
    if (kg == 0) {
#pragma unroll
      for (int a = 0; a < TPW; a++)
#pragma unroll
        for (int b = 0; b < TQW; b++)
#pragma unroll
          for (int e = 0; e < 2; e++)
            acc[a][b][e] += red[(wp * 8 * TPW + a * 8 + gid) + (wq * 8 * TQW + b * 8 + 2 * tig + e) * P];
    }
    __syncthreads();
  }
  if (kg == 0) {
    double* out = part + (int64_t)blockIdx.x * P * Q;
#pragma unroll
    for (int a = 0; a < TPW; a++)
#pragma unroll
      for (int b = 0; b < TQW; b++)
#pragma unroll
        for (int e = 0; e < 2; e++)
          out[(wp * 8 * TPW + a * 8 + gid) + (wq * 8 * TQW + b * 8 + 2 * tig + e) * P] = acc[a][b][e];
  }
}

// one warp per output element: lane l adds the partial sums l, l + 32, ... (ascending), then the
// 32 lane sums are added in a fixed butterfly order
template <typename T>
__global__ void __launch_bounds__(256)
tall_gram_reduce_kernel(int p, int q, int P, int PQ, int G, const T* __restrict__ part,
                        T* __restrict__ S, int64_t lds) {
  const int idx = blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (idx >= p * q) return;
  const int i = idx % p, j = idx / p;
  T s = zero<T>();
  for (int g = lane; g < G; g += 32) s = add(s, part[(int64_t)g * PQ + i + j * P]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    T t;
    if constexpr (Traits<T>::is_complex) {
      t.x = __shfl_xor_sync(0xffffffffu, s.x, o);
      t.y = __shfl_xor_sync(0xffffffffu, s.y, o);
    } else {
      t = __shfl_xor_sync(0xffffffffu, s, o);
    }
    s = add(s, t);
  }
  if (lane == 0) S[i + (int64_t)j * lds] = s;
}

// common launch logic: a grid of co-resident CTAs over the rows, then the fixed-order reduction
template <typename T, typename K>
void launch_common(Context& ctx, K kern, DeviceOnce& attr, int* per_sm, size_t smem, int P, int Q,
                   int R, int64_t m, int p, int q, const T* X, int64_t ldx, const T* Y, int64_t ldy,
                   T* S, int64_t lds) {
  const int dev = ctx.device >= 0 && ctx.device < DeviceOnce::kMax ? ctx.device : 0;
  if (attr.pending(ctx.device) || per_sm[dev] == 0) {
    LA_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int nb = 1;
    LA_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kern, TG_NT, smem));
    per_sm[dev] = nb < 1 ? 1 : nb;
    attr.mark(ctx.device);
  }
  const int sms = ctx.sm_limit > 0 && ctx.sm_limit < ctx.sm_count ? ctx.sm_limit : ctx.sm_count;
  int64_t G = (int64_t)sms * per_sm[dev];
  const int64_t max_by_rows = ceil_div(m, (int64_t)4 * R);
  if (G > max_by_rows) G = max_by_rows;
  if (G < 1) G = 1;
  const int64_t rows_per_cta = round_up(ceil_div(m, G), (int64_t)R);
  G = ceil_div(m, rows_per_cta);
  Arena::Mark mk = ctx.arena.mark();
  T* part = ctx.alloc<T>((size_t)G * P * Q);
  kern<<<(unsigned)G, TG_NT, smem, ctx.stream>>>(m, p, q, X, ldx, Y, ldy, part, rows_per_cta);
  tall_gram_reduce_kernel<T><<<(unsigned)ceil_div((int64_t)p * q, 8), 256, 0, ctx.stream>>>(
      p, q, P, P * Q, (int)G, part, S, lds);
  LA_CUDA_CHECK(cudaGetLastError());
  ctx.arena.release(mk);
}

// rows per stage: about 32 KB per stage, a power of two, at least 16
constexpr int tg_rows(int pq, int elem_bytes) {
  int r = 16;
  while (2 * r * pq * elem_bytes <= 36864 && r < 512) r *= 2;
  return r;
}

// thread tile (T) and thread-grid extent (NT) of the generic kernel for a tile width
template <int W> struct TgSide;
template <> struct TgSide<8> { static constexpr int T = 2, NT = 4; };
template <> struct TgSide<16> { static constexpr int T = 4, NT = 4; };
template <> struct TgSide<32> { static constexpr int T = 4, NT = 8; };
template <> struct TgSide<64> { static constexpr int T = 8, NT = 8; };

template <typename T, int P, int Q>
void launch_tall_gram(Context& ctx, int64_t m, int p, int q, const T* X, int64_t ldx, const T* Y,
                      int64_t ldy, T* S, int64_t lds) {
  static DeviceOnce attr;
  static int per_sm[DeviceOnce::kMax] = {};
  if constexpr (std::is_same<T, double>::value && P >= 32 && Q >= 32) {
    constexpr int R = tg_rows(P + Q, 8);
    constexpr size_t smem = (size_t)TG_STAGES * (P + Q) * (R + 4) * sizeof(double);
    launch_common<double>(ctx, tall_gram_dmma_kernel<P, Q, 2, (Q == 64 ? 4 : 2), R>, attr, per_sm,
                          smem, P, Q, R, m, p, q, X, ldx, Y, ldy, S, lds);
  } else {
    constexpr int R = tg_rows(P + Q, (int)sizeof(T));
    constexpr size_t smem = (size_t)TG_STAGES * (P + Q) * (R + 16 / (int)sizeof(T)) * sizeof(T);
    launch_common<T>(ctx,
                     tall_gram_kernel<T, TgSide<P>::T, TgSide<Q>::T, TgSide<P>::NT, TgSide<Q>::NT, R>,
                     attr, per_sm, smem, P, Q, R, m, p, q, X, ldx, Y, ldy, S, lds);
  }
}

template <typename T, int P>
bool dispatch_q(Context& ctx, int64_t m, int p, int q, const T* X, int64_t ldx, const T* Y,
                int64_t ldy, T* S, int64_t lds) {
  if (q <= 8) launch_tall_gram<T, P, 8>(ctx, m, p, q, X, ldx, Y, ldy, S, lds);
  else if (q <= 16) launch_tall_gram<T, P, 16>(ctx, m, p, q, X, ldx, Y, ldy, S, lds);
  else if (q <= 32) launch_tall_gram<T, P, 32>(ctx, m, p, q, X, ldx, Y, ldy, S, lds);
  else {
    // (a complex accumulator beyond 32 x 32 per CTA does not fit the register file)
    if constexpr (Traits<T>::is_complex && P > 16) return false;
    else launch_tall_gram<T, P, 64>(ctx, m, p, q, X, ldx, Y, ldy, S, lds);
  }
  return true;
}

// C (m x nc) -= V (m x h) * Z^H with Z: nc x h, h and nc <= 64, m tall: the rank-h updates
// inside a QR block.  One matrix row per thread (coalesced column accesses), V's row in
// registers, Z^H in shared memory (broadcast reads), four columns of C in flight per thread.
// Algorithmic bytes (h + 2 nc) * m * sizeof(T): a streaming kernel.
template <typename T, int H>
__global__ void __launch_bounds__(256)
tall_update_kernel(int64_t m, int h, int nc, const T* __restrict__ V, int64_t ldv,
                   const T* __restrict__ Z, int64_t ldz, T* __restrict__ C, int64_t ldc) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T* zs = reinterpret_cast<T*>(smem_raw);                      // zs[c * H + j] = conj(Z(c, j))
  for (int i = threadIdx.x; i < nc * H; i += 256) {
    const int c = i / H, j = i % H;
    zs[i] = j < h ? conj_(Z[c + (int64_t)j * ldz]) : zero<T>();
  }
  __syncthreads();
  for (int64_t row = (int64_t)blockIdx.x * 256 + threadIdx.x; row < m; row += (int64_t)gridDim.x * 256) {
    T v[H];
#pragma unroll
    for (int j = 0; j < H; j++) v[j] = j < h ? V[row + (int64_t)j * ldv] : zero<T>();
    for (int c0 = 0; c0 < nc; c0 += 4) {
      T cv[4], acc[4];
#pragma unroll
      for (int u = 0; u < 4; u++) {
        cv[u] = c0 + u < nc ? C[row + (int64_t)(c0 + u) * ldc] : zero<T>();
        acc[u] = zero<T>();
      }
      // (16 bytes of Z per shared-memory load: with one load per multiply-add the kernel is bound
      // by the load pipe, at a third of the memory bandwidth for 32 reflectors)
      constexpr int EPV = 16 / (int)sizeof(T);
      struct alignas(16) ZV { T e[EPV]; };
#pragma unroll
      for (int j = 0; j < H; j += EPV) {
#pragma unroll
        for (int u = 0; u < 4; u++)
          if (c0 + u < nc) {
            const ZV z = *reinterpret_cast<const ZV*>(&zs[(c0 + u) * H + j]);
#pragma unroll
            for (int i = 0; i < EPV; i++) acc[u] = fma_(v[j + i], z.e[i], acc[u]);
          }
      }
#pragma unroll
      for (int u = 0; u < 4; u++)
        if (c0 + u < nc) C[row + (int64_t)(c0 + u) * ldc] = sub(cv[u], acc[u]);
    }
  }
}

template <typename T, int H>
void launch_tall_update(Context& ctx, int64_t m, int h, int nc, const T* V, int64_t ldv, const T* Z,
                        int64_t ldz, T* C, int64_t ldc) {
  const size_t smem = (size_t)nc * H * sizeof(T);
  static DeviceOnce attr;
  if (attr.pending(ctx.device)) {
    LA_CUDA_CHECK(cudaFuncSetAttribute(tall_update_kernel<T, H>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)(64 * H * sizeof(T))));
    attr.mark(ctx.device);
  }
  const int sms = ctx.sm_limit > 0 && ctx.sm_limit < ctx.sm_count ? ctx.sm_limit : ctx.sm_count;
  int64_t G = ceil_div(m, (int64_t)256);
  if (G > (int64_t)sms * 4) G = (int64_t)sms * 4;
  tall_update_kernel<T, H><<<(unsigned)G, 256, smem, ctx.stream>>>(m, h, nc, V, ldv, Z, ldz, C, ldc);
  LA_CUDA_CHECK(cudaGetLastError());
}

}  // namespace

// S (p x q, leading dimension lds) = X^H Y.  Returns false (nothing enqueued) when the shape or
// the alignment is not for this kernel; the caller then uses the general GEMM.
template <typename T>
bool tall_gram(Context& ctx, int64_t p, int64_t q, int64_t m, const T* X, int64_t ldx, const T* Y,
               int64_t ldy, T* S, int64_t lds) {
  if (p < 1 || q < 1 || p > 64 || q > 64 || m < 8192) return false;
  if (!ctx.allow_side_stream) return false;          // needs workspace for the partial sums
  constexpr int EPV = 16 / (int)sizeof(T);
  if (((uintptr_t)X & 15) || ((uintptr_t)Y & 15) || ldx % EPV || ldy % EPV) return false;
  if (p <= 8) return dispatch_q<T, 8>(ctx, m, (int)p, (int)q, X, ldx, Y, ldy, S, lds);
  if (p <= 16) return dispatch_q<T, 16>(ctx, m, (int)p, (int)q, X, ldx, Y, ldy, S, lds);
  if (p <= 32) return dispatch_q<T, 32>(ctx, m, (int)p, (int)q, X, ldx, Y, ldy, S, lds);
  if constexpr (Traits<T>::is_complex) {
    if (q > 16) return false;
  }
  return dispatch_q<T, 64>(ctx, m, (int)p, (int)q, X, ldx, Y, ldy, S, lds);
}

// C -= V Z^H (see tall_update_kernel).  Returns false when the shape is not for this kernel.
template <typename T>
bool tall_update(Context& ctx, int64_t m, int64_t h, int64_t nc, const T* V, int64_t ldv, const T* Z,
                 int64_t ldz, T* C, int64_t ldc) {
  if (h < 1 || nc < 1 || nc > 64 || m < 8192 || h > 64) return false;
  // V's row lives in registers: at most 32 reflectors per pass (a 64-wide pass was measured
  // three times slower than two 32-wide ones)
  for (int64_t h0 = 0; h0 < h; h0 += 32) {
    const int hh = (int)((h - h0) < 32 ? (h - h0) : 32);
    const T* Vp = V + h0 * ldv;
    const T* Zp = Z + h0 * ldz;
    if (hh <= 8) launch_tall_update<T, 8>(ctx, m, hh, (int)nc, Vp, ldv, Zp, ldz, C, ldc);
    else if (hh <= 16) launch_tall_update<T, 16>(ctx, m, hh, (int)nc, Vp, ldv, Zp, ldz, C, ldc);
    else launch_tall_update<T, 32>(ctx, m, hh, (int)nc, Vp, ldv, Zp, ldz, C, ldc);
  }
  return true;
}

template bool tall_gram<float>(Context&, int64_t, int64_t, int64_t, const float*, int64_t,
                               const float*, int64_t, float*, int64_t);
template bool tall_gram<double>(Context&, int64_t, int64_t, int64_t, const double*, int64_t,
                                const double*, int64_t, double*, int64_t);
template bool tall_gram<zcomplex>(Context&, int64_t, int64_t, int64_t, const zcomplex*, int64_t,
                                  const zcomplex*, int64_t, zcomplex*, int64_t);

}  // namespace la

namespace la {
template bool tall_update<float>(Context&, int64_t, int64_t, int64_t, const float*, int64_t,
                                 const float*, int64_t, float*, int64_t);
template bool tall_update<double>(Context&, int64_t, int64_t, int64_t, const double*, int64_t,
                                  const double*, int64_t, double*, int64_t);
template bool tall_update<zcomplex>(Context&, int64_t, int64_t, int64_t, const zcomplex*, int64_t,
                                    const zcomplex*, int64_t, zcomplex*, int64_t);
}  // namespace la
