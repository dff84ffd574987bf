// la_b200 — S = X^H Y for tall operands (X: m x p, Y: m x q, p, q <= 64, m in the hundred
// thousands): the products behind the block reflectors of a tall-skinny QR (?geqrf / ?gels,
// src/linalg/linalg.cr:493-506): Gram matrices V^H V, cross terms V1^H V2 of merged T factors
// and W = V^H C of the updates inside a block.
//
// On the general GEMM kernel these products put 8..64 useful rows into 128-row tiles; they are
// bound by the HBM stream of X and Y (p, q <= 32) or by the FP64 pipe (p = q = 64), not by
// anything a large tile helps with.  Here every CTA streams its row range of X and Y through a
// 3-stage cp.async ring in shared memory and keeps a P x Q register-blocked accumulator spread
// over its threads (thread tile TP x TQ, the rest of the threads split the rows); the per-CTA
// partial sums go to workspace and are added in a fixed order, so the result does not depend on
// scheduling.  Algorithmic bytes (p + q) * m * sizeof(T), flops 2 * p * q * m.
#include "la_internal.h"

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

// P = TP * NTP columns of X, Q = TQ * NTQ columns of Y, KG = 256 / (NTP * NTQ) row groups,
// R rows per stage (a multiple of 16 / sizeof(T)-independent: R % 16 == 0), row pitch R + 2
// elements (keeps 16-byte alignment and spreads the columns of a warp's loads over the banks).
template <typename T, int TP, int TQ, int NTP, int NTQ, int R>
__global__ void __launch_bounds__(TG_NT)
tall_gram_kernel(int64_t m, int p, int q, const T* __restrict__ X, int64_t ldx,
                 const T* __restrict__ Y, int64_t ldy, T* __restrict__ part, int64_t rows_per_cta) {
  constexpr int P = TP * NTP, Q = TQ * NTQ;
  constexpr int KG = TG_NT / (NTP * NTQ);
  constexpr int RP = R + 16 / (int)sizeof(T) * 1;          // pitch in elements (16-byte multiple)
  constexpr int EPV = 16 / (int)sizeof(T);                 // elements per 16-byte piece
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
    if (chunk < nchunks) {
      T* st = smem + (chunk % TG_STAGES) * STAGE_ELEMS;
      const int64_t r0 = r_begin + (int64_t)chunk * R;
      constexpr int PIECES_PER_COL = R / EPV;
      for (int i = tid; i < (P + Q) * PIECES_PER_COL; i += TG_NT) {
        const int col = i / PIECES_PER_COL, piece = i % PIECES_PER_COL;
        const int64_t r = r0 + (int64_t)piece * EPV;
        const bool isx = col < P;
        const int c = isx ? col : col - P;
        const bool col_ok = isx ? c < p : c < q;
        int64_t left = r_end - r;                            // valid rows of this piece
        int bytes = (!col_ok || left <= 0) ? 0 : (left >= EPV ? 16 : (int)left * (int)sizeof(T));
        const T* src = isx ? X + (int64_t)c * ldx : Y + (int64_t)c * ldy;
        cp_async16(st + col * RP + piece * EPV, bytes ? (const void*)(src + r) : (const void*)X, bytes);
      }
    }
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
#pragma unroll 2
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
  // reduction over the KG row groups, in shared memory (reusing the stages), fixed order
  T* red = smem;                                               // [KG][P * Q]
  constexpr int CAP = TG_STAGES * STAGE_ELEMS / (P * Q);       // row groups that fit at once
  static_assert(CAP >= 1, "reduction buffer");
  // (group 0 keeps its own sums in registers and adds the others to them)
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

template <typename T>
__global__ void tall_gram_reduce_kernel(int p, int q, int P, int PQ, int G, const T* __restrict__ part,
                                        T* __restrict__ S, int64_t lds) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= p * q) return;
  const int i = idx % p, j = idx / p;
  T s = zero<T>();
  for (int g = 0; g < G; g++) s = add(s, part[(int64_t)g * PQ + i + j * P]);
  S[i + (int64_t)j * lds] = s;
}

template <typename T, int TP, int TQ, int NTP, int NTQ, int R>
void launch_tall_gram(Context& ctx, int64_t m, int p, int q, const T* X, int64_t ldx, const T* Y,
                      int64_t ldy, T* S, int64_t lds) {
  constexpr int P = TP * NTP, Q = TQ * NTQ;
  constexpr int RP = R + 16 / (int)sizeof(T);
  constexpr size_t smem = (size_t)TG_STAGES * (P + Q) * RP * sizeof(T);
  auto kern = tall_gram_kernel<T, TP, TQ, NTP, NTQ, R>;
  static DeviceOnce attr;
  static int per_sm[DeviceOnce::kMax] = {};
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
  tall_gram_reduce_kernel<T><<<(unsigned)ceil_div((int64_t)p * q, 128), 128, 0, ctx.stream>>>(
      p, q, P, P * Q, (int)G, part, S, lds);
  LA_CUDA_CHECK(cudaGetLastError());
  ctx.arena.release(mk);
}

template <typename T> struct TgRows;   // rows per stage for the four tile sizes (8, 16, 32, 64)
template <> struct TgRows<double> { static constexpr int r8 = 256, r16 = 128, r32 = 64, r64 = 32; };
template <> struct TgRows<float> { static constexpr int r8 = 512, r16 = 256, r32 = 128, r64 = 64; };
template <> struct TgRows<zcomplex> { static constexpr int r8 = 128, r16 = 64, r32 = 32, r64 = 16; };

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
  const int64_t w = p > q ? p : q;
  using RR = TgRows<T>;
  if (w <= 8) launch_tall_gram<T, 2, 2, 4, 4, RR::r8>(ctx, m, (int)p, (int)q, X, ldx, Y, ldy, S, lds);
  else if (w <= 16) launch_tall_gram<T, 4, 4, 4, 4, RR::r16>(ctx, m, (int)p, (int)q, X, ldx, Y, ldy, S, lds);
  else if (w <= 32) launch_tall_gram<T, 4, 4, 8, 8, RR::r32>(ctx, m, (int)p, (int)q, X, ldx, Y, ldy, S, lds);
  else {
    // (a 64 x 64 complex accumulator does not fit the register file: the general GEMM takes it)
    if constexpr (Traits<T>::is_complex) return false;
    else launch_tall_gram<T, 8, 8, 8, 8, RR::r64>(ctx, m, (int)p, (int)q, X, ldx, Y, ldy, S, lds);
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
