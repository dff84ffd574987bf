// la_b200 — FP64 GEMM on the tensor cores (DMMA) with TMA-staged shared-memory tiles.
//
// Hot path of cblas_dgemm (src/linalg/mult.cr:64-69 -> libCBLAS.cr:175) and of every
// trailing update inside dgetrf / dpotrf / dgeqrf / dtrsm.
//
// sm_100a has no tcgen05 kind for f64; the FP64 tensor path is mma.sync.m8n8k4.f64
// (SASS DMMA.8x8x4, 64 FMA/clk/SM = 37.2 TFLOP/s at 1965 MHz, measured 37.0 by
// tools/microbench/peak_fp64.cu; the larger PTX shapes only expand to the same instruction).
// Design:
//   * persistent CTAs (one per SM), grouped tile order for L2 reuse;
//   * CTA tile 128x128x16, 4-stage ring of TMA loads (cp.async.bulk.tensor.2d, 128-byte
//     swizzle, zero fill at ragged edges) signalled through mbarriers;
//   * 1 producer warpgroup (one elected lane issues TMA; a whole warpgroup because register
//     allocation is per 4 warps: 384 threads x 168 registers fill the SM) + 8 consumer warps (2x4), warp tile 64x32 = 8x4 DMMA tiles, 64 FP64
//     accumulators per thread;
//   * both operand layouts per side (contiguous along M/N or along K), so NN/NT/TN/TT need
//     no transposition pass;
//   * k-slot permutation: the 16 k of a stage are assigned to (tig, step) as
//     k = (tig&1) + 4*(tig>>1) + {0,2,8,10}[step].  Two of a warp's four k have bit 0 set and
//     two have bit 2 set, which spreads a fragment load over all 32 banks in BOTH swizzled
//     layouts (K-major: chunk = (k>>1)^(mn&7), half = k&1; MN-major: chunk = (mn>>1)^(k&7)),
//     i.e. the minimum two wavefronts per 64-bit warp load;
//   * epilogue straight from registers: alpha/beta, optional triangle mask (SYRK-style).
// Algorithmic work: 2*m*n*k flops per call; HBM traffic ~ (m*k + k*n)*8*(reuse factor) +
// m*n*8*(1 or 2) bytes, tensor-pipe bound for k >= 64.
#include "la_internal.h"
#include "tma_utils.cuh"
#include <mutex>

namespace la {

// ------------------------------------------------------------------ tensor-map helper
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                    const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                    const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode_fn() {
  static PFN_encodeTiled fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres);
    if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !p) {
      fprintf(stderr, "la_b200: cuTensorMapEncodeTiled is not available from the driver\n");
      abort();
    }
    fn = (PFN_encodeTiled)p;
  });
  return fn;
}

bool make_tensor_map_2d(CUtensorMap* map, CUtensorMapDataType dtype, size_t elem_bytes,
                        const void* base, uint64_t inner_extent, uint64_t outer_extent,
                        uint64_t outer_stride_bytes, uint32_t box_inner, uint32_t box_outer,
                        CUtensorMapSwizzle swizzle) {
  if (((uintptr_t)base & 15) || (outer_stride_bytes & 15)) return false;
  if (inner_extent == 0 || outer_extent == 0) return false;
  cuuint64_t dims[2] = {inner_extent, outer_extent};
  cuuint64_t strides[1] = {outer_stride_bytes};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = get_encode_fn()(map, dtype, 2, const_cast<void*>(base), dims, strides, box, estr,
                               CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle,
                               CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS;
}

namespace {

constexpr int BM = 128, BN = 128, BK = 16, STAGES = 4;
constexpr int CONSUMER_WARPS = 8;
constexpr int PRODUCER_WARPS = 4;  // warps are allocated in groups of four
constexpr int NTHREADS = (CONSUMER_WARPS + PRODUCER_WARPS) * 32;
constexpr int A_STAGE_BYTES = BM * BK * 8;  // 16 KB
constexpr int B_STAGE_BYTES = BN * BK * 8;  // 16 KB
constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;
constexpr int GROUP_M = 16;  // tile rows per rasterisation group

struct GemmParams {
  int64_t m, n, k;
  double alpha, beta;
  double* C;
  int64_t ldc;
  int tri;
  int tiles_m, tiles_n, num_tiles;
  int ksplit;     // > 1: split-K, raw partial sums go to `partial` (one m x n slab per split)
  int kchunk;     // k-blocks per split
  double* partial;
  int64_t ld_part;
  int zmode;      // 1: rows are interleaved (re, im) parts of a complex matrix (gemm_z.cu)
  int herm_diag;  // zmode only: store 0 in the imaginary part of the diagonal
};

__device__ __forceinline__ void dmma884(double (&c)[2], double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c[0]), "+d"(c[1])
               : "d"(a), "d"(b));
}

// byte offset of element (mn, k) inside one operand stage
template <bool MN_MAJOR>
__device__ __forceinline__ uint32_t frag_off(int mn, int k) {
  if (MN_MAJOR)  // [mn/16][k][16 mn], 128-byte rows, 128B swizzle
    return (uint32_t)((mn >> 4) * (BK * 128) + k * 128 + ((((mn & 15) >> 1) ^ (k & 7)) << 4) +
                      ((mn & 1) << 3));
  else           // [mn][16 k]
    return (uint32_t)(mn * 128 + ((((k >> 1) ^ (mn & 7))) << 4) + ((k & 1) << 3));
}

// Tile t -> (tm, tn).  Full products walk groups of GROUP_M tile rows (L2 reuse of the A rows of
// a group and of the B columns it sweeps).  Masked (SYRK-style) products enumerate only the
// tiles that intersect the stored triangle, in the same banded order, so that the persistent
// CTAs stay balanced.  Lower: valid tiles have tm >= tn (BM == BN).
__host__ __device__ inline int tri_band_count(int g0, int gsz, int TN, int& cnt_full) {
  const int cols_full = (g0 + 1 < TN) ? g0 + 1 : TN;
  cnt_full = cols_full * gsz;
  int cnt_part = 0;
  const int tend = (g0 + gsz < TN) ? g0 + gsz : TN;
  for (int tn = g0 + 1; tn < tend; tn++) cnt_part += g0 + gsz - tn;
  return cnt_full + cnt_part;
}
__host__ __device__ inline int tri_tile_count(int TM, int TN) {
  int total = 0;
  for (int g0 = 0; g0 < TM; g0 += GROUP_M) {
    int cf;
    total += tri_band_count(g0, (TM - g0 < GROUP_M) ? TM - g0 : GROUP_M, TN, cf);
  }
  return total;
}
__device__ __forceinline__ void tri_tile_coords(int t, int TM, int TN, int& tm, int& tn) {
  int g0 = 0, gsz, cnt_full;
  while (true) {
    gsz = min(GROUP_M, TM - g0);
    const int cnt = tri_band_count(g0, gsz, TN, cnt_full);
    if (t < cnt) break;
    t -= cnt;
    g0 += gsz;
  }
  if (t < cnt_full) {
    tn = t / gsz;
    tm = g0 + t % gsz;
  } else {
    t -= cnt_full;
    for (tn = g0 + 1;; tn++) {
      const int c = g0 + gsz - tn;
      if (t < c) { tm = tn + t; break; }
      t -= c;
    }
  }
}

__device__ __forceinline__ void tile_coords(int tri, int t, int tiles_m, int tiles_n, int& tm,
                                            int& tn) {
  if (tri == TriLower) { tri_tile_coords(t, tiles_m, tiles_n, tm, tn); return; }
  if (tri == TriUpper) { tri_tile_coords(t, tiles_n, tiles_m, tn, tm); return; }
  const int per_group = GROUP_M * tiles_n;
  const int g = t / per_group;
  const int first_m = g * GROUP_M;
  const int gsz = min(tiles_m - first_m, GROUP_M);
  const int r = t - g * per_group;
  tm = first_m + r % gsz;
  tn = r / gsz;
}

// zmode products enumerate all tiles and skip those entirely outside the stored triangle of the
// COMPLEX matrix (complex row = real row >> 1).
__device__ __forceinline__ bool tile_active(const GemmParams& p, int tm, int tn) {
  if (!p.zmode || p.tri == TriFull) return true;
  if (p.tri == TriLower) return ((tm * BM + BM - 1) >> 1) >= tn * BN;
  return ((tm * BM) >> 1) <= tn * BN + BN - 1;
}

template <bool A_MN, bool B_MN>
__global__ void __launch_bounds__(NTHREADS, 1)
dgemm_dmma_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                  const GemmParams p) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = (unsigned char*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint64_t* full_bar = (uint64_t*)(smem + STAGES * STAGE_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; s++) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], CONSUMER_WARPS);
    }
    mbar_fence_init();
  }
  __syncthreads();

  const int num_tiles = p.num_tiles;
  const int kblocks = (int)((p.k + BK - 1) / BK);

  if (warp >= CONSUMER_WARPS) {
    // ===================== TMA producer (one elected lane of the producer warpgroup) =======
    // The producer warpgroup hands registers to the consumers.  Its own code must then fit in
    // what it keeps (a larger footprint silently corrupts other warps' registers), hence the
    // generous 56 and the rolled issue loops below.
    asm volatile("setmaxnreg.dec.sync.aligned.u32 56;");
    if (warp == CONSUMER_WARPS && lane == 0) {
      tma_prefetch_desc(&tmA);
      tma_prefetch_desc(&tmB);
      int stage = 0;
      uint32_t phase = 0;
      for (int w = blockIdx.x; w < num_tiles * p.ksplit; w += gridDim.x) {
        const int t = w % num_tiles, split = w / num_tiles;
        int tm, tn;
        tile_coords(p.zmode ? TriFull : p.tri, t, p.tiles_m, p.tiles_n, tm, tn);
        if (!tile_active(p, tm, tn)) continue;
        const int m0 = tm * BM, n0 = tn * BN;
        const int kb0 = split * p.kchunk, kb1 = min(kblocks, kb0 + p.kchunk);
        for (int kb = kb0; kb < kb1; kb++) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          mbar_expect_tx(&full_bar[stage], STAGE_BYTES);
          unsigned char* sA = smem + stage * STAGE_BYTES;
          unsigned char* sB = sA + A_STAGE_BYTES;
          const int k0 = kb * BK;
          if (A_MN) {
#pragma unroll 1
            for (int j = 0; j < BM / 16; j++)
              tma_load_2d(sA + j * (BK * 128), &tmA, &full_bar[stage], m0 + 16 * j, k0);
          } else {
            tma_load_2d(sA, &tmA, &full_bar[stage], k0, m0);
          }
          if (B_MN) {
#pragma unroll 1
            for (int j = 0; j < BN / 16; j++)
              tma_load_2d(sB + j * (BK * 128), &tmB, &full_bar[stage], n0 + 16 * j, k0);
          } else {
            tma_load_2d(sB, &tmB, &full_bar[stage], k0, n0);
          }
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
    return;
  }

  // ===================== DMMA consumers =====================
  asm volatile("setmaxnreg.inc.sync.aligned.u32 224;");
  const int gid = lane >> 2, tig = lane & 3;
  const int wm = warp & 1, wn = warp >> 1;     // 2 x 4 warps -> 64 x 32 warp tiles
  const int kslot = (tig & 1) + 4 * (tig >> 1);  // + kstep(ks) = {0, 2, 8, 10}
  int stage = 0;
  uint32_t phase = 0;
  const bool beta_zero = p.beta == 0.0;

  for (int w = blockIdx.x; w < num_tiles * p.ksplit; w += gridDim.x) {
    const int t = w % num_tiles, split = w / num_tiles;
    int tm, tn;
    tile_coords(p.zmode ? TriFull : p.tri, t, p.tiles_m, p.tiles_n, tm, tn);
    if (!tile_active(p, tm, tn)) continue;
    const int kb0 = split * p.kchunk, kb1 = min(kblocks, kb0 + p.kchunk);

    if (!beta_zero && p.ksplit == 1) {
      // pull the C tile into L2 while the main loop runs: thread t covers column t/2, four
      // 128-byte lines of its 128 rows
      const int64_t col = (int64_t)tn * BN + (threadIdx.x >> 1);
      if (col < p.n) {
        const double* cp = p.C + col * p.ldc + (int64_t)tm * BM + (threadIdx.x & 1) * 64;
#pragma unroll
        for (int q = 0; q < 4; q++)
          if ((int64_t)tm * BM + (threadIdx.x & 1) * 64 + q * 16 < p.m)
            asm volatile("prefetch.global.L2 [%0];" ::"l"(cp + q * 16));
      }
    }
    double acc[8][4][2];
#pragma unroll
    for (int i = 0; i < 8; i++)
#pragma unroll
      for (int j = 0; j < 4; j++) acc[i][j][0] = acc[i][j][1] = 0.0;

    for (int kb = kb0; kb < kb1; kb++) {
      mbar_wait(&full_bar[stage], phase);
      const unsigned char* sA = smem + stage * STAGE_BYTES;
      const unsigned char* sB = sA + A_STAGE_BYTES;
#pragma unroll
      for (int ks = 0; ks < 4; ks++) {
        const int kk = kslot + 2 * (ks & 1) + 8 * (ks >> 1);
        double a[8], b[4];
#pragma unroll
        for (int i = 0; i < 8; i++)
          a[i] = *reinterpret_cast<const double*>(sA + frag_off<A_MN>(wm * 64 + i * 8 + gid, kk));
#pragma unroll
        for (int j = 0; j < 4; j++)
          b[j] = *reinterpret_cast<const double*>(sB + frag_off<B_MN>(wn * 32 + j * 8 + gid, kk));
#pragma unroll
        for (int i = 0; i < 8; i++)
#pragma unroll
          for (int j = 0; j < 4; j++) dmma884(acc[i][j], a[i], b[j]);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty_bar[stage]);
      if (++stage == STAGES) { stage = 0; phase ^= 1; }
    }

    // ---- epilogue: C(row, col..col+1) for row = gid, col = 2*tig of each 8x8 tile
    const int64_t row_base = (int64_t)tm * BM + wm * 64 + gid;
    const int64_t col_base = (int64_t)tn * BN + wn * 32 + 2 * tig;
#pragma unroll
    for (int j = 0; j < 4; j++) {
#pragma unroll
      for (int c = 0; c < 2; c++) {
        const int64_t col = col_base + j * 8 + c;
        if (col >= p.n) continue;
        if (p.ksplit > 1) {  // raw partial sums; alpha/beta are applied by the reduction pass
          double* Pcol = p.partial + ((int64_t)split * p.n + col) * p.ld_part;
#pragma unroll
          for (int i = 0; i < 8; i++) {
            const int64_t row = row_base + i * 8;
            if (row < p.m) Pcol[row] = acc[i][j][c];
          }
          continue;
        }
        double* Ccol = p.C + col * p.ldc;
        // all loads of a column first (independent, in flight together), then the stores: a
        // load-store chain per element would expose one memory round trip per row
        bool act[8];
        double cin[8];
#pragma unroll
        for (int i = 0; i < 8; i++) {
          const int64_t row = row_base + i * 8;
          const int64_t mrow = p.zmode ? (row >> 1) : row;
          act[i] = row < p.m && !(p.tri == TriLower && mrow < col) &&
                   !(p.tri == TriUpper && mrow > col);
          cin[i] = (act[i] && !beta_zero) ? __ldcg(Ccol + row) : 0.0;
        }
#pragma unroll
        for (int i = 0; i < 8; i++) {
          if (!act[i]) continue;
          const int64_t row = row_base + i * 8;
          double v = p.alpha * acc[i][j][c];
          if (!beta_zero) v = fma(p.beta, cin[i], v);
          if (p.herm_diag && (p.zmode ? (row >> 1) : row) == col && (row & 1)) v = 0.0;
          Ccol[row] = v;
        }
      }
    }
  }
}

// C = alpha * sum_s partial_s + beta * C   (split-K reduction)
__global__ void splitk_reduce_kernel(int64_t m, int64_t n, int ksplit, const double* __restrict__ part,
                                     int64_t ld_part, double alpha, double beta,
                                     double* __restrict__ C, int64_t ldc) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  for (int64_t j = blockIdx.y; j < n; j += gridDim.y) {
    double s = 0.0;
    for (int sp = 0; sp < ksplit; sp++) s += part[((int64_t)sp * n + j) * ld_part + i];
    double v = alpha * s;
    if (beta != 0.0) v = fma(beta, C[i + j * ldc], v);
    C[i + j * ldc] = v;
  }
}

template <bool A_MN, bool B_MN>
void launch(Context& ctx, const CUtensorMap& tmA, const CUtensorMap& tmB, const GemmParams& p) {
  static bool attr = false;
  auto kern = dgemm_dmma_kernel<A_MN, B_MN>;
  if (!attr) {
    LA_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    attr = true;
  }
  int tiles = p.num_tiles * p.ksplit;
  const int sms = ctx.sm_limit > 0 && ctx.sm_limit < ctx.sm_count ? ctx.sm_limit : ctx.sm_count;
  int grid = tiles < sms ? tiles : sms;
  kern<<<grid, NTHREADS, SMEM_BYTES, ctx.stream>>>(tmA, tmB, p);
  LA_CUDA_CHECK(cudaGetLastError());
}

}  // namespace

bool gemm_dmma_f64(Context& ctx, Op ta, Op tb, int64_t m, int64_t n, int64_t k, double alpha,
                   const double* A, int64_t lda, const double* B, int64_t ldb, double beta,
                   double* C, int64_t ldc, int tri) {
  return gemm_dmma_f64_ex(ctx, ta, tb, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, tri, 0, false);
}

bool gemm_dmma_f64_ex(Context& ctx, Op ta, Op tb, int64_t m, int64_t n, int64_t k, double alpha,
                      const double* A, int64_t lda, const double* B, int64_t ldb, double beta,
                      double* C, int64_t ldc, int tri, int zmode, bool herm_diag) {
  if (ta == OpC) ta = OpT;
  if (tb == OpC) tb = OpT;
  if (m <= 0 || n <= 0) return true;
  if (k <= 0 || alpha == 0.0) return false;  // pure scaling: the generic kernel handles it
  // tiny problems are latency bound either way; keep them on the simple kernel
  if (m * n * k < (int64_t)64 * 64 * 64) return false;
  if (m > (int64_t)1 << 30 || n > (int64_t)1 << 30 || k > (int64_t)1 << 30) return false;
  const bool a_mn = (ta == OpN), b_mn = (tb != OpN);
  CUtensorMap tmA, tmB;
  // operand A as op(A) (m x k): stored m x k (lda) when ta == N, else k x m
  if (a_mn) {
    if (!make_tensor_map_2d(&tmA, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 8, A, m, k, lda * 8, 16, BK,
                            CU_TENSOR_MAP_SWIZZLE_128B)) return false;
  } else {
    if (!make_tensor_map_2d(&tmA, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 8, A, k, m, lda * 8, BK, BM,
                            CU_TENSOR_MAP_SWIZZLE_128B)) return false;
  }
  // operand B as op(B) (k x n): stored k x n (ldb) when tb == N, else n x k
  if (b_mn) {
    if (!make_tensor_map_2d(&tmB, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 8, B, n, k, ldb * 8, 16, BK,
                            CU_TENSOR_MAP_SWIZZLE_128B)) return false;
  } else {
    if (!make_tensor_map_2d(&tmB, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 8, B, k, n, ldb * 8, BK, BN,
                            CU_TENSOR_MAP_SWIZZLE_128B)) return false;
  }
  GemmParams p;
  p.m = m; p.n = n; p.k = k;
  p.alpha = alpha; p.beta = beta;
  p.C = C; p.ldc = ldc; p.tri = tri;
  p.tiles_m = (int)ceil_div(m, BM);
  p.tiles_n = (int)ceil_div(n, BN);
  if ((int64_t)p.tiles_m * p.tiles_n > (int64_t)1 << 30) return false;
  p.zmode = zmode;
  p.herm_diag = herm_diag ? 1 : 0;
  if (zmode) p.num_tiles = p.tiles_m * p.tiles_n;
  else if (tri == TriLower) p.num_tiles = tri_tile_count(p.tiles_m, p.tiles_n);
  else if (tri == TriUpper) p.num_tiles = tri_tile_count(p.tiles_n, p.tiles_m);
  else p.num_tiles = p.tiles_m * p.tiles_n;
  if (p.num_tiles <= 0) return true;
  // split-K for products with few output tiles and a long k (V^H C in QR, Gram matrices)
  p.ksplit = 1;
  p.kchunk = (int)ceil_div(k, BK);
  p.partial = nullptr;
  p.ld_part = 0;
  Arena::Mark mk = ctx.arena.mark();
  {
    const int kblocks = (int)ceil_div(k, BK);
    const int sms = ctx.sm_limit > 0 && ctx.sm_limit < ctx.sm_count ? ctx.sm_limit : ctx.sm_count;
    // (contexts that must not allocate — allow_side_stream == false — never split)
    if (ctx.allow_side_stream && tri == TriFull && p.num_tiles * 4 <= sms && kblocks >= 128) {
      int S = sms / p.num_tiles;
      if (S > kblocks / 32) S = kblocks / 32;
      if (S >= 2) {
        p.kchunk = (int)ceil_div(kblocks, S);
        p.ksplit = (int)ceil_div(kblocks, p.kchunk);
        p.ld_part = round_up(m, 2);
        p.partial = ctx.alloc<double>((size_t)p.ksplit * (size_t)n * (size_t)p.ld_part);
      }
    }
  }
  if (a_mn && b_mn) launch<true, true>(ctx, tmA, tmB, p);
  else if (a_mn && !b_mn) launch<true, false>(ctx, tmA, tmB, p);
  else if (!a_mn && b_mn) launch<false, true>(ctx, tmA, tmB, p);
  else launch<false, false>(ctx, tmA, tmB, p);
  if (p.ksplit > 1) {
    dim3 g((unsigned)ceil_div(m, 128), (unsigned)(n < 1024 ? n : 1024));
    splitk_reduce_kernel<<<g, 128, 0, ctx.stream>>>(m, n, p.ksplit, p.partial, p.ld_part, alpha,
                                                    beta, C, ldc);
    LA_CUDA_CHECK(cudaGetLastError());
  }
  ctx.arena.release(mk);
  return true;
}

}  // namespace la
