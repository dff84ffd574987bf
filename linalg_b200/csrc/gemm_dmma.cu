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
//     allocation is per 4 warps) + 1 epilogue warpgroup + 8 consumer warps (2x4), warp tile 64x32 = 8x4 DMMA tiles, 64 FP64
//     accumulators per thread;
//   * both operand layouts per side (contiguous along M/N or along K), so NN/NT/TN/TT need
//     no transposition pass;
//   * k-slot permutation: the 16 k of a stage are assigned to (tig, step) as
//     k = (tig&1) + 4*(tig>>1) + {0,2,8,10}[step].  Two of a warp's four k have bit 0 set and
//     two have bit 2 set, which spreads a fragment load over all 32 banks in BOTH swizzled
//     layouts (K-major: chunk = (k>>1)^(mn&7), half = k&1; MN-major: chunk = (mn>>1)^(k&7)),
//     i.e. the minimum two wavefronts per 64-bit warp load;
//   * epilogue off the critical path: the consumers park the accumulators of a finished tile in
//     TENSOR MEMORY (tcgen05.st, 128 KB per tile, two buffers = the SM's whole 256 KB) and start
//     the next tile at once; four epilogue warps read them back (tcgen05.ld) and do the
//     alpha/beta read-modify-write of C (optional triangle mask, SYRK-style) meanwhile.  In the
//     rank-128..1024 updates of the factorisations the C round trip was 10-30% of a tile.
// Algorithmic work: 2*m*n*k flops per call; HBM traffic ~ (m*k + k*n)*8*(reuse factor) +
// m*n*8*(1 or 2) bytes, tensor-pipe bound for k >= 64.
#include "la_internal.h"
#include "tma_utils.cuh"
#include <mutex>
#include <type_traits>

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
constexpr int EPILOGUE_WARPS = 4;  // one per tensor-memory lane quarter
constexpr int NTHREADS = (CONSUMER_WARPS + PRODUCER_WARPS + EPILOGUE_WARPS) * 32;
constexpr int TMEM_COLS = 512;     // 2 accumulator buffers x 2 consumer warps per quarter x 128
constexpr int A_STAGE_BYTES = BM * BK * 8;  // 16 KB
constexpr int B_STAGE_BYTES = BN * BK * 8;  // 16 KB
constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers, tmem*/;
constexpr int GROUP_M = 16;  // tile rows per rasterisation group

struct GemmParams {
  int64_t m, n, k;
  double alpha, beta;
  double* C;
  int64_t ldc;
  int tri;
  int tiles_m, tiles_n, num_tiles;
  int ksplit;     // > 1: split-K: the tiles from `whole_tiles` on are cut into ksplit k-ranges whose
                  // raw sums go to `partial` (a dense BM x BN tile per (split, tile))
  int kchunk;     // k-blocks per split
  int whole_tiles;  // leading tiles (in enumeration order) that are not split
  double* partial;
  int zmode;      // 1: rows are interleaved (re, im) parts of a complex matrix (gemm_z.cu)
  int herm_diag;  // zmode only: store 0 in the imaginary part of the diagonal
  int full_enum;  // masked product, few tiles: enumerate all tiles and skip the masked-out ones
  int use_red;    // beta == 1: accumulate into C with red.global.add instead of load/fma/store
  // block-staircase mask (StairMask, la_internal.h); st_nb == 0: none
  int st_nb, st_rf, st_rs, st_cf, st_cs;
};

__device__ __forceinline__ void dmma884(double (&c)[2], double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c[0]), "+d"(c[1])
               : "d"(a), "d"(b));
}

// ---- tensor memory as a staging buffer for the accumulators (no tcgen05.mma involved: FP64
// has no tcgen05 kind).  A warp may touch the 32 TMEM lanes of its quarter (warp index % 4);
// lane l of the warp owns TMEM lane 32 * quarter + l, a 32x32b access moves one 32-bit column
// per register.
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
#define LA_R8(v, o) "r"(v[o]), "r"(v[o + 1]), "r"(v[o + 2]), "r"(v[o + 3]), "r"(v[o + 4]), "r"(v[o + 5]), "r"(v[o + 6]), "r"(v[o + 7])
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
      :
      : "r"(taddr), LA_R8(v, 0), LA_R8(v, 8), LA_R8(v, 16), LA_R8(v, 24)
      : "memory");
}
#undef LA_R8
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]),
        "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]),
        "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// Fragment loads are explicit ld.shared with a 32-bit address and an immediate offset: through
// a generic pointer the compiler could not tell that the staged tiles live in shared memory and
// emitted generic 64-bit loads with two integer adds each.
template <int I, int N, typename F>
__device__ __forceinline__ void static_for_impl(F&& f) {
  if constexpr (I < N) {
    f(std::integral_constant<int, I>{});
    static_for_impl<I + 1, N>(f);
  }
}
template <typename F> __device__ __forceinline__ void static_for4(F&& f) { static_for_impl<0, 4>(f); }
template <typename F> __device__ __forceinline__ void static_for8(F&& f) { static_for_impl<0, 8>(f); }

template <int IMM>
__device__ __forceinline__ double lds_f64(uint32_t addr) {
  double v;
  asm volatile("ld.shared.f64 %0, [%1+%2];" : "=d"(v) : "r"(addr), "n"(IMM));
  return v;
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

// Work unit w of a launch: whole tiles first, then (tile, k-range) pieces of the split tiles.
struct WorkUnit { int t, split, kb0, kb1; bool part; };
__device__ __forceinline__ WorkUnit decode_unit(const GemmParams& p, int w, int num_tiles, int kblocks) {
  WorkUnit u;
  if (w < p.whole_tiles) {
    u.t = w; u.split = 0; u.kb0 = 0; u.kb1 = kblocks; u.part = false;
  } else {
    const int x = w - p.whole_tiles, ns = num_tiles - p.whole_tiles;
    u.t = p.whole_tiles + x % ns;
    u.split = x / ns;
    u.kb0 = u.split * p.kchunk;
    u.kb1 = min(kblocks, u.kb0 + p.kchunk);
    u.part = p.ksplit > 1;
  }
  return u;
}
__host__ __device__ __forceinline__ int work_units(const GemmParams& p) {
  return p.whole_tiles + (p.num_tiles - p.whole_tiles) * p.ksplit;
}

// zmode products enumerate all tiles and skip those entirely outside the stored triangle of the
// COMPLEX matrix (complex row = real row >> 1).
template <int BN_>
__device__ __forceinline__ bool tile_active(const GemmParams& p, int tm, int tn) {
  if (p.st_nb) {
    // last row of the tile against its first column, both as GLOBAL block indices (the first
    // column's block is the smallest of the tile, the last row's block the largest)
    int64_t r = (int64_t)tm * BM + BM - 1;
    if (r >= p.m) r = p.m - 1;
    const int64_t rb = p.st_rf + (r / p.st_nb) * p.st_rs;
    const int64_t cb = p.st_cf + ((int64_t)tn * BN_ / p.st_nb) * p.st_cs;
    return rb >= cb;
  }
  if (p.tri == TriFull) return true;
  if (p.zmode) {
    if (p.tri == TriLower) return ((tm * BM + BM - 1) >> 1) >= tn * BN_;
    return ((tm * BM) >> 1) <= tn * BN_ + BN_ - 1;
  }
  if (!p.full_enum) return true;   // the triangular enumeration only visits active tiles
  if (p.tri == TriLower) return tm * BM + BM - 1 >= tn * BN_;
  return tm * BM <= tn * BN_ + BN_ - 1;
}

// NJ = DMMA tiles per warp along N: 4 -> CTA tile 128 x 128; 2 / 1 -> CTA tile 128 x 64 / 128 x 32
// for products with at most 64 / 32 columns (W^H = C^H V in QR, Gram matrices), which would otherwise spend 3/4 of
// the tensor work on padding.
template <bool A_MN, bool B_MN, int NJ>
__global__ void __launch_bounds__(NTHREADS, 1)
dgemm_dmma_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                  const GemmParams p) {
  constexpr int BN = 32 * NJ;
  constexpr int B_STAGE_BYTES = BN * BK * 8;
  constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = (unsigned char*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint64_t* full_bar = (uint64_t*)(smem + STAGES * STAGE_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* acc_full = empty_bar + STAGES;   // [2] accumulators of a tile are in tensor memory
  uint64_t* acc_empty = acc_full + 2;        // [2] the epilogue has read them back
  uint32_t* tmem_ptr = (uint32_t*)(acc_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; s++) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], CONSUMER_WARPS);
    }
    for (int a = 0; a < 2; a++) {
      mbar_init(&acc_full[a], CONSUMER_WARPS);
      mbar_init(&acc_empty[a], EPILOGUE_WARPS);
    }
    mbar_fence_init();
  }
  if (warp == CONSUMER_WARPS + 2) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     smem_u32(tmem_ptr)),
                 "n"(TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  const int num_tiles = p.num_tiles;
  const int kblocks = (int)((p.k + BK - 1) / BK);

  if (warp >= CONSUMER_WARPS + PRODUCER_WARPS) {
    // ===================== epilogue warps: tensor memory -> alpha/beta -> C ===============
    // They run one tile behind the consumers, so the read-modify-write of C (an HBM round trip
    // per tile that nothing else can hide in a short-k product) overlaps the next main loop.
    asm volatile("setmaxnreg.dec.sync.aligned.u32 88;");
    const int q = warp & 3;
    const int gid = lane >> 2, tig = lane & 3;
    const bool beta_zero = p.beta == 0.0, beta_one = p.beta == 1.0 && p.use_red;
    int abuf = 0;
    uint32_t aphase = 0;
    const int nunits = work_units(p);
    for (int w = blockIdx.x; w < nunits; w += gridDim.x) {
      const WorkUnit wu = decode_unit(p, w, num_tiles, kblocks);
      const int t = wu.t, split = wu.split;
      int tm, tn;
      tile_coords((p.zmode || p.full_enum) ? TriFull : p.tri, t, p.tiles_m, p.tiles_n, tm, tn);
      if (!tile_active<BN>(p, tm, tn)) continue;
      mbar_wait(&acc_full[abuf], aphase);
      tc_fence_after();
      // 16 chunks (2 consumer warps x 8 columns of C each), rolled: this code runs beside the
      // main loop and must stay small in registers and instructions.
      //  * beta == 1 (every trailing update of the factorisations): C += alpha * acc is sent
      //    to the L2 as fire-and-forget red.global.add.f64 — one rounding, the same result as
      //    the load / fma / store sequence, every element is touched once so the order does not
      //    matter, and no load latency is exposed at all;
      //  * otherwise the 8 loads of a column are issued together, then the stores.
#pragma unroll 1
      for (int n = 0; n < 4 * NJ; n++) {
        const int h = n / (2 * NJ);               // which of the quarter's two consumer warps
        const int cw = q + 4 * h;                 // the consumer warp whose fragment this is
        const int jc = n - h * (2 * NJ);          // jc = 2 * j + c
        const int64_t row_base = (int64_t)tm * BM + (cw & 1) * 64 + gid;
        const int64_t col = (int64_t)tn * BN + (cw >> 1) * (8 * NJ) + 2 * tig + (jc >> 1) * 8 + (jc & 1);
        uint32_t v[16];
        tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + abuf * 256 + h * 128 + jc * 16, v);
        if (col >= p.n) continue;
        if (wu.part) {  // raw partial sums; alpha/beta are applied by the reduction pass
          const int64_t slot = (int64_t)split * (num_tiles - p.whole_tiles) + (t - p.whole_tiles);
          double* Pcol = p.partial + (slot * BN + (col - (int64_t)tn * BN)) * BM + (cw & 1) * 64 + gid;
#pragma unroll
          for (int i = 0; i < 8; i++)
            Pcol[i * 8] = __hiloint2double((int)v[2 * i + 1], (int)v[2 * i]);
          continue;
        }
        double* Ccol = p.C + col * p.ldc;
        if (beta_one) {
#pragma unroll
          for (int i = 0; i < 8; i++) {
            const int64_t row = row_base + i * 8;
            const int64_t mrow = p.zmode ? (row >> 1) : row;
            if (row >= p.m || (p.tri == TriLower && mrow < col) || (p.tri == TriUpper && mrow > col))
              continue;
            const double r = p.alpha * __hiloint2double((int)v[2 * i + 1], (int)v[2 * i]);
            if (p.herm_diag && mrow == col && (row & 1)) { Ccol[row] = 0.0; continue; }
            asm volatile("red.global.add.f64 [%0], %1;" ::"l"(Ccol + row), "d"(r) : "memory");
          }
          continue;
        }
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
          double r = p.alpha * __hiloint2double((int)v[2 * i + 1], (int)v[2 * i]);
          if (!beta_zero) r = fma(p.beta, cin[i], r);
          if (p.herm_diag && (p.zmode ? (row >> 1) : row) == col && (row & 1)) r = 0.0;
          Ccol[row] = r;
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_empty[abuf]);
      if (++abuf == 2) { abuf = 0; aphase ^= 1; }
    }
  } else if (warp >= CONSUMER_WARPS) {
    // ===================== TMA producer (one elected lane of the producer warpgroup) =======
    // The producer warpgroup hands registers to the consumers.  Its own code must then fit in
    // what it keeps, hence the generous 56 and the rolled issue loops below.
    asm volatile("setmaxnreg.dec.sync.aligned.u32 56;");
    if (warp == CONSUMER_WARPS + 1 && lane == 0) {
      tma_prefetch_desc(&tmA);
      tma_prefetch_desc(&tmB);
      int stage = 0;
      uint32_t phase = 0;
      const int nunits = work_units(p);
      for (int w = blockIdx.x; w < nunits; w += gridDim.x) {
        const WorkUnit wu = decode_unit(p, w, num_tiles, kblocks);
        int tm, tn;
        tile_coords((p.zmode || p.full_enum) ? TriFull : p.tri, wu.t, p.tiles_m, p.tiles_n, tm, tn);
        if (!tile_active<BN>(p, tm, tn)) continue;
        const int m0 = tm * BM, n0 = tn * BN;
        const int kb0 = wu.kb0, kb1 = wu.kb1;
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
  } else {
    // ===================== DMMA consumers =====================
    asm volatile("setmaxnreg.inc.sync.aligned.u32 184;");
    const int gid = lane >> 2, tig = lane & 3;
    const int wm = warp & 1, wn = warp >> 1;     // 2 x 4 warps -> 64 x 32 warp tiles
    const int kslot = (tig & 1) + 4 * (tig >> 1);  // + kstep(ks) = {0, 2, 8, 10}
    int stage = 0;
    uint32_t phase = 0;
    int abuf = 0;
    uint32_t aphase = 0;
    const bool beta_zero = p.beta == 0.0;
    // Per-thread parts of the fragment addresses.  With mn = w*.. + 8*i + gid and
    // k = kslot + {0, 2, 8, 10}[ks], frag_off() splits into a compile-time part (i, ks) and a
    // thread part that only depends on the parities the 128-byte swizzle mixes in:
    //   MN-major: (i & 1, ks & 1)      K-major: ks
    uint32_t offA[4], offB[4];
#pragma unroll
    for (int v = 0; v < 4; v++) {
      if (A_MN) offA[v] = frag_off<true>(wm * 64 + (v >> 1) * 8 + gid, kslot + 2 * (v & 1));
      else      offA[v] = frag_off<false>(wm * 64 + gid, kslot + 2 * (v & 1) + 8 * (v >> 1));
      if (B_MN) offB[v] = frag_off<true>(wn * (8 * NJ) + (v >> 1) * 8 + gid, kslot + 2 * (v & 1));
      else      offB[v] = frag_off<false>(wn * (8 * NJ) + gid, kslot + 2 * (v & 1) + 8 * (v >> 1));
    }
    const uint32_t smem_base = smem_u32(smem);
    // this warp's slot in tensor memory: lane quarter warp % 4, columns (warp / 4) * 128
    const uint32_t tslot = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + (warp >> 2) * 128;

    const int nunits = work_units(p);
    for (int w = blockIdx.x; w < nunits; w += gridDim.x) {
      const WorkUnit wu = decode_unit(p, w, num_tiles, kblocks);
      int tm, tn;
      tile_coords((p.zmode || p.full_enum) ? TriFull : p.tri, wu.t, p.tiles_m, p.tiles_n, tm, tn);
      if (!tile_active<BN>(p, tm, tn)) continue;
      const int kb0 = wu.kb0, kb1 = wu.kb1;

      if (!beta_zero && !wu.part) {
        // pull the C tile into L2 while the main loop runs: thread t covers column t/2, four
        // 128-byte lines of its 128 rows
        const int64_t col = (int64_t)tn * BN + (threadIdx.x >> 1);
        if (col < p.n && (int)(threadIdx.x >> 1) < BN) {
          const double* cp = p.C + col * p.ldc + (int64_t)tm * BM + (threadIdx.x & 1) * 64;
#pragma unroll
          for (int qq = 0; qq < 4; qq++)
            if ((int64_t)tm * BM + (threadIdx.x & 1) * 64 + qq * 16 < p.m)
              asm volatile("prefetch.global.L2 [%0];" ::"l"(cp + qq * 16));
        }
      }
      double acc[8][NJ][2];
#pragma unroll
      for (int i = 0; i < 8; i++)
#pragma unroll
        for (int j = 0; j < NJ; j++) acc[i][j][0] = acc[i][j][1] = 0.0;

      for (int kb = kb0; kb < kb1; kb++) {
        mbar_wait(&full_bar[stage], phase);
        const uint32_t sA = smem_base + stage * STAGE_BYTES;
        const uint32_t sB = sA + A_STAGE_BYTES;
        static_for4([&](auto KS) {
          constexpr int ks = decltype(KS)::value;
          double a[8], b[NJ];
          static_for8([&](auto I) {
            constexpr int i = decltype(I)::value;
            // MN-major: k adds (2*(ks&1) + 8*(ks>>1)) * 128 bytes, the 8-row group i adds
            // (i>>1) * 2048; K-major: row group i adds i * 8 * 128 bytes
            if (A_MN) a[i] = lds_f64<(8 * (ks >> 1)) * 128 + (i >> 1) * (BK * 128)>(sA + offA[(i & 1) * 2 + (ks & 1)]);
            else      a[i] = lds_f64<i * 8 * 128>(sA + offA[ks]);
          });
          static_for_impl<0, NJ>([&](auto J) {
            constexpr int j = decltype(J)::value;
            if (B_MN) b[j] = lds_f64<(8 * (ks >> 1)) * 128 + (j >> 1) * (BK * 128)>(sB + offB[(j & 1) * 2 + (ks & 1)]);
            else      b[j] = lds_f64<j * 8 * 128>(sB + offB[ks]);
          });
#pragma unroll
          for (int i = 0; i < 8; i++)
#pragma unroll
            for (int j = 0; j < NJ; j++) dmma884(acc[i][j], a[i], b[j]);
        });
        // The stage may only be handed back once every fragment load of this warp has been
        // PERFORMED, not merely issued: ptxas schedules the last loads of the k-block right
        // before the arrive and the DMMAs that consume them after it, and the refill by the
        // TMA unit (async proxy) was observed to overtake those loads (wrong A fragments in the
        // last k-step of a stage, in the warps that share a scheduler with the producer).
        // The proxy fence orders this warp's generic-proxy reads before the async-proxy write
        // that the arrive enables.
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty_bar[stage]);
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }

      // ---- hand the accumulators to the epilogue warps through tensor memory: column
      // (2 * j + c) * 16 + 2 * i (+1) of this warp's slot holds the low (high) word of
      // acc[i][j][c], so that a 16-column read returns the 8 rows of one column of C
      mbar_wait(&acc_empty[abuf], aphase ^ 1);
      tc_fence_after();
#pragma unroll
      for (int j = 0; j < NJ; j++) {
        uint32_t v[32];
#pragma unroll
        for (int c = 0; c < 2; c++)
#pragma unroll
          for (int i = 0; i < 8; i++) {
            v[(c * 8 + i) * 2] = (uint32_t)__double2loint(acc[i][j][c]);
            v[(c * 8 + i) * 2 + 1] = (uint32_t)__double2hiint(acc[i][j][c]);
          }
        tmem_st32(tslot + abuf * 256 + j * 32, v);
      }
      asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_full[abuf]);
      if (++abuf == 2) { abuf = 0; aphase ^= 1; }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == CONSUMER_WARPS + 2) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base),
                 "n"(TMEM_COLS)
                 : "memory");
  }
}

// C = alpha * sum_s partial_s + beta * C over the split tiles (split-K reduction): one CTA per
// tile, the sums in a fixed order (results do not depend on how the pieces were scheduled).
__global__ void __launch_bounds__(256)
splitk_reduce_kernel(GemmParams p, int bn) {
  // blockIdx.x: split tile; blockIdx.y: a group of columns of the tile (two columns per pass)
  const int ns = p.num_tiles - p.whole_tiles;
  const int o = blockIdx.x;
  int tm, tn;
  tile_coords(TriFull, p.whole_tiles + o, p.tiles_m, p.tiles_n, tm, tn);
  const int i = threadIdx.x % BM;
  for (int j = blockIdx.y * 2 + threadIdx.x / BM; j < bn; j += gridDim.y * 2) {
    const int64_t row = (int64_t)tm * BM + i, col = (int64_t)tn * bn + j;
    if (row >= p.m || col >= p.n) continue;
    double sum = 0.0;
    for (int sp = 0; sp < p.ksplit; sp++)
      sum += p.partial[(((int64_t)sp * ns + o) * bn + j) * BM + i];
    double v = p.alpha * sum;
    double* c = p.C + row + col * p.ldc;
    if (p.beta != 0.0) v = fma(p.beta, *c, v);
    *c = v;
  }
}

template <bool A_MN, bool B_MN, int NJ>
void launch(Context& ctx, const CUtensorMap& tmA, const CUtensorMap& tmB, const GemmParams& p) {
  static DeviceOnce attr;
  auto kern = dgemm_dmma_kernel<A_MN, B_MN, NJ>;
  if (attr.pending(ctx.device)) {
    LA_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    attr.mark(ctx.device);
  }
  int tiles = work_units(p);
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
                      double* C, int64_t ldc, int tri, int zmode, bool herm_diag,
                      const StairMask* stair) {
  if (stair && (tri != TriFull || zmode || stair->nb % BN != 0)) return false;
  if (ta == OpC) ta = OpT;
  if (tb == OpC) tb = OpT;
  if (m <= 0 || n <= 0) return true;
  if (k <= 0 || alpha == 0.0) return false;  // pure scaling: the generic kernel handles it
  // tiny problems are latency bound either way; keep them on the simple kernel
  if (m * n * k < (int64_t)64 * 64 * 64) return false;
  if (m > (int64_t)1 << 30 || n > (int64_t)1 << 30 || k > (int64_t)1 << 30) return false;
  const bool a_mn = (ta == OpN), b_mn = (tb != OpN);
  // Long-k products of large matrices run as a sequence of k-chunks accumulating into C.  With
  // the whole k range per tile, a wave of 148 tiles streams ~400 MB of operands and the CTAs of
  // a wave drift apart in k, so the 126 MB L2 caught only ~3/4 of the operand re-reads (ncu:
  // 108 GB of DRAM traffic per N = 16384 product against 8.6 GB algorithmic).  A 4096-wide
  // chunk keeps a wave's operands (~100 MB) and the A panel of a tile-row group resident, and
  // the extra passes over C cost nothing now that the epilogue is off the critical path.
  constexpr int64_t KCHUNK = 4096;
  if (k >= 2 * KCHUNK && (double)m * (double)n >= 4096.0 * 4096.0) {
    for (int64_t l0 = 0; l0 < k; l0 += KCHUNK) {
      const int64_t kc = KCHUNK < k - l0 ? KCHUNK : k - l0;
      const double* Ac = a_mn ? A + l0 * lda : A + l0;
      const double* Bc = b_mn ? B + l0 * ldb : B + l0;
      if (!gemm_dmma_f64_ex(ctx, ta, tb, m, n, kc, alpha, Ac, lda, Bc, ldb, l0 == 0 ? beta : 1.0, C,
                            ldc, tri, zmode, herm_diag, stair)) {
        if (l0 == 0) return false;
        fprintf(stderr, "la_b200: internal error: k-chunked GEMM fell back after the first chunk\n");
        abort();
      }
    }
    return true;
  }
  // narrow CTA tile (128 x 32) for products with at most 32 columns
  // CTA tile width: 128 unless a narrower tile keeps more SMs busy (few output tiles: the single
  // SM that owns a 128 x 128 tile needs 2.1 us per 16 k whatever the rest of the machine does) or
  // the product has at most 64 columns.  Score = fraction of the SMs' time spent on useful tiles
  // x relative efficiency of the tile shape (fragment loads per DMMA).
  int bn = BN;
  if (tri == TriFull && !zmode && !stair) {
    const int sms_ = ctx.sm_limit > 0 && ctx.sm_limit < ctx.sm_count ? ctx.sm_limit : ctx.sm_count;
    const int64_t tm_ = ceil_div(m, (int64_t)BM);
    double best = -1.0;
    const int cand[3] = {128, 64, 32};
    const double eff[3] = {1.0, 0.92, 0.8};
    for (int c = 0; c < 3; c++) {
      if (cand[c] < 128 && n > 64 && k > 1024) continue;   // long k: split-K / big tiles do better
      const int64_t tiles = tm_ * ceil_div(n, (int64_t)cand[c]);
      const int64_t waves = ceil_div(tiles, (int64_t)sms_);
      const double used = (double)n / (double)(ceil_div(n, (int64_t)cand[c]) * cand[c]);  // padding
      const double score = (double)tiles / (double)(waves * sms_) * used * eff[c];
      if (score > best * 1.05) { best = score; bn = cand[c]; }
    }
  }
  // masked (SYRK-style) products with few tiles: narrow tiles too, enumerating every tile and
  // skipping those outside the stored triangle (the triangular enumeration needs square tiles)
  bool full_enum = false;
  if (tri != TriFull && !zmode) {
    const int sms_ = ctx.sm_limit > 0 && ctx.sm_limit < ctx.sm_count ? ctx.sm_limit : ctx.sm_count;
    const int64_t tm_ = ceil_div(m, (int64_t)BM);
    const int64_t t128 = tri == TriLower ? tri_tile_count((int)tm_, (int)ceil_div(n, (int64_t)BN))
                                         : tri_tile_count((int)ceil_div(n, (int64_t)BN), (int)tm_);
    if (t128 * 2 <= sms_ && k <= 1024) {
      bn = t128 * 4 <= sms_ ? 32 : 64;
      full_enum = true;
    }
  }
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
    if (!make_tensor_map_2d(&tmB, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 8, B, k, n, ldb * 8, BK, bn,
                            CU_TENSOR_MAP_SWIZZLE_128B)) return false;
  }
  GemmParams p;
  p.m = m; p.n = n; p.k = k;
  p.alpha = alpha; p.beta = beta;
  p.C = C; p.ldc = ldc; p.tri = tri;
  p.tiles_m = (int)ceil_div(m, BM);
  p.tiles_n = (int)ceil_div(n, bn);
  if ((int64_t)p.tiles_m * p.tiles_n > (int64_t)1 << 30) return false;
  p.zmode = zmode;
  p.full_enum = (full_enum || stair) ? 1 : 0;
  p.st_nb = stair ? (int)stair->nb : 0;
  p.st_rf = stair ? (int)stair->row_first : 0;
  p.st_rs = stair ? (int)stair->row_step : 0;
  p.st_cf = stair ? (int)stair->col_first : 0;
  p.st_cs = stair ? (int)stair->col_step : 0;
  p.herm_diag = herm_diag ? 1 : 0;
  static const bool no_red = getenv("LA_B200_NO_RED") != nullptr;
  p.use_red = no_red ? 0 : 1;
  if (zmode || full_enum || stair) p.num_tiles = p.tiles_m * p.tiles_n;
  else if (tri == TriLower) p.num_tiles = tri_tile_count(p.tiles_m, p.tiles_n);
  else if (tri == TriUpper) p.num_tiles = tri_tile_count(p.tiles_n, p.tiles_m);
  else p.num_tiles = p.tiles_m * p.tiles_n;
  if (p.num_tiles <= 0) return true;
  // split-K: for products with few output tiles and a long k (V^H C in QR, Gram matrices) every
  // tile is cut; for products whose tiles fill the SMs a non-integer number of times (N = 2048:
  // 256 tiles on 148 SMs) only the tiles of the last, partial wave are, so that its pieces fill
  // the machine again.  Partial sums are reduced in a fixed order by a second kernel.
  p.ksplit = 1;
  p.kchunk = (int)ceil_div(k, BK);
  p.whole_tiles = 0;
  p.partial = nullptr;
  Arena::Mark mk = ctx.arena.mark();
  {
    const int kblocks = (int)ceil_div(k, BK);
    const int sms = ctx.sm_limit > 0 && ctx.sm_limit < ctx.sm_count ? ctx.sm_limit : ctx.sm_count;
    static const bool hybrid_on = !(getenv("LA_B200_GEMM_HYBRID") && atoi(getenv("LA_B200_GEMM_HYBRID")) == 0);
    // (contexts that must not allocate — allow_side_stream == false — never split)
    const bool may_split = ctx.allow_side_stream && tri == TriFull && !stair;
    int S = 1, whole = 0;
    if (may_split && p.num_tiles * 4 <= sms && kblocks >= 128) {
      S = sms / p.num_tiles;
      if (S > kblocks / 32) S = kblocks / 32;
    } else if (may_split && !zmode && hybrid_on && kblocks >= 64 && p.num_tiles < 24 * sms) {
      // cost in k-block times of one tile; a work unit costs its k-blocks plus a fixed part
      // (pipeline fill, accumulator hand-over); units are dealt round-robin to the CTAs
      static const double ovh = getenv("LA_B200_GEMM_OVH") ? atof(getenv("LA_B200_GEMM_OVH")) : 4.0;
      static const double red_bw = getenv("LA_B200_GEMM_REDBW") ? atof(getenv("LA_B200_GEMM_REDBW")) : 4.0e6;
      auto makespan = [&](int wh, int sp) {
        const int units = wh + (p.num_tiles - wh) * sp;
        const int G = units < sms ? units : sms;
        const double piece = (double)ceil_div(kblocks, sp) + ovh;
        double worst = 0.0;
        for (int b = 0; b < G; b++) {
          double t = 0.0;
          for (int w = b; w < units; w += G) t += w < wh ? (double)kblocks + ovh : piece;
          if (t > worst) worst = t;
        }
        // the reduction pass: partial tiles read once, C read and written
        if (sp > 1) {
          const double bytes = (double)(p.num_tiles - wh) * BM * bn * 8.0 * (sp + 2);
          worst += (8.0 + bytes / red_bw) / (2.18 * bn / 128.0);   // us -> k-block times
        }
        return worst;
      };
      const double base = makespan(p.num_tiles, 1);
      double best = base * 0.95;
      const int wh_full = (p.num_tiles / sms) * sms;
      const int cand_s[5] = {2, 3, 4, 6, 8};
      for (int ci = 0; ci < 5; ci++) {
        if (kblocks / cand_s[ci] < 16) break;
        for (int pass = 0; pass < 2; pass++) {
          const int wh = pass == 0 ? wh_full : 0;
          if (pass == 1 && wh_full == 0) break;
          if (wh >= p.num_tiles) continue;
          const double c = makespan(wh, cand_s[ci]);
          if (c < best) { best = c; S = cand_s[ci]; whole = wh; }
        }
      }
    }
    if (S >= 2) {
      p.kchunk = (int)ceil_div(kblocks, S);
      p.ksplit = (int)ceil_div(kblocks, p.kchunk);
      p.whole_tiles = whole;
      p.partial = ctx.alloc<double>((size_t)p.ksplit * (size_t)(p.num_tiles - whole) * BM * bn);
    }
  }
  if (bn == 32) {
    if (a_mn && b_mn) launch<true, true, 1>(ctx, tmA, tmB, p);
    else if (a_mn && !b_mn) launch<true, false, 1>(ctx, tmA, tmB, p);
    else if (!a_mn && b_mn) launch<false, true, 1>(ctx, tmA, tmB, p);
    else launch<false, false, 1>(ctx, tmA, tmB, p);
  } else if (bn == 64) {
    if (a_mn && b_mn) launch<true, true, 2>(ctx, tmA, tmB, p);
    else if (a_mn && !b_mn) launch<true, false, 2>(ctx, tmA, tmB, p);
    else if (!a_mn && b_mn) launch<false, true, 2>(ctx, tmA, tmB, p);
    else launch<false, false, 2>(ctx, tmA, tmB, p);
  } else {
    if (a_mn && b_mn) launch<true, true, 4>(ctx, tmA, tmB, p);
    else if (a_mn && !b_mn) launch<true, false, 4>(ctx, tmA, tmB, p);
    else if (!a_mn && b_mn) launch<false, true, 4>(ctx, tmA, tmB, p);
    else launch<false, false, 4>(ctx, tmA, tmB, p);
  }
  if (p.ksplit > 1) {
    const int ns = p.num_tiles - p.whole_tiles;
    // enough CTAs to fill the machine whatever the number of split tiles
    int gy = (int)ceil_div((int64_t)8 * ctx.sm_count, (int64_t)ns);
    if (gy > bn / 2) gy = bn / 2;
    if (gy < 1) gy = 1;
    splitk_reduce_kernel<<<dim3((unsigned)ns, (unsigned)gy), 256, 0, ctx.stream>>>(p, bn);
    LA_CUDA_CHECK(cudaGetLastError());
  }
  ctx.arena.release(mk);
  return true;
}

}  // namespace la
