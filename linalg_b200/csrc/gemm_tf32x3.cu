// la_b200 — FP32 GEMM on the 5th-generation tensor cores with FP32-level accuracy (3xTF32).
//
// cblas_sgemm (src/linalg/mult.cr:64-69 -> libCBLAS.cr:174) and every FP32 trailing update.
// tcgen05.mma.kind::tf32 keeps 10 mantissa bits of its inputs; each operand is therefore split
// once into hi = tf32(x) and lo = tf32(x - hi) and the product is accumulated in FP32 (TMEM) as
//        A*B ~= Ahi*Bhi + Ahi*Blo + Alo*Bhi            (the lo*lo term is below FP32 rounding).
// Design:
//   * a packing pass writes the four K-major ("k contiguous") operands Ahi, Alo (m x k) and
//     Bhi, Blo (n x k) into workspace, absorbing op(A)/op(B); the main kernel is a pure
//     K-major x K-major product;
//   * persistent kernel, one CTA per SM, tile 128 x 256 (128 / 64 columns for products with few
//     tiles), a ring of four 48 KB stages of 16 floats of k (64-byte rows, 64-byte swizzle) fed by
//     TMA -> six tcgen05.mma (M=128, N=256, K=8) per stage issued by one thread -> accumulator in
//     TMEM, double buffered (2 x 256 columns) so that the epilogue of one chunk overlaps the
//     MMAs of the next;
//   * the tensor core truncates when it accumulates: k is cut into chunks (256 terms) whose sums
//     are added in FP32 in the epilogue warps' registers, C is touched once per tile;
//   * warp roles: 0 = TMA producer, 1 = MMA issuer, 2 = TMEM allocator, 4..11 = epilogue (two
//     warps per TMEM lane quarter; tcgen05.ld 32 lanes x 32 columns, alpha/beta, coalesced
//     column-major loads/stores issued 32 at a time);
//   * optional (LA_B200_SGEMM_CLUSTER): clusters of two CTAs sharing their B tile by TMA
//     multicast.
// Algorithmic work 2*m*n*k flops; the tensor pipe executes 3x that in TF32.
#include "la_internal.h"
#include "tma_utils.cuh"

namespace la {

namespace {

constexpr int SM_ = 128, SN_ = 256, SK_ = 32;         // CTA tile (K in floats = 128 bytes)
constexpr int S_STAGES = 2;
constexpr int S_A_BYTES = SM_ * SK_ * 4;               // 16 KB per (hi or lo)
constexpr int S_B_BYTES = SN_ * SK_ * 4;               // 32 KB per (hi or lo)
constexpr int S_STAGE_BYTES = 2 * S_A_BYTES + 2 * S_B_BYTES;  // 96 KB
constexpr int S_THREADS = 384;                         // 4 control warps + 8 epilogue warps
constexpr int S_TMEM_COLS = 512;                       // two 256-column accumulators
// The tensor core truncates when it adds into its FP32 accumulator; on same-signed data that is a
// bias of ~1e-8 per accumulated term.  k is therefore processed in chunks of S_KCHUNK: each chunk
// is accumulated in TMEM from zero and added to C in FP32 (round to nearest) by the epilogue,
// which the double-buffered accumulator overlaps with the next chunk's MMAs.

struct SParams {
  int64_t m, n, k;
  float alpha, beta;
  float* C;
  int64_t ldc;
  int tri;
  int tiles_m, tiles_n;
  int kchunk_terms;   // terms accumulated in tensor memory before the FP32 add in registers
};

__device__ __forceinline__ uint32_t to_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return r;
}

// hi/lo split of op(X) into K-major storage: out(r, c) at out[r * ldo + c], r = row of op(X)
// for A (m rows) or column of op(B) for B (n rows), c = k index.  `kcontig` says whether the
// stored matrix already has k contiguous for a fixed r (no transposition needed).
__global__ void split_kernel(int kcontig, int64_t rows, int64_t k, float scale,
                             const float* __restrict__ X, int64_t ldx, float* __restrict__ hi,
                             float* __restrict__ lo, int64_t ldo) {
  __shared__ float tile[32][33];
  const int64_t nbr = ceil_div(rows, 32), nbk = ceil_div(k, 32);
  for (int64_t blk = blockIdx.x; blk < nbr * nbk; blk += gridDim.x) {
    const int64_t r0 = (blk % nbr) * 32, c0 = (blk / nbr) * 32;
    if (kcontig) {  // element (r, c) stored at X[c + r * ldx]
      for (int rr = threadIdx.y; rr < 32; rr += blockDim.y) {
        const int64_t r = r0 + rr, c = c0 + threadIdx.x;
        if (r < rows && c < k) tile[rr][threadIdx.x] = X[c + r * ldx];
      }
    } else {        // element (r, c) stored at X[r + c * ldx]
      for (int cc = threadIdx.y; cc < 32; cc += blockDim.y) {
        const int64_t r = r0 + threadIdx.x, c = c0 + cc;
        if (r < rows && c < k) tile[threadIdx.x][cc] = X[r + c * ldx];
      }
    }
    __syncthreads();
    for (int rr = threadIdx.y; rr < 32; rr += blockDim.y) {
      const int64_t r = r0 + rr, c = c0 + threadIdx.x;
      if (r < rows && c < k) {
        const float x = tile[rr][threadIdx.x] * scale;
        const float h = __uint_as_float(to_tf32(x));
        const float l = __uint_as_float(to_tf32(x - h));
        hi[r * ldo + c] = h;
        lo[r * ldo + c] = l;
      }
    }
    __syncthreads();
  }
}

// The same with 64 x 64 tiles and 16-byte accesses on both sides (operands whose base and leading
// dimension are multiples of four floats): the packing pass is pure memory traffic (2 reads +
// 4 writes of the operands), 9 % of an N = 16384 product with the 4-byte form above.
__device__ __forceinline__ void split4(const float4 x, float4& h, float4& l) {
  h.x = __uint_as_float(to_tf32(x.x)); l.x = __uint_as_float(to_tf32(x.x - h.x));
  h.y = __uint_as_float(to_tf32(x.y)); l.y = __uint_as_float(to_tf32(x.y - h.y));
  h.z = __uint_as_float(to_tf32(x.z)); l.z = __uint_as_float(to_tf32(x.z - h.z));
  h.w = __uint_as_float(to_tf32(x.w)); l.w = __uint_as_float(to_tf32(x.w - h.w));
}
__global__ void __launch_bounds__(256)
split_vec_kernel(int kcontig, int64_t rows, int64_t k, const float* __restrict__ X, int64_t ldx,
                 float* __restrict__ hi, float* __restrict__ lo, int64_t ldo) {
  __shared__ float tile[64][65];
  const int t = threadIdx.x, q4 = (t & 15) * 4, sub = t >> 4;
  const int64_t nbr = ceil_div(rows, 64), nbk = ceil_div(k, 64);
  for (int64_t blk = blockIdx.x; blk < nbr * nbk; blk += gridDim.x) {
    const int64_t r0 = (blk % nbr) * 64, c0 = (blk / nbr) * 64;
    if (kcontig) {  // element (r, c) stored at X[c + r * ldx]: straight through
#pragma unroll
      for (int p = 0; p < 4; p++) {
        const int64_t r = r0 + p * 16 + sub, c = c0 + q4;
        if (r >= rows || c >= k) continue;
        if (c + 3 < k) {
          float4 h, l;
          split4(*reinterpret_cast<const float4*>(&X[c + r * ldx]), h, l);
          *reinterpret_cast<float4*>(&hi[r * ldo + c]) = h;
          *reinterpret_cast<float4*>(&lo[r * ldo + c]) = l;
        } else {
          for (int64_t cc = c; cc < k; cc++) {
            const float x = X[cc + r * ldx];
            const float h = __uint_as_float(to_tf32(x));
            hi[r * ldo + cc] = h;
            lo[r * ldo + cc] = __uint_as_float(to_tf32(x - h));
          }
        }
      }
    } else {        // element (r, c) stored at X[r + c * ldx]: transposed through shared memory
#pragma unroll
      for (int p = 0; p < 4; p++) {
        const int cc = p * 16 + sub;
        const int64_t c = c0 + cc, r = r0 + q4;
        float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
        if (c < k && r < rows) {
          if (r + 3 < rows) {
            x = *reinterpret_cast<const float4*>(&X[r + c * ldx]);
          } else {
            x.x = X[r + c * ldx];
            if (r + 1 < rows) x.y = X[r + 1 + c * ldx];
            if (r + 2 < rows) x.z = X[r + 2 + c * ldx];
          }
        }
        tile[cc][q4] = x.x; tile[cc][q4 + 1] = x.y; tile[cc][q4 + 2] = x.z; tile[cc][q4 + 3] = x.w;
      }
      __syncthreads();
#pragma unroll
      for (int p = 0; p < 4; p++) {
        const int rr = p * 16 + sub;
        const int64_t r = r0 + rr, c = c0 + q4;
        if (r >= rows || c >= k) continue;
        const float4 x = make_float4(tile[q4][rr], tile[q4 + 1][rr], tile[q4 + 2][rr], tile[q4 + 3][rr]);
        float4 h, l;
        split4(x, h, l);
        if (c + 3 < k) {
          *reinterpret_cast<float4*>(&hi[r * ldo + c]) = h;
          *reinterpret_cast<float4*>(&lo[r * ldo + c]) = l;
        } else {
          const float hv[4] = {h.x, h.y, h.z, h.w}, lv[4] = {l.x, l.y, l.z, l.w};
          for (int i = 0; c + i < k; i++) { hi[r * ldo + c + i] = hv[i]; lo[r * ldo + c + i] = lv[i]; }
        }
      }
      __syncthreads();
    }
  }
}

// hi/lo packing of one operand (see split_kernel)
static void launch_split(Context& ctx, bool kcontig, int64_t rows, int64_t k, const float* X,
                         int64_t ldx, float* hi, float* lo, int64_t ldo) {
  static const bool old_form = getenv("LA_B200_SGEMM_PACK4B") != nullptr;
  if (!old_form && ((uintptr_t)X & 15) == 0 && ldx % 4 == 0 && ldo % 4 == 0) {
    const int64_t nblk = ceil_div(rows, 64) * ceil_div(k, 64);
    const int64_t cap = (int64_t)ctx.sm_count * 16;
    split_vec_kernel<<<(unsigned)(nblk < cap ? nblk : cap), 256, 0, ctx.stream>>>(
        kcontig ? 1 : 0, rows, k, X, ldx, hi, lo, ldo);
  } else {
    const int64_t nblk = ceil_div(rows, 32) * ceil_div(k, 32);
    const unsigned g = (unsigned)(nblk < 65535 * 8 ? nblk : 65535 * 8);
    split_kernel<<<g, dim3(32, 8), 0, ctx.stream>>>(kcontig ? 1 : 0, rows, k, 1.0f, X, ldx, hi, lo, ldo);
  }
}

// ---- tcgen05 helpers -------------------------------------------------------------------------
template <int SWZ_BYTES>
__device__ __forceinline__ uint64_t make_kmajor_desc(uint32_t smem_addr) {
  // start address (>>4), LBO = 1 (unused for swizzled K-major), SBO = 8 rows of SWZ_BYTES
  // between 8-row groups, version = 1 (sm_100), layout = SWIZZLE_128B (2) or SWIZZLE_64B (4)
  static_assert(SWZ_BYTES == 128 || SWZ_BYTES == 64, "swizzle span");
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3fff);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)((8 * SWZ_BYTES) >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)(SWZ_BYTES == 128 ? 2 : 4) << 61;
  return d;
}
constexpr uint32_t make_idesc_tf32(int M, int N) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void umma_tf32(uint32_t tmem_c, uint64_t adesc, uint64_t bdesc,
                                          uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(tmem_c),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                   smem_u32(bar))
               : "memory");
}
// the same arrive, delivered to the barrier at this offset in every CTA of `mask` (cluster)
__device__ __forceinline__ void umma_commit_multicast(uint64_t* bar, uint16_t mask) {
  asm volatile(
      "tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
          smem_u32(bar)),
      "h"(mask)
      : "memory");
}
// TMA tile load whose bytes (and complete_tx) go to the same shared-memory offsets of every CTA
// in `mask`.  The tensor map covers THIS GPU's memory only (workspace of the calling context).
__device__ __forceinline__ void tma_load_2d_multicast(void* smem_dst, const CUtensorMap* map,
                                                      uint64_t* bar, int c0, int c1, uint16_t mask) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster "
      "[%0], [%1, {%3, %4}], [%2], %5;" ::"r"(smem_u32(smem_dst)),
      "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "h"(mask)
      : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]),
        "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]),
        "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]),
        "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
        "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

template <int SN>
__device__ __forceinline__ bool s_tile_skipped(int tri, int tm, int tn) {
  if (tri == TriLower) return tm * SM_ + SM_ - 1 < tn * SN;
  if (tri == TriUpper) return tm * SM_ > tn * SN + SN - 1;
  return false;
}
// stages of the TMA ring that fit the shared memory for a tile width
__host__ __device__ constexpr int s_stage_bytes(int sn, int skt) { return 2 * (SM_ + sn) * skt * 4; }
__host__ __device__ constexpr int s_stages(int sn, int skt) {
  return (200 * 1024) / s_stage_bytes(sn, skt) > 6 ? 6 : (200 * 1024) / s_stage_bytes(sn, skt);
}
__host__ __device__ constexpr int s_smem_bytes(int sn, int skt) {
  return s_stages(sn, skt) * s_stage_bytes(sn, skt) + 1024 + 256;
}

__device__ __forceinline__ void s_tile_coords(int t, int tiles_m, int tiles_n, int& tm, int& tn) {
  constexpr int GROUP = 8;
  const int per_group = GROUP * tiles_n;
  const int g = t / per_group;
  const int first_m = g * GROUP;
  const int gsz = min(tiles_m - first_m, GROUP);
  const int r = t - g * per_group;
  tm = first_m + r % gsz;
  tn = r / gsz;
}

// CL = true: clusters of two CTAs that own vertically adjacent tiles (same tile column).  Both
// need the same B tile: each loads half of it and multicasts it to the pair, which cuts the
// L2 -> shared-memory traffic per tile from 96 to 64 KB per stage (the four operand streams of
// the split product are what bounds the single-CTA kernel at large N).  A stage is recycled when
// BOTH tensor cores have consumed it (the release is multicast to the pair as well).
// SN = tile width (256, 128 or 64 columns; narrower tiles for products with few tiles: more CTAs
// busy, and their smaller stages allow a deeper TMA ring).
// SKT = k extent of a stage in floats: 32 (128-byte rows, 128B swizzle) or 16 (64-byte rows, 64B
// swizzle).  The ring holds as many stages as fit the shared memory: what the tensor core can be
// fed is (bytes in flight) / (TMA latency), and with two 96 KB stages only one is ever in flight.
template <bool CL, int SN, int SKT>
__global__ void __launch_bounds__(S_THREADS, 1)
sgemm_tf32x3_kernel(const __grid_constant__ CUtensorMap tmAhi, const __grid_constant__ CUtensorMap tmAlo,
                    const __grid_constant__ CUtensorMap tmBhi, const __grid_constant__ CUtensorMap tmBlo,
                    const SParams p) {
  constexpr int SA = SM_ * SKT * 4;                         // bytes of one A tile (hi or lo)
  constexpr int SB = SN * SKT * 4;                          // bytes of one B tile (hi or lo)
  constexpr int STAGE = 2 * SA + 2 * SB;
  constexpr int NST = s_stages(SN, SKT);
  constexpr int SWZ = SKT * 4;                              // swizzle span = row length in bytes
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = (unsigned char*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint64_t* full_bar = (uint64_t*)(smem + NST * STAGE);
  uint64_t* empty_bar = full_bar + NST;
  uint64_t* tfull_bar = empty_bar + NST;   // accumulator ready for the epilogue
  uint64_t* tempty_bar = tfull_bar + 2;         // accumulator drained by the epilogue
  uint32_t* tmem_ptr = (uint32_t*)(tempty_bar + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int crank = CL ? (int)(blockIdx.x & 1) : 0;        // rank in the pair
  if (threadIdx.x == 0) {
    for (int s = 0; s < NST; s++) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], CL ? 2 : 1); }
    for (int a = 0; a < 2; a++) { mbar_init(&tfull_bar[a], 1); mbar_init(&tempty_bar[a], 8); }
    mbar_fence_init();
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     smem_u32(tmem_ptr)),
                 "n"(S_TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  if (CL) cluster_sync_all();          // the peer's barriers exist before anything targets them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  // work items: tiles, or (CL) pairs of vertically adjacent tiles, one pair per cluster
  const int rows_w = CL ? (p.tiles_m + 1) / 2 : p.tiles_m;
  const int num_tiles = rows_w * p.tiles_n;
  const int w_first = CL ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int w_step = CL ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  const int kblocks = (int)((p.k + SKT - 1) / SKT);
  const int kchunk = p.kchunk_terms / SKT;                 // k-blocks per accumulation chunk

  // warps 0-3 (producer, MMA issuer, TMEM allocator) hand registers to the two epilogue
  // warpgroups, which keep a 128-float running sum per thread
  if (warp == 0) {
    // ===================== TMA producer =====================
    asm volatile("setmaxnreg.dec.sync.aligned.u32 56;");
    if (lane == 0) {
      tma_prefetch_desc(&tmAhi); tma_prefetch_desc(&tmAlo);
      tma_prefetch_desc(&tmBhi); tma_prefetch_desc(&tmBlo);
      int stage = 0;
      uint32_t phase = 0;
      for (int t = w_first; t < num_tiles; t += w_step) {
        int tm, tn;
        s_tile_coords(t, rows_w, p.tiles_n, tm, tn);
        if (CL) tm = 2 * tm + crank;
        if (s_tile_skipped<SN>(p.tri, tm, tn)) continue;
        const int m0 = tm * SM_, n0 = tn * SN;
        for (int kb = 0; kb < kblocks; kb++) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          mbar_expect_tx(&full_bar[stage], STAGE);
          unsigned char* s = smem + stage * STAGE;
          const int k0 = kb * SKT;
          tma_load_2d(s, &tmAhi, &full_bar[stage], k0, m0);
          tma_load_2d(s + SA, &tmAlo, &full_bar[stage], k0, m0);
          if (CL) {   // my half of the B tile, delivered to both CTAs of the pair
            const int half = crank * (SB / 2), nh = n0 + crank * (SN / 2);
            tma_load_2d_multicast(s + 2 * SA + half, &tmBhi, &full_bar[stage], k0, nh, 3);
            tma_load_2d_multicast(s + 2 * SA + SB + half, &tmBlo, &full_bar[stage], k0, nh, 3);
          } else {
            tma_load_2d(s + 2 * SA, &tmBhi, &full_bar[stage], k0, n0);
            tma_load_2d(s + 2 * SA + SB, &tmBlo, &full_bar[stage], k0, n0);
          }
          if (++stage == NST) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (one thread) =====================
    asm volatile("setmaxnreg.dec.sync.aligned.u32 56;");
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_tf32(SM_, SN);
      int stage = 0, as = 0;
      uint32_t phase = 0, aphase = 0;
      for (int t = w_first; t < num_tiles; t += w_step) {
        int tm, tn;
        s_tile_coords(t, rows_w, p.tiles_n, tm, tn);
        if (CL) tm = 2 * tm + crank;
        if (s_tile_skipped<SN>(p.tri, tm, tn)) continue;
        for (int kb0 = 0; kb0 < kblocks; kb0 += kchunk) {
          const int kb1 = min(kblocks, kb0 + kchunk);
          mbar_wait(&tempty_bar[as], aphase ^ 1);
          tc_fence_after();
          const uint32_t tacc = tmem_base + (uint32_t)(as * SN);
          for (int kb = kb0; kb < kb1; kb++) {
            mbar_wait(&full_bar[stage], phase);
            tc_fence_after();
            const uint32_t s = smem_u32(smem + stage * STAGE);
#pragma unroll
            for (int ks = 0; ks < SKT / 8; ks++) {
              const uint32_t ko = ks * 32;  // 8 floats
              const uint64_t ahi = make_kmajor_desc<SWZ>(s + ko);
              const uint64_t alo = make_kmajor_desc<SWZ>(s + SA + ko);
              const uint64_t bhi = make_kmajor_desc<SWZ>(s + 2 * SA + ko);
              const uint64_t blo = make_kmajor_desc<SWZ>(s + 2 * SA + SB + ko);
              umma_tf32(tacc, ahi, blo, idesc, (kb > kb0 || ks) ? 1u : 0u);
              umma_tf32(tacc, alo, bhi, idesc, 1u);
              umma_tf32(tacc, ahi, bhi, idesc, 1u);
            }
            // frees the smem stage when these MMAs retire (in both CTAs of a pair: the peer's
            // producer writes half of MY next B tile)
            if (CL) umma_commit_multicast(&empty_bar[stage], 3);
            else umma_commit(&empty_bar[stage]);
            if (++stage == NST) { stage = 0; phase ^= 1; }
          }
          umma_commit(&tfull_bar[as]);        // chunk accumulator complete
          if (++as == 2) { as = 0; aphase ^= 1; }
        }
      }
    }
  } else if (warp < 4) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 56;");   // warps 2, 3: the whole warpgroup must
  } else {
    // ===================== epilogue: TMEM -> registers -> C =====================
    asm volatile("setmaxnreg.inc.sync.aligned.u32 216;");
    const int q = warp & 3;                 // TMEM lane quarter of this warp
    const int half = (warp - 4) >> 2;       // which half of the column chunks it drains
    int as = 0;
    uint32_t aphase = 0;
    const bool beta_zero = p.beta == 0.f;
    for (int t = w_first; t < num_tiles; t += w_step) {
      int tm, tn;
      s_tile_coords(t, rows_w, p.tiles_n, tm, tn);
      if (CL) tm = 2 * tm + crank;
      if (s_tile_skipped<SN>(p.tri, tm, tn)) continue;
      const int64_t row = (int64_t)tm * SM_ + q * 32 + lane;
      // The tensor core truncates when it accumulates, so k is cut into chunks of 512 terms; the
      // chunk sums are added in FP32 (round to nearest) in REGISTERS — this thread owns one row
      // and four of the eight 32-column groups of the tile — and C is touched once per tile.
      // (Summing the chunks into C through memory cost 64 GB of extra traffic at N = 16384.)
      float run[SN / 64][32];
      for (int kb0 = 0; kb0 < kblocks; kb0 += kchunk) {
        const bool first = kb0 == 0;
        mbar_wait(&tfull_bar[as], aphase);
        tc_fence_after();
        const uint32_t tbase = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * SN);
#pragma unroll
        for (int ci = 0; ci < SN / 64; ci++) {
          uint32_t v[32];
          tmem_ld32(tbase + (half + 2 * ci) * 32, v);
#pragma unroll
          for (int j = 0; j < 32; j++)
            run[ci][j] = first ? __uint_as_float(v[j]) : run[ci][j] + __uint_as_float(v[j]);
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tempty_bar[as]);   // the accumulator is free again
        if (++as == 2) { as = 0; aphase ^= 1; }
      }
      if (row < p.m) {
#pragma unroll
        for (int ci = 0; ci < SN / 64; ci++) {
          const int64_t col0 = (int64_t)tn * SN + (half + 2 * ci) * 32;
          float* dst0 = p.C + row + col0 * p.ldc;
          float cold[32];
          if (!beta_zero) {  // all loads first: 32 independent requests in flight
#pragma unroll
            for (int j = 0; j < 32; j++)
              cold[j] = (col0 + j < p.n) ? dst0[(int64_t)j * p.ldc] : 0.f;
          }
#pragma unroll
          for (int j = 0; j < 32; j++) {
            const int64_t col = col0 + j;
            if (col >= p.n) break;
            if (p.tri == TriLower && row < col) continue;
            if (p.tri == TriUpper && row > col) continue;
            float r = p.alpha * run[ci][j];
            if (!beta_zero) r = fmaf(p.beta, cold[j], r);
            dst0[(int64_t)j * p.ldc] = r;
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (CL) cluster_sync_all();          // no CTA leaves while its peer may still signal it
  if (warp == 2) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base),
                 "n"(S_TMEM_COLS)
                 : "memory");
  }
}

}  // namespace

bool gemm_tf32x3(Context& ctx, Op ta, Op tb, int64_t m, int64_t n, int64_t k, float alpha,
                 const float* A, int64_t lda, const float* B, int64_t ldb, float beta, float* C,
                 int64_t ldc, int tri) {
  if (ta == OpC) ta = OpT;
  if (tb == OpC) tb = OpT;
  if (m <= 0 || n <= 0) return true;
  if (k <= 0 || alpha == 0.f) return false;
  // below this the packing passes and the 128x256 tiles do not pay off
  if ((double)m * (double)n * (double)k < 256.0 * 256.0 * 256.0) return false;
  if (m > (int64_t)1 << 30 || n > (int64_t)1 << 30 || k > (int64_t)1 << 30) return false;
  static DeviceOnce attr;
  if (attr.pending(ctx.device)) {
    LA_CUDA_CHECK(cudaFuncSetAttribute(sgemm_tf32x3_kernel<false, 256, 16>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, s_smem_bytes(256, 16)));
    LA_CUDA_CHECK(cudaFuncSetAttribute(sgemm_tf32x3_kernel<true, 256, 16>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, s_smem_bytes(256, 16)));
    LA_CUDA_CHECK(cudaFuncSetAttribute(sgemm_tf32x3_kernel<false, 256, 32>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, s_smem_bytes(256, 32)));
    LA_CUDA_CHECK(cudaFuncSetAttribute(sgemm_tf32x3_kernel<true, 256, 32>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, s_smem_bytes(256, 32)));
    LA_CUDA_CHECK(cudaFuncSetAttribute(sgemm_tf32x3_kernel<false, 128, 32>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, s_smem_bytes(128, 32)));
    LA_CUDA_CHECK(cudaFuncSetAttribute(sgemm_tf32x3_kernel<false, 64, 32>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, s_smem_bytes(64, 32)));
    attr.mark(ctx.device);
  }
  Arena::Mark mk = ctx.arena.mark();
  const int64_t ldk = round_up(k, 32);
  float* Ahi = ctx.alloc<float>((size_t)ldk * (size_t)m);
  float* Alo = ctx.alloc<float>((size_t)ldk * (size_t)m);
  float* Bhi = ctx.alloc<float>((size_t)ldk * (size_t)n);
  float* Blo = ctx.alloc<float>((size_t)ldk * (size_t)n);
  {
    // op(A)(r, c): ta == N -> A[r + c*lda] (r contiguous); else A[c + r*lda] (k contiguous)
    launch_split(ctx, ta != OpN, m, k, A, lda, Ahi, Alo, ldk);
    // op(B)(c, r) with r the column: tb == N -> B[c + r*ldb] (k contiguous); else B[r + c*ldb]
    launch_split(ctx, tb == OpN, n, k, B, ldb, Bhi, Blo, ldk);
    LA_CUDA_CHECK(cudaGetLastError());
  }
  // pairs of CTAs sharing B tiles: full products with at least two tile rows and enough tiles to
  // fill the machine with pairs
  // (measured with the four-stage ring: 240 TFLOP/s paired against 260 unpaired at N = 16384 —
  // the pair's lock-step costs more than the L2 traffic it saves; kept behind LA_B200_SGEMM_CLUSTER)
  static const bool use_cluster = getenv("LA_B200_SGEMM_CLUSTER") != nullptr;
  const bool cluster = use_cluster && tri == TriFull && ceil_div(m, SM_) >= 2 &&
                       ceil_div(m, SM_) * ceil_div(n, 256) >= 2 * (int64_t)ctx.sm_count;
  // tile width: 256 unless a narrower tile keeps more SMs busy (score = fraction of the SMs'
  // time spent on tiles x relative efficiency of the shape, as for the FP64 kernel)
  int sn = 256;
  if (!cluster) {
    const int sms_ = ctx.sm_limit > 0 && ctx.sm_limit < ctx.sm_count ? ctx.sm_limit : ctx.sm_count;
    const int cand[3] = {256, 128, 64};
    const double eff[3] = {1.0, 0.9, 0.7};
    double best = -1.0;
    for (int c = 0; c < 3; c++) {
      const int64_t tl = ceil_div(m, SM_) * ceil_div(n, (int64_t)cand[c]);
      const int64_t waves = ceil_div(tl, (int64_t)sms_);
      const double used = (double)n / (double)(ceil_div(n, (int64_t)cand[c]) * cand[c]);
      const double score = (double)tl / (double)(waves * sms_) * used * eff[c];
      if (score > best * 1.05) { best = score; sn = cand[c]; }
    }
  }
  const uint32_t bbox = cluster ? 128 : (uint32_t)sn;
  // 256-wide tiles: 16-float stages (48 KB, four in the ring); LA_B200_SGEMM_SK=32 keeps two 96 KB ones
  static const int sk_env = getenv("LA_B200_SGEMM_SK") ? atoi(getenv("LA_B200_SGEMM_SK")) : 16;
  const int skt = sn == 256 ? (sk_env == 32 ? 32 : 16) : 32;
  const CUtensorMapSwizzle swz = skt == 32 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B;
  CUtensorMap tAh, tAl, tBh, tBl;
  bool ok = make_tensor_map_2d(&tAh, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, Ahi, k, m, ldk * 4, skt, SM_, swz) &&
            make_tensor_map_2d(&tAl, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, Alo, k, m, ldk * 4, skt, SM_, swz) &&
            make_tensor_map_2d(&tBh, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, Bhi, k, n, ldk * 4, skt, bbox, swz) &&
            make_tensor_map_2d(&tBl, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, Blo, k, n, ldk * 4, skt, bbox, swz);
  if (!ok) {
    ctx.arena.release(mk);
    return false;
  }
  SParams p;
  p.m = m; p.n = n; p.k = k;
  p.alpha = alpha; p.beta = beta;
  p.C = C; p.ldc = ldc; p.tri = tri;
  p.tiles_m = (int)ceil_div(m, SM_);
  p.tiles_n = (int)ceil_div(n, (int64_t)sn);
  // terms per tensor-memory accumulation chunk (the tensor core truncates when it accumulates,
  // a bias that grows with the chunk): measured at N = 16384, error against FP64 / TFLOP/s:
  // 512 -> 3.6e-6 / 262, 256 -> 1.8e-6 / 252, 128 -> 9.3e-7 / 245, 64 -> 5.4e-7 / 232
  static const int kc_env = getenv("LA_B200_SGEMM_KCHUNK") ? atoi(getenv("LA_B200_SGEMM_KCHUNK")) : 256;
  p.kchunk_terms = kc_env >= 32 ? kc_env / 32 * 32 : 256;
  const int tiles = p.tiles_m * p.tiles_n;
  const int sms = ctx.sm_limit > 0 && ctx.sm_limit < ctx.sm_count ? ctx.sm_limit : ctx.sm_count;
  if (cluster) {
    const int pairs = (int)(ceil_div(p.tiles_m, 2) * p.tiles_n);
    int grid = 2 * (pairs < sms / 2 ? pairs : sms / 2);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(S_THREADS);
    cfg.dynamicSmemBytes = s_smem_bytes(256, skt);
    cfg.stream = ctx.stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    if (skt == 16)
      LA_CUDA_CHECK(cudaLaunchKernelEx(&cfg, sgemm_tf32x3_kernel<true, 256, 16>, tAh, tAl, tBh, tBl, p));
    else
      LA_CUDA_CHECK(cudaLaunchKernelEx(&cfg, sgemm_tf32x3_kernel<true, 256, 32>, tAh, tAl, tBh, tBl, p));
  } else {
    const int grid = tiles < sms ? tiles : sms;
    if (sn == 256 && skt == 16)
      sgemm_tf32x3_kernel<false, 256, 16><<<grid, S_THREADS, s_smem_bytes(256, 16), ctx.stream>>>(tAh, tAl, tBh, tBl, p);
    else if (sn == 256)
      sgemm_tf32x3_kernel<false, 256, 32><<<grid, S_THREADS, s_smem_bytes(256, 32), ctx.stream>>>(tAh, tAl, tBh, tBl, p);
    else if (sn == 128)
      sgemm_tf32x3_kernel<false, 128, 32><<<grid, S_THREADS, s_smem_bytes(128, 32), ctx.stream>>>(tAh, tAl, tBh, tBl, p);
    else
      sgemm_tf32x3_kernel<false, 64, 32><<<grid, S_THREADS, s_smem_bytes(64, 32), ctx.stream>>>(tAh, tAl, tBh, tBl, p);
  }
  LA_CUDA_CHECK(cudaGetLastError());
  ctx.arena.release(mk);
  return true;
}

}  // namespace la
