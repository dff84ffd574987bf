// la_b200 — FP64 TRSM base case (triangular block of order <= 64) on the DMMA tensor cores.
//
// The recursive TRSM of level3.cu bottoms out in "solve against a 64x64 triangle for many
// independent vectors" (rows of B for side = Right, columns for side = Left).  One thread per
// vector leaves the FP64 pipe almost idle (a 16384-row panel is only 512 warps), so the base case
// is recast as tensor-core work.  With the vectors as ROWS of X (X U = alpha B, U the effective
// upper/lower triangular matrix) and U cut in 8x8 blocks:
//      X_j = ( alpha B_j - sum_{i before j} X_i U_ij ) * inv(U_jj)
// every term is an 8x8x8 product = two mma.sync.m8n8k4.f64.  The accumulator fragment of one
// DMMA (thread holds row gid, columns 2*tig, 2*tig+1) IS the A fragment of the next one when the
// k-slots are numbered k = 2*tig + step, so X never leaves registers and no shuffle is needed.
// The 8x8 diagonal blocks are inverted once per CTA in shared memory (tiny, well conditioned
// relative to the whole triangle); off-diagonal blocks are stored negated.
// Work: na^2/2 * nvec FMAs (+ the diagonal-block products); HBM traffic 2 * na * nvec * 8 B.
#include "la_internal.h"

namespace la {

namespace {

constexpr int TB = 8;            // block order
constexpr int TGROUPS = 2;       // 8-vector groups per warp (independent chains)
// Two instances: triangles up to 64 (4 warps, 64 vectors per CTA) and up to 128 (8 warps, 128
// vectors per CTA, triangle in dynamic shared memory); the larger one saves a level of the
// recursion (two base launches and a thin GEMM) wherever there are enough vectors to fill it.

__device__ __forceinline__ void dmma(double (&c)[2], double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c[0]), "+d"(c[1])
               : "d"(a), "d"(b));
}

// X (nvec x na, element (v,i) at X[v*vs + i*is]) <- alpha * X * U^-1, U = na x na triangular
// (upper != 0: upper), given as Umat(i,j) = U[i*ur + j*uc] with optional unit diagonal.
template <int TNB, int TWARPS>
__global__ void __launch_bounds__(TWARPS * 32)
trsm_dmma_kernel(int upper, int unit, int na, int64_t nvec, double alpha,
                 const double* __restrict__ U, int64_t ur, int64_t uc, double* __restrict__ X,
                 int64_t vs, int64_t is) {
  constexpr int TLD = TNB + 1;       // shared-memory pitch
  extern __shared__ __align__(16) unsigned char trsm_smem[];
  double* Us = reinterpret_cast<double*>(trsm_smem);   // [TNB * TLD] -U off the diagonal blocks, inv(U_jj) on them
  double* Dg = Us + TNB * TLD;                         // [TNB * TB] scratch copy of the diagonal blocks
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int gid = lane >> 2, tig = lane & 3;
  const int nblk = (na + TB - 1) / TB;
  const int npad = nblk * TB;
  // ---- stage U (padded with the identity), negated.  All global loads of a thread are issued
  // before the first shared-memory store (one memory round trip instead of one per element),
  // and the fast thread index follows the contiguous direction of the stored matrix.
  {
    constexpr int TOTAL = TNB * TNB / (TWARPS * 32);   // elements per thread
    constexpr int PER = 32;                            // of which 32 are in flight at a time
    constexpr int SH = TNB == 64 ? 6 : 7;
    const bool row_contig = (uc == 1);
#pragma unroll 1
    for (int round = 0; round < TOTAL / PER; round++) {
      double v[PER];
#pragma unroll
      for (int q = 0; q < PER; q++) {
        const int idx = tid + (round * PER + q) * (TWARPS * 32);
        const int fa = idx & (TNB - 1), sl = idx >> SH;
        const int i = row_contig ? sl : fa, j = row_contig ? fa : sl;
        double x = (i == j) ? 1.0 : 0.0;
        if (i < na && j < na) {
          const bool stored = upper ? (i <= j) : (i >= j);
          x = 0.0;
          if (stored) x = (i == j && unit) ? 1.0 : U[i * ur + j * uc];
        }
        v[q] = x;
      }
#pragma unroll
      for (int q = 0; q < PER; q++) {
        const int idx = tid + (round * PER + q) * (TWARPS * 32);
        const int fa = idx & (TNB - 1), sl = idx >> SH;
        const int i = row_contig ? sl : fa, j = row_contig ? fa : sl;
        if (i < npad && j < npad) {
          Us[i * TLD + j] = -v[q];
          if (i / TB == j / TB) Dg[(i / TB) * (TB * TB) + (i % TB) * TB + (j % TB)] = v[q];
        }
      }
    }
  }
  __syncthreads();
  // ---- invert the diagonal blocks: thread (b, c) computes column c of inv(U_bb)
  if (tid < nblk * TB) {
    const int b = tid / TB, c = tid % TB;
    const double* D = Dg + b * (TB * TB);
    double x[TB];
#pragma unroll
    for (int i = 0; i < TB; i++) x[i] = (i == c) ? 1.0 : 0.0;
    if (upper) {
#pragma unroll
      for (int jj = 0; jj < TB; jj++) {
        const int j = TB - 1 - jj;
        x[j] = x[j] / D[j * TB + j];
#pragma unroll
        for (int i = 0; i < TB; i++)
          if (i < j) x[i] = fma(-D[i * TB + j], x[j], x[i]);
      }
    } else {
#pragma unroll
      for (int j = 0; j < TB; j++) {
        x[j] = x[j] / D[j * TB + j];
#pragma unroll
        for (int i = 0; i < TB; i++)
          if (i > j) x[i] = fma(-D[i * TB + j], x[j], x[i]);
      }
    }
#pragma unroll
    for (int i = 0; i < TB; i++) Us[(b * TB + i) * TLD + b * TB + c] = x[i];
  }
  __syncthreads();

  // ---- solve: each warp owns TGROUPS groups of 8 vectors
  const int64_t v0 = ((int64_t)blockIdx.x * TWARPS + warp) * (TGROUPS * 8);
  double xb[TGROUPS][TNB / TB][2];   // X blocks: [group][block][2 columns of this thread]
#pragma unroll
  for (int g = 0; g < TGROUPS; g++) {
    const int64_t v = v0 + g * 8 + gid;
#pragma unroll
    for (int b = 0; b < TNB / TB; b++) {
#pragma unroll
      for (int s = 0; s < 2; s++) {
        const int i = b * TB + 2 * tig + s;
        xb[g][b][s] = (v < nvec && i < na) ? alpha * X[v * vs + i * is] : 0.0;
      }
    }
  }
  // block j (static index, so that xb stays in registers): acc = alpha B_j - sum_i X_i U_ij over
  // the blocks solved before it, then X_j = acc * inv(U_jj)
  auto solve_block = [&](const int j, const bool up) {
#pragma unroll
    for (int ii = 0; ii < TNB / TB; ii++) {
      const bool before = up ? (ii < j) : (ii > j);
      if (before && ii < nblk) {
        // B fragments of (-U_ij): thread holds -U[8*ii + 2*tig + s][8*j + gid]
        const double b0 = Us[(ii * TB + 2 * tig + 0) * TLD + j * TB + gid];
        const double b1 = Us[(ii * TB + 2 * tig + 1) * TLD + j * TB + gid];
#pragma unroll
        for (int g = 0; g < TGROUPS; g++) {
          dmma(xb[g][j], xb[g][ii][0], b0);
          dmma(xb[g][j], xb[g][ii][1], b1);
        }
      }
    }
    const double d0 = Us[(j * TB + 2 * tig + 0) * TLD + j * TB + gid];
    const double d1 = Us[(j * TB + 2 * tig + 1) * TLD + j * TB + gid];
#pragma unroll
    for (int g = 0; g < TGROUPS; g++) {
      double r[2] = {0.0, 0.0};
      dmma(r, xb[g][j][0], d0);
      dmma(r, xb[g][j][1], d1);
      xb[g][j][0] = r[0];
      xb[g][j][1] = r[1];
    }
  };
  if (upper) {   // forward over the column blocks
#pragma unroll
    for (int j = 0; j < TNB / TB; j++)
      if (j < nblk) solve_block(j, true);
  } else {       // backward
#pragma unroll
    for (int jj = 0; jj < TNB / TB; jj++) {
      const int j = TNB / TB - 1 - jj;
      if (j < nblk) solve_block(j, false);
    }
  }
#pragma unroll
  for (int g = 0; g < TGROUPS; g++) {
    const int64_t v = v0 + g * 8 + gid;
#pragma unroll
    for (int b = 0; b < TNB / TB; b++) {
#pragma unroll
      for (int s = 0; s < 2; s++) {
        const int i = b * TB + 2 * tig + s;
        if (v < nvec && i < na) X[v * vs + i * is] = xb[g][b][s];
      }
    }
  }
}

}  // namespace

// Base case hook used by trsm<double>: returns false when the block is too large.
bool trsm_base_dmma_f64(Context& ctx, Side side, Uplo uplo, Op op, Diag diag, int64_t m, int64_t n,
                        double alpha, const double* A, int64_t lda, double* B, int64_t ldb) {
  const int64_t na = side == Left ? m : n;
  const int64_t nvec = side == Left ? n : m;
  // the 128 instance only where its CTAs (128 vectors each) fill the machine's SMs about once
  if (na <= 0 || nvec <= 0 || na > 128 || (na > 64 && nvec < 4096)) return false;
  // Right: X op(A) = alpha B, vectors are rows of B:         U = op(A),   vs = 1,   is = ldb
  // Left : op(A) X = alpha B  <=>  X^T op(A)^T = alpha B^T:  U = op(A)^T, vs = ldb, is = 1
  const bool trans = op != OpN;
  bool u_is_transposed_A = (side == Right) ? trans : !trans;   // U = A^T ?
  // U(i,j) = A(i,j) -> ur = 1, uc = lda ;  U(i,j) = A(j,i) -> ur = lda, uc = 1
  const int64_t ur = u_is_transposed_A ? lda : 1;
  const int64_t uc = u_is_transposed_A ? 1 : lda;
  const bool a_upper = uplo == Upper;
  const bool u_upper = u_is_transposed_A ? !a_upper : a_upper;
  const int64_t vs = side == Right ? 1 : ldb;
  const int64_t is = side == Right ? ldb : 1;
  if (na <= 64) {
    constexpr int W = 4;
    constexpr size_t smem = sizeof(double) * (64 * 65 + 64 * TB);
    const unsigned grid = (unsigned)ceil_div(nvec, (int64_t)W * TGROUPS * 8);
    trsm_dmma_kernel<64, W><<<grid, W * 32, smem, ctx.stream>>>(
        u_upper ? 1 : 0, diag == Unit ? 1 : 0, (int)na, nvec, alpha, A, ur, uc, B, vs, is);
  } else {
    constexpr int W = 8;
    constexpr size_t smem = sizeof(double) * (128 * 129 + 128 * TB);
    static DeviceOnce attr;
    if (attr.pending(ctx.device)) {
      LA_CUDA_CHECK(cudaFuncSetAttribute(trsm_dmma_kernel<128, W>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      attr.mark(ctx.device);
    }
    const unsigned grid = (unsigned)ceil_div(nvec, (int64_t)W * TGROUPS * 8);
    trsm_dmma_kernel<128, W><<<grid, W * 32, smem, ctx.stream>>>(
        u_upper ? 1 : 0, diag == Unit ? 1 : 0, (int)na, nvec, alpha, A, ur, uc, B, vs, is);
  }
  LA_CUDA_CHECK(cudaGetLastError());
  return true;
}

}  // namespace la
