// la_b200 — internal device-level API (all pointers are DEVICE pointers, column-major).
//
// One function per library routine the reference reaches through its FFI
// (SURVEY.md §8a).  Every routine only enqueues work on ctx.stream; LAPACK `info`
// words live on the device (d_info) and are read back by the C-ABI layer.
#pragma once
#include "common.cuh"
#include "context.h"

namespace la {

enum TriMask : int { TriFull = 0, TriLower = 1, TriUpper = 2 };

// C = alpha*op(A)*op(B) + beta*C.  `tri` restricts the stored part of C to one
// triangle (syrk/herk style); `herm_diag` forces Im(C_ii)=0 (herk).
template <typename T>
void gemm(Context& ctx, Op ta, Op tb, int64_t m, int64_t n, int64_t k, T alpha, const T* A,
          int64_t lda, const T* B, int64_t ldb, T beta, T* C, int64_t ldc, int tri = TriFull,
          bool herm_diag = false);

// Block-staircase mask for products on block-cyclic LOCAL matrices (multi-GPU factorisations):
// local rows [i*nb, (i+1)*nb) of C are global block row_first + i*row_step, local columns
// [j*nb, (j+1)*nb) global block col_first + j*col_step; only blocks whose global row block is
// >= the global column block are computed (whole blocks: no element mask inside them).
struct StairMask {
  int64_t nb = 0;
  int64_t row_first = 0, row_step = 1, col_first = 0, col_step = 1;
};
template <typename T>
void gemm_stair(Context& ctx, Op ta, Op tb, int64_t m, int64_t n, int64_t k, T alpha, const T* A,
                int64_t lda, const T* B, int64_t ldb, T beta, T* C, int64_t ldc,
                const StairMask& mask);

// Which GEMM path the calling thread used last (0 = SIMT, 1 = DMMA f64, 2 = DMMA complex,
// 3 = 3xTF32 tcgen05); written by the dispatcher, read by la_b200_last_gemm_path().
int& last_gemm_path();

// Generic SIMT implementation (any alignment / leading dimension / type).
template <typename T>
void gemm_simt(Context& ctx, Op ta, Op tb, int64_t m, int64_t n, int64_t k, T alpha, const T* A,
               int64_t lda, const T* B, int64_t ldb, T beta, T* C, int64_t ldc, int tri,
               bool herm_diag);

// Fast paths; return false when the shape/alignment is not supported.
bool gemm_dmma_f64(Context& ctx, Op ta, Op tb, int64_t m, int64_t n, int64_t k, double alpha,
                   const double* A, int64_t lda, const double* B, int64_t ldb, double beta,
                   double* C, int64_t ldc, int tri);
// same kernel with the complex-interleaved row mask / diagonal rule used by gemm_z.cu
bool gemm_dmma_f64_ex(Context& ctx, Op ta, Op tb, int64_t m, int64_t n, int64_t k, double alpha,
                      const double* A, int64_t lda, const double* B, int64_t ldb, double beta,
                      double* C, int64_t ldc, int tri, int zmode, bool herm_diag,
                      const StairMask* stair = nullptr);
bool gemm_dmma_z(Context& ctx, Op ta, Op tb, int64_t m, int64_t n, int64_t k, zcomplex alpha,
                 const zcomplex* A, int64_t lda, const zcomplex* B, int64_t ldb, zcomplex beta,
                 zcomplex* C, int64_t ldc, int tri, bool herm_diag);
bool gemm_tf32x3(Context& ctx, Op ta, Op tb, int64_t m, int64_t n, int64_t k, float alpha,
                 const float* A, int64_t lda, const float* B, int64_t ldb, float beta, float* C,
                 int64_t ldc, int tri);

// S (p x q) = X^H Y for tall X (m x p), Y (m x q), p, q <= 64 (tall_gram.cu); false: not taken.
template <typename T>
bool tall_gram(Context& ctx, int64_t p, int64_t q, int64_t m, const T* X, int64_t ldx, const T* Y,
               int64_t ldy, T* S, int64_t lds);

// C (m x nc) -= V (m x h) Z^H for tall m and h, nc <= 64 (tall_gram.cu); false: not taken.
template <typename T>
bool tall_update(Context& ctx, int64_t m, int64_t h, int64_t nc, const T* V, int64_t ldv, const T* Z,
                 int64_t ldz, T* C, int64_t ldc);

// C = alpha*op(A)*op(A)^H + beta*C on one triangle (trans = OpN: A is n x k; else k x n).
template <typename T>
void herk(Context& ctx, Uplo uplo, Op trans, int64_t n, int64_t k, double alpha, const T* A,
          int64_t lda, double beta, T* C, int64_t ldc);

// op(A) X = alpha B for nrhs <= 8 right-hand sides in one cooperative launch (trsv.cu); returns
// false when the shape is not for that path.
template <typename T>
bool trsv_small(Context& ctx, Uplo uplo, Op op, Diag diag, int64_t n, int64_t nrhs, T alpha,
                const T* A, int64_t lda, T* B, int64_t ldb);

template <typename T>
void trsm(Context& ctx, Side side, Uplo uplo, Op op, Diag diag, int64_t m, int64_t n, T alpha,
          const T* A, int64_t lda, T* B, int64_t ldb);

template <typename T>
void trmm(Context& ctx, Side side, Uplo uplo, Op op, Diag diag, int64_t m, int64_t n, T alpha,
          const T* A, int64_t lda, T* B, int64_t ldb);

// symm (herm=false) / hemm (herm=true)
template <typename T>
void symm(Context& ctx, bool herm, Side side, Uplo uplo, int64_t m, int64_t n, T alpha,
          const T* A, int64_t lda, const T* B, int64_t ldb, T beta, T* C, int64_t ldc);

// elementwise helpers
template <typename T>
void copy_matrix(Context& ctx, int64_t m, int64_t n, const T* A, int64_t lda, T* B, int64_t ldb);
template <typename T>  // B = op(A); A is m x n, B is n x m
void transpose_matrix(Context& ctx, bool conj, int64_t m, int64_t n, const T* A, int64_t lda,
                      T* B, int64_t ldb);
template <typename T>
void set_matrix(Context& ctx, int64_t m, int64_t n, T offdiag, T diag, T* A, int64_t lda);
// copy the `uplo` triangle of A (n x n) to B, zero the other, optional unit diagonal
template <typename T>
void copy_triangle(Context& ctx, Uplo uplo, bool unit, int64_t m, int64_t n, const T* A,
                   int64_t lda, T* B, int64_t ldb);
// B = full symmetric/hermitian matrix from the `uplo` triangle of A
template <typename T>
void symmetrize(Context& ctx, bool herm, Uplo uplo, int64_t n, const T* A, int64_t lda, T* B,
                int64_t ldb);

// row interchanges: for k in [k1,k2) (or reversed) swap rows k and ipiv[k]-1 of the n columns
template <typename T>
void laswp(Context& ctx, int64_t n, T* A, int64_t lda, int64_t k1, int64_t k2, const int* d_ipiv,
           bool reverse);

// the same interchanges as one net permutation per 1024 pivots (parallel over rows and columns);
// m = number of rows of A (upper bound of the pivot indices)
template <typename T>
void laswp_rows(Context& ctx, int64_t m, int64_t n, T* A, int64_t lda, int64_t k1, int64_t k2,
                const int* d_ipiv, bool reverse = false);

// factorizations / solves
// Width of the block columns / rows in which potrf reports its result final through
// Context::on_block_final (n itself when it does not report).
int64_t potrf_block_width(int64_t n);
template <typename T>
void potrf(Context& ctx, Uplo uplo, int64_t n, T* A, int64_t lda, int* d_info);
template <typename T>
void potrs(Context& ctx, Uplo uplo, int64_t n, int64_t nrhs, const T* A, int64_t lda, T* B,
           int64_t ldb);
template <typename T>
void getrf(Context& ctx, int64_t m, int64_t n, T* A, int64_t lda, int* d_ipiv, int* d_info);
template <typename T>
void getrs(Context& ctx, Op trans, int64_t n, int64_t nrhs, const T* A, int64_t lda,
           const int* d_ipiv, T* B, int64_t ldb);
template <typename T>
void trtrs(Context& ctx, Uplo uplo, Op trans, Diag diag, int64_t n, int64_t nrhs, const T* A,
           int64_t lda, T* B, int64_t ldb, int* d_info);
template <typename T>
void geqrf(Context& ctx, int64_t m, int64_t n, T* A, int64_t lda, T* d_tau);
template <typename T>  // multiply C by Q or Q^H from geqrf reflectors (ormqr / unmqr)
void ormqr(Context& ctx, Side side, Op trans, int64_t m, int64_t n, int64_t k, const T* A,
           int64_t lda, const T* d_tau, T* C, int64_t ldc);
template <typename T>  // generate the leading n columns of Q (orgqr / ungqr)
void orgqr(Context& ctx, int64_t m, int64_t n, int64_t k, T* A, int64_t lda, const T* d_tau);
// matrix norms (aux_ops.cu): the result lands in a device double
enum NormKind : int { NormMax = 0, NormOne = 1, NormInf = 2, NormFro = 3 };
enum NormShape : int { ShapeGeneral = 0, ShapeSym = 1, ShapeHerm = 2, ShapeTri = 3 };
int norm_kind_from_char(char c);    // 'M', 'O' / '1', 'I', 'F' / 'E' -> NormKind, else -1
template <typename T>
void matrix_norm(Context& ctx, int kind, int shape, Uplo uplo, bool unit, int64_t m, int64_t n,
                 const T* A, int64_t lda, double* d_out);

// RQ ('R'), LQ ('L'), QL ('Q') factorisations in LAPACK's storage, and their Q generators
template <typename T>
void geqrf_variant(Context& ctx, char which, int64_t m, int64_t n, T* A, int64_t lda, T* d_tau);
template <typename T>
void orgqr_variant(Context& ctx, char which, int64_t m, int64_t n, int64_t k, T* A, int64_t lda,
                   const T* d_tau);
// QR with column pivoting; d_jpvt holds the starting permutation (fixed columns first)
template <typename T>
void geqp3(Context& ctx, int64_t m, int64_t n, T* A, int64_t lda, int* d_jpvt, T* d_tau,
           int64_t nfixed);
template <typename T>
void swap_columns(Context& ctx, int64_t m, T* a, T* b);
template <typename T>
void gels(Context& ctx, Op trans, int64_t m, int64_t n, int64_t nrhs, T* A, int64_t lda, T* B,
          int64_t ldb, int* d_info);
// inverses built on the factors
template <typename T>
void trtri(Context& ctx, Uplo uplo, Diag diag, int64_t n, T* A, int64_t lda, int* d_info);
template <typename T>
void getri(Context& ctx, int64_t n, T* A, int64_t lda, const int* d_ipiv, int* d_info);
template <typename T>
void potri(Context& ctx, Uplo uplo, int64_t n, T* A, int64_t lda, int* d_info);

}  // namespace la
