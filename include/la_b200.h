/* la_b200.h — C ABI of libla_b200.so, the B200 (sm_100a) backend behind konovod/linalg's
 * BLAS/LAPACK binding points.
 *
 * Part 1 and 2 are DROP-IN symbols: same names, argument order, types and in/out rules as
 * the `fun` declarations in the reference's `lib LibCBLAS` / `lib LibLAPACK` blocks (each
 * entry cites the declaration it replaces, path relative to the reference root).  Installing
 * the library under the sonames the reference links (`-lcblas`, `-llapack`:
 * src/linalg/libCBLAS.cr:1-7, src/linalg/libLAPACK_linux.cr:5-9 — its CI does exactly this
 * with symlinks, .github/workflows/linux.yml:26-29) needs no Crystal change.
 *   - all matrix pointers are HOST pointers, column-major unless `order` says otherwise;
 *   - `int` is 32-bit (LibCBLAS::Blasint = LibC::Int, libCBLAS.cr:26);
 *   - LAPACK symbols use the Fortran-77 convention: every argument by pointer, character
 *     flags as `char*` WITHOUT hidden length arguments, `info` last;
 *   - lwork = -1 is a workspace query (work[0] receives the size to allocate, real part for
 *     complex: src/linalg/lapack_helper.cr:215-280); `work` is never written beyond work[0];
 *   - every call is synchronous and thread-safe (spec/linalg_spec.cr:519-539).
 * Part 3 is the device-resident API (device pointers, caller's stream, no host copies) used
 * by the benchmark, the multi-GPU drivers and by hosts that keep operands on the GPU.
 */
#ifndef LA_B200_H
#define LA_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- enum values (src/linalg/libCBLAS.cr:19-48) --------------------------------------- */
enum { LA_ROW_MAJOR = 101, LA_COL_MAJOR = 102 };
enum { LA_NO_TRANS = 111, LA_TRANS = 112, LA_CONJ_TRANS = 113 };
enum { LA_UPPER = 121, LA_LOWER = 122 };
enum { LA_NON_UNIT = 131, LA_UNIT = 132 };
enum { LA_LEFT = 141, LA_RIGHT = 142 };

/* ===================== Part 1: CBLAS drop-in symbols (host pointers) ==================== */
/* src/linalg/libCBLAS.cr:174 */
void cblas_sgemm(int order, int transa, int transb, int m, int n, int k, float alpha,
                 const float* a, int lda, const float* b, int ldb, float beta, float* c, int ldc);
/* src/linalg/libCBLAS.cr:175 */
void cblas_dgemm(int order, int transa, int transb, int m, int n, int k, double alpha,
                 const double* a, int lda, const double* b, int ldb, double beta, double* c,
                 int ldc);
/* src/linalg/libCBLAS.cr:178 (alpha, beta by pointer to {re, im}) */
void cblas_zgemm(int order, int transa, int transb, int m, int n, int k, const void* alpha,
                 const void* a, int lda, const void* b, int ldb, const void* beta, void* c,
                 int ldc);
/* src/linalg/libCBLAS.cr:180, :181, :183 */
void cblas_ssymm(int order, int side, int uplo, int m, int n, float alpha, const float* a,
                 int lda, const float* b, int ldb, float beta, float* c, int ldc);
void cblas_dsymm(int order, int side, int uplo, int m, int n, double alpha, const double* a,
                 int lda, const double* b, int ldb, double beta, double* c, int ldc);
void cblas_zsymm(int order, int side, int uplo, int m, int n, const void* alpha, const void* a,
                 int lda, const void* b, int ldb, const void* beta, void* c, int ldc);
/* src/linalg/libCBLAS.cr:201 */
void cblas_zhemm(int order, int side, int uplo, int m, int n, const void* alpha, const void* a,
                 int lda, const void* b, int ldb, const void* beta, void* c, int ldc);
/* src/linalg/libCBLAS.cr:192, :193, :195 */
void cblas_strmm(int order, int side, int uplo, int transa, int diag, int m, int n, float alpha,
                 const float* a, int lda, float* b, int ldb);
void cblas_dtrmm(int order, int side, int uplo, int transa, int diag, int m, int n, double alpha,
                 const double* a, int lda, double* b, int ldb);
void cblas_ztrmm(int order, int side, int uplo, int transa, int diag, int m, int n,
                 const void* alpha, const void* a, int lda, void* b, int ldb);
/* src/linalg/libCBLAS.cr:196, :197, :199 */
void cblas_strsm(int order, int side, int uplo, int transa, int diag, int m, int n, float alpha,
                 const float* a, int lda, float* b, int ldb);
void cblas_dtrsm(int order, int side, int uplo, int transa, int diag, int m, int n, double alpha,
                 const double* a, int lda, double* b, int ldb);
void cblas_ztrsm(int order, int side, int uplo, int transa, int diag, int m, int n,
                 const void* alpha, const void* a, int lda, void* b, int ldb);
/* src/linalg/libCBLAS.cr:184, :185, :203 */
void cblas_ssyrk(int order, int uplo, int trans, int n, int k, float alpha, const float* a,
                 int lda, float beta, float* c, int ldc);
void cblas_dsyrk(int order, int uplo, int trans, int n, int k, double alpha, const double* a,
                 int lda, double beta, double* c, int ldc);
void cblas_zherk(int order, int uplo, int trans, int n, int k, double alpha, const void* a,
                 int lda, double beta, void* c, int ldc);
/* src/linalg/libCBLAS.cr:50-55 — accepted and ignored (there is no host thread pool) */
void openblas_set_num_threads(int num_threads);
int openblas_get_num_threads(void);
int openblas_get_num_procs(void);
char* openblas_get_config(void);
char* openblas_get_corename(void);
int openblas_get_parallel(void);

/* ===================== Part 2: LAPACK drop-in symbols (host pointers) =================== */
/* src/linalg/libLAPACK_linux.cr:1126 — reports LAPACK 3.12.0, the version whose algorithms
 * and conventions this backend follows */
void ilaver_(int* vers_major, int* vers_minor, int* vers_patch);

/* ?getrf: src/linalg/libLAPACK_linux.cr:1213 (s), :645 (d), :1755 (z) */
void sgetrf_(const int* m, const int* n, float* a, const int* lda, int* ipiv, int* info);
void dgetrf_(const int* m, const int* n, double* a, const int* lda, int* ipiv, int* info);
void zgetrf_(const int* m, const int* n, void* a, const int* lda, int* ipiv, int* info);
/* ?getrs: :1216, :648, :1758 */
void sgetrs_(const char* trans, const int* n, const int* nrhs, const float* a, const int* lda,
             const int* ipiv, float* b, const int* ldb, int* info);
void dgetrs_(const char* trans, const int* n, const int* nrhs, const double* a, const int* lda,
             const int* ipiv, double* b, const int* ldb, int* info);
void zgetrs_(const char* trans, const int* n, const int* nrhs, const void* a, const int* lda,
             const int* ipiv, void* b, const int* ldb, int* info);
/* ?gesv: :1204, :636, :1746 */
void sgesv_(const int* n, const int* nrhs, float* a, const int* lda, int* ipiv, float* b,
            const int* ldb, int* info);
void dgesv_(const int* n, const int* nrhs, double* a, const int* lda, int* ipiv, double* b,
            const int* ldb, int* info);
void zgesv_(const int* n, const int* nrhs, void* a, const int* lda, int* ipiv, void* b,
            const int* ldb, int* info);
/* ?potrf: :1512, :944, :2065 */
void spotrf_(const char* uplo, const int* n, float* a, const int* lda, int* info);
void dpotrf_(const char* uplo, const int* n, double* a, const int* lda, int* info);
void zpotrf_(const char* uplo, const int* n, void* a, const int* lda, int* info);
/* ?potrs: :1515, :947, :2068 */
void spotrs_(const char* uplo, const int* n, const int* nrhs, const float* a, const int* lda,
             float* b, const int* ldb, int* info);
void dpotrs_(const char* uplo, const int* n, const int* nrhs, const double* a, const int* lda,
             double* b, const int* ldb, int* info);
void zpotrs_(const char* uplo, const int* n, const int* nrhs, const void* a, const int* lda,
             void* b, const int* ldb, int* info);
/* ?posv: :1508, :940, :2061 */
void sposv_(const char* uplo, const int* n, const int* nrhs, float* a, const int* lda, float* b,
            const int* ldb, int* info);
void dposv_(const char* uplo, const int* n, const int* nrhs, double* a, const int* lda,
            double* b, const int* ldb, int* info);
void zposv_(const char* uplo, const int* n, const int* nrhs, void* a, const int* lda, void* b,
            const int* ldb, int* info);
/* ?geqrf: :1193, :625, :1735 */
void sgeqrf_(const int* m, const int* n, float* a, const int* lda, float* tau, float* work,
             const int* lwork, int* info);
void dgeqrf_(const int* m, const int* n, double* a, const int* lda, double* tau, double* work,
             const int* lwork, int* info);
void zgeqrf_(const int* m, const int* n, void* a, const int* lda, void* tau, void* work,
             const int* lwork, int* info);
/* ?ormqr / zunmqr: :1485, :917, :2216 */
void sormqr_(const char* side, const char* trans, const int* m, const int* n, const int* k,
             const float* a, const int* lda, const float* tau, float* c, const int* ldc,
             float* work, const int* lwork, int* info);
void dormqr_(const char* side, const char* trans, const int* m, const int* n, const int* k,
             const double* a, const int* lda, const double* tau, double* c, const int* ldc,
             double* work, const int* lwork, int* info);
void zunmqr_(const char* side, const char* trans, const int* m, const int* n, const int* k,
             const void* a, const int* lda, const void* tau, void* c, const int* ldc, void* work,
             const int* lwork, int* info);
/* ?gels: :1176, :608, :1718 */
void sgels_(const char* trans, const int* m, const int* n, const int* nrhs, float* a,
            const int* lda, float* b, const int* ldb, float* work, const int* lwork, int* info);
void dgels_(const char* trans, const int* m, const int* n, const int* nrhs, double* a,
            const int* lda, double* b, const int* ldb, double* work, const int* lwork, int* info);
void zgels_(const char* trans, const int* m, const int* n, const int* nrhs, void* a,
            const int* lda, void* b, const int* ldb, void* work, const int* lwork, int* info);
/* ?trtrs: :1670, :1106, :2180 */
void strtrs_(const char* uplo, const char* trans, const char* diag, const int* n, const int* nrhs,
             const float* a, const int* lda, float* b, const int* ldb, int* info);
void dtrtrs_(const char* uplo, const char* trans, const char* diag, const int* n, const int* nrhs,
             const double* a, const int* lda, double* b, const int* ldb, int* info);
void ztrtrs_(const char* uplo, const char* trans, const char* diag, const int* n, const int* nrhs,
             const void* a, const int* lda, void* b, const int* ldb, int* info);
/* "next" row 1 of SURVEY.md §8(f): the inv! family on the same factors
 * ?getri: :1215, :647, :1757   ?potri: :1514, :946, :2067   ?trtri: :1669, :1105, :2179 */
void sgetri_(const int* n, float* a, const int* lda, const int* ipiv, float* work,
             const int* lwork, int* info);
void dgetri_(const int* n, double* a, const int* lda, const int* ipiv, double* work,
             const int* lwork, int* info);
void zgetri_(const int* n, void* a, const int* lda, const int* ipiv, void* work, const int* lwork,
             int* info);
void spotri_(const char* uplo, const int* n, float* a, const int* lda, int* info);
void dpotri_(const char* uplo, const int* n, double* a, const int* lda, int* info);
void zpotri_(const char* uplo, const int* n, void* a, const int* lda, int* info);
void strtri_(const char* uplo, const char* diag, const int* n, float* a, const int* lda, int* info);
void dtrtri_(const char* uplo, const char* diag, const int* n, double* a, const int* lda, int* info);
void ztrtri_(const char* uplo, const char* diag, const int* n, void* a, const int* lda, int* info);
/* "next" row 2: explicit Q — ?orgqr / zungqr: :1470, :902, :2201 (src/linalg/qr.cr:107) */
void sorgqr_(const int* m, const int* n, const int* k, float* a, const int* lda, const float* tau,
             float* work, const int* lwork, int* info);
void dorgqr_(const int* m, const int* n, const int* k, double* a, const int* lda,
             const double* tau, double* work, const int* lwork, int* info);
void zungqr_(const int* m, const int* n, const int* k, void* a, const int* lda, const void* tau,
             void* work, const int* lwork, int* info);
/* "next" row 2, rest: the RQ / LQ / QL siblings (src/linalg/rq_lq_ql.cr:3-39)
 * ?gerqf: :1201, :633, :1743   ?gelqf: :1173, :605, :1715   ?geqlf: :1187, :619, :1729
 * ?orgrq / zungrq: :1472, :904, :2203   ?orglq / zunglq: :1468, :900, :2199
 * ?orgql / zungql: :1469, :901, :2200 */
void sgerqf_(const int* m, const int* n, float* a, const int* lda, float* tau, float* work,
             const int* lwork, int* info);
void sgelqf_(const int* m, const int* n, float* a, const int* lda, float* tau, float* work,
             const int* lwork, int* info);
void sgeqlf_(const int* m, const int* n, float* a, const int* lda, float* tau, float* work,
             const int* lwork, int* info);
void sorgrq_(const int* m, const int* n, const int* k, float* a, const int* lda,
             const float* tau, float* work, const int* lwork, int* info);
void sorglq_(const int* m, const int* n, const int* k, float* a, const int* lda,
             const float* tau, float* work, const int* lwork, int* info);
void sorgql_(const int* m, const int* n, const int* k, float* a, const int* lda,
             const float* tau, float* work, const int* lwork, int* info);
void dgerqf_(const int* m, const int* n, double* a, const int* lda, double* tau, double* work,
             const int* lwork, int* info);
void dgelqf_(const int* m, const int* n, double* a, const int* lda, double* tau, double* work,
             const int* lwork, int* info);
void dgeqlf_(const int* m, const int* n, double* a, const int* lda, double* tau, double* work,
             const int* lwork, int* info);
void dorgrq_(const int* m, const int* n, const int* k, double* a, const int* lda,
             const double* tau, double* work, const int* lwork, int* info);
void dorglq_(const int* m, const int* n, const int* k, double* a, const int* lda,
             const double* tau, double* work, const int* lwork, int* info);
void dorgql_(const int* m, const int* n, const int* k, double* a, const int* lda,
             const double* tau, double* work, const int* lwork, int* info);
void zgerqf_(const int* m, const int* n, void* a, const int* lda, void* tau, void* work,
             const int* lwork, int* info);
void zgelqf_(const int* m, const int* n, void* a, const int* lda, void* tau, void* work,
             const int* lwork, int* info);
void zgeqlf_(const int* m, const int* n, void* a, const int* lda, void* tau, void* work,
             const int* lwork, int* info);
void zungrq_(const int* m, const int* n, const int* k, void* a, const int* lda,
             const void* tau, void* work, const int* lwork, int* info);
void zunglq_(const int* m, const int* n, const int* k, void* a, const int* lda,
             const void* tau, void* work, const int* lwork, int* info);
void zungql_(const int* m, const int* n, const int* k, void* a, const int* lda,
             const void* tau, void* work, const int* lwork, int* info);
/* QR with column pivoting (src/linalg/qr.cr:38-41; rank(method: QRP), src/linalg/linalg.cr:772-774)
 * ?geqp3: :1188, :620, :1730 (the complex form carries rwork) */
void sgeqp3_(const int* m, const int* n, float* a, const int* lda, int* jpvt, float* tau,
             float* work, const int* lwork, int* info);
void dgeqp3_(const int* m, const int* n, double* a, const int* lda, int* jpvt, double* tau,
             double* work, const int* lwork, int* info);
void zgeqp3_(const int* m, const int* n, void* a, const int* lda, int* jpvt, void* tau, void* work,
             const int* lwork, double* rwork, int* info);
/* "next" row 4: norms (src/linalg/linalg.cr:721-748) and the copy / fill helpers
 * ?lange: :1335, :767, :1959   ?lansy: :1342, :774, :1969   zlanhe: :1962
 * ?lantr: :1345, :777, :1972   ?lacpy: :1288, :720, :1929   ?laset: :1421, :853, :2019
 * (REAL functions return float, FloatReturn = Float, libLAPACK_linux.cr:17; `work` is unused) */
float slange_(const char* norm, const int* m, const int* n, const float* a, const int* lda, float* work);
double dlange_(const char* norm, const int* m, const int* n, const double* a, const int* lda, double* work);
double zlange_(const char* norm, const int* m, const int* n, const void* a, const int* lda, double* work);
float slansy_(const char* norm, const char* uplo, const int* n, const float* a, const int* lda, float* work);
double dlansy_(const char* norm, const char* uplo, const int* n, const double* a, const int* lda, double* work);
double zlansy_(const char* norm, const char* uplo, const int* n, const void* a, const int* lda, double* work);
double zlanhe_(const char* norm, const char* uplo, const int* n, const void* a, const int* lda, double* work);
float slantr_(const char* norm, const char* uplo, const char* diag, const int* m, const int* n,
              const float* a, const int* lda, float* work);
double dlantr_(const char* norm, const char* uplo, const char* diag, const int* m, const int* n,
               const double* a, const int* lda, double* work);
double zlantr_(const char* norm, const char* uplo, const char* diag, const int* m, const int* n,
               const void* a, const int* lda, double* work);
void slacpy_(const char* uplo, const int* m, const int* n, const float* a, const int* lda, float* b, const int* ldb);
void dlacpy_(const char* uplo, const int* m, const int* n, const double* a, const int* lda, double* b, const int* ldb);
void zlacpy_(const char* uplo, const int* m, const int* n, const void* a, const int* lda, void* b, const int* ldb);
void slaset_(const char* uplo, const int* m, const int* n, const float* alpha, const float* beta, float* a, const int* lda);
void dlaset_(const char* uplo, const int* m, const int* n, const double* alpha, const double* beta, double* a, const int* lda);
void zlaset_(const char* uplo, const int* m, const int* n, const void* alpha, const void* beta, void* a, const int* lda);

/* ===================== Part 3: device-resident API ======================================
 * Same mathematics as the symbols above, but every matrix/ipiv/tau/info pointer is a DEVICE
 * pointer, dimensions are 64-bit (N = 65536 overflows the reference's Int32 element count,
 * src/matrix/general_matrix.cr:26), and the work is only ENQUEUED on `stream` (a
 * cudaStream_t; NULL = legacy default stream).  `info` is a device int written by the
 * kernels.  Scratch memory comes from a per-thread, per-device arena that grows on first use
 * and is reused by that thread's next call: a call on another stream than the thread's previous
 * call is ordered behind it with an event (calls on ONE stream, the normal case, are ordered
 * anyway); different host threads have separate scratch and run concurrently.
 * Complex scalars are passed as pointers to {re, im}.  Return value: 0, or -i if argument i
 * is invalid (nothing is enqueued).  `X` in the names below is s, d or z.
 */
int la_b200_sgemm_dev(int transa, int transb, int64_t m, int64_t n, int64_t k, float alpha,
                      const float* a, int64_t lda, const float* b, int64_t ldb, float beta,
                      float* c, int64_t ldc, void* stream);
int la_b200_dgemm_dev(int transa, int transb, int64_t m, int64_t n, int64_t k, double alpha,
                      const double* a, int64_t lda, const double* b, int64_t ldb, double beta,
                      double* c, int64_t ldc, void* stream);
int la_b200_zgemm_dev(int transa, int transb, int64_t m, int64_t n, int64_t k, const void* alpha,
                      const void* a, int64_t lda, const void* b, int64_t ldb, const void* beta,
                      void* c, int64_t ldc, void* stream);
/* C <- alpha*op(A)*op(B) + beta*C restricted to a block staircase: for the multi-GPU
 * factorisations (no counterpart in the reference, which has no distributed path).
 * C is the LOCAL part of a block-cyclic matrix: its rows [i*nb, (i+1)*nb) are global block
 * row_first + i*row_step, its columns [j*nb, (j+1)*nb) global block col_first + j*col_step;
 * only the blocks whose global row block >= global column block are computed (the trailing
 * update of a lower Cholesky on a process grid).  Extra error codes: -14 .. -18 for nb,
 * row_first, row_step, col_first, col_step. */
int la_b200_sgemm_stair_dev(int transa, int transb, int64_t m, int64_t n, int64_t k, float alpha,
                            const float* a, int64_t lda, const float* b, int64_t ldb, float beta,
                            float* c, int64_t ldc, int64_t nb, int64_t row_first, int64_t row_step,
                            int64_t col_first, int64_t col_step, void* stream);
int la_b200_dgemm_stair_dev(int transa, int transb, int64_t m, int64_t n, int64_t k, double alpha,
                            const double* a, int64_t lda, const double* b, int64_t ldb,
                            double beta, double* c, int64_t ldc, int64_t nb, int64_t row_first,
                            int64_t row_step, int64_t col_first, int64_t col_step, void* stream);
int la_b200_zgemm_stair_dev(int transa, int transb, int64_t m, int64_t n, int64_t k,
                            const void* alpha, const void* a, int64_t lda, const void* b,
                            int64_t ldb, const void* beta, void* c, int64_t ldc, int64_t nb,
                            int64_t row_first, int64_t row_step, int64_t col_first,
                            int64_t col_step, void* stream);
/* The calling thread's later device-API calls run their persistent kernels on at most `sms`
 * SMs (0 = all), leaving the rest to kernels of other streams (communication beside a long
 * product); returns the previous limit. */
int la_b200_set_sm_limit(int sms);
#define LA_B200_DECL_DEV(X, T)                                                                    \
  int la_b200_##X##trsm_dev(int side, int uplo, int transa, int diag, int64_t m, int64_t n,       \
                            const void* alpha, const T* a, int64_t lda, T* b, int64_t ldb,        \
                            void* stream);                                                        \
  int la_b200_##X##herk_dev(int uplo, int trans, int64_t n, int64_t k, double alpha, const T* a,  \
                            int64_t lda, double beta, T* c, int64_t ldc, void* stream);           \
  int la_b200_##X##getrf_dev(int64_t m, int64_t n, T* a, int64_t lda, int* ipiv, int* info,       \
                             void* stream);                                                       \
  /* rows k (k1 <= k < k2, 0-based) of the m x n matrix swapped with rows ipiv[k]-1, in that  \
     order: the ?laswp step inside ?getrf/?getrs, exported for the multi-GPU LU */               \
  int la_b200_##X##laswp_dev(int64_t m, int64_t n, T* a, int64_t lda, int64_t k1, int64_t k2,     \
                             const int* ipiv, void* stream);                                      \
  /* ?lange / ?lacpy / ?laset on resident matrices: norm = 'M', 'O', 'I' or 'F'; the norm is    \
     written to the DEVICE double `result`; uplo = 'U', 'L' or anything else for the whole  \
     matrix; alpha / beta by pointer */                                                         \
  int la_b200_##X##lange_dev(int norm, int64_t m, int64_t n, const T* a, int64_t lda,             \
                             double* result, void* stream);                                       \
  int la_b200_##X##lacpy_dev(int uplo, int64_t m, int64_t n, const T* a, int64_t lda, T* b,       \
                             int64_t ldb, void* stream);                                          \
  int la_b200_##X##laset_dev(int uplo, int64_t m, int64_t n, const void* alpha, const void* beta, \
                             T* a, int64_t lda, void* stream);                                    \
  /* the element-wise steps between the products of a resident chain (the reference's          \
     Matrix#add!, scalar multiples and map(&.abs): src/linalg/matrix.cr, expm.cr:112-166):     \
     B <- alpha A + beta B (B is not read when beta == 0; alpha / beta by pointer), and          \
     B <- |A| (modulus, zero imaginary part for complex) */                                      \
  int la_b200_##X##geadd_dev(int64_t m, int64_t n, const void* alpha, const T* a, int64_t lda,    \
                             const void* beta, T* b, int64_t ldb, void* stream);                  \
  int la_b200_##X##geabs_dev(int64_t m, int64_t n, const T* a, int64_t lda, T* b, int64_t ldb,    \
                             void* stream);                                                       \
  int la_b200_##X##getrs_dev(int trans, int64_t n, int64_t nrhs, const T* a, int64_t lda,         \
                             const int* ipiv, T* b, int64_t ldb, void* stream);                   \
  int la_b200_##X##potrf_dev(int uplo, int64_t n, T* a, int64_t lda, int* info, void* stream);    \
  int la_b200_##X##potrs_dev(int uplo, int64_t n, int64_t nrhs, const T* a, int64_t lda, T* b,    \
                             int64_t ldb, void* stream);                                          \
  int la_b200_##X##geqrf_dev(int64_t m, int64_t n, T* a, int64_t lda, T* tau, void* stream);      \
  int la_b200_##X##ormqr_dev(int side, int trans, int64_t m, int64_t n, int64_t k, const T* a,    \
                             int64_t lda, const T* tau, T* c, int64_t ldc, void* stream);         \
  int la_b200_##X##gels_dev(int trans, int64_t m, int64_t n, int64_t nrhs, T* a, int64_t lda,     \
                            T* b, int64_t ldb, int* info, void* stream);                          \
  int la_b200_##X##trtrs_dev(int uplo, int trans, int diag, int64_t n, int64_t nrhs, const T* a,  \
                             int64_t lda, T* b, int64_t ldb, int* info, void* stream);
LA_B200_DECL_DEV(s, float)
LA_B200_DECL_DEV(d, double)
LA_B200_DECL_DEV(z, void)
#undef LA_B200_DECL_DEV

/* No counterpart in the reference.  FP64 LU factorisation (same arguments, results and pivot
 * vector as la_b200_dgetrf_dev) for hosts that must not allocate GPU memory: every byte of
 * scratch comes from one 384 KiB array inside the library image, which every GPU that loads
 * the library holds until exit.  m, n <= 32768, else -100 is returned (message on stderr) and
 * nothing is enqueued.  The first call on a device does a one-time set-up and should precede
 * any stream capture.  All calls share the array: order them on one stream; calls that can
 * overlap on the GPU race.  -101: set-up failed.  -102: a scratch request did not fit the
 * array (nothing is allocated in that case either). */
int la_b200_dgetrf_dev_noalloc(int64_t m, int64_t n, double* a, int64_t lda, int* ipiv, int* info,
                               void* stream);

/* Device-matrix handles (no counterpart in the reference): a host that runs
 * a CHAIN of products and solves — `**` (src/linalg/power.cr:19-30), expm-style code
 * (src/linalg/expm.cr:37-111) — creates its operands once, uploads once, calls the *_dev entry
 * points on la_b200_mat_ptr / la_b200_mat_ld, and downloads the result once, instead of paying
 * an upload and a download per call of the drop-in symbols.  type = 's', 'd' or 'z'; the
 * leading dimension is the row count rounded up to 16.  upload / download are synchronous;
 * NULL on allocation failure; -1 / -3: bad handle or pointer / leading dimension. */
void* la_b200_mat_create(int type, int64_t rows, int64_t cols);
void la_b200_mat_free(void* handle);
void* la_b200_mat_ptr(void* handle);
int64_t la_b200_mat_ld(void* handle);
int64_t la_b200_mat_rows(void* handle);
int64_t la_b200_mat_cols(void* handle);
int la_b200_mat_upload(void* handle, const void* host, int64_t ldh);
int la_b200_mat_download(void* handle, void* host, int64_t ldh);

/* library identification / diagnostics */
const char* la_b200_version(void);
/* bytes of device workspace currently held by the calling thread's arena */
int64_t la_b200_workspace_bytes(void);
/* 1 if the last f64/z/f32 GEMM issued by this thread used the tensor-core path */
int la_b200_last_gemm_path(void);
/* Give the device workspaces back to the driver: the calling thread's device-API arenas and
 * those of every idle context of the host-pointer symbols (the pool otherwise keeps each
 * arena at its high-water mark, the analogue of the reference's WORK_POOL,
 * src/linalg/workpool.cr:73).  Synchronises the devices concerned; returns the bytes freed. */
int64_t la_b200_release_workspace(void);

#ifdef __cplusplus
}
#endif
#endif /* LA_B200_H */
