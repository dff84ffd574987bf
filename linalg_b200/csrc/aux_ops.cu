// la_b200 — the O(N^2) helpers around the hot path:
//   * matrix norms ?lange_ / ?lansy_ / zlanhe_ / ?lantr_ (src/linalg/linalg.cr:721-748), one pass
//     over the matrix at HBM bandwidth (host pointers: staged upload first), plus a device entry
//     point for matrices that are already resident;
//   * ?lacpy_ / ?laset_ on HOST matrices (the copies and fills the reference does with
//     per-element Crystal blocks: to_general / clone / zeros / tril! / triu!,
//     src/matrix/matrix.cr:107-111,650-663, src/matrix/general_matrix.cr:24-39): pure host-memory
//     traffic, done by a few host threads — a trip over PCIe would only add to it — and their
//     device forms for resident matrices;
//   * device-matrix handles, so that a chain of products and solves (`**`, expm-style code,
//     src/linalg/power.cr:19-30, src/linalg/expm.cr:37-111) uploads its operands once and
//     downloads its result once instead of paying both per call.
#include "shim_common.h"
#include <thread>
#include <vector>

namespace la {

namespace {

template <typename T>
__device__ __forceinline__ double abs2(T v) {
  return (double)real_(v) * (double)real_(v) + (double)imag_(v) * (double)imag_(v);
}
template <typename T>
__device__ __forceinline__ double absv(T v) {
  if constexpr (Traits<T>::is_complex) return sqrt(abs2(v));
  else return fabs((double)v);
}

// logical element (i, j) of the matrix the norm is taken of
template <typename T>
__device__ __forceinline__ T logical_elem(int shape, int lower, int unit, int64_t i, int64_t j,
                                          const T* A, int64_t lda) {
  if (shape == ShapeGeneral) return A[i + j * lda];
  const bool stored = lower ? i >= j : i <= j;
  if (shape == ShapeTri) {
    if (!stored) return zero<T>();
    if (unit && i == j) return one<T>();
    return A[i + j * lda];
  }
  T v = stored ? A[i + j * lda] : A[j + i * lda];
  if (shape == ShapeHerm) {
    if (i == j) v = from_real<T>((double)real_(v));
    else if (!stored) v = conj_(v);
  }
  return v;
}

// per column j: sum |a|, max |a|, sum |a|^2 over the rows  (grid-stride over columns)
template <typename T>
__global__ void __launch_bounds__(256)
norm_cols_kernel(int shape, int lower, int unit, int64_t m, int64_t n, const T* __restrict__ A,
                 int64_t lda, double* __restrict__ csum, double* __restrict__ cmax,
                 double* __restrict__ cssq) {
  __shared__ double r1[8], r2[8], r3[8];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int64_t j = blockIdx.x; j < n; j += gridDim.x) {
    double s = 0.0, mx = 0.0, q = 0.0;
    bool nan = false;
    for (int64_t i = tid; i < m; i += 256) {
      const double a = absv<T>(logical_elem<T>(shape, lower, unit, i, j, A, lda));
      s += a;
      q += a * a;
      if (a != a) nan = true;
      mx = a > mx ? a : mx;
    }
    if (nan) mx = s;                                 // NaN propagates (as ?lange does)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      s += __shfl_xor_sync(0xffffffffu, s, o);
      q += __shfl_xor_sync(0xffffffffu, q, o);
      const double om = __shfl_xor_sync(0xffffffffu, mx, o);
      mx = (om > mx || om != om) ? om : mx;
    }
    if (lane == 0) { r1[warp] = s; r2[warp] = mx; r3[warp] = q; }
    __syncthreads();
    if (tid == 0) {
      double ts = 0.0, tm = 0.0, tq = 0.0;
      for (int w = 0; w < 8; w++) {
        ts += r1[w];
        tq += r3[w];
        tm = (r2[w] > tm || r2[w] != r2[w]) ? r2[w] : tm;
      }
      csum[j] = ts; cmax[j] = tm; cssq[j] = tq;
    }
    __syncthreads();
  }
}

// per row i: sum over the columns of |a| (one thread per row: coalesced in column-major storage)
template <typename T>
__global__ void norm_rows_kernel(int shape, int lower, int unit, int64_t m, int64_t n,
                                 const T* __restrict__ A, int64_t lda, double* __restrict__ rsum) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  double s = 0.0;
  for (int64_t j = 0; j < n; j++) s += absv<T>(logical_elem<T>(shape, lower, unit, i, j, A, lda));
  rsum[i] = s;
}

// final reduction by one CTA: kind Max -> max of cmax, One -> max of csum, Inf -> max of rsum,
// Fro -> sqrt(sum cssq)
__global__ void __launch_bounds__(256)
norm_final_kernel(int kind, int64_t cnt, const double* __restrict__ v, double* __restrict__ out) {
  __shared__ double red[8];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  double acc = 0.0;
  for (int64_t i = tid; i < cnt; i += 256) {
    const double x = v[i];
    if (kind == NormFro) acc += x;
    else acc = (x > acc || x != x) ? x : acc;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double ov = __shfl_xor_sync(0xffffffffu, acc, o);
    if (kind == NormFro) acc += ov;
    else acc = (ov > acc || ov != ov) ? ov : acc;
  }
  if (lane == 0) red[warp] = acc;
  __syncthreads();
  if (tid == 0) {
    double t = 0.0;
    for (int w = 0; w < 8; w++) {
      if (kind == NormFro) t += red[w];
      else t = (red[w] > t || red[w] != red[w]) ? red[w] : t;
    }
    *out = kind == NormFro ? sqrt(t) : t;
  }
}

template <typename T>
__global__ void lacpy_kernel(int part, int64_t m, int64_t n, const T* __restrict__ A, int64_t lda,
                             T* __restrict__ B, int64_t ldb) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  for (int64_t j = blockIdx.y; j < n; j += gridDim.y)
    if (part == 0 || (part == 1 && i <= j) || (part == 2 && i >= j)) B[i + j * ldb] = A[i + j * lda];
}
// B <- alpha A + beta B (beta == 0: B is not read) and B <- |A|: the element-wise steps between
// the products of a resident chain (Matrix#add!, scalar multiples, map(&.abs) in the reference)
template <typename T>
__global__ void geadd_kernel(int64_t m, int64_t n, T alpha, const T* __restrict__ A, int64_t lda,
                             T beta, int beta_zero, T* __restrict__ B, int64_t ldb) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  for (int64_t j = blockIdx.y; j < n; j += gridDim.y) {
    const T x = mul(alpha, A[i + j * lda]);
    B[i + j * ldb] = beta_zero ? x : add(x, mul(beta, B[i + j * ldb]));
  }
}
template <typename T>
__global__ void geabs_kernel(int64_t m, int64_t n, const T* __restrict__ A, int64_t lda,
                             T* __restrict__ B, int64_t ldb) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  for (int64_t j = blockIdx.y; j < n; j += gridDim.y) {
    const T a = A[i + j * lda];
    double r;
    if constexpr (Traits<T>::is_complex) r = hypot(a.x, a.y);
    else r = fabs((double)a);
    B[i + j * ldb] = from_real<T>(r);
  }
}
template <typename T>
__global__ void laset_kernel(int part, int64_t m, int64_t n, T alpha, T beta, T* __restrict__ A,
                             int64_t lda) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  for (int64_t j = blockIdx.y; j < n; j += gridDim.y) {
    if (i == j) A[i + j * lda] = beta;
    else if (part == 0 || (part == 1 && i < j) || (part == 2 && i > j)) A[i + j * lda] = alpha;
  }
}

}  // namespace
int norm_kind_from_char(char c) {
  switch (c) {
    case 'M': case 'm': return NormMax;
    case 'O': case 'o': case '1': return NormOne;
    case 'I': case 'i': return NormInf;
    case 'F': case 'f': case 'E': case 'e': return NormFro;
    default: return -1;
  }
}
namespace {
int part_from_char(char c) { return (c == 'U' || c == 'u') ? 1 : ((c == 'L' || c == 'l') ? 2 : 0); }

// run fn(j0, j1) over column ranges on a few host threads (big matrices only)
template <typename F>
void host_parallel_columns(int64_t n, size_t bytes, F fn) {
  unsigned nt = std::thread::hardware_concurrency();
  nt = nt >= 16 ? 8 : (nt >= 4 ? 4 : 1);
  if (bytes < (size_t(8) << 20) || n < (int64_t)nt * 4) { fn(0, n); return; }
  std::vector<std::thread> ts;
  for (unsigned t = 0; t < nt; t++)
    ts.emplace_back([=] { fn(n * t / nt, n * (t + 1) / nt); });
  for (auto& t : ts) t.join();
}

}  // namespace

// norm of the logical matrix into *d_out (device double); enqueues only
template <typename T>
void matrix_norm(Context& ctx, int kind, int shape, Uplo uplo, bool unit, int64_t m, int64_t n,
                 const T* A, int64_t lda, double* d_out) {
  if (m <= 0 || n <= 0) {
    LA_CUDA_CHECK(cudaMemsetAsync(d_out, 0, sizeof(double), ctx.stream));
    return;
  }
  Arena::Mark mk = ctx.arena.mark();
  const int lower = uplo == Lower;
  if (kind == NormInf && shape != ShapeSym && shape != ShapeHerm) {
    double* rs = ctx.alloc<double>(m);
    norm_rows_kernel<T><<<(unsigned)ceil_div(m, 128), 128, 0, ctx.stream>>>(shape, lower, unit, m, n,
                                                                           A, lda, rs);
    norm_final_kernel<<<1, 256, 0, ctx.stream>>>(NormInf, m, rs, d_out);
  } else {
    double* cs = ctx.alloc<double>(3 * n);
    const int64_t g = n < (int64_t)ctx.sm_count * 8 ? n : (int64_t)ctx.sm_count * 8;
    norm_cols_kernel<T><<<(unsigned)g, 256, 0, ctx.stream>>>(shape, lower, unit, m, n, A, lda, cs,
                                                              cs + n, cs + 2 * n);
    const double* v = kind == NormMax ? cs + n : (kind == NormFro ? cs + 2 * n : cs);
    norm_final_kernel<<<1, 256, 0, ctx.stream>>>(kind == NormInf ? NormOne : kind, n, v, d_out);
  }
  LA_CUDA_CHECK(cudaGetLastError());
  ctx.arena.release(mk);
}

#define LA_INST_NORM(T)                                                                          \
  template void matrix_norm<T>(Context&, int, int, Uplo, bool, int64_t, int64_t, const T*,       \
                               int64_t, double*);
LA_INST_NORM(float)
LA_INST_NORM(double)
LA_INST_NORM(zcomplex)

template <typename T>
static double h_norm(const char* pnorm, int shape, const char* puplo, const char* pdiag, int64_t m,
                     int64_t n, const T* A, int64_t lda) {
  const int kind = norm_kind_from_char(*pnorm);
  if (kind < 0 || m <= 0 || n <= 0) return 0.0;
  const int uplo = puplo ? uplo_from_char(*puplo) : Upper;
  const bool unit = pdiag && (*pdiag == 'U' || *pdiag == 'u');
  ContextLease ctx;
  Arena::Mark mk = ctx->arena.mark();
  DevMat<T> dA = dev_from_host<T>(*ctx, m, n, A, lda);
  double* d_out = ctx->alloc<double>(1);
  matrix_norm<T>(*ctx, kind, shape, uplo ? (Uplo)uplo : Upper, unit, m, n, dA.p, dA.ld, d_out);
  double* h = (double*)pinned_scratch(*ctx, sizeof(double));
  LA_CUDA_CHECK(cudaMemcpyAsync(h, d_out, sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  sync(*ctx);
  const double r = *h;
  ctx->arena.release(mk);
  return r;
}

template <typename T>
static void h_lacpy(const char* puplo, const int* pm, const int* pn, const T* A, const int* plda,
                    T* B, const int* pldb) {
  const int64_t m = *pm, n = *pn, lda = *plda, ldb = *pldb;
  if (m <= 0 || n <= 0) return;
  const int part = part_from_char(*puplo);
  host_parallel_columns(n, (size_t)m * n * sizeof(T), [=](int64_t j0, int64_t j1) {
    for (int64_t j = j0; j < j1; j++) {
      const int64_t i0 = part == 2 ? (j < m ? j : m) : 0;
      const int64_t i1 = part == 1 ? ((j + 1) < m ? (j + 1) : m) : m;
      if (i1 > i0) memcpy(B + i0 + j * ldb, A + i0 + j * lda, (size_t)(i1 - i0) * sizeof(T));
    }
  });
}

template <typename T>
static void h_laset(const char* puplo, const int* pm, const int* pn, const T* palpha, const T* pbeta,
                    T* A, const int* plda) {
  const int64_t m = *pm, n = *pn, lda = *plda;
  if (m <= 0 || n <= 0) return;
  const int part = part_from_char(*puplo);
  const T alpha = *palpha, beta = *pbeta;
  host_parallel_columns(n, (size_t)m * n * sizeof(T), [=](int64_t j0, int64_t j1) {
    for (int64_t j = j0; j < j1; j++) {
      T* col = A + j * lda;
      const int64_t d = j < m ? j : m;               // rows above the diagonal in this column
      if (part != 2) for (int64_t i = 0; i < d; i++) col[i] = alpha;
      if (part != 1) for (int64_t i = j + 1; i < m; i++) col[i] = alpha;
      if (j < m) col[j] = beta;
    }
  });
}

// ------------------------------------------------------------------ device-matrix handles
struct MatHandle {
  int letter;          // 's', 'd', 'z'
  int device;
  int64_t rows, cols, ld;
  void* p;
};
static size_t elem_size(int letter) { return letter == 's' ? 4 : (letter == 'd' ? 8 : 16); }

}  // namespace la

using namespace la;

extern "C" {

float slange_(const char* norm, const int* m, const int* n, const float* a, const int* lda, float*) {
  return (float)h_norm<float>(norm, ShapeGeneral, nullptr, nullptr, *m, *n, a, *lda);
}
double dlange_(const char* norm, const int* m, const int* n, const double* a, const int* lda, double*) {
  return h_norm<double>(norm, ShapeGeneral, nullptr, nullptr, *m, *n, a, *lda);
}
double zlange_(const char* norm, const int* m, const int* n, const void* a, const int* lda, double*) {
  return h_norm<zcomplex>(norm, ShapeGeneral, nullptr, nullptr, *m, *n, (const zcomplex*)a, *lda);
}
float slansy_(const char* norm, const char* uplo, const int* n, const float* a, const int* lda, float*) {
  return (float)h_norm<float>(norm, ShapeSym, uplo, nullptr, *n, *n, a, *lda);
}
double dlansy_(const char* norm, const char* uplo, const int* n, const double* a, const int* lda, double*) {
  return h_norm<double>(norm, ShapeSym, uplo, nullptr, *n, *n, a, *lda);
}
double zlansy_(const char* norm, const char* uplo, const int* n, const void* a, const int* lda, double*) {
  return h_norm<zcomplex>(norm, ShapeSym, uplo, nullptr, *n, *n, (const zcomplex*)a, *lda);
}
double zlanhe_(const char* norm, const char* uplo, const int* n, const void* a, const int* lda, double*) {
  return h_norm<zcomplex>(norm, ShapeHerm, uplo, nullptr, *n, *n, (const zcomplex*)a, *lda);
}
float slantr_(const char* norm, const char* uplo, const char* diag, const int* m, const int* n,
              const float* a, const int* lda, float*) {
  return (float)h_norm<float>(norm, ShapeTri, uplo, diag, *m, *n, a, *lda);
}
double dlantr_(const char* norm, const char* uplo, const char* diag, const int* m, const int* n,
               const double* a, const int* lda, double*) {
  return h_norm<double>(norm, ShapeTri, uplo, diag, *m, *n, a, *lda);
}
double zlantr_(const char* norm, const char* uplo, const char* diag, const int* m, const int* n,
               const void* a, const int* lda, double*) {
  return h_norm<zcomplex>(norm, ShapeTri, uplo, diag, *m, *n, (const zcomplex*)a, *lda);
}

void slacpy_(const char* uplo, const int* m, const int* n, const float* a, const int* lda, float* b,
             const int* ldb) { h_lacpy<float>(uplo, m, n, a, lda, b, ldb); }
void dlacpy_(const char* uplo, const int* m, const int* n, const double* a, const int* lda,
             double* b, const int* ldb) { h_lacpy<double>(uplo, m, n, a, lda, b, ldb); }
void zlacpy_(const char* uplo, const int* m, const int* n, const void* a, const int* lda, void* b,
             const int* ldb) { h_lacpy<zcomplex>(uplo, m, n, (const zcomplex*)a, lda, (zcomplex*)b, ldb); }
void slaset_(const char* uplo, const int* m, const int* n, const float* alpha, const float* beta,
             float* a, const int* lda) { h_laset<float>(uplo, m, n, alpha, beta, a, lda); }
void dlaset_(const char* uplo, const int* m, const int* n, const double* alpha, const double* beta,
             double* a, const int* lda) { h_laset<double>(uplo, m, n, alpha, beta, a, lda); }
void zlaset_(const char* uplo, const int* m, const int* n, const void* alpha, const void* beta,
             void* a, const int* lda) {
  h_laset<zcomplex>(uplo, m, n, (const zcomplex*)alpha, (const zcomplex*)beta, (zcomplex*)a, lda);
}

// ---- device forms (resident matrices, caller's stream; the norm lands in a device double)
#define LA_AUX_DEV(X, T, VT)                                                                      \
  int la_b200_##X##lacpy_dev(int uplo, int64_t m, int64_t n, const VT* a, int64_t lda, VT* b,     \
                             int64_t ldb, void* stream) {                                         \
    if (m < 0) return -2;                                                                         \
    if (n < 0) return -3;                                                                         \
    if (lda < (m > 1 ? m : 1)) return -5;                                                         \
    if (ldb < (m > 1 ? m : 1)) return -7;                                                         \
    if (m == 0 || n == 0) return 0;                                                               \
    dim3 g((unsigned)ceil_div(m, (int64_t)256), (unsigned)(n < 4096 ? n : 4096));                 \
    lacpy_kernel<T><<<g, 256, 0, (cudaStream_t)stream>>>(part_from_char((char)uplo), m, n,        \
                                                         (const T*)a, lda, (T*)b, ldb);           \
    return cudaGetLastError() == cudaSuccess ? 0 : -100;                                          \
  }                                                                                               \
  int la_b200_##X##laset_dev(int uplo, int64_t m, int64_t n, const void* alpha, const void* beta, \
                             VT* a, int64_t lda, void* stream) {                                  \
    if (m < 0) return -2;                                                                         \
    if (n < 0) return -3;                                                                         \
    if (lda < (m > 1 ? m : 1)) return -7;                                                         \
    if (m == 0 || n == 0) return 0;                                                               \
    dim3 g((unsigned)ceil_div(m, (int64_t)256), (unsigned)(n < 4096 ? n : 4096));                 \
    laset_kernel<T><<<g, 256, 0, (cudaStream_t)stream>>>(part_from_char((char)uplo), m, n,        \
                                                         *(const T*)alpha, *(const T*)beta,       \
                                                         (T*)a, lda);                             \
    return cudaGetLastError() == cudaSuccess ? 0 : -100;                                          \
  }                                                                                               \
  int la_b200_##X##geadd_dev(int64_t m, int64_t n, const void* alpha, const VT* a, int64_t lda,   \
                             const void* beta, VT* b, int64_t ldb, void* stream) {                \
    if (m < 0) return -1;                                                                         \
    if (n < 0) return -2;                                                                         \
    if (lda < (m > 1 ? m : 1)) return -5;                                                         \
    if (ldb < (m > 1 ? m : 1)) return -8;                                                         \
    if (m == 0 || n == 0) return 0;                                                               \
    const T be = *(const T*)beta;                                                                 \
    dim3 g((unsigned)ceil_div(m, (int64_t)256), (unsigned)(n < 4096 ? n : 4096));                 \
    geadd_kernel<T><<<g, 256, 0, (cudaStream_t)stream>>>(m, n, *(const T*)alpha, (const T*)a,     \
                                                         lda, be, is_zero(be) ? 1 : 0, (T*)b, ldb); \
    return cudaGetLastError() == cudaSuccess ? 0 : -100;                                          \
  }                                                                                               \
  int la_b200_##X##geabs_dev(int64_t m, int64_t n, const VT* a, int64_t lda, VT* b, int64_t ldb,  \
                             void* stream) {                                                      \
    if (m < 0) return -1;                                                                         \
    if (n < 0) return -2;                                                                         \
    if (lda < (m > 1 ? m : 1)) return -4;                                                         \
    if (ldb < (m > 1 ? m : 1)) return -6;                                                         \
    if (m == 0 || n == 0) return 0;                                                               \
    dim3 g((unsigned)ceil_div(m, (int64_t)256), (unsigned)(n < 4096 ? n : 4096));                 \
    geabs_kernel<T><<<g, 256, 0, (cudaStream_t)stream>>>(m, n, (const T*)a, lda, (T*)b, ldb);     \
    return cudaGetLastError() == cudaSuccess ? 0 : -100;                                          \
  }
LA_AUX_DEV(s, float, float)
LA_AUX_DEV(d, double, double)
LA_AUX_DEV(z, zcomplex, void)

// ---- handles
void* la_b200_mat_create(int letter, int64_t rows, int64_t cols) {
  if ((letter != 's' && letter != 'd' && letter != 'z') || rows < 0 || cols < 0) return nullptr;
  MatHandle* h = new MatHandle();
  h->letter = letter;
  h->rows = rows;
  h->cols = cols;
  h->ld = round_up(rows > 0 ? rows : 1, 16);
  if (cudaGetDevice(&h->device) != cudaSuccess ||
      cudaMalloc(&h->p, (size_t)h->ld * (size_t)(cols > 0 ? cols : 1) * elem_size(letter)) != cudaSuccess) {
    cudaGetLastError();
    delete h;
    return nullptr;
  }
  return h;
}
void la_b200_mat_free(void* handle) {
  MatHandle* h = (MatHandle*)handle;
  if (!h) return;
  cudaFree(h->p);
  delete h;
}
void* la_b200_mat_ptr(void* handle) { return handle ? ((MatHandle*)handle)->p : nullptr; }
int64_t la_b200_mat_ld(void* handle) { return handle ? ((MatHandle*)handle)->ld : 0; }
int64_t la_b200_mat_rows(void* handle) { return handle ? ((MatHandle*)handle)->rows : 0; }
int64_t la_b200_mat_cols(void* handle) { return handle ? ((MatHandle*)handle)->cols : 0; }
// host (ldh) -> handle; synchronous (pageable sources go through the pinned staging ring)
int la_b200_mat_upload(void* handle, const void* host, int64_t ldh) {
  MatHandle* h = (MatHandle*)handle;
  if (!h || !host) return -1;
  if (ldh < (h->rows > 1 ? h->rows : 1)) return -3;
  if (h->rows == 0 || h->cols == 0) return 0;
  ContextLease ctx;
  const size_t es = elem_size(h->letter);
  h2d_2d(*ctx, h->p, (size_t)h->ld * es, host, (size_t)ldh * es, (size_t)h->rows * es, (size_t)h->cols,
         ctx->stream);
  sync(*ctx);
  return 0;
}
int la_b200_mat_download(void* handle, void* host, int64_t ldh) {
  MatHandle* h = (MatHandle*)handle;
  if (!h || !host) return -1;
  if (ldh < (h->rows > 1 ? h->rows : 1)) return -3;
  if (h->rows == 0 || h->cols == 0) return 0;
  ContextLease ctx;
  const size_t es = elem_size(h->letter);
  // the matrix may have been written on other streams of this device
  LA_CUDA_CHECK(cudaDeviceSynchronize());
  d2h_2d(*ctx, host, (size_t)ldh * es, h->p, (size_t)h->ld * es, (size_t)h->rows * es, (size_t)h->cols,
         ctx->stream);
  sync(*ctx);
  return 0;
}

}  // extern "C"
