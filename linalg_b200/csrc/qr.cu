// la_b200 — Householder QR (geqrf), application/generation of Q (ormqr/unmqr, orgqr/ungqr)
// and the least-squares driver gels.
//
// Reference call sites replaced: ?geqrf_ (src/linalg/qr.cr:42-46), ?gels_
// (src/linalg/linalg.cr:493-506, :541-543 -> geqrf + ormqr/unmqr + trtrs inside LAPACK),
// ?orgqr_/zungqr_ (src/linalg/qr.cr:107).  Conventions follow LAPACK 3.12 ?geqr2/?larfg/
// ?larft/?larfb: H(i) = I - tau_i v_i v_i^H, v_i(i) = 1, beta = -sign(Re alpha)*norm, R's
// diagonal is real; Q = H(1) H(2) ... H(k).
//
// Panel kernel: one cooperative launch per panel (<= 32 columns).  Rows are spread over the
// CTAs; each column costs ONE grid barrier: the partial sums  s_c = sum_r conj(a(r,j)) a(r,c)
// for c = j..jb-1 give both the column norm (c = j) and, after scaling, v^H A for the
// update, so the reflector and its application are derived redundantly by every CTA from
// the same reduced vector.  The CTA's row slice is cached in shared memory when it fits,
// otherwise streamed from global memory (tall-skinny panels).
#include "la_internal.h"
#include <cooperative_groups.h>

namespace la {

namespace {

constexpr int QR_NT = 256;
constexpr int QR_MAX_JB = 32;
constexpr int QR_MAX_CTAS = 640;
constexpr size_t QR_SMEM_BUDGET = 200 * 1024;

struct __align__(16) QrLL {
  unsigned long long data;
  unsigned long long tag;
};

template <typename T>
struct QrExchange {
  unsigned int barrier;
  unsigned int pad[63];
  T part[2][QR_MAX_CTAS][QR_MAX_JB];  // per-CTA partial sums s_c
  T rowj[2][QR_MAX_JB];               // row j of the panel
  // streamed panels that take two reflectors per pass: the sums against column j+1 and row j+1
  T part2[2][QR_MAX_CTAS][QR_MAX_JB];
  T rowj2[2][QR_MAX_JB];
  // shared-memory panels: the same sums and row j as 16-byte (payload, tag) packets that the
  // readers poll — no grid barrier between a column's sums and their reduction (two words per
  // value: complex); cleared once per factorisation, tags grow from launch to launch
  QrLL pll[2][QR_MAX_CTAS][QR_MAX_JB * 2];
  QrLL rowll[2][QR_MAX_JB * 2];
};

__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void grid_barrier(unsigned* counter, unsigned target) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(counter, 1u);
    while (ld_acquire_u32(counter) < target) { }
    __threadfence();
  }
  __syncthreads();
}

// tagged packets (the protocol of lu_panel.cu): a packet is valid when the upper half of its tag
// is the step's sequence number; one vector store, polled with volatile loads
__device__ __forceinline__ void qll_store(QrLL* p, unsigned long long data, unsigned seq) {
  asm volatile("st.global.cg.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(data),
               "l"((unsigned long long)seq << 32) : "memory");
}
__device__ __forceinline__ unsigned long long qll_wait(const QrLL* p, unsigned seq) {
  unsigned long long d, t;
  do {
    asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(d), "=l"(t) : "l"(p) : "memory");
  } while ((unsigned)(t >> 32) != seq);
  return d;
}
template <typename T> struct QllWords { static constexpr int N = 1; };
template <> struct QllWords<zcomplex> { static constexpr int N = 2; };
template <typename T> __device__ __forceinline__ void qll_put(QrLL* p, T v, unsigned seq);
template <> __device__ __forceinline__ void qll_put<double>(QrLL* p, double v, unsigned seq) {
  qll_store(p, (unsigned long long)__double_as_longlong(v), seq);
}
template <> __device__ __forceinline__ void qll_put<float>(QrLL* p, float v, unsigned seq) {
  qll_store(p, (unsigned long long)__float_as_uint(v), seq);
}
template <> __device__ __forceinline__ void qll_put<zcomplex>(QrLL* p, zcomplex v, unsigned seq) {
  qll_store(p, (unsigned long long)__double_as_longlong(v.x), seq);
  qll_store(p + 1, (unsigned long long)__double_as_longlong(v.y), seq);
}
template <typename T> __device__ __forceinline__ T qll_get(const QrLL* p, unsigned seq);
template <> __device__ __forceinline__ double qll_get<double>(const QrLL* p, unsigned seq) {
  return __longlong_as_double((long long)qll_wait(p, seq));
}
template <> __device__ __forceinline__ float qll_get<float>(const QrLL* p, unsigned seq) {
  return __uint_as_float((unsigned)qll_wait(p, seq));
}
template <> __device__ __forceinline__ zcomplex qll_get<zcomplex>(const QrLL* p, unsigned seq) {
  zcomplex v;
  v.x = __longlong_as_double((long long)qll_wait(p, seq));
  v.y = __longlong_as_double((long long)qll_wait(p + 1, seq));
  return v;
}

template <typename T> __device__ __forceinline__ void st_cg(T* p, T v) { __stcg(p, v); }
template <> __device__ __forceinline__ void st_cg<zcomplex>(zcomplex* p, zcomplex v) {
  __stcg(&p->x, v.x); __stcg(&p->y, v.y);
}
template <typename T> __device__ __forceinline__ T ld_cg(const T* p) { return __ldcg(p); }
template <> __device__ __forceinline__ zcomplex ld_cg<zcomplex>(const zcomplex* p) {
  zcomplex v; v.x = __ldcg(&p->x); v.y = __ldcg(&p->y); return v;
}

template <typename T>
__device__ __forceinline__ T warp_sum(T v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
template <>
__device__ __forceinline__ zcomplex warp_sum<zcomplex>(zcomplex v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    v.x += __shfl_xor_sync(0xffffffffu, v.x, o);
    v.y += __shfl_xor_sync(0xffffffffu, v.y, o);
  }
  return v;
}

// ?larfg from the reduced quantities: alpha = a(j,j), xnorm2 = sum |x|^2.
// Returns beta (new a(j,j)), tau and the scale applied to x.
template <typename T>
__device__ __forceinline__ void householder(T alpha, double xnorm2, T& beta_out, T& tau, T& scale) {
  if constexpr (Traits<T>::is_complex) {
    if (xnorm2 == 0.0 && alpha.y == 0.0) {
      tau = zero<T>(); scale = zero<T>(); beta_out = alpha;
      return;
    }
    double nrm = sqrt(alpha.x * alpha.x + alpha.y * alpha.y + xnorm2);
    double beta = -copysign(nrm, alpha.x);
    tau = zcomplex{(beta - alpha.x) / beta, -alpha.y / beta};
    scale = cdiv(zcomplex{1.0, 0.0}, zcomplex{alpha.x - beta, alpha.y});
    beta_out = zcomplex{beta, 0.0};
  } else {
    if (xnorm2 == 0.0) {
      tau = zero<T>(); scale = zero<T>(); beta_out = alpha;
      return;
    }
    double a = (double)alpha;
    double nrm = sqrt(a * a + xnorm2);
    double beta = -copysign(nrm, a);
    tau = (T)((beta - a) / beta);
    scale = (T)(1.0 / (a - beta));
    beta_out = (T)beta;
  }
}

#ifdef LA_PANEL_TRACE
__device__ long long g_qr_trace[64 * 8];
#define QTRACE(slot) do { if (blockIdx.x == 0 && threadIdx.x == 0 && j < 64) g_qr_trace[j * 8 + (slot)] = clock64(); } while (0)
#else
#define QTRACE(slot) do { } while (0)
#endif

// QR of an mp x jb panel (mp >= 1, jb <= 32), reflectors left in place, tau[0..kmax).
// SMEM: the CTA's row slice is held in shared memory for the whole launch (a compile-time switch,
// so that the panel is addressed with shared-memory loads and stores, not generic ones)
template <typename T, bool SMEM>
__global__ void __launch_bounds__(QR_NT)
qr_panel_kernel(int64_t mp, int jb, T* __restrict__ A, int64_t lda, T* __restrict__ tau,
                QrExchange<T>* xch, int64_t rpc, unsigned seq_base) {
  constexpr int use_smem = SMEM ? 1 : 0;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ T s_f[QR_MAX_JB];       // f_c = conj(tau) * w_c
  __shared__ T s_red[QR_NT / 32][QR_MAX_JB];
  __shared__ T s_scale;
  __shared__ T s_beta;

  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);   // warp-uniform for the compiler
  const int G = gridDim.x;
  const int64_t rs = (int64_t)blockIdx.x * rpc;
  int64_t nr64 = mp - rs; if (nr64 > rpc) nr64 = rpc; if (nr64 < 0) nr64 = 0;
  const int64_t nr = nr64;

  T* base;
  int64_t pitch;
  if (use_smem) {
    base = reinterpret_cast<T*>(smem_raw);
    pitch = rpc;
    // (eight columns per trip, all loads of a trip in flight before the first store: column by
    // column this loop was up to 64 dependent rounds of global-memory latency per launch)
    for (int64_t r = tid; r < nr; r += QR_NT)
      for (int c0 = 0; c0 < jb; c0 += 8) {
        T v[8];
#pragma unroll
        for (int i = 0; i < 8; i++) v[i] = c0 + i < jb ? A[rs + r + (int64_t)(c0 + i) * lda] : zero<T>();
#pragma unroll
        for (int i = 0; i < 8; i++)
          if (c0 + i < jb) base[(c0 + i) * pitch + r] = v[i];
      }
    __syncthreads();
  } else {
    base = A + rs;
    pitch = lda;
  }

  const int kmax = (int)(mp < jb ? mp : jb);
  constexpr int QW = QllWords<T>::N;
  for (int j = 0; j < kmax; j++) {
    const int par = j & 1;
    const unsigned seq = seq_base + (unsigned)j + 1u;
    QTRACE(0);
    // ---- partial sums s_c over owned rows r > j, one warp per column, published as packets
    // (a warp's up to four columns c0, c0+8, ... are summed together: the element of column j is
    // loaded once per row and the four chains of multiply-adds and their shuffle reductions
    // overlap — one column after the other this phase was 40 % of a step)
    {
      constexpr int NWQ = QR_NT / 32, CPW = QR_MAX_JB / NWQ;
      const int c0 = j + warp;
      T acc[CPW];
#pragma unroll
      for (int q = 0; q < CPW; q++) acc[q] = zero<T>();
      int64_t rbeg = j + 1 - rs; if (rbeg < 0) rbeg = 0;
#pragma unroll 4
      for (int64_t r = rbeg + lane; r < nr; r += 32) {
        const T x = conj_(base[j * pitch + r]);
#pragma unroll
        for (int q = 0; q < CPW; q++)
          if (c0 + q * NWQ < jb) acc[q] = fma_(x, base[(c0 + q * NWQ) * pitch + r], acc[q]);
      }
#pragma unroll
      for (int q = 0; q < CPW; q++) acc[q] = warp_sum<T>(acc[q]);
      if (lane == 0) {
#pragma unroll
        for (int q = 0; q < CPW; q++)
          if (c0 + q * NWQ < jb) qll_put<T>(&xch->pll[par][blockIdx.x][(c0 + q * NWQ) * QW], acc[q], seq);
      }
    }
    if (j >= rs && j < rs + nr && tid >= j && tid < jb)
      qll_put<T>(&xch->rowll[par][tid * QW], base[tid * pitch + (j - rs)], seq);
    QTRACE(1);
    // ---- reduce the partials of all CTAs as they arrive: 8 groups x 32 columns.  (A CTA can be
    // two columns ahead of no other: it publishes column j+1 only after it has read every CTA's
    // column j, so the two buffers of a parity never hold live data of different columns.)
    {
      const int c = tid & 31, grp = tid >> 5;
      T s = zero<T>();
      if (c >= j && c < jb)
        for (int g = grp; g < G; g += QR_NT / 32) s = add(s, qll_get<T>(&xch->pll[par][g][c * QW], seq));
      s_red[grp][c] = s;
    }
    QTRACE(2);
    __syncthreads();
    QTRACE(3);
    if (tid < 32) {
      const int c = tid;
      T s = zero<T>();
      if (c >= j && c < jb)
        for (int g = 0; g < QR_NT / 32; g++) s = add(s, s_red[g][c]);
      // every lane needs alpha, s_j
      T rj = (c >= j && c < jb) ? qll_get<T>(&xch->rowll[par][c * QW], seq) : zero<T>();
      T alpha, sj;
      if constexpr (Traits<T>::is_complex) {
        alpha.x = __shfl_sync(0xffffffffu, rj.x, j); alpha.y = __shfl_sync(0xffffffffu, rj.y, j);
        sj.x = __shfl_sync(0xffffffffu, s.x, j); sj.y = 0.0;
      } else {
        alpha = __shfl_sync(0xffffffffu, rj, j);
        sj = __shfl_sync(0xffffffffu, s, j);
      }
      T beta, tj, scale;
      householder<T>(alpha, (double)real_(sj), beta, tj, scale);
      // w_c = a(j,c) + conj(scale) * s_c ;  f_c = conj(tau) * w_c
      T w = fma_(conj_(scale), s, rj);
      s_f[c] = mul(conj_(tj), w);
      if (c == j) {
        s_scale = scale;
        s_beta = beta;
        if (blockIdx.x == 0) tau[j] = tj;
      }
    }
    __syncthreads();
    QTRACE(4);
    // ---- apply: rows r > j: v_r = x_r*scale, a(r,c) -= v_r f_c; row j: a(j,j)=beta, a(j,c) -= f_c
    {
      const T scale = s_scale;
      const bool identity = is_zero(scale) && is_zero(s_f[j]);  // tau == 0: H = I
      if (!identity || true) {
        int64_t rbeg = j + 1 - rs; if (rbeg < 0) rbeg = 0;
        for (int64_t r = rbeg + tid; r < nr; r += QR_NT) {
          T v = mul(base[j * pitch + r], scale);
          base[j * pitch + r] = v;
          // (eight independent load / multiply-add / store chains in flight)
#pragma unroll 8
          for (int c = j + 1; c < jb; c++)
            base[c * pitch + r] = fnma_(v, s_f[c], base[c * pitch + r]);
        }
        if (j >= rs && j < rs + nr) {
          if (tid == j) base[j * pitch + (j - rs)] = s_beta;
          else if (tid > j && tid < jb)
            base[tid * pitch + (j - rs)] = sub(base[tid * pitch + (j - rs)], s_f[tid]);
        }
      }
    }
    __syncthreads();
    QTRACE(5);
  }
  if (use_smem) {
    for (int64_t r = tid; r < nr; r += QR_NT)
      for (int c0 = 0; c0 < jb; c0 += 8) {
#pragma unroll
        for (int i = 0; i < 8; i++)
          if (c0 + i < jb) A[rs + r + (int64_t)(c0 + i) * lda] = base[(c0 + i) * pitch + r];
      }
  }
}

// One pass of the streaming panel kernel over the rows of a CTA (see qr_panel_stream_kernel):
// apply reflectors ja0 .. ja0+na-1, store, accumulate the sums against columns js = ja0+na and
// js+1, publish rows js and js+1.  With JA0 >= 0 the pass is compiled for that (ja0, na): every
// column index is then a constant and the selects / predicates of the generic form (which cost
// several times the useful instructions) fold away.
template <typename T, int JB, int U, bool PAIR, int JA0, int NA>
__device__ __forceinline__ void qr_stream_sweep(
    T* __restrict__ base, int64_t lda, int64_t nr, int64_t rs, int jb, int kmax, int tid, int ja0_rt,
    int na_rt, const T (*s_f)[QR_MAX_JB], const T (&scale)[2], const T (&betav)[2],
    T* __restrict__ rowj_out, T* __restrict__ rowj2_out, T (&acc)[JB], T (&acc2)[PAIR ? JB : 1]) {
  const int ja0 = JA0 >= 0 ? JA0 : ja0_rt;
  const int na = JA0 >= 0 ? NA : na_rt;
  const int js = ja0 + na;
  const int js2 = js + 1;
  const bool two = PAIR && js2 < kmax;
  for (int64_t r0 = tid; r0 < nr; r0 += (int64_t)U * QR_NT) {
    T a[U][JB];
    bool live[U];
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int64_t r = r0 + (int64_t)u * QR_NT;
      live[u] = r < nr && rs + r >= ja0;        // rows above ja0 are finished rows of R
#pragma unroll
      for (int c = 0; c < JB; c++)
        a[u][c] = (live[u] && c >= ja0 && c < jb) ? base[(int64_t)c * lda + r] : zero<T>();
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      if (!live[u]) continue;
      const int64_t r = r0 + (int64_t)u * QR_NT;
      const int64_t gr = rs + r;
#pragma unroll
      for (int i = 0; i < (PAIR ? 2 : 1); i++) {
        if (i >= na) continue;
        const int q = ja0 + i;
        if (gr == q) {                            // the reflector's own row: v = 1
#pragma unroll
          for (int c = 0; c < JB; c++) {
            const T fc = s_f[i][c];
            if (c == q) a[u][c] = betav[i];
            else if (c > q && c < jb) a[u][c] = sub(a[u][c], fc);
          }
        } else if (gr > q) {                      // v = x * scale
          T v = zero<T>();
#pragma unroll
          for (int c = 0; c < JB; c++)
            if (c == q) { v = mul(a[u][c], scale[i]); a[u][c] = v; }
#pragma unroll
          for (int c = 0; c < JB; c++) {
            if (c > q && c < jb) a[u][c] = fnma_(v, s_f[i][c], a[u][c]);
          }
        }
      }
      if (na > 0) {
#pragma unroll
        for (int c = 0; c < JB; c++)
          if (c >= ja0 && c < jb) base[(int64_t)c * lda + r] = a[u][c];
      }
      if (js < kmax) {
        if (gr == js) {                           // next reflector's own row: publish it
#pragma unroll
          for (int c = 0; c < JB; c++)
            if (c >= js && c < jb) st_cg<T>(&rowj_out[c], a[u][c]);
        } else if (gr > js) {
          T x = zero<T>(), x2 = zero<T>();
#pragma unroll
          for (int c = 0; c < JB; c++) {
            if (c == js) x = conj_(a[u][c]);
            if (PAIR && c == js2) x2 = conj_(a[u][c]);
          }
#pragma unroll
          for (int c = 0; c < JB; c++)
            if (c >= js && c < jb) acc[c] = fma_(x, a[u][c], acc[c]);
          if (PAIR && two) {
#pragma unroll
            for (int c = 0; c < JB; c++)
              if (c >= js2 && c < jb) acc2[PAIR ? c : 0] = fma_(x2, a[u][c], acc2[PAIR ? c : 0]);
            if (gr == js2) {                      // the row of the reflector after next
#pragma unroll
              for (int c = 0; c < JB; c++)
                if (c >= js && c < jb) st_cg<T>(&rowj2_out[c], a[u][c]);
            }
          }
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Tall panels (the row slice of a CTA does not fit shared memory, e.g. 2^20 x 8): streaming
// variant.  Same mathematics and the same one-grid-barrier-per-column exchange, but the
// application of reflector j and the partial sums that define reflector j+1 are FUSED into a
// single pass over the CTA's rows: each row is read once and written once per column.  A thread
// keeps U rows of the (at most JB-column) panel in registers with all their loads in flight, and
// several CTAs share an SM, so that a pass runs at memory bandwidth.  The HBM/L2 traffic of a
// panel is 2 * (jb + 1) * jb / 2 * rows * sizeof(T) bytes, quadratic in its width: geqrf therefore
// factors tall matrices in 8-column panels inside a recursive block structure.
template <typename T, int JB, int U>
__global__ void __launch_bounds__(QR_NT, ((JB <= 8 && !Traits<T>::is_complex) ? 2 : 1))
qr_panel_stream_kernel(int64_t mp, int jb, T* __restrict__ A, int64_t lda, T* __restrict__ tau,
                       QrExchange<T>* xch, int64_t rpc, T* __restrict__ Vc, int64_t ldv,
                       T* __restrict__ Tout, int64_t ldt) {
  // Two reflectors per pass where the panel is narrow (JB <= 8): a pass applies reflectors
  // ja0, ja0+1 and accumulates the sums against BOTH of the next two columns; the second
  // reflector's sums after the first one has been applied follow algebraically (the reflector
  // is a rank-1 change of the columns), so a panel of 8 needs 5 passes and grid barriers, not 9.
  // The derived norm is a difference: when it cancels badly the second reflector is left to a
  // pass of its own.
  constexpr bool PAIR = JB <= 8 && !Traits<T>::is_complex;   // (complex: the registers do not allow it)
  __shared__ T s_f[2][QR_MAX_JB];
  __shared__ T s_tau[QR_MAX_JB];
  __shared__ T s_red[2][QR_NT / 32][QR_MAX_JB];
  __shared__ T s_scale[2], s_beta[2];
  __shared__ int s_napply;
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);   // warp-uniform for the compiler
  const int G = gridDim.x;
  const int64_t rs = (int64_t)blockIdx.x * rpc;
  int64_t nr = mp - rs; if (nr > rpc) nr = rpc; if (nr < 0) nr = 0;
  T* base = A + rs;
  const int kmax = (int)(mp < jb ? mp : jb);
  unsigned target = 0;

  // pass: apply reflectors ja0 .. ja0+na-1 (scale/f/beta in shared memory) to the owned rows and
  // accumulate, for js = ja0+na (and js2 = js+1), the sums s_c = sum_{r > js} conj(a(r,js)) a(r,c)
  // (s2_c = sum_{r > js} conj(a(r,js2)) a(r,c)); rows js and js2 are published as they are.
  int ja0 = 0, na = 0, pass = 0;
  while (true) {
    const int js = ja0 + na;
    const int js2 = js + 1;
    const bool two = PAIR && js2 < kmax;
    T acc[JB], acc2[PAIR ? JB : 1];
#pragma unroll
    for (int c = 0; c < JB; c++) acc[c] = zero<T>();
#pragma unroll
    for (int c = 0; c < (PAIR ? JB : 1); c++) acc2[c] = zero<T>();
    T scale[2], betav[2];
    scale[0] = na > 0 ? s_scale[0] : zero<T>();
    scale[1] = na > 1 ? s_scale[1] : zero<T>();
    betav[0] = na > 0 ? s_beta[0] : zero<T>();
    betav[1] = na > 1 ? s_beta[1] : zero<T>();
    T* rowj_out = &xch->rowj[pass & 1][0];
    T* rowj2_out = &xch->rowj2[pass & 1][0];
    if constexpr (JB <= 8) {
      // one compiled pass per (first reflector, number of reflectors)
#define LA_QR_SWEEP(J, N)                                                                         \
  case (J) * 3 + (N):                                                                             \
    qr_stream_sweep<T, JB, U, PAIR, (J) + (N) <= JB ? (J) : 0, (J) + (N) <= JB ? (N) : 0>(         \
        base, lda, nr, rs, jb, kmax, tid, ja0, na, s_f, scale, betav, rowj_out, rowj2_out, acc,    \
        acc2);                                                                                    \
    break;
      switch (ja0 * 3 + na) {
        LA_QR_SWEEP(0, 0)
        LA_QR_SWEEP(0, 1) LA_QR_SWEEP(1, 1) LA_QR_SWEEP(2, 1) LA_QR_SWEEP(3, 1)
        LA_QR_SWEEP(4, 1) LA_QR_SWEEP(5, 1) LA_QR_SWEEP(6, 1) LA_QR_SWEEP(7, 1)
        LA_QR_SWEEP(0, 2) LA_QR_SWEEP(1, 2) LA_QR_SWEEP(2, 2) LA_QR_SWEEP(3, 2)
        LA_QR_SWEEP(4, 2) LA_QR_SWEEP(5, 2) LA_QR_SWEEP(6, 2)
        default:
          qr_stream_sweep<T, JB, U, PAIR, -1, 0>(base, lda, nr, rs, jb, kmax, tid, ja0, na, s_f,
                                                 scale, betav, rowj_out, rowj2_out, acc, acc2);
      }
#undef LA_QR_SWEEP
    } else {
      qr_stream_sweep<T, JB, U, PAIR, -1, 0>(base, lda, nr, rs, jb, kmax, tid, ja0, na, s_f, scale,
                                             betav, rowj_out, rowj2_out, acc, acc2);
    }
    if (js >= kmax) break;
    // ---- block reduction of the partial sums, publication, grid barrier, global reduction
    const int par = pass & 1;
#pragma unroll
    for (int c = 0; c < JB; c++) {
      const T sa = warp_sum<T>(acc[c]);
      if (lane == 0) s_red[0][warp][c] = sa;
      if (PAIR) {
        const T sb = warp_sum<T>(acc2[PAIR ? c : 0]);
        if (lane == 0) s_red[1][warp][c] = sb;
      }
    }
    __syncthreads();
    if (tid < jb && tid >= js) {
      T sa = zero<T>(), sb = zero<T>();
      for (int w = 0; w < QR_NT / 32; w++) {
        sa = add(sa, s_red[0][w][tid]);
        if (PAIR) sb = add(sb, s_red[1][w][tid]);
      }
      st_cg<T>(&xch->part[par][blockIdx.x][tid], sa);
      if (PAIR && two) st_cg<T>(&xch->part2[par][blockIdx.x][tid], sb);
    }
    target += (unsigned)G;
    grid_barrier(&xch->barrier, target);
    {
      const int c = tid & 31, grp = tid >> 5;
      T sa = zero<T>(), sb = zero<T>();
      if (c >= js && c < jb)
        for (int g = grp; g < G; g += QR_NT / 32) {
          sa = add(sa, ld_cg<T>(&xch->part[par][g][c]));
          if (PAIR && two) sb = add(sb, ld_cg<T>(&xch->part2[par][g][c]));
        }
      s_red[0][grp][c] = sa;
      if (PAIR) s_red[1][grp][c] = sb;
    }
    __syncthreads();
    if (tid < 32) {
      const int c = tid;
      const bool in = c >= js && c < jb;
      T s1 = zero<T>(), s2 = zero<T>();
      if (in)
        for (int g = 0; g < QR_NT / 32; g++) {
          s1 = add(s1, s_red[0][g][c]);
          if (PAIR) s2 = add(s2, s_red[1][g][c]);
        }
      const T r1 = in ? ld_cg<T>(&xch->rowj[par][c]) : zero<T>();
      auto bcast = [](T v, int src) {
        T o;
        if constexpr (Traits<T>::is_complex) {
          o.x = __shfl_sync(0xffffffffu, v.x, src); o.y = __shfl_sync(0xffffffffu, v.y, src);
        } else {
          o = __shfl_sync(0xffffffffu, v, src);
        }
        return o;
      };
      const T alpha1 = bcast(r1, js);
      const double sigma1 = (double)real_(bcast(s1, js));
      T beta1, tau1, sc1;
      householder<T>(alpha1, sigma1, beta1, tau1, sc1);
      const T w1 = fma_(conj_(sc1), s1, r1);        // w_c = a(js,c) + conj(scale) * s_c
      const T f1 = mul(conj_(tau1), w1);            // f_c = conj(tau) * w_c
      s_f[0][c] = f1;
      int napply = 1;
      if (PAIR && two) {
        // column sums after reflector js, over the rows > js:
        //   P_c = S2_c - conj(f_js2) conj(sc) S1_c - f_c sc conj(S1_js2) + conj(f_js2) f_c |sc|^2 S1_js
        // row js2 after reflector js: r2'_c = r2_c - (r2_js sc) f_c; its own term leaves the sums.
        const T r2 = in ? ld_cg<T>(&xch->rowj2[par][c]) : zero<T>();
        const T f1_2 = bcast(f1, js2);
        const T s1_2 = bcast(s1, js2);
        const T v1_2 = mul(bcast(r2, js), sc1);
        T pc = s2;
        pc = sub(pc, mul(mul(conj_(f1_2), conj_(sc1)), s1));
        pc = sub(pc, mul(mul(f1, sc1), conj_(s1_2)));
        const double sc_abs2 = (double)real_(mul(conj_(sc1), sc1));
        pc = add(pc, mul(mul(conj_(f1_2), f1), from_real<T>(sc_abs2 * sigma1)));
        const T r2p = fnma_(v1_2, f1, r2);
        const T alpha2 = bcast(r2p, js2);
        const T s2p = sub(pc, mul(conj_(alpha2), r2p));
        const double sigma2 = (double)real_(bcast(s2p, js2));
        const double orig2 = (double)real_(bcast(s2, js2));
        if (sigma2 >= 0.0625 * orig2 && sigma2 == sigma2) {
          T beta2, tau2, sc2;
          householder<T>(alpha2, sigma2, beta2, tau2, sc2);
          const T w2 = fma_(conj_(sc2), s2p, r2p);
          s_f[1][c] = mul(conj_(tau2), w2);
          napply = 2;
          if (c == js2) {
            s_scale[1] = sc2;
            s_beta[1] = beta2;
            s_tau[js2] = tau2;
            if (blockIdx.x == 0) tau[js2] = tau2;
          }
        }
      }
      if (c == js) {
        s_scale[0] = sc1;
        s_beta[0] = beta1;
        s_tau[js] = tau1;
        if (blockIdx.x == 0) tau[js] = tau1;
        s_napply = napply;
      }
    }
    __syncthreads();
    ja0 = js;
    na = s_napply;
    pass++;
  }
  // ---- optional epilogue (8-column panels): one more pass over the finished panel writes the
  // clean copy of the reflectors (unit diagonal, zeros above) and accumulates G = V^H V; after a
  // last grid barrier CTA 0 turns G and tau into the panel's T factor (?larft).  This replaces a
  // copy kernel, a Gram product, its reduction and a ?larft launch per panel.
  if constexpr (JB <= 8) {
    if (Vc != nullptr) {
      constexpr int NG = JB * (JB + 1) / 2;
      T g[NG];
#pragma unroll
      for (int i = 0; i < NG; i++) g[i] = zero<T>();
      for (int64_t r = tid; r < nr; r += QR_NT) {
        const int64_t gr = rs + r;
        T v[JB];
#pragma unroll
        for (int c = 0; c < JB; c++) {
          const T a = c < jb ? base[(int64_t)c * lda + r] : zero<T>();
          v[c] = (c < jb && gr > c) ? a : ((c < jb && gr == c) ? one<T>() : zero<T>());
        }
#pragma unroll
        for (int c = 0; c < JB; c++)
          if (c < jb) Vc[(int64_t)c * ldv + gr] = v[c];
        int idx = 0;
#pragma unroll
        for (int j = 0; j < JB; j++)
#pragma unroll
          for (int i = 0; i <= j; i++) {
            g[idx] = fma_(conj_(v[i]), v[j], g[idx]);
            idx++;
          }
      }
      // block reduction through s_red (8 warps x 32 slots; NG <= 36 needs two rounds)
      // (partial sums go to the parity half the last column pass did not use: every CTA has left
      // that half behind before any CTA can be here)
      T* gp = &xch->part[pass & 1][0][0];
      for (int base_i = 0; base_i < NG; base_i += 32) {
#pragma unroll
        for (int i = 0; i < NG; i++) {
          if (i >= base_i && i < base_i + 32) {
            const T sgm = warp_sum<T>(g[i]);
            if (lane == 0) s_red[0][warp][i - base_i] = sgm;
          }
        }
        __syncthreads();
        if (tid < 32 && base_i + tid < NG) {
          T sgm = zero<T>();
          for (int w = 0; w < QR_NT / 32; w++) sgm = add(sgm, s_red[0][w][tid]);
          st_cg<T>(&gp[(int64_t)blockIdx.x * NG + base_i + tid], sgm);
        }
        __syncthreads();
      }
      target += (unsigned)G;
      grid_barrier(&xch->barrier, target);
      if (blockIdx.x == 0) {
        __shared__ T s_g[NG];
        if (tid < NG) {
          T sgm = zero<T>();
          for (int cta = 0; cta < G; cta++) sgm = add(sgm, ld_cg<T>(&gp[(int64_t)cta * NG + tid]));
          s_g[tid] = sgm;
        }
        __syncthreads();
        if (tid == 0) {
          // G(i, j), i <= j, sits at j * (j + 1) / 2 + i
          T Tl[JB][JB];
          for (int i = 0; i < jb; i++) {
            for (int r2 = 0; r2 < jb; r2++) Tl[r2][i] = zero<T>();
            const T ti = s_tau[i];
            for (int r2 = 0; r2 < i; r2++) {
              T sgm = zero<T>();
              for (int c = r2; c < i; c++) sgm = fma_(Tl[r2][c], s_g[i * (i + 1) / 2 + c], sgm);
              Tl[r2][i] = mul(neg(ti), sgm);
            }
            Tl[i][i] = ti;
          }
          for (int i = 0; i < jb; i++)
            for (int r2 = 0; r2 < jb; r2++) Tout[r2 + (int64_t)i * ldt] = Tl[r2][i];
        }
      }
    }
  }
}

// T factor of a block reflector from G = V^H V and tau (forward, columnwise ?larft):
//   T(i,i) = tau_i,  T(0:i,i) = -tau_i * T(0:i,0:i) * G(0:i,i);  strictly lower part = 0.
template <typename T>
__global__ void __launch_bounds__(64)
larft_kernel(int k, const T* __restrict__ Gm, int64_t ldg, const T* __restrict__ tau,
             T* __restrict__ Tm, int64_t ldt) {
  // T(0:i, i) = -tau_i T(0:i, 0:i) G(0:i, i): row t of T depends on row t alone, so thread t walks
  // its row from the diagonal to the right without any barrier; G and tau are staged in shared
  // memory first (the earlier form read them from global memory inside the recurrence and
  // synchronised the block once per column: 38 us for a 32 x 32 factor).
  __shared__ T Ts[QR_MAX_JB][QR_MAX_JB + 1];
  __shared__ T Gs[QR_MAX_JB][QR_MAX_JB + 1];    // Gs[i][c] = G(c, i)
  __shared__ T taus[QR_MAX_JB];
  const int t = threadIdx.x;
  for (int idx = t; idx < k * k; idx += blockDim.x) {
    const int c = idx % k, i = idx / k;
    Gs[i][c] = Gm[c + (int64_t)i * ldg];
    Ts[c][i] = zero<T>();
  }
  if (t < k) taus[t] = tau[t];
  __syncthreads();
  if (t < k) {
    Ts[t][t] = taus[t];
    for (int i = t + 1; i < k; i++) {
      T s = zero<T>();
      for (int c = t; c < i; c++) s = fma_(Ts[t][c], Gs[i][c], s);
      Ts[t][i] = mul(neg(taus[i]), s);
    }
  }
  __syncthreads();
  for (int idx = t; idx < k * k; idx += blockDim.x) Tm[idx % k + (int64_t)(idx / k) * ldt] = Ts[idx % k][idx / k];
}

template <typename T>
struct QrWork {
  QrExchange<T>* xch;
  T* V;     // mp x nb clean copy of the reflectors (unit diagonal, zeros above)
  int64_t ldv;
  T* Tm;    // nb x nb
  T* Gm;    // nb x nb
  T* W;     // nb x ncols  (Left)  or  mrows x nb (Right)
  T* W2;
  int64_t ldw;
};

// launch the streaming kernel on as many co-resident CTAs as the occupancy allows (at least
// 1024 rows each): the more loads in flight, the closer a pass gets to the HBM bandwidth
template <typename T, int JB, int U>
void launch_stream(Context& ctx, int64_t mp, int jb, T* A, int64_t lda, T* tau, QrExchange<T>* xch,
                   T* Vc = nullptr, int64_t ldv = 0, T* Tout = nullptr, int64_t ldt = 0) {
  static int per_sm[DeviceOnce::kMax] = {};
  const int dev = ctx.device >= 0 && ctx.device < DeviceOnce::kMax ? ctx.device : 0;
  if (per_sm[dev] == 0) {
    int nb = 1;
    LA_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(
        &nb, qr_panel_stream_kernel<T, JB, U>, QR_NT, 0));
    static const int cap = getenv("LA_B200_QR_CTAS_PER_SM") ? atoi(getenv("LA_B200_QR_CTAS_PER_SM")) : 4;
    per_sm[dev] = nb < 1 ? 1 : (nb > cap ? cap : nb);
  }
  const int sms = ctx.sm_limit > 0 && ctx.sm_limit < ctx.sm_count ? ctx.sm_limit : ctx.sm_count;
  int64_t G = (int64_t)sms * per_sm[dev];
  if (G > QR_MAX_CTAS) G = QR_MAX_CTAS;
  const int64_t by_rows = ceil_div(mp, (int64_t)1024);
  if (G > by_rows) G = by_rows;
  if (G < 1) G = 1;
  int64_t rpc = round_up(ceil_div(mp, G), 2);
  G = ceil_div(mp, rpc);
  if (G * (JB * (JB + 1) / 2) > (int64_t)QR_MAX_CTAS * QR_MAX_JB) Vc = nullptr;   // (never with JB = 8)
  void* sargs[] = {&mp, &jb, &A, &lda, &tau, &xch, &rpc, &Vc, &ldv, &Tout, &ldt};
  LA_CUDA_CHECK(cudaLaunchCooperativeKernel((void*)qr_panel_stream_kernel<T, JB, U>,
                                            dim3((unsigned)G), dim3(QR_NT), sargs, 0, ctx.stream));
}

// true when an mp x QR_MAX_JB panel does not fit the shared memory of all SMs together: such
// panels are streamed from global memory
template <typename T>
bool qr_panel_is_tall(Context& ctx, int64_t mp) {
  return (size_t)round_up(ceil_div(mp, (int64_t)ctx.sm_count), 2) * QR_MAX_JB * sizeof(T) >
         QR_SMEM_BUDGET;
}

// Returns true when the panel kernel also produced the clean reflectors in Vc and the panel's T
// factor in Tout (streamed 8-column panels); otherwise the caller builds them.
template <typename T>
bool qr_panel(Context& ctx, int64_t mp, int jb, T* A, int64_t lda, T* tau, QrExchange<T>* xch,
              unsigned& seq, T* Vc = nullptr, int64_t ldv = 0, T* Tout = nullptr, int64_t ldt = 0) {
  int64_t G = ceil_div(mp, 512);
  if (G > ctx.sm_count) G = ctx.sm_count;
  if (G < 1) G = 1;
  int64_t rpc = round_up(ceil_div(mp, G), 2);
  G = ceil_div(mp, rpc);
  size_t smem = (size_t)rpc * jb * sizeof(T);
  int use_smem = 1;
  if (smem > QR_SMEM_BUDGET) {
    // try all SMs
    G = ctx.sm_count;
    rpc = round_up(ceil_div(mp, G), 2);
    G = ceil_div(mp, rpc);
    smem = (size_t)rpc * jb * sizeof(T);
    if (smem > QR_SMEM_BUDGET) { use_smem = 0; smem = 0; }
  }
  LA_CUDA_CHECK(cudaMemsetAsync(&xch->barrier, 0, sizeof(unsigned), ctx.stream));
  if (!use_smem) {
    if (jb <= 8) {
      const bool fuse = Vc != nullptr && Tout != nullptr && mp >= jb;
      launch_stream<T, 8, 4>(ctx, mp, jb, A, lda, tau, xch, fuse ? Vc : nullptr, ldv, Tout, ldt);
      return fuse;
    }
    else if (jb <= 16) launch_stream<T, 16, 2>(ctx, mp, jb, A, lda, tau, xch);
    else {
      if constexpr (Traits<T>::is_complex) {
        fprintf(stderr, "la_b200: internal error: complex streaming QR panel wider than 16\n");
        abort();
      } else {
        launch_stream<T, 32, 1>(ctx, mp, jb, A, lda, tau, xch);
      }
    }
    return false;
  }
  unsigned seq_base = seq;
  seq += 64;
  void* args[] = {&mp, &jb, &A, &lda, &tau, &xch, &rpc, &seq_base};
  LA_CUDA_CHECK(cudaLaunchCooperativeKernel((void*)qr_panel_kernel<T, true>, dim3((unsigned)G),
                                            dim3(QR_NT), args, smem, ctx.stream));
  return false;
}

template <typename T>
void ensure_attrs(Context& ctx) {
  static DeviceOnce done;
  if (done.pending(ctx.device)) {
    LA_CUDA_CHECK(cudaFuncSetAttribute(qr_panel_kernel<T, true>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)QR_SMEM_BUDGET));
    done.mark(ctx.device);
  }
}

// Build the clean V (mp x jb) and T (jb x jb) of one block of reflectors stored in A.
template <typename T>
void form_block_reflector(Context& ctx, int64_t mp, int jb, const T* A, int64_t lda, const T* tau,
                          QrWork<T>& w) {
  copy_triangle<T>(ctx, Lower, true, mp, jb, A, lda, w.V, w.ldv);
  gemm<T>(ctx, OpC, OpN, jb, jb, mp, one<T>(), w.V, w.ldv, w.V, w.ldv, zero<T>(), w.Gm,
          QR_MAX_JB);
  larft_kernel<T><<<1, 64, 0, ctx.stream>>>(jb, w.Gm, QR_MAX_JB, tau, w.Tm, QR_MAX_JB);
  LA_CUDA_CHECK(cudaGetLastError());
}

// C <- op(H) C (Left) or C op(H) (Right), H = I - V T V^H, op = N or C.
template <typename T>
void apply_block_reflector(Context& ctx, Side side, Op trans, int64_t m, int64_t n, int jb,
                           QrWork<T>& w, T* C, int64_t ldc) {
  const T onev = one<T>(), zerov = zero<T>(), mone = from_real<T>(-1.0);
  if (side == Left) {
    // With W = V^H C (jb x n) and W2 = op(T) W the update is C -= V W2.  The products are
    // formed transposed, Wt = W^H = C^H V (n x jb) and W2t = W2^H = Wt op(T)^H, so that the long
    // reduction over the m rows has the jb <= 32 columns on the N side of the GEMM, where the
    // narrow 128 x 32 tile applies (with jb rows on the M side 3/4 of a 128-row tile is padding).
    const int64_t ldt = round_up(n, 16);
    gemm<T>(ctx, OpC, OpN, n, jb, m, onev, C, ldc, w.V, w.ldv, zerov, w.W, ldt);
    gemm<T>(ctx, OpN, trans == OpN ? OpC : OpN, n, jb, jb, onev, w.W, ldt, w.Tm, QR_MAX_JB, zerov,
            w.W2, ldt);
    gemm<T>(ctx, OpN, OpC, m, n, jb, mone, w.V, w.ldv, w.W2, ldt, onev, C, ldc);
  } else {
    // W = C V (m x jb); W2 = W op(T); C -= W2 V^H
    gemm<T>(ctx, OpN, OpN, m, jb, n, onev, C, ldc, w.V, w.ldv, zerov, w.W, w.ldw);
    gemm<T>(ctx, OpN, trans == OpN ? OpN : OpC, m, jb, jb, onev, w.W, w.ldw, w.Tm, QR_MAX_JB,
            zerov, w.W2, w.ldw);
    gemm<T>(ctx, OpN, OpC, m, n, jb, mone, w.W2, w.ldw, w.V, w.ldv, onev, C, ldc);
  }
}

template <typename T>
QrWork<T> make_work(Context& ctx, int64_t vrows, int64_t w_rows, int64_t w_cols) {
  QrWork<T> w;
  w.xch = (QrExchange<T>*)ctx.arena.alloc(sizeof(QrExchange<T>));
  w.ldv = round_up(vrows, 16);
  w.V = ctx.alloc<T>(w.ldv * QR_MAX_JB);
  w.Tm = ctx.alloc<T>(QR_MAX_JB * QR_MAX_JB);
  w.Gm = ctx.alloc<T>(QR_MAX_JB * QR_MAX_JB);
  w.ldw = round_up(w_rows, 16);
  // (+16 columns: the Left update stores its jb x n workspaces transposed, n rounded up to 16)
  w.W = ctx.alloc<T>(w.ldw * ((w_cols > 0 ? w_cols : 1) + 16));
  w.W2 = ctx.alloc<T>(w.ldw * ((w_cols > 0 ? w_cols : 1) + 16));
  return w;
}

}  // namespace

namespace {

// Workspace of one block of at most QR_NB reflectors (geqrf_ex): a clean copy of the reflectors
// (unit diagonal, zeros above), the block's T factor, and the small matrices of the updates.
constexpr int QR_NB = 128;

template <typename T>
struct QrBlock {
  QrExchange<T>* xch;
  T* Vc; int64_t ldv;      // (rows of the block) x QR_NB
  T* Tb;                   // QR_NB x QR_NB, upper triangular (ld QR_NB)
  T* X; T* X2;             // QR_NB x QR_NB scratch
  T* G;                    // QR_MAX_JB x QR_MAX_JB Gram matrix of a leaf
  T* W; T* W2; int64_t ldw;  // (columns to update) x QR_NB
  int64_t b0;              // first column (= first row) of the block
  int leaf;                // leaf panel width
  unsigned seq;            // packet sequence of the shared-memory panel launches (this call)
};

// C <- H^H C with H = I - V T V^H (V: mrows x jb, unit lower trapezoidal, clean; T: jb x jb).
// The products with the long reduction are formed transposed (W^H = C^H V is ncols x jb), so
// that the jb reflectors sit on the N side of the GEMM, where narrow tiles / split-K apply.
template <typename T>
void apply_block_left_conj(Context& ctx, int64_t mrows, int64_t ncols, int64_t jb, const T* V,
                           int64_t ldv, const T* Tm, int64_t ldt, T* C, int64_t ldc, QrBlock<T>& w) {
  if (mrows <= 0 || ncols <= 0 || jb <= 0) return;
  const T onev = one<T>(), zerov = zero<T>(), mone = from_real<T>(-1.0);
  if (ncols <= 64 && jb > 64 && mrows >= 8192) {
    // few columns against a wide block (right-hand sides): 64 reflectors at a time, so that the
    // streaming X^H Y kernel takes the product instead of 128-row tensor-core tiles
    for (int64_t c0 = 0; c0 < jb; c0 += 64)
      gemm<T>(ctx, OpC, OpN, ncols, (jb - c0) < 64 ? (jb - c0) : 64, mrows, onev, C, ldc,
              V + c0 * ldv, ldv, zerov, w.W + c0 * w.ldw, w.ldw);
  } else {
    gemm<T>(ctx, OpC, OpN, ncols, jb, mrows, onev, C, ldc, V, ldv, zerov, w.W, w.ldw);
  }
  gemm<T>(ctx, OpN, OpN, ncols, jb, jb, onev, w.W, w.ldw, Tm, ldt, zerov, w.W2, w.ldw);
  if (!tall_update<T>(ctx, mrows, jb, ncols, V, ldv, w.W2, w.ldw, C, ldc))
    gemm<T>(ctx, OpN, OpC, mrows, ncols, jb, mone, V, ldv, w.W2, w.ldw, onev, C, ldc);
}

// Factor columns [j0, j0 + jn) of A (rows j0..m), all earlier reflectors applied, and leave the
// clean reflectors in w.Vc and their T factor in w.Tb.  Recursive halving: the left half is
// factored, applied to the right half, the right half is factored, and the two T factors are
// merged,  T = [T1, -T1 (V1^H V2) T2; 0, T2].
template <typename T>
void qr_rec(Context& ctx, int64_t m, T* A, int64_t lda, T* tau, int64_t j0, int64_t jn,
            QrBlock<T>& w) {
  const int64_t o = j0 - w.b0;                 // offset inside the block
  T* Vn = w.Vc + o + o * w.ldv;                // reflectors of this node, rows from j0
  T* Tn = w.Tb + o + o * QR_NB;
  const int64_t mp = m - j0;
  if (jn <= w.leaf) {
    T* P = A + j0 + j0 * lda;
    if (qr_panel<T>(ctx, mp, (int)jn, P, lda, tau + j0, w.xch, w.seq, Vn, w.ldv, Tn, QR_NB)) return;
    const int64_t kk = mp < jn ? mp : jn;      // reflectors actually produced
    copy_triangle<T>(ctx, Lower, true, mp, jn, P, lda, Vn, w.ldv);
    gemm<T>(ctx, OpC, OpN, kk, kk, mp, one<T>(), Vn, w.ldv, Vn, w.ldv, zero<T>(), w.G, QR_MAX_JB);
    larft_kernel<T><<<1, 64, 0, ctx.stream>>>((int)kk, w.G, QR_MAX_JB, tau + j0, Tn, QR_NB);
    LA_CUDA_CHECK(cudaGetLastError());
    return;
  }
  int64_t h = round_up(jn / 2, (int64_t)w.leaf);
  if (h >= jn) h = jn - w.leaf;
  qr_rec<T>(ctx, m, A, lda, tau, j0, h, w);
  apply_block_left_conj<T>(ctx, mp, jn - h, h, Vn, w.ldv, Tn, QR_NB, A + j0 + (j0 + h) * lda, lda, w);
  if (mp - h <= 0) return;                     // (only when m < n: nothing left below)
  qr_rec<T>(ctx, m, A, lda, tau, j0 + h, jn - h, w);
  // merge: rows j0+h.. of V1 against V2 (V2 is zero above them)
  const T* V1s = Vn + h;
  const T* V2 = Vn + h + h * w.ldv;
  const T* T2 = Tn + h + h * QR_NB;
  T* T12 = Tn + h * QR_NB;
  const int64_t r2 = jn - h;
  gemm<T>(ctx, OpC, OpN, h, r2, mp - h, one<T>(), V1s, w.ldv, V2, w.ldv, zero<T>(), w.X, QR_NB);
  gemm<T>(ctx, OpN, OpN, h, r2, h, from_real<T>(-1.0), Tn, QR_NB, w.X, QR_NB, zero<T>(), w.X2, QR_NB);
  gemm<T>(ctx, OpN, OpN, h, r2, r2, one<T>(), w.X2, QR_NB, T2, QR_NB, zero<T>(), T12, QR_NB);
}

template <typename T>
__global__ void zero_square_kernel(int n, T* A, int64_t lda) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  for (int j = blockIdx.y; j < n; j += gridDim.y) A[i + (int64_t)j * lda] = zero<T>();
}

}  // namespace

// Householder QR of A with the reflectors also applied to nrhs extra columns B (B <- Q^H B;
// nrhs = 0: plain ?geqrf).  Blocks of QR_NB columns, each factored recursively from leaf panels
// (8 columns when the panel has to be streamed from HBM, 32 when it fits shared memory); the
// columns right of a block, and B, receive one rank-QR_NB update per block.
template <typename T>
void geqrf_ex(Context& ctx, int64_t m, int64_t n, T* A, int64_t lda, T* d_tau, int64_t nrhs, T* B,
              int64_t ldb) {
  const int64_t k = m < n ? m : n;
  if (k <= 0) return;
  ensure_attrs<T>(ctx);
  Arena::Mark mk = ctx.arena.mark();
  QrBlock<T> w;
  w.xch = (QrExchange<T>*)ctx.arena.alloc(sizeof(QrExchange<T>));
  // the packet area comes out of the arena with whatever was there: cleared in stream order, and
  // the tags of this call start above every cleared value
  LA_CUDA_CHECK(cudaMemsetAsync(&w.xch->pll[0][0][0], 0,
                                sizeof(w.xch->pll) + sizeof(w.xch->rowll), ctx.stream));
  w.seq = 64;
  w.ldv = round_up(m, 16);
  w.Vc = ctx.alloc<T>(w.ldv * QR_NB);
  w.Tb = ctx.alloc<T>(QR_NB * QR_NB);
  w.X = ctx.alloc<T>(QR_NB * QR_NB);
  w.X2 = ctx.alloc<T>(QR_NB * QR_NB);
  w.G = ctx.alloc<T>(QR_MAX_JB * QR_MAX_JB);
  const int64_t wc = (n > nrhs ? n : nrhs);
  w.ldw = round_up(wc > 0 ? wc : 1, 16);
  w.W = ctx.alloc<T>(w.ldw * QR_NB);
  w.W2 = ctx.alloc<T>(w.ldw * QR_NB);
  w.leaf = qr_panel_is_tall<T>(ctx, m) ? 8 : QR_MAX_JB;
  for (int64_t b0 = 0; b0 < k; b0 += QR_NB) {
    const int64_t bw = (k - b0) < QR_NB ? (k - b0) : QR_NB;
    const int64_t mp = m - b0;
    w.b0 = b0;
    // the part of Vc above the diagonal of later leaves and the lower part of Tb must be zero
    {
      const int sq = (int)(mp < QR_NB ? mp : QR_NB);
      dim3 g((unsigned)ceil_div(sq, 128), (unsigned)sq);
      zero_square_kernel<T><<<g, 128, 0, ctx.stream>>>(sq, w.Vc, w.ldv);
      dim3 g2(1, QR_NB);
      zero_square_kernel<T><<<g2, 128, 0, ctx.stream>>>(QR_NB, w.Tb, QR_NB);
      LA_CUDA_CHECK(cudaGetLastError());
    }
    qr_rec<T>(ctx, m, A, lda, d_tau, b0, bw, w);
    const int64_t kk = mp < bw ? mp : bw;
    const int64_t nc = n - b0 - bw;
    if (nc > 0)
      apply_block_left_conj<T>(ctx, mp, nc, kk, w.Vc, w.ldv, w.Tb, QR_NB, A + b0 + (b0 + bw) * lda,
                               lda, w);
    if (nrhs > 0)
      apply_block_left_conj<T>(ctx, mp, nrhs, kk, w.Vc, w.ldv, w.Tb, QR_NB, B + b0, ldb, w);
  }
  ctx.arena.release(mk);
}

template <typename T>
void geqrf(Context& ctx, int64_t m, int64_t n, T* A, int64_t lda, T* d_tau) {
  geqrf_ex<T>(ctx, m, n, A, lda, d_tau, 0, nullptr, 0);
}

template <typename T>
void ormqr(Context& ctx, Side side, Op trans, int64_t m, int64_t n, int64_t k, const T* A,
           int64_t lda, const T* d_tau, T* C, int64_t ldc) {
  if (m <= 0 || n <= 0 || k <= 0) return;
  if (trans == OpT && !Traits<T>::is_complex) trans = OpC;
  const int64_t nq = side == Left ? m : n;  // order of Q
  Arena::Mark mk = ctx.arena.mark();
  QrWork<T> w = side == Left ? make_work<T>(ctx, nq, QR_MAX_JB, n)
                             : make_work<T>(ctx, nq, m, QR_MAX_JB);
  // Q = Hb_1 ... Hb_p.  Left/N and Right/C apply the last block first.
  const bool forward = (side == Left) == (trans != OpN);
  const int64_t nblk = ceil_div(k, QR_MAX_JB);
  for (int64_t b = 0; b < nblk; b++) {
    const int64_t bi = forward ? b : nblk - 1 - b;
    const int64_t j0 = bi * QR_MAX_JB;
    const int jb = (int)((k - j0) < QR_MAX_JB ? (k - j0) : QR_MAX_JB);
    const int64_t mp = nq - j0;
    form_block_reflector<T>(ctx, mp, jb, A + j0 + j0 * lda, lda, d_tau + j0, w);
    if (side == Left)
      apply_block_reflector<T>(ctx, Left, trans, mp, n, jb, w, C + j0, ldc);
    else
      apply_block_reflector<T>(ctx, Right, trans, m, mp, jb, w, C + j0 * ldc, ldc);
  }
  ctx.arena.release(mk);
}

template <typename T>
void orgqr(Context& ctx, int64_t m, int64_t n, int64_t k, T* A, int64_t lda, const T* d_tau) {
  if (m <= 0 || n <= 0) return;
  Arena::Mark mk = ctx.arena.mark();
  const int64_t ldq = round_up(m, 16);
  T* Q = ctx.alloc<T>(ldq * n);
  set_matrix<T>(ctx, m, n, zero<T>(), one<T>(), Q, ldq);
  ormqr<T>(ctx, Left, OpN, m, n, k, A, lda, d_tau, Q, ldq);
  copy_matrix<T>(ctx, m, n, Q, ldq, A, lda);
  ctx.arena.release(mk);
}

namespace {
template <typename T>
__global__ void zero_rows_kernel(int64_t r0, int64_t r1, int64_t n, T* B, int64_t ldb) {
  int64_t i = r0 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= r1) return;
  for (int64_t j = blockIdx.y; j < n; j += gridDim.y) B[i + j * ldb] = zero<T>();
}
template <typename T>
void zero_rows(Context& ctx, int64_t r0, int64_t r1, int64_t n, T* B, int64_t ldb) {
  if (r1 <= r0 || n <= 0) return;
  dim3 g((unsigned)ceil_div(r1 - r0, 256), (unsigned)(n < 1024 ? n : 1024));
  zero_rows_kernel<T><<<g, 256, 0, ctx.stream>>>(r0, r1, n, B, ldb);
  LA_CUDA_CHECK(cudaGetLastError());
}
}  // namespace

// ?gels: full-rank least squares / minimum norm.  The LQ branches are run as the QR of A^H
// held in workspace; A is overwritten with the LQ factors in LAPACK's ?gelqf layout (the
// conjugate transpose of the QR result).  The ?lascl rescaling of badly scaled inputs that
// LAPACK performs first is not replicated.
template <typename T>
void gels(Context& ctx, Op trans, int64_t m, int64_t n, int64_t nrhs, T* A, int64_t lda, T* B,
          int64_t ldb, int* d_info) {
  LA_CUDA_CHECK(cudaMemsetAsync(d_info, 0, sizeof(int), ctx.stream));
  const int64_t mn = m < n ? m : n;
  if (mn <= 0 || nrhs < 0) {
    if (nrhs > 0) zero_rows<T>(ctx, 0, m > n ? m : n, nrhs, B, ldb);
    return;
  }
  const bool tpose = trans != OpN;
  Arena::Mark mk = ctx.arena.mark();
  T* tau = ctx.alloc<T>(mn);
  if (m >= n) {
    if (!tpose) {
      // least squares: B <- Q^H B (applied block by block while A is factored), solve R X = B(0:n)
      geqrf_ex<T>(ctx, m, n, A, lda, tau, nrhs, B, ldb);
      trtrs<T>(ctx, Upper, OpN, NonUnit, n, nrhs, A, lda, B, ldb, d_info);
    } else {
      geqrf<T>(ctx, m, n, A, lda, tau);
      // minimum norm of A^H X = B: X = Q [R^-H B; 0]
      trtrs<T>(ctx, Upper, OpC, NonUnit, n, nrhs, A, lda, B, ldb, d_info);
      zero_rows<T>(ctx, n, m, nrhs, B, ldb);
      ormqr<T>(ctx, Left, OpN, m, nrhs, n, A, lda, tau, B, ldb);
    }
  } else {
    // A = L Q  <=>  A^H = Q' R with L = R^H
    const int64_t ldt = round_up(n, 16);
    T* At = ctx.alloc<T>(ldt * m);
    transpose_matrix<T>(ctx, true, m, n, A, lda, At, ldt);  // At = A^H (n x m)
    geqrf<T>(ctx, n, m, At, ldt, tau);
    if (!tpose) {
      // minimum norm: X = Q' [R^-H B; 0]   (L y = b with L = R^H)
      trtrs<T>(ctx, Upper, OpC, NonUnit, m, nrhs, At, ldt, B, ldb, d_info);
      zero_rows<T>(ctx, m, n, nrhs, B, ldb);
      ormqr<T>(ctx, Left, OpN, n, nrhs, m, At, ldt, tau, B, ldb);
    } else {
      // least squares of A^H X = B (B is n x nrhs): B <- Q'^H B, solve R X = B(0:m)
      ormqr<T>(ctx, Left, OpC, n, nrhs, m, At, ldt, tau, B, ldb);
      trtrs<T>(ctx, Upper, OpN, NonUnit, m, nrhs, At, ldt, B, ldb, d_info);
    }
    transpose_matrix<T>(ctx, true, n, m, At, ldt, A, lda);  // LQ layout back into A
  }
  ctx.arena.release(mk);
}

// =============================================================================================
// RQ / LQ / QL factorisations and their Q generators (src/linalg/rq_lq_ql.cr:3-39 ->
// ?gerqf_/?gelqf_/?geqlf_, ?orgrq_/?orglq_/?orgql_ and the zung* forms).  Each is the QR
// factorisation of a mirrored and / or conjugate-transposed copy of A, which gives exactly
// LAPACK's storage: ?gelq2 / ?gerq2 / ?geql2 generate the same ?larfg reflectors as ?geqr2 does
// on that copy (rows instead of columns, last-to-first instead of first-to-last).
//   LQ:  A^H            = Q R   ->  A = R^H Q^H            tau as returned
//   QL:  J A J          = Q R   ->  A = (J Q J)(J R J)     tau reversed
//   RQ:  (J A J)^H      = Q R   ->  A = (J R^H J)(J Q^H J) tau reversed
namespace {

// B (rows x cols of the RESULT) from A: optional mirror of both index ranges, optional conjugate
// transposition.  mode bit 0: conjugate-transpose, bit 1: mirror.  A is ma x na.
template <typename T>
__global__ void mirror_copy_kernel(int mode, int64_t ma, int64_t na, const T* __restrict__ A,
                                   int64_t lda, T* __restrict__ B, int64_t ldb) {
  const bool tr = mode & 1, mir = mode & 2;
  const int64_t mb = tr ? na : ma, nb = tr ? ma : na;
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= mb) return;
  for (int64_t c = blockIdx.y; c < nb; c += gridDim.y) {
    int64_t sr = tr ? c : r, sc = tr ? r : c;
    if (mir) { sr = ma - 1 - sr; sc = na - 1 - sc; }
    const T v = A[sr + sc * lda];
    B[r + c * ldb] = tr ? conj_(v) : v;
  }
}
template <typename T>
void mirror_copy(Context& ctx, int mode, int64_t ma, int64_t na, const T* A, int64_t lda, T* B,
                 int64_t ldb) {
  if (ma <= 0 || na <= 0) return;
  const int64_t mb = (mode & 1) ? na : ma, nb = (mode & 1) ? ma : na;
  dim3 g((unsigned)ceil_div(mb, 256), (unsigned)(nb < 2048 ? nb : 2048));
  mirror_copy_kernel<T><<<g, 256, 0, ctx.stream>>>(mode, ma, na, A, lda, B, ldb);
  LA_CUDA_CHECK(cudaGetLastError());
}
template <typename T>
__global__ void reverse_copy_kernel(int64_t n, const T* __restrict__ x, T* __restrict__ y) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) y[i] = x[n - 1 - i];
}
template <typename T>
void reverse_copy(Context& ctx, int64_t n, const T* x, T* y) {
  if (n <= 0) return;
  reverse_copy_kernel<T><<<(unsigned)ceil_div(n, 256), 256, 0, ctx.stream>>>(n, x, y);
  LA_CUDA_CHECK(cudaGetLastError());
}

}  // namespace

// which: 'L' = gelqf, 'R' = gerqf, 'Q' = geqlf
template <typename T>
void geqrf_variant(Context& ctx, char which, int64_t m, int64_t n, T* A, int64_t lda, T* d_tau) {
  const int64_t k = m < n ? m : n;
  if (k <= 0) return;
  const int mode = which == 'L' ? 1 : (which == 'Q' ? 2 : 3);
  const int64_t mw = (mode & 1) ? n : m, nw = (mode & 1) ? m : n;
  Arena::Mark mk = ctx.arena.mark();
  const int64_t ldw = round_up(mw, 16);
  T* W = ctx.alloc<T>(ldw * nw);
  T* tw = ctx.alloc<T>(k);
  mirror_copy<T>(ctx, mode, m, n, A, lda, W, ldw);
  geqrf<T>(ctx, mw, nw, W, ldw, mode == 1 ? d_tau : tw);
  if (mode != 1) reverse_copy<T>(ctx, k, tw, d_tau);
  mirror_copy<T>(ctx, mode, mw, nw, W, ldw, A, lda);        // the mapping is its own inverse
  ctx.arena.release(mk);
}

// which: 'L' = orglq (m x n, orthonormal rows), 'R' = orgrq, 'Q' = orgql (m x n, orthonormal columns)
template <typename T>
void orgqr_variant(Context& ctx, char which, int64_t m, int64_t n, int64_t k, T* A, int64_t lda,
                   const T* d_tau) {
  if (m <= 0 || n <= 0) return;
  const int mode = which == 'L' ? 1 : (which == 'Q' ? 2 : 3);
  const int64_t mw = (mode & 1) ? n : m, nw = (mode & 1) ? m : n;
  Arena::Mark mk = ctx.arena.mark();
  const int64_t ldw = round_up(mw, 16);
  T* W = ctx.alloc<T>(ldw * nw);
  T* tw = ctx.alloc<T>(k > 0 ? k : 1);
  mirror_copy<T>(ctx, mode, m, n, A, lda, W, ldw);
  const T* tq = d_tau;
  if (mode != 1 && k > 0) { reverse_copy<T>(ctx, k, d_tau, tw); tq = tw; }
  // after mirroring, the k reflectors sit in the FIRST k columns of W, first reflector first
  orgqr<T>(ctx, mw, nw, k, W, ldw, tq);
  mirror_copy<T>(ctx, mode, mw, nw, W, ldw, A, lda);
  ctx.arena.release(mk);
}

// =============================================================================================
// QR with column pivoting, ?geqp3_ (src/linalg/qr.cr:38-41; rank(method: QRP),
// src/linalg/linalg.cr:772-774).  Column by column as ?laqp2 does: the free column with the
// largest partial norm (first maximum) is swapped into place, its ?larfg reflector is generated
// and applied to the columns to the right, whose partial norms are then downdated with LAPACK's
// formula and recomputed when the downdate has lost its accuracy (tol3z = sqrt(eps)).  Two
// launches per column: this routine is about parity of pivots and factors, not speed.
namespace {

constexpr int QP_NT = 512;

template <typename T>
__global__ void __launch_bounds__(QP_NT)
qp3_norms_kernel(int64_t m, int64_t n, const T* __restrict__ A, int64_t lda, double* vn1, double* vn2) {
  __shared__ double red[QP_NT / 32];
  for (int64_t c = blockIdx.x; c < n; c += gridDim.x) {
    double s = 0.0;
    for (int64_t i = threadIdx.x; i < m; i += QP_NT) {
      const T v = A[i + c * lda];
      s += (double)real_(v) * (double)real_(v) + (double)imag_(v) * (double)imag_(v);
    }
    s = warp_sum<double>(s);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
      double t = 0.0;
      for (int w = 0; w < QP_NT / 32; w++) t += red[w];
      vn1[c] = vn2[c] = sqrt(t);
    }
    __syncthreads();
  }
}

// one CTA: choose the pivot of step j, swap it into column j, generate the reflector
template <typename T>
__global__ void __launch_bounds__(QP_NT)
qp3_pivot_kernel(int64_t m, int64_t n, int64_t j, int64_t nfixed, T* __restrict__ A, int64_t lda,
                 int* __restrict__ jpvt, T* __restrict__ tau, double* vn1, double* vn2) {
  __shared__ double sv[QP_NT / 32];
  __shared__ int si[QP_NT / 32];
  __shared__ int s_pvt;
  __shared__ T s_scale;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (j >= nfixed) {
    double bv = -1.0;
    int bi = 0x7fffffff;
    for (int64_t c = j + tid; c < n; c += QP_NT) {
      const double v = vn1[c];
      if (v > bv) { bv = v; bi = (int)c; }            // ascending c: the first maximum stays
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const double ov = __shfl_xor_sync(0xffffffffu, bv, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
    }
    if (lane == 0) { sv[warp] = bv; si[warp] = bi; }
    __syncthreads();
    if (tid == 0) {
      for (int w = 1; w < QP_NT / 32; w++)
        if (sv[w] > bv || (sv[w] == bv && si[w] < bi)) { bv = sv[w]; bi = si[w]; }
      s_pvt = bi == 0x7fffffff ? (int)j : bi;
    }
    __syncthreads();
  } else {
    if (tid == 0) s_pvt = (int)j;
    __syncthreads();
  }
  const int64_t pvt = s_pvt;
  if (pvt != j) {
    for (int64_t i = tid; i < m; i += QP_NT) {
      const T t = A[i + pvt * lda];
      A[i + pvt * lda] = A[i + j * lda];
      A[i + j * lda] = t;
    }
    if (tid == 0) {
      const int t = jpvt[pvt]; jpvt[pvt] = jpvt[j]; jpvt[j] = t;
      vn1[pvt] = vn1[j];
      vn2[pvt] = vn2[j];
    }
  }
  __syncthreads();
  // ?larfg on A(j:m, j)
  T* col = A + j * lda;
  double s = 0.0;
  for (int64_t i = j + 1 + tid; i < m; i += QP_NT) {
    const T v = col[i];
    s += (double)real_(v) * (double)real_(v) + (double)imag_(v) * (double)imag_(v);
  }
  s = warp_sum<double>(s);
  __syncthreads();
  if (lane == 0) sv[warp] = s;
  __syncthreads();
  if (tid == 0) {
    double t = 0.0;
    for (int w = 0; w < QP_NT / 32; w++) t += sv[w];
    T beta, tj, scale;
    householder<T>(col[j], t, beta, tj, scale);
    col[j] = beta;
    tau[j] = tj;
    s_scale = scale;
  }
  __syncthreads();
  const T scale = s_scale;
  for (int64_t i = j + 1 + tid; i < m; i += QP_NT) col[i] = mul(col[i], scale);
}

// one CTA per column c > j: apply H(j)^H, then downdate (or recompute) the partial norm
template <typename T>
__global__ void __launch_bounds__(256)
qp3_apply_kernel(int64_t m, int64_t n, int64_t j, T* __restrict__ A, int64_t lda,
                 const T* __restrict__ tau, double* vn1, double* vn2, double tol3z) {
  __shared__ T red[8];
  __shared__ double dred[8];
  __shared__ T s_w;
  __shared__ int s_recompute;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const T* v = A + j * lda;
  const T tj = tau[j];
  for (int64_t c = j + 1 + blockIdx.x; c < n; c += gridDim.x) {
    T* col = A + c * lda;
    T acc = zero<T>();
    for (int64_t i = j + tid; i < m; i += 256) {
      const T vi = i == j ? one<T>() : v[i];
      acc = fma_(conj_(vi), col[i], acc);
    }
    acc = warp_sum<T>(acc);
    if (lane == 0) red[warp] = acc;
    __syncthreads();
    if (tid == 0) {
      T t = zero<T>();
      for (int w = 0; w < 8; w++) t = add(t, red[w]);
      s_w = mul(conj_(tj), t);                      // conj(tau) * v^H a_c
    }
    __syncthreads();
    const T f = s_w;
    for (int64_t i = j + tid; i < m; i += 256) {
      const T vi = i == j ? one<T>() : v[i];
      col[i] = fnma_(vi, f, col[i]);
    }
    __syncthreads();
    if (tid == 0) {
      int rec = 0;
      if (vn1[c] != 0.0) {
        const T ajc = col[j];
        const double a = sqrt((double)real_(ajc) * (double)real_(ajc) + (double)imag_(ajc) * (double)imag_(ajc));
        double temp = 1.0 - (a / vn1[c]) * (a / vn1[c]);
        temp = temp > 0.0 ? temp : 0.0;
        const double r = vn1[c] / vn2[c];
        const double temp2 = temp * r * r;
        if (temp2 <= tol3z) rec = 1;
        else vn1[c] *= sqrt(temp);
      }
      s_recompute = rec;
    }
    __syncthreads();
    if (s_recompute) {
      double s = 0.0;
      for (int64_t i = j + 1 + tid; i < m; i += 256) {
        const T x = col[i];
        s += (double)real_(x) * (double)real_(x) + (double)imag_(x) * (double)imag_(x);
      }
      s = warp_sum<double>(s);
      if (lane == 0) dred[warp] = s;
      __syncthreads();
      if (tid == 0) {
        double t = 0.0;
        for (int w = 0; w < 8; w++) t += dred[w];
        vn1[c] = vn2[c] = j + 1 < m ? sqrt(t) : 0.0;
      }
    }
    __syncthreads();
  }
}

template <typename T>
__global__ void swap_columns_kernel(int64_t m, T* a, T* b) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < m) { const T t = a[i]; a[i] = b[i]; b[i] = t; }
}

}  // namespace

// d_jpvt: device array of n ints, on entry LAPACK's "fixed column" markers already resolved by
// the caller into a permutation (jpvt[c] = 1-based original index of the column now at c) with
// the `nfixed` fixed columns in front; on exit the final permutation.
template <typename T>
void geqp3(Context& ctx, int64_t m, int64_t n, T* A, int64_t lda, int* d_jpvt, T* d_tau,
           int64_t nfixed) {
  const int64_t k = m < n ? m : n;
  if (k <= 0) return;
  Arena::Mark mk = ctx.arena.mark();
  double* vn1 = ctx.alloc<double>(n);
  double* vn2 = ctx.alloc<double>(n);
  const double eps = sizeof(typename Traits<T>::real) == 4 ? 5.9604644775390625e-08 : 1.1102230246251565e-16;
  const double tol3z = sqrt(eps);
  const int64_t gn = n < (int64_t)ctx.sm_count * 4 ? n : (int64_t)ctx.sm_count * 4;
  qp3_norms_kernel<T><<<(unsigned)gn, QP_NT, 0, ctx.stream>>>(m, n, A, lda, vn1, vn2);
  for (int64_t j = 0; j < k; j++) {
    qp3_pivot_kernel<T><<<1, QP_NT, 0, ctx.stream>>>(m, n, j, nfixed, A, lda, d_jpvt, d_tau, vn1, vn2);
    const int64_t nc = n - j - 1;
    if (nc > 0) {
      const int64_t g = nc < (int64_t)ctx.sm_count * 8 ? nc : (int64_t)ctx.sm_count * 8;
      qp3_apply_kernel<T><<<(unsigned)g, 256, 0, ctx.stream>>>(m, n, j, A, lda, d_tau, vn1, vn2, tol3z);
    }
  }
  LA_CUDA_CHECK(cudaGetLastError());
  ctx.arena.release(mk);
}

template <typename T>
void swap_columns(Context& ctx, int64_t m, T* a, T* b) {
  if (m <= 0 || a == b) return;
  swap_columns_kernel<T><<<(unsigned)ceil_div(m, 256), 256, 0, ctx.stream>>>(m, a, b);
  LA_CUDA_CHECK(cudaGetLastError());
}

#define LA_INST_QR(T)                                                                            \
  template void geqrf<T>(Context&, int64_t, int64_t, T*, int64_t, T*);                           \
  template void ormqr<T>(Context&, Side, Op, int64_t, int64_t, int64_t, const T*, int64_t,       \
                         const T*, T*, int64_t);                                                 \
  template void orgqr<T>(Context&, int64_t, int64_t, int64_t, T*, int64_t, const T*);            \
  template void gels<T>(Context&, Op, int64_t, int64_t, int64_t, T*, int64_t, T*, int64_t, int*); \
  template void geqrf_variant<T>(Context&, char, int64_t, int64_t, T*, int64_t, T*);             \
  template void orgqr_variant<T>(Context&, char, int64_t, int64_t, int64_t, T*, int64_t,         \
                                 const T*);                                                      \
  template void geqp3<T>(Context&, int64_t, int64_t, T*, int64_t, int*, T*, int64_t);            \
  template void swap_columns<T>(Context&, int64_t, T*, T*);
LA_INST_QR(float)
LA_INST_QR(double)
LA_INST_QR(zcomplex)

}  // namespace la

#ifdef LA_PANEL_TRACE
extern "C" void la_b200_debug_qr_trace(long long* out) {
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(out, la::g_qr_trace, sizeof(long long) * 64 * 8);
}
#endif
