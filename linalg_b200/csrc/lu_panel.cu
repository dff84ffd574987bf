// la_b200 — LU panel kernel: partial-pivoting LU of an m x jb (jb <= 32) panel in ONE
// cooperative launch (the leaf of ?getrf, src/linalg/lu.cr:74-80 -> ?getrf_).
//
// The panel is distributed by contiguous row slices over G CTAs and lives in their shared
// memory for the whole launch.  Partial pivoting needs one global decision per column; the
// serial chain of N such decisions bounds ?getrf, so the exchange is built for latency:
//   * every CTA publishes its best candidate (|value|, row index) together with the candidate
//     row's current contents; the owner of row j publishes row j;
//   * publication uses 16-byte (payload, tag) packets written with single vector stores and
//     polled by the readers ("LL" protocol): a packet is valid when its tag equals the step's
//     sequence number, so there is no barrier, no fence and no atomic on the critical path —
//     one L2 write->read per column;
//   * all CTAs derive the same winner (max |a|, ties -> smallest row index = i?amax's "first
//     maximal element"), the two owners swap rows in place, everyone scales and updates its
//     rows and folds the arg-max of the next column into the same pass.
// Buffers are double-buffered by column parity; tags grow monotonically across launches, so
// the exchange area (persistent per context) never needs clearing.
// Bytes per column step (algorithmic): (jb-j) * rows * sizeof(T) read+written in SHARED memory;
// HBM traffic is one read and one write of the panel per launch: 2*m*jb*sizeof(T).
#include "la_internal.h"
#include <new>

namespace la {

namespace {

constexpr int LU_NT = 512;
constexpr int LU_MAX_JB = 32;
constexpr int LU_MAX_CTAS = 160;
constexpr size_t LU_SMEM_BUDGET = 136 * 1024;   // row slice; the landing areas come on top
constexpr size_t LU_SMEM_TOTAL = 224 * 1024;

struct __align__(16) LLPacket {
  unsigned long long data;
  unsigned long long tag;
};

template <typename T> struct Words { static constexpr int N = 1; };
template <> struct Words<zcomplex> { static constexpr int N = 2; };
// widest leaf per element type (the row is staged in registers during the update)
template <typename T> struct MaxJb { static constexpr int N = LU_MAX_JB; };
template <> struct MaxJb<zcomplex> { static constexpr int N = 16; };

struct LuExchange {
  LLPacket hdr[2][LU_MAX_CTAS];                       // data = |a| (double bits); tag = seq<<32 | row
  LLPacket row[2][LU_MAX_CTAS][LU_MAX_JB * 2];        // candidate row contents
  LLPacket rowj[2][LU_MAX_JB * 2];                    // contents of row j
};

__device__ __forceinline__ void ll_store(LLPacket* p, unsigned long long data,
                                         unsigned long long tag) {
  asm volatile("st.global.cg.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(data), "l"(tag) : "memory");
}
__device__ __forceinline__ void ll_load(const LLPacket* p, unsigned long long& data,
                                        unsigned long long& tag) {
  asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(data), "=l"(tag) : "l"(p) : "memory");
}
// spin until the packet carries `seq` in the upper half of its tag
__device__ __forceinline__ void ll_wait(const LLPacket* p, unsigned seq, unsigned long long& data,
                                        unsigned long long& tag) {
  do { ll_load(p, data, tag); } while ((unsigned)(tag >> 32) != seq);
}

template <typename T> __device__ __forceinline__ void ll_put(LLPacket* p, T v, unsigned seq);
template <> __device__ __forceinline__ void ll_put<double>(LLPacket* p, double v, unsigned seq) {
  ll_store(p, (unsigned long long)__double_as_longlong(v), (unsigned long long)seq << 32);
}
template <> __device__ __forceinline__ void ll_put<float>(LLPacket* p, float v, unsigned seq) {
  ll_store(p, (unsigned long long)__float_as_uint(v), (unsigned long long)seq << 32);
}
template <> __device__ __forceinline__ void ll_put<zcomplex>(LLPacket* p, zcomplex v, unsigned seq) {
  ll_store(p, (unsigned long long)__double_as_longlong(v.x), (unsigned long long)seq << 32);
  ll_store(p + 1, (unsigned long long)__double_as_longlong(v.y), (unsigned long long)seq << 32);
}
template <typename T> __device__ __forceinline__ T ll_get(const LLPacket* p, unsigned seq);
template <> __device__ __forceinline__ double ll_get<double>(const LLPacket* p, unsigned seq) {
  unsigned long long d, t;
  ll_wait(p, seq, d, t);
  return __longlong_as_double((long long)d);
}
template <> __device__ __forceinline__ float ll_get<float>(const LLPacket* p, unsigned seq) {
  unsigned long long d, t;
  ll_wait(p, seq, d, t);
  return __uint_as_float((unsigned)d);
}
template <> __device__ __forceinline__ zcomplex ll_get<zcomplex>(const LLPacket* p, unsigned seq) {
  unsigned long long d, t;
  zcomplex v;
  ll_wait(p, seq, d, t);
  v.x = __longlong_as_double((long long)d);
  ll_wait(p + 1, seq, d, t);
  v.y = __longlong_as_double((long long)d);
  return v;
}

// candidates are valid iff their index is; NaNs never win a comparison but never dead-lock
__device__ __forceinline__ void argmax_combine(double& v, int& i, double ov, int oi) {
  if (oi != 0x7fffffff && (i == 0x7fffffff || ov > v || (ov == v && oi < i))) { v = ov; i = oi; }
}
__device__ __forceinline__ void warp_argmax(double& v, int& i) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    double ov = __shfl_xor_sync(0xffffffffu, v, o);
    int oi = __shfl_xor_sync(0xffffffffu, i, o);
    argmax_combine(v, i, ov, oi);
  }
}

template <typename T>
__global__ void __launch_bounds__(LU_NT)
lu_panel_kernel(int m, int jb, T* __restrict__ A, int64_t lda, int* __restrict__ ipiv, int row0,
                int* info, LuExchange* xch, int rpc, unsigned seq_base) {
  using R = typename Traits<T>::real;
  constexpr int W = Words<T>::N;
  constexpr int NW = LU_NT / 32;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int G = gridDim.x;
  T* S = reinterpret_cast<T*>(smem_raw);         // S[c * rpc + r]: this CTA's row slice
  T* rows_all = S + (size_t)jb * rpc;            // rows_all[g * jb + c]: every CTA's candidate row
  T* rj = rows_all + (size_t)G * jb;             // rj[c]: row j of the panel
  double* hv = reinterpret_cast<double*>(((uintptr_t)(rj + LU_MAX_JB) + 7) & ~(uintptr_t)7);  // headers
  int* hi = reinterpret_cast<int*>(hv + G);
  __shared__ double sval[NW];
  __shared__ int sidx[NW];
  __shared__ int s_pg;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int rs = blockIdx.x * rpc;
  const int nr = max(0, min(m - rs, rpc));

  // each thread stages its rows with several independent loads in flight (coalesced over r)
  for (int r = tid; r < nr; r += LU_NT) {
    const T* src = A + (int64_t)(rs + r);
#pragma unroll 8
    for (int c = 0; c < jb; c++) S[c * rpc + r] = src[(int64_t)c * lda];
  }
  // arg-max of column 0 over owned rows (each thread only ever touches its own rows of S)
  double bv = -1.0;
  int bi = 0x7fffffff;
  for (int r = tid; r < nr; r += LU_NT) {
    double v = (double)abs1(S[r]);
    if (v > bv || bi == 0x7fffffff) { bv = v; bi = rs + r; }  // NaN-safe: first row always taken
  }

  for (int j = 0; j < jb; j++) {
    const int par = j & 1;
    const unsigned seq = seq_base + (unsigned)j + 1u;
    // ---- CTA-level arg-max: warps reduce, warp 0 finishes and publishes
    warp_argmax(bv, bi);
    if (lane == 0) { sval[warp] = bv; sidx[warp] = bi; }
    __syncthreads();
    if (warp == 0) {
      double cv = lane < NW ? sval[lane] : -2.0;
      int ci = lane < NW ? sidx[lane] : 0x7fffffff;
      warp_argmax(cv, ci);
      const bool have = ci != 0x7fffffff;
      const T crow = (have && lane < jb) ? S[lane * rpc + (ci - rs)] : zero<T>();
      const bool own_j = j >= rs && j < rs + nr;
      if (G == 1) {  // single CTA: no global exchange
        if (lane == 0) { hv[0] = cv; hi[0] = ci; }
        if (lane < jb) { rows_all[lane] = crow; rj[lane] = S[lane * rpc + (j - rs)]; }
      } else {
        if (lane == 0)
          ll_store(&xch->hdr[par][blockIdx.x], (unsigned long long)__double_as_longlong(cv),
                   ((unsigned long long)seq << 32) | (unsigned)(have ? ci : 0xffffffffu));
        if (lane < jb) ll_put<T>(&xch->row[par][blockIdx.x][lane * W], crow, seq);
        if (own_j && lane < jb) ll_put<T>(&xch->rowj[par][lane * W], S[lane * rpc + (j - rs)], seq);
      }
    }
    // ---- all warps gather: warp w lands the packets of CTAs w, w+NW, ... in shared memory
    if (G > 1) {
      for (int g = warp; g < G; g += NW) {
        if (lane < jb) rows_all[g * jb + lane] = ll_get<T>(&xch->row[par][g][lane * W], seq);
        if (lane == 31) {
          unsigned long long d, t;
          ll_wait(&xch->hdr[par][g], seq, d, t);
          hv[g] = __longlong_as_double((long long)d);
          const unsigned uidx = (unsigned)(t & 0xffffffffull);
          hi[g] = uidx == 0xffffffffu ? 0x7fffffff : (int)uidx;
        }
      }
      if (warp == NW - 1 && lane < jb) rj[lane] = ll_get<T>(&xch->rowj[par][lane * W], seq);
    }
    __syncthreads();
    // ---- warp 0 derives the winner from the landed headers, interchanges rows j <-> pi inside
    //      the slices that own them (one lane per column) and posts (pi, pg) for the others
    if (warp == 0) {
      double pv = -1.0;
      int pi = 0x7fffffff, pg = 0;
      for (int g = lane; g < G; g += 32) {
        const double v = hv[g];
        const int idx = hi[g];
        if (idx != 0x7fffffff && (pi == 0x7fffffff || v > pv || (v == pv && idx < pi))) {
          pv = v; pi = idx; pg = g;
        }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const double ov = __shfl_xor_sync(0xffffffffu, pv, o);
        const int oi = __shfl_xor_sync(0xffffffffu, pi, o);
        const int og = __shfl_xor_sync(0xffffffffu, pg, o);
        if (oi != 0x7fffffff && (pi == 0x7fffffff || ov > pv || (ov == pv && oi < pi))) {
          pv = ov; pi = oi; pg = og;
        }
      }
      const T* pr = rows_all + pg * jb;
      if (pi != j && lane < jb) {
        if (pi >= rs && pi < rs + nr) S[lane * rpc + (pi - rs)] = rj[lane];
        if (j >= rs && j < rs + nr) S[lane * rpc + (j - rs)] = pr[lane];
      }
      if (lane == 0) {
        s_pg = pg;
        if (blockIdx.x == 0) ipiv[j] = row0 + pi + 1;
      }
    }
    __syncthreads();
    const T* prow = rows_all + s_pg * jb;
    // ---- scale column j, rank-1 update of the columns to its right, and the arg-max of
    //      column j+1 in the same pass; each thread works on its own rows only
    const T pivot = prow[j];
    const bool zero_pivot = is_zero(pivot);
    if (zero_pivot && blockIdx.x == 0 && tid == 0) atomicCAS(info, 0, row0 + j + 1);
    const bool use_recip = (R)abs1(pivot) >= sfmin<R>();
    const T rinv = zero_pivot ? zero<T>() : recip(pivot);
    bv = -1.0;
    bi = 0x7fffffff;
    for (int r = tid; r < nr; r += LU_NT) {
      const int gr = rs + r;
      if (gr <= j) continue;
      T* row = S + r;
      if (zero_pivot) {  // as ?getf2: nothing to scale, the update is a no-op
        if (j + 1 < jb) {
          const double av = (double)abs1(row[(j + 1) * rpc]);
          if (av > bv || bi == 0x7fffffff) { bv = av; bi = gr; }
        }
        continue;
      }
      const T a = row[j * rpc];
      const T l = use_recip ? mul(a, rinv) : div_(a, pivot);
      row[j * rpc] = l;
      int c = j + 1;
      if (c < jb) {
        const T x = fnma_(l, prow[c], row[c * rpc]);
        row[c * rpc] = x;
        const double av = (double)abs1(x);
        if (av > bv || bi == 0x7fffffff) { bv = av; bi = gr; }
        c++;
      }
      for (; c + 3 < jb; c += 4) {  // loads first, so that four updates are in flight
        const T x0 = row[c * rpc], x1 = row[(c + 1) * rpc], x2 = row[(c + 2) * rpc],
                x3 = row[(c + 3) * rpc];
        const T p0 = prow[c], p1 = prow[c + 1], p2 = prow[c + 2], p3 = prow[c + 3];
        row[c * rpc] = fnma_(l, p0, x0);
        row[(c + 1) * rpc] = fnma_(l, p1, x1);
        row[(c + 2) * rpc] = fnma_(l, p2, x2);
        row[(c + 3) * rpc] = fnma_(l, p3, x3);
      }
      for (; c < jb; c++) row[c * rpc] = fnma_(l, prow[c], row[c * rpc]);
    }
    // landing areas and sval are rewritten only after the next column's first __syncthreads
  }
  __syncthreads();
  for (int r = tid; r < nr; r += LU_NT) {
    T* dst = A + (int64_t)(rs + r);
#pragma unroll 8
    for (int c = 0; c < jb; c++) dst[(int64_t)c * lda] = S[c * rpc + r];
  }
}

}  // namespace

// persistent exchange area + launch sequence, one per context (kernels of one context are
// stream-ordered, so consecutive launches never overlap)
struct LuPanelState {
  LuExchange* xch = nullptr;
  unsigned seq = 64;
};

static LuPanelState* panel_state(Context& ctx) {
  if (!ctx.lu_state) {
    LuPanelState* st = new LuPanelState();
    LA_CUDA_CHECK(cudaMalloc(&st->xch, sizeof(LuExchange)));
    LA_CUDA_CHECK(cudaMemset(st->xch, 0, sizeof(LuExchange)));
    ctx.lu_state = st;
  }
  return (LuPanelState*)ctx.lu_state;
}

// For callers that bring their own exchange memory (getrf_noalloc.cu): size of the area, and a
// way to install caller-owned, zero-initialised device memory as the exchange area of `ctx`
// with host-side bookkeeping in caller-owned storage, so that nothing is allocated here.
size_t lu_panel_exchange_bytes() { return sizeof(LuExchange); }
size_t lu_panel_state_bytes() { return sizeof(LuPanelState); }
void lu_panel_adopt_exchange(Context& ctx, void* device_mem, void* state_storage) {
  LuPanelState* st = new (state_storage) LuPanelState();
  st->xch = (LuExchange*)device_mem;
  ctx.lu_state = st;
}

int lu_panel_max_width(Context& ctx, int64_t m, size_t elem_bytes) {
  int64_t rpc = round_up(ceil_div(m, ctx.sm_count), 2);
  int64_t nb0 = (int64_t)(LU_SMEM_BUDGET / (rpc * elem_bytes));
  const int64_t cap = elem_bytes > 8 ? MaxJb<zcomplex>::N : LU_MAX_JB;
  if (nb0 > cap) nb0 = cap;
  return (int)nb0;
}

template <typename T>
void lu_panel(Context& ctx, int64_t m, int jb, T* A, int64_t lda, int* ipiv, int64_t row0,
              int* d_info) {
  static bool attr = false;
  if (!attr) {
    LA_CUDA_CHECK(cudaFuncSetAttribute(lu_panel_kernel<T>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)LU_SMEM_TOTAL));
    attr = true;
  }
  LuPanelState* st = panel_state(ctx);
  // as few CTAs as keep the slice resident and the per-column work short
  int64_t G = ceil_div(m, 512);
  if (G > ctx.sm_count) G = ctx.sm_count;
  if (G < 1) G = 1;
  int64_t rpc = ceil_div(m, G);
  while ((size_t)rpc * jb * sizeof(T) > LU_SMEM_BUDGET && G < ctx.sm_count) {
    G++;
    rpc = ceil_div(m, G);
  }
  rpc = round_up(rpc, 2);
  G = ceil_div(m, rpc);
  size_t smem = (size_t)rpc * jb * sizeof(T);
  if (smem > LU_SMEM_BUDGET || G > LU_MAX_CTAS) {
    fprintf(stderr, "la_b200: internal error: LU panel %lld x %d does not fit\n", (long long)m, jb);
    abort();
  }
  // landing areas: every CTA's candidate row, row j, headers
  smem += ((size_t)G * jb + LU_MAX_JB) * sizeof(T) + (size_t)G * (sizeof(double) + sizeof(int)) + 16;
  int mi = (int)m, row0i = (int)row0, rpci = (int)rpc;
  unsigned seq_base = st->seq;
  st->seq += 64;
  LuExchange* xch = st->xch;
  void* args[] = {&mi, &jb, &A, &lda, &ipiv, &row0i, &d_info, &xch, &rpci, &seq_base};
  LA_CUDA_CHECK(cudaLaunchCooperativeKernel((void*)lu_panel_kernel<T>, dim3((unsigned)G),
                                            dim3(LU_NT), args, smem, ctx.stream));
}

template void lu_panel<float>(Context&, int64_t, int, float*, int64_t, int*, int64_t, int*);
template void lu_panel<double>(Context&, int64_t, int, double*, int64_t, int*, int64_t, int*);
template void lu_panel<zcomplex>(Context&, int64_t, int, zcomplex*, int64_t, int*, int64_t, int*);

}  // namespace la
