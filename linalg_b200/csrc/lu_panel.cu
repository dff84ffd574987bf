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
// Buffers are double-buffered by column parity; tags grow monotonically from launch to launch
// within one factorisation; every factorisation clears the area and restarts the sequence
// (lu_panel_begin), so a call enqueues the same work whatever ran before it and the 32-bit
// sequence cannot wrap.
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

// Pivot candidates are compared through integer keys: for non-negative doubles the IEEE bit
// pattern orders like the value, and integer compares are several times shorter than FP64 ones
// on the serial path of every column.  NaN magnitudes get key 0 (they never win, as in i?amax,
// where every comparison with a NaN is false).  A candidate is valid iff its index is.
typedef unsigned long long ukey;
__device__ __forceinline__ ukey key_of(double absval) {
  const ukey k = (ukey)__double_as_longlong(absval);
  return k > 0x7ff0000000000000ull ? 0ull : k;
}
__device__ __forceinline__ void argmax_combine(ukey& v, int& i, ukey ov, int oi) {
  if (oi != 0x7fffffff && (i == 0x7fffffff || ov > v || (ov == v && oi < i))) { v = ov; i = oi; }
}
__device__ __forceinline__ void warp_argmax(ukey& v, int& i) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const ukey ov = __shfl_xor_sync(0xffffffffu, v, o);
    const int oi = __shfl_xor_sync(0xffffffffu, i, o);
    argmax_combine(v, i, ov, oi);
  }
}

#ifdef LA_PANEL_TRACE
__device__ long long g_panel_trace[64 * 8];
#define TRACE(slot) do { if (blockIdx.x == 0 && threadIdx.x == 0 && j < 64) g_panel_trace[j * 8 + (slot)] = clock64(); } while (0)
#else
#define TRACE(slot) do { } while (0)
#endif

// 64-bit key arg-max over a warp with three REDUX steps (max of the high word, max of the low
// word among the lanes that tie on it, smallest index among the lanes that tie on both).
__device__ __forceinline__ void warp_argmax_redux(ukey& v, int& i) {
  const unsigned hi = (unsigned)(v >> 32), lo = (unsigned)v;
  const bool valid = i != 0x7fffffff;
  const unsigned mh = __reduce_max_sync(0xffffffffu, valid ? hi : 0u);
  const bool t1 = valid && hi == mh;
  const unsigned ml = __reduce_max_sync(0xffffffffu, t1 ? lo : 0u);
  const bool t2 = t1 && lo == ml;
  const unsigned mi = __reduce_min_sync(0xffffffffu, t2 ? (unsigned)i : 0x7fffffffu);
  v = ((ukey)mh << 32) | ml;
  i = (int)mi;
}

// One elimination step on a row held in registers: l = a[J] / pivot, a[c] -= l * prow[c] for
// c > J, and the key of the next column's entry.  J is a compile-time constant so that a[]
// never leaves the register file.
template <typename T, int JBM, int J>
__device__ __forceinline__ void eliminate(T (&a)[JBM], const T* __restrict__ prow, T rinv, T pivot,
                                          bool use_recip, ukey& next_key) {
  const T l = use_recip ? mul(a[J], rinv) : div_(a[J], pivot);
  a[J] = l;
#pragma unroll
  for (int c = J + 1; c < JBM; c++) a[c] = fnma_(l, prow[c], a[c]);
  if (J + 1 < JBM) next_key = key_of((double)abs1(a[J + 1 < JBM ? J + 1 : J]));
}

template <typename T, int JBM, int J>
struct Dispatch {
  __device__ __forceinline__ static void run(int j, T (&a)[JBM], const T* prow, T rinv, T pivot,
                                             bool use_recip, ukey& nk) {
    if (j == J) eliminate<T, JBM, J>(a, prow, rinv, pivot, use_recip, nk);
    else Dispatch<T, JBM, J + 1>::run(j, a, prow, rinv, pivot, use_recip, nk);
  }
};
template <typename T, int JBM>
struct Dispatch<T, JBM, JBM> {
  __device__ __forceinline__ static void run(int, T (&)[JBM], const T*, T, T, bool, ukey&) {}
};

template <typename T, int JBM, int J>
struct KeyAt {
  __device__ __forceinline__ static ukey get(int j, const T (&a)[JBM]) {
    return j == J ? key_of((double)abs1(a[J])) : KeyAt<T, JBM, J + 1>::get(j, a);
  }
};
template <typename T, int JBM>
struct KeyAt<T, JBM, JBM> {
  __device__ __forceinline__ static ukey get(int, const T (&)[JBM]) { return 0; }
};

// LU with partial pivoting of an m x jb panel (m >= jb, jb <= JBM), R rows per thread (rows
// rs + r * LU_NT + tid of the CTA's slice of R * LU_NT rows), rows in REGISTERS for the whole
// launch (the elimination would otherwise be bound by the shared-memory bandwidth of the few SMs
// that hold the panel).  R * JBM is constant (MaxJb<T>): the common case is one row of 32 columns
// per thread; panels taller than sm_count * LU_NT rows trade leaf width for rows per thread.
template <typename T, int R>
__global__ void __launch_bounds__(LU_NT)
lu_panel_kernel(int m, int jb, T* __restrict__ A, int64_t lda, int* __restrict__ ipiv, int row0,
                int* info, LuExchange* xch, int rpc, unsigned seq_base) {
  using Real = typename Traits<T>::real;
  constexpr int W = Words<T>::N;
  constexpr int NW = LU_NT / 32;
  constexpr int JBM = MaxJb<T>::N / R;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int G = gridDim.x;
  T* cand = reinterpret_cast<T*>(smem_raw);      // cand[w * JBM + c]: each warp's best row
  T* prow2 = cand + NW * JBM;                    // prow2[par][c]: the pivot row of column j
  T* rj2 = prow2 + 2 * JBM;                      // rj2[par][c]: row j of the panel
  __shared__ ukey sval[NW];
  __shared__ int sidx[NW];
  __shared__ int s_pi[2];

  const int tid = threadIdx.x, lane = tid & 31;
  // broadcast from lane 0: tells the compiler that the warp index is warp-uniform, so that the
  // reductions / ballots / shuffles under `if (warp == 0)` are emitted inline and not as
  // re-convergence (WARPSYNC.COLLECTIVE) sequences
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const int rs = blockIdx.x * rpc;                 // rpc == R * LU_NT rows per CTA
  const int gr0 = rs + tid;                        // this thread's rows: gr0 + r * LU_NT

  T a[R][JBM];
  ukey nk[R];
#pragma unroll
  for (int r = 0; r < R; r++) {
    const int gr = gr0 + r * LU_NT;
#pragma unroll
    for (int c = 0; c < JBM; c++)
      a[r][c] = (gr < m && c < jb) ? A[(int64_t)gr + (int64_t)c * lda] : zero<T>();
    nk[r] = gr < m ? key_of((double)abs1(a[r][0])) : 0;
  }

  // Two block barriers per column.  Everything the barriers separate is double-buffered by
  // column parity (prow2, rj2, s_pi, the exchange packets) or is only touched by warp 0 between
  // them (cand, sval, sidx), so column j+1's first phase may start while slower warps are still
  // eliminating column j.
  for (int j = 0; j < jb; j++) {
    const int par = j & 1;
    const unsigned seq = seq_base + (unsigned)j + 1u;
    T* prow = prow2 + par * JBM;
    T* rj = rj2 + par * JBM;
    const bool own_j = j >= rs && j < rs + R * LU_NT;
    TRACE(0);
    // ---- arg-max of column j over the rows gr >= j: per warp, then across warps
    ukey bv = 0;
    int bi = 0x7fffffff;
#pragma unroll
    for (int r = 0; r < R; r++) {   // this thread's best row (rows ascend with r: ties keep the first)
      const int gr = gr0 + r * LU_NT;
      if (gr < m && gr >= j && (bi == 0x7fffffff || nk[r] > bv)) { bv = nk[r]; bi = gr; }
    }
    const int mine = bi;
    warp_argmax_redux(bv, bi);
#pragma unroll
    for (int r = 0; r < R; r++) {
      const int gr = gr0 + r * LU_NT;
      if (gr == bi && mine == bi && gr < m) {   // this lane owns the warp's best row: stage it
#pragma unroll
        for (int c = 0; c < JBM; c++) cand[warp * JBM + c] = a[r][c];
      }
      if (gr == j) {                            // this lane owns row j
#pragma unroll
        for (int c = 0; c < JBM; c++) rj[c] = a[r][c];
      }
    }
    if (lane == 0) { sval[warp] = bv; sidx[warp] = bi; }
    __syncthreads();
    TRACE(1);
    if (warp == 0) {
      ukey cv = lane < NW ? sval[lane] : 0;
      int ci = lane < NW ? sidx[lane] : 0x7fffffff;
      const int myw = lane;
      warp_argmax_redux(cv, ci);
      // the warp whose entry won (smallest warp among equals holds the smallest row)
      const bool have = ci != 0x7fffffff;
      const unsigned wm = __ballot_sync(0xffffffffu, lane < NW && sidx[myw < NW ? myw : 0] == ci && have);
      const int wwin = have ? __ffs(wm) - 1 : 0;
      const T crow = (have && lane < JBM) ? cand[wwin * JBM + lane] : zero<T>();
      if (G == 1) {  // single CTA: no global exchange
        if (lane < JBM) prow[lane] = crow;
        if (lane == 0) { s_pi[par] = ci; ipiv[j] = row0 + ci + 1; }
      } else {
        // publish this CTA's candidate (|value|, row) with the row's contents, and row j
        if (lane == 0)
          ll_store(&xch->hdr[par][blockIdx.x], cv,
                   ((unsigned long long)seq << 32) | (unsigned)(have ? ci : 0xffffffffu));
        if (lane < JBM) ll_put<T>(&xch->row[par][blockIdx.x][lane * W], crow, seq);
        if (own_j && lane < JBM) ll_put<T>(&xch->rowj[par][lane * W], rj[lane], seq);
        TRACE(2);
        // every CTA's header, one (or a few) per lane -> the same winner everywhere
        ukey pv = 0;
        int pi = 0x7fffffff, pg = 0;
        for (int g = lane; g < G; g += 32) {
          unsigned long long d, t;
          ll_wait(&xch->hdr[par][g], seq, d, t);
          const unsigned uidx = (unsigned)(t & 0xffffffffull);
          const int idx = uidx == 0xffffffffu ? 0x7fffffff : (int)uidx;
          if (idx != 0x7fffffff && (pi == 0x7fffffff || d > pv || (d == pv && idx < pi))) {
            pv = d; pi = idx; pg = g;
          }
        }
        int pi_w = pi;
        warp_argmax_redux(pv, pi_w);
        const unsigned gm = __ballot_sync(0xffffffffu, pi == pi_w && pi_w != 0x7fffffff);
        const int src = gm ? __ffs(gm) - 1 : 0;
        pg = __shfl_sync(0xffffffffu, pg, src);
        TRACE(3);
        // only the winner's row is fetched (it was published together with its header)
        if (lane < JBM)
          prow[lane] = pg == (int)blockIdx.x ? crow : ll_get<T>(&xch->row[par][pg][lane * W], seq);
        if (lane == 0) {
          s_pi[par] = pi_w;
          if (blockIdx.x == 0) ipiv[j] = row0 + pi_w + 1;
        }
      }
    } else if (warp == 1 && G > 1 && !own_j) {
      // row j comes from its owner CTA
      if (lane < JBM) rj[lane] = ll_get<T>(&xch->rowj[par][lane * W], seq);
    }
    __syncthreads();
    TRACE(4);
    const int pi = s_pi[par];
    // ---- row interchange j <-> pi in the registers of the two owners
    if (pi != j) {
#pragma unroll
      for (int r = 0; r < R; r++) {
        const int gr = gr0 + r * LU_NT;
        if (gr == pi) {
#pragma unroll
          for (int c = 0; c < JBM; c++) a[r][c] = rj[c];
        } else if (gr == j) {
#pragma unroll
          for (int c = 0; c < JBM; c++) a[r][c] = prow[c];
        }
      }
    }
    // ---- eliminate: scale column j, rank-1 update, key of the next column's entry
    const T pivot = prow[j];
    const bool zero_pivot = is_zero(pivot);
    if (zero_pivot && blockIdx.x == 0 && tid == 0) atomicCAS(info, 0, row0 + j + 1);
    const bool use_recip = (Real)abs1(pivot) >= sfmin<Real>();
    const T rinv = zero_pivot ? zero<T>() : recip(pivot);
#pragma unroll
    for (int r = 0; r < R; r++) {
      const int gr = gr0 + r * LU_NT;
      nk[r] = 0;
      if (gr < m && gr > j) {
        if (!zero_pivot) Dispatch<T, JBM, 0>::run(j, a[r], prow, rinv, pivot, use_recip, nk[r]);
        else nk[r] = KeyAt<T, JBM, 0>::get(j + 1, a[r]);   // as ?getf2: no scaling, no update
      }
    }
    TRACE(5);
  }
#pragma unroll
  for (int r = 0; r < R; r++) {
    const int gr = gr0 + r * LU_NT;
    if (gr < m) {
#pragma unroll
      for (int c = 0; c < JBM; c++)
        if (c < jb) A[(int64_t)gr + (int64_t)c * lda] = a[r][c];
    }
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

// Make sure the context owns its exchange area (allocates on first use).
void lu_panel_prepare(Context& ctx) { (void)panel_state(ctx); }

// Start of a factorisation of an m-row matrix on ctx.stream: stale packets of earlier calls are
// wiped (in stream order) and the launch sequence restarts, so tags never repeat within a call.
void lu_panel_begin(Context& ctx, int64_t m) {
  if (m <= LU_NT) return;          // single-CTA leaves exchange nothing
  LuPanelState* st = panel_state(ctx);
  LA_CUDA_CHECK(cudaMemsetAsync(st->xch, 0, sizeof(LuExchange), ctx.stream));
  st->seq = 64;
}

// Rows per thread needed for an m-row panel (a power of two; 0: too tall even for the narrowest
// leaf).  One row per thread while the panel fits sm_count CTAs of LU_NT rows.
static int lu_rows_per_thread(Context& ctx, int64_t m, int max_jb) {
  const int64_t cap = ctx.sm_count < LU_MAX_CTAS ? ctx.sm_count : LU_MAX_CTAS;
  for (int R = 1; R <= max_jb; R *= 2)
    if (ceil_div(m, (int64_t)R * LU_NT) <= cap) return R;
  return 0;
}

// Widest leaf the panel kernel takes for an m-row panel (0: the panel is too tall for one
// cooperative launch: more than sm_count * LU_NT * MaxJb rows).
int lu_panel_max_width(Context& ctx, int64_t m, size_t elem_bytes) {
  const int base = elem_bytes > 8 ? MaxJb<zcomplex>::N : LU_MAX_JB;
  const int R = lu_rows_per_thread(ctx, m, base);
  return R ? base / R : 0;
}

template <typename T, int R>
static void lu_panel_launch(Context& ctx, int64_t m, int jb, T* A, int64_t lda, int* ipiv,
                            int64_t row0, int* d_info) {
  static DeviceOnce attr;
  if (attr.pending(ctx.device)) {
    LA_CUDA_CHECK(cudaFuncSetAttribute(lu_panel_kernel<T, R>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)LU_SMEM_TOTAL));
    attr.mark(ctx.device);
  }
  LuPanelState* st = panel_state(ctx);
  constexpr int JBM = MaxJb<T>::N / R;
  const int64_t rpc = (int64_t)R * LU_NT;
  const int64_t G = ceil_div(m, rpc);
  if (G > LU_MAX_CTAS || G > ctx.sm_count || jb > JBM) {
    fprintf(stderr, "la_b200: internal error: LU panel %lld x %d does not fit\n", (long long)m, jb);
    abort();
  }
  size_t smem = ((size_t)(LU_NT / 32) * JBM + 4 * (size_t)JBM) * sizeof(T) + 16;
  int mi = (int)m, row0i = (int)row0, rpci = (int)rpc;
  unsigned seq_base = st->seq;
  st->seq += 64;
  LuExchange* xch = st->xch;
  void* args[] = {&mi, &jb, &A, &lda, &ipiv, &row0i, &d_info, &xch, &rpci, &seq_base};
  LA_CUDA_CHECK(cudaLaunchCooperativeKernel((void*)lu_panel_kernel<T, R>, dim3((unsigned)G),
                                            dim3(LU_NT), args, smem, ctx.stream));
}

template <typename T>
void lu_panel(Context& ctx, int64_t m, int jb, T* A, int64_t lda, int* ipiv, int64_t row0,
              int* d_info) {
  constexpr int base = MaxJb<T>::N;
  // the narrowest R that fits the rows AND still takes jb columns
  int R = lu_rows_per_thread(ctx, m, base);
  if (R == 0 || jb > base / R) {
    fprintf(stderr, "la_b200: internal error: LU panel %lld x %d does not fit\n", (long long)m, jb);
    abort();
  }
  switch (R) {
    case 1: lu_panel_launch<T, 1>(ctx, m, jb, A, lda, ipiv, row0, d_info); break;
    case 2: lu_panel_launch<T, 2>(ctx, m, jb, A, lda, ipiv, row0, d_info); break;
    case 4: lu_panel_launch<T, 4>(ctx, m, jb, A, lda, ipiv, row0, d_info); break;
    case 8: lu_panel_launch<T, 8>(ctx, m, jb, A, lda, ipiv, row0, d_info); break;
    default:
      if constexpr (base >= 32) {
        if (R == 16) { lu_panel_launch<T, 16>(ctx, m, jb, A, lda, ipiv, row0, d_info); break; }
        if (R == 32) { lu_panel_launch<T, 32>(ctx, m, jb, A, lda, ipiv, row0, d_info); break; }
      } else {
        if (R == 16) { lu_panel_launch<T, 16>(ctx, m, jb, A, lda, ipiv, row0, d_info); break; }
      }
      fprintf(stderr, "la_b200: internal error: LU panel rows-per-thread %d\n", R);
      abort();
  }
}

#ifdef LA_PANEL_TRACE
extern "C" void la_b200_debug_panel_trace(long long* out) {
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(out, g_panel_trace, sizeof(long long) * 64 * 8);
}
#endif

template void lu_panel<float>(Context&, int64_t, int, float*, int64_t, int*, int64_t, int*);
template void lu_panel<double>(Context&, int64_t, int, double*, int64_t, int*, int64_t, int*);
template void lu_panel<zcomplex>(Context&, int64_t, int, zcomplex*, int64_t, int*, int64_t, int*);

}  // namespace la
