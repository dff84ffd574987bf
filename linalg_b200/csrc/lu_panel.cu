// la_b200 — LU panel kernel: partial-pivoting LU of an m x jb (jb <= 32) panel in ONE
// cooperative launch (the leaf of ?getrf, src/linalg/lu.cr:74-80 -> ?getrf_).
//
// The panel is distributed by contiguous row slices over G CTAs; every thread keeps its row(s)
// in registers for the whole launch.  Partial pivoting needs one global decision per column, and
// the serial chain of N such decisions bounds ?getrf, so everything is arranged around the
// length of that chain:
//   * rows never move.  A row carries its current LAPACK position; an interchange swaps two
//     positions, the pivot row freezes (it then holds U(j, j:)) and every row is written to its
//     final position once, at the end;
//   * the registers of a row rotate by one column per step (the rank-1 update writes column c
//     into the register of column c-1), so the current column is always register 0 and the code
//     of a step does not depend on the column: no dispatch, nothing that is executed only once;
//     finished L entries go to shared memory;
//   * the update of column j is split: its head (the entries of column j+1, all the next
//     arg-max needs) runs as soon as the pivot row is known, its tail runs while warp 0 waits
//     for the other CTAs;
//   * the CTA first agrees on its candidate, then only that row is staged (a single lane
//     writing a row costs as much as a full warp doing so); a row staged while its tail is
//     pending is completed by warp 0, one entry per lane;
//   * every CTA publishes its candidate (|value|, position) together with the row's contents in
//     16-byte (payload, tag) packets written with single vector stores and polled by the readers
//     ("LL" protocol): a packet is valid when its tag equals the step's sequence number, so there
//     is no barrier, no fence and no atomic on the critical path — one L2 write->read per column;
//     1 / pivot is computed by the candidate's CTA after the row has left and travels behind it;
//   * all CTAs derive the same winner (max |a|, ties -> smallest position = i?amax's "first
//     maximal element").
// Buffers are double-buffered by column parity; tags grow monotonically from launch to launch
// within one factorisation; every factorisation clears the area and restarts the sequence
// (lu_panel_begin), so a call enqueues the same work whatever ran before it and the 32-bit
// sequence cannot wrap.
// Bytes per column step (algorithmic): (jb-j) * rows * sizeof(T) read+written in registers;
// HBM traffic is one read and one write of the panel per launch: 2*m*jb*sizeof(T).
#include "la_internal.h"
#include <new>

namespace la {

namespace {

constexpr int LU_NT = 512;        // threads (= rows per CTA with one row per thread) of the standard leaf
constexpr int LU_NT_WIDE = 256;   // the wide leaf: half the threads, rows twice as long
constexpr int LU_WIDE_MAX_CTAS = 32;
constexpr int LU_MAX_JB = 32;
constexpr int LU_MAX_CTAS = 160;
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
  LLPacket hdr[2][LU_MAX_CTAS];                       // data = |a| (double bits); tag = seq<<32 | position
  LLPacket row[2][LU_MAX_CTAS * LU_MAX_JB];           // candidate rows, columns j.. (rotated): CTA g's
                                                      // starts at packet g * (leaf width in words)
  LLPacket rinv[2][LU_MAX_CTAS][2];                   // 1 / (the candidate's entry in column j)
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
template <typename T> __device__ __forceinline__ T from_words(const unsigned long long* d);
template <> __device__ __forceinline__ double from_words<double>(const unsigned long long* d) {
  return __longlong_as_double((long long)d[0]);
}
template <> __device__ __forceinline__ float from_words<float>(const unsigned long long* d) {
  return __uint_as_float((unsigned)d[0]);
}
template <> __device__ __forceinline__ zcomplex from_words<zcomplex>(const unsigned long long* d) {
  zcomplex v;
  v.x = __longlong_as_double((long long)d[0]);
  v.y = __longlong_as_double((long long)d[1]);
  return v;
}
// This lane's EPL entries (lane, lane + 32, ...) of a published row of `width` entries together
// with the reciprocal that follows it: all packets are polled in one loop, so that their L2
// round trips overlap.
template <typename T, int EPL>
__device__ __forceinline__ void ll_get_row(const LLPacket* row, int lane, int width,
                                           const LLPacket* rinvp, unsigned seq, T (&v)[EPL], T& rv) {
  constexpr int W = Words<T>::N;
  constexpr int NP = (EPL + 1) * W;
  unsigned long long d[NP], t[NP];
  bool ok;
  do {
    ok = true;
#pragma unroll
    for (int e = 0; e < EPL; e++) {
      const int c = lane + 32 * e < width ? lane + 32 * e : 0;
#pragma unroll
      for (int w = 0; w < W; w++) ll_load(row + c * W + w, d[e * W + w], t[e * W + w]);
    }
#pragma unroll
    for (int w = 0; w < W; w++) ll_load(rinvp + w, d[EPL * W + w], t[EPL * W + w]);
#pragma unroll
    for (int i = 0; i < NP; i++) ok = ok && (unsigned)(t[i] >> 32) == seq;
  } while (!ok);
#pragma unroll
  for (int e = 0; e < EPL; e++) v[e] = from_words<T>(d + e * W);
  rv = from_words<T>(d + EPL * W);
}

// 16 bytes of a row at a time (the staging store of a candidate row), through an explicit 32-bit
// shared-memory address: with generic pointers the compiler re-derived the shared window
// (S2R SR_CgaCtaId + LEA) in front of every one of the 16 predicated stores, ~40 cycles each on
// the serial path of a column.
__device__ __forceinline__ uint32_t lu_smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
template <typename T> __device__ __forceinline__ void sts16(uint32_t addr, const T* v);
template <> __device__ __forceinline__ void sts16<double>(uint32_t addr, const double* v) {
  asm volatile("st.shared.v2.f64 [%0], {%1, %2};" ::"r"(addr), "d"(v[0]), "d"(v[1]) : "memory");
}
template <> __device__ __forceinline__ void sts16<float>(uint32_t addr, const float* v) {
  asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(v[0]), "f"(v[1]),
               "f"(v[2]), "f"(v[3]) : "memory");
}
template <> __device__ __forceinline__ void sts16<zcomplex>(uint32_t addr, const zcomplex* v) {
  asm volatile("st.shared.v2.f64 [%0], {%1, %2};" ::"r"(addr), "d"(v[0].x), "d"(v[0].y) : "memory");
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

// LU with partial pivoting of an m x jb panel (m >= jb, jb <= JBM), R rows per thread (rows
// rs + r * LU_NT + tid of the CTA's slice of R * LU_NT rows), rows in REGISTERS for the whole
// launch (the elimination would otherwise be bound by the shared-memory bandwidth of the few SMs
// that hold the panel).  R * JBM is constant (MaxJb<T>): the common case is one row of 32 columns
// per thread; panels taller than sm_count * LU_NT rows trade leaf width for rows per thread.
template <typename T, int R, int NT>
__global__ void __launch_bounds__(NT)
lu_panel_kernel(int m, int jb, T* __restrict__ A, int64_t lda, int* __restrict__ ipiv, int row0,
                int* info, LuExchange* xch, int rpc, unsigned seq_base) {
  using Real = typename Traits<T>::real;
  constexpr int W = Words<T>::N;
  constexpr int NW = NT / 32;
  constexpr int JBM = MaxJb<T>::N * (LU_NT / NT) / R;   // R * NT * JBM is the same for all variants
  constexpr int EPL = (JBM + 31) / 32;             // row entries per lane of warp 0
  constexpr int PS = JBM + 2;                      // stride of a pivot-row buffer
  constexpr int VE = 16 / (int)sizeof(T);          // entries per 16-byte store
  constexpr int NONE = 0x7fffffff;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int G = gridDim.x;
  T* Ls = reinterpret_cast<T*>(smem_raw);          // Ls[c * rpc + local row]: finished L entries
  T* cand = Ls + (size_t)JBM * R * NT;          // cand[c]: the CTA's candidate row as staged
  T* prow2 = cand + PS;                            // prow2[par][c]: pivot row of column j, columns j + c
  T* misc = prow2 + 2 * PS;                        // [par]: 1 / pivot; [2]: l of the candidate's pending tail
  __shared__ ukey sval[NW];
  __shared__ int sidx[NW];
  __shared__ int s_pi[2];
  __shared__ int s_cand_raw;                       // the staged candidate lacks the tail of column j-1

  uint32_t cand_s = lu_smem_u32(cand);
  asm volatile("" : "+r"(cand_s));   // opaque: keep it in a register instead of re-deriving it per store
  const int tid = threadIdx.x, lane = tid & 31;
  // broadcast from lane 0: tells the compiler that the warp index is warp-uniform, so that the
  // reductions / ballots / shuffles under `if (warp == 0)` are emitted inline and not as
  // re-convergence (WARPSYNC.COLLECTIVE) sequences
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const int rs = blockIdx.x * rpc;                 // rpc == R * NT rows per CTA
  const int gr0 = rs + tid;                        // this thread's rows: gr0 + r * NT

  T a[R][JBM];     // a[r][c]: column (current column + c) of the row
  T lprev[R];      // the row's multiplier of the column whose tail is pending
  ukey nk[R];      // key of a[r][0]
  int pos[R];      // the row's current position in the panel (LAPACK's row index)
  unsigned act = 0;   // bit r: row r is in play (valid and not yet a pivot row)
  unsigned pend = 0;  // bit r: the tail of the previous column is pending
#pragma unroll
  for (int r = 0; r < R; r++) {
    const int gr = gr0 + r * NT;
#pragma unroll
    for (int c = 0; c < JBM; c++)
      a[r][c] = (gr < m && c < jb) ? A[(int64_t)gr + (int64_t)c * lda] : zero<T>();
    nk[r] = gr < m ? key_of((double)abs1(a[r][0])) : 0;
    lprev[r] = zero<T>();
    pos[r] = gr;
    if (gr < m) act |= 1u << r;
  }
  if (tid < 2 * PS) prow2[tid] = zero<T>();

  for (int j = 0; j < jb; j++) {
    const int par = j & 1;
    const unsigned seq = seq_base + (unsigned)j + 1u;
    T* prow = prow2 + par * PS;
    const T* pp = prow2 + (par ^ 1) * PS;          // pivot row of column j-1 (pending tails)
    TRACE(0);
    // ---- arg-max of column j over the rows in play: per thread, per warp, per CTA
    ukey bv = 0;
    int bi = NONE;
#pragma unroll
    for (int r = 0; r < R; r++)
      if (((act >> r) & 1u) && (bi == NONE || nk[r] > bv || (nk[r] == bv && pos[r] < bi))) {
        bv = nk[r]; bi = pos[r];
      }
    warp_argmax_redux(bv, bi);
    if (lane == 0) { sval[warp] = bv; sidx[warp] = bi; }
    __syncthreads();
    TRACE(1);
    ukey cv = lane < NW ? sval[lane] : 0;
    int ci = lane < NW ? sidx[lane] : NONE;
    warp_argmax_redux(cv, ci);                      // every warp: the CTA's candidate
    // ---- its owner stages it as it is (registers 0..JBM-1 and the pending multiplier)
#pragma unroll
    for (int r = 0; r < R; r++)
      if (((act >> r) & 1u) && pos[r] == ci) {
        // (columns j .. jb-1 only, one register further with a pending tail)
        const int live = jb - j + 1;
        if (JBM >= VE) {
#pragma unroll
          for (int c = 0; c + VE <= JBM; c += VE)
            if (c < live) sts16<T>(cand_s + c * (uint32_t)sizeof(T), &a[r][c]);
        } else {
#pragma unroll
          for (int c = 0; c < JBM; c++) cand[c] = a[r][c];
        }
        misc[2] = lprev[r];
        s_cand_raw = (int)((pend >> r) & 1u);
      }
    __syncthreads();
    TRACE(2);
    if (warp == 0) {
      const bool have = ci != NONE;
      // the candidate in the coordinates of column j: entry 0 is final (the head wrote it); with
      // a pending tail entry c >= 1 is still in register c+1 and lacks l * (pivot row j-1)[c+1]
      T crow[EPL];
#pragma unroll
      for (int e = 0; e < EPL; e++) {
        const int c = lane + 32 * e;
        crow[e] = zero<T>();
        if (have && c < JBM) {
          if (!s_cand_raw || c == 0) crow[e] = cand[c];
          else if (c + 1 < JBM) crow[e] = fnma_(misc[2], pp[c + 1], cand[c + 1]);
        }
      }
      const T cpiv = have ? cand[0] : zero<T>();
      // a zero pivot leaves the trailing columns alone (?getf2): its row goes out as zeros
      if (is_zero(cpiv)) {
#pragma unroll
        for (int e = 0; e < EPL; e++)
          if (lane + 32 * e >= 1) crow[e] = zero<T>();
      }
      TRACE(7);
      if (G == 1) {  // single CTA: no global exchange
#pragma unroll
        for (int e = 0; e < EPL; e++)
          if (lane + 32 * e < JBM) prow[lane + 32 * e] = crow[e];
        const T crinv = is_zero(cpiv) ? zero<T>() : recip(cpiv);
        if (lane == 0) { s_pi[par] = ci; ipiv[j] = row0 + ci + 1; misc[par] = crinv; }
      } else {
        // publish this CTA's candidate (|value|, position) with the row's contents; the
        // reciprocal is worked out once they are on their way
        if (lane == 0)
          ll_store(&xch->hdr[par][blockIdx.x], cv,
                   ((unsigned long long)seq << 32) | (unsigned)(have ? ci : 0xffffffffu));
        LLPacket* mine = &xch->row[par][(size_t)blockIdx.x * (JBM * W)];
#pragma unroll
        for (int e = 0; e < EPL; e++)
          if (lane + 32 * e < JBM) ll_put<T>(&mine[(lane + 32 * e) * W], crow[e], seq);
        TRACE(3);
        const T crinv = is_zero(cpiv) ? zero<T>() : recip(cpiv);
        if (lane == 0) ll_put<T>(&xch->rinv[par][blockIdx.x][0], crinv, seq);
      }
    }
    // ---- pending tails (column j-1), while the packets travel
#pragma unroll
    for (int r = 0; r < R; r++)
      if ((pend >> r) & 1u) {
#pragma unroll
        for (int c = 2; c < JBM; c++) a[r][c - 1] = fnma_(lprev[r], pp[c], a[r][c]);
        a[r][JBM - 1] = zero<T>();
      }
    pend = 0;
    if (warp == 0 && G > 1) {
      // every CTA's header, one (or a few) per lane -> the same winner everywhere
      ukey pv = 0;
      int pi = NONE, pg = 0;
      for (int g = lane; g < G; g += 32) {
        unsigned long long d, t;
        ll_wait(&xch->hdr[par][g], seq, d, t);
        const unsigned uidx = (unsigned)(t & 0xffffffffull);
        const int idx = uidx == 0xffffffffu ? NONE : (int)uidx;
        if (idx != NONE && (pi == NONE || d > pv || (d == pv && idx < pi))) {
          pv = d; pi = idx; pg = g;
        }
      }
      int pi_w = pi;
      warp_argmax_redux(pv, pi_w);
      const unsigned gm = __ballot_sync(0xffffffffu, pi == pi_w && pi_w != NONE);
      const int src = gm ? __ffs(gm) - 1 : 0;
      pg = __shfl_sync(0xffffffffu, pg, src);
      TRACE(4);
      // only the winner's row is fetched (it was published together with its header), and the
      // reciprocal that followed it
      T prv[EPL], rv;
      ll_get_row<T, EPL>(&xch->row[par][(size_t)pg * (JBM * W)], lane, JBM, &xch->rinv[par][pg][0],
                         seq, prv, rv);
#pragma unroll
      for (int e = 0; e < EPL; e++)
        if (lane + 32 * e < JBM) prow[lane + 32 * e] = prv[e];
      if (lane == 0) {
        misc[par] = rv;
        s_pi[par] = pi_w;
        if (blockIdx.x == 0) ipiv[j] = row0 + pi_w + 1;
      }
    }
    __syncthreads();
    TRACE(5);
    // ---- the interchange (positions only), then the head of the elimination: scale column j,
    //      update column j+1 (into register 0), key of its entry
    const int pi = s_pi[par];
    const T pivot = prow[0];
    const T p1 = prow[1];
    const T rinv = misc[par];
    const bool zero_pivot = is_zero(pivot);
    if (zero_pivot && blockIdx.x == 0 && tid == 0) atomicCAS(info, 0, row0 + j + 1);
    const bool use_recip = (Real)abs1(pivot) >= sfmin<Real>();
#pragma unroll
    for (int r = 0; r < R; r++)
      if ((act >> r) & 1u) {
        if (pos[r] == pi) {           // the pivot row: now at position j, holding U(j, j:)
          pos[r] = j;
          act &= ~(1u << r);
        } else {
          if (pos[r] == j) pos[r] = pi;
          const T l = zero_pivot ? a[r][0]
                                 : (use_recip ? mul(a[r][0], rinv) : div_(a[r][0], pivot));
          Ls[(size_t)j * rpc + r * NT + tid] = l;
          lprev[r] = l;
          if (JBM > 1) {
            a[r][0] = fnma_(l, p1, a[r][JBM > 1 ? 1 : 0]);
            nk[r] = key_of((double)abs1(a[r][0]));
            pend |= 1u << r;
          }
        }
      }
    TRACE(6);
  }
  // ---- every row goes to its final position: L entries from shared memory; a pivot row (frozen
  //      at column p = its position) adds U(p, p:) from its registers.  The tail still pending here
  //      belongs to column jb-1 and would touch columns >= jb only.
#pragma unroll
  for (int r = 0; r < R; r++) {
    const int gr = gr0 + r * NT;
    if (gr < m) {
      const int p = pos[r];
      const bool frozen = !((act >> r) & 1u);
      const int nl = frozen ? p : jb;
      T* dst = A + (int64_t)p;
      for (int c = 0; c < nl; c++) dst[(int64_t)c * lda] = Ls[(size_t)c * rpc + r * NT + tid];
      if (frozen) {
#pragma unroll
        for (int c = 0; c < JBM; c++)
          if (p + c < jb) dst[(int64_t)(p + c) * lda] = a[r][c];
      }
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
  if (m <= LU_NT_WIDE) return;     // single-CTA leaves exchange nothing
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

// The wide leaf (twice the columns on half the threads per CTA) halves the number of leaves and
// removes the interchange / TRSM / GEMM launches between the two halves; it needs twice the CTAs
// per row, so it is for panels that are short enough to leave the other SMs to the trailing
// update.  LA_B200_LU_WIDE=0 turns it off.
bool lu_panel_wide_ok(int64_t m) {
  static const bool enabled = !(getenv("LA_B200_LU_WIDE") && atoi(getenv("LA_B200_LU_WIDE")) == 0);
  return enabled && ceil_div(m, (int64_t)LU_NT_WIDE) <= LU_WIDE_MAX_CTAS;
}

// Widest leaf the panel kernel takes for an m-row panel (0: the panel is too tall for one
// cooperative launch: more than sm_count * LU_NT * MaxJb rows).
int lu_panel_max_width(Context& ctx, int64_t m, size_t elem_bytes, bool wide) {
  const int base = elem_bytes > 8 ? MaxJb<zcomplex>::N : LU_MAX_JB;
  if (wide && lu_panel_wide_ok(m)) return base * (LU_NT / LU_NT_WIDE);
  const int R = lu_rows_per_thread(ctx, m, base);
  return R ? base / R : 0;
}

// CTAs (= SMs) the leaves of an m-row panel occupy.
int lu_panel_ctas(Context& ctx, int64_t m, size_t elem_bytes, bool wide) {
  if (wide && lu_panel_wide_ok(m)) return (int)ceil_div(m, (int64_t)LU_NT_WIDE);
  const int base = elem_bytes > 8 ? MaxJb<zcomplex>::N : LU_MAX_JB;
  const int R = lu_rows_per_thread(ctx, m, base);
  return R ? (int)ceil_div(m, (int64_t)R * LU_NT) : 0;
}

template <typename T, int R, int NT>
static void lu_panel_launch(Context& ctx, int64_t m, int jb, T* A, int64_t lda, int* ipiv,
                            int64_t row0, int* d_info) {
  static DeviceOnce attr;
  if (attr.pending(ctx.device)) {
    LA_CUDA_CHECK(cudaFuncSetAttribute(lu_panel_kernel<T, R, NT>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)LU_SMEM_TOTAL));
    attr.mark(ctx.device);
  }
  LuPanelState* st = panel_state(ctx);
  constexpr int JBM = MaxJb<T>::N * (LU_NT / NT) / R;
  static_assert(JBM * Words<T>::N <= 2 * LU_MAX_JB, "row packets");
  const int64_t rpc = (int64_t)R * NT;
  const int64_t G = ceil_div(m, rpc);
  if (G > LU_MAX_CTAS || G > ctx.sm_count || jb > JBM ||
      G * JBM * Words<T>::N > (int64_t)LU_MAX_CTAS * LU_MAX_JB) {
    fprintf(stderr, "la_b200: internal error: LU panel %lld x %d does not fit\n", (long long)m, jb);
    abort();
  }
  size_t smem = ((size_t)MaxJb<T>::N * LU_NT + 3 * (size_t)(JBM + 2) + 4) * sizeof(T) + 16;
  int mi = (int)m, row0i = (int)row0, rpci = (int)rpc;
  unsigned seq_base = st->seq;
  st->seq += 128;
  LuExchange* xch = st->xch;
  void* args[] = {&mi, &jb, &A, &lda, &ipiv, &row0i, &d_info, &xch, &rpci, &seq_base};
  LA_CUDA_CHECK(cudaLaunchCooperativeKernel((void*)lu_panel_kernel<T, R, NT>, dim3((unsigned)G),
                                            dim3(NT), args, smem, ctx.stream));
}

template <typename T>
void lu_panel(Context& ctx, int64_t m, int jb, T* A, int64_t lda, int* ipiv, int64_t row0,
              int* d_info) {
  constexpr int base = MaxJb<T>::N;
  if (jb > base) {   // only the wide leaf takes this many columns
    if (!lu_panel_wide_ok(m) || jb > base * (LU_NT / LU_NT_WIDE)) {
      fprintf(stderr, "la_b200: internal error: LU panel %lld x %d does not fit\n", (long long)m, jb);
      abort();
    }
    lu_panel_launch<T, 1, LU_NT_WIDE>(ctx, m, jb, A, lda, ipiv, row0, d_info);
    return;
  }
  // the narrowest R that fits the rows AND still takes jb columns
  int R = lu_rows_per_thread(ctx, m, base);
  if (R == 0 || jb > base / R) {
    fprintf(stderr, "la_b200: internal error: LU panel %lld x %d does not fit\n", (long long)m, jb);
    abort();
  }
  switch (R) {
    case 1: lu_panel_launch<T, 1, LU_NT>(ctx, m, jb, A, lda, ipiv, row0, d_info); break;
    case 2: lu_panel_launch<T, 2, LU_NT>(ctx, m, jb, A, lda, ipiv, row0, d_info); break;
    case 4: lu_panel_launch<T, 4, LU_NT>(ctx, m, jb, A, lda, ipiv, row0, d_info); break;
    case 8: lu_panel_launch<T, 8, LU_NT>(ctx, m, jb, A, lda, ipiv, row0, d_info); break;
    default:
      if constexpr (base >= 32) {
        if (R == 16) { lu_panel_launch<T, 16, LU_NT>(ctx, m, jb, A, lda, ipiv, row0, d_info); break; }
        if (R == 32) { lu_panel_launch<T, 32, LU_NT>(ctx, m, jb, A, lda, ipiv, row0, d_info); break; }
      } else {
        if (R == 16) { lu_panel_launch<T, 16, LU_NT>(ctx, m, jb, A, lda, ipiv, row0, d_info); break; }
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
