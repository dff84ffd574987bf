// Micro-benchmark of the primitives on the serial path of the panel kernels (B200).
#include <cstdio>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s line %d\n", cudaGetErrorString(e_), __LINE__); return 1; } } while (0)

__device__ unsigned long long g_pkt[4];

__global__ void lat(long long* out, double* din, int iters) {
  __shared__ double sm[1024];
  __shared__ unsigned long long sk[32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  double x = din[tid % 8] + 1.5, y = din[(tid + 1) % 8] + 2.5;
  unsigned long long key = __double_as_longlong(x) + tid;
  int idx = tid;
  long long t0, t1;
  // (a) dependent DFMA chain
  __syncthreads(); t0 = clock64();
  for (int i = 0; i < iters; i++) x = fma(x, y, 1.0);
  t1 = clock64(); if (tid == 0) out[0] = (t1 - t0) / iters;
  // (b) double division chain
  __syncthreads(); t0 = clock64();
  for (int i = 0; i < iters; i++) x = 1.0 / (x + 3.0);
  t1 = clock64(); if (tid == 0) out[1] = (t1 - t0) / iters;
  // (c) warp argmax on 64-bit integer keys (5 shuffle rounds)
  __syncthreads(); t0 = clock64();
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      unsigned long long ok = __shfl_xor_sync(0xffffffffu, key, o);
      int oi = __shfl_xor_sync(0xffffffffu, idx, o);
      if (ok > key || (ok == key && oi < idx)) { key = ok; idx = oi; }
    }
    key += i;
  }
  t1 = clock64(); if (tid == 0) out[2] = (t1 - t0) / iters;
  // (d) __syncthreads, all warps
  __syncthreads(); t0 = clock64();
  for (int i = 0; i < iters; i++) __syncthreads();
  t1 = clock64(); if (tid == 0) out[3] = (t1 - t0) / iters;
  // (e) smem write -> sync -> read by another warp -> sync
  __syncthreads(); t0 = clock64();
  for (int i = 0; i < iters; i++) {
    sm[tid] = x; __syncthreads();
    x += sm[(tid + 32) % blockDim.x]; __syncthreads();
  }
  t1 = clock64(); if (tid == 0) out[4] = (t1 - t0) / iters;
  // (f) global packet round trip by one thread: st.cg then poll with ld.volatile
  __syncthreads(); t0 = clock64();
  if (tid == 0) {
    for (int i = 1; i <= iters; i++) {
      asm volatile("st.global.cg.v2.u64 [%0], {%1, %2};" ::"l"(g_pkt), "l"((unsigned long long)i), "l"((unsigned long long)i) : "memory");
      unsigned long long a, b;
      do { asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "l"(g_pkt) : "memory"); } while (b != (unsigned long long)i);
    }
  }
  t1 = clock64(); if (tid == 0) out[5] = (t1 - t0) / iters;
  // (g) sqrt chain
  __syncthreads(); t0 = clock64();
  for (int i = 0; i < iters; i++) x = sqrt(x + 2.0);
  t1 = clock64(); if (tid == 0) out[6] = (t1 - t0) / iters;
  // (h) smem LDS -> DFMA -> STS dependent chain (one element)
  __syncthreads(); t0 = clock64();
  for (int i = 0; i < iters; i++) { double v = sm[tid]; v = fma(v, y, x); sm[tid] = v; }
  t1 = clock64(); if (tid == 0) out[7] = (t1 - t0) / iters;
  // (i) warp argmax on 64-bit keys with three REDUX steps
  __syncthreads(); t0 = clock64();
  for (int i = 0; i < iters; i++) {
    const unsigned hi = (unsigned)(key >> 32), lo = (unsigned)key;
    const unsigned mh = __reduce_max_sync(0xffffffffu, hi);
    const bool t1_ = hi == mh;
    const unsigned ml = __reduce_max_sync(0xffffffffu, t1_ ? lo : 0u);
    const bool t2_ = t1_ && lo == ml;
    const unsigned mi = __reduce_min_sync(0xffffffffu, t2_ ? (unsigned)idx : 0x7fffffffu);
    key = (((unsigned long long)mh << 32) | ml) + i + lane;
    idx = (int)mi + lane;
  }
  t1 = clock64(); if (tid == 0) out[8] = (t1 - t0) / iters;
  // (j) one lane stages 32 doubles (16 x STS.128), then a block barrier
  double r[32];
#pragma unroll
  for (int c = 0; c < 32; c++) r[c] = x + c;
  __syncthreads(); t0 = clock64();
  for (int i = 0; i < iters; i++) {
    if (lane == (i & 31)) {
#pragma unroll
      for (int c = 0; c < 32; c++) sm[warp * 32 + c] = r[c] + i;
    }
    __syncthreads();
  }
  t1 = clock64(); if (tid == 0) out[9] = (t1 - t0) / iters;
  // (j2) the same without arithmetic; (j3) one warp only; (j4) 64-bit stores
  __syncthreads(); t0 = clock64();
  for (int i = 0; i < iters; i++) {
    if (lane == (i & 31)) {
#pragma unroll
      for (int c = 0; c < 32; c++) sm[warp * 32 + c] = r[c];
    }
    __syncthreads();
  }
  t1 = clock64(); if (tid == 0) out[13] = (t1 - t0) / iters;
  __syncthreads(); t0 = clock64();
  for (int i = 0; i < iters; i++) {
    if (tid == (i & 31)) {
#pragma unroll
      for (int c = 0; c < 32; c++) sm[warp * 32 + c] = r[c];
    }
    __syncthreads();
  }
  t1 = clock64(); if (tid == 0) out[14] = (t1 - t0) / iters;
  {
    volatile double* vsm = sm;
    __syncthreads(); t0 = clock64();
    for (int i = 0; i < iters; i++) {
      if (lane == (i & 31)) {
#pragma unroll
        for (int c = 0; c < 32; c++) vsm[warp * 32 + c] = r[c];
      }
      __syncthreads();
    }
    t1 = clock64(); if (tid == 0) out[16] = (t1 - t0) / iters;
  }
  // (j5) every lane holds a row; the row of lane `src` goes to shared memory through shuffles:
  // lane c takes element c
  __syncthreads(); t0 = clock64();
  for (int i = 0; i < iters; i++) {
    const int src = i & 31;
    double mine = 0.0;
#pragma unroll
    for (int c = 0; c < 32; c++) {
      const double v = __shfl_sync(0xffffffffu, r[c], src);
      if (lane == c) mine = v;
    }
    sm[warp * 32 + lane] = mine;
    __syncthreads();
  }
  t1 = clock64(); if (tid == 0) out[17] = (t1 - t0) / iters;
  // (k) code executed once: 32 different blocks of ~30 DFMAs each, cold then warm
  for (int pass = 0; pass < 2; pass++) {
    __syncthreads(); t0 = clock64();
    for (int i = 0; i < 32; i++) {
      switch (i) {
#define BLK(J) case J: { _Pragma("unroll") for (int c = 0; c < 32; c++) if (c > J % 3) r[c] = fma(r[(c + J) & 31], y + J, r[c]); } break;
        BLK(0) BLK(1) BLK(2) BLK(3) BLK(4) BLK(5) BLK(6) BLK(7) BLK(8) BLK(9) BLK(10) BLK(11) BLK(12)
        BLK(13) BLK(14) BLK(15) BLK(16) BLK(17) BLK(18) BLK(19) BLK(20) BLK(21) BLK(22) BLK(23) BLK(24)
        BLK(25) BLK(26) BLK(27) BLK(28) BLK(29) BLK(30) BLK(31)
      }
      __syncthreads();
    }
    t1 = clock64(); if (tid == 0) out[10 + pass] = (t1 - t0) / 32;
  }
  // (l) the same amount of FP64 work from ONE block of code, 32 times
  __syncthreads(); t0 = clock64();
  for (int i = 0; i < 32; i++) {
#pragma unroll
    for (int c = 0; c < 32; c++) if (c > 0) r[c] = fma(r[(c + 5) & 31], y, r[c]);
    __syncthreads();
  }
  t1 = clock64(); if (tid == 0) out[12] = (t1 - t0) / 32;
  for (int c = 0; c < 32; c++) x += r[c];
  if (x == 123.0 && key == 5 && idx == 7) out[15] = 1;
  (void)warp; (void)lane; (void)sk;
}

int main() {
  long long* out; double* din; CK(cudaMalloc(&out, 18 * 8)); CK(cudaMalloc(&din, 64)); CK(cudaMemset(din, 0, 64));
  const char* names[] = {"dfma_dep", "ddiv_dep", "warp_argmax_u64", "syncthreads", "smem_xchg_2sync", "global_pkt_roundtrip", "dsqrt_dep", "lds_dfma_sts",
                         "warp_argmax_redux", "stage_row_1lane+sync", "once_code_cold(31dfma+sync)",
                         "once_code_warm", "same_code(31dfma+sync)", "stage_noarith", "stage_one_warp", "unused",
                         "stage_sts64", "stage_shuffle"};
  for (int threads : {32, 512}) {
    lat<<<1, threads>>>(out, din, 200); CK(cudaDeviceSynchronize());
    lat<<<1, threads>>>(out, din, 200); CK(cudaDeviceSynchronize());
    long long h[18]; CK(cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost));
    for (int i = 0; i < 18; i++) if (i != 15) printf("threads=%d %s: %lld cycles\n", threads, names[i], h[i]);
  }
  return 0;
}
