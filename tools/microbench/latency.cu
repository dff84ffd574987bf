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
  if (x == 123.0 && key == 5 && idx == 7) out[15] = 1;
  (void)warp; (void)lane; (void)sk;
}

int main() {
  long long* out; double* din; CK(cudaMalloc(&out, 16 * 8)); CK(cudaMalloc(&din, 64)); CK(cudaMemset(din, 0, 64));
  const char* names[] = {"dfma_dep", "ddiv_dep", "warp_argmax_u64", "syncthreads", "smem_xchg_2sync", "global_pkt_roundtrip", "dsqrt_dep", "lds_dfma_sts"};
  for (int threads : {32, 512}) {
    lat<<<1, threads>>>(out, din, 200); CK(cudaDeviceSynchronize());
    lat<<<1, threads>>>(out, din, 200); CK(cudaDeviceSynchronize());
    long long h[16]; CK(cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost));
    for (int i = 0; i < 8; i++) printf("threads=%d %s: %lld cycles\n", threads, names[i], h[i]);
  }
  return 0;
}
