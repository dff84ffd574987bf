// Micro-benchmark: tcgen05.mma.kind::tf32 peak on B200 (sm_100a).
//
// One CTA per SM issues back-to-back M=128 x N=256 x K=8 TF32 MMAs (cta_group::1) from one
// thread, operands resident in shared memory (never reloaded), accumulators in tensor memory
// (two 256-column buffers used alternately).  This is the denominator of the 3xTF32 SGEMM's
// roofline fraction (MEASURED_PEAKS.json carries bf16 only).  Also run with N=128/64 and with
// cta_group::1 M=64 to show the shape dependence.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o peak_tf32 peak_tf32.cu && ./peak_tf32
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { \
  fprintf(stderr, "CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1);} } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ uint64_t kmajor_sw128_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3fff);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
__host__ __device__ constexpr uint32_t idesc_tf32(int M, int N) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void umma_tf32(uint32_t tmem_c, uint64_t a, uint64_t b, uint32_t idesc,
                                          uint32_t acc) {
  asm volatile(
      "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n}\n" ::"r"(tmem_c),
      "l"(a), "l"(b), "r"(idesc), "r"(acc)
      : "memory");
}

template <int M, int N>
__global__ void __launch_bounds__(128, 1) tf32_peak_kernel(int iters, float* out) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = (unsigned char*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_ptr;
  // operands: A tile 128 x 32 floats (16 KB), B tile 256 x 32 floats (32 KB), small finite values
  float* f = (float*)smem;
  for (int i = threadIdx.x; i < (16384 + 32768) / 4; i += blockDim.x) f[i] = 1.0f / (float)(1 + (i & 15));
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(
                     smem_u32(&tmem_ptr))
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = tmem_ptr;
  if (threadIdx.x == 0) {
    const uint32_t s = smem_u32(smem);
    constexpr uint32_t idesc = idesc_tf32(M, N);
    for (int it = 0; it < iters; it++) {
      const uint32_t tacc = tmem_base + (uint32_t)((it & 1) * 256);
#pragma unroll
      for (int ks = 0; ks < 4; ks++) {
        const uint64_t a = kmajor_sw128_desc(s + ks * 32);
        const uint64_t b = kmajor_sw128_desc(s + 16384 + ks * 32);
        umma_tf32(tacc, a, b, idesc, (it > 1 || ks) ? 1u : 0u);
      }
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                     smem_u32(&bar))
                 : "memory");
    uint32_t ok = 0;
    while (!ok) {
      asm volatile(
          "{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n"
          "selp.u32 %0, 1, 0, p;\n}\n"
          : "=r"(ok)
          : "r"(smem_u32(&bar))
          : "memory");
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  if (threadIdx.x < 32) {
    uint32_t v;
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(v) : "r"(tmem_base) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    if (__uint_as_float(v) == 123.456f) out[0] = __uint_as_float(v);
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base) : "memory");
  }
}

template <int M, int N>
static void run(int sms, float* out, double seconds) {
  const int smem = 16384 + 32768 + 1024;
  CK(cudaFuncSetAttribute(tf32_peak_kernel<M, N>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  const int iters = 40000;
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  tf32_peak_kernel<M, N><<<sms, 128, smem>>>(iters, out); CK(cudaDeviceSynchronize());
  double best = 1e30;
  for (int r = 0; r < 5; r++) {
    CK(cudaEventRecord(e0));
    tf32_peak_kernel<M, N><<<sms, 128, smem>>>(iters, out);
    CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); if (ms < best) best = ms;
  }
  const double flop = 2.0 * M * N * 32.0 * iters * sms;
  printf("{\"kind\":\"tf32\",\"M\":%d,\"N\":%d,\"ms\":%.4f,\"tflops_burst\":%.1f", M, N, best, flop / best * 1e-9);
  if (seconds > 0) {   // sustained: back to back for `seconds`
    int launches = 0; float ms = 0;
    CK(cudaEventRecord(e0));
    while (true) {
      for (int i = 0; i < 10; i++) tf32_peak_kernel<M, N><<<sms, 128, smem>>>(iters, out);
      launches += 10;
      CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); CK(cudaEventElapsedTime(&ms, e0, e1));
      if (ms > seconds * 1e3) break;
    }
    printf(",\"tflops_sustained\":%.1f,\"sustained_s\":%.1f", flop * launches / ms * 1e-9, ms * 1e-3);
  }
  printf("}\n");
  CK(cudaGetLastError());
}

int main() {
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
  const int sms = p.multiProcessorCount;
  printf("{\"device\":\"%s\",\"sms\":%d}\n", p.name, sms);
  float* out; CK(cudaMalloc(&out, 4));
  run<128, 256>(sms, out, 3.0);
  run<128, 128>(sms, out, 0);
  run<128, 64>(sms, out, 0);
  run<64, 256>(sms, out, 0);
  return 0;
}
