// Micro-benchmark: register-resident FP64 peak on B200 (sm_100a).
// Measures DMMA (mma.sync f64) for the four PTX shapes and plain DFMA so that
// roofline fractions for the f64 GEMM/trailing-update kernels have a measured
// denominator (MEASURED_PEAKS.json carries no FP64 figure).
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { \
  fprintf(stderr, "CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1);} } while (0)

template <int SHAPE> struct Dmma;

template <> struct Dmma<0> {  // m8n8k4
  static constexpr int FMAS = 8 * 8 * 4;
  static constexpr int NA = 1, NB = 1, NC = 2;
  __device__ static void run(double (&c)[2], const double (&a)[1], const double (&b)[1]) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c[0]), "+d"(c[1]) : "d"(a[0]), "d"(b[0]));
  }
};
template <> struct Dmma<1> {  // m16n8k4
  static constexpr int FMAS = 16 * 8 * 4;
  static constexpr int NA = 2, NB = 1, NC = 4;
  __device__ static void run(double (&c)[4], const double (&a)[2], const double (&b)[1]) {
    asm volatile("mma.sync.aligned.m16n8k4.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
                 : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3]) : "d"(a[0]), "d"(a[1]), "d"(b[0]));
  }
};
template <> struct Dmma<2> {  // m16n8k8
  static constexpr int FMAS = 16 * 8 * 8;
  static constexpr int NA = 4, NB = 2, NC = 4;
  __device__ static void run(double (&c)[4], const double (&a)[4], const double (&b)[2]) {
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3])
                 : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(b[0]), "d"(b[1]));
  }
};
template <> struct Dmma<3> {  // m16n8k16
  static constexpr int FMAS = 16 * 8 * 16;
  static constexpr int NA = 8, NB = 4, NC = 4;
  __device__ static void run(double (&c)[4], const double (&a)[8], const double (&b)[4]) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};"
                 : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3])
                 : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]), "d"(a[6]), "d"(a[7]),
                   "d"(b[0]), "d"(b[1]), "d"(b[2]), "d"(b[3]));
  }
};

template <int SHAPE, int NACC>
__global__ void __launch_bounds__(1024) dmma_kernel(double* out, int iters, double seed) {
  using M = Dmma<SHAPE>;
  double a[M::NA], b[M::NB], c[NACC][M::NC];
  for (int i = 0; i < M::NA; i++) a[i] = seed + threadIdx.x * 1e-9 + i;
  for (int i = 0; i < M::NB; i++) b[i] = seed * 0.5 + i;
  for (int j = 0; j < NACC; j++) for (int i = 0; i < M::NC; i++) c[j][i] = 0.0;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int j = 0; j < NACC; j++) M::run(c[j], a, b);
  }
  double s = 0;
  for (int j = 0; j < NACC; j++) for (int i = 0; i < M::NC; i++) s += c[j][i];
  if (s == 123.456) out[0] = s;
}

template <int NACC>
__global__ void __launch_bounds__(1024) dfma_kernel(double* out, int iters, double seed) {
  double c[NACC];
  double a = seed + threadIdx.x * 1e-9, b = seed * 0.5;
  for (int j = 0; j < NACC; j++) c[j] = j;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int j = 0; j < NACC; j++) c[j] = fma(a, b, c[j]);
  }
  double s = 0;
  for (int j = 0; j < NACC; j++) s += c[j];
  if (s == 123.456) out[0] = s;
}

template <typename F>
static double time_ms(F launch, int reps) {
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  launch(); CK(cudaDeviceSynchronize());
  double best = 1e30;
  for (int r = 0; r < reps; r++) {
    CK(cudaEventRecord(e0)); launch(); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); if (ms < best) best = ms;
  }
  return best;
}

template <int SHAPE, int NACC>
static void bench_dmma(const char* name, int sms, int threads, int blocks_per_sm, double* out) {
  int iters = 20000 / NACC * 8 / (Dmma<SHAPE>::FMAS / 256);
  if (iters < 64) iters = 64;
  int grid = sms * blocks_per_sm;
  double ms = time_ms([&] { dmma_kernel<SHAPE, NACC><<<grid, threads>>>(out, iters, 1.0); }, 5);
  CK(cudaGetLastError());
  double flops = 2.0 * Dmma<SHAPE>::FMAS * NACC * (double)iters * (threads / 32) * grid;
  printf("{\"kind\":\"dmma\",\"shape\":\"%s\",\"nacc\":%d,\"threads\":%d,\"blocks_per_sm\":%d,\"ms\":%.4f,\"tflops\":%.3f}\n",
         name, NACC, threads, blocks_per_sm, ms, flops / ms * 1e-9);
}

int main() {
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
  int sms = p.multiProcessorCount;
  printf("{\"device\":\"%s\",\"sms\":%d,\"clock_khz\":%d}\n", p.name, sms, p.clockRate);
  double* out; CK(cudaMalloc(&out, 8));
  for (int threads : {128, 256, 512}) {
    bench_dmma<0, 8>("m8n8k4", sms, threads, 1, out);
    bench_dmma<1, 8>("m16n8k4", sms, threads, 1, out);
    bench_dmma<2, 8>("m16n8k8", sms, threads, 1, out);
    bench_dmma<3, 8>("m16n8k16", sms, threads, 1, out);
  }
  bench_dmma<3, 1>("m16n8k16", sms, 128, 1, out);
  bench_dmma<3, 2>("m16n8k16", sms, 128, 1, out);
  bench_dmma<3, 4>("m16n8k16", sms, 128, 1, out);
  bench_dmma<3, 16>("m16n8k16", sms, 128, 1, out);
  bench_dmma<3, 16>("m16n8k16", sms, 256, 1, out);
  bench_dmma<0, 16>("m8n8k4", sms, 256, 2, out);
  for (int threads : {128, 256, 512, 1024}) {
    int iters = 40000;
    double ms = time_ms([&] { dfma_kernel<16><<<sms, threads>>>(out, iters, 1.0); }, 5);
    double flops = 2.0 * 16 * (double)iters * threads * sms;
    printf("{\"kind\":\"dfma\",\"threads\":%d,\"ms\":%.4f,\"tflops\":%.3f}\n", threads, ms, flops / ms * 1e-9);
  }
  // sustained: ~3 s of m16n8k16 back to back
  {
    int iters = 4000;
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    dmma_kernel<3, 8><<<sms, 256>>>(out, iters, 1.0); CK(cudaDeviceSynchronize());
    int launches = 0; float ms = 0;
    CK(cudaEventRecord(e0));
    while (true) {
      for (int i = 0; i < 20; i++) dmma_kernel<3, 8><<<sms, 256>>>(out, iters, 1.0);
      launches += 20;
      CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); CK(cudaEventElapsedTime(&ms, e0, e1));
      if (ms > 3000) break;
    }
    double flops = 2.0 * Dmma<3>::FMAS * 8 * (double)iters * 8 * sms * launches;
    printf("{\"kind\":\"dmma_sustained\",\"shape\":\"m16n8k16\",\"seconds\":%.2f,\"tflops\":%.3f}\n", ms * 1e-3, flops / ms * 1e-9);
  }
  return 0;
}
