// la_b200 — common device/host helpers shared by every kernel file.
//
// Scalar types on the hot path follow the reference's three element types
// (src/linalg/lapack_helper.cr:187-193 picks s/d/z from Float32/Float64/Complex):
//   float, double, zcomplex (two doubles, same layout as LibCBLAS::ComplexDouble,
//   src/linalg/libCBLAS.cr:9-12).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <atomic>

namespace la {

struct __align__(16) zcomplex {
  double x, y;
};

#define LA_HD __host__ __device__ __forceinline__

// ---- scalar traits -------------------------------------------------------
template <typename T> struct Traits;
template <> struct Traits<float> {
  using real = float;
  static constexpr bool is_complex = false;
  static constexpr char letter = 's';
};
template <> struct Traits<double> {
  using real = double;
  static constexpr bool is_complex = false;
  static constexpr char letter = 'd';
};
template <> struct Traits<zcomplex> {
  using real = double;
  static constexpr bool is_complex = true;
  static constexpr char letter = 'z';
};

LA_HD float  make_zero(float*)  { return 0.f; }
LA_HD double make_zero(double*) { return 0.0; }
LA_HD zcomplex make_zero(zcomplex*) { return zcomplex{0.0, 0.0}; }
template <typename T> LA_HD T zero() { return make_zero((T*)nullptr); }
LA_HD float  make_real(float*, double r)  { return (float)r; }
LA_HD double make_real(double*, double r) { return r; }
LA_HD zcomplex make_real(zcomplex*, double r) { return zcomplex{r, 0.0}; }
template <typename T> LA_HD T from_real(double r) { return make_real((T*)nullptr, r); }
template <typename T> LA_HD T one() { return from_real<T>(1.0); }

LA_HD float  add(float a, float b) { return a + b; }
LA_HD double add(double a, double b) { return a + b; }
LA_HD zcomplex add(zcomplex a, zcomplex b) { return zcomplex{a.x + b.x, a.y + b.y}; }
LA_HD float  sub(float a, float b) { return a - b; }
LA_HD double sub(double a, double b) { return a - b; }
LA_HD zcomplex sub(zcomplex a, zcomplex b) { return zcomplex{a.x - b.x, a.y - b.y}; }
LA_HD float  mul(float a, float b) { return a * b; }
LA_HD double mul(double a, double b) { return a * b; }
LA_HD zcomplex mul(zcomplex a, zcomplex b) {
  return zcomplex{a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x};
}
LA_HD float  neg(float a) { return -a; }
LA_HD double neg(double a) { return -a; }
LA_HD zcomplex neg(zcomplex a) { return zcomplex{-a.x, -a.y}; }
LA_HD float  conj_(float a) { return a; }
LA_HD double conj_(double a) { return a; }
LA_HD zcomplex conj_(zcomplex a) { return zcomplex{a.x, -a.y}; }
// c + a*b
LA_HD float  fma_(float a, float b, float c) { return fmaf(a, b, c); }
LA_HD double fma_(double a, double b, double c) { return fma(a, b, c); }
LA_HD zcomplex fma_(zcomplex a, zcomplex b, zcomplex c) {
  zcomplex r;
  r.x = fma(a.x, b.x, c.x); r.x = fma(-a.y, b.y, r.x);
  r.y = fma(a.x, b.y, c.y); r.y = fma(a.y, b.x, r.y);
  return r;
}
// c - a*b
template <typename T> LA_HD T fnma_(T a, T b, T c) { return fma_(neg(a), b, c); }

LA_HD float  real_(float a) { return a; }
LA_HD double real_(double a) { return a; }
LA_HD double real_(zcomplex a) { return a.x; }
LA_HD float  imag_(float) { return 0.f; }
LA_HD double imag_(double) { return 0.0; }
LA_HD double imag_(zcomplex a) { return a.y; }

LA_HD bool is_zero(float a) { return a == 0.f; }
LA_HD bool is_zero(double a) { return a == 0.0; }
LA_HD bool is_zero(zcomplex a) { return a.x == 0.0 && a.y == 0.0; }
LA_HD bool is_one(float a) { return a == 1.f; }
LA_HD bool is_one(double a) { return a == 1.0; }
LA_HD bool is_one(zcomplex a) { return a.x == 1.0 && a.y == 0.0; }

// |re|+|im| — the "cabs1" magnitude i?amax uses for pivot search (LAPACK izamax).
LA_HD float  abs1(float a) { return fabsf(a); }
LA_HD double abs1(double a) { return fabs(a); }
LA_HD double abs1(zcomplex a) { return fabs(a.x) + fabs(a.y); }

LA_HD float  scale_real(float a, float s) { return a * s; }
LA_HD double scale_real(double a, double s) { return a * s; }
LA_HD zcomplex scale_real(zcomplex a, double s) { return zcomplex{a.x * s, a.y * s}; }

// reciprocal / division
LA_HD float  recip(float a) { return 1.f / a; }
LA_HD double recip(double a) { return 1.0 / a; }
// complex reciprocal following LAPACK's zladiv-style scaled (Smith) algorithm
LA_HD zcomplex cdiv(zcomplex a, zcomplex b) {
  zcomplex r;
  if (fabs(b.y) < fabs(b.x)) {
    double e = b.y / b.x, f = b.x + b.y * e;
    r.x = (a.x + a.y * e) / f;
    r.y = (a.y - a.x * e) / f;
  } else {
    double e = b.x / b.y, f = b.y + b.x * e;
    r.x = (a.y + a.x * e) / f;
    r.y = (-a.x + a.y * e) / f;
  }
  return r;
}
LA_HD zcomplex recip(zcomplex a) { return cdiv(zcomplex{1.0, 0.0}, a); }
LA_HD float  div_(float a, float b) { return a / b; }
LA_HD double div_(double a, double b) { return a / b; }
LA_HD zcomplex div_(zcomplex a, zcomplex b) { return cdiv(a, b); }

// safe minimum (dlamch('S')) used by getf2's "scale by reciprocal" rule
template <typename R> LA_HD R sfmin();
template <> LA_HD float  sfmin<float>()  { return 1.17549435e-38f; }
template <> LA_HD double sfmin<double>() { return 2.2250738585072014e-308; }

// ---- BLAS/LAPACK option codes (values of the CBLAS enums, libCBLAS.cr:19-48)
enum Op : int { OpN = 111, OpT = 112, OpC = 113 };
enum Uplo : int { Upper = 121, Lower = 122 };
enum Diag : int { NonUnit = 131, Unit = 132 };
enum Side : int { Left = 141, Right = 142 };

template <typename T> LA_HD T apply_conj(T a, bool c) { return c ? conj_(a) : a; }

// ---- error handling --------------------------------------------------------
#define LA_CUDA_CHECK(expr)                                                          \
  do {                                                                               \
    cudaError_t err__ = (expr);                                                      \
    if (err__ != cudaSuccess) {                                                      \
      fprintf(stderr, "la_b200: CUDA error '%s' at %s:%d (%s)\n",                    \
              cudaGetErrorString(err__), __FILE__, __LINE__, #expr);                 \
      abort();                                                                       \
    }                                                                                \
  } while (0)

// One-time set-up that CUDA keeps PER DEVICE (cudaFuncSetAttribute and the like): a flag per
// device ordinal, safe to test from any number of host threads.  Two threads that race on the
// first use both run the (idempotent) set-up; devices beyond the table repeat it every call.
struct DeviceOnce {
  static constexpr int kMax = 64;
  std::atomic<bool> done[kMax] = {};
  bool pending(int dev) const {
    return dev < 0 || dev >= kMax || !done[dev].load(std::memory_order_acquire);
  }
  void mark(int dev) {
    if (dev >= 0 && dev < kMax) done[dev].store(true, std::memory_order_release);
  }
};

LA_HD int64_t round_up(int64_t a, int64_t b) { return (a + b - 1) / b * b; }
LA_HD int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }

}  // namespace la
