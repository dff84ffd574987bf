// la_b200 — small triangular solves with the right-hand side held in registers.
// Shared by the TRSM base case (level3.cu) and the Cholesky leaf kernel (factor.cu).
#pragma once
#include "common.cuh"

namespace la {

// Solve M x = alpha*b for ONE vector owned by the calling thread.  M is the effective
// (already op()-ed) NA x NA triangular matrix, dense in shared memory with row pitch LDM,
// other triangle ignored.  x[] holds b on entry and the solution on exit; the loops are fully
// unrolled so x[] stays in registers; `na` (<= NA) is the run-time order.
template <typename T, int NA, int LDM>
__device__ __forceinline__ void solve_vec_regs(const T* __restrict__ Ms, int na, bool lower,
                                               bool unit, T alpha, T (&x)[NA]) {
  const bool scale = !is_one(alpha);
  if (lower) {
#pragma unroll
    for (int i = 0; i < NA; i++) {
      if (i < na) {
        T s = scale ? mul(alpha, x[i]) : x[i];
#pragma unroll
        for (int j = 0; j < i; j++) s = fnma_(Ms[i * LDM + j], x[j], s);
        x[i] = unit ? s : div_(s, Ms[i * LDM + i]);
      }
    }
  } else {
#pragma unroll
    for (int ii = 0; ii < NA; ii++) {
      const int i = NA - 1 - ii;
      if (i < na) {
        T s = scale ? mul(alpha, x[i]) : x[i];
#pragma unroll
        for (int j = i + 1; j < NA; j++)
          if (j < na) s = fnma_(Ms[i * LDM + j], x[j], s);
        x[i] = unit ? s : div_(s, Ms[i * LDM + i]);
      }
    }
  }
}

}  // namespace la
