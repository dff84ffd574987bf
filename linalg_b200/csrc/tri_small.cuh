// la_b200 — small triangular solves with the right-hand side held in registers.
// Shared by the TRSM base case (level3.cu) and the Cholesky leaf kernel (cholesky.cu).
#pragma once
#include "common.cuh"

namespace la {

// Solve M x = alpha*b for ONE vector owned by the calling thread.  M is the effective
// (already op()-ed) NA x NA triangular matrix, dense in shared memory with row pitch LDM (other
// triangle ignored); rdiag[i] = 1 / M(i,i) (ignored when `unit`).  x[] holds b on entry and
// the solution on exit.  The loops are fully unrolled so that x[] stays in registers, and the
// elimination is column oriented (x_j is final, then every later entry is updated
// independently), which keeps the dependent chain at NA steps instead of NA^2/2.
// `na` (<= NA) is the run-time order.
template <typename T, int NA, int LDM>
__device__ __forceinline__ void solve_vec_regs(const T* __restrict__ Ms, const T* __restrict__ rdiag,
                                               int na, bool lower, bool unit, T alpha, T (&x)[NA]) {
  if (!is_one(alpha)) {
#pragma unroll
    for (int i = 0; i < NA; i++) x[i] = mul(alpha, x[i]);
  }
  if (lower) {
#pragma unroll
    for (int j = 0; j < NA; j++) {
      if (j < na) {
        if (!unit) x[j] = mul(x[j], rdiag[j]);
        const T xj = x[j];
#pragma unroll
        for (int i = j + 1; i < NA; i++)
          if (i < na) x[i] = fnma_(Ms[i * LDM + j], xj, x[i]);
      }
    }
  } else {
#pragma unroll
    for (int jj = 0; jj < NA; jj++) {
      const int j = NA - 1 - jj;
      if (j < na) {
        if (!unit) x[j] = mul(x[j], rdiag[j]);
        const T xj = x[j];
#pragma unroll
        for (int i = 0; i < j; i++) x[i] = fnma_(Ms[i * LDM + j], xj, x[i]);
      }
    }
  }
}

}  // namespace la
