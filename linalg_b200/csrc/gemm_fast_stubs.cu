// la_b200 — placeholders for tensor-core GEMM paths that are not built yet.  Each returns
// false so that the dispatcher in level3.cu uses the generic SIMT kernel.
#include "la_internal.h"

namespace la {

#ifndef LA_HAVE_DMMA_Z
bool gemm_dmma_z(Context&, Op, Op, int64_t, int64_t, int64_t, zcomplex, const zcomplex*, int64_t,
                 const zcomplex*, int64_t, zcomplex, zcomplex*, int64_t, int, bool) {
  return false;
}
#endif

#ifndef LA_HAVE_TF32X3
bool gemm_tf32x3(Context&, Op, Op, int64_t, int64_t, int64_t, float, const float*, int64_t,
                 const float*, int64_t, float, float*, int64_t, int) {
  return false;
}
#endif

#ifndef LA_HAVE_DMMA_F64
bool gemm_dmma_f64(Context&, Op, Op, int64_t, int64_t, int64_t, double, const double*, int64_t,
                   const double*, int64_t, double, double*, int64_t, int) {
  return false;
}
#endif

}  // namespace la
