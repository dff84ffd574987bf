/* oracle/la_ref.c — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C CPU oracle for the hot path: instantiates la_ref_impl.h for float (ref_s*),
 * double (ref_d*) and double complex (ref_z*).  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline leg may load the resulting oracle/_build/libla_ref.so; the
 * product (linalg_b200/) never does.  See la_ref_impl.h for what is restated and from where.
 *
 * Pinning: tests/test_oracle.py checks this file against the reference's own golden vectors
 *  and against OpenBLAS 0.3.31.dev + LAPACK 3.12.0 (the library the
 * reference links, shipped here inside SciPy).
 */
#include <complex.h>
#include <math.h>
#include <stdlib.h>
#include <stddef.h>

#define S float
#define R float
#define FN(name) ref_s##name
#define IS_COMPLEX 0
#define CONJ(x) (x)
#define CREAL(x) (x)
#define CIMAG(x) (0.0f)
#define ABS1(x) fabsf(x)
#define SFMIN 1.17549435e-38f
#include "la_ref_impl.h"
#undef S
#undef R
#undef FN
#undef IS_COMPLEX
#undef CONJ
#undef CREAL
#undef CIMAG
#undef ABS1
#undef SFMIN

#define S double
#define R double
#define FN(name) ref_d##name
#define IS_COMPLEX 0
#define CONJ(x) (x)
#define CREAL(x) (x)
#define CIMAG(x) (0.0)
#define ABS1(x) fabs(x)
#define SFMIN 2.2250738585072014e-308
#include "la_ref_impl.h"
#undef S
#undef R
#undef FN
#undef IS_COMPLEX
#undef CONJ
#undef CREAL
#undef CIMAG
#undef ABS1
#undef SFMIN

#define S double complex
#define R double
#define FN(name) ref_z##name
#define IS_COMPLEX 1
#define CONJ(x) conj(x)
#define CREAL(x) creal(x)
#define CIMAG(x) cimag(x)
#define ABS1(x) (fabs(creal(x)) + fabs(cimag(x)))
#define SFMIN 2.2250738585072014e-308
#include "la_ref_impl.h"
