"""oracle/ — TEST INFRASTRUCTURE, NOT PRODUCT CODE.

CPU oracles for the hot path.  Only tests/, __graft_entry__.smoke() and bench.py's CPU legs
may import this package; nothing under linalg_b200/ does.

  oracle.c_ref        plain-C restatement of the reference BLAS/LAPACK 3.12.0 algorithms
                      (la_ref.c / la_ref_impl.h), loaded through ctypes.
  oracle.openblas_ref the library the reference actually links at run time (OpenBLAS +
                      LAPACK 3.12.0; here the copy SciPy ships, symbols prefixed `scipy_`),
                      bound with the same `BlasLapackABI` the product uses, so the host-side
                      mirror (linalg_b200.la) can run unchanged against it.
  oracle.checks       residual / pivot comparison helpers with the tolerances of
                      BASELINE.json's north_star.

Pinning status: PINNED.  tests/test_oracle.py checks both oracles against the reference's
own golden vectors (spec/linalg_spec.cr, spec/matrix_spec.cr: K1-K17 of SURVEY.md §4) and
against each other on seeded random inputs.
"""
