/* oracle/la_ref_impl.h — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Type-generic body of the CPU oracle; la_ref.c includes it three times (float, double,
 * double complex).  It restates, in plain unblocked C99, the published algorithms of the
 * library the reference delegates its arithmetic to (reference BLAS + LAPACK 3.12.0, the
 * version the reference's CI installs, /root/reference/.github/workflows/linux.yml:24-29;
 * that library is NOT under /root/reference and is not version-pinned by shard.yml):
 *   xGEMM, xTRSM (reference BLAS), xGETF2/xGETRS, xPOTF2/xPOTRS, xGEQR2 + xLARFG + xLARF,
 *   xORM2R/xUNM2R, xGELQ2 + xORML2/xUNML2, xTRTRS, xGELS, xGETRI/xPOTRI/xTRTRI (by solves).
 * Call sites in the reference that reach each routine are cited in include/la_b200.h.
 * Conventions kept: column-major, 1-based ipiv, i?amax picks the FIRST maximal |re|+|im|,
 * xGETF2 scales by the reciprocal pivot when |pivot| >= sfmin, xLARFG's
 * beta = -sign(Re alpha)*norm, info semantics.
 *
 * Required macros: S (scalar type), R (real type), FN(name), IS_COMPLEX, CONJ(x), CREAL(x),
 * CIMAG(x), ABS1(x), SFMIN.
 */

#define IDX(A, ld, i, j) (A)[(size_t)(i) + (size_t)(j) * (size_t)(ld)]

static S FN(opget)(const S* A, int lda, int trans, int i, int j) {
  /* op(A)(i,j) */
  if (trans == 'N') return IDX(A, lda, i, j);
  if (trans == 'T') return IDX(A, lda, j, i);
  return CONJ(IDX(A, lda, j, i));
}

void FN(gemm)(int ta, int tb, int m, int n, int k, const S* alpha, const S* A, int lda,
              const S* B, int ldb, const S* beta, S* C, int ldc) {
  for (int j = 0; j < n; j++)
    for (int i = 0; i < m; i++) {
      S acc = 0;
      for (int l = 0; l < k; l++) acc += FN(opget)(A, lda, ta, i, l) * FN(opget)(B, ldb, tb, l, j);
      S c = (*beta == 0) ? 0 : (*beta) * IDX(C, ldc, i, j);
      IDX(C, ldc, i, j) = (*alpha) * acc + c;
    }
}

/* B <- alpha * op(A)^-1 B (side 'L') or alpha * B op(A)^-1 (side 'R') */
void FN(trsm)(int side, int uplo, int trans, int diag, int m, int n, const S* alpha, const S* A,
              int lda, S* B, int ldb) {
  int na = side == 'L' ? m : n;
  int eff_lower = (uplo == 'L') != (trans != 'N');
  for (int j = 0; j < n; j++)
    for (int i = 0; i < m; i++) IDX(B, ldb, i, j) *= *alpha;
  if (side == 'L') {
    for (int j = 0; j < n; j++) {
      if (eff_lower) {
        for (int i = 0; i < na; i++) {
          S s = IDX(B, ldb, i, j);
          for (int l = 0; l < i; l++) s -= FN(opget)(A, lda, trans, i, l) * IDX(B, ldb, l, j);
          IDX(B, ldb, i, j) = diag == 'U' ? s : s / FN(opget)(A, lda, trans, i, i);
        }
      } else {
        for (int i = na - 1; i >= 0; i--) {
          S s = IDX(B, ldb, i, j);
          for (int l = i + 1; l < na; l++) s -= FN(opget)(A, lda, trans, i, l) * IDX(B, ldb, l, j);
          IDX(B, ldb, i, j) = diag == 'U' ? s : s / FN(opget)(A, lda, trans, i, i);
        }
      }
    }
  } else {
    /* X M = B, M = op(A): column j of X needs columns l with M(l,j) != 0 */
    for (int i = 0; i < m; i++) {
      if (!eff_lower) { /* M upper: forward over columns */
        for (int j = 0; j < na; j++) {
          S s = IDX(B, ldb, i, j);
          for (int l = 0; l < j; l++) s -= IDX(B, ldb, i, l) * FN(opget)(A, lda, trans, l, j);
          IDX(B, ldb, i, j) = diag == 'U' ? s : s / FN(opget)(A, lda, trans, j, j);
        }
      } else {
        for (int j = na - 1; j >= 0; j--) {
          S s = IDX(B, ldb, i, j);
          for (int l = j + 1; l < na; l++) s -= IDX(B, ldb, i, l) * FN(opget)(A, lda, trans, l, j);
          IDX(B, ldb, i, j) = diag == 'U' ? s : s / FN(opget)(A, lda, trans, j, j);
        }
      }
    }
  }
}

/* xGETF2: right-looking unblocked LU with partial pivoting */
void FN(getrf)(int m, int n, S* A, int lda, int* ipiv, int* info) {
  *info = 0;
  int kmax = m < n ? m : n;
  for (int j = 0; j < kmax; j++) {
    int p = j;
    R best = ABS1(IDX(A, lda, j, j));
    for (int i = j + 1; i < m; i++) {
      R v = ABS1(IDX(A, lda, i, j));
      if (v > best) { best = v; p = i; }
    }
    ipiv[j] = p + 1;
    if (IDX(A, lda, p, j) != 0) {
      if (p != j)
        for (int c = 0; c < n; c++) {
          S t = IDX(A, lda, j, c); IDX(A, lda, j, c) = IDX(A, lda, p, c); IDX(A, lda, p, c) = t;
        }
      S piv = IDX(A, lda, j, j);
      if (ABS1(piv) >= SFMIN) {
        S r = (S)1 / piv;
        for (int i = j + 1; i < m; i++) IDX(A, lda, i, j) *= r;
      } else {
        for (int i = j + 1; i < m; i++) IDX(A, lda, i, j) /= piv;
      }
    } else if (*info == 0) {
      *info = j + 1;
    }
    for (int c = j + 1; c < n; c++) {
      S u = IDX(A, lda, j, c);
      for (int i = j + 1; i < m; i++) IDX(A, lda, i, c) -= IDX(A, lda, i, j) * u;
    }
  }
}

static void FN(laswp)(int n, S* B, int ldb, int k1, int k2, const int* ipiv, int reverse) {
  if (!reverse) {
    for (int k = k1; k < k2; k++) {
      int p = ipiv[k] - 1;
      if (p != k) for (int c = 0; c < n; c++) {
        S t = IDX(B, ldb, k, c); IDX(B, ldb, k, c) = IDX(B, ldb, p, c); IDX(B, ldb, p, c) = t;
      }
    }
  } else {
    for (int k = k2 - 1; k >= k1; k--) {
      int p = ipiv[k] - 1;
      if (p != k) for (int c = 0; c < n; c++) {
        S t = IDX(B, ldb, k, c); IDX(B, ldb, k, c) = IDX(B, ldb, p, c); IDX(B, ldb, p, c) = t;
      }
    }
  }
}

void FN(getrs)(int trans, int n, int nrhs, const S* A, int lda, const int* ipiv, S* B, int ldb) {
  S one = 1;
  if (trans == 'N') {
    FN(laswp)(nrhs, B, ldb, 0, n, ipiv, 0);
    FN(trsm)('L', 'L', 'N', 'U', n, nrhs, &one, A, lda, B, ldb);
    FN(trsm)('L', 'U', 'N', 'N', n, nrhs, &one, A, lda, B, ldb);
  } else {
    FN(trsm)('L', 'U', trans, 'N', n, nrhs, &one, A, lda, B, ldb);
    FN(trsm)('L', 'L', trans, 'U', n, nrhs, &one, A, lda, B, ldb);
    FN(laswp)(nrhs, B, ldb, 0, n, ipiv, 1);
  }
}

/* xPOTF2 (unblocked Cholesky), only the `uplo` triangle is referenced */
void FN(potrf)(int uplo, int n, S* A, int lda, int* info) {
  *info = 0;
  for (int j = 0; j < n; j++) {
    R d = CREAL(IDX(A, lda, j, j));
    for (int l = 0; l < j; l++) {
      S v = uplo == 'L' ? IDX(A, lda, j, l) : IDX(A, lda, l, j);
      d -= CREAL(v * CONJ(v));
    }
    if (!(d > 0)) { IDX(A, lda, j, j) = d; *info = j + 1; return; }
    d = (R)sqrt((double)d);
    IDX(A, lda, j, j) = d;
    for (int i = j + 1; i < n; i++) {
      if (uplo == 'L') {
        S s = IDX(A, lda, i, j);
        for (int l = 0; l < j; l++) s -= IDX(A, lda, i, l) * CONJ(IDX(A, lda, j, l));
        IDX(A, lda, i, j) = s / d;
      } else {
        S s = IDX(A, lda, j, i);
        for (int l = 0; l < j; l++) s -= CONJ(IDX(A, lda, l, j)) * IDX(A, lda, l, i);
        IDX(A, lda, j, i) = s / d;
      }
    }
  }
}

void FN(potrs)(int uplo, int n, int nrhs, const S* A, int lda, S* B, int ldb) {
  S one = 1;
  if (uplo == 'L') {
    FN(trsm)('L', 'L', 'N', 'N', n, nrhs, &one, A, lda, B, ldb);
    FN(trsm)('L', 'L', 'C', 'N', n, nrhs, &one, A, lda, B, ldb);
  } else {
    FN(trsm)('L', 'U', 'C', 'N', n, nrhs, &one, A, lda, B, ldb);
    FN(trsm)('L', 'U', 'N', 'N', n, nrhs, &one, A, lda, B, ldb);
  }
}

void FN(trtrs)(int uplo, int trans, int diag, int n, int nrhs, const S* A, int lda, S* B,
               int ldb, int* info) {
  *info = 0;
  if (diag == 'N')
    for (int i = 0; i < n; i++)
      if (IDX(A, lda, i, i) == 0) { *info = i + 1; return; }
  S one = 1;
  FN(trsm)('L', uplo, trans, diag, n, nrhs, &one, A, lda, B, ldb);
}

/* xLARFG on (alpha, x[0..n-1) with stride incx): returns tau, overwrites alpha with beta and
 * x with v(2:n).  The safmin rescaling loop of LAPACK is omitted. */
static S FN(larfg)(int n, S* alpha, S* x, int incx) {
  double xn2 = 0;
  for (int i = 0; i < n - 1; i++) {
    S v = x[(size_t)i * incx];
    xn2 += (double)CREAL(v) * (double)CREAL(v) + (double)CIMAG(v) * (double)CIMAG(v);
  }
  double ar = (double)CREAL(*alpha), ai = (double)CIMAG(*alpha);
  if (xn2 == 0 && ai == 0) return 0;
  double nrm = sqrt(ar * ar + ai * ai + xn2);
  double beta = ar >= 0 ? -nrm : nrm;
  if (ar == 0 && signbit(ar)) beta = nrm;
  S tau;
#if IS_COMPLEX
  tau = (beta - ar) / beta + I * (-ai / beta);
#else
  tau = (S)((beta - ar) / beta);
#endif
  S scale = (S)1 / (*alpha - (S)beta);
  for (int i = 0; i < n - 1; i++) x[(size_t)i * incx] *= scale;
  *alpha = (S)beta;
  return tau;
}

/* xGEQR2 */
void FN(geqrf)(int m, int n, S* A, int lda, S* tau) {
  int k = m < n ? m : n;
  for (int j = 0; j < k; j++) {
    tau[j] = FN(larfg)(m - j, &IDX(A, lda, j, j), &IDX(A, lda, (j + 1 < m ? j + 1 : j), j), 1);
    /* apply H(j)^H to A(j:m, j+1:n) from the left */
    S ct = CONJ(tau[j]);
    for (int c = j + 1; c < n; c++) {
      S w = IDX(A, lda, j, c);
      for (int i = j + 1; i < m; i++) w += CONJ(IDX(A, lda, i, j)) * IDX(A, lda, i, c);
      w *= ct;
      IDX(A, lda, j, c) -= w;
      for (int i = j + 1; i < m; i++) IDX(A, lda, i, c) -= IDX(A, lda, i, j) * w;
    }
  }
}

/* xORM2R / xUNM2R: C <- op(Q) C (side 'L') or C op(Q) (side 'R'), trans 'N' or 'C' */
void FN(ormqr)(int side, int trans, int m, int n, int k, const S* A, int lda, const S* tau, S* C,
               int ldc) {
  int left = side == 'L', notran = trans == 'N';
  int forward = (left && !notran) || (!left && notran);
  for (int q = 0; q < k; q++) {
    int j = forward ? q : k - 1 - q;
    S t = notran ? tau[j] : CONJ(tau[j]);
    int nq = left ? m : n;
    if (left) {
      for (int c = 0; c < n; c++) {
        S w = IDX(C, ldc, j, c);
        for (int i = j + 1; i < nq; i++) w += CONJ(IDX(A, lda, i, j)) * IDX(C, ldc, i, c);
        w *= t;
        IDX(C, ldc, j, c) -= w;
        for (int i = j + 1; i < nq; i++) IDX(C, ldc, i, c) -= IDX(A, lda, i, j) * w;
      }
    } else {
      for (int r = 0; r < m; r++) {
        S w = IDX(C, ldc, r, j);
        for (int i = j + 1; i < nq; i++) w += IDX(C, ldc, r, i) * IDX(A, lda, i, j);
        w *= t;
        IDX(C, ldc, r, j) -= w;
        for (int i = j + 1; i < nq; i++) IDX(C, ldc, r, i) -= w * CONJ(IDX(A, lda, i, j));
      }
    }
  }
}

/* xGELQ2: A = L Q, reflector i stored (conjugated for complex) in row i right of the diagonal */
static void FN(gelq2)(int m, int n, S* A, int lda, S* tau) {
  int k = m < n ? m : n;
  for (int i = 0; i < k; i++) {
#if IS_COMPLEX
    for (int c = i; c < n; c++) IDX(A, lda, i, c) = CONJ(IDX(A, lda, i, c));
#endif
    tau[i] = FN(larfg)(n - i, &IDX(A, lda, i, i), &IDX(A, lda, i, (i + 1 < n ? i + 1 : i)), lda);
    /* apply H(i) to A(i+1:m, i:n) from the right: row r <- row r - tau (row r . v) v^H */
    for (int r = i + 1; r < m; r++) {
      S w = IDX(A, lda, r, i);
      for (int c = i + 1; c < n; c++) w += IDX(A, lda, r, c) * IDX(A, lda, i, c);
      w *= tau[i];
      IDX(A, lda, r, i) -= w;
      for (int c = i + 1; c < n; c++) IDX(A, lda, r, c) -= w * CONJ(IDX(A, lda, i, c));
    }
#if IS_COMPLEX
    for (int c = i + 1; c < n; c++) IDX(A, lda, i, c) = CONJ(IDX(A, lda, i, c));
#endif
  }
}

/* xORML2 / xUNML2, side 'L' only: C <- op(Q) C with Q = H(k)^H ... H(1)^H from xGELQ2 */
static void FN(ormlq_left)(int trans, int m, int n, int k, const S* A, int lda, const S* tau, S* C,
                           int ldc) {
  int notran = trans == 'N';
  /* Q C = H(k)^H ... H(1)^H C: apply H(1)^H first -> forward; Q^H C = H(1) ... H(k) C -> reverse */
  for (int q = 0; q < k; q++) {
    int i = notran ? q : k - 1 - q;
    S t = notran ? CONJ(tau[i]) : tau[i];
    /* v = [1, conj(A(i, i+1:m))] for complex storage (stored conj(v)), real: A(i, i+1:m) */
    for (int c = 0; c < n; c++) {
      S w = IDX(C, ldc, i, c);
      for (int r = i + 1; r < m; r++) w += IDX(A, lda, i, r) * IDX(C, ldc, r, c);
      /* w = v^H C(:,c) with v_r = conj(stored) -> conj(v_r) = stored */
      w *= t;
      IDX(C, ldc, i, c) -= w;
      for (int r = i + 1; r < m; r++) IDX(C, ldc, r, c) -= CONJ(IDX(A, lda, i, r)) * w;
    }
  }
}

/* xGELS (full rank; no xLASCL rescaling).  B is max(m,n) x nrhs. */
void FN(gels)(int trans, int m, int n, int nrhs, S* A, int lda, S* B, int ldb, int* info) {
  *info = 0;
  int mn = m < n ? m : n;
  if (mn == 0) return;
  S* tau = (S*)malloc(sizeof(S) * (size_t)mn);
  int tpose = trans != 'N';
  if (m >= n) {
    FN(geqrf)(m, n, A, lda, tau);
    if (!tpose) {
      FN(ormqr)('L', 'C', m, nrhs, n, A, lda, tau, B, ldb);
      FN(trtrs)('U', 'N', 'N', n, nrhs, A, lda, B, ldb, info);
    } else {
      FN(trtrs)('U', 'C', 'N', n, nrhs, A, lda, B, ldb, info);
      if (*info == 0) {
        for (int c = 0; c < nrhs; c++) for (int i = n; i < m; i++) IDX(B, ldb, i, c) = 0;
        FN(ormqr)('L', 'N', m, nrhs, n, A, lda, tau, B, ldb);
      }
    }
  } else {
    FN(gelq2)(m, n, A, lda, tau);
    if (!tpose) {
      FN(trtrs)('L', 'N', 'N', m, nrhs, A, lda, B, ldb, info);
      if (*info == 0) {
        for (int c = 0; c < nrhs; c++) for (int i = m; i < n; i++) IDX(B, ldb, i, c) = 0;
        FN(ormlq_left)('C', n, nrhs, m, A, lda, tau, B, ldb);
      }
    } else {
      FN(ormlq_left)('N', n, nrhs, m, A, lda, tau, B, ldb);
      FN(trtrs)('L', 'C', 'N', m, nrhs, A, lda, B, ldb, info);
    }
  }
  free(tau);
}

/* inverses by solving against the identity (same results as xGETRI/xPOTRI/xTRTRI up to rounding) */
void FN(getri)(int n, S* A, int lda, const int* ipiv, int* info) {
  *info = 0;
  for (int i = 0; i < n; i++) if (IDX(A, lda, i, i) == 0) { *info = i + 1; return; }
  S* X = (S*)calloc((size_t)n * (size_t)n, sizeof(S));
  for (int i = 0; i < n; i++) X[(size_t)i + (size_t)i * n] = 1;
  FN(getrs)('N', n, n, A, lda, ipiv, X, n);
  for (int j = 0; j < n; j++) for (int i = 0; i < n; i++) IDX(A, lda, i, j) = X[(size_t)i + (size_t)j * n];
  free(X);
}

void FN(potri)(int uplo, int n, S* A, int lda, int* info) {
  *info = 0;
  for (int i = 0; i < n; i++) if (IDX(A, lda, i, i) == 0) { *info = i + 1; return; }
  S* X = (S*)calloc((size_t)n * (size_t)n, sizeof(S));
  for (int i = 0; i < n; i++) X[(size_t)i + (size_t)i * n] = 1;
  FN(potrs)(uplo, n, n, A, lda, X, n);
  for (int j = 0; j < n; j++)
    for (int i = 0; i < n; i++)
      if (uplo == 'L' ? i >= j : i <= j) IDX(A, lda, i, j) = X[(size_t)i + (size_t)j * n];
  free(X);
}

void FN(trtri)(int uplo, int diag, int n, S* A, int lda, int* info) {
  *info = 0;
  if (diag == 'N') for (int i = 0; i < n; i++) if (IDX(A, lda, i, i) == 0) { *info = i + 1; return; }
  S* X = (S*)calloc((size_t)n * (size_t)n, sizeof(S));
  for (int i = 0; i < n; i++) X[(size_t)i + (size_t)i * n] = 1;
  S one = 1;
  FN(trsm)('L', uplo, 'N', diag, n, n, &one, A, lda, X, n);
  for (int j = 0; j < n; j++)
    for (int i = 0; i < n; i++)
      if (uplo == 'L' ? i >= j : i <= j) IDX(A, lda, i, j) = X[(size_t)i + (size_t)j * n];
  free(X);
}

#undef IDX
