/* Minimal LP64 Fortran-ABI BLAS/LAPACK shim for the serial oracle build of the
 * d4est reference.  TEST INFRASTRUCTURE ONLY - never linked into the product.
 * The reference calls these through LinearAlgebra/d4est_blas.h:4-19 on operands
 * of at most a few hundred entries (1-D DG matrices), so textbook loops are
 * adequate; summation order differs from OpenBLAS at the 1e-16 level only. */
#include <stdlib.h>
#include <stdio.h>
#include <math.h>
#include <string.h>

static int is_t(const char *c) { return *c == 'T' || *c == 't' || *c == 'C' || *c == 'c'; }

/* C(m x n) = alpha op(A) op(B) + beta C, column-major */
void dgemm_(const char *ta, const char *tb, const int *m, const int *n, const int *k,
            const double *alpha, const double *A, const int *lda, const double *B,
            const int *ldb, const double *beta, double *C, const int *ldc)
{
  int M = *m, N = *n, K = *k, tA = is_t(ta), tB = is_t(tb);
  for (int j = 0; j < N; j++)
    for (int i = 0; i < M; i++) {
      double s = 0.;
      for (int l = 0; l < K; l++) {
        double a = tA ? A[l + (size_t)i * *lda] : A[i + (size_t)l * *lda];
        double b = tB ? B[j + (size_t)l * *ldb] : B[l + (size_t)j * *ldb];
        s += a * b;
      }
      double *c = &C[i + (size_t)j * *ldc];
      *c = (*beta == 0.) ? *alpha * s : *alpha * s + *beta * *c;
    }
}

void dgemv_(const char *t, const int *m, const int *n, const double *alpha, const double *A,
            const int *lda, const double *x, const int *incx, const double *beta, double *y,
            const int *incy)
{
  int M = *m, N = *n, tA = is_t(t);
  int ly = tA ? N : M, lx = tA ? M : N;
  for (int i = 0; i < ly; i++) {
    double s = 0.;
    for (int l = 0; l < lx; l++) {
      double a = tA ? A[l + (size_t)i * *lda] : A[i + (size_t)l * *lda];
      s += a * x[(size_t)l * *incx];
    }
    double *c = &y[(size_t)i * *incy];
    *c = (*beta == 0.) ? *alpha * s : *alpha * s + *beta * *c;
  }
}

void daxpy_(const int *n, const double *a, const double *x, const int *incx, double *y, const int *incy)
{ for (int i = 0; i < *n; i++) y[(size_t)i * *incy] += *a * x[(size_t)i * *incx]; }

void dscal_(const int *n, const double *a, double *x, const int *incx)
{ for (int i = 0; i < *n; i++) x[(size_t)i * *incx] *= *a; }

double ddot_(const int *n, const double *x, const int *incx, const double *y, const int *incy)
{ double s = 0.; for (int i = 0; i < *n; i++) s += x[(size_t)i * *incx] * y[(size_t)i * *incy]; return s; }

/* LU with partial pivoting, column-major, LAPACK conventions (1-based ipiv) */
void dgetrf_(const int *m, const int *n, double *A, const int *lda, int *ipiv, int *info)
{
  int M = *m, N = *n, L = *lda; *info = 0;
  int mn = M < N ? M : N;
  for (int j = 0; j < mn; j++) {
    int p = j; double mx = fabs(A[j + (size_t)j * L]);
    for (int i = j + 1; i < M; i++) if (fabs(A[i + (size_t)j * L]) > mx) { mx = fabs(A[i + (size_t)j * L]); p = i; }
    ipiv[j] = p + 1;
    if (mx == 0.) { if (!*info) *info = j + 1; continue; }
    if (p != j) for (int c = 0; c < N; c++) { double t = A[j + (size_t)c * L]; A[j + (size_t)c * L] = A[p + (size_t)c * L]; A[p + (size_t)c * L] = t; }
    for (int i = j + 1; i < M; i++) A[i + (size_t)j * L] /= A[j + (size_t)j * L];
    for (int c = j + 1; c < N; c++) {
      double f = A[j + (size_t)c * L];
      for (int i = j + 1; i < M; i++) A[i + (size_t)c * L] -= A[i + (size_t)j * L] * f;
    }
  }
}

/* inverse from LU: solve A X = I column by column */
void dgetri_(const int *n, double *A, const int *lda, const int *ipiv, double *work, const int *lwork, int *info)
{
  int N = *n, L = *lda; *info = 0; (void)work; (void)lwork;
  double *X = (double *)malloc(sizeof(double) * N * N);
  for (int c = 0; c < N; c++) {
    double *b = &X[(size_t)c * N];
    for (int i = 0; i < N; i++) b[i] = (i == c) ? 1. : 0.;
    for (int i = 0; i < N; i++) { int p = ipiv[i] - 1; if (p != i) { double t = b[i]; b[i] = b[p]; b[p] = t; } }
    for (int i = 0; i < N; i++) for (int l = 0; l < i; l++) b[i] -= A[i + (size_t)l * L] * b[l];
    for (int i = N - 1; i >= 0; i--) { for (int l = i + 1; l < N; l++) b[i] -= A[i + (size_t)l * L] * b[l]; b[i] /= A[i + (size_t)i * L]; }
  }
  for (int c = 0; c < N; c++) for (int i = 0; i < N; i++) A[i + (size_t)c * L] = X[i + (size_t)c * N];
  free(X);
}

/* symmetric eigen-solver: cyclic Jacobi, eigenvalues ascending; jobz 'V' returns vectors in A */
void dsyev_(const char *jobz, const char *uplo, const int *n, double *A, const int *lda, double *w,
            double *work, const int *lwork, int *info)
{
  int N = *n, L = *lda; *info = 0; (void)work;
  if (*lwork == -1) { work[0] = 3. * N; return; }
  double *S = (double *)malloc(sizeof(double) * N * N), *V = (double *)malloc(sizeof(double) * N * N);
  int up = (*uplo == 'U' || *uplo == 'u');
  for (int i = 0; i < N; i++) for (int j = 0; j < N; j++) {
    int r = i, c = j; if ((up && r > c) || (!up && r < c)) { r = j; c = i; }
    S[i + (size_t)j * N] = A[r + (size_t)c * L]; V[i + (size_t)j * N] = (i == j);
  }
  for (int sweep = 0; sweep < 100; sweep++) {
    double off = 0.; for (int i = 0; i < N; i++) for (int j = 0; j < i; j++) off += S[i + (size_t)j * N] * S[i + (size_t)j * N];
    if (off < 1e-300) break;
    for (int p = 0; p < N; p++) for (int q = p + 1; q < N; q++) {
      double apq = S[p + (size_t)q * N]; if (apq == 0.) continue;
      double th = (S[q + (size_t)q * N] - S[p + (size_t)p * N]) / (2. * apq);
      double t = (th >= 0 ? 1. : -1.) / (fabs(th) + sqrt(th * th + 1.)), c = 1. / sqrt(t * t + 1.), s = t * c;
      for (int k = 0; k < N; k++) { double a = S[k + (size_t)p * N], b = S[k + (size_t)q * N]; S[k + (size_t)p * N] = c * a - s * b; S[k + (size_t)q * N] = s * a + c * b; }
      for (int k = 0; k < N; k++) { double a = S[p + (size_t)k * N], b = S[q + (size_t)k * N]; S[p + (size_t)k * N] = c * a - s * b; S[q + (size_t)k * N] = s * a + c * b; }
      for (int k = 0; k < N; k++) { double a = V[k + (size_t)p * N], b = V[k + (size_t)q * N]; V[k + (size_t)p * N] = c * a - s * b; V[k + (size_t)q * N] = s * a + c * b; }
    }
  }
  int *idx = (int *)malloc(sizeof(int) * N);
  for (int i = 0; i < N; i++) idx[i] = i;
  for (int i = 0; i < N; i++) for (int j = i + 1; j < N; j++) if (S[idx[j] + (size_t)idx[j] * N] < S[idx[i] + (size_t)idx[i] * N]) { int t = idx[i]; idx[i] = idx[j]; idx[j] = t; }
  for (int i = 0; i < N; i++) w[i] = S[idx[i] + (size_t)idx[i] * N];
  if (*jobz == 'V' || *jobz == 'v') for (int j = 0; j < N; j++) for (int i = 0; i < N; i++) A[i + (size_t)j * L] = V[i + (size_t)idx[j] * N];
  free(S); free(V); free(idx);
}
