/* Forwards the Fortran-ABI BLAS/LAPACK names the reference calls (LinearAlgebra/d4est_blas.h:4-19) to the LP64
 * OpenBLAS bundled with scipy in this image (symbols carry a scipy_ prefix).  TEST INFRASTRUCTURE ONLY: used by
 * the TIMED reference binary (oracle/_ref/d4b_harness_blas: bench.py cpu_baseline / --impl reference), so that the
 * CPU baseline runs on a real BLAS as SURVEY 8(d) asks; the parity fixtures keep the naive shim (blas_shim.c). */
void scipy_dgemm_(const char *, const char *, const int *, const int *, const int *, const double *, const double *, const int *, const double *, const int *, const double *, double *, const int *);
void scipy_dgemv_(const char *, const int *, const int *, const double *, const double *, const int *, const double *, const int *, const double *, double *, const int *);
void scipy_daxpy_(const int *, const double *, const double *, const int *, double *, const int *);
void scipy_dscal_(const int *, const double *, double *, const int *);
double scipy_ddot_(const int *, const double *, const int *, const double *, const int *);
void scipy_dgetrf_(const int *, const int *, double *, const int *, int *, int *);
void scipy_dgetri_(const int *, double *, const int *, const int *, double *, const int *, int *);
void scipy_dsyev_(const char *, const char *, const int *, double *, const int *, double *, double *, const int *, int *);
void dgemm_(const char *ta, const char *tb, const int *m, const int *n, const int *k, const double *al, const double *A, const int *lda, const double *B, const int *ldb, const double *be, double *C, const int *ldc)
{ scipy_dgemm_(ta, tb, m, n, k, al, A, lda, B, ldb, be, C, ldc); }
void dgemv_(const char *t, const int *m, const int *n, const double *al, const double *A, const int *lda, const double *x, const int *ix, const double *be, double *y, const int *iy)
{ scipy_dgemv_(t, m, n, al, A, lda, x, ix, be, y, iy); }
void daxpy_(const int *n, const double *a, const double *x, const int *ix, double *y, const int *iy) { scipy_daxpy_(n, a, x, ix, y, iy); }
void dscal_(const int *n, const double *a, double *x, const int *ix) { scipy_dscal_(n, a, x, ix); }
double ddot_(const int *n, const double *x, const int *ix, const double *y, const int *iy) { return scipy_ddot_(n, x, ix, y, iy); }
void dgetrf_(const int *m, const int *n, double *A, const int *lda, int *piv, int *info) { scipy_dgetrf_(m, n, A, lda, piv, info); }
void dgetri_(const int *n, double *A, const int *lda, const int *piv, double *w, const int *lw, int *info) { scipy_dgetri_(n, A, lda, piv, w, lw, info); }
void dsyev_(const char *j, const char *u, const int *n, double *A, const int *lda, double *w, double *wk, const int *lw, int *info) { scipy_dsyev_(j, u, n, A, lda, w, wk, lw, info); }
