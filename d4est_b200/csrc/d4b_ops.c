/* d4b_ops.c -- 1-D reference-element tables (host, plain C, long-double build).
 *
 * Same mathematical objects as the reference's lazily built operator cache
 * (dGMath/d4est_operators.c: D :855, GLL->GL interp :411, p-prolong :995,
 * hp-prolong :944-993; nodes/weights tables fetched :727-829), but built from
 * closed forms (Newton on Legendre polynomials + barycentric Lagrange) instead of
 * Vandermonde inverses through LAPACK, so no BLAS/LAPACK dependency exists on the
 * product side.  tests/test_oracle.py::test_product_1d_tables_match_the_reference_tables checks
 * them against the reference's own tables (tests/golden/ref_ops_tables.fix, dumped by oracle/_ref/d4b_ops_dump).
 */
#include "d4b_ops.h"
#include <math.h>
#include <string.h>

typedef long double ld;

/* Legendre P_n(x) and derivative by the three-term recurrence */
static void legendre(int n, ld x, ld *p, ld *dp)
{
  ld p0 = 1.0L, p1 = x;
  if (n == 0) { *p = 1.0L; *dp = 0.0L; return; }
  for (int k = 2; k <= n; k++) {
    ld pk = ((2 * k - 1) * x * p1 - (k - 1) * p0) / k;
    p0 = p1; p1 = pk;
  }
  *p = p1;
  /* (1-x^2) P_n' = n (P_{n-1} - x P_n) */
  if (fabsl(x) == 1.0L) *dp = (x > 0 ? 1.0L : ((n % 2) ? 1.0L : -1.0L)) * n * (n + 1) / 2.0L;
  else *dp = n * (p0 - x * p1) / (1.0L - x * x);
}

void d4b_ops_gauss(int n, double *x, double *w)
{ /* n points, ascending */
  for (int i = 0; i < n; i++) {
    ld z = -cosl(3.14159265358979323846264338327950288L * (i + 0.75L) / (n + 0.5L));
    ld p, dp;
    for (int it = 0; it < 100; it++) {
      legendre(n, z, &p, &dp);
      ld dz = p / dp; z -= dz;
      if (fabsl(dz) < 1e-19L) break;
    }
    legendre(n, z, &p, &dp);
    x[i] = (double)z;
    w[i] = (double)(2.0L / ((1.0L - z * z) * dp * dp));
  }
  for (int i = 0; i < n / 2; i++) { /* enforce exact symmetry */
    double s = 0.5 * (x[n - 1 - i] - x[i]); x[i] = -s; x[n - 1 - i] = s;
    double ww = 0.5 * (w[i] + w[n - 1 - i]); w[i] = w[n - 1 - i] = ww;
  }
  if (n % 2) x[n / 2] = 0.0;
}

void d4b_ops_gll(int deg, double *x, double *w)
{ /* deg+1 points, ascending: -1, roots of P_deg', +1 */
  int n = deg;
  x[0] = -1.0; x[n] = 1.0;
  for (int i = 1; i < n; i++) {
    ld z = -cosl(3.14159265358979323846264338327950288L * i / n);
    for (int it = 0; it < 100; it++) {
      /* q = P_n', q' from Legendre ODE: (1-z^2) P'' = 2 z P' - n(n+1) P */
      ld p, dp; legendre(n, z, &p, &dp);
      ld ddp = (2 * z * dp - n * (n + 1) * p) / (1.0L - z * z);
      ld dz = dp / ddp; z -= dz;
      if (fabsl(dz) < 1e-19L) break;
    }
    x[i] = (double)z;
  }
  for (int i = 0; i <= n / 2; i++) {
    double s = 0.5 * (x[n - i] - x[i]); x[i] = -s; x[n - i] = s;
  }
  if (n % 2 == 0) x[n / 2] = 0.0;
  for (int i = 0; i <= n; i++) {
    ld p, dp; legendre(n, (ld)x[i], &p, &dp);
    w[i] = (double)(2.0L / (n * (n + 1) * p * p));
  }
}

static void bary_weights(int n, const double *x, ld *lam)
{
  for (int j = 0; j < n; j++) {
    ld p = 1.0L;
    for (int k = 0; k < n; k++) if (k != j) p *= ((ld)x[j] - (ld)x[k]);
    lam[j] = 1.0L / p;
  }
}

/* row[j] = l_j(y) for the Lagrange basis on nodes x[0..n) */
static void lagrange_row(int n, const double *x, const ld *lam, ld y, double *row)
{
  for (int j = 0; j < n; j++)
    if (y == (ld)x[j]) { memset(row, 0, n * sizeof(double)); row[j] = 1.0; return; }
  ld den = 0.0L; ld t[D4B_OPS_MAXN];
  for (int j = 0; j < n; j++) { t[j] = lam[j] / (y - (ld)x[j]); den += t[j]; }
  for (int j = 0; j < n; j++) row[j] = (double)(t[j] / den);
}

void d4b_ops_diff(int deg, double *D)
{ /* D[i*N+j] = l_j'(x_i) */
  int N = deg + 1; double x[D4B_OPS_MAXN], w[D4B_OPS_MAXN]; ld lam[D4B_OPS_MAXN];
  d4b_ops_gll(deg, x, w); bary_weights(N, x, lam);
  for (int i = 0; i < N; i++) {
    ld s = 0.0L;
    for (int j = 0; j < N; j++) if (j != i) {
      ld v = (lam[j] / lam[i]) / ((ld)x[i] - (ld)x[j]);
      D[i * N + j] = (double)v; s += v;
    }
    D[i * N + i] = (double)(-s);
  }
}

void d4b_ops_quad(int degq, int quad_type, double *x, double *w)
{
  if (quad_type == 1) d4b_ops_gll(degq, x, w); else d4b_ops_gauss(degq + 1, x, w);
}

/* I[q*N+j] = l_j(y_q): y_q = quad point (child<0), or (quad point -/+ 1)/2 for child 0/1 */
void d4b_ops_interp(int deg, int degq, int quad_type, int child, double *I)
{
  int N = deg + 1, Nq = degq + 1;
  double x[D4B_OPS_MAXN], w[D4B_OPS_MAXN], y[D4B_OPS_MAXN], wy[D4B_OPS_MAXN]; ld lam[D4B_OPS_MAXN];
  d4b_ops_gll(deg, x, w); bary_weights(N, x, lam);
  d4b_ops_quad(degq, quad_type, y, wy);
  for (int q = 0; q < Nq; q++) {
    ld yy = (ld)y[q];
    if (child == 0) yy = 0.5L * yy - 0.5L; else if (child == 1) yy = 0.5L * yy + 0.5L;
    lagrange_row(N, x, lam, yy, &I[q * N]);
  }
}

/* P[i*NH+j] = l^H_j(x^h_i) (child<0) or l^H_j((x^h_i -/+ 1)/2) (child 0/1);
 * reference: p-prolong dGMath/d4est_operators.c:995, hp-prolong :944-993 */
void d4b_ops_prolong(int degH, int degh, int child, double *P)
{
  int NH = degH + 1, Nh = degh + 1;
  double xH[D4B_OPS_MAXN], xh[D4B_OPS_MAXN], w[D4B_OPS_MAXN]; ld lam[D4B_OPS_MAXN];
  d4b_ops_gll(degH, xH, w); bary_weights(NH, xH, lam); d4b_ops_gll(degh, xh, w);
  for (int i = 0; i < Nh; i++) {
    ld yy = (ld)xh[i];
    if (child == 0) yy = 0.5L * yy - 0.5L; else if (child == 1) yy = 0.5L * yy + 0.5L;
    lagrange_row(NH, xH, lam, yy, &P[i * NH]);
  }
}

/* Lagrange basis ON the quadrature points (valid representation of a degree-deg polynomial
 * whenever degq >= deg): Dq[a*Nq+m] = l~_m'(y_a), e0[m] = l~_m(-1), e1[m] = l~_m(+1).
 * Used by the register-blocked kernel (d4b_apply_v3.cu): face traces and lifts become
 * rank-1 operations along the face normal. */
void d4b_ops_quadbasis(int degq, int quad_type, double *Dq, double *e0, double *e1)
{
  int Nq = degq + 1; double y[D4B_OPS_MAXN], wy[D4B_OPS_MAXN]; ld lam[D4B_OPS_MAXN];
  d4b_ops_quad(degq, quad_type, y, wy); bary_weights(Nq, y, lam);
  for (int i = 0; i < Nq; i++) {
    ld s = 0.0L;
    for (int j = 0; j < Nq; j++) if (j != i) {
      ld v = (lam[j] / lam[i]) / ((ld)y[i] - (ld)y[j]);
      Dq[i * Nq + j] = (double)v; s += v;
    }
    Dq[i * Nq + i] = (double)(-s);
  }
  lagrange_row(Nq, y, lam, -1.0L, e0);
  lagrange_row(Nq, y, lam, 1.0L, e1);
}
