/* d4b_ops.h -- 1-D reference-element tables (host side). See d4b_ops.c. */
#ifndef D4B_OPS_H
#define D4B_OPS_H
#ifdef __cplusplus
extern "C" {
#endif
#define D4B_OPS_MAXN 32
void d4b_ops_gauss(int n, double *x, double *w);
void d4b_ops_gll(int deg, double *x, double *w);
void d4b_ops_quad(int degq, int quad_type, double *x, double *w);
void d4b_ops_diff(int deg, double *D);
void d4b_ops_interp(int deg, int degq, int quad_type, int child, double *I);
void d4b_ops_prolong(int degH, int degh, int child, double *P);
void d4b_ops_quadbasis(int degq, int quad_type, double *Dq, double *e0, double *e1);
#ifdef __cplusplus
}
#endif
#endif
