/* d4b_solver.cu -- device-resident Krylov/Chebyshev vector work and drivers.
 *   BLAS-1 (LinearAlgebra/d4est_linalg.c:179-235), Chebyshev smoother
 *   (Solver/d4est_solver_multigrid_smoother_cheby.c:81-176), CG spectral bound
 *   (Solver/d4est_solver_cg_eigs.c:115-276), flexible CG (Solver/d4est_solver_fcg_improved.c:97-346),
 *   plain CG (Solver/d4est_solver_cg.c:76).  All vectors are device pointers; scalars that steer the
 *   iteration come back to the host exactly where the reference does an sc_allreduce.
 */
#include "d4b_internal.h"
#include <cmath>
#include <cstdlib>
#include <cstring>

#define VB 256
static inline int vgrid(int64_t n) { int64_t g = (n + VB * 4 - 1) / (VB * 4); return (int)(g < 1 ? 1 : (g > 148 * 16 ? 148 * 16 : g)); }

__global__ void k_axpy(int64_t n, double a, const double *__restrict__ x, double *__restrict__ y)
{ for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) y[i] = fma(a, x[i], y[i]); }
__global__ void k_xpby(int64_t n, const double *__restrict__ x, double b, double *__restrict__ y)
{ for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) y[i] = fma(b, y[i], x[i]); }
/* z may alias x or y (each thread reads its element before it writes it): no __restrict__ here */
__global__ void k_axpyeqz(int64_t n, double a, const double *x, const double *y, double *z)
{ for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) z[i] = fma(a, x[i], y[i]); }
__global__ void k_scaleto2(int64_t n, double a, const double *__restrict__ x1, double *__restrict__ z1,
                           const double *__restrict__ x2, double *__restrict__ z2)
{ for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) { z1[i] = a * x1[i]; z2[i] = a * x2[i]; } }
__global__ void k_scale(int64_t n, double a, double *__restrict__ x)
{ for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) x[i] *= a; }
/* fused Chebyshev update (smoother_cheby.c:139-153): p = alpha (rhs - Au) + beta p;  u += p */
__global__ void k_cheby_update(int64_t n, double alpha, double beta, const double *__restrict__ rhs,
                               const double *__restrict__ Au, double *__restrict__ p, double *__restrict__ u)
{
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    double pi = fma(beta, p[i], alpha * (rhs[i] - Au[i]));
    p[i] = pi; u[i] += pi;
  }
}
/* up to 4 dot products in one sweep; deterministic two-stage reduction (no atomics) */
struct dot_args { const double *x[4], *y[4]; int k; };
__global__ void k_dots_partial(int64_t n, dot_args a, double *__restrict__ part)
{
  __shared__ double sh[4][VB / 32];
  double s[4] = {0., 0., 0., 0.};
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
#pragma unroll
    for (int k = 0; k < 4; k++) if (k < a.k) s[k] = fma(a.x[k][i], a.y[k][i], s[k]);
#pragma unroll
  for (int k = 0; k < 4; k++) {
    for (int o = 16; o > 0; o >>= 1) s[k] += __shfl_down_sync(0xffffffffu, s[k], o);
    if ((threadIdx.x & 31) == 0) sh[k][threadIdx.x >> 5] = s[k];
  }
  __syncthreads();
  if (threadIdx.x < 4) {
    double t = 0.;
    for (int w = 0; w < VB / 32; w++) t += sh[threadIdx.x][w];
    part[threadIdx.x * gridDim.x + blockIdx.x] = t;
  }
}
__global__ void k_dots_final(int nblk, int k, const double *__restrict__ part, double *__restrict__ out)
{
  __shared__ double sh[VB / 32];
  for (int j = 0; j < k; j++) {
    double s = 0.;
    for (int i = threadIdx.x; i < nblk; i += blockDim.x) s += part[j * nblk + i];
    for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) { double t = 0.; for (int w = 0; w < VB / 32; w++) t += sh[w]; out[j] = t; }
    __syncthreads();
  }
}

static double *g_part = nullptr, *g_out = nullptr, *g_out_h = nullptr;
static int ensure_scratch()
{
  if (g_part) return D4B_OK;
  D4B_CUDA(cudaMalloc((void **)&g_part, 4 * 148 * 16 * sizeof(double)));
  D4B_CUDA(cudaMalloc((void **)&g_out, 4 * sizeof(double)));
  D4B_CUDA(cudaMallocHost((void **)&g_out_h, 4 * sizeof(double)));
  return D4B_OK;
}
int d4b_allreduce_sum_host(double *vals, int k, cudaStream_t st); /* d4b_halo.cu: no-op on one rank */

/* k <= 4 dots, globally reduced, returned on the host */
static int dots(int64_t n, int k, const double **x, const double **y, double *out, cudaStream_t st)
{
  int rc = ensure_scratch(); if (rc) return rc;
  dot_args a; a.k = k;
  for (int j = 0; j < 4; j++) { a.x[j] = x[j < k ? j : 0]; a.y[j] = y[j < k ? j : 0]; }
  int g = vgrid(n);
  k_dots_partial<<<g, VB, 0, st>>>(n, a, g_part);
  k_dots_final<<<1, VB, 0, st>>>(g, k, g_part, g_out);
  d4b_count_launch(2);
  D4B_CUDA(cudaMemcpyAsync(g_out_h, g_out, k * sizeof(double), cudaMemcpyDeviceToHost, st));
  D4B_CUDA(cudaStreamSynchronize(st));
  for (int j = 0; j < k; j++) out[j] = g_out_h[j];
  return d4b_allreduce_sum_host(out, k, st);
}

#define ST ((cudaStream_t)stream)
#define LAUNCH1() do { D4B_CUDA(cudaGetLastError()); d4b_count_launch(1); } while (0)

extern "C" {

/* y += a x            (d4est_linalg_vec_axpy  linalg.c:208) */
int d4b_vec_axpy(int64_t n, double a, const double *x, double *y, void *stream)
{ k_axpy<<<vgrid(n), VB, 0, ST>>>(n, a, x, y); LAUNCH1(); return D4B_OK; }
/* y = x + b y         (d4est_linalg_vec_xpby  linalg.c:223) */
int d4b_vec_xpby(int64_t n, const double *x, double b, double *y, void *stream)
{ k_xpby<<<vgrid(n), VB, 0, ST>>>(n, x, b, y); LAUNCH1(); return D4B_OK; }
/* z = a x + y         (d4est_linalg_vec_axpyeqz) */
int d4b_vec_axpyeqz(int64_t n, double a, const double *x, const double *y, double *z, void *stream)
{ k_axpyeqz<<<vgrid(n), VB, 0, ST>>>(n, a, x, y, z); LAUNCH1(); return D4B_OK; }
/* x *= a              (d4est_linalg_vec_scale linalg.c:179) */
int d4b_vec_scale(int64_t n, double a, double *x, void *stream)
{ k_scale<<<vgrid(n), VB, 0, ST>>>(n, a, x); LAUNCH1(); return D4B_OK; }
int d4b_vec_copy(int64_t n, const double *x, double *y, void *stream)
{ D4B_CUDA(cudaMemcpyAsync(y, x, n * sizeof(double), cudaMemcpyDeviceToDevice, ST)); return D4B_OK; }
int d4b_vec_zero(int64_t n, double *x, void *stream)
{ D4B_CUDA(cudaMemsetAsync(x, 0, n * sizeof(double), ST)); return D4B_OK; }
/* x.y summed over all ranks, on the host (d4est_linalg_vec_dot linalg.c:230 + sc_allreduce) */
int d4b_vec_dot(int64_t n, const double *x, const double *y, double *result, void *stream)
{ const double *xs[1] = {x}, *ys[1] = {y}; return dots(n, 1, xs, ys, result, ST); }

/* Chebyshev semi-iteration: restates d4est_solver_multigrid_smoother_cheby_iterate_aux
 * (Solver/d4est_solver_multigrid_smoother_cheby.c:81-176).  work: 2 n doubles (Au, p).
 * On return r = rhs - A u (the residual the V-cycle restricts, multigrid.c:1054). */
int d4b_cheby_iterate(d4b_mesh_t *m, double *u, const double *rhs, double *r, int iter, double lmin,
                      double lmax, int residual_at_end, double *work, void *stream)
{
  const int64_t n = m->local_nodes;
  double *Au = work, *p = work + n;
  const double d = (lmax + lmin) * .5, c = (lmax - lmin) * .5;
  double alpha = 0., beta;
  int rc;
  D4B_CUDA(cudaMemsetAsync(p, 0, n * sizeof(double), ST));
  for (int i = 0; i < iter; i++) {
    if ((rc = d4b_apply_dev(m, u, Au, stream)) != D4B_OK) return rc;
    if (i == 0) alpha = 1. / d;
    else if (i == 1) alpha = 2. * d / (2 * d * d - c * c);
    else alpha = 1. / (d - (alpha * c * c / 4.));
    beta = alpha * d - 1.;
    k_cheby_update<<<vgrid(n), VB, 0, ST>>>(n, alpha, beta, rhs, Au, p, u); LAUNCH1();
  }
  if (residual_at_end) {
    if ((rc = d4b_apply_dev(m, u, Au, stream)) != D4B_OK) return rc;
    k_axpyeqz<<<vgrid(n), VB, 0, ST>>>(n, -1., Au, rhs, r); LAUNCH1();
  }
  return D4B_OK;
}

static void gershgorin(int i, int64_t local_nodes, double a0, double b0, double a1, double b1, int use_new,
                       double *mx)
{ /* Solver/d4est_solver_cg_eigs.c:7-59 (the default, use_new = 0, keeps the reference's bound) */
  double diag, off;
  if (!use_new) {
    if (i != 0 && i < local_nodes - 1) { diag = 1. / a1 + b0 / a0; off = fabs(sqrt(b1) / a1) + fabs(sqrt(b0) / a0); }
    else if (i == 0) { diag = 1. / a1; off = sqrt(b1) / a1; }
    else { diag = 1. / a1 + b0 / a0; off = fabs(sqrt(b0) / a0); }
  } else {
    if (i != 0) { diag = 1. / a1 + b0 / a0; off = fabs(sqrt(b0) / a0); }
    else { diag = 1. / a1; off = sqrt(b1) / a1; }
  }
  *mx = diag + off;
}

/* CG spectral bound: restates cg_eigs (Solver/d4est_solver_cg_eigs.c:115-276), including its side
 * effect of advancing u by the CG steps (:230).  work: 3 n doubles (Au, d, r). */
int d4b_cg_eigs(d4b_mesh_t *m, double *u, const double *rhs, int imax, int use_new, double *spectral_bound,
                double *work, void *stream)
{
  const int64_t n = m->local_nodes;
  double *Au = work, *d = work + n, *r = work + 2 * n;
  double delta_new, delta_old, alpha = -1., beta = -1., alpha_old, beta_old, dAu, tmax;
  int rc;
  if ((rc = d4b_apply_dev(m, u, Au, stream)) != D4B_OK) return rc;
  k_axpyeqz<<<vgrid(n), VB, 0, ST>>>(n, -1., Au, rhs, r); LAUNCH1();
  if ((rc = d4b_vec_copy(n, r, d, stream)) != D4B_OK) return rc;
  if ((rc = d4b_vec_dot(n, r, r, &delta_new, stream)) != D4B_OK) return rc;
  for (int i = 0; i < imax; i++) {
    if ((rc = d4b_apply_dev(m, d, Au, stream)) != D4B_OK) return rc;
    if ((rc = d4b_vec_dot(n, d, Au, &dAu, stream)) != D4B_OK) return rc;
    alpha_old = alpha; alpha = delta_new / dAu;
    k_axpy<<<vgrid(n), VB, 0, ST>>>(n, alpha, d, u); LAUNCH1();
    k_axpy<<<vgrid(n), VB, 0, ST>>>(n, -alpha, Au, r); LAUNCH1();
    delta_old = delta_new;
    if ((rc = d4b_vec_dot(n, r, r, &delta_new, stream)) != D4B_OK) return rc;
    beta_old = beta; beta = delta_new / delta_old;
    k_xpby<<<vgrid(n), VB, 0, ST>>>(n, r, beta, d); LAUNCH1();
    gershgorin(i, n, alpha_old, beta_old, alpha, beta, use_new, &tmax);
    *spectral_bound = (i > 0) ? (tmax > *spectral_bound ? tmax : *spectral_bound) : tmax;
  }
  return D4B_OK;
}

} /* extern C */

extern "C" {

/* Plain CG: restates d4est_solver_cg_solve (Solver/d4est_solver_cg.c:76-190).  work: 3 n doubles. */
int d4b_cg_solve(d4b_mesh_t *m, double *u, const double *rhs, int imax, double atol, double rtol,
                 int *iters, double *res2, double *work, void *stream)
{
  const int64_t n = m->local_nodes;
  double *Au = work, *d = work + n, *r = work + 2 * n;
  double delta_new, delta_0, delta_old, dAu;
  int rc, i;
  if ((rc = d4b_apply_dev(m, u, Au, stream)) != D4B_OK) return rc;
  k_axpyeqz<<<vgrid(n), VB, 0, ST>>>(n, -1., Au, rhs, r); LAUNCH1();
  if ((rc = d4b_vec_copy(n, r, d, stream)) != D4B_OK) return rc;
  if ((rc = d4b_vec_dot(n, r, r, &delta_new, stream)) != D4B_OK) return rc;
  delta_0 = delta_new;
  for (i = 0; i < imax && (delta_new > atol * atol + delta_0 * rtol * rtol); i++) {
    if ((rc = d4b_apply_dev(m, d, Au, stream)) != D4B_OK) return rc;
    if ((rc = d4b_vec_dot(n, d, Au, &dAu, stream)) != D4B_OK) return rc;
    double alpha = delta_new / dAu;
    k_axpy<<<vgrid(n), VB, 0, ST>>>(n, alpha, d, u); LAUNCH1();
    k_axpy<<<vgrid(n), VB, 0, ST>>>(n, -alpha, Au, r); LAUNCH1();
    delta_old = delta_new;
    if ((rc = d4b_vec_dot(n, r, r, &delta_new, stream)) != D4B_OK) return rc;
    k_xpby<<<vgrid(n), VB, 0, ST>>>(n, r, delta_new / delta_old, d); LAUNCH1();
  }
  if (iters) *iters = i;
  if (res2) *res2 = delta_new;
  return D4B_OK;
}

} /* extern C */
int g_fcg_textbook = 0;   /* d4b_set_option("fcg_textbook_update", ...) in d4b_api.cu */
extern "C" {

/* Flexible CG with a preconditioner callback on DEVICE vectors: restates d4est_solver_fcg_solve
 * (Solver/d4est_solver_fcg_improved.c:97-346); pc(ctx, r, z) plays d4est_krylov_pc_t.pc_apply
 * (Solver/d4est_krylov_pc.h:27-50); pc == NULL copies r (do_not_use_preconditioner).
 * The 2-4 dots of one iteration travel in ONE reduction (:245).  work: 6 n doubles.
 * Direction update: the reference calls d4est_linalg_vec_axpyeqz(-gamma/rho, d, v, d) (:262; :266 for q), and
 * that routine copies its third argument into the output BEFORE adding a*x (LinearAlgebra/d4est_linalg.c:217-221),
 * so with the output aliasing x the old direction is gone by the time it is used: what the reference computes is
 * d_k = (1 - gamma/rho) v_k, q_k = (1 - gamma/rho) w_k.  The iteration counts of the reference (the +-1 contract
 * of north_star) come from that arithmetic, so it is the default here; d4b_set_option("fcg_textbook_update", 1)
 * selects the update the source comments describe (d_k = v_k - gamma/rho d_{k-1}). */
int d4b_fcg_solve(d4b_mesh_t *m, double *u, const double *rhs, int imax, double atol, double rtol,
                  int (*pc)(void *ctx, const double *r_dev, double *z_dev, void *stream), void *pc_ctx,
                  int *iters, double *final_res, double *work, void *stream)
{
  const int64_t n = m->local_nodes;
  double *Au = work, *r = work + n, *w = work + 2 * n, *d = work + 3 * n, *v = work + 4 * n, *q = work + 5 * n;
  double r0, g[4] = {0, 0, 0, 0}, rho = 0.;
  int rc, k;
  if ((rc = d4b_apply_dev(m, u, Au, stream)) != D4B_OK) return rc;
  k_axpyeqz<<<vgrid(n), VB, 0, ST>>>(n, -1., Au, rhs, r); LAUNCH1();
  if ((rc = d4b_vec_dot(n, r, r, &r0, stream)) != D4B_OK) return rc;
  const double tol = atol + rtol * sqrt(r0);
  double last = sqrt(r0);
  for (k = 0; k < imax; k++) {
    if (pc) { if ((rc = pc(pc_ctx, r, v, stream)) != D4B_OK) return rc; }
    else if ((rc = d4b_vec_copy(n, r, v, stream)) != D4B_OK) return rc;
    if ((rc = d4b_apply_dev(m, v, w, stream)) != D4B_OK) return rc;
    const double *xs[4] = {v, v, v, r}, *ys[4] = {r, w, q, r};
    if ((rc = dots(n, k > 0 ? 4 : 2, xs, ys, g, ST)) != D4B_OK) return rc;
    const double alpha = g[0], beta = g[1];
    if (k > 0) {
      const double gamma = g[2];
      if (g_fcg_textbook) {
        k_axpyeqz<<<vgrid(n), VB, 0, ST>>>(n, -gamma / rho, d, v, d); LAUNCH1();
        k_axpyeqz<<<vgrid(n), VB, 0, ST>>>(n, -gamma / rho, q, w, q); LAUNCH1();
      } else {
        k_scaleto2<<<vgrid(n), VB, 0, ST>>>(n, 1. - gamma / rho, v, d, w, q); LAUNCH1();
      }
      rho = beta - (gamma * gamma) / rho;
    } else {
      rho = beta;
      if ((rc = d4b_vec_copy(n, v, d, stream)) != D4B_OK) return rc;
      if ((rc = d4b_vec_copy(n, w, q, stream)) != D4B_OK) return rc;
    }
    k_axpy<<<vgrid(n), VB, 0, ST>>>(n, alpha / rho, d, u); LAUNCH1();
    k_axpy<<<vgrid(n), VB, 0, ST>>>(n, -alpha / rho, q, r); LAUNCH1();
    if (k > 0) last = sqrt(g[3]);
    if (k > 0 && sqrt(g[3]) <= tol) break;
  }
  if (iters) *iters = k;
  if (final_res) *final_res = last;
  return D4B_OK;
}

} /* extern C */
