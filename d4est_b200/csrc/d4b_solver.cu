/* d4b_solver.cu -- device-resident Krylov/Chebyshev vector work and drivers.
 *   BLAS-1 (LinearAlgebra/d4est_linalg.c:179-235), Chebyshev smoother
 *   (Solver/d4est_solver_multigrid_smoother_cheby.c:81-176), CG spectral bound
 *   (Solver/d4est_solver_cg_eigs.c:115-276), flexible CG (Solver/d4est_solver_fcg_improved.c:97-346),
 *   plain CG (Solver/d4est_solver_cg.c:76).
 *
 * The scalars that steer an iteration (alpha, beta, the Gershgorin bound, the Chebyshev coefficients) stay in
 * device memory: a reduction ends in a per-level scalar block, the next kernel reads it from there, and -- with
 * more than one rank -- an in-stream ncclAllReduce on that block sits in between (d4b_allreduce_sum_dev).  cg_eigs,
 * the Chebyshev sweep and therefore a whole V-cycle enqueue WITHOUT a host synchronisation; the host reads a
 * scalar back only where the caller asks for it (the spectral bound of the public d4b_cg_eigs, the convergence
 * test of the Krylov drivers: once per iteration, where the reference does its sc_allreduce).
 * Update pairs are fused (u and r of a CG step with the partial sums of the next r.r; d and q, u and r of FCG;
 * the five BLAS-1 sweeps of a Chebyshev iteration).  Reductions are two-stage and deterministic (no atomics);
 * their scratch belongs to the level, so solves on different levels / streams do not share state.
 */
#include "d4b_internal.h"
#include <cmath>
#include <cstdlib>
#include <cstring>

#define VB 256
#define VGRID_MAX (148 * 16)
static inline int vgrid(int64_t n) { int64_t g = (n + VB * 4 - 1) / (VB * 4); return (int)(g < 1 ? 1 : (g > VGRID_MAX ? VGRID_MAX : g)); }
#define GS_LOOP(i, n) for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < (n); i += (int64_t)gridDim.x * blockDim.x)

/* slots of a level's scalar block (device doubles) */
enum { S_DOT = 0 /* 4 raw dot results */, S_DELTA = 4 /* 2-slot ring of r.r */, S_DAU = 6, S_ALPHA, S_ALPHA_OLD, S_BETA,
       S_BETA_OLD, S_TMAX, S_R2, S_N };

__global__ void k_axpy(int64_t n, double a, const double *__restrict__ x, double *__restrict__ y)
{ GS_LOOP(i, n) y[i] = fma(a, x[i], y[i]); }
__global__ void k_xpby(int64_t n, const double *__restrict__ x, double b, double *__restrict__ y)
{ GS_LOOP(i, n) y[i] = fma(b, y[i], x[i]); }
/* z may alias x or y (each thread reads its element before it writes it): no __restrict__ here */
__global__ void k_axpyeqz(int64_t n, double a, const double *x, const double *y, double *z)
{ GS_LOOP(i, n) z[i] = fma(a, x[i], y[i]); }
__global__ void k_scale(int64_t n, double a, double *__restrict__ x)
{ GS_LOOP(i, n) x[i] *= a; }
/* FCG direction update, both vectors in one sweep.  textbook = 0: d = c v, q = c w (what the reference's aliased
 * axpyeqz calls compute, c = 1 - gamma/rho); textbook = 1: d = v + a d, q = w + a q (a = -gamma/rho) */
__global__ void k_fcg_dir(int64_t n, int textbook, double c, const double *__restrict__ v, const double *__restrict__ w,
                          double *__restrict__ d, double *__restrict__ q)
{
  if (textbook) { GS_LOOP(i, n) { d[i] = fma(c, d[i], v[i]); q[i] = fma(c, q[i], w[i]); } }
  else { GS_LOOP(i, n) { d[i] = c * v[i]; q[i] = c * w[i]; } }
}
/* u += a d; r -= a q */
__global__ void k_fcg_step(int64_t n, double a, const double *__restrict__ d, const double *__restrict__ q,
                           double *__restrict__ u, double *__restrict__ r)
{ GS_LOOP(i, n) { u[i] = fma(a, d[i], u[i]); r[i] = fma(-a, q[i], r[i]); } }
/* fused Chebyshev update (smoother_cheby.c:139-153): p = alpha (rhs - Au) + beta p;  u += p; coefficients from
 * device memory (coef[0] = alpha_i, coef[1] = beta_i) or, when coef == NULL, from the arguments */
__global__ void k_cheby_update(int64_t n, const double *__restrict__ coef, double alpha, double beta,
                               const double *__restrict__ rhs, const double *__restrict__ Au, double *__restrict__ p,
                               double *__restrict__ u)
{
  if (coef) { alpha = coef[0]; beta = coef[1]; }
  GS_LOOP(i, n) {
    double pi = fma(beta, p[i], alpha * (rhs[i] - Au[i]));
    p[i] = pi; u[i] += pi;
  }
}
/* Chebyshev coefficients of one smoother call, on the device: eig = *src * mult is stored in *dst_eig, then
 * alpha_i, beta_i for i < iter with [lmin, lmax] = [eig / ratio, eig] (smoother_cheby.c:100-137) */
__global__ void k_cheby_coefs(const double *src, double mult, double *dst_eig, double ratio, int iter, double *coef)
{
  if (threadIdx.x || blockIdx.x) return;
  const double eig = *src * mult;
  *dst_eig = eig;
  const double lmax = eig, lmin = eig / ratio, d = (lmax + lmin) * .5, c = (lmax - lmin) * .5;
  double alpha = 0.;
  for (int i = 0; i < iter; i++) {
    if (i == 0) alpha = 1. / d;
    else if (i == 1) alpha = 2. * d / (2 * d * d - c * c);
    else alpha = 1. / (d - (alpha * c * c / 4.));
    coef[2 * i] = alpha; coef[2 * i + 1] = alpha * d - 1.;
  }
}

/* ---- reductions: up to 4 dot products in one sweep; two-stage, deterministic */
struct dot_args { const double *x[4], *y[4]; int k; };
__device__ __forceinline__ void block_partials(double (&s)[4], int k, double *__restrict__ part)
{
  __shared__ double sh[4][VB / 32];
#pragma unroll
  for (int j = 0; j < 4; j++) {
    for (int o = 16; o > 0; o >>= 1) s[j] += __shfl_down_sync(0xffffffffu, s[j], o);
    if ((threadIdx.x & 31) == 0) sh[j][threadIdx.x >> 5] = s[j];
  }
  __syncthreads();
  if (threadIdx.x < k) {
    double t = 0.;
    for (int w = 0; w < VB / 32; w++) t += sh[threadIdx.x][w];
    part[threadIdx.x * gridDim.x + blockIdx.x] = t;
  }
}
__global__ void k_dots_partial(int64_t n, dot_args a, double *__restrict__ part)
{
  double s[4] = {0., 0., 0., 0.};
  GS_LOOP(i, n)
#pragma unroll
    for (int k = 0; k < 4; k++) if (k < a.k) s[k] = fma(a.x[k][i], a.y[k][i], s[k]);
  block_partials(s, a.k, part);
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
/* r = rhs - Au, d = r, and the partial sums of r.r in one sweep (start of cg_eigs / cg_solve) */
__global__ void k_cg_init(int64_t n, const double *__restrict__ rhs, const double *__restrict__ Au, double *__restrict__ r,
                          double *__restrict__ d, double *__restrict__ part)
{
  double s[4] = {0., 0., 0., 0.};
  GS_LOOP(i, n) { const double ri = rhs[i] - Au[i]; r[i] = ri; if (d) d[i] = ri; s[0] = fma(ri, ri, s[0]); }
  block_partials(s, 1, part);
}
/* one CG step with alpha = delta / (d.Au) read from the scalar block: u += alpha d, r -= alpha Au, and the partial
 * sums of the new r.r (cg_eigs.c:222-236, cg.c:160-178).  Thread 0 keeps alpha for the Gershgorin bound. */
__global__ void k_cg_step(int64_t n, double *__restrict__ sc, int ring, const double *__restrict__ d,
                          const double *__restrict__ Au, double *__restrict__ u, double *__restrict__ r,
                          double *__restrict__ part)
{
  const double alpha = sc[S_DELTA + ring] / sc[S_DAU];
  double s[4] = {0., 0., 0., 0.};
  GS_LOOP(i, n) { u[i] = fma(alpha, d[i], u[i]); const double ri = fma(-alpha, Au[i], r[i]); r[i] = ri; s[0] = fma(ri, ri, s[0]); }
  block_partials(s, 1, part);
  if (blockIdx.x == 0 && threadIdx.x == 0) { sc[S_ALPHA_OLD] = sc[S_ALPHA]; sc[S_ALPHA] = alpha; }
}
/* d = r + beta d with beta = delta_new / delta_old from the ring; thread 0 updates the Gershgorin bound of the
 * Lanczos matrix (cg_eigs.c:7-59, :251-262; use_new = 0 keeps the reference's default variant) */
__global__ void k_cg_dir(int64_t n, double *__restrict__ sc, int ring, const double *__restrict__ r, double *__restrict__ d,
                         int it, int use_new, int track_bound)
{
  const double beta = sc[S_DELTA + (ring ^ 1)] / sc[S_DELTA + ring];
  GS_LOOP(i, n) d[i] = fma(beta, d[i], r[i]);
  if (track_bound && blockIdx.x == 0 && threadIdx.x == 0) {
    const double a0 = sc[S_ALPHA_OLD], b0 = sc[S_BETA], a1 = sc[S_ALPHA], b1 = beta;   /* S_BETA still holds beta_{i-1} */
    double diag, off;
    if (!use_new) {
      if (it != 0 && it < n - 1) { diag = 1. / a1 + b0 / a0; off = fabs(sqrt(b1) / a1) + fabs(sqrt(b0) / a0); }
      else if (it == 0) { diag = 1. / a1; off = sqrt(b1) / a1; }
      else { diag = 1. / a1 + b0 / a0; off = fabs(sqrt(b0) / a0); }
    } else {
      if (it != 0) { diag = 1. / a1 + b0 / a0; off = fabs(sqrt(b0) / a0); }
      else { diag = 1. / a1; off = sqrt(b1) / a1; }
    }
    const double t = diag + off;
    sc[S_TMAX] = it > 0 ? (t > sc[S_TMAX] ? t : sc[S_TMAX]) : t;
    sc[S_BETA_OLD] = b0; sc[S_BETA] = beta;
  }
}

int d4b_allreduce_sum_dev(double *dev, int k, cudaStream_t st); /* d4b_halo.cu: no-op on one rank */

static int ensure_scal(d4b_mesh *m)
{
  if (m->sc) return D4B_OK;
  D4B_CUDA(cudaMalloc((void **)&m->sc, (S_N + 2 * 64) * sizeof(double)));   /* scalar block + Chebyshev coefficients */
  D4B_CUDA(cudaMemset(m->sc, 0, (S_N + 2 * 64) * sizeof(double)));
  D4B_CUDA(cudaMalloc((void **)&m->dot_part, 4 * VGRID_MAX * sizeof(double)));
  D4B_CUDA(cudaMallocHost((void **)&m->sc_h, S_N * sizeof(double)));
  return D4B_OK;
}

#define ST ((cudaStream_t)stream)
#define LAUNCH(n_) do { D4B_CUDA(cudaGetLastError()); d4b_count_launch(n_); } while (0)

/* k <= 4 dots of length n into out[0..k) (device, inside the level's scalar block), reduced over all ranks */
static int dots_dev(d4b_mesh *m, int64_t n, int k, const double **x, const double **y, double *out, cudaStream_t st)
{
  dot_args a; a.k = k;
  for (int j = 0; j < 4; j++) { a.x[j] = x[j < k ? j : 0]; a.y[j] = y[j < k ? j : 0]; }
  const int g = vgrid(n);
  k_dots_partial<<<g, VB, 0, st>>>(n, a, m->dot_part);
  k_dots_final<<<1, VB, 0, st>>>(g, k, m->dot_part, out);
  LAUNCH(2);
  return d4b_allreduce_sum_dev(out, k, st);
}
static int finish_partials(d4b_mesh *m, int64_t n, double *out, cudaStream_t st)
{
  k_dots_final<<<1, VB, 0, st>>>(vgrid(n), 1, m->dot_part, out);
  LAUNCH(1);
  return d4b_allreduce_sum_dev(out, 1, st);
}
/* copy k scalars of the block to the host and wait (the one place a driver synchronises) */
static int read_scalars(d4b_mesh *m, int first, int k, double *host, cudaStream_t st)
{
  D4B_CUDA(cudaMemcpyAsync(m->sc_h + first, m->sc + first, k * sizeof(double), cudaMemcpyDeviceToHost, st));
  D4B_CUDA(cudaStreamSynchronize(st));
  for (int j = 0; j < k; j++) host[j] = m->sc_h[first + j];
  return D4B_OK;
}

/* ---- global (mesh-less) dot for d4b_vec_dot: its own scratch; one stream at a time */
static double *g_part = nullptr, *g_out = nullptr, *g_out_h = nullptr;
static int ensure_scratch()
{
  if (g_part) return D4B_OK;
  D4B_CUDA(cudaMalloc((void **)&g_part, 4 * VGRID_MAX * sizeof(double)));
  D4B_CUDA(cudaMalloc((void **)&g_out, 4 * sizeof(double)));
  D4B_CUDA(cudaMallocHost((void **)&g_out_h, 4 * sizeof(double)));
  return D4B_OK;
}

int g_fcg_textbook = 0;   /* d4b_set_option("fcg_textbook_update", ...) in d4b_api.cu */

/* ---- internal, sync-free forms used by the multigrid (d4b_mg.cu) -------------------------------------------- */
/* cg_eigs with the bound left in the level's scalar block (S_TMAX); u is advanced by the CG steps (cg_eigs.c:230) */
int d4b_cg_eigs_dev(d4b_mesh *m, double *u, const double *rhs, int imax, int use_new, double *work, cudaStream_t st)
{
  int rc = ensure_scal(m); if (rc) return rc;
  const int64_t n = m->local_nodes;
  double *Au = work, *d = work + n, *r = work + 2 * n, *sc = m->sc;
  if ((rc = d4b_apply_dev(m, u, Au, st)) != D4B_OK) return rc;
  k_cg_init<<<vgrid(n), VB, 0, st>>>(n, rhs, Au, r, d, m->dot_part); LAUNCH(1);
  if ((rc = finish_partials(m, n, sc + S_DELTA, st)) != D4B_OK) return rc;
  int ring = 0;
  for (int i = 0; i < imax; i++) {
    if ((rc = d4b_apply_dev(m, d, Au, st)) != D4B_OK) return rc;
    const double *xs[1] = {d}, *ys[1] = {Au};
    if ((rc = dots_dev(m, n, 1, xs, ys, sc + S_DAU, st)) != D4B_OK) return rc;
    k_cg_step<<<vgrid(n), VB, 0, st>>>(n, sc, ring, d, Au, u, r, m->dot_part); LAUNCH(1);
    if ((rc = finish_partials(m, n, sc + S_DELTA + (ring ^ 1), st)) != D4B_OK) return rc;
    k_cg_dir<<<vgrid(n), VB, 0, st>>>(n, sc, ring, r, d, i, use_new, 1); LAUNCH(1);
    ring ^= 1;
  }
  return D4B_OK;
}
/* Chebyshev sweep whose interval comes from a DEVICE scalar: eig = *eig_src * mult is stored to *eig_dst (both
 * device pointers), [lmin, lmax] = [eig / ratio, eig].  On return r = rhs - A u.  work: 2 n doubles. */
int d4b_cheby_dev(d4b_mesh *m, double *u, const double *rhs, double *r, int iter, const double *eig_src, double mult,
                  double *eig_dst, double ratio, double *work, cudaStream_t st)
{
  int rc = ensure_scal(m); if (rc) return rc;
  if (iter > 64) { d4b_set_error("cheby_imax %d > 64", iter); return D4B_EINVAL; }
  const int64_t n = m->local_nodes;
  double *Au = work, *p = work + n, *coef = m->sc + S_N;
  k_cheby_coefs<<<1, 1, 0, st>>>(eig_src, mult, eig_dst, ratio, iter, coef); LAUNCH(1);
  D4B_CUDA(cudaMemsetAsync(p, 0, n * sizeof(double), st));
  for (int i = 0; i < iter; i++) {
    if ((rc = d4b_apply_dev(m, u, Au, st)) != D4B_OK) return rc;
    k_cheby_update<<<vgrid(n), VB, 0, st>>>(n, coef + 2 * i, 0., 0., rhs, Au, p, u); LAUNCH(1);
  }
  if ((rc = d4b_apply_dev(m, u, Au, st)) != D4B_OK) return rc;
  k_axpyeqz<<<vgrid(n), VB, 0, st>>>(n, -1., Au, rhs, r); LAUNCH(1);
  return D4B_OK;
}
double *d4b_scalar_tmax(d4b_mesh *m) { return ensure_scal(m) == D4B_OK ? m->sc + S_TMAX : nullptr; }
/* r.r into the level's scalar block (S_R2), no synchronisation; read with d4b_read_r2 */
int d4b_norm2_dev(d4b_mesh *m, const double *r, cudaStream_t st)
{
  int rc = ensure_scal(m); if (rc) return rc;
  const double *xs[1] = {r}, *ys[1] = {r};
  return dots_dev(m, m->local_nodes, 1, xs, ys, m->sc + S_R2, st);
}
int d4b_read_r2(d4b_mesh *m, double *r2, cudaStream_t st) { return read_scalars(m, S_R2, 1, r2, st); }

extern "C" {

/* y += a x            (d4est_linalg_vec_axpy  linalg.c:208) */
int d4b_vec_axpy(int64_t n, double a, const double *x, double *y, void *stream)
{ k_axpy<<<vgrid(n), VB, 0, ST>>>(n, a, x, y); LAUNCH(1); return D4B_OK; }
/* y = x + b y         (d4est_linalg_vec_xpby  linalg.c:223) */
int d4b_vec_xpby(int64_t n, const double *x, double b, double *y, void *stream)
{ k_xpby<<<vgrid(n), VB, 0, ST>>>(n, x, b, y); LAUNCH(1); return D4B_OK; }
/* z = a x + y         (d4est_linalg_vec_axpyeqz; z may alias x or y) */
int d4b_vec_axpyeqz(int64_t n, double a, const double *x, const double *y, double *z, void *stream)
{ k_axpyeqz<<<vgrid(n), VB, 0, ST>>>(n, a, x, y, z); LAUNCH(1); return D4B_OK; }
/* x *= a              (d4est_linalg_vec_scale linalg.c:179) */
int d4b_vec_scale(int64_t n, double a, double *x, void *stream)
{ k_scale<<<vgrid(n), VB, 0, ST>>>(n, a, x); LAUNCH(1); return D4B_OK; }
int d4b_vec_copy(int64_t n, const double *x, double *y, void *stream)
{ D4B_CUDA(cudaMemcpyAsync(y, x, n * sizeof(double), cudaMemcpyDeviceToDevice, ST)); return D4B_OK; }
int d4b_vec_zero(int64_t n, double *x, void *stream)
{ D4B_CUDA(cudaMemsetAsync(x, 0, n * sizeof(double), ST)); return D4B_OK; }
/* x.y summed over all ranks, on the host (d4est_linalg_vec_dot linalg.c:230 + sc_allreduce) */
int d4b_vec_dot(int64_t n, const double *x, const double *y, double *result, void *stream)
{
  int rc = ensure_scratch(); if (rc) return rc;
  dot_args a; a.k = 1;
  for (int j = 0; j < 4; j++) { a.x[j] = x; a.y[j] = y; }
  const int g = vgrid(n);
  k_dots_partial<<<g, VB, 0, ST>>>(n, a, g_part);
  k_dots_final<<<1, VB, 0, ST>>>(g, 1, g_part, g_out);
  LAUNCH(2);
  if ((rc = d4b_allreduce_sum_dev(g_out, 1, ST)) != D4B_OK) return rc;
  D4B_CUDA(cudaMemcpyAsync(g_out_h, g_out, sizeof(double), cudaMemcpyDeviceToHost, ST));
  D4B_CUDA(cudaStreamSynchronize(ST));
  *result = g_out_h[0];
  return D4B_OK;
}

/* ---- allocation-free dot product ---------------------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  Same two kernels as d4b_vec_dot
 * (k_dots_partial + k_dots_final, same grid, so the same bits), but the block partials live in a file-scope device
 * array instead of cudaMalloc'ed scratch and the result is written to a caller-supplied DEVICE double.  The array
 * is 4 * VGRID_MAX doubles = 75 776 bytes and stays resident on every GPU that loads the library.  In host code
 * `g_noalloc_part` is only the host-side stand-in of the symbol: the device address comes from
 * cudaGetSymbolAddress, once per device, with that device current.  Every call shares the array: calls that
 * overlap on the GPU (different streams) race on it; calls on one stream are ordered.  Nothing in it has to start
 * at zero -- the first kernel overwrites every partial the second one reads. */
#define D4B_NOALLOC_MAX_DEV 64
static __device__ __align__(256) double g_noalloc_part[4 * VGRID_MAX];
static double *g_noalloc_ptr[D4B_NOALLOC_MAX_DEV];

int d4b_vec_dot_noalloc(int64_t n, const double *x, const double *y, double *result_dev, void *stream)
{
  if (n < 0 || n > D4B_NOALLOC_DOT_MAX_N) {
    d4b_set_error("d4b_vec_dot_noalloc: n = %lld is outside [0, %lld] (D4B_NOALLOC_DOT_MAX_N)", (long long)n,
                  (long long)D4B_NOALLOC_DOT_MAX_N);
    return D4B_EINVAL;
  }
  if (!result_dev || (n > 0 && (!x || !y))) { d4b_set_error("d4b_vec_dot_noalloc: NULL vector or result pointer"); return D4B_EINVAL; }
  int dev = -1;
  D4B_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= D4B_NOALLOC_MAX_DEV) { d4b_set_error("d4b_vec_dot_noalloc: device ordinal %d is not supported", dev); return D4B_EINVAL; }
  if (!g_noalloc_ptr[dev]) {   /* set-up, once per device: resolve the symbol's device address */
    double *ptr = nullptr;
    D4B_CUDA(cudaGetSymbolAddress((void **)&ptr, g_noalloc_part));
    if (!ptr) { d4b_set_error("d4b_vec_dot_noalloc: cudaGetSymbolAddress returned NULL"); return D4B_ECUDA; }
    g_noalloc_ptr[dev] = ptr;
  }
  double *part = g_noalloc_ptr[dev];
  dot_args a; a.k = 1;
  for (int j = 0; j < 4; j++) { a.x[j] = x; a.y[j] = y; }
  const int g = vgrid(n);
  k_dots_partial<<<g, VB, 0, ST>>>(n, a, part);
  k_dots_final<<<1, VB, 0, ST>>>(g, 1, part, result_dev);
  LAUNCH(2);
  return D4B_OK;
}

/* Chebyshev semi-iteration with a HOST interval: restates d4est_solver_multigrid_smoother_cheby_iterate_aux
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
    k_cheby_update<<<vgrid(n), VB, 0, ST>>>(n, nullptr, alpha, beta, rhs, Au, p, u); LAUNCH(1);
  }
  if (residual_at_end) {
    if ((rc = d4b_apply_dev(m, u, Au, stream)) != D4B_OK) return rc;
    k_axpyeqz<<<vgrid(n), VB, 0, ST>>>(n, -1., Au, rhs, r); LAUNCH(1);
  }
  return D4B_OK;
}

/* CG spectral bound: restates cg_eigs (Solver/d4est_solver_cg_eigs.c:115-276), including its side
 * effect of advancing u by the CG steps (:230).  work: 3 n doubles (Au, d, r).  The bound is read back once. */
int d4b_cg_eigs(d4b_mesh_t *m, double *u, const double *rhs, int imax, int use_new, double *spectral_bound,
                double *work, void *stream)
{
  int rc = d4b_cg_eigs_dev(m, u, rhs, imax, use_new, work, ST);
  if (rc != D4B_OK) return rc;
  return read_scalars(m, S_TMAX, 1, spectral_bound, ST);
}

/* Plain CG: restates d4est_solver_cg_solve (Solver/d4est_solver_cg.c:76-190).  work: 3 n doubles.  The stopping
 * test needs r.r on the host once per iteration (where the reference has its sc_allreduce). */
int d4b_cg_solve(d4b_mesh_t *m, double *u, const double *rhs, int imax, double atol, double rtol,
                 int *iters, double *res2, double *work, void *stream)
{
  int rc = ensure_scal(m); if (rc) return rc;
  const int64_t n = m->local_nodes;
  double *Au = work, *d = work + n, *r = work + 2 * n, *sc = m->sc;
  double delta_new, delta_0;
  int i, ring = 0;
  if ((rc = d4b_apply_dev(m, u, Au, stream)) != D4B_OK) return rc;
  k_cg_init<<<vgrid(n), VB, 0, ST>>>(n, rhs, Au, r, d, m->dot_part); LAUNCH(1);
  if ((rc = finish_partials(m, n, sc + S_DELTA, ST)) != D4B_OK) return rc;
  if ((rc = read_scalars(m, S_DELTA, 1, &delta_new, ST)) != D4B_OK) return rc;
  delta_0 = delta_new;
  for (i = 0; i < imax && (delta_new > atol * atol + delta_0 * rtol * rtol); i++) {
    if ((rc = d4b_apply_dev(m, d, Au, stream)) != D4B_OK) return rc;
    const double *xs[1] = {d}, *ys[1] = {Au};
    if ((rc = dots_dev(m, n, 1, xs, ys, sc + S_DAU, ST)) != D4B_OK) return rc;
    k_cg_step<<<vgrid(n), VB, 0, ST>>>(n, sc, ring, d, Au, u, r, m->dot_part); LAUNCH(1);
    if ((rc = finish_partials(m, n, sc + S_DELTA + (ring ^ 1), ST)) != D4B_OK) return rc;
    k_cg_dir<<<vgrid(n), VB, 0, ST>>>(n, sc, ring, r, d, i, 0, 0); LAUNCH(1);
    ring ^= 1;
    if ((rc = read_scalars(m, S_DELTA + ring, 1, &delta_new, ST)) != D4B_OK) return rc;
  }
  if (iters) *iters = i;
  if (res2) *res2 = delta_new;
  return D4B_OK;
}

/* Flexible CG with a preconditioner callback on DEVICE vectors: restates d4est_solver_fcg_solve
 * (Solver/d4est_solver_fcg_improved.c:97-346); pc(ctx, r, z) plays d4est_krylov_pc_t.pc_apply
 * (Solver/d4est_krylov_pc.h:27-50); pc == NULL copies r (do_not_use_preconditioner).
 * The 2-4 dots of one iteration travel in ONE reduction (:245) and come to the host once per iteration for the
 * stopping test.  work: 6 n doubles.
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
  int rc = ensure_scal(m); if (rc) return rc;
  const int64_t n = m->local_nodes;
  double *Au = work, *r = work + n, *w = work + 2 * n, *d = work + 3 * n, *v = work + 4 * n, *q = work + 5 * n;
  double r0, g[4] = {0, 0, 0, 0}, rho = 0.;
  int k;
  if ((rc = d4b_apply_dev(m, u, Au, stream)) != D4B_OK) return rc;
  k_cg_init<<<vgrid(n), VB, 0, ST>>>(n, rhs, Au, r, nullptr, m->dot_part); LAUNCH(1);
  if ((rc = finish_partials(m, n, m->sc + S_DOT, ST)) != D4B_OK) return rc;
  if ((rc = read_scalars(m, S_DOT, 1, &r0, ST)) != D4B_OK) return rc;
  const double tol = atol + rtol * sqrt(r0);
  double last = sqrt(r0);
  for (k = 0; k < imax; k++) {
    if (pc) { if ((rc = pc(pc_ctx, r, v, stream)) != D4B_OK) return rc; }
    else if ((rc = d4b_vec_copy(n, r, v, stream)) != D4B_OK) return rc;
    if ((rc = d4b_apply_dev(m, v, w, stream)) != D4B_OK) return rc;
    const double *xs[4] = {v, v, v, r}, *ys[4] = {r, w, q, r};
    const int nd = k > 0 ? 4 : 2;
    if ((rc = dots_dev(m, n, nd, xs, ys, m->sc + S_DOT, ST)) != D4B_OK) return rc;
    if ((rc = read_scalars(m, S_DOT, nd, g, ST)) != D4B_OK) return rc;
    const double alpha = g[0], beta = g[1];
    if (k > 0) {
      const double gamma = g[2];
      k_fcg_dir<<<vgrid(n), VB, 0, ST>>>(n, g_fcg_textbook, g_fcg_textbook ? -gamma / rho : 1. - gamma / rho, v, w, d, q); LAUNCH(1);
      rho = beta - (gamma * gamma) / rho;
    } else {
      rho = beta;
      if ((rc = d4b_vec_copy(n, v, d, stream)) != D4B_OK) return rc;
      if ((rc = d4b_vec_copy(n, w, q, stream)) != D4B_OK) return rc;
    }
    k_fcg_step<<<vgrid(n), VB, 0, ST>>>(n, alpha / rho, d, q, u, r); LAUNCH(1);
    if (k > 0) last = sqrt(g[3]);
    if (k > 0 && sqrt(g[3]) <= tol) break;
  }
  if (iters) *iters = k;
  if (final_res) *final_res = last;
  return D4B_OK;
}

} /* extern C */
