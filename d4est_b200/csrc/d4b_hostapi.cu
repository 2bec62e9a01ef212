/* d4b_hostapi.cu -- entry points that take HOST vectors at the granularity d4est's solver stack calls them:
 * one residual, one smoother call (cg_eigs + Chebyshev sweep), one bottom solve, one V-cycle preconditioner
 * application.  Each stages its vectors once (H2D before, D2H after) around MANY operator applies, which is
 * where the host boundary belongs: a smoother call is 15 + 15 + 1 applies for two uploads and two downloads.
 * Used by the d4est-facing glue (d4est_b200/glue/d4est_b200_solver_glue.c). */
#include "d4b_internal.h"
#include <cstdlib>

static int ensure_host_work(d4b_mesh *m)
{
  if (m->host_work) return D4B_OK;
  D4B_CUDA(cudaMalloc((void **)&m->host_work, (size_t)(6 * (m->local_nodes + 2)) * sizeof(double)));
  return D4B_OK;
}

extern "C" {

/* r = rhs - A u on device vectors: the body of every build_residual callback of the Laplacian problems
 * (e.g. Problems/Poisson/poisson_sinx_fcns.h:133-148: apply_lhs, then Au = rhs - Au) */
int d4b_residual_dev(d4b_mesh_t *m, const double *u, const double *rhs, double *r, void *stream)
{
  int rc = d4b_apply_dev(m, u, r, stream);
  if (rc != D4B_OK) return rc;
  return d4b_vec_axpyeqz(m->local_nodes, -1., r, rhs, r, stream);
}

int d4b_residual_host(d4b_mesh_t *m, const double *u_host, const double *rhs_host, double *r_host)
{
  int rc = ensure_host_work(m); if (rc) return rc;
  const int64_t n = m->local_nodes; const size_t nb = (size_t)n * sizeof(double);
  double *u = m->host_work, *rhs = u + n, *r = rhs + n;
  D4B_CUDA(cudaMemcpyAsync(u, u_host, nb, cudaMemcpyHostToDevice, m->stream));
  D4B_CUDA(cudaMemcpyAsync(rhs, rhs_host, nb, cudaMemcpyHostToDevice, m->stream));
  if ((rc = d4b_residual_dev(m, u, rhs, r, m->stream)) != D4B_OK) return rc;
  D4B_CUDA(cudaMemcpyAsync(r_host, r, nb, cudaMemcpyDeviceToHost, m->stream));
  D4B_CUDA(cudaStreamSynchronize(m->stream));
  return D4B_OK;
}

/* One call of the reference's Chebyshev smoother on host vectors (d4est_solver_multigrid_smoother_cheby,
 * Solver/d4est_solver_multigrid_smoother_cheby.c:264-360): optional cg_eigs (which also advances u, cg_eigs.c:230),
 * *eig *= max_multiplier, then cheby_imax sweeps with [eig / ratio, eig]; on return u is smoothed and
 * r = rhs - A u.  The bottom solver (bottom_solver_cheby.c:59-113) is the same call with compute_eig = 1. */
int d4b_cheby_smooth_host(d4b_mesh_t *m, double *u_host, const double *rhs_host, double *r_host, int cheby_imax,
                          int eigs_cg_imax, double lmax_lmin_ratio, double max_multiplier, int use_new_cg_eigs,
                          int compute_eig, double *eig)
{
  int rc = ensure_host_work(m); if (rc) return rc;
  const int64_t n = m->local_nodes; const size_t nb = (size_t)n * sizeof(double);
  double *u = m->host_work, *rhs = u + n, *r = rhs + n, *work = r + n;
  D4B_CUDA(cudaMemcpyAsync(u, u_host, nb, cudaMemcpyHostToDevice, m->stream));
  D4B_CUDA(cudaMemcpyAsync(rhs, rhs_host, nb, cudaMemcpyHostToDevice, m->stream));
  if (compute_eig) {
    if ((rc = d4b_cg_eigs(m, u, rhs, eigs_cg_imax, use_new_cg_eigs, eig, work, m->stream)) != D4B_OK) return rc;
    *eig *= max_multiplier;
  }
  if ((rc = d4b_cheby_iterate(m, u, rhs, r, cheby_imax, *eig / lmax_lmin_ratio, *eig, 1, work, m->stream)) != D4B_OK) return rc;
  D4B_CUDA(cudaMemcpyAsync(u_host, u, nb, cudaMemcpyDeviceToHost, m->stream));
  D4B_CUDA(cudaMemcpyAsync(r_host, r, nb, cudaMemcpyDeviceToHost, m->stream));
  D4B_CUDA(cudaStreamSynchronize(m->stream));
  return D4B_OK;
}


/* ---- mass-type terms of apply_jac / build_residual: Au += M_c v with
 *      M_c = I^T diag(w J f(x, u0)) I, the operator d4est_quadrature_apply_fofufofvlilj applies element by element
 *      (Quadrature/d4est_quadrature.c:593; caller e.g. Problems/ConstantDensityStar/constant_density_star_fcns.h:484-605).
 *      The coefficient c = J f(x, u0) at the volume quadrature points changes only with the linearisation point: it
 *      is set once (host array of quad_off[n_local] doubles in element order) and stays resident. */
int d4b_mass_set_coef(d4b_mesh_t *m, const double *coef_quad_host)
{
  if (!coef_quad_host) { cudaFree(m->mass_coef); m->mass_coef = nullptr; return D4B_OK; }
  if (!m->mass_coef) D4B_CUDA(cudaMalloc((void **)&m->mass_coef, (size_t)(m->total_quad ? m->total_quad : 1) * sizeof(double)));
  D4B_CUDA(cudaMemcpy(m->mass_coef, coef_quad_host, (size_t)m->total_quad * sizeof(double), cudaMemcpyHostToDevice));
  return D4B_OK;
}
int d4b_mass_apply_dev(d4b_mesh_t *m, const double *v_dev, double *out_dev, int accumulate, void *stream)
{
  if (!m->mass_coef) { d4b_set_error("d4b_mass_apply: no coefficient set (d4b_mass_set_coef)"); return D4B_EINVAL; }
  return d4b_launch_mass(m, m->mass_coef, v_dev, out_dev, accumulate, (cudaStream_t)stream);
}
/* Au = A u + M_c u: the Jacobian apply of a semilinear problem (Laplacian + f'(u0) mass term) in one call */
int d4b_apply_jac_dev(d4b_mesh_t *m, const double *u_dev, double *Au_dev, void *stream)
{
  int rc = d4b_apply_dev(m, u_dev, Au_dev, stream);
  if (rc != D4B_OK) return rc;
  return d4b_mass_apply_dev(m, u_dev, Au_dev, 1, stream);
}
int d4b_mass_apply_host(d4b_mesh_t *m, const double *v_host, double *out_host, int accumulate)
{
  int rc = ensure_host_work(m); if (rc) return rc;
  const int64_t n = m->local_nodes; const size_t nb = (size_t)n * sizeof(double);
  double *v = m->host_work, *o = v + n;
  D4B_CUDA(cudaMemcpyAsync(v, v_host, nb, cudaMemcpyHostToDevice, m->stream));
  if (accumulate) D4B_CUDA(cudaMemcpyAsync(o, out_host, nb, cudaMemcpyHostToDevice, m->stream));
  if ((rc = d4b_mass_apply_dev(m, v, o, accumulate, m->stream)) != D4B_OK) return rc;
  D4B_CUDA(cudaMemcpyAsync(out_host, o, nb, cudaMemcpyDeviceToHost, m->stream));
  D4B_CUDA(cudaStreamSynchronize(m->stream));
  return D4B_OK;
}

} /* extern C */
