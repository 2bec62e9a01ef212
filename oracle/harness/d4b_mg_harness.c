/* d4b_mg_harness.c -- TEST INFRASTRUCTURE ONLY (oracle side).  Links the UNMODIFIED reference solver stack
 * (Solver/d4est_solver_multigrid.c, d4est_solver_fcg_improved.c, d4est_krylov_pc_multigrid.c, the Chebyshev
 * smoother / bottom solver, the element-data updater) and
 *   1. builds a curved hp mesh the way the reference's unit tests do,
 *   2. solves A u = rhs with the reference's own multigrid-preconditioned flexible CG on the CPU, recording the
 *      iteration count, the residual norm of every iteration and the first preconditioner application (r0 -> z0),
 *   3. captures the reference's multigrid hierarchy through the product glue (d4est_b200_mg_capture_tables) and
 *      writes every level's flat tables, the transfer tables, the solver parameters and the reference results
 *      to a fixture (read by tests/test_mg_reference_gpu.py),
 *   4. with --gpu-check (needs a CUDA device): runs the same things through the d4est-facing entry points of the
 *      product glue IN THIS PROCESS, next to the reference, and prints the differences as one JSON line.
 *
 * usage: d4b_mg_harness <input> <amr: none|random_hp|uniform_h> <steps> <out.fix|-> [--gpu-check [ref_iterations] | --capture-only]
 *   With --gpu-check the minutes-long reference solve is not repeated: the reference side of the comparison is
 *   one reference V-cycle (z0 = PC(rhs), the first preconditioner application of the solve) and the reference
 *   operator applies, all run in this process; the reference iteration count comes from the fixture made by
 *   the run without --gpu-check (passed as ref_iterations).
 */
#include <pXest.h>
#include <d4est_quadrature.h>
#include <d4est_element_data.h>
#include <d4est_geometry.h>
#include <d4est_mesh.h>
#include <d4est_linalg.h>
#include <d4est_amr.h>
#include <d4est_laplacian.h>
#include <d4est_laplacian_flux.h>
#include <d4est_laplacian_with_opt.h>
#include <d4est_util.h>
#include <d4est_solver_multigrid.h>
#include <d4est_solver_fcg_improved.h>
#include <d4est_krylov_pc_multigrid.h>
#include <d4est_solver_cg_eigs.h>
#include <d4est_solver_cg.h>
#include <time.h>
#include <math.h>
#include "d4est_b200_glue.h"
#include "fixture_io.h"

static double sin_fcn(double x, double y, double z, void *user)
{ (void)user; return sin(M_PI * x) * sin(M_PI * y) * sin(M_PI * z); }
static double poly_fcn(double x, double y, double z, void *user) { (void)user; return x * x + y * y + z * z; }

static void ref_apply_lhs(p4est_t *p4est, d4est_ghost_t *ghost, d4est_ghost_data_t *gd, d4est_elliptic_data_t *v,
                          d4est_operators_t *ops, d4est_geometry_t *geom, d4est_quadrature_t *quad,
                          d4est_mesh_data_t *factors, void *user)
{ d4est_laplacian_with_opt_apply_aij(p4est, ghost, gd, v, (d4est_laplacian_with_opt_flux_data_t *)user, ops, geom, quad, factors, 0); }

static double test_fofu(double x, double y, double z, double u, void *ctx)
{ (void)ctx; return 1. + 0.5 * x - 0.25 * y * z + u * u / (1. + u * u); }

static double now_s(void) { struct timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return t.tv_sec + 1e-9 * t.tv_nsec; }
static double nrm2(const double *a, int n) { double s = 0.; for (int i = 0; i < n; i++) s += a[i] * a[i]; return sqrt(s); }
static double rel_l2(const double *a, const double *b, int n)
{ double s = 0., t = 0.; for (int i = 0; i < n; i++) { s += (a[i] - b[i]) * (a[i] - b[i]); t += b[i] * b[i]; } return sqrt(s / (t > 0 ? t : 1)); }

static d4est_mesh_local_sizes_t update_mesh(p4est_t *p4est, d4est_ghost_t **ghost, d4est_operators_t *ops,
                                            d4est_geometry_t *geom, d4est_quadrature_t *quad,
                                            d4est_mesh_data_t *factors, d4est_mesh_initial_extents_t *ext, int initial)
{
  return d4est_mesh_update(p4est, ghost, ops, geom, quad, factors, ext, INITIALIZE_GHOST, INITIALIZE_QUADRATURE_DATA,
                           INITIALIZE_GEOMETRY_DATA, INITIALIZE_GEOMETRY_ALIASES,
                           initial ? d4est_mesh_set_initial_extents : d4est_mesh_set_quadratures_after_amr, (void *)ext);
}

/* preconditioner wrapper: counts applications, keeps |r_k| and the first (r0, z0) pair */
#define MAXIT 200
typedef struct { d4est_krylov_pc_t *inner; int n, count; double hist[MAXIT]; double *r0, *z0; } rec_pc_t;
static void rec_pc_apply(d4est_krylov_pc_t *kpc, double *x, double *y)
{
  rec_pc_t *R = (rec_pc_t *)kpc->pc_data;
  R->inner->pc_ctx = kpc->pc_ctx;
  if (R->count < MAXIT) R->hist[R->count] = nrm2(x, R->n);
  R->inner->pc_apply(R->inner, x, y);
  if (R->count == 0 && R->r0) { memcpy(R->r0, x, (size_t)R->n * 8); memcpy(R->z0, y, (size_t)R->n * 8); }
  R->count++;
}

int main(int argc, char **argv)
{
  if (argc < 5) { fprintf(stderr, "usage: see header of d4b_mg_harness.c\n"); return 2; }
  const char *input = argv[1], *amr = argv[2]; int steps = atoi(argv[3]); const char *out = argv[4];
  int gpu_check = argc > 5 && !strcmp(argv[5], "--gpu-check");
  int capture_only = argc > 5 && !strcmp(argv[5], "--capture-only");   /* hierarchy + rhs only: no reference solve */
  int ref_iters_arg = argc > 6 ? atoi(argv[6]) : -1;

  sc_MPI_Comm comm = sc_MPI_COMM_WORLD;
  sc_MPI_Init(&argc, &argv);
  p4est_init(NULL, SC_LP_ERROR);
  d4est_geometry_t *geom = d4est_geometry_new(0, input, "geometry", NULL);
  d4est_mesh_initial_extents_t *ext = d4est_mesh_initial_extents_parse(input, geom);
  p4est_t *p4est = p4est_new_ext(comm, geom->p4est_conn, ext->min_quadrants, ext->min_level, ext->fill_uniform,
                                 sizeof(d4est_element_data_t), NULL, NULL);
  p4est_partition(p4est, 1, NULL);
  p4est_balance(p4est, P4EST_CONNECT_FULL, NULL);
  d4est_ghost_t *ghost = d4est_ghost_init(p4est);
  d4est_operators_t *ops = d4est_ops_init(20);
  d4est_mesh_data_t *factors = d4est_mesh_data_init(p4est);
  d4est_quadrature_t *quad = d4est_quadrature_new(p4est, ops, geom, input, "quadrature");
  d4est_mesh_local_sizes_t ls = update_mesh(p4est, &ghost, ops, geom, quad, factors, ext, 1);
  ext->initial_nodes = ls.local_nodes;
  if (strcmp(amr, "none") != 0 && steps > 0) {
    d4est_amr_t *a = !strcmp(amr, "random_hp") ? d4est_amr_init_random_hp(p4est, steps) : d4est_amr_init_uniform_h(p4est, steps);
    for (int i = 0; i < steps; i++) {
      d4est_amr_step(p4est, ops, a, NULL, NULL, NULL, NULL);
      ls = update_mesh(p4est, &ghost, ops, geom, quad, factors, ext, 0);
    }
    d4est_amr_destroy(a);
  }
  const int n = ls.local_nodes;
  d4est_field_type_t ft = NODAL;
  d4est_ghost_data_t *gd = d4est_ghost_data_init(p4est, ghost, &ft, 1);

  /* operator of the solve: homogeneous Dirichlet data (what every Jacobian / preconditioner apply of the
   * reference's drivers uses) */
  d4est_laplacian_dirichlet_bc_t bcd; bcd.dirichlet_fcn = zero_fcn; bcd.eval_method = EVAL_BNDRY_FCN_ON_LOBATTO; bcd.user = NULL;
  d4est_laplacian_with_opt_flux_data_t *fl1 = d4est_laplacian_with_opt_flux_new(p4est, input, BC_DIRICHLET, &bcd);
  d4est_laplacian_flux_data_t *fl0 = d4est_laplacian_flux_new(p4est, input, BC_DIRICHLET, &bcd);
  d4est_elliptic_eqns_t fcns; fcns.apply_lhs = ref_apply_lhs; fcns.build_residual = NULL; fcns.user = fl1;

  /* manufactured problem: rhs = A u*, u* = sin sin sin + rand (speedup.c:429-432 convention) */
  double *ustar = P4EST_ALLOC(double, n), *rhs = P4EST_ALLOC_ZERO(double, n), *u = P4EST_ALLOC_ZERO(double, n),
         *Au = P4EST_ALLOC_ZERO(double, n);
  d4est_mesh_init_field(p4est, ustar, sin_fcn, ops, geom, factors, INIT_FIELD_ON_LOBATTO, NULL);
  srand(102321);
  for (int i = 0; i < n; i++) ustar[i] += (double)rand() / (double)RAND_MAX;
  d4est_elliptic_data_t vecs; memset(&vecs, 0, sizeof vecs);
  vecs.local_nodes = n; vecs.field_types = &ft; vecs.num_of_fields = 1;
  vecs.u = ustar; vecs.Au = rhs;
  double t_apply = now_s();
  d4est_laplacian_with_opt_apply_aij(p4est, ghost, gd, &vecs, fl1, ops, geom, quad, factors, 0);
  t_apply = now_s() - t_apply;
  vecs.u = u; vecs.Au = Au; vecs.rhs = rhs;

  d4est_solver_multigrid_t *mg = d4est_solver_multigrid_data_init(p4est, ops, geom, quad, &ghost, &gd, factors, ext, input);
  d4est_krylov_pc_t *pc_ref = d4est_krylov_pc_multigrid_create(mg, NULL);
  d4est_solver_fcg_params_t fp; d4est_solver_fcg_input(p4est, input, "fcg", &fp);

  /* ---- the reference solve (CPU) */
  rec_pc_t R; memset(&R, 0, sizeof R);
  R.inner = pc_ref; R.n = n; R.r0 = P4EST_ALLOC(double, n); R.z0 = P4EST_ALLOC(double, n);
  d4est_krylov_pc_t pc_rec; pc_rec.pc_apply = rec_pc_apply; pc_rec.pc_setup = NULL; pc_rec.pc_data = &R; pc_rec.pc_ctx = NULL;
  double t0 = now_s(), t_ref = 0., res_ref = -1.;
  int it_ref = ref_iters_arg;
  if (capture_only) {
    memcpy(R.r0, rhs, (size_t)n * 8); memset(R.z0, 0, (size_t)n * 8);
  } else if (!gpu_check) {
    d4est_solver_fcg_solve(p4est, &vecs, &fcns, &ghost, &gd, ops, geom, quad, factors, &fp, &pc_rec, -1, -1);
    t_ref = now_s() - t0;
    /* iterations as d4b_fcg_solve counts them: the loop index at the break (fcg_improved.c:283) */
    it_ref = R.count - 1 + (R.count == fp.imax ? 1 : 0);
    vecs.u = u; vecs.Au = Au;
    d4est_laplacian_with_opt_apply_aij(p4est, ghost, gd, &vecs, fl1, ops, geom, quad, factors, 0);
    double s = 0.; for (int i = 0; i < n; i++) s += (rhs[i] - Au[i]) * (rhs[i] - Au[i]); res_ref = sqrt(s);
  } else {   /* one reference V-cycle: the first preconditioner application of that solve (u0 = 0 => r0 = rhs) */
    krylov_ctx_t k0; memset(&k0, 0, sizeof k0);
    k0.p4est = p4est; k0.vecs = &vecs; k0.fcns = &fcns; k0.ghost = &ghost; k0.ghost_data = &gd;
    k0.d4est_ops = ops; k0.d4est_geom = geom; k0.d4est_quad = quad; k0.d4est_factors = factors;
    pc_rec.pc_ctx = &k0;
    rec_pc_apply(&pc_rec, rhs, Au);
    t_ref = now_s() - t0;
  }

  /* ---- the hierarchy, through the product glue (host only: no device needed) */
  d4b_mesh_desc_t **desc; d4est_b200_mg_transfer_table_t *tt;
  d4est_b200_mg_capture_tables(p4est, mg, fl1, &vecs, &desc, &tt);
  d4b_mg_params_t par; d4est_b200_mg_params_from_d4est(mg, &par);
  const int nl = mg->num_of_levels;

  if (strcmp(out, "-") != 0) {
    FILE *f = fix_open(out);
    int32_t meta[4] = { nl, it_ref, R.count, fp.imax };
    fix_write(f, "mg_meta", FIX_I32, 4, meta);
    double pd[20] = { par.vcycle_imax, par.vcycle_rtol, par.vcycle_atol, par.cheby_imax, par.cheby_eigs_cg_imax,
                      par.cheby_eigs_lmax_lmin_ratio, par.cheby_eigs_max_multiplier, par.eigs_reuse_fromdownvcycle,
                      par.eigs_reuse_fromlastvcycle, par.use_new_cg_eigs, par.bottom_kind, par.bottom_imax,
                      par.bottom_eigs_cg_imax, par.bottom_eigs_lmax_lmin_ratio, par.bottom_eigs_max_multiplier,
                      par.bottom_atol, par.bottom_rtol, fp.atol, fp.rtol, t_ref };
    fix_write(f, "mg_params", FIX_F64, 20, pd);
    fix_write(f, "mg_rhs", FIX_F64, n, rhs);
    fix_write(f, "mg_u_ref", FIX_F64, n, u);
    fix_write(f, "mg_ustar", FIX_F64, n, ustar);
    fix_write(f, "mg_res_hist", FIX_F64, R.count < MAXIT ? R.count : MAXIT, R.hist);
    fix_write(f, "mg_r0", FIX_F64, n, R.r0);
    fix_write(f, "mg_z0", FIX_F64, n, R.z0);
    if (!capture_only) { /* the reference's Krylov drivers WITHOUT a preconditioner, a fixed 25 iterations from u = 0 on the same
         system: pins the restated drivers (oracle/solver_oracle.py, d4b_fcg_solve, d4b_cg_solve) step by step */
      double *uk = P4EST_ALLOC_ZERO(double, n), *Ak = P4EST_ALLOC_ZERO(double, n);
      d4est_elliptic_data_t vk; d4est_elliptic_data_copy_ptrs(&vecs, &vk); vk.u = uk; vk.Au = Ak; vk.rhs = rhs;
      d4est_solver_fcg_params_t f25 = fp; f25.imax = 25; f25.rtol = 1e-300; f25.atol = 0.; f25.do_not_use_preconditioner = 1;
      d4est_solver_fcg_solve(p4est, &vk, &fcns, &ghost, &gd, ops, geom, quad, factors, &f25, NULL, -1, -1);
      fix_write(f, "krylov_fcg25_u", FIX_F64, n, uk);
      memset(uk, 0, (size_t)n * 8);
      d4est_solver_cg_params_t c25; memset(&c25, 0, sizeof c25); c25.iter = 25; c25.rtol = 1e-300; c25.atol = 0.; c25.monitor = 0;
      d4est_solver_cg_solve(p4est, &vk, &fcns, &ghost, &gd, ops, geom, quad, factors, &c25, NULL);
      fix_write(f, "krylov_cg25_u", FIX_F64, n, uk);
      P4EST_FREE(uk); P4EST_FREE(Ak);
    }
    for (int l = 0; l < nl; l++) {
      char pre[16]; snprintf(pre, sizeof pre, "L%d.", l); fix_prefix = pre;
      fix_write_desc(f, desc[l]);
      if (l + 1 < nl) {
        snprintf(pre, sizeof pre, "T%d.", l);
        fix_write_p(f, "kind", FIX_I32, tt[l].n_coarse, tt[l].kind);
        fix_write_p(f, "degH", FIX_I32, tt[l].n_coarse, tt[l].degH);
        fix_write_p(f, "coarse_off", FIX_I64, tt[l].n_coarse, tt[l].coarse_off);
        fix_write_p(f, "child_deg", FIX_I32, 8 * (int64_t)tt[l].n_coarse, tt[l].child_deg);
        fix_write_p(f, "child_off", FIX_I64, 8 * (int64_t)tt[l].n_coarse, tt[l].child_off);
        /* the reference's own restriction / prolongation of probe vectors over this table
         * (d4est_operators_apply_{p,hp}_prolong[_transpose], as its callbacks call them,
         * Solver/d4est_solver_multigrid_callbacks.h:139,162,266,287) */
        int64_t nf = tt[l].fine_nodes, nc = tt[l].coarse_nodes;
        double *pf = P4EST_ALLOC(double, nf), *pc = P4EST_ALLOC(double, nc), *rf = P4EST_ALLOC_ZERO(double, nc),
               *pr = P4EST_ALLOC_ZERO(double, nf);
        for (int64_t i = 0; i < nf; i++) pf[i] = sin(0.37 * (double)i + 0.1 * l) + 0.01 * (double)(i % 7);
        for (int64_t i = 0; i < nc; i++) pc[i] = cos(0.53 * (double)i + 0.2 * l) - 0.02 * (double)(i % 5);
        for (int e = 0; e < tt[l].n_coarse; e++) {
          int dH = tt[l].degH[e]; int *dh = &tt[l].child_deg[8 * e];
          double *xH_r = rf + tt[l].coarse_off[e], *xH_p = pc + tt[l].coarse_off[e];
          int64_t fo = tt[l].child_off[8 * e];
          if (tt[l].kind[e] == 1) {
            d4est_operators_apply_hp_prolong_transpose(ops, pf + fo, dh, 3, dH, xH_r);
            d4est_operators_apply_hp_prolong(ops, xH_p, dH, 3, dh, pr + fo);
          } else {
            d4est_operators_apply_p_prolong_transpose(ops, pf + fo, dh[0], 3, dH, xH_r);
            d4est_operators_apply_p_prolong(ops, xH_p, dH, 3, dh[0], pr + fo);
          }
        }
        fix_write_p(f, "probe_fine", FIX_F64, nf, pf); fix_write_p(f, "ref_restricted", FIX_F64, nc, rf);
        fix_write_p(f, "probe_coarse", FIX_F64, nc, pc); fix_write_p(f, "ref_prolonged", FIX_F64, nf, pr);
        P4EST_FREE(pf); P4EST_FREE(pc); P4EST_FREE(rf); P4EST_FREE(pr);
      }
      fix_prefix = "";
    }
    fclose(f);
  }
  printf("{\"what\": \"reference_mg_fcg\", \"elements\": %d, \"nodes\": %d, \"levels\": %d, \"fcg_iterations\": %d, "
         "\"pc_applications\": %d, \"true_residual\": %.6e, \"rhs_norm\": %.6e, \"seconds_cpu\": %.3f, "
         "\"err_vs_ustar\": %.3e, \"seconds_reference_apply_top_level\": %.4f",
         (int)p4est->local_num_quadrants, n, nl, it_ref, R.count, res_ref, nrm2(rhs, n), t_ref, rel_l2(u, ustar, n), t_apply);
  for (int l = 0; l < nl; l++) printf(", \"L%d_elements\": %d, \"L%d_nodes\": %lld", l, desc[l]->n_local, l, (long long)desc[l]->node_off[desc[l]->n_local]);
  printf("}\n");
  fflush(stdout);

  if (gpu_check) {
    /* (a) the operator callbacks: d4est_apply_operator_fcn_t, the baseline twin, build_residual */
    double *y0 = P4EST_ALLOC_ZERO(double, n), *y1 = P4EST_ALLOC_ZERO(double, n), *y2 = P4EST_ALLOC_ZERO(double, n);
    d4est_elliptic_data_t v; d4est_elliptic_data_copy_ptrs(&vecs, &v);
    v.u = ustar; v.Au = y0;
    d4est_laplacian_with_opt_apply_aij(p4est, ghost, gd, &v, fl1, ops, geom, quad, factors, 0);
    v.Au = y1; d4est_b200_apply_lhs(p4est, ghost, gd, &v, ops, geom, quad, factors, fl1);
    double e_apply = rel_l2(y1, y0, n);
    v.Au = y2; d4est_laplacian_apply_aij(p4est, ghost, gd, &v, fl0, ops, geom, quad, factors, 0);
    d4est_b200_level_invalidate(factors);   /* force the baseline-twin flatten */
    v.Au = y1; d4est_b200_laplacian_baseline_apply_aij(p4est, ghost, gd, &v, fl0, ops, geom, quad, factors, 0);
    double e_twin = rel_l2(y1, y2, n);
    d4est_b200_level_invalidate(factors);
    v.rhs = rhs; v.Au = y1; d4est_b200_build_residual(p4est, ghost, gd, &v, ops, geom, quad, factors, fl1);
    for (int i = 0; i < n; i++) y2[i] = rhs[i] - y0[i];
    double e_resid = 0.; { double s = 0.; for (int i = 0; i < n; i++) s += (y1[i] - y2[i]) * (y1[i] - y2[i]); e_resid = sqrt(s) / nrm2(y0, n); }

    /* (a') the mass-type term: the reference's per-element d4est_quadrature_apply_fofufofvlilj loop against the glue
     *      (coefficient tabulated with the same callback, one device call) */
    double e_mass;
    {
      double *mref = P4EST_ALLOC_ZERO(double, n), *mdev = P4EST_ALLOC_ZERO(double, n);
      for (p4est_topidx_t tt = p4est->first_local_tree; tt <= p4est->last_local_tree; ++tt) {
        sc_array_t *tq = &p4est_tree_array_index(p4est->trees, tt)->quadrants;
        for (size_t q = 0; q < tq->elem_count; ++q) {
          d4est_element_data_t *ed = (d4est_element_data_t *)p4est_quadrant_array_index(tq, q)->p.user_data;
          d4est_quadrature_volume_t mo;
          mo.dq = ed->dq; mo.tree = ed->tree; mo.q[0] = ed->q[0]; mo.q[1] = ed->q[1]; mo.q[2] = ed->q[2]; mo.element_id = ed->id;
          d4est_mesh_data_on_element_t md = d4est_mesh_data_on_element(factors, ed);
          d4est_quadrature_apply_fofufofvlilj(ops, geom, quad, &mo, QUAD_OBJECT_VOLUME, QUAD_INTEGRAND_UNKNOWN,
                                              &ustar[ed->nodal_stride], &ustar[ed->nodal_stride], NULL, ed->deg, md.xyz_quad,
                                              d4est_mesh_get_jacobian_on_quadrature_points(factors, ed), ed->deg_quad,
                                              &mref[ed->nodal_stride], test_fofu, NULL, NULL, NULL, QUAD_APPLY_MATRIX, 0, NULL);
        }
      }
      d4est_b200_mass_set_fofu(p4est, ghost, fl1, ops, geom, quad, factors, ustar, test_fofu, NULL);
      d4est_elliptic_data_t vm; d4est_elliptic_data_copy_ptrs(&vecs, &vm); vm.u = ustar; vm.Au = mdev;
      d4est_b200_mass_apply_add(p4est, ghost, fl1, &vm, ops, geom, quad, factors);
      e_mass = rel_l2(mdev, mref, n);
      P4EST_FREE(mref); P4EST_FREE(mdev);
    }

    /* (b) the resident multigrid behind d4est_krylov_pc_t and d4est_solver_fcg_solve's argument list */
    d4est_b200_mg_t *G = d4est_b200_mg_capture(p4est, mg, fl1, &vecs);
    d4est_krylov_pc_t *pc_dev = d4est_b200_krylov_pc_multigrid_create(G);
    pc_dev->pc_apply(pc_dev, R.r0, y1);
    double e_pc = rel_l2(y1, R.z0, n);
    memset(u, 0, (size_t)n * 8); vecs.u = u; vecs.Au = Au; vecs.rhs = rhs;
    t0 = now_s();
    int it_dev = d4est_b200_fcg_solve(p4est, &vecs, &fcns, &ghost, &gd, ops, geom, quad, factors, &fp, pc_dev, -1, -1);
    double t_dev = now_s() - t0;
    v.u = u; v.Au = y1; d4est_laplacian_with_opt_apply_aij(p4est, ghost, gd, &v, fl1, ops, geom, quad, factors, 0);
    double res_dev; { double s = 0.; for (int i = 0; i < n; i++) s += (rhs[i] - y1[i]) * (rhs[i] - y1[i]); res_dev = sqrt(s); }

    /* (c) d4est's own V-cycle with the smoother / bottom solver bodies on the device */
    d4est_b200_multigrid_use_device(mg, fl1);
    pc_ref->pc_ctx = NULL;
    krylov_ctx_t kctx; memset(&kctx, 0, sizeof kctx);
    kctx.p4est = p4est; kctx.vecs = &vecs; kctx.fcns = &fcns; kctx.ghost = &ghost; kctx.ghost_data = &gd;
    kctx.d4est_ops = ops; kctx.d4est_geom = geom; kctx.d4est_quad = quad; kctx.d4est_factors = factors;
    pc_ref->pc_ctx = &kctx;
    t0 = now_s();
    pc_ref->pc_apply(pc_ref, R.r0, y2);
    double t_hyb = now_s() - t0;
    double e_hyb = rel_l2(y2, R.z0, n);

    printf("{\"what\": \"gpu_check\", \"apply_lhs_rel_l2\": %.3e, \"baseline_twin_rel_l2\": %.3e, \"build_residual_rel_l2\": %.3e, \"mass_term_rel_l2\": %.3e, "
           "\"pc_apply_resident_rel_l2\": %.3e, \"pc_apply_d4est_vcycle_device_smoother_rel_l2\": %.3e, "
           "\"fcg_iterations_reference\": %d, \"fcg_iterations_device\": %d, \"true_residual_reference\": %.6e, "
           "\"true_residual_device\": %.6e, \"seconds_reference_solve_cpu\": %.3f, \"seconds_device_solve\": %.4f, "
           "\"seconds_d4est_vcycle_device_smoother\": %.4f, \"kernel_launches\": %lld}\n",
           e_apply, e_twin, e_resid, e_mass, e_pc, e_hyb, it_ref, it_dev, res_ref, res_dev, t_ref, t_dev, t_hyb,
           (long long)d4b_kernel_launches());
    d4est_b200_krylov_pc_multigrid_destroy(pc_dev);
    d4est_b200_mg_destroy(G);
  }
  return 0;
}
