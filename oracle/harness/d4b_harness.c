/* d4b_harness.c -- TEST INFRASTRUCTURE ONLY (oracle side, never shipped or timed as
 * the product).  Links the UNMODIFIED reference (oracle/_ref/libd8est_ref.a) and:
 *   1. builds a mesh the way the reference's own unit tests do
 *      (Tests/Unit/d4est_test_laplacian_speedup.c:163-330),
 *   2. runs the reference CPU apply, both variants
 *      (d4est_laplacian_apply_aij dGMath/d4est_laplacian.c:317 and
 *       d4est_laplacian_with_opt_apply_aij dGMath/d4est_laplacian_with_opt.c:289),
 *   3. flattens the mesh through the product glue (d4est_b200_flatten) and writes the
 *      flat tables + u + reference Au to a fixture file (format: oracle/fixture.py),
 *   4. optionally times the reference apply (bench.py cpu_baseline kind "reference").
 *
 * usage: d4b_harness <input> <amr: none|random_hp|uniform_h> <steps> <bc: zero|poly|polyquad|robin>
 *                    <u: polyrand|sinrand|poly> <out.fix|-> [timing_reps]
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
#include <time.h>
#include <math.h>
#include <d4est_solver_cg_eigs.h>
#include <d4est_solver_multigrid_smoother_cheby.h>
#include "d4est_b200_glue.h"
#include "fixture_io.h"

static double poly_fcn(double x, double y, double z, void *user) { (void)user; return x * x + y * y + z * z; }
static double sin_fcn(double x, double y, double z, void *user)
{ (void)user; return sin(M_PI * x) * sin(M_PI * y) * sin(M_PI * z); }

/* Robin data alpha u + du/dn-type condition: sj (alpha u - beta), alpha(x) as in the reference's Lorentzian problem
 * (Problems/Poisson/poisson_lorentzian_fcns.h:22-37), beta(x) nonzero so that the right-hand-side part is exercised */
static double robin_coeff_opt(double x, double y, double z, void *user, d4est_laplacian_with_opt_flux_boundary_data_t *bd, int i)
{ (void)user; (void)bd; (void)i; double r2 = x * x + y * y + z * z; return sqrt(r2) / (1. + r2); }
static double robin_rhs_opt(double x, double y, double z, void *user, d4est_laplacian_with_opt_flux_boundary_data_t *bd, int i)
{ (void)user; (void)bd; (void)i; return 0.3 + 0.1 * x - 0.2 * y * z; }
static double robin_coeff_base(double x, double y, double z, void *user, d4est_laplacian_flux_boundary_data_t *bd, int i)
{ return robin_coeff_opt(x, y, z, user, NULL, i); }
static double robin_rhs_base(double x, double y, double z, void *user, d4est_laplacian_flux_boundary_data_t *bd, int i)
{ return robin_rhs_opt(x, y, z, user, NULL, i); }

/* coefficient of the mass-type term test: a smooth function of position and of the linearisation point */
static double test_fofu(double x, double y, double z, double u, void *ctx)
{ (void)ctx; return 1. + 0.5 * x - 0.25 * y * z + u * u / (1. + u * u); }

static void ref_apply_lhs(p4est_t *p4est, d4est_ghost_t *ghost, d4est_ghost_data_t *gd, d4est_elliptic_data_t *v,
                          d4est_operators_t *ops, d4est_geometry_t *geom, d4est_quadrature_t *quad,
                          d4est_mesh_data_t *factors, void *user)
{ d4est_laplacian_with_opt_apply_aij(p4est, ghost, gd, v, (d4est_laplacian_with_opt_flux_data_t *)user, ops, geom, quad, factors, 0); }

static double now_s(void)
{ struct timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return t.tv_sec + 1e-9 * t.tv_nsec; }

static d4est_mesh_local_sizes_t update_mesh(p4est_t *p4est, d4est_ghost_t **ghost, d4est_operators_t *ops,
                                            d4est_geometry_t *geom, d4est_quadrature_t *quad,
                                            d4est_mesh_data_t *factors, d4est_mesh_initial_extents_t *ext,
                                            int initial)
{
  return d4est_mesh_update(p4est, ghost, ops, geom, quad, factors, ext, INITIALIZE_GHOST,
                           INITIALIZE_QUADRATURE_DATA, INITIALIZE_GEOMETRY_DATA, INITIALIZE_GEOMETRY_ALIASES,
                           initial ? d4est_mesh_set_initial_extents : d4est_mesh_set_quadratures_after_amr,
                           (void *)ext);
}

int main(int argc, char **argv)
{
  if (argc < 7) { fprintf(stderr, "usage: see header of d4b_harness.c\n"); return 2; }
  const char *input = argv[1], *amr = argv[2]; int steps = atoi(argv[3]);
  const char *bc = argv[4], *ukind = argv[5], *out = argv[6];
  int reps = (argc > 7) ? atoi(argv[7]) : 0;

  sc_MPI_Comm comm = sc_MPI_COMM_WORLD;
  sc_MPI_Init(&argc, &argv);
  p4est_init(NULL, SC_LP_ERROR);
  d4est_geometry_t *geom = d4est_geometry_new(0, input, "geometry", NULL);
  d4est_mesh_initial_extents_t *ext = d4est_mesh_initial_extents_parse(input, geom);
  p4est_t *p4est = p4est_new_ext(comm, geom->p4est_conn, ext->min_quadrants, ext->min_level,
                                 ext->fill_uniform, sizeof(d4est_element_data_t), NULL, NULL);
  p4est_partition(p4est, 1, NULL);
  p4est_balance(p4est, P4EST_CONNECT_FULL, NULL);
  d4est_ghost_t *ghost = d4est_ghost_init(p4est);
  d4est_operators_t *ops = d4est_ops_init(20);
  d4est_mesh_data_t *factors = d4est_mesh_data_init(p4est);
  d4est_quadrature_t *quad = d4est_quadrature_new(p4est, ops, geom, input, "quadrature");
  d4est_mesh_local_sizes_t ls = update_mesh(p4est, &ghost, ops, geom, quad, factors, ext, 1);
  ext->initial_nodes = ls.local_nodes;

  if (strcmp(amr, "none") != 0 && steps > 0) {
    d4est_amr_t *a = !strcmp(amr, "random_hp") ? d4est_amr_init_random_hp(p4est, steps)
                   : !strcmp(amr, "uniform_h") ? d4est_amr_init_uniform_h(p4est, steps) : NULL;
    if (!a) { fprintf(stderr, "unknown amr kind %s\n", amr); return 2; }
    for (int i = 0; i < steps; i++) {
      d4est_amr_step(p4est, ops, a, NULL, NULL, NULL, NULL);
      ls = update_mesh(p4est, &ghost, ops, geom, quad, factors, ext, 0);
      if (i + 1 < steps && !strcmp(amr, "random_hp")) {
        /* the reference test perturbs u with rand() between AMR steps
         * (d4est_test_laplacian_speedup.c:429-432), which advances the generator the next
         * random-hp step draws from; replay that so the level-1 golden sum is reproduced */
        srand(102321);
        for (int k = 0; k < ls.local_nodes; k++) (void)rand();
      }
    }
    d4est_amr_destroy(a);
  }
  int n = ls.local_nodes;
  d4est_field_type_t ft = NODAL;
  d4est_ghost_data_t *gd = d4est_ghost_data_init(p4est, ghost, &ft, 1);

  d4est_laplacian_dirichlet_bc_t bcd;
  bcd.dirichlet_fcn = !strcmp(bc, "zero") ? zero_fcn : poly_fcn;
  /* "polyquad": the same polynomial data evaluated directly at the mortar quadrature points (..._flux_sipg.c:125-138) */
  bcd.eval_method = !strcmp(bc, "polyquad") ? EVAL_BNDRY_FCN_ON_QUAD : EVAL_BNDRY_FCN_ON_LOBATTO;
  bcd.user = NULL;
  d4est_laplacian_robin_bc_t rb0; rb0.robin_coeff = robin_coeff_base; rb0.robin_rhs = robin_rhs_base; rb0.user = NULL;
  d4est_laplacian_with_opt_robin_bc_t rb1; rb1.robin_coeff = robin_coeff_opt; rb1.robin_rhs = robin_rhs_opt; rb1.user = NULL;
  const int robin = !strcmp(bc, "robin");
  d4est_laplacian_flux_data_t *fl0 = robin ? d4est_laplacian_flux_new(p4est, input, BC_ROBIN, &rb0)
                                           : d4est_laplacian_flux_new(p4est, input, BC_DIRICHLET, &bcd);
  d4est_laplacian_with_opt_flux_data_t *fl1 = robin ? d4est_laplacian_with_opt_flux_new(p4est, input, BC_ROBIN, &rb1)
                                                    : d4est_laplacian_with_opt_flux_new(p4est, input, BC_DIRICHLET, &bcd);

  double *u = P4EST_ALLOC(double, n), *Au0 = P4EST_ALLOC_ZERO(double, n), *Au1 = P4EST_ALLOC_ZERO(double, n);
  d4est_mesh_init_field(p4est, u, !strcmp(ukind, "sinrand") ? sin_fcn : poly_fcn, ops, geom, factors,
                        INIT_FIELD_ON_LOBATTO, NULL);
  if (strcmp(ukind, "poly") != 0) {   /* "poly": the pure polynomial of Tests/Unit/d4est_test_laplacian_consistency.c */
    srand(102321); /* Tests/Unit/d4est_test_laplacian_speedup.c:429-432 */
    for (int i = 0; i < n; i++) u[i] += (double)rand() / (double)RAND_MAX;
  }

  d4est_elliptic_data_t vecs; memset(&vecs, 0, sizeof vecs);
  vecs.u = u; vecs.local_nodes = n; vecs.field_types = &ft; vecs.num_of_fields = 1;
  vecs.Au = Au0;
  double t0 = now_s();
  d4est_laplacian_apply_aij(p4est, ghost, gd, &vecs, fl0, ops, geom, quad, factors, 0);
  double t_base = now_s() - t0;
  vecs.Au = Au1; t0 = now_s();
  d4est_laplacian_with_opt_apply_aij(p4est, ghost, gd, &vecs, fl1, ops, geom, quad, factors, 0);
  double t_opt = now_s() - t0;
  double maxdiff = 0., s0 = 0., s1 = 0.;
  for (int i = 0; i < n; i++) { double d = fabs(Au0[i] - Au1[i]); if (d > maxdiff) maxdiff = d; s0 += Au0[i]; s1 += Au1[i]; }

  double t_best = t_opt;
  for (int r = 0; r < reps; r++) {
    t0 = now_s();
    d4est_laplacian_with_opt_apply_aij(p4est, ghost, gd, &vecs, fl1, ops, geom, quad, factors, 0);
    double t = now_s() - t0; if (t < t_best) t_best = t;
  }

  if (strcmp(out, "-") != 0) {
    d4b_mesh_desc_t *D = d4est_b200_flatten(p4est, ghost, fl1, ops, geom, quad, factors);
    FILE *f = fix_open(out);
    fix_write_desc(f, D);
    fix_write(f, "u", FIX_F64, n, u);
    fix_write(f, "Au_ref_baseline", FIX_F64, n, Au0);
    fix_write(f, "Au_ref_with_opt", FIX_F64, n, Au1);
    { /* mass-type term: the reference's d4est_quadrature_apply_fofufofvlilj element by element (u0 = u, vec = u),
         and the coefficient table J f(x, u0) it is equivalent to */
      double *mref = P4EST_ALLOC_ZERO(double, n);
      int64_t nqt = D->quad_off[D->n_local];
      double *coef = P4EST_ALLOC(double, nqt), *uq = P4EST_ALLOC(double, 32 * 32 * 32);
      for (p4est_topidx_t tt = p4est->first_local_tree; tt <= p4est->last_local_tree; ++tt) {
        sc_array_t *tq = &p4est_tree_array_index(p4est->trees, tt)->quadrants;
        for (size_t q = 0; q < tq->elem_count; ++q) {
          d4est_element_data_t *ed = (d4est_element_data_t *)p4est_quadrant_array_index(tq, q)->p.user_data;
          d4est_quadrature_volume_t mo;
          mo.dq = ed->dq; mo.tree = ed->tree; mo.q[0] = ed->q[0]; mo.q[1] = ed->q[1]; mo.q[2] = ed->q[2]; mo.element_id = ed->id;
          d4est_mesh_data_on_element_t md = d4est_mesh_data_on_element(factors, ed);
          double *J = d4est_mesh_get_jacobian_on_quadrature_points(factors, ed);
          d4est_quadrature_apply_fofufofvlilj(ops, geom, quad, &mo, QUAD_OBJECT_VOLUME, QUAD_INTEGRAND_UNKNOWN,
                                              &u[ed->nodal_stride], &u[ed->nodal_stride], NULL, ed->deg, md.xyz_quad, J,
                                              ed->deg_quad, &mref[ed->nodal_stride], test_fofu, NULL, NULL, NULL,
                                              QUAD_APPLY_MATRIX, 0, NULL);
          int n3 = d4est_lgl_get_nodes(3, ed->deg_quad);
          d4est_quadrature_interpolate(ops, quad, geom, &mo, QUAD_OBJECT_VOLUME, QUAD_INTEGRAND_UNKNOWN,
                                       &u[ed->nodal_stride], ed->deg, uq, ed->deg_quad);
          for (int i = 0; i < n3; i++)
            coef[ed->quad_stride + i] = J[i] * test_fofu(md.xyz_quad[0][i], md.xyz_quad[1][i], md.xyz_quad[2][i], uq[i], NULL);
        }
      }
      fix_write(f, "mass_coef", FIX_F64, nqt, coef);
      fix_write(f, "mass_ref", FIX_F64, n, mref);
      P4EST_FREE(mref); P4EST_FREE(coef); P4EST_FREE(uq);
    }
    if (!getenv("D4B_HARNESS_NO_SOLVER")) { /* reference smoother pieces on the same operator: cg_eigs (Solver/d4est_solver_cg_eigs.c:115) from a
         zero guess with rhs = A u, then d4est_solver_multigrid_smoother_cheby_iterate_aux (:81) */
      double *su = P4EST_ALLOC_ZERO(double, n), *sAu = P4EST_ALLOC_ZERO(double, n), *sr = P4EST_ALLOC_ZERO(double, n);
      d4est_elliptic_eqns_t fc; fc.apply_lhs = ref_apply_lhs; fc.build_residual = NULL; fc.user = fl1;
      d4est_elliptic_data_t sv; memset(&sv, 0, sizeof sv);
      sv.u = su; sv.Au = sAu; sv.rhs = Au1; sv.local_nodes = n; sv.field_types = &ft; sv.num_of_fields = 1;
      double eig = -1.;
      cg_eigs(p4est, &sv, &fc, ghost, gd, ops, geom, quad, factors, 10, 0, 0, &eig);
      fix_write(f, "sol_eig", FIX_F64, 1, &eig);
      fix_write(f, "sol_u_after_eigs", FIX_F64, n, su);
      d4est_solver_multigrid_smoother_cheby_iterate_aux(p4est, ops, geom, quad, factors, ghost, gd, &sv, &fc, sr, 8,
                                                        eig / 30., eig, 0, 0, 1);
      fix_write(f, "sol_u_after_cheby", FIX_F64, n, su);
      fix_write(f, "sol_r_after_cheby", FIX_F64, n, sr);
      P4EST_FREE(su); P4EST_FREE(sAu); P4EST_FREE(sr);
    }
    fclose(f);
    d4est_b200_flatten_free(D);
  }
  printf("{\"Au_baseline_33010\": %.17g, \"Au_with_opt_33010\": %.17g, ", n > 33010 ? Au0[33010] : 0., n > 33010 ? Au1[33010] : 0.);
  printf("\"elements\": %d, \"nodes\": %d, \"Au_sum_baseline\": %.17g, \"Au_sum_with_opt\": %.17g, "
         "\"max_abs_diff_variants\": %.3e, \"t_baseline_s\": %.6f, \"t_with_opt_s\": %.6f, "
         "\"t_with_opt_best_s\": %.6f, \"reps\": %d}\n",
         (int)p4est->local_num_quadrants, n, s0, s1, maxdiff, t_base, t_opt, t_best, reps);
  return 0;
}
