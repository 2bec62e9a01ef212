/* d4est_b200_solver_glue.c -- the reference's SOLVER interfaces on the device path (host C, compiled against
 * d4est's own headers; see include/d4est_b200_glue.h):
 *
 *   d4est_b200_multigrid_use_device   keeps d4est's own V-cycle (Solver/d4est_solver_multigrid.c:751) and replaces
 *                                     the bodies behind d4est_solver_multigrid_smoother_t.smooth and
 *                                     d4est_solver_multigrid_bottom_solver_t.solve (multigrid.h:129-144) by one
 *                                     device call per smoother call (d4b_cheby_smooth_host)
 *   d4est_b200_mg_capture             walks d4est's hierarchy ONCE with its own coarsen / balance / updater code
 *                                     and turns it into resident device levels + transfer tables (d4b_mg_t)
 *   d4est_b200_krylov_pc_multigrid_create   d4est_krylov_pc_t (Solver/d4est_krylov_pc.h:13-50) whose pc_apply is the
 *                                     resident device V-cycle
 *   d4est_b200_fcg_solve              d4est_solver_fcg_solve's argument list (Solver/d4est_solver_fcg_improved.c:97);
 *                                     the whole solve on the device when the pc is the one above
 *   PETSc MatShell / PCShell thunks   under D4EST_B200_HAVE_PETSC (PETSc is absent from this snapshot)
 */
#include "d4est_b200_glue.h"
#include <d4est_solver_multigrid.h>
#include <d4est_solver_multigrid_smoother_cheby.h>
#include <d4est_solver_multigrid_bottom_solver_cheby.h>
#include <d4est_solver_multigrid_bottom_solver_cg.h>
#include <d4est_solver_multigrid_element_data_updater.h>
#include <d4est_solver_fcg_improved.h>
#include <d4est_krylov_pc.h>
#include <d4est_lgl.h>
#include <d4est_util.h>
#include <stdlib.h>
#include <string.h>
#include <stdio.h>

static void check_rc(int rc, const char *what)
{
  if (rc == D4B_OK) return;
  fprintf(stderr, "d4est_b200: %s: %s\n", what, d4b_last_error());
  D4EST_ABORT("d4est_b200: device call failed (no CPU fallback)");
}

/* ------------------------------------------------------------------ smoother / bottom solver inside d4est's V-cycle */
static d4est_laplacian_with_opt_flux_data_t *g_mg_flux = NULL;   /* operator of the multigrid in use */

static d4b_mesh_t *current_level(p4est_t *p4est)
{
  d4est_solver_multigrid_t *mg = (d4est_solver_multigrid_t *)p4est->user_pointer;
  d4est_solver_multigrid_element_data_updater_t *up = mg->elem_data_updater;
  return d4est_b200_level_get(p4est, up->current_d4est_ghost, g_mg_flux, mg->d4est_ops, mg->d4est_geom,
                              mg->d4est_quad, up->current_d4est_factors);
}

/* d4est_solver_multigrid_smoother_fcn_t; body of d4est_solver_multigrid_smoother_cheby
 * (Solver/d4est_solver_multigrid_smoother_cheby.c:264-360) with the eigs_compute flag maintained by the
 * reference's own update callback (:221-262), which stays installed */
static void b200_cheby_smooth(p4est_t *p4est, d4est_elliptic_data_t *vecs, d4est_elliptic_eqns_t *fcns, double *r, int level)
{
  d4est_solver_multigrid_t *mg = (d4est_solver_multigrid_t *)p4est->user_pointer;
  d4est_solver_multigrid_smoother_cheby_t *ch = (d4est_solver_multigrid_smoother_cheby_t *)mg->smoother->user;
  if (ch->cheby_use_zero_guess_for_eigs) D4EST_ABORT("d4est_b200: cheby_use_zero_guess_for_eigs is not supported");
  check_rc(d4b_cheby_smooth_host(current_level(p4est), vecs->u, vecs->rhs, r, ch->cheby_imax, ch->cheby_eigs_cg_imax,
                                 ch->cheby_eigs_lmax_lmin_ratio, ch->cheby_eigs_max_multiplier, ch->cheby_use_new_cg_eigs,
                                 ch->cheby_eigs_compute, &ch->eigs[level]), "smoother");
  (void)fcns;
}

/* d4est_solver_multigrid_bottom_solver_fcn_t; body of d4est_solver_multigrid_bottom_solver_cheby
 * (Solver/d4est_solver_multigrid_bottom_solver_cheby.c:59-113) */
static void b200_bottom_cheby(p4est_t *p4est, d4est_elliptic_data_t *vecs, d4est_elliptic_eqns_t *fcns, double *r)
{
  d4est_solver_multigrid_t *mg = (d4est_solver_multigrid_t *)p4est->user_pointer;
  d4est_solver_multigrid_bottom_solver_cheby_t *ch = (d4est_solver_multigrid_bottom_solver_cheby_t *)mg->bottom_solver->user;
  check_rc(d4b_cheby_smooth_host(current_level(p4est), vecs->u, vecs->rhs, r, ch->cheby_imax, ch->cheby_eigs_cg_imax,
                                 ch->cheby_eigs_lmax_lmin_ratio, ch->cheby_eigs_max_multiplier, ch->cheby_use_new_cg_eigs,
                                 1, &ch->eig), "bottom solver");
  (void)fcns;
}

void d4est_b200_multigrid_use_device(d4est_solver_multigrid_t *mg_data, d4est_laplacian_with_opt_flux_data_t *flux_data)
{
  if (mg_data->smoother->type != MG_SMOOTHER_CHEBY)
    D4EST_ABORT("d4est_b200: only mg_smoother_cheby has a device body");
  if (!d4est_util_match(mg_data->bottom_solver_name, "mg_bottom_solver_cheby"))
    D4EST_ABORT("d4est_b200: only mg_bottom_solver_cheby has a device body inside d4est's V-cycle");
  g_mg_flux = flux_data;
  mg_data->smoother->smooth = b200_cheby_smooth;
  mg_data->bottom_solver->solve = b200_bottom_cheby;
}

/* ------------------------------------------------------------------ hierarchy capture */
struct d4est_b200_mg {
  int n_levels;
  d4b_mesh_t **levels;
  d4b_mg_transfer_t **transfers;
  d4b_mg_t *mg;
  int64_t top_nodes;
};

typedef struct {
  d4est_laplacian_with_opt_flux_data_t *flux;
  int n_levels;
  d4b_mesh_desc_t **desc;            /* [n_levels] */
  d4est_b200_mg_transfer_table_t *tt; /* [n_levels - 1]: tt[l] joins l (coarse) and l + 1 (fine) */
} capture_t;

/* transfer table between the CURRENT (coarse, balanced) p4est and the finer level, read from the reference's
 * coarse_grid_refinement[] exactly as its restriction callback walks it
 * (Solver/d4est_solver_multigrid_callbacks.h:109-216): hrefine 0 = same cell (p-prolongation), 1 = eight children,
 * 2 = a family that 2:1 balance split again: eight coarse elements that are copies of their fine twins */
static void capture_transfer(p4est_t *p4est, d4est_solver_multigrid_t *mg, d4est_b200_mg_transfer_table_t *T)
{
  int nc = (int)p4est->local_num_quadrants;
  T->n_coarse = nc;
  T->kind = (int32_t *)malloc((size_t)nc * sizeof(int32_t));
  T->degH = (int32_t *)malloc((size_t)nc * sizeof(int32_t));
  T->coarse_off = (int64_t *)malloc((size_t)nc * sizeof(int64_t));
  T->child_deg = (int32_t *)calloc((size_t)8 * nc, sizeof(int32_t));
  T->child_off = (int64_t *)calloc((size_t)8 * nc, sizeof(int64_t));
  const d4est_solver_multigrid_refine_data_t *cgr = mg->coarse_grid_refinement;
  int stride = mg->stride, temp = 0;
  int64_t fine = 0, coarse = 0;
  for (int e = 0; e < nc; e++) {
    const d4est_solver_multigrid_refine_data_t *R = &cgr[stride];
    T->coarse_off[e] = coarse;
    if (R->hrefine == 0) {
      T->kind[e] = 0; T->degH[e] = R->degH; T->child_deg[8 * e] = R->degh[0]; T->child_off[8 * e] = fine;
      fine += d4est_lgl_get_nodes(3, R->degh[0]); coarse += d4est_lgl_get_nodes(3, R->degH); stride++;
    } else if (R->hrefine == 1) {
      T->kind[e] = 1; T->degH[e] = R->degH;
      for (int c = 0; c < 8; c++) { T->child_deg[8 * e + c] = R->degh[c]; T->child_off[8 * e + c] = fine; fine += d4est_lgl_get_nodes(3, R->degh[c]); }
      coarse += d4est_lgl_get_nodes(3, R->degH); stride++;
    } else {
      int deg = R->degh[temp], nodes = d4est_lgl_get_nodes(3, deg);
      T->kind[e] = 0; T->degH[e] = deg; T->child_deg[8 * e] = deg; T->child_off[8 * e] = fine;
      fine += nodes; coarse += nodes;
      if (temp == 7) { temp = 0; stride++; } else temp++;
    }
  }
  T->fine_nodes = fine; T->coarse_nodes = coarse;
}

static void capture_update(p4est_t *p4est, int level, d4est_elliptic_data_t *vecs)
{
  d4est_solver_multigrid_t *mg = (d4est_solver_multigrid_t *)p4est->user_pointer;
  capture_t *C = (capture_t *)mg->user_callbacks->user;
  d4est_solver_multigrid_element_data_updater_t *up = mg->elem_data_updater;
  int which = -1;
  if (mg->mg_state == PRE_V) which = mg->num_of_levels - 1;
  else if (mg->mg_state == DOWNV_POST_BALANCE) which = level - 1;   /* the updater has just switched to it (:115-163) */
  if (which >= 0 && !C->desc[which])
    C->desc[which] = d4est_b200_flatten(p4est, up->current_d4est_ghost, C->flux, mg->d4est_ops, mg->d4est_geom,
                                        mg->d4est_quad, up->current_d4est_factors);
  if (mg->mg_state == DOWNV_PRE_RESTRICTION && !C->tt[level - 1].kind) capture_transfer(p4est, mg, &C->tt[level - 1]);
  (void)vecs;
}

static void noop_smooth(p4est_t *p4est, d4est_elliptic_data_t *vecs, d4est_elliptic_eqns_t *fcns, double *r, int level)
{ memset(r, 0, (size_t)vecs->local_nodes * sizeof(double)); (void)p4est; (void)fcns; (void)level; }
static void noop_bottom(p4est_t *p4est, d4est_elliptic_data_t *vecs, d4est_elliptic_eqns_t *fcns, double *r)
{ memset(r, 0, (size_t)vecs->local_nodes * sizeof(double)); (void)p4est; (void)fcns; }
static void noop_apply(p4est_t *p4est, d4est_ghost_t *ghost, d4est_ghost_data_t *gd, d4est_elliptic_data_t *vecs,
                       d4est_operators_t *ops, d4est_geometry_t *geom, d4est_quadrature_t *quad, d4est_mesh_data_t *f, void *user)
{ memset(vecs->Au, 0, (size_t)vecs->local_nodes * sizeof(double)); (void)p4est; (void)ghost; (void)gd; (void)ops; (void)geom; (void)quad; (void)f; (void)user; }

void d4est_b200_mg_params_from_d4est(const d4est_solver_multigrid_t *mg_data, d4b_mg_params_t *par)
{
  memset(par, 0, sizeof *par);
  par->vcycle_imax = mg_data->vcycle_imax; par->vcycle_rtol = mg_data->vcycle_rtol; par->vcycle_atol = mg_data->vcycle_atol;
  if (mg_data->smoother->type != MG_SMOOTHER_CHEBY) D4EST_ABORT("d4est_b200: the device V-cycle smooths with mg_smoother_cheby");
  const d4est_solver_multigrid_smoother_cheby_t *ch = (const d4est_solver_multigrid_smoother_cheby_t *)mg_data->smoother->user;
  if (ch->cheby_use_zero_guess_for_eigs) D4EST_ABORT("d4est_b200: cheby_use_zero_guess_for_eigs is not supported");
  par->cheby_imax = ch->cheby_imax; par->cheby_eigs_cg_imax = ch->cheby_eigs_cg_imax;
  par->cheby_eigs_lmax_lmin_ratio = ch->cheby_eigs_lmax_lmin_ratio; par->cheby_eigs_max_multiplier = ch->cheby_eigs_max_multiplier;
  par->eigs_reuse_fromdownvcycle = ch->cheby_eigs_reuse_fromdownvcycle; par->eigs_reuse_fromlastvcycle = ch->cheby_eigs_reuse_fromlastvcycle;
  par->use_new_cg_eigs = ch->cheby_use_new_cg_eigs;
  if (d4est_util_match(mg_data->bottom_solver_name, "mg_bottom_solver_cheby")) {
    const d4est_solver_multigrid_bottom_solver_cheby_t *b = (const d4est_solver_multigrid_bottom_solver_cheby_t *)mg_data->bottom_solver->user;
    par->bottom_kind = 0; par->bottom_imax = b->cheby_imax; par->bottom_eigs_cg_imax = b->cheby_eigs_cg_imax;
    par->bottom_eigs_lmax_lmin_ratio = b->cheby_eigs_lmax_lmin_ratio; par->bottom_eigs_max_multiplier = b->cheby_eigs_max_multiplier;
    /* the bottom solver reads its own use_new flag (bottom_solver_cheby.c:82); the device driver has one flag */
    if (b->cheby_use_new_cg_eigs != ch->cheby_use_new_cg_eigs) D4EST_ABORT("d4est_b200: smoother and bottom solver disagree on cheby_use_new_cg_eigs");
  } else if (d4est_util_match(mg_data->bottom_solver_name, "mg_bottom_solver_cg")) {
    const d4est_solver_multigrid_bottom_solver_cg_t *b = (const d4est_solver_multigrid_bottom_solver_cg_t *)mg_data->bottom_solver->user;
    par->bottom_kind = 1; par->bottom_imax = b->bottom_imax; par->bottom_atol = b->bottom_atol; par->bottom_rtol = b->bottom_rtol;
  } else D4EST_ABORT("d4est_b200: bottom solver must be mg_bottom_solver_cheby or mg_bottom_solver_cg");
}

/* Walk d4est's hierarchy once.  The walk IS d4est_solver_multigrid_solve with one V-cycle, the smoother, bottom
 * solver and operator replaced by no-ops: coarsening (p4est_coarsen_ext + the min-degree rule,
 * multigrid_callbacks.h:23-75), 2:1 balance with its bookkeeping (:77-105), the per-level element data and
 * geometric factors (multigrid_element_data_updater.c:88) and the refinement back to the fine mesh all run in the
 * reference's own code; the capture callback only reads. */
void d4est_b200_mg_capture_tables(p4est_t *p4est, d4est_solver_multigrid_t *mg_data,
                                  d4est_laplacian_with_opt_flux_data_t *flux_data, d4est_elliptic_data_t *vecs_template,
                                  d4b_mesh_desc_t ***desc_out, d4est_b200_mg_transfer_table_t **tt_out)
{
  int nl = mg_data->num_of_levels, n = vecs_template->local_nodes;
  capture_t C; C.flux = flux_data; C.n_levels = nl;
  C.desc = (d4b_mesh_desc_t **)calloc((size_t)nl, sizeof(void *));
  C.tt = (d4est_b200_mg_transfer_table_t *)calloc((size_t)(nl > 1 ? nl - 1 : 1), sizeof *C.tt);
  d4est_solver_multigrid_user_callbacks_t cb; memset(&cb, 0, sizeof cb);
  cb.update = capture_update; cb.user = &C;
  /* save what the walk overrides */
  d4est_solver_multigrid_smoother_fcn_t s_smooth = mg_data->smoother->smooth;
  d4est_solver_multigrid_bottom_solver_fcn_t s_solve = mg_data->bottom_solver->solve;
  d4est_solver_multigrid_user_callbacks_t *s_cb = mg_data->user_callbacks;
  d4est_solver_multigrid_logger_t *s_log = mg_data->logger;
  int s_imax = mg_data->vcycle_imax; double s_rtol = mg_data->vcycle_rtol, s_atol = mg_data->vcycle_atol;
  mg_data->smoother->smooth = noop_smooth; mg_data->bottom_solver->solve = noop_bottom;
  mg_data->user_callbacks = &cb; mg_data->logger = NULL;
  mg_data->vcycle_imax = 1; mg_data->vcycle_rtol = 0.; mg_data->vcycle_atol = 0.;
  double *u = P4EST_ALLOC_ZERO(double, n), *Au = P4EST_ALLOC_ZERO(double, n), *rhs = P4EST_ALLOC(double, n);
  for (int i = 0; i < n; i++) rhs[i] = 1.;   /* nonzero initial residual: the V-cycle loop runs once (multigrid.c:1463) */
  d4est_elliptic_data_t v; d4est_elliptic_data_copy_ptrs(vecs_template, &v);
  v.u = u; v.Au = Au; v.rhs = rhs; v.local_nodes = n;
  d4est_elliptic_eqns_t noop; noop.apply_lhs = noop_apply; noop.build_residual = noop_apply; noop.user = NULL;
  d4est_solver_multigrid_solve(p4est, &v, &noop, mg_data);
  P4EST_FREE(u); P4EST_FREE(Au); P4EST_FREE(rhs);
  mg_data->smoother->smooth = s_smooth; mg_data->bottom_solver->solve = s_solve;
  mg_data->user_callbacks = s_cb; mg_data->logger = s_log;
  mg_data->vcycle_imax = s_imax; mg_data->vcycle_rtol = s_rtol; mg_data->vcycle_atol = s_atol;
  for (int l = 0; l < nl; l++) if (!C.desc[l]) D4EST_ABORT("d4est_b200: hierarchy capture missed a level");
  for (int l = 0; l + 1 < nl; l++) {
    if (!C.tt[l].kind) D4EST_ABORT("d4est_b200: hierarchy capture missed a transfer");
    if (C.tt[l].coarse_nodes != C.desc[l]->node_off[C.desc[l]->n_local] || C.tt[l].fine_nodes != C.desc[l + 1]->node_off[C.desc[l + 1]->n_local])
      D4EST_ABORT("d4est_b200: transfer table does not match the level sizes");
  }
  *desc_out = C.desc; *tt_out = C.tt;
}

void d4est_b200_mg_transfer_table_free(d4est_b200_mg_transfer_table_t *T)
{ free(T->kind); free(T->degH); free(T->coarse_off); free(T->child_deg); free(T->child_off); memset(T, 0, sizeof *T); }

d4est_b200_mg_t *d4est_b200_mg_capture(p4est_t *p4est, d4est_solver_multigrid_t *mg_data,
                                       d4est_laplacian_with_opt_flux_data_t *flux_data, d4est_elliptic_data_t *vecs_template)
{
  d4b_mesh_desc_t **desc; d4est_b200_mg_transfer_table_t *tt;
  d4est_b200_mg_capture_tables(p4est, mg_data, flux_data, vecs_template, &desc, &tt);
  int nl = mg_data->num_of_levels;
  d4est_b200_mg_t *G = (d4est_b200_mg_t *)calloc(1, sizeof *G);
  G->n_levels = nl;
  G->levels = (d4b_mesh_t **)calloc((size_t)nl, sizeof(void *));
  G->transfers = (d4b_mg_transfer_t **)calloc((size_t)nl, sizeof(void *));
  for (int l = 0; l < nl; l++) {
    check_rc(d4b_mesh_create(desc[l], &G->levels[l]), "mg level");
    if (l + 1 < nl)
      check_rc(d4b_mg_transfer_create(tt[l].n_coarse, tt[l].kind, tt[l].degH, tt[l].coarse_off, tt[l].child_deg,
                                      tt[l].child_off, &G->transfers[l]), "mg transfer");
  }
  G->top_nodes = desc[nl - 1]->node_off[desc[nl - 1]->n_local];
  for (int l = 0; l < nl; l++) d4est_b200_flatten_free(desc[l]);
  for (int l = 0; l + 1 < nl; l++) d4est_b200_mg_transfer_table_free(&tt[l]);
  free(desc); free(tt);
  d4b_mg_params_t par; d4est_b200_mg_params_from_d4est(mg_data, &par);
  check_rc(d4b_mg_create(nl, G->levels, G->transfers, &par, &G->mg), "mg create");
  return G;
}

void d4est_b200_mg_destroy(d4est_b200_mg_t *G)
{
  if (!G) return;
  d4b_mg_destroy(G->mg);
  for (int l = 0; l < G->n_levels; l++) { d4b_mesh_destroy(G->levels[l]); if (l + 1 < G->n_levels) d4b_mg_transfer_destroy(G->transfers[l]); }
  free(G->levels); free(G->transfers); free(G);
}
d4b_mg_t *d4est_b200_mg_handle(d4est_b200_mg_t *G) { return G->mg; }

/* ------------------------------------------------------------------ d4est_krylov_pc_t on the resident V-cycle */
static void b200_pc_apply(d4est_krylov_pc_t *kpc, double *xp, double *yp)
{ /* = d4est_krylov_pc_multigrid_apply (Solver/d4est_krylov_pc_multigrid.c:41-75): y = MG(x) from a zero guess */
  d4est_b200_mg_t *G = (d4est_b200_mg_t *)kpc->pc_data;
  check_rc(d4b_mg_pc_apply_host(G->mg, xp, yp), "pc_apply");
}
d4est_krylov_pc_t *d4est_b200_krylov_pc_multigrid_create(d4est_b200_mg_t *G)
{
  d4est_krylov_pc_t *pc = P4EST_ALLOC(d4est_krylov_pc_t, 1);
  pc->pc_apply = b200_pc_apply; pc->pc_setup = NULL; pc->pc_data = G; pc->pc_ctx = NULL;
  return pc;
}
void d4est_b200_krylov_pc_multigrid_destroy(d4est_krylov_pc_t *pc) { P4EST_FREE(pc); }

/* ------------------------------------------------------------------ flexible CG, d4est_solver_fcg_solve's argument list */
typedef struct { d4est_krylov_pc_t *pc; int64_t n; double *x, *y; } host_pc_t;
static int host_pc_thunk(void *ctx, const double *r_dev, double *z_dev, void *stream)
{ /* any other d4est_krylov_pc_t: its pc_apply wants host arrays, so this iteration's r and z are staged */
  host_pc_t *H = (host_pc_t *)ctx;
  int rc;
  if ((rc = d4b_d2h(H->x, r_dev, (size_t)H->n * sizeof(double), stream)) || (rc = d4b_sync(stream))) return rc;
  H->pc->pc_apply(H->pc, H->x, H->y);
  if ((rc = d4b_h2d(z_dev, H->y, (size_t)H->n * sizeof(double), stream))) return rc;
  return d4b_sync(stream);
}

int d4est_b200_fcg_solve(p4est_t *p4est, d4est_elliptic_data_t *vecs, d4est_elliptic_eqns_t *fcns,
                         d4est_ghost_t **ghost, d4est_ghost_data_t **ghost_data, d4est_operators_t *ops,
                         d4est_geometry_t *geom, d4est_quadrature_t *quad, d4est_mesh_data_t *factors, void *fcg_params,
                         d4est_krylov_pc_t *pc, int amr_level, int newton_iteration)
{
  d4est_solver_fcg_params_t *par = (d4est_solver_fcg_params_t *)fcg_params;
  d4b_mesh_t *m = d4est_b200_level_get(p4est, *ghost, (d4est_laplacian_with_opt_flux_data_t *)fcns->user, ops, geom, quad, factors);
  const int64_t n = vecs->local_nodes;
  int iters = 0; double res = 0.;
  const int use_pc = pc != NULL && !par->do_not_use_preconditioner;
  if (use_pc && pc->pc_apply == b200_pc_apply) {   /* everything resident: one upload, one download */
    d4est_b200_mg_t *G = (d4est_b200_mg_t *)pc->pc_data;
    if (G->top_nodes != n) D4EST_ABORT("d4est_b200: captured hierarchy does not match this mesh (re-capture after d4est_mesh_update)");
    check_rc(d4b_mg_fcg_solve_host(G->mg, vecs->u, vecs->rhs, par->imax, par->atol, par->rtol, &iters, &res), "fcg");
  } else {
    void *du = NULL, *drhs = NULL, *dwork = NULL;
    check_rc(d4b_dev_alloc((size_t)n * 8, &du), "alloc"); check_rc(d4b_dev_alloc((size_t)n * 8, &drhs), "alloc");
    check_rc(d4b_dev_alloc((size_t)n * 48, &dwork), "alloc");
    check_rc(d4b_h2d(du, vecs->u, (size_t)n * 8, NULL), "h2d"); check_rc(d4b_h2d(drhs, vecs->rhs, (size_t)n * 8, NULL), "h2d");
    host_pc_t H; H.pc = pc; H.n = n; H.x = NULL; H.y = NULL;
    if (use_pc) {
      H.x = P4EST_ALLOC(double, n); H.y = P4EST_ALLOC(double, n);
      krylov_ctx_t kctx; memset(&kctx, 0, sizeof kctx);   /* what fcg_improved.c:140-160 hands to the pc */
      kctx.p4est = p4est; kctx.vecs = vecs; kctx.fcns = fcns; kctx.ghost = ghost; kctx.ghost_data = ghost_data;
      kctx.d4est_ops = ops; kctx.d4est_geom = geom; kctx.d4est_quad = quad; kctx.d4est_factors = factors;
      kctx.amr_level = amr_level; kctx.using_newton = newton_iteration != -1; kctx.newton_iteration = newton_iteration;
      pc->pc_ctx = &kctx;
      if (pc->pc_setup) pc->pc_setup(pc);
      check_rc(d4b_fcg_solve(m, (double *)du, (const double *)drhs, par->imax, par->atol, par->rtol, host_pc_thunk, &H, &iters, &res, (double *)dwork, NULL), "fcg");
      P4EST_FREE(H.x); P4EST_FREE(H.y);
    } else
      check_rc(d4b_fcg_solve(m, (double *)du, (const double *)drhs, par->imax, par->atol, par->rtol, NULL, NULL, &iters, &res, (double *)dwork, NULL), "fcg");
    check_rc(d4b_d2h(vecs->u, du, (size_t)n * 8, NULL), "d2h"); check_rc(d4b_sync(NULL), "sync");
    d4b_dev_free(du); d4b_dev_free(drhs); d4b_dev_free(dwork);
  }
  return iters;
}

/* ------------------------------------------------------------------ PETSc hooks (PETSc is not in this snapshot) */
/* ---- AMR field transfer: what d4est_amr_interpolate_field (hpAMR/d4est_amr.c:406-485) does with the logs of the
 * step that has just run, on the GPU.  *field is reallocated the way the reference does it (D4EST_REALLOC there). */
int d4est_b200_amr_interpolate_field(d4est_amr_t *d4est_amr, double **field, int post_balance_nodes)
{
  int64_t n_in = 0, n_aux = 0, n_out = 0;
  int rc = d4b_amr_interpolate_field_sizes(d4est_amr->initial_log_size, d4est_amr->initial_log, d4est_amr->refinement_log,
                                           d4est_amr->balance_log_size, d4est_amr->balance_log, &n_in, &n_aux, &n_out);
  if (rc != D4B_OK) return rc;
  if (n_out != post_balance_nodes) {
    fprintf(stderr, "d4est_b200_amr_interpolate_field: logs give %lld nodes, the mesh has %d\n", (long long)n_out, post_balance_nodes);
    return D4B_EINVAL;
  }
  double *out = D4EST_ALLOC(double, post_balance_nodes);
  rc = d4b_amr_interpolate_field_host(d4est_amr->initial_log_size, d4est_amr->initial_log, d4est_amr->refinement_log,
                                      d4est_amr->balance_log_size, d4est_amr->balance_log, *field, n_in, out, n_out);
  if (rc != D4B_OK) { D4EST_FREE(out); return rc; }
  D4EST_FREE(*field);
  *field = out;
  return D4B_OK;
}

#ifdef D4EST_B200_HAVE_PETSC
#include <petscsnes.h>
/* MatShell multiply, installed with MatShellSetOperation(A, MATOP_MULT, ...) in place of
 * d4est_solver_krylov_petsc_apply_aij (Solver/d4est_solver_krylov_petsc.c:273-315, installed :428): same body,
 * the operator call going through the device entry point */
PetscErrorCode d4est_b200_krylov_petsc_apply_aij(Mat A, Vec x, Vec y)
{
  void *ctx; PetscErrorCode ierr = MatShellGetContext(A, &ctx); CHKERRQ(ierr);
  krylov_ctx_t *k = (krylov_ctx_t *)ctx;
  const double *px; double *py;
  ierr = VecGetArrayRead(x, &px); CHKERRQ(ierr); ierr = VecGetArray(y, &py); CHKERRQ(ierr);
  d4est_elliptic_data_t v; d4est_elliptic_data_copy_ptrs(k->vecs, &v);
  v.u = (double *)px; v.Au = py;
  d4est_b200_apply_lhs(k->p4est, *k->ghost, *k->ghost_data, &v, k->d4est_ops, k->d4est_geom, k->d4est_quad,
                       k->d4est_factors, k->fcns->user);
  ierr = VecRestoreArrayRead(x, &px); CHKERRQ(ierr); ierr = VecRestoreArray(y, &py); CHKERRQ(ierr);
  return 0;
}
/* PCShell apply: the reference's own thunk d4est_solver_krylov_petsc_pc_apply (Solver/d4est_krylov_pc.h:71-90)
 * already calls kp->pc_apply; with kp = d4est_b200_krylov_pc_multigrid_create(...) nothing else changes. */
#endif
