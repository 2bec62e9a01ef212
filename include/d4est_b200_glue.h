/* d4est_b200_glue.h -- d4est-facing host C layer (compiled INSIDE the d4est tree,
 * against d4est's own headers).  It keeps the reference's C operator API:
 *
 *   d4est_b200_laplacian_apply_aij        <-> d4est_laplacian_with_opt_apply_aij
 *                                             (dGMath/d4est_laplacian_with_opt.h:23, .c:289)
 *   d4est_b200_level_create / _destroy    <-> per-level residency; hook is every
 *                                             d4est_mesh_update (Mesh/d4est_mesh.c:2790)
 *   d4est_b200_flatten / _flatten_free    <-> one walk of the reference's own face
 *                                             iterator (Mesh/d4est_mortars_with_opt.c:829)
 *                                             with recording callbacks
 *
 * Below this layer everything is plain pointers (include/d4est_b200.h).
 */
#ifndef D4EST_B200_GLUE_H
#define D4EST_B200_GLUE_H

#include <pXest.h>
#include <d4est_ghost.h>
#include <d4est_ghost_data.h>
#include <d4est_elliptic_data.h>
#include <d4est_operators.h>
#include <d4est_geometry.h>
#include <d4est_quadrature.h>
#include <d4est_mesh.h>
#include <d4est_laplacian_with_opt_flux.h>
#include <d4est_amr.h>
#include "d4est_b200.h"

/* Flatten one mesh level into the pointer-and-size description.  All arrays in the
 * returned descriptor are owned by it; release with d4est_b200_flatten_free. */
d4b_mesh_desc_t *d4est_b200_flatten(p4est_t *p4est, d4est_ghost_t *ghost,
                                    d4est_laplacian_with_opt_flux_data_t *flux_data,
                                    d4est_operators_t *ops, d4est_geometry_t *geom,
                                    d4est_quadrature_t *quad, d4est_mesh_data_t *factors);
void d4est_b200_flatten_free(d4b_mesh_desc_t *desc);

/* Device residency for one (p4est level, factors) pair, keyed on the factors pointer
 * (the multigrid passes a different d4est_mesh_data_t per level,
 * Solver/d4est_solver_multigrid_smoother_cheby.c:196-209). */
d4b_mesh_t *d4est_b200_level_get(p4est_t *p4est, d4est_ghost_t *ghost,
                                 d4est_laplacian_with_opt_flux_data_t *flux_data,
                                 d4est_operators_t *ops, d4est_geometry_t *geom,
                                 d4est_quadrature_t *quad, d4est_mesh_data_t *factors);
void d4est_b200_level_invalidate(d4est_mesh_data_t *factors); /* call after d4est_mesh_update */
void d4est_b200_level_invalidate_all(void);

/* Drop-in for d4est_laplacian_with_opt_apply_aij: same argument list, same contract
 * (reads vecs->u, overwrites vecs->Au for field which_field). */
void d4est_b200_laplacian_apply_aij(p4est_t *p4est, d4est_ghost_t *ghost,
                                    d4est_ghost_data_t *ghost_data,
                                    d4est_elliptic_data_t *vecs,
                                    d4est_laplacian_with_opt_flux_data_t *flux_data,
                                    d4est_operators_t *ops, d4est_geometry_t *geom,
                                    d4est_quadrature_t *quad, d4est_mesh_data_t *factors,
                                    int which_field);

/* The other entry points of the reference's C operator API for this path (all host C, all ending in the C ABI):
 *   d4est_b200_apply_lhs / d4est_b200_build_residual      d4est_apply_operator_fcn_t, EllipticSystem/d4est_elliptic_eqns.h:12-37
 *   d4est_b200_laplacian_baseline_apply_aij                d4est_laplacian_apply_aij, dGMath/d4est_laplacian.c:317
 *   d4est_b200_laplacian_build_rhs_with_strong_bc          dGMath/d4est_laplacian.c:16 */
#include <d4est_laplacian_flux.h>
#include <d4est_xyz_functions.h>
void d4est_b200_apply_lhs(p4est_t *p4est, d4est_ghost_t *ghost, d4est_ghost_data_t *ghost_data,
                          d4est_elliptic_data_t *vecs, d4est_operators_t *ops, d4est_geometry_t *geom,
                          d4est_quadrature_t *quad, d4est_mesh_data_t *factors, void *user);
void d4est_b200_build_residual(p4est_t *p4est, d4est_ghost_t *ghost, d4est_ghost_data_t *ghost_data,
                               d4est_elliptic_data_t *vecs, d4est_operators_t *ops, d4est_geometry_t *geom,
                               d4est_quadrature_t *quad, d4est_mesh_data_t *factors, void *user);
d4b_mesh_desc_t *d4est_b200_flatten_baseline(p4est_t *p4est, d4est_ghost_t *ghost, d4est_laplacian_flux_data_t *fd,
                                             d4est_operators_t *ops, d4est_geometry_t *geom,
                                             d4est_quadrature_t *quad, d4est_mesh_data_t *factors);
d4b_mesh_t *d4est_b200_level_get_baseline(p4est_t *p4est, d4est_ghost_t *ghost, d4est_laplacian_flux_data_t *fd,
                                          d4est_operators_t *ops, d4est_geometry_t *geom,
                                          d4est_quadrature_t *quad, d4est_mesh_data_t *factors);
void d4est_b200_laplacian_baseline_apply_aij(p4est_t *p4est, d4est_ghost_t *ghost, d4est_ghost_data_t *ghost_data,
                                             d4est_elliptic_data_t *vecs, d4est_laplacian_flux_data_t *flux_fcn_data,
                                             d4est_operators_t *ops, d4est_geometry_t *geom, d4est_quadrature_t *quad,
                                             d4est_mesh_data_t *factors, int which_field);
void d4est_b200_laplacian_build_rhs_with_strong_bc(p4est_t *p4est, d4est_ghost_t *ghost, d4est_ghost_data_t *ghost_data,
                                                   d4est_operators_t *ops, d4est_geometry_t *geom,
                                                   d4est_quadrature_t *quad, d4est_mesh_data_t *factors,
                                                   d4est_elliptic_data_t *prob_vecs,
                                                   d4est_laplacian_flux_data_t *flux_fcn_data_for_build_rhs,
                                                   double *rhs, d4est_xyz_fcn_t problem_rhs_fcn,
                                                   d4est_mesh_init_field_option_t init_option, void *ctx, int which_field);

/* Mass-type terms of apply_jac / build_residual (d4est_quadrature_apply_fofufofvlilj, Quadrature/d4est_quadrature.c:593;
 * callers e.g. Problems/ConstantDensityStar/constant_density_star_fcns.h:484-605): the coefficient J f(x, u0) is
 * tabulated once per linearisation point with the user's callback; Au += M_c u is then one device call. */
void d4est_b200_mass_set_fofu(p4est_t *p4est, d4est_ghost_t *ghost, d4est_laplacian_with_opt_flux_data_t *flux_data,
                              d4est_operators_t *ops, d4est_geometry_t *geom, d4est_quadrature_t *quad,
                              d4est_mesh_data_t *factors, double *u0, d4est_xyzu_fcn_t fofu_fcn, void *fofu_ctx);
void d4est_b200_mass_apply_add(p4est_t *p4est, d4est_ghost_t *ghost, d4est_laplacian_with_opt_flux_data_t *flux_data,
                               d4est_elliptic_data_t *vecs, d4est_operators_t *ops, d4est_geometry_t *geom,
                               d4est_quadrature_t *quad, d4est_mesh_data_t *factors);

/* ------------------------------------------------------------------ solver interfaces (d4est_b200_solver_glue.c)
 * Compiled only where d4est's Solver/ headers are on the include path (define D4EST_B200_NO_SOLVER_GLUE to skip). */
#ifndef D4EST_B200_NO_SOLVER_GLUE
#include <d4est_solver_multigrid.h>
#include <d4est_krylov_pc.h>

/* (1) Keep d4est's own V-cycle, run its smoother and bottom solver on the device: after
 *     d4est_solver_multigrid_data_init(...) (smoother_name = mg_smoother_cheby, bottom_solver_name =
 *     mg_bottom_solver_cheby) one call re-points d4est_solver_multigrid_smoother_t.smooth and
 *     d4est_solver_multigrid_bottom_solver_t.solve (Solver/d4est_solver_multigrid.h:129-144).  Per smoother call:
 *     two uploads, two downloads, cheby_eigs_cg_imax + cheby_imax + 1 operator applies on the device. */
void d4est_b200_multigrid_use_device(d4est_solver_multigrid_t *mg_data, d4est_laplacian_with_opt_flux_data_t *flux_data);

/* (2) Fully resident multigrid: walk d4est's hierarchy once (its own coarsen / balance / element-data updater code
 *     runs; see d4est_b200_solver_glue.c) and keep every level and transfer table in HBM. */
typedef struct d4est_b200_mg d4est_b200_mg_t;
typedef struct {   /* one entry per COARSE element: the reference's d4est_solver_multigrid_refine_data_t, flattened */
  int32_t n_coarse; int32_t *kind, *degH, *child_deg; int64_t *coarse_off, *child_off; int64_t fine_nodes, coarse_nodes;
} d4est_b200_mg_transfer_table_t;
d4est_b200_mg_t *d4est_b200_mg_capture(p4est_t *p4est, d4est_solver_multigrid_t *mg_data,
                                       d4est_laplacian_with_opt_flux_data_t *flux_data, d4est_elliptic_data_t *vecs_template);
void d4est_b200_mg_capture_tables(p4est_t *p4est, d4est_solver_multigrid_t *mg_data,
                                  d4est_laplacian_with_opt_flux_data_t *flux_data, d4est_elliptic_data_t *vecs_template,
                                  d4b_mesh_desc_t ***desc_out, d4est_b200_mg_transfer_table_t **tt_out);
void d4est_b200_mg_transfer_table_free(d4est_b200_mg_transfer_table_t *T);
void d4est_b200_mg_params_from_d4est(const d4est_solver_multigrid_t *mg_data, d4b_mg_params_t *par);
void d4est_b200_mg_destroy(d4est_b200_mg_t *G);
d4b_mg_t *d4est_b200_mg_handle(d4est_b200_mg_t *G);

/* (3) d4est_krylov_pc_t (Solver/d4est_krylov_pc.h:13-50) = one resident V-cycle per application; drop-in for
 *     d4est_krylov_pc_multigrid_create (Solver/d4est_krylov_pc_multigrid.c:77) in FCG, CG or the PETSc PCShell. */
d4est_krylov_pc_t *d4est_b200_krylov_pc_multigrid_create(d4est_b200_mg_t *G);
void d4est_b200_krylov_pc_multigrid_destroy(d4est_krylov_pc_t *pc);

/* (4) d4est_solver_fcg_solve's argument list (Solver/d4est_solver_fcg_improved.c:97); fcns->user must be the
 *     d4est_laplacian_with_opt_flux_data_t of the operator.  Returns the iteration count (the reference returns
 *     nothing).  With the pc of (3) the whole solve is resident: one upload of u and rhs, one download of u. */
int d4est_b200_fcg_solve(p4est_t *p4est, d4est_elliptic_data_t *vecs, d4est_elliptic_eqns_t *fcns,
                         d4est_ghost_t **ghost, d4est_ghost_data_t **ghost_data, d4est_operators_t *ops,
                         d4est_geometry_t *geom, d4est_quadrature_t *quad, d4est_mesh_data_t *factors, void *fcg_params,
                         d4est_krylov_pc_t *pc, int amr_level, int newton_iteration);

/* (4b) AMR field transfer: same effect as the static d4est_amr_interpolate_field(p4est, ops, d4est_amr, field,
 *      post_balance_nodes) (hpAMR/d4est_amr.c:406, called from d4est_amr_step :1052) -- *field is replaced by the
 *      field on the refined and balanced grid, computed on the GPU from the logs d4est_amr_step has just filled.
 *      Returns 0 or a D4B_E* code (d4b_last_error()). */
int d4est_b200_amr_interpolate_field(d4est_amr_t *d4est_amr, double **field, int post_balance_nodes);

/* (5) with -DD4EST_B200_HAVE_PETSC: PetscErrorCode d4est_b200_krylov_petsc_apply_aij(Mat, Vec, Vec), the MatShell
 *     multiply (Solver/d4est_solver_krylov_petsc.c:273, installed :428). */
#endif
#endif
