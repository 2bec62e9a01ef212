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
#endif
