/* d4b_amr_harness.c -- TEST INFRASTRUCTURE ONLY.  Runs the UNMODIFIED reference's AMR step with a nodal field
 * attached (hpAMR/d4est_amr.c:868 d4est_amr_step -> :406 d4est_amr_interpolate_field: initial grid -> refined grid
 * by p- or hp-prolongation per element, refined grid -> 2:1-balanced grid by hp-prolongation) and dumps, per step,
 * the field before, the three logs that drive the interpolation and the field after, so that the device transfer
 * (d4b_amr_interpolate_field_*, d4est_b200/csrc/d4b_mg.cu) is pinned to the reference itself.
 *
 * usage: d4b_amr_harness <input> <steps> <out.fix>
 * writes: steps; per step k: S<k>.initial_log, S<k>.refinement_log, S<k>.balance_log (int32), S<k>.field_pre,
 *         S<k>.field_post (double).  With D4EST_B200_GPU_CHECK=1 in the environment the same transfer is also run
 *         through the glue entry d4est_b200_amr_interpolate_field on the GPU, in process, and compared. */
#include <pXest.h>
#include <d4est_operators.h>
#include <d4est_geometry.h>
#include <d4est_mesh.h>
#include <d4est_amr.h>
#include <d4est_amr_random.h>
#include <d4est_quadrature.h>
#include <d4est_ghost.h>
#include <math.h>
#include <stdlib.h>
#include "fixture_io.h"
#include "d4est_b200_glue.h"

static d4est_mesh_local_sizes_t update_mesh(p4est_t *p4est, d4est_ghost_t **ghost, d4est_operators_t *ops,
                                            d4est_geometry_t *geom, d4est_quadrature_t *quad,
                                            d4est_mesh_data_t *factors, d4est_mesh_initial_extents_t *ext, int initial)
{
  return d4est_mesh_update(p4est, ghost, ops, geom, quad, factors, ext, INITIALIZE_GHOST,
                           INITIALIZE_QUADRATURE_DATA, INITIALIZE_GEOMETRY_DATA, INITIALIZE_GEOMETRY_ALIASES,
                           initial ? d4est_mesh_set_initial_extents : d4est_mesh_set_quadratures_after_amr,
                           (void *)ext);
}

int main(int argc, char **argv)
{
  if (argc < 4) { fprintf(stderr, "usage: d4b_amr_harness <input> <steps> <out.fix>\n"); return 2; }
  const char *input = argv[1]; const int steps = atoi(argv[2]); const char *out = argv[3];
  const int gpu_check = getenv("D4EST_B200_GPU_CHECK") != NULL;
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

  int n = ls.local_nodes;
  double *field = D4EST_ALLOC(double, n);   /* d4est_amr_interpolate_field reallocates it */
  for (int i = 0; i < n; i++) field[i] = sin(0.37 * i) + 0.25 * cos(0.011 * i) + 1e-3 * i / (double)n;

  FILE *f = fix_open(out);
  if (!f) return 3;
  int32_t st32 = steps;
  fix_write(f, "steps", FIX_I32, 1, &st32);
  d4est_amr_t *a = d4est_amr_init_random_hp(p4est, steps);
  char nm[64];
  double worst = 0.;
  for (int k = 0; k < steps; k++) {
    double *pre = P4EST_ALLOC(double, n);
    memcpy(pre, field, (size_t)n * sizeof(double));
    const int n_pre = n;
    d4est_amr_step(p4est, ops, a, &field, NULL, NULL, NULL);
    ls = update_mesh(p4est, &ghost, ops, geom, quad, factors, ext, 0);
    n = ls.local_nodes;
    snprintf(nm, sizeof nm, "S%d.initial_log", k); fix_write(f, nm, FIX_I32, a->initial_log_size, a->initial_log);
    snprintf(nm, sizeof nm, "S%d.refinement_log", k); fix_write(f, nm, FIX_I32, a->initial_log_size, a->refinement_log);
    snprintf(nm, sizeof nm, "S%d.balance_log", k); fix_write(f, nm, FIX_I32, a->balance_log_size, a->balance_log);
    snprintf(nm, sizeof nm, "S%d.field_pre", k); fix_write(f, nm, FIX_F64, n_pre, pre);
    snprintf(nm, sizeof nm, "S%d.field_post", k); fix_write(f, nm, FIX_F64, n, field);
    if (gpu_check) {   /* the glue entry a d4est build would call instead of d4est_amr_interpolate_field */
      double *dev_field = D4EST_ALLOC(double, n_pre);
      memcpy(dev_field, pre, (size_t)n_pre * sizeof(double));
      if (d4est_b200_amr_interpolate_field(a, &dev_field, n) != 0) { fprintf(stderr, "gpu check: %s\n", d4b_last_error()); return 4; }
      double num = 0., den = 0.;
      for (int i = 0; i < n; i++) { num += (dev_field[i] - field[i]) * (dev_field[i] - field[i]); den += field[i] * field[i]; }
      const double rel = sqrt(num / den);
      if (rel > worst) worst = rel;
      D4EST_FREE(dev_field);
    }
    P4EST_FREE(pre);
    printf("{\"step\": %d, \"elements_before\": %d, \"elements_after\": %d, \"nodes_before\": %d, \"nodes_after\": %d}\n",
           k, a->initial_log_size, (int)p4est->local_num_quadrants, n_pre, n);
  }
  if (gpu_check) printf("{\"gpu_check_amr_transfer_rel_l2\": %.3e}\n", worst);
  fclose(f);
  d4est_amr_destroy(a);
  D4EST_FREE(field);
  return (gpu_check && !(worst < 1e-13)) ? 5 : 0;
}
