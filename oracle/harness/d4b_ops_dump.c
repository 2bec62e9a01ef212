/* d4b_ops_dump.c -- TEST INFRASTRUCTURE ONLY.  Dumps the UNMODIFIED reference's 1-D operator tables
 * (dGMath/d4est_operators.c: GLL nodes/weights :727-829, D = V_r V^-1 :855, GLL->Gauss interpolation :411,
 * p-prolongation :995, hp-prolongation :944-993) into a small fixture, so that the product's closed-form tables
 * (d4est_b200/csrc/d4b_ops.c) are pinned to the reference itself (tests/test_oracle.py).
 * Tables without a public fetch function are obtained column by column by applying the public
 * d4est_operators_apply_* functions to unit vectors in 1-D.
 * usage: d4b_ops_dump <out.fix> */
#include <pXest.h>
#include <d4est_operators.h>
#include <d4est_lgl.h>
#include "fixture_io.h"

int main(int argc, char **argv)
{
  if (argc < 2) return 2;
  d4est_operators_t *ops = d4est_ops_init(20);
  FILE *f = fix_open(argv[1]);
  char nm[32];
  double buf[32 * 32], col[64], e[32];
  for (int deg = 1; deg <= 8; deg++) {
    int N = deg + 1;
    snprintf(nm, sizeof nm, "gll_x%d", deg); fix_write(f, nm, FIX_F64, N, d4est_operators_fetch_lobatto_nodes_1d(ops, deg));
    snprintf(nm, sizeof nm, "gll_w%d", deg); fix_write(f, nm, FIX_F64, N, d4est_operators_fetch_lobatto_weights_1d(ops, deg));
    snprintf(nm, sizeof nm, "gauss_x%d", deg); fix_write(f, nm, FIX_F64, N, d4est_operators_fetch_gauss_nodes_1d(ops, deg));
    snprintf(nm, sizeof nm, "gauss_w%d", deg); fix_write(f, nm, FIX_F64, N, d4est_operators_fetch_gauss_weights_1d(ops, deg));
    for (int j = 0; j < N; j++) {   /* D[i*N + j] = l_j'(x_i) */
      for (int k = 0; k < N; k++) e[k] = (k == j);
      d4est_operators_apply_dij(ops, e, 1, deg, 0, col);
      for (int i = 0; i < N; i++) buf[i * N + j] = col[i];
    }
    snprintf(nm, sizeof nm, "D%d", deg); fix_write(f, nm, FIX_F64, N * N, buf);
    for (int degq = deg; degq <= deg + 3 && degq <= 10; degq++) {
      snprintf(nm, sizeof nm, "I%d_%d", deg, degq);
      fix_write(f, nm, FIX_F64, (degq + 1) * N, d4est_operators_fetch_lobatto_to_gauss_interp_1d(ops, deg, degq));
    }
    for (int degh = deg; degh <= 8; degh++) {   /* prolongation degH = deg -> degh */
      int Nh = degh + 1;
      for (int j = 0; j < N; j++) {
        for (int k = 0; k < N; k++) e[k] = (k == j);
        d4est_operators_apply_p_prolong(ops, e, deg, 1, degh, col);
        for (int i = 0; i < Nh; i++) buf[i * N + j] = col[i];
      }
      snprintf(nm, sizeof nm, "P%d_%d", deg, degh); fix_write(f, nm, FIX_F64, Nh * N, buf);
      int dh2[2] = { degh, degh };
      for (int j = 0; j < N; j++) {   /* two stacked children: rows 0..Nh-1 lower half, Nh..2Nh-1 upper half */
        for (int k = 0; k < N; k++) e[k] = (k == j);
        d4est_operators_apply_hp_prolong(ops, e, deg, 1, dh2, col);
        for (int i = 0; i < 2 * Nh; i++) buf[i * N + j] = col[i];
      }
      snprintf(nm, sizeof nm, "HP%d_%d", deg, degh); fix_write(f, nm, FIX_F64, 2 * Nh * N, buf);
    }
  }
  fclose(f);
  return 0;
}
