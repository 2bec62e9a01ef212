#!/bin/sh
# Regenerates the committed golden fixtures by running the UNMODIFIED reference
# (oracle/_ref/d4b_harness, built by oracle/ref_build/Makefile from /root/reference).
# Each .fix holds the flat mesh tables, the input u and the reference's own Au
# (d4est_laplacian_apply_aij and d4est_laplacian_with_opt_apply_aij).
set -e
cd "$(dirname "$0")/../.."
H=oracle/_ref/d4b_harness
$H oracle/inputs/brick_p3_l2.input none 0 zero sinrand tests/golden/brick_p3_l2.fix
$H oracle/inputs/cubed_sphere_hp.input random_hp 1 poly polyrand tests/golden/cubed_sphere_hp1.fix
oracle/_ref/d4b_amr_harness oracle/inputs/cubed_sphere_hp.input 2 tests/golden/amr_transfer_cubed_sphere_hp2.fix
