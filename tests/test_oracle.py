"""Pins the numpy oracle (oracle/dg_oracle.py) against the UNMODIFIED reference:
Au written by oracle/_ref/d4b_harness (tests/golden/make_golden.sh) and the golden sums the
reference's own unit test prints."""
import os
import numpy as np
import pytest
import fixture
import dg_oracle

TOL = 1e-12  # north_star: relative L2 per matvec


def rel(a, b):
    return np.linalg.norm(a - b) / np.linalg.norm(b)


@pytest.mark.parametrize("name", ["brick_p3_l2", "cubed_sphere_hp1"])
def test_oracle_matches_reference(golden_dir, name):
    fx = fixture.load(os.path.join(golden_dir, name + ".fix"))
    Au = dg_oracle.apply(fx, fx["u"])
    assert rel(Au, fx["Au_ref_with_opt"]) < TOL
    assert rel(Au, fx["Au_ref_baseline"]) < TOL


def test_reference_variants_agree(golden_dir):
    # Tests/Unit/d4est_test_laplacian_speedup.c:485: |baseline - with_opt| <= 1e-13 per node
    fx = fixture.load(os.path.join(golden_dir, "cubed_sphere_hp1.fix"))
    assert np.abs(fx["Au_ref_baseline"] - fx["Au_ref_with_opt"]).max() < 1e-13


def test_golden_sum_speedup_test(golden_dir):
    # level-0 value printed by the reference's d4est_test_laplacian_speedup
    fx = fixture.load(os.path.join(golden_dir, "cubed_sphere_hp1.fix"))
    assert int(fx["hdr"][0]) == 300 and fx["u"].size == 5807
    assert abs(fx["Au_ref_baseline"].sum() - 9.04698104235131) < 1e-10
    assert abs(dg_oracle.apply(fx, fx["u"]).sum() - 9.04698104235131) < 1e-10


def test_fixture_covers_hard_cases(golden_dir):
    fx = fixture.load(os.path.join(golden_dir, "cubed_sphere_hp1.fix"))
    hm = fx["hm"]
    assert (hm["nbr"] < 0).any() and (hm["g_off"] >= 0).any()      # inhomogeneous Dirichlet
    assert (hm["child_m"] >= 0).any() and (hm["child_p"] >= 0).any()  # hanging, both roles
    assert len(np.unique(hm["tcode"])) > 2                          # inter-tree orientations
    assert len(np.unique(fx["deg"])) > 1                            # mixed p
    assert (fx["degq"] > fx["deg"][: fx["degq"].size]).all()         # over-integration


def test_operator_symmetry_homogeneous(golden_dir):
    # Tests/Unit/d4est_test_laplacian_symmetry.c: A = A^T for zero Dirichlet data
    fx = fixture.load(os.path.join(golden_dir, "brick_p3_l2.fix"))
    rng = np.random.default_rng(0)
    x, y = rng.standard_normal(fx["u"].size), rng.standard_normal(fx["u"].size)
    a, b = y @ dg_oracle.apply(fx, x), x @ dg_oracle.apply(fx, y)
    assert abs(a - b) < 1e-11 * max(abs(a), 1.0)


@pytest.mark.parametrize("name", ["brick_p3_l2", "cubed_sphere_hp1"])
def test_solver_oracle_matches_reference_cg_eigs_and_cheby(golden_dir, name):
    """cg_eigs (Solver/d4est_solver_cg_eigs.c:115) and the Chebyshev iteration
    (Solver/d4est_solver_multigrid_smoother_cheby.c:81) run by the UNMODIFIED reference in the harness."""
    import solver_oracle as so
    fx = fixture.load(os.path.join(golden_dir, name + ".fix"))
    A = lambda v: dg_oracle.apply(fx, v)
    rhs = fx["Au_ref_with_opt"]
    eig, u1 = so.cg_eigs(A, np.zeros_like(rhs), rhs, 10)
    assert abs(eig - fx["sol_eig"][0]) < 1e-10 * abs(eig)
    assert rel(u1, fx["sol_u_after_eigs"]) < 1e-10
    u2, r2 = so.cheby_iterate(A, fx["sol_u_after_eigs"].copy(), rhs, 8, fx["sol_eig"][0] / 30., fx["sol_eig"][0])
    assert rel(u2, fx["sol_u_after_cheby"]) < 1e-11
    assert rel(r2, fx["sol_r_after_cheby"]) < 1e-10
