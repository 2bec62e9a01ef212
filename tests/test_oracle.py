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


def test_product_1d_tables_match_the_reference_tables(golden_dir):
    """d4est_b200/csrc/d4b_ops.c (closed forms) against the UNMODIFIED reference's own operator tables
    (d4est_operators_fetch_* / apply_* on unit vectors, dumped by oracle/_ref/d4b_ops_dump into
    tests/golden/ref_ops_tables.fix): GLL and Gauss rules, D, GLL->Gauss interpolation, p- and hp-prolongation."""
    import ctypes
    import d4est_b200
    fx = fixture.load(os.path.join(golden_dir, "ref_ops_tables.fix"))
    L = ctypes.CDLL(d4est_b200.LIB_PATH)
    P = ctypes.c_void_p
    for deg in range(1, 9):
        N = deg + 1
        D = np.zeros((N, N)); L.d4b_ops_diff(deg, D.ctypes.data_as(P))
        assert np.abs(D.ravel() - fx[f"D{deg}"]).max() < 1e-13 * np.abs(D).max()
        assert np.abs(dg_oracle.diff_matrix(deg).ravel() - fx[f"D{deg}"]).max() < 1e-12 * np.abs(D).max()
        x = np.zeros(N); w = np.zeros(N)
        L.d4b_ops_gll(deg, x.ctypes.data_as(P), w.ctypes.data_as(P))
        assert np.abs(x - fx[f"gll_x{deg}"]).max() < 1e-15 and np.abs(w - fx[f"gll_w{deg}"]).max() < 1e-15
        L.d4b_ops_gauss(N, x.ctypes.data_as(P), w.ctypes.data_as(P))
        assert np.abs(x - fx[f"gauss_x{deg}"]).max() < 1e-15 and np.abs(w - fx[f"gauss_w{deg}"]).max() < 1e-15
        for degq in range(deg, min(deg + 3, 10) + 1):
            I = np.zeros((degq + 1, N)); L.d4b_ops_interp(deg, degq, 0, -1, I.ctypes.data_as(P))
            assert np.abs(I.ravel() - fx[f"I{deg}_{degq}"]).max() < 1e-14
            assert np.abs(dg_oracle.interp(deg, degq, 0).ravel() - fx[f"I{deg}_{degq}"]).max() < 1e-13
        for degh in range(deg, 9):
            Nh = degh + 1
            Pm = np.zeros((Nh, N)); L.d4b_ops_prolong(deg, degh, -1, Pm.ctypes.data_as(P))
            assert np.abs(Pm.ravel() - fx[f"P{deg}_{degh}"]).max() < 1e-14
            for c in (0, 1):
                L.d4b_ops_prolong(deg, degh, c, Pm.ctypes.data_as(P))
                assert np.abs(Pm.ravel() - fx[f"HP{deg}_{degh}"].reshape(2, Nh * N)[c]).max() < 1e-14


def _mg_fixture(ref_fixture_dir):
    p = os.path.join(ref_fixture_dir, "mg_cubed_sphere_hp1.fix")
    if not os.path.exists(p):
        pytest.skip("multigrid fixture not built (needs /root/reference at build time)")
    return fixture.load(p)


def test_transfer_oracle_matches_reference_prolongation(ref_fixture_dir):
    """solver_oracle.transfer over the hierarchy captured from the reference's own coarsening against the
    reference's d4est_operators_apply_{p,hp}_prolong[_transpose] on probe vectors (hp case: 8 children of mixed
    degree, p case: same cell, copies of families that 2:1 balance split again)."""
    import solver_oracle as so
    fx = _mg_fixture(ref_fixture_dir)
    nl = int(fx["mg_meta"][0])
    for l in range(nl - 1):
        T = fixture.sub(fx, f"T{l}.")
        r = so.transfer(T, T["probe_fine"], T["probe_coarse"].size, True)
        assert np.linalg.norm(r - T["ref_restricted"]) < 1e-13 * np.linalg.norm(T["ref_restricted"])
        p = so.transfer(T, T["probe_coarse"], T["probe_fine"].size, False)
        assert np.linalg.norm(p - T["ref_prolonged"]) < 1e-13 * np.linalg.norm(T["ref_prolonged"])


def test_krylov_oracles_match_reference_drivers(ref_fixture_dir):
    """solver_oracle.fcg_solve / cg_solve against d4est_solver_fcg_solve / d4est_solver_cg_solve run by the
    reference itself for exactly 25 iterations (no preconditioner) on the curved hp mesh."""
    import solver_oracle as so
    fx = _mg_fixture(ref_fixture_dir)
    nl = int(fx["mg_meta"][0])
    top = fixture.sub(fx, f"L{nl - 1}.")
    A = lambda v: dg_oracle.apply(top, v)
    rhs = fx["mg_rhs"]
    u, k = so.fcg_solve(A, np.zeros(rhs.size), rhs, 25, 0., 1e-300)
    assert np.linalg.norm(u - fx["krylov_fcg25_u"]) < 1e-9 * np.linalg.norm(fx["krylov_fcg25_u"])
    u, k, _ = so.cg_solve(A, np.zeros(rhs.size), rhs, 25, 0., 1e-300)
    assert k == 25 and np.linalg.norm(u - fx["krylov_cg25_u"]) < 1e-9 * np.linalg.norm(fx["krylov_cg25_u"])


def test_vcycle_oracle_matches_reference_preconditioner(ref_fixture_dir):
    """solver_oracle.mg_pc (V-cycle with Chebyshev smoother + cg_eigs + Chebyshev bottom solver on the captured
    hierarchy) against the reference's own d4est_krylov_pc_multigrid_apply: first preconditioner application of
    its MG-FCG solve."""
    import solver_oracle as so
    fx = _mg_fixture(ref_fixture_dir)
    nl = int(fx["mg_meta"][0])
    lv = [fixture.sub(fx, f"L{l}.") for l in range(nl)]
    A = [(lambda v, t=t: dg_oracle.apply(t, v)) for t in lv]
    tabs = [fixture.sub(fx, f"T{l}.") for l in range(nl - 1)]
    sizes = [int(t["node_off"][int(t["hdr"][0])]) for t in lv]
    z = so.mg_pc(A, tabs, sizes, fx["mg_r0"], so.mg_params(fx["mg_params"]))
    assert np.linalg.norm(z - fx["mg_z0"]) < 1e-9 * np.linalg.norm(fx["mg_z0"])


def test_reference_golden_numbers_are_reproduced_by_the_oracle_build(ref_fixture_dir):
    """The numbers the reference's own unit tests print / assert, reproduced by oracle/_ref (the unmodified reference
    compiled here): level-1 Au sum of d4est_test_laplacian_speedup (938.36171996774783 baseline, ...774669 with_opt,
    SURVEY 8c) and Au[33010] = -0.0034181705065613 of d4est_test_laplacian_consistency.c:415-426."""
    import json
    p2 = os.path.join(ref_fixture_dir, "cubed_sphere_hp2.json")
    pw = os.path.join(ref_fixture_dir, "wedge_consistency.json")
    if not (os.path.exists(p2) and os.path.exists(pw)):
        pytest.skip("reference fixtures not built")
    j2 = json.load(open(p2)); jw = json.load(open(pw))
    assert j2["elements"] == 1294 and j2["nodes"] == 35558
    assert abs(j2["Au_sum_baseline"] - 938.36171996774783) < 1e-9 and abs(j2["Au_sum_with_opt"] - 938.36171996774669) < 1e-9
    assert abs(jw["Au_baseline_33010"] - (-0.0034181705065613010)) < 1e-12     # the reference test itself allows 1e-3
    fw = fixture.load(os.path.join(ref_fixture_dir, "wedge_consistency.fix"))
    assert fw["Au_ref_baseline"][33010] == jw["Au_baseline_33010"]


def test_mass_term_oracle_matches_reference(ref_fixture_dir):
    """dg_oracle.mass_apply against the reference's d4est_quadrature_apply_fofufofvlilj run element by element on the
    curved hp mesh (mass-type term of apply_jac)."""
    p = os.path.join(ref_fixture_dir, "cubed_sphere_hp2.fix")
    if not os.path.exists(p):
        pytest.skip("reference fixture not built")
    fx = fixture.load(p)
    if "mass_ref" not in fx:
        pytest.skip("fixture predates the mass-term reference")
    assert rel(dg_oracle.mass_apply(fx, fx["mass_coef"], fx["u"]), fx["mass_ref"]) < TOL


def test_amr_field_transfer_oracle_matches_reference(golden_dir):
    """d4est_amr_step with a field attached (unmodified reference, oracle/harness/d4b_amr_harness.c): the numpy
    restatement of d4est_amr_interpolate_field reproduces the reference's field after every step."""
    import solver_oracle as so
    fx = fixture.load(os.path.join(golden_dir, "amr_transfer_cubed_sphere_hp2.fix"))
    steps = int(fx["steps"][0])
    assert steps == 2
    kinds = set()
    for k in range(steps):
        il, rl, bl = fx[f"S{k}.initial_log"], fx[f"S{k}.refinement_log"], fx[f"S{k}.balance_log"]
        kinds |= {("h" if c < 0 else ("p" if c > d else "keep")) for d, c in zip(il, rl)}
        kinds |= {"balance_h"} if (bl < 0).any() else set()
        out = so.amr_interpolate_field(il, rl, bl, fx[f"S{k}.field_pre"])
        ref = fx[f"S{k}.field_post"]
        assert out.shape == ref.shape
        assert np.abs(out - ref).max() <= 1e-13 * np.abs(ref).max()
        if k + 1 < steps:   # the reference chains the steps: its output is the next step's input, bit for bit
            assert np.array_equal(ref, fx[f"S{k + 1}.field_pre"])
    assert {"h", "p", "balance_h"} <= kinds   # the fixture exercises h-, p-refinement and the 2:1 balance sweep


@pytest.mark.parametrize("name", ["cubed_sphere_hp1", "brick_p3_l2"])
def test_twin_trace_formulation_equals_the_reference(golden_dir, name):
    """The identity the two-pass device kernels rest on, on the CPU: taking the neighbour side of every interior
    half-mortar from the own-side trace of the neighbour's twin record (u+ = W[twin].u, sum a+ du+ = -W[twin].q)
    reproduces the reference's Au; every interior record of a one-rank mesh has a twin, hanging and mixed-p included."""
    fx = fixture.load(os.path.join(golden_dir, name + ".fix"))
    nl = int(fx["hdr"][0])
    twin = dg_oracle.twin_records(fx)
    interior = np.array([0 <= int(h["nbr"]) < nl for h in fx["hm"][:int(fx["hm_begin"][nl])]])
    assert (twin[interior] >= 0).all() and (twin[~interior] < 0).all()
    assert (twin[twin[interior]] == np.nonzero(interior)[0]).all()      # an involution
    Au = dg_oracle.apply(fx, fx["u"], twin_traces=True)
    ref = fx["Au_ref_with_opt"]
    assert np.linalg.norm(Au - ref) / np.linalg.norm(ref) < 1e-12
    assert np.linalg.norm(Au - dg_oracle.apply(fx, fx["u"])) / np.linalg.norm(ref) < 1e-13
