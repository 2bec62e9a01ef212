"""Parity of the CUDA path, called through the C ABI, against (a) Au computed by the UNMODIFIED
reference on the same mesh/input and (b) the numpy oracle.  Tolerance: relative L2 <= 1e-12 per
matvec (BASELINE.json north_star)."""
import glob
import os
import numpy as np
import pytest
import d4est_b200
import fixture
import dg_oracle

pytestmark = pytest.mark.gpu
TOL = 1e-12
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def rel(a, b):
    return np.linalg.norm(a - b) / np.linalg.norm(b)


def all_fixtures():
    fs = sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "*.fix")))
    fs += sorted(glob.glob(os.path.join(ROOT, "oracle", "_ref", "fixtures", "*.fix")))
    # single-level apply fixtures only (not the 1-D table dump, not the multigrid fixtures)
    return [f for f in fs if not os.path.basename(f).startswith(("ref_ops", "mg_", "amr_"))]


@pytest.mark.parametrize("path", all_fixtures(), ids=lambda p: os.path.basename(p)[:-4])
def test_apply_matches_reference(path):
    fx = fixture.load(path)
    lvl = d4est_b200.MeshLevel(fx)
    Au = lvl.apply_host(fx["u"])
    assert np.isfinite(Au).all()
    assert rel(Au, fx["Au_ref_with_opt"]) < TOL
    assert rel(Au, fx["Au_ref_baseline"]) < TOL
    lvl.close()


@pytest.mark.parametrize("path", all_fixtures(), ids=lambda p: os.path.basename(p)[:-4])
def test_hp_kernel_matches_reference_on_every_element_class(path):
    """The templated hanging / mixed-p kernel (d4b_apply_hp.cu) run on EVERY element of every fixture
    (D4B_ALL_HP=1: regular, mixed-p, hanging in both roles, boundary), and the default split (regular elements in
    the register-blocked / line-task kernels, the others in the hp kernel) against the completeness kernel."""
    fx = fixture.load(path)
    out, split = {}, {}
    L = d4est_b200.lib()
    for key, env in (("default", {}), ("all_hp", {"D4B_ALL_HP": "1"}), ("no_hp", {"D4B_NO_HP": "1"}),
                     ("two_pass", {}), ("all_hp_two_pass", {"D4B_ALL_HP": "1"})):
        os.environ.update(env)
        # two-pass (twin-trace) hp kernels on (the default) or off (single-pass kernel for every hp element)
        d4est_b200.capi.check(L.d4b_set_option(b"two_pass_min_elements", 0 if "two_pass" in key else 1 << 30))
        try:
            lvl = d4est_b200.MeshLevel(fx)
            split[key] = lvl.kernel_split
            out[key] = lvl.apply_host(fx["u"])
            if key == "all_hp":
                assert np.array_equal(out[key], lvl.apply_host(fx["u"]))   # fixed accumulation order
            lvl.close()
        finally:
            d4est_b200.capi.check(L.d4b_set_option(b"two_pass_min_elements", -1))   # back to the default rule
            for k in env:
                del os.environ[k]
    nl = int(fx["hdr"][0])
    assert split["all_hp_two_pass"][1] == nl
    assert rel(out["all_hp_two_pass"], out["all_hp"]) < 1e-13   # a+ = -a-(twin): equal to rounding, not bitwise
    assert split["all_hp"][0] == 0 and split["no_hp"][1] == 0
    assert split["all_hp"][1] == nl, split        # every (N, Nq) of the fixtures has an instance
    assert split["default"][2] == 0, split        # nothing is left to the completeness kernel
    for key in out:
        assert rel(out[key], fx["Au_ref_with_opt"]) < TOL, (key, split[key])
    assert rel(out["all_hp"], out["no_hp"]) < 1e-13


@pytest.mark.parametrize("name", ["brick_p3_l2", "cubed_sphere_hp1"])
def test_apply_matches_oracle_random_input(golden_dir, name):
    fx = fixture.load(os.path.join(golden_dir, name + ".fix"))
    lvl = d4est_b200.MeshLevel(fx)
    u = np.random.default_rng(7).uniform(-1, 1, fx["u"].size)
    assert rel(lvl.apply_host(u), dg_oracle.apply(fx, u)) < TOL


def test_linearity_and_symmetry_homogeneous(golden_dir):
    fx = fixture.load(os.path.join(golden_dir, "brick_p3_l2.fix"))  # zero Dirichlet data -> linear, A = A^T
    lvl = d4est_b200.MeshLevel(fx)
    rng = np.random.default_rng(1)
    x, y = rng.standard_normal(fx["u"].size), rng.standard_normal(fx["u"].size)
    Ax, Ay = lvl.apply_host(x), lvl.apply_host(y)
    assert rel(lvl.apply_host(2.0 * x - 3.0 * y), 2.0 * Ax - 3.0 * Ay) < 1e-13
    assert abs(y @ Ax - x @ Ay) < 1e-11 * abs(y @ Ax)
    assert x @ Ax > 0  # SPD


def test_repeatable_bitwise(golden_dir):
    fx = fixture.load(os.path.join(golden_dir, "cubed_sphere_hp1.fix"))
    lvl = d4est_b200.MeshLevel(fx)
    a, b = lvl.apply_host(fx["u"]), lvl.apply_host(fx["u"])
    assert np.array_equal(a, b)  # no atomics anywhere on the path
    assert d4est_b200.lib().d4b_kernel_launches() > 0


@pytest.mark.parametrize("name", ["brick_p3_l3_poly", "cubed_sphere_p3_uniform"])
def test_register_blocked_kernel_equals_line_kernel_and_generic(name, ref_fixture_dir):
    """The three element kernels (v3 register-blocked, v2 line tasks, generic) are interchangeable on regular
    elements: same Au to round-off on a brick with inhomogeneous Dirichlet data and on the curved cubed sphere with
    uniform p=3 (all 832 elements regular, five inter-tree orientation codes)."""
    p = os.path.join(ref_fixture_dir, name + ".fix")
    if not os.path.exists(p):
        pytest.skip("reference fixture not built")
    fx = fixture.load(p)
    out = {}
    for key, env in (("v3", {}), ("v3_three_rounds", {"D4B_V3_NO_SHARE": "1"}), ("v2", {"D4B_NO_V3": "1"}),
                     ("generic", {"D4B_NO_FAST": "1"})):
        for k, v in env.items():
            os.environ[k] = v
        try:
            n0 = d4est_b200.lib().d4b_kernel_launches()
            lvl = d4est_b200.MeshLevel(fx)
            out[key] = lvl.apply_host(fx["u"])
            lvl.close()
        finally:
            for k in env:
                del os.environ[k]
    for key in out:
        assert rel(out[key], fx["Au_ref_with_opt"]) < TOL, key
    assert rel(out["v3"], out["v2"]) < 1e-13 and rel(out["v3"], out["generic"]) < 1e-13
    assert rel(out["v3"], out["v3_three_rounds"]) < 1e-13


@pytest.mark.parametrize("dims,deg", [((5, 3, 7), 3), ((9, 9, 9), 3), ((17, 4, 3), 1), ((6, 10, 2), 1), ((2, 2, 2), 3),
                                      ((7, 5, 3), 2), ((8, 8, 8), 2)])
def test_register_blocked_kernel_on_odd_bricks(dims, deg):
    """Bricks whose element count is not a multiple of the warp batch and whose Morton families are incomplete
    (odd extents): the shuffle exchange / one-round staging plan, the three-round plan and the line-task kernel
    agree to round-off, and the result is bitwise repeatable."""
    out = {}
    u = None
    for key, env in (("v3", {}), ("v3_three_rounds", {"D4B_V3_NO_SHARE": "1"}), ("v2", {"D4B_NO_V3": "1"})):
        for k, v in env.items():
            os.environ[k] = v
        try:
            lvl = d4est_b200.MeshLevel.brick(dims, deg=deg)
            if u is None:
                u = np.random.default_rng(5).uniform(-1, 1, lvl.local_nodes)
            out[key] = lvl.apply_host(u)
            if key == "v3":
                assert np.array_equal(out[key], lvl.apply_host(u))
            lvl.close()
        finally:
            for k in env:
                del os.environ[k]
    assert np.isfinite(out["v3"]).all() and np.abs(out["v3"]).max() > 0
    assert rel(out["v3"], out["v2"]) < 1e-13 and rel(out["v3"], out["v3_three_rounds"]) < 1e-13


def test_back_to_back_device_applies_on_a_multi_bucket_mesh(golden_dir):
    """Many asynchronous applies in a row on a mesh with several degree buckets (each launch re-points the
    constant-memory tables): the last result equals the synchronous host-path result bit for bit."""
    from d4est_b200.capi import DevVec
    fx = fixture.load(os.path.join(golden_dir, "cubed_sphere_hp1.fix"))
    lvl = d4est_b200.MeshLevel(fx)
    ref = lvl.apply_host(fx["u"])
    U, V = DevVec(lvl.local_nodes, fx["u"]), DevVec(lvl.local_nodes)
    for _ in range(60):
        lvl.apply_dev(U.ptr.value, V.ptr.value, 0)
    assert np.array_equal(V.get(), ref)


def test_reference_golden_numbers_on_the_device(ref_fixture_dir):
    """The golden numbers of the reference's own unit tests, from the CUDA path: Au sum 938.3617199677 on the level-1
    mesh of d4est_test_laplacian_speedup, Au[33010] = -0.0034181705065613 of d4est_test_laplacian_consistency."""
    p2, pw = os.path.join(ref_fixture_dir, "cubed_sphere_hp2.fix"), os.path.join(ref_fixture_dir, "wedge_consistency.fix")
    if not (os.path.exists(p2) and os.path.exists(pw)):
        pytest.skip("reference fixtures not built")
    fx = fixture.load(p2)
    Au = d4est_b200.MeshLevel(fx).apply_host(fx["u"])
    assert abs(Au.sum() - 938.36171996774) < 1e-8
    fw = fixture.load(pw)
    Aw = d4est_b200.MeshLevel(fw).apply_host(fw["u"])
    assert abs(Aw[33010] - (-0.0034181705065613010)) < 1e-12


@pytest.mark.parametrize("name", ["cubed_sphere_hp2", "cubed_sphere_hp_p7", "brick_p3_l3"])
def test_mass_term_matches_reference(name, ref_fixture_dir):
    """d4b_mass_apply (I^T diag(w J f(x, u0)) I v, coefficient resident) against the reference's
    d4est_quadrature_apply_fofufofvlilj run element by element; and A u + M u in one call (d4b_apply_jac_dev)."""
    from d4est_b200.capi import DevVec, lib, check
    p = os.path.join(ref_fixture_dir, name + ".fix")
    if not os.path.exists(p):
        pytest.skip("reference fixture not built")
    fx = fixture.load(p)
    if "mass_ref" not in fx:
        pytest.skip("fixture predates the mass-term reference")
    lvl = d4est_b200.MeshLevel(fx)
    coef = np.ascontiguousarray(fx["mass_coef"])
    check(lib().d4b_mass_set_coef(lvl.handle, coef.ctypes.data_as(ctypes_voidp())))
    out = np.zeros_like(fx["u"]); u = np.ascontiguousarray(fx["u"])
    check(lib().d4b_mass_apply_host(lvl.handle, u.ctypes.data_as(ctypes_voidp()), out.ctypes.data_as(ctypes_voidp()), 0))
    assert rel(out, fx["mass_ref"]) < TOL
    U, V = DevVec(lvl.local_nodes, u), DevVec(lvl.local_nodes)
    check(lib().d4b_apply_jac_dev(lvl.handle, U.ptr, V.ptr, None))
    assert rel(V.get(), fx["Au_ref_with_opt"] + fx["mass_ref"]) < TOL
    lvl.close()


def ctypes_voidp():
    import ctypes
    return ctypes.c_void_p
