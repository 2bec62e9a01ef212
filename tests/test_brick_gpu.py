"""The device-side synthetic brick generator (d4b_brick_create, the bench workload) must give the
same operator as the tables flattened from the reference's own brick mesh: apply it to the
fixture's u and compare with the reference's Au."""
import os
import numpy as np
import pytest
import d4est_b200
import fixture

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CASES = [("tests/golden/brick_p3_l2.fix", 4, 3), ("oracle/_ref/fixtures/brick_p3_l3.fix", 8, 3),
         ("oracle/_ref/fixtures/brick_p3_l4.fix", 16, 3), ("oracle/_ref/fixtures/brick_p5_l2.fix", 4, 5)]


@pytest.mark.parametrize("path,n,deg", CASES, ids=[c[0].split("/")[-1][:-4] for c in CASES])
def test_brick_generator_matches_reference(path, n, deg):
    p = os.path.join(ROOT, path)
    if not os.path.exists(p):
        pytest.skip("reference-made fixture not built")
    fx = fixture.load(p)
    if (fx["hm"]["g_off"] >= 0).any():   # fixture with inhomogeneous data: compare the linear part
        lvl_f = d4est_b200.MeshLevel(fx)
        zero = lvl_f.apply_host(np.zeros_like(fx["u"]))
        ref = fx["Au_ref_with_opt"] - zero
    else:
        ref = fx["Au_ref_with_opt"]
    lvl = d4est_b200.MeshLevel.brick(n, deg=deg)
    assert lvl.local_nodes == fx["u"].size
    Au = lvl.apply_host(fx["u"])
    assert np.linalg.norm(Au - ref) / np.linalg.norm(ref) < 1e-12


def test_fast_and_generic_kernels_agree(golden_dir):
    """D4B_NO_FAST=1 routes every element through the generic kernel."""
    fx = fixture.load(os.path.join(golden_dir, "cubed_sphere_hp1.fix"))
    a = d4est_b200.MeshLevel(fx).apply_host(fx["u"])
    os.environ["D4B_NO_FAST"] = "1"
    try:
        b = d4est_b200.MeshLevel(fx).apply_host(fx["u"])
    finally:
        del os.environ["D4B_NO_FAST"]
    assert np.linalg.norm(a - b) / np.linalg.norm(b) < 1e-13


def test_non_power_of_two_brick_is_spd_and_symmetric():
    lvl = d4est_b200.MeshLevel.brick((5, 3, 6), deg=2, degq=3)
    rng = np.random.default_rng(3)
    x, y = rng.standard_normal(lvl.local_nodes), rng.standard_normal(lvl.local_nodes)
    Ax, Ay = lvl.apply_host(x), lvl.apply_host(y)
    assert abs(y @ Ax - x @ Ay) < 1e-11 * abs(y @ Ax) and x @ Ax > 0


def test_host_pipeline_equals_plain_host_path():
    """d4b_apply_host on a uniform level runs as a 3-stream pipeline over element ranges (chunked H2D, kernel per
    range once the chunks it reads have arrived, D2H per finished range); option no_pipeline is the plain
    H2D -> kernel -> D2H path.  Same bits."""
    import os
    lvl = d4est_b200.MeshLevel.brick((24, 20, 28), deg=3)   # 13 440 elements: above the pipeline threshold
    u = np.random.default_rng(3).uniform(-1, 1, lvl.local_nodes)
    a = lvl.apply_host(u)
    from d4est_b200 import capi
    capi.check(capi.lib().d4b_set_option(b"no_pipeline", 1))
    try:
        b = lvl.apply_host(u)
    finally:
        capi.check(capi.lib().d4b_set_option(b"no_pipeline", 0))
    assert np.array_equal(a, b) and np.isfinite(a).all() and np.abs(a).max() > 0


def test_host_pipeline_late_pass_with_pinned_buffers():
    """With pinned (device-mapped) caller buffers the pipeline defers the elements that read a neighbour more than
    one chunk ahead to a last launch that stores straight into the caller's Au (d4b_apply_host); option no_late_pass
    makes every range wait for the last chunk it reads instead.  The deferred elements go through the kernel's
    three-round staging plan (a+ of their own record instead of -a- of the neighbour): equal to rounding."""
    import ctypes
    import os
    from d4est_b200 import capi
    L = capi.lib()
    lvl = d4est_b200.MeshLevel.brick((40, 36, 44), deg=3)   # 63 360 elements, 16 ranges, Morton seams inside
    n = lvl.local_nodes
    hp, hq = ctypes.c_void_p(), ctypes.c_void_p()
    capi.check(L.d4b_host_alloc_pinned(n * 8, ctypes.byref(hp)))
    capi.check(L.d4b_host_alloc_pinned(n * 8, ctypes.byref(hq)))
    try:
        hu = np.ctypeslib.as_array(ctypes.cast(hp, ctypes.POINTER(ctypes.c_double)), shape=(n,))
        hAu = np.ctypeslib.as_array(ctypes.cast(hq, ctypes.POINTER(ctypes.c_double)), shape=(n,))
        hu[:] = np.random.default_rng(11).uniform(-1, 1, n)
        for _ in range(3):                       # repeated: stale data of the previous apply must not leak
            hAu[:] = np.nan
            lvl.apply_host(hu, out=hAu)
            a = hAu.copy()
            assert np.isfinite(a).all()
        capi.check(L.d4b_set_option(b"no_late_pass", 1))
        try:
            hAu[:] = np.nan
            lvl.apply_host(hu, out=hAu)
            b = hAu.copy()
        finally:
            capi.check(L.d4b_set_option(b"no_late_pass", 0))
        ref = lvl.apply_host(np.array(hu))       # pageable buffers: no late pass
        assert np.array_equal(b, ref)
        assert np.linalg.norm(a - ref) <= 1e-13 * np.linalg.norm(ref)
    finally:
        capi.check(L.d4b_host_free_pinned(hp)); capi.check(L.d4b_host_free_pinned(hq))
