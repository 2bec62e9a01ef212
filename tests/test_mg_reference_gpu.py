"""The device multigrid / Krylov drivers against the UNMODIFIED reference solver stack
(Solver/d4est_solver_multigrid.c, d4est_solver_fcg_improved.c, d4est_solver_cg.c, d4est_krylov_pc_multigrid.c,
Chebyshev smoother + bottom solver) on a curved hp mesh: the fixture oracle/_ref/fixtures/mg_cubed_sphere_hp1.fix is
written by oracle/_ref/d4b_mg_harness, which runs the reference's own MG-preconditioned flexible CG on the CPU and
captures its hierarchy through the product glue.  All device calls go through the C ABI."""
import ctypes as C
import json
import os
import subprocess
import numpy as np
import pytest
import d4est_b200
from d4est_b200 import capi
from d4est_b200.capi import DevVec, lib, check
import fixture
import solver_oracle as so

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FIX = os.path.join(ROOT, "oracle", "_ref", "fixtures", "mg_cubed_sphere_hp1.fix")


def rel(a, b):
    return np.linalg.norm(a - b) / np.linalg.norm(b)


@pytest.fixture(scope="module")
def mgfx():
    if not os.path.exists(FIX):
        pytest.skip("multigrid fixture not built")
    fx = fixture.load(FIX)
    nl = int(fx["mg_meta"][0])
    levels = [d4est_b200.MeshLevel(fixture.sub(fx, f"L{l}.")) for l in range(nl)]
    tabs = [fixture.sub(fx, f"T{l}.") for l in range(nl - 1)]
    trs = [capi.Transfer(t) for t in tabs]
    par = so.mg_params(fx["mg_params"])
    kw = {k: par[k] for k in ("vcycle_imax", "vcycle_rtol", "vcycle_atol", "cheby_imax", "cheby_eigs_cg_imax",
                              "cheby_eigs_lmax_lmin_ratio", "cheby_eigs_max_multiplier", "eigs_reuse_fromdownvcycle",
                              "eigs_reuse_fromlastvcycle", "use_new_cg_eigs", "bottom_kind", "bottom_imax",
                              "bottom_eigs_cg_imax", "bottom_eigs_lmax_lmin_ratio", "bottom_eigs_max_multiplier",
                              "bottom_atol", "bottom_rtol")}
    mg = capi.Multigrid(levels, trs, **kw)
    return fx, levels, tabs, trs, mg, par


def test_transfers_match_reference_prolongation(mgfx):
    """k_mg_transfer over the captured hierarchy against the reference's d4est_operators_apply_{p,hp}_prolong and
    their transposes (restriction = P^T, Solver/d4est_solver_multigrid_callbacks.h:139,162; prolongation :266,287)."""
    fx, levels, tabs, trs, mg, par = mgfx
    for T, tr in zip(tabs, trs):
        F, Cc = DevVec(T["probe_fine"].size, T["probe_fine"]), DevVec(T["probe_coarse"].size)
        check(lib().d4b_mg_restrict(tr.handle, F.ptr, Cc.ptr, None))
        assert rel(Cc.get(), T["ref_restricted"]) < 1e-12
        Cc.set(T["probe_coarse"]); F2 = DevVec(T["probe_fine"].size)
        check(lib().d4b_mg_prolong(tr.handle, Cc.ptr, F2.ptr, None))
        assert rel(F2.get(), T["ref_prolonged"]) < 1e-12


def test_krylov_drivers_match_reference_25_iterations(mgfx):
    """d4b_fcg_solve / d4b_cg_solve against d4est_solver_fcg_solve / d4est_solver_cg_solve run by the reference for
    exactly 25 iterations without a preconditioner."""
    fx, levels, tabs, trs, mg, par = mgfx
    top = levels[-1]; n = top.local_nodes
    R, W = DevVec(n, fx["mg_rhs"]), DevVec(6 * n)
    it, res = C.c_int(), C.c_double()
    U = DevVec(n)
    check(lib().d4b_fcg_solve(top.handle, U.ptr, R.ptr, 25, 0., 1e-300, None, None, C.byref(it), C.byref(res), W.ptr, None))
    assert it.value == 25 and rel(U.get(), fx["krylov_fcg25_u"]) < 1e-9
    U2 = DevVec(n)
    check(lib().d4b_cg_solve(top.handle, U2.ptr, R.ptr, 25, 0., 1e-300, C.byref(it), C.byref(res), W.ptr, None))
    assert it.value == 25 and rel(U2.get(), fx["krylov_cg25_u"]) < 1e-9


def test_vcycle_preconditioner_matches_reference(mgfx):
    """z = MG(r): one V-cycle on the resident hierarchy (Chebyshev smoother with cg_eigs on every level, down and
    up, Chebyshev bottom solver) against the reference's d4est_krylov_pc_multigrid_apply on the same r."""
    fx, levels, tabs, trs, mg, par = mgfx
    n = levels[-1].local_nodes
    z = np.empty(n)
    r0 = np.ascontiguousarray(fx["mg_r0"])
    check(lib().d4b_mg_pc_apply_host(mg.handle, r0.ctypes.data_as(C.c_void_p), z.ctypes.data_as(C.c_void_p)))
    assert rel(z, fx["mg_z0"]) < 1e-9


def test_mg_fcg_solve_matches_reference_iteration_count(mgfx):
    """BASELINE config 4 on the curved hp mesh: MG-preconditioned flexible CG to the reference's tolerance; the
    reference's own solve of the same system took mg_meta[1] iterations (north_star: +-1 iteration) and ended at
    the same true residual."""
    fx, levels, tabs, trs, mg, par = mgfx
    n = levels[-1].local_nodes
    it_ref, imax = int(fx["mg_meta"][1]), int(fx["mg_meta"][3])
    u = np.zeros(n); rhs = np.ascontiguousarray(fx["mg_rhs"])
    it, res = C.c_int(), C.c_double()
    check(lib().d4b_mg_fcg_solve_host(mg.handle, u.ctypes.data_as(C.c_void_p), rhs.ctypes.data_as(C.c_void_p), imax,
                                      par["fcg_atol"], par["fcg_rtol"], C.byref(it), C.byref(res)))
    assert abs(it.value - it_ref) <= 1, (it.value, it_ref)
    true_res = np.linalg.norm(rhs - levels[-1].apply_host(u))
    ref_res = np.linalg.norm(rhs - levels[-1].apply_host(fx["mg_u_ref"]))
    assert true_res < 10 * ref_res + 1e-12 * np.linalg.norm(rhs)
    assert rel(u, fx["mg_u_ref"]) < 1e-6


def test_cuda_graph_replay_of_the_preconditioner_is_bitwise_the_plain_path(mgfx):
    """option "mg_graph": the single-V-cycle preconditioner application captured into a CUDA graph (second call on)
    replays the same kernels in the same order: same iterates, bit for bit, and far fewer host launches."""
    fx, levels, tabs, trs, mg, par = mgfx
    n = levels[-1].local_nodes
    rhs = np.ascontiguousarray(fx["mg_rhs"])
    L = lib()
    out = {}
    for key, on in (("plain", 0), ("graph", 1)):
        check(L.d4b_set_option(b"mg_graph", on))
        try:
            u = np.zeros(n); it, res = C.c_int(), C.c_double()
            n0 = L.d4b_kernel_launches()
            check(L.d4b_mg_fcg_solve_host(mg.handle, u.ctypes.data_as(C.c_void_p), rhs.ctypes.data_as(C.c_void_p), 12,
                                          0., 1e-300, C.byref(it), C.byref(res)))
            out[key] = (u, it.value, L.d4b_kernel_launches() - n0)
        finally:
            check(L.d4b_set_option(b"mg_graph", 0))
    assert out["plain"][1] == out["graph"][1] == 12
    assert np.array_equal(out["plain"][0], out["graph"][0])
    assert out["graph"][2] < out["plain"][2] // 4, (out["graph"][2], out["plain"][2])


def test_d4est_facing_entry_points_in_one_process_with_the_reference():
    """oracle/_ref/d4b_mg_harness --gpu-check: ONE process that holds the unmodified reference and calls, next to
    it, the d4est-facing entry points of the product glue on the GPU: d4est_b200_apply_lhs (d4est_apply_operator_fcn_t),
    the baseline-twin entry, d4est_b200_build_residual, the d4est_krylov_pc_t on the resident hierarchy captured
    from d4est's own coarsening, d4est_b200_fcg_solve (d4est_solver_fcg_solve's argument list), and d4est's own
    V-cycle with the smoother / bottom-solver bodies on the device (d4est_b200_multigrid_use_device)."""
    harness = os.path.join(ROOT, "oracle", "_ref", "d4b_mg_harness")
    if not os.path.exists(harness) or not os.path.exists(FIX):
        pytest.skip("oracle harness not built")
    it_ref = int(fixture.load(FIX)["mg_meta"][1])
    r = subprocess.run([harness, "oracle/inputs/cubed_sphere_hp_mg.input", "random_hp", "1", "-", "--gpu-check", str(it_ref)],
                       cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=1200)
    assert r.returncode == 0, r.stdout[-3000:]
    out = [json.loads(l) for l in r.stdout.splitlines() if l.startswith("{")]
    chk = [o for o in out if o.get("what") == "gpu_check"][0]
    print(json.dumps(chk))
    assert chk["apply_lhs_rel_l2"] < 1e-12 and chk["baseline_twin_rel_l2"] < 1e-12 and chk["build_residual_rel_l2"] < 1e-12
    assert chk["mass_term_rel_l2"] < 1e-12
    assert chk["pc_apply_resident_rel_l2"] < 1e-9
    assert chk["pc_apply_d4est_vcycle_device_smoother_rel_l2"] < 1e-9
    assert abs(chk["fcg_iterations_device"] - chk["fcg_iterations_reference"]) <= 1
    assert chk["kernel_launches"] > 0
