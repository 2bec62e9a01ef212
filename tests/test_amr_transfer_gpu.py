"""AMR field transfer on the device (d4b_amr_* in include/d4est_b200.h) against the UNMODIFIED reference's
d4est_amr_step with a field attached (tests/golden/amr_transfer_cubed_sphere_hp2.fix, made by
oracle/harness/d4b_amr_harness.c) and against the numpy restatement."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import fixture
import solver_oracle as so
from d4est_b200 import capi
from d4est_b200.capi import DevVec, check, lib

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _i32(a):
    a = np.ascontiguousarray(a, np.int32)
    return a, a.ctypes.data_as(C.c_void_p)


def _steps(golden_dir):
    fx = fixture.load(os.path.join(golden_dir, "amr_transfer_cubed_sphere_hp2.fix"))
    for k in range(int(fx["steps"][0])):
        yield (fx[f"S{k}.initial_log"], fx[f"S{k}.refinement_log"], fx[f"S{k}.balance_log"],
               fx[f"S{k}.field_pre"], fx[f"S{k}.field_post"])


def test_host_entry_matches_reference(golden_dir):
    L = lib()
    for il, rl, bl, pre, post in _steps(golden_dir):
        (il, pil), (rl, prl), (bl, pbl) = _i32(il), _i32(rl), _i32(bl)
        pre = np.ascontiguousarray(pre); out = np.empty(post.size)
        check(L.d4b_amr_interpolate_field_host(il.size, pil, prl, bl.size, pbl, pre.ctypes.data_as(C.c_void_p), pre.size,
                                               out.ctypes.data_as(C.c_void_p), out.size))
        assert np.abs(out - post).max() <= 1e-13 * np.abs(post).max()
        ora = so.amr_interpolate_field(il, rl, bl, pre)
        assert np.abs(out - ora).max() <= 1e-13 * np.abs(post).max()


def test_device_entry_and_single_sweep_tables(golden_dir):
    L = lib()
    il, rl, bl, pre, post = next(_steps(golden_dir))
    (il, pil), (rl, prl), (bl, pbl) = _i32(il), _i32(rl), _i32(bl)
    n_in, n_aux, n_out = C.c_int64(), C.c_int64(), C.c_int64()
    check(L.d4b_amr_interpolate_field_sizes(il.size, pil, prl, bl.size, pbl, C.byref(n_in), C.byref(n_aux), C.byref(n_out)))
    assert (n_in.value, n_out.value) == (pre.size, post.size)
    X, A, Y = DevVec(n_in.value, pre), DevVec(n_aux.value), DevVec(n_out.value)
    check(L.d4b_amr_interpolate_field_dev(il.size, pil, prl, bl.size, pbl, X.ptr, A.ptr, Y.ptr, None))
    assert np.abs(Y.get() - post).max() <= 1e-13 * np.abs(post).max()
    # the first sweep alone through its transfer table + d4b_mg_prolong
    T = C.c_void_p(); a, b = C.c_int64(), C.c_int64()
    check(L.d4b_amr_transfer_create(il.size, pil, prl, C.byref(T), C.byref(a), C.byref(b)))
    assert (a.value, b.value) == (n_in.value, n_aux.value)
    A2 = DevVec(n_aux.value)
    check(L.d4b_mg_prolong(T, X.ptr, A2.ptr, None)); check(L.d4b_sync(None))
    t1, _, _ = so.amr_sweep_table(il, rl)
    assert np.abs(A2.get() - so.transfer(t1, pre, n_aux.value, restrict=False)).max() <= 1e-13 * np.abs(pre).max()
    L.d4b_mg_transfer_destroy(T)


def test_bad_logs_are_refused(golden_dir):
    """coarsening (code below the degree) aborts in the reference (d4est_amr.c:395): D4B_EINVAL here; so do logs
    whose sizes do not chain and h-refinement codes that change the degree"""
    L = lib()
    x = np.zeros(27 + 64); out = np.zeros(8 * 27 + 64)
    cases = [([2, 3], [-2, 2], [2] * 8 + [3]),        # p-coarsening 3 -> 2
             ([2, 3], [-3, 3], [2] * 8 + [3]),        # h-refinement with another degree
             ([2, 3], [-2, 3], [2] * 7 + [3])]        # balance log does not consume what the refinement log produced
    for il, rl, bl in cases:
        (il, pil), (rl, prl), (bl, pbl) = _i32(il), _i32(rl), _i32(bl)
        rc = L.d4b_amr_interpolate_field_host(il.size, pil, prl, bl.size, pbl, x.ctypes.data_as(C.c_void_p), x.size,
                                              out.ctypes.data_as(C.c_void_p), out.size)
        assert rc == 1, (il, rl, bl)
        assert b"amr transfer" in L.d4b_last_error()


def test_glue_entry_in_process_with_the_reference():
    """d4est_b200_amr_interpolate_field (the d4est-facing glue) on the GPU next to the reference's own
    d4est_amr_interpolate_field in one process, every step of a 3-step random-hp refinement."""
    exe = os.path.join(ROOT, "oracle", "_ref", "d4b_amr_harness")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref not built")
    env = dict(os.environ, D4EST_B200_GPU_CHECK="1")
    r = subprocess.run([exe, os.path.join(ROOT, "oracle", "inputs", "cubed_sphere_hp.input"), "3", "/tmp/d4b_amr_check.fix"],
                       cwd=ROOT, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:]
    assert "gpu_check_amr_transfer_rel_l2" in r.stdout
