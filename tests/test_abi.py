"""The C-ABI library loads, exports every symbol include/d4est_b200.h declares, and refuses to
compute without a CUDA device (no CPU fallback).  No GPU needed."""
import ctypes
import os
import re
import numpy as np
import pytest
import d4est_b200
from d4est_b200 import capi
import fixture

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    txt = open(os.path.join(ROOT, "include", "d4est_b200.h")).read()
    return sorted(set(re.findall(r"\b(d4b_[a-z0-9_]+)\s*\(", txt)))


def test_exports_every_declared_symbol():
    L = ctypes.CDLL(d4est_b200.LIB_PATH)
    names = declared_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(L, n), f"{n} declared in d4est_b200.h but not exported"
    assert set(names) == set(capi.SYMBOLS), "capi.SYMBOLS out of sync with the header"


def test_struct_layout_matches_header():
    assert ctypes.sizeof(capi.HM) == 40 == fixture.HM_DTYPE.itemsize
    assert capi.HM.fac_off.offset == 16 and capi.HM.g_off.offset == 24


def test_no_cpu_fallback(golden_dir):
    n = ctypes.c_int(-1)
    d4est_b200.lib().d4b_device_count(ctypes.byref(n))
    if n.value > 0:
        pytest.skip("CUDA device present")
    fx = fixture.load(os.path.join(golden_dir, "brick_p3_l2.fix"))
    with pytest.raises(d4est_b200.D4BError, match="no CUDA device"):
        d4est_b200.MeshLevel(fx)


def test_ops_tables_match_independent_numpy_tables():
    """d4b_ops.c (closed forms, long double) vs oracle/dg_oracle.py (numpy.polynomial)."""
    import dg_oracle
    L = ctypes.CDLL(d4est_b200.LIB_PATH)
    P = ctypes.c_void_p
    for deg in range(1, 9):
        N = deg + 1
        D = np.zeros((N, N)); L.d4b_ops_diff(deg, D.ctypes.data_as(P))
        assert np.abs(D - dg_oracle.diff_matrix(deg)).max() < 2e-13
        x = np.zeros(N); w = np.zeros(N); L.d4b_ops_gll(deg, x.ctypes.data_as(P), w.ctypes.data_as(P))
        xo, wo = dg_oracle.gll(deg)
        assert np.abs(x - xo).max() < 1e-15 and np.abs(w - wo).max() < 1e-14
        for degq in range(deg, min(deg + 4, 12)):
            for qt in (0, 1):
                for child in (-1, 0, 1):
                    I = np.zeros((degq + 1, N))
                    L.d4b_ops_interp(deg, degq, qt, child, I.ctypes.data_as(P))
                    assert np.abs(I - dg_oracle.interp(deg, degq, qt, child)).max() < 1e-13
                    assert np.abs(I.sum(1) - 1).max() < 1e-14  # partition of unity
