"""Allocation-free dot product (d4b_vec_dot_noalloc, include/d4est_b200.h): same result as d4b_vec_dot, twice in a
row, through the C ABI; sizes beyond the stated maximum are refused.  Kept in a file pytest collects last."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RTOL = 1e-12   # the tolerance of the matvec / BLAS-1 parity tests


def _lib():
    import torch
    from d4est_b200 import capi
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    return torch, capi, capi.lib()


@pytest.mark.parametrize("n", [1, 1000, 148 * 16 * 1024 + 17, 5_000_003])
def test_noalloc_dot_matches_vec_dot_twice(n):
    torch, capi, L = _lib()
    g = torch.Generator(device="cuda"); g.manual_seed(1234 + n % 97)
    st = torch.cuda.current_stream().cuda_stream
    out = torch.full((1,), float("nan"), dtype=torch.float64, device="cuda")
    for call in range(2):   # the static scratch keeps the previous call's partials: the second call must not see them
        x = torch.rand(n, dtype=torch.float64, device="cuda", generator=g) * 2 - 1
        y = torch.rand(n, dtype=torch.float64, device="cuda", generator=g) * 2 - 1
        ref = ctypes.c_double(0.)
        capi.check(L.d4b_vec_dot(n, x.data_ptr(), y.data_ptr(), ctypes.byref(ref), st))
        out.fill_(float("nan"))
        capi.check(L.d4b_vec_dot_noalloc(n, x.data_ptr(), y.data_ptr(), out.data_ptr(), st))
        got = float(out.item())
        scale = float((x.abs() * y.abs()).sum().item())
        assert abs(got - ref.value) <= RTOL * scale, (call, n, got, ref.value)
        exact = float(np.dot(x.cpu().numpy(), y.cpu().numpy()))
        assert abs(got - exact) <= 1e-11 * scale


def test_noalloc_dot_shorter_vector_after_a_longer_one():
    """fewer blocks than the previous call: stale partials beyond the new grid stay unread"""
    torch, capi, L = _lib()
    st = torch.cuda.current_stream().cuda_stream
    out = torch.zeros(1, dtype=torch.float64, device="cuda")
    big = torch.ones(3_000_000, dtype=torch.float64, device="cuda")
    capi.check(L.d4b_vec_dot_noalloc(big.numel(), big.data_ptr(), big.data_ptr(), out.data_ptr(), st))
    assert float(out.item()) == 3_000_000.0
    small = torch.full((100,), 2.0, dtype=torch.float64, device="cuda")
    capi.check(L.d4b_vec_dot_noalloc(small.numel(), small.data_ptr(), small.data_ptr(), out.data_ptr(), st))
    assert float(out.item()) == 400.0
    capi.check(L.d4b_vec_dot_noalloc(0, small.data_ptr(), small.data_ptr(), out.data_ptr(), st))
    assert float(out.item()) == 0.0


def test_noalloc_dot_allocates_nothing_and_rejects_oversize():
    torch, capi, L = _lib()
    st = torch.cuda.current_stream().cuda_stream
    n = 1 << 20
    x = torch.ones(n, dtype=torch.float64, device="cuda")
    out = torch.zeros(1, dtype=torch.float64, device="cuda")
    capi.check(L.d4b_vec_dot_noalloc(n, x.data_ptr(), x.data_ptr(), out.data_ptr(), st))   # set-up call
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(10):
        capi.check(L.d4b_vec_dot_noalloc(n, x.data_ptr(), x.data_ptr(), out.data_ptr(), st))
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0
    assert float(out.item()) == float(n)
    too_big = (1 << 28) + 1
    rc = L.d4b_vec_dot_noalloc(too_big, x.data_ptr(), x.data_ptr(), out.data_ptr(), st)
    assert rc == 1   # D4B_EINVAL
    assert b"D4B_NOALLOC_DOT_MAX_N" in L.d4b_last_error()
    assert L.d4b_vec_dot_noalloc(-1, x.data_ptr(), x.data_ptr(), out.data_ptr(), st) == 1
    assert float(out.item()) == float(n)   # a refused call launches nothing
