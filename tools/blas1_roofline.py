"""Achieved HBM bandwidth of the BLAS-1 / reduction kernels of the solver (d4b_solver.cu) at 64 M doubles per vector,
as a fraction of MEASURED_PEAKS.json hbm_gbs.  Bytes counted: 8 per double read or written, each vector once."""
import ctypes as C, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import d4est_b200
from d4est_b200.capi import lib, check
L = lib(); check(L.d4b_set_device(0))
n = 64_000_000
x, y, z = (torch.rand(n, dtype=torch.float64, device="cuda") for _ in range(3))
st = torch.cuda.current_stream().cuda_stream
peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
d = C.c_double()
cases = {
    "k_axpy (y += a x)": (lambda: L.d4b_vec_axpy(n, 0.5, x.data_ptr(), y.data_ptr(), st), 3),
    "k_xpby (y = x + b y)": (lambda: L.d4b_vec_xpby(n, x.data_ptr(), 0.5, y.data_ptr(), st), 3),
    "k_axpyeqz (z = a x + y)": (lambda: L.d4b_vec_axpyeqz(n, 0.5, x.data_ptr(), y.data_ptr(), z.data_ptr(), st), 3),
    "k_scale (x *= a)": (lambda: L.d4b_vec_scale(n, 1.0000001, x.data_ptr(), st), 2),
    "k_dots_partial+final (x.y, incl. D2H of the scalar)": (lambda: L.d4b_vec_dot(n, x.data_ptr(), y.data_ptr(), C.byref(d), st), 2),
}
out = {}
for name, (f, vecs) in cases.items():
    for _ in range(3):
        check(f())
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        check(f())
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    gbs = vecs * n * 8 / ms / 1e6
    out[name] = {"ms": ms, "GBps": gbs, "frac_of_measured_hbm": gbs / peak}
print(json.dumps(out, indent=1))
