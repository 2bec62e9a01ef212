"""Secondary measurements (BASELINE configs 3 and 4), printed as JSON lines; not the headline bench.

  python tools/bench_solver.py [--top 64] [--deg 3]

1. MG-preconditioned flexible CG (Chebyshev smoother, Chebyshev bottom solver; CDS options.input:101-132
   values) to relative residual 1e-10 on a uniform brick hierarchy 4^3 .. top^3, everything device-resident.
2. Apply throughput on the curved mixed-hp cubed-sphere tables produced by the reference harness
   (hanging faces, 13 trees, inter-tree orientations), if those table files are present.
"""
import argparse
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import d4est_b200  # noqa: E402
from d4est_b200 import capi, tables  # noqa: E402
from d4est_b200.capi import DevVec, check, lib  # noqa: E402


def transfer_table(n_coarse, deg):
    nH, N3 = n_coarse ** 3, (deg + 1) ** 3
    return {"kind": np.ones(nH, np.int32), "degH": np.full(nH, deg, np.int32),
            "coarse_off": np.arange(nH, dtype=np.int64) * N3, "child_deg": np.full(8 * nH, deg, np.int32),
            "child_off": np.arange(8 * nH, dtype=np.int64) * N3}


def mg_solve(top, deg):
    sizes = []
    n = 4
    while n <= top:
        sizes.append(n); n *= 2
    levels = [d4est_b200.MeshLevel.brick(s, deg=deg) for s in sizes]
    trs = [capi.Transfer(transfer_table(s, deg)) for s in sizes[:-1]]
    mg = capi.Multigrid(levels, trs, cheby_imax=15, cheby_eigs_cg_imax=15, bottom_imax=15, bottom_eigs_cg_imax=15)
    nd = levels[-1].local_nodes
    rhs = np.random.default_rng(11).standard_normal(nd)
    U, R, W = DevVec(nd), DevVec(nd, rhs), DevVec(6 * nd)
    it, res = C.c_int(), C.c_double()
    out = {}
    for rep in range(2):  # first solve also estimates the spectral bounds; time the second
        check(lib().d4b_vec_zero(nd, U.ptr, None)); check(lib().d4b_sync(None))
        l0 = lib().d4b_kernel_launches(); t0 = time.perf_counter()
        check(lib().d4b_mg_fcg_solve(mg.handle, U.ptr, R.ptr, 100, 0., 1e-10, C.byref(it), C.byref(res), W.ptr, None))
        check(lib().d4b_sync(None))
        out = {"seconds": time.perf_counter() - t0, "iterations": it.value, "launches": int(lib().d4b_kernel_launches() - l0)}
    r = rhs - levels[-1].apply_host(U.get())
    out.update({"what": "mg_fcg_solve_to_1e-10", "levels": sizes, "deg": deg, "dof": nd,
                "true_rel_residual": float(np.linalg.norm(r) / np.linalg.norm(rhs)),
                "mdof_per_s_to_solution": nd / out["seconds"] / 1e6})
    print(json.dumps(out), flush=True)


def hp_apply(path):
    import torch
    fx = tables.load(path)
    lvl = d4est_b200.MeshLevel(fx)
    n = lvl.local_nodes
    parity = None
    if "Au_ref_with_opt" in fx:  # the table file carries the reference's own Au for its u
        ref = fx["Au_ref_with_opt"]
        parity = float(np.linalg.norm(lvl.apply_host(fx["u"]) - ref) / np.linalg.norm(ref))
    u = torch.rand(n, dtype=torch.float64, device="cuda"); Au = torch.empty_like(u)
    st = torch.cuda.current_stream().cuda_stream
    for _ in range(5):
        lvl.apply_dev(u.data_ptr(), Au.data_ptr(), st)
    torch.cuda.synchronize()
    reps = 200
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        lvl.apply_dev(u.data_ptr(), Au.data_ptr(), st)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    deg = fx["deg"][:lvl.n_local]
    print(json.dumps({"what": "hp_curved_apply", "tables": os.path.basename(path), "elements": lvl.n_local, "dof": n,
                      "deg_min": int(deg.min()), "deg_max": int(deg.max()), "ms_per_apply": ms,
                      "gdof_per_s": n / ms / 1e6, "algorithmic_GBps": lvl.algorithmic_bytes / ms / 1e6,
                      "hanging_records": int((fx["hm"]["child_m"] >= 0).sum() + (fx["hm"]["child_p"] >= 0).sum()),
                      "kernel_split_regular_hp_generic": lvl.kernel_split,
                      "rel_l2_vs_reference_Au": parity}), flush=True)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--top", type=int, default=64)
    ap.add_argument("--deg", type=int, default=3)
    ap.add_argument("--hp", default=None, help="flat-table file of a (large) curved hp mesh: only time its apply")
    a = ap.parse_args()
    check(lib().d4b_set_device(0))
    if a.hp:
        hp_apply(a.hp)
        sys.exit(0)
    mg_solve(a.top, a.deg)
    for nm in ("cubed_sphere_hp3", "cubed_sphere_hp2"):
        p = os.path.join(ROOT, "oracle", "_ref", "fixtures", nm + ".fix")
        if os.path.exists(p):
            hp_apply(p)
