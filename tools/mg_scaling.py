"""BASELINE config 5: the multigrid-preconditioned flexible-CG solve on the curved hp mesh on N GPUs (one rank per
GPU, strong scaling: the mesh is fixed).  Not the headline metric (bench.py is).

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/mg_scaling.py

The hierarchy is captured once from the reference's own coarsening (oracle/_ref/d4b_mg_harness --capture-only: mesh,
reference apply for the right-hand side, hierarchy walk; no reference solve), then cut into rank-local levels and
transfer tables the way d4est partitions (families stay together, d4est_b200/partition.py); every apply on every
level exchanges its halo with ncclSend/ncclRecv, every reduction of cg_eigs / FCG is an in-stream ncclAllReduce.
Prints one JSON line (rank 0): seconds of the solve (max over ranks), iterations, true relative residual."""
import ctypes as C
import json
import os
import subprocess
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import d4est_b200  # noqa: E402
from d4est_b200 import capi, partition as P  # noqa: E402
import fixture  # noqa: E402   (table reader only)
import solver_oracle as so  # noqa: E402   (parameter names only)


def main():
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    steps = sys.argv[1] if len(sys.argv) > 1 else "3"
    inp = sys.argv[2] if len(sys.argv) > 2 else "oracle/inputs/cubed_sphere_hp_l2_mg.input"
    path = f"/tmp/d4b_mgscale_{steps}.fix"
    torch.cuda.set_device(local)
    L = d4est_b200.lib()
    capi.check(L.d4b_set_device(local))
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if rank == 0 and not os.path.exists(path):
        subprocess.run([os.path.join(ROOT, "oracle", "_ref", "d4b_mg_harness"), inp, "random_hp", steps, path, "--capture-only"],
                       cwd=ROOT, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    if world > 1:
        dist.barrier()
        idbuf = C.create_string_buffer(128)
        if rank == 0:
            capi.check(L.d4b_comm_unique_id(idbuf))
        t = torch.tensor(list(idbuf.raw), dtype=torch.uint8, device="cuda")
        dist.broadcast(t, 0)
        idbuf = C.create_string_buffer(bytes(t.cpu().tolist()), 128)
        capi.check(L.d4b_comm_init(rank, world, idbuf))
    fx = fixture.load(path)
    nl = int(fx["mg_meta"][0])
    levels = [fixture.sub(fx, f"L{l}.") for l in range(nl)]
    tabs = [fixture.sub(fx, f"T{l}.") for l in range(nl - 1)]
    # the same level set for every N (d4est drops coarse levels as ranks are added, multigrid.c:253-277): the base
    # level is the coarsest one with at least 64 elements
    base = next(l for l in range(nl) if int(levels[l]["hdr"][0]) >= 64)
    t0 = time.time()
    lv, tr = P.partition_hierarchy(levels, tabs, world, base, only=rank)[rank]
    t_part = time.time() - t0
    dl = [d4est_b200.MeshLevel(t) for t in lv]
    dt = [capi.Transfer(t) for t in tr]
    par = so.mg_params(fx["mg_params"])
    kw = {k: par[k] for k in ("vcycle_imax", "vcycle_rtol", "vcycle_atol", "cheby_imax", "cheby_eigs_cg_imax",
                              "cheby_eigs_lmax_lmin_ratio", "cheby_eigs_max_multiplier", "eigs_reuse_fromdownvcycle",
                              "eigs_reuse_fromlastvcycle", "use_new_cg_eigs", "bottom_kind", "bottom_imax",
                              "bottom_eigs_cg_imax", "bottom_eigs_lmax_lmin_ratio", "bottom_eigs_max_multiplier",
                              "bottom_atol", "bottom_rtol")}
    mg = capi.Multigrid(dl, dt, **kw)
    a, b = lv[-1]["node_range"]
    rhs = np.ascontiguousarray(fx["mg_rhs"][a:b]); n = rhs.size
    capi.check(L.d4b_set_option(b"fcg_textbook_update", 1))   # see bench.py secondary_measurements
    it, res = C.c_int(), C.c_double()
    secs = []
    for rep in range(2):
        u = np.zeros(n)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(); t0 = time.perf_counter()
        capi.check(L.d4b_mg_fcg_solve_host(mg.handle, u.ctypes.data_as(C.c_void_p), rhs.ctypes.data_as(C.c_void_p), 200,
                                           par["fcg_atol"], par["fcg_rtol"], C.byref(it), C.byref(res)))
        torch.cuda.synchronize(); secs.append(time.perf_counter() - t0)
    r = rhs - dl[-1].apply_host(u)
    red = torch.tensor([secs[-1], float(r @ r), float(rhs @ rhs)], dtype=torch.float64, device="cuda")
    if world > 1:
        tmax = red[:1].clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        sums = red[1:].clone(); dist.all_reduce(sums)
        red = torch.cat([tmax, sums])
    if rank == 0:
        print(json.dumps({"what": "mg_fcg_curved_hp_strong_scaling", "n_gpus": world, "seconds": float(red[0]),
                          "iterations": it.value, "true_relative_residual": float((red[1] / red[2]).sqrt()),
                          "levels_elements_global": [int(levels[l]["hdr"][0]) for l in range(base, nl)],
                          "dof_global": int(levels[-1]["node_off"][int(levels[-1]["hdr"][0])]),
                          "local_elements_rank0": [int(t["hdr"][0]) for t in lv], "ghost_elements_rank0": [int(t["hdr"][1]) for t in lv],
                          "partition_seconds_host": t_part, "fcg_update": "textbook", "launches_rank0": int(L.d4b_kernel_launches())}),
              flush=True)
    if world > 1:
        for m in dl:
            m.close()
        capi.check(L.d4b_comm_finalize())
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
