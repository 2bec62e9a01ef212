"""Worker of the world_size>1 tests (launched by torch.distributed.run).

  python -m torch.distributed.run --nproc-per-node 2 ... tests/mp_halo_worker.py <fixture> <cpu|gpu>

cpu: gloo halo (d4est_b200.partition.exchange_halo_torch) + numpy oracle apply on the rank-local
     tables must equal the serial oracle apply on the rank's node range.
gpu: NCCL halo inside d4b_apply_dev (d4b_comm_init + d4b_halo_exchange) must equal the reference Au.
"""
import ctypes
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import fixture  # noqa: E402
import dg_oracle  # noqa: E402
from d4est_b200 import partition as part_mod  # noqa: E402


def main():
    fx = fixture.load(sys.argv[1]); mode = sys.argv[2]
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    parts = part_mod.partition(fx, world)
    me = parts[rank]
    u = fx["u"]
    ul = part_mod.local_vector(fx, me, u)
    a, b = me["node_range"]
    if mode == "cpu":
        dist.init_process_group("gloo")
        ug = part_mod.exchange_halo_torch(me, ul, dist)
        assert np.array_equal(ug, part_mod.ghost_vector(fx, me, u)), "halo delivered wrong ghost values"
        Au = dg_oracle.apply(me, ul, ug)
        ref = dg_oracle.apply(fx, u)[a:b]
    else:
        import d4est_b200
        from d4est_b200 import capi
        local = int(os.environ.get("LOCAL_RANK", rank))
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        L = d4est_b200.lib()
        capi.check(L.d4b_set_device(local))
        idbuf = ctypes.create_string_buffer(128)
        if rank == 0:
            capi.check(L.d4b_comm_unique_id(idbuf))
        t = torch.tensor(list(idbuf.raw), dtype=torch.uint8, device="cuda")
        dist.broadcast(t, 0)
        idbuf = ctypes.create_string_buffer(bytes(t.cpu().tolist()), 128)
        capi.check(L.d4b_comm_init(rank, world, idbuf))
        # small fixtures: lift the size threshold so that the apply overlaps its exchange with the ghost-free batches
        capi.check(L.d4b_set_option(b"overlap_min_elements", 0))
        lvl = d4est_b200.MeshLevel(me)
        Au = lvl.apply_host(ul)          # halo exchange over NCCL inside the apply
        capi.check(L.d4b_set_option(b"no_overlap", 1))
        Au_serial = lvl.apply_host(ul)   # exchange-then-apply: same kernels on the same data, same bits
        capi.check(L.d4b_set_option(b"no_overlap", 0))
        assert np.array_equal(Au, Au_serial), "overlapped and exchange-then-apply results differ"
        ref = fx["Au_ref_with_opt"][a:b]
        krylov_ok = True
        if len(sys.argv) > 3 and sys.argv[3] == "krylov" and fx["bc_g"].size == 0:
            # the reductions of the Krylov drivers over ncclAllReduce (device scalars, in stream): a fixed number of
            # CG / flexible-CG iterations on the partitioned level must reproduce the 1-rank iterates on this rank's
            # node range (sum order of the dots differs: agreement to rounding), and the stopping test must fire in the
            # same iteration on every rank
            from d4est_b200.capi import DevVec
            n = lvl.local_nodes
            rhs_g = np.sin(0.3 * np.arange(u.size)) + 0.1
            R, U, W = DevVec(n, rhs_g[a:b]), DevVec(n), DevVec(6 * n)
            it, res = ctypes.c_int(), ctypes.c_double()
            capi.check(L.d4b_cg_solve(lvl.handle, U.ptr, R.ptr, 30, 0., 1e-300, ctypes.byref(it), ctypes.byref(res), W.ptr, None))
            ucg = U.get()
            U2 = DevVec(n)
            capi.check(L.d4b_fcg_solve(lvl.handle, U2.ptr, R.ptr, 12, 0., 1e-300, None, None, ctypes.byref(it), ctypes.byref(res), W.ptr, None))
            ufcg = U2.get()
            U3 = DevVec(n); it3 = ctypes.c_int()
            capi.check(L.d4b_cg_solve(lvl.handle, U3.ptr, R.ptr, 2000, 0., 1e-8, ctypes.byref(it3), ctypes.byref(res), W.ptr, None))
            # serial references on rank 0's GPU... every rank computes them on its own GPU from the unpartitioned tables
            capi.check(L.d4b_comm_finalize())
            whole = d4est_b200.MeshLevel(fx); nw = whole.local_nodes
            Rw, Uw, Ww = DevVec(nw, rhs_g), DevVec(nw), DevVec(6 * nw)
            capi.check(L.d4b_cg_solve(whole.handle, Uw.ptr, Rw.ptr, 30, 0., 1e-300, ctypes.byref(it), ctypes.byref(res), Ww.ptr, None))
            e1 = np.linalg.norm(ucg - Uw.get()[a:b]) / np.linalg.norm(Uw.get()[a:b])
            Uw2 = DevVec(nw)
            capi.check(L.d4b_fcg_solve(whole.handle, Uw2.ptr, Rw.ptr, 12, 0., 1e-300, None, None, ctypes.byref(it), ctypes.byref(res), Ww.ptr, None))
            e2 = np.linalg.norm(ufcg - Uw2.get()[a:b]) / np.linalg.norm(Uw2.get()[a:b])
            Uw3 = DevVec(nw); itw = ctypes.c_int()
            capi.check(L.d4b_cg_solve(whole.handle, Uw3.ptr, Rw.ptr, 2000, 0., 1e-8, ctypes.byref(itw), ctypes.byref(res), Ww.ptr, None))
            krylov_ok = e1 < 1e-9 and e2 < 1e-9 and abs(it3.value - itw.value) <= 1
            print(f"rank {rank}: CG 30 its rel diff vs 1 rank {e1:.2e}, FCG 12 its {e2:.2e}, CG to 1e-8: {it3.value} iterations "
                  f"(1 rank: {itw.value})", flush=True)
            whole.close()
        else:
            capi.check(L.d4b_comm_finalize())
        lvl.close()
    err = np.linalg.norm(Au - ref) / np.linalg.norm(ref)
    ok = torch.tensor([1.0 if (err < 1e-12 and (mode != "gpu" or krylov_ok)) else 0.0])
    if mode == "gpu":
        ok = ok.cuda()
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    print(f"rank {rank}/{world}: {me['hdr'][0]} local, {me['hdr'][1]} ghost elements, rel L2 {err:.2e}", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if float(ok.item()) == 1.0 else 1)


if __name__ == "__main__":
    main()
