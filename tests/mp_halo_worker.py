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
        lvl = d4est_b200.MeshLevel(me)
        Au = lvl.apply_host(ul)          # halo exchange over NCCL inside the apply
        ref = fx["Au_ref_with_opt"][a:b]
        lvl.close()
        capi.check(L.d4b_comm_finalize())
    err = np.linalg.norm(Au - ref) / np.linalg.norm(ref)
    ok = torch.tensor([1.0 if err < 1e-12 else 0.0])
    if mode == "gpu":
        ok = ok.cuda()
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    print(f"rank {rank}/{world}: {me['hdr'][0]} local, {me['hdr'][1]} ghost elements, rel L2 {err:.2e}", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if float(ok.item()) == 1.0 else 1)


if __name__ == "__main__":
    main()
