#!/usr/bin/env python
"""bench.py -- GDoF/s of the matrix-free SIPG/DG Laplacian apply (FP64) on B200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--n 100] [--deg 3]

A "step" is one operator apply Au = A u over the whole resident mesh (BASELINE.json configs[1]:
uniform brick, p = 3, ~1M elements).  One JSON line on stdout (rank 0).  See DESIGN.md section 6.
"""
import argparse
import ctypes
import json
import os
import re
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
METRIC = "dg_laplacian_apply_gdof_per_s"
UNIT = "GDoF/s"


def workload_name(n, deg):
    return f"brick_{n}^3_p{deg}_gauss_legendre_degq{deg}"


def run_reference_copies(harness, inp, reps, copies):
    """`copies` concurrent processes of the reference harness, one per host core, each applying the operator on its
    own brick -- what d4est's MPI path does with one rank per core, minus its halo traffic (the snapshot has no
    MPI).  Returns (aggregate DoF/s, nodes per copy, slowest best-time)."""
    procs = [subprocess.Popen([harness, inp, "none", "0", "zero", "sinrand", "-", str(reps)], cwd=ROOT,
                              stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True) for _ in range(copies)]
    rate, nodes, tmax = 0.0, 0, 0.0
    for p in procs:
        out = p.communicate()[0].strip().splitlines()[-1]
        r = json.loads(out)
        rate += r["nodes"] / r["t_with_opt_best_s"]; nodes = r["nodes"]; tmax = max(tmax, r["t_with_opt_best_s"])
    return rate, nodes, tmax


def host_cores():
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except Exception:
        return max(1, os.cpu_count() or 1)


def reference_arm(args, rank):
    """The reference's own CPU apply (d4est_laplacian_with_opt_apply_aij, unmodified) timed by oracle/_ref/d4b_harness
    on a bounded sample of the workload, on all host cores (one serial copy per core)."""
    if rank != 0:
        return
    harness = os.path.join(ROOT, "oracle", "_ref", "d4b_harness")
    level = 4  # 16^3 = 4096 elements, 262144 DoF per copy: ~0.3 s per apply per core
    inp = os.path.join(ROOT, "oracle", "inputs", f"brick_p{args.deg}_l{level}.input")
    if not (os.path.exists(harness) and os.path.exists(inp)):
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/d4b_harness or its input not built"}))
        return
    reps = max(1, args.warmup + args.steps - 1)
    reps = min(reps, 20)
    cores = host_cores()
    rate, nodes, t = run_reference_copies(harness, inp, reps, cores)
    v = rate / 1e9
    sample = (f"{cores} concurrent serial copies of the reference (one per host core; the snapshot has no MPI), each brick 16^3 "
              f"p{args.deg} ({nodes} DoF), best of {reps + 1} applies; per-DoF rate, the reference apply is O(N_elem)")
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": t * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(args.n, args.deg), "sample": sample},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "reference", "sample": sample},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


class ClockSampler:
    """Streams `nvidia-smi -lms 20` for one GPU while the timed region runs (clocks + throttle reasons)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, dev):
        self.rows, self.proc, self.thr = [], None, None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(dev), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thr = threading.Thread(target=self._read, daemon=True)
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.rows.append((time.perf_counter(), [c.strip() for c in ln.strip().split(",")]))

    def start(self):
        if self.thr:
            self.thr.start()
            time.sleep(0.3)  # let the first samples arrive before the timed region starts

    def mark(self):
        return time.perf_counter()

    def summary(self, t0=None, t1=None):
        if self.proc:
            time.sleep(0.05)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=3)
            except Exception:
                self.proc.kill()
        rows = [r for (ts, r) in self.rows if (t0 is None or ts >= t0) and (t1 is None or ts <= t1 + 0.05)]
        if not rows:
            rows = [r for (_, r) in self.rows]
        sm = sorted(int(r[0]) for r in rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in rows if len(r) >= 6 for i in range(4) if r[2 + i] == "Active"})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def ncu_traffic(workload):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch of the apply kernel on this workload, from the
    committed `ncu --set full` capture (profiles/ncu_traffic.json); None when no capture exists for it."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json"))).get(workload, {}).get("dram_bytes")
    except Exception:
        return None


def cpu_baseline(args):
    """Reference CPU apply on a bounded sample (rank 0, N=1 only)."""
    harness = os.path.join(ROOT, "oracle", "_ref", "d4b_harness")
    inp = os.path.join(ROOT, "oracle", "inputs", f"brick_p{args.deg}_l4.input")
    if os.path.exists(harness) and os.path.exists(inp):
        cores = host_cores()
        rate, nodes, _ = run_reference_copies(harness, inp, 8, cores)
        return {"value": rate / 1e9, "unit": UNIT, "cores": cores, "kind": "reference",
                "sample": f"d4est_laplacian_with_opt_apply_aij, {cores} concurrent serial copies (one per host core; the reference "
                          f"snapshot has no MPI), each brick 16^3 p{args.deg} ({nodes} DoF), best of 9 applies"}
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import fixture, dg_oracle  # noqa: E401
    fx = fixture.load(os.path.join(ROOT, "tests", "golden", "brick_p3_l2.fix"))
    t0 = time.time(); dg_oracle.apply(fx, fx["u"]); t = time.time() - t0
    return {"value": fx["u"].size / t / 1e9, "unit": UNIT, "cores": 1, "kind": "port",
            "sample": "numpy oracle, brick 4^3 p3 (4096 DoF), one apply"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--n", type=int, default=100, help="brick elements per direction (per GPU)")
    ap.add_argument("--deg", type=int, default=3)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        reference_arm(args, rank)
        return
    args.warmup = max(args.warmup, 3)

    import numpy as np
    import torch
    import torch.distributed as dist
    import d4est_b200
    from d4est_b200 import capi
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local)
    capi.check(d4est_b200.lib().d4b_set_device(local))
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    L = d4est_b200.lib()
    if world > 1:
        # one communicator inside the library (NCCL, created from an id broadcast by the host program as d4est
        # would with MPI_Bcast): every apply then exchanges the halo of u before the element kernels
        idbuf = ctypes.create_string_buffer(128)
        if rank == 0:
            capi.check(L.d4b_comm_unique_id(idbuf))
        idt = torch.tensor(list(idbuf.raw), dtype=torch.uint8, device="cuda")
        dist.broadcast(idt, 0)
        idbuf = ctypes.create_string_buffer(bytes(idt.cpu().tolist()), 128)
        capi.check(L.d4b_comm_init(rank, world, idbuf))
    # weak scaling: rank r owns the z-slab r of the brick n x n x (n*world) over [0,1]^2 x [0,world]
    lvl = d4est_b200.MeshLevel.brick(args.n, deg=args.deg, box=(0., 1., 0., 1., 0., float(world)), rank=rank, nranks=world)
    ndof = lvl.local_nodes
    g = torch.Generator(device="cuda"); g.manual_seed(0x5EED + rank)
    u = torch.rand(ndof, dtype=torch.float64, device="cuda", generator=g) * 2 - 1
    Au = torch.empty_like(u)
    st = torch.cuda.current_stream().cuda_stream

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        lvl.apply_dev(u.data_ptr(), Au.data_ptr(), st)
    barrier()
    sampler = ClockSampler(local); sampler.start()
    barrier()
    n0 = L.d4b_kernel_launches()
    t_begin = sampler.mark()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    ev[0].record()
    for i in range(args.steps):
        lvl.apply_dev(u.data_ptr(), Au.data_ptr(), st)
        ev[i + 1].record()
    barrier()
    t_end = sampler.mark()
    launches = int(L.d4b_kernel_launches() - n0)
    total_ms = ev[0].elapsed_time(ev[-1])
    per = sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(args.steps))
    clocks = sampler.summary(t_begin, t_end)
    tmax = torch.tensor([total_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    total_ms_max = float(tmax.item())
    value = ndof * world * args.steps / (total_ms_max * 1e-3) / 1e9

    # end to end through the host-buffer entry point (the call d4est makes): pinned host u/Au,
    # H2D + kernel + D2H inside the timed region
    hp = ctypes.c_void_p(); hq = ctypes.c_void_p()
    capi.check(L.d4b_host_alloc_pinned(ndof * 8, ctypes.byref(hp))); capi.check(L.d4b_host_alloc_pinned(ndof * 8, ctypes.byref(hq)))
    hu = np.ctypeslib.as_array(ctypes.cast(hp, ctypes.POINTER(ctypes.c_double)), shape=(ndof,))
    hAu = np.ctypeslib.as_array(ctypes.cast(hq, ctypes.POINTER(ctypes.c_double)), shape=(ndof,))
    hu[:] = u.cpu().numpy()
    e2e_steps = max(3, min(args.steps, 10))
    lvl.apply_host(hu, out=hAu)
    barrier(); t0 = time.perf_counter()
    for _ in range(e2e_steps):
        lvl.apply_host(hu, out=hAu)
    barrier(); e2e_s = time.perf_counter() - t0
    te = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_val = ndof * world * e2e_steps / float(te.item()) / 1e9
    chk = float(np.abs(hAu - Au.cpu().numpy()).max())

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        alg = lvl.algorithmic_bytes
        kern_ms = per[len(per) // 2]  # one kernel launch per step: median step duration
        achieved = alg / (kern_ms * 1e-3) / 1e9
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": total_ms_max / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f64", "data": "synthetic",
                "config": {"workload": workload_name(args.n, args.deg), "elements_per_gpu": lvl.n_local, "dof_per_gpu": ndof,
                           "l2": f"inputs larger than L2 (resident tables {lvl.resident_bytes / 1e9:.2f} GB, vectors {2 * ndof * 8 / 1e9:.2f} GB)",
                           "parallelism": (f"{world} ranks, one z-slab of the brick {args.n}x{args.n}x{args.n * world} each; every apply = pack kernel + "
                                           f"NCCL send/recv of the mirror planes ({args.n * args.n * (args.deg + 1) ** 3 * 8 / 1e6:.1f} MB per neighbour) + element kernel"
                                           if world > 1 else "1 GPU")},
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak,
                             "unit": "GB/s", "frac": achieved / peak, "traffic": ncu_traffic(workload_name(args.n, args.deg)),
                             "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650 GB/s",
                             "algorithmic_bytes_per_launch": alg, "bytes_per_dof": alg / ndof, "kernel_ms": kern_ms},
                "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": ndof * 8, "d2h_bytes_per_step": ndof * 8,
                        "steps": e2e_steps, "max_abs_diff_vs_device_path": chk},
                "gpu_launches": launches, "clocks": clocks}
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(args)
        print(json.dumps(line))
    if world > 1:
        lvl.close()
        capi.check(L.d4b_comm_finalize())
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
