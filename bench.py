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
    env = dict(os.environ, OPENBLAS_NUM_THREADS="1", OMP_NUM_THREADS="1")   # one serial copy per core
    procs = [subprocess.Popen([harness, inp, "none", "0", "zero", "sinrand", "-", str(reps)], cwd=ROOT, env=env,
                              stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True) for _ in range(copies)]
    rate, nodes, tmax = 0.0, 0, 0.0
    for p in procs:
        out = p.communicate()[0].strip().splitlines()[-1]
        r = json.loads(out)
        rate += r["nodes"] / r["t_with_opt_best_s"]; nodes = r["nodes"]; tmax = max(tmax, r["t_with_opt_best_s"])
    return rate, nodes, tmax


def timed_reference_binary():
    """The reference binary that is TIMED: built -O3 and linked against OpenBLAS (the LP64 one bundled with scipy in
    this image) when that build exists and runs, else the parity build with the naive BLAS shim."""
    blas = os.path.join(ROOT, "oracle", "_ref", "d4b_harness_blas")
    inp = os.path.join(ROOT, "oracle", "inputs", "brick_p3_l2.input")
    if os.path.exists(blas):
        try:
            r = subprocess.run([blas, inp, "none", "0", "zero", "sinrand", "-", "0"], cwd=ROOT, stdout=subprocess.PIPE,
                               stderr=subprocess.DEVNULL, text=True, timeout=120)
            if r.returncode == 0 and r.stdout.strip().startswith("{"):
                return blas, "gcc -O3, OpenBLAS (scipy's LP64 build), 1 BLAS thread per copy"
        except Exception:
            pass
    return os.path.join(ROOT, "oracle", "_ref", "d4b_harness"), "gcc -O3, naive triple-loop BLAS shim (oracle/ref_build/blas_shim.c)"


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
    harness, build = timed_reference_binary()
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
    sample = (f"{cores} concurrent serial copies of the reference ({build}; one per host core; the snapshot has no MPI), each brick 16^3 "
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
    harness, build = timed_reference_binary()
    inp = os.path.join(ROOT, "oracle", "inputs", f"brick_p{args.deg}_l4.input")
    if os.path.exists(harness) and os.path.exists(inp):
        cores = host_cores()
        rate, nodes, _ = run_reference_copies(harness, inp, 8, cores)
        return {"value": rate / 1e9, "unit": UNIT, "cores": cores, "kind": "reference",
                "sample": f"d4est_laplacian_with_opt_apply_aij ({build}), {cores} concurrent serial copies (one per host core; the "
                          f"reference snapshot has no MPI), each brick 16^3 p{args.deg} ({nodes} DoF), best of 9 applies"}
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import fixture, dg_oracle  # noqa: E401
    fx = fixture.load(os.path.join(ROOT, "tests", "golden", "brick_p3_l2.fix"))
    t0 = time.time(); dg_oracle.apply(fx, fx["u"]); t = time.time() - t0
    return {"value": fx["u"].size / t / 1e9, "unit": UNIT, "cores": 1, "kind": "port",
            "sample": "numpy oracle, brick 4^3 p3 (4096 DoF), one apply"}


def morton_order(nx, ny, nz):
    """Element coordinates in the order of d4b_brick_create (Morton key over ix, iy, iz; d4b_brick.cu morton3)."""
    import numpy as np
    ix, iy, iz = np.meshgrid(np.arange(nx), np.arange(ny), np.arange(nz), indexing="ij")
    ix, iy, iz = ix.reshape(-1), iy.reshape(-1), iz.reshape(-1)
    key = np.zeros(ix.size, np.int64)
    for b in range(10):
        key |= ((ix >> b) & 1) << (3 * b) | ((iy >> b) & 1) << (3 * b + 1) | ((iz >> b) & 1) << (3 * b + 2)
    o = np.argsort(key, kind="stable")
    return ix[o], iy[o], iz[o]


def halo_check(d4est_b200, rank, world, deg, n=32):   # 32^3 per slab: above the overlap threshold, so the overlapped exchange is what is checked
    """Correctness of the NCCL data plane inside the bench: every rank applies the operator to ITS slab of a small
    brick n x n x (n world) through d4b_apply_dev (pack kernel + ncclSend/ncclRecv of the mirror planes + element
    kernel) and compares with the same elements of a whole-brick apply computed on its own GPU without any halo.
    u is a deterministic function of the global element / node index, so no data is exchanged outside NCCL."""
    import numpy as np
    import torch
    N3 = (deg + 1) ** 3
    box = (0., 1., 0., 1., 0., float(world))
    whole = d4est_b200.MeshLevel.brick((n, n, n * world), deg=deg, box=box)
    wx, wy, wz = morton_order(n, n, n * world)
    gid = {(int(a), int(b), int(c)): e for e, (a, b, c) in enumerate(zip(wx, wy, wz))}
    idx = np.arange(whole.local_nodes, dtype=np.float64)
    u = np.sin(0.37 * idx) + 0.25 * np.cos(0.011 * idx)
    Au = whole.apply_host(u)
    whole.close()
    slab = d4est_b200.MeshLevel.brick((n, n, n), deg=deg, box=box, rank=rank, nranks=world)
    sx, sy, sz = morton_order(n, n, n)
    el = np.array([gid[(int(a), int(b), int(c) + rank * n)] for a, b, c in zip(sx, sy, sz)])
    take = (el[:, None] * N3 + np.arange(N3)[None, :]).reshape(-1)
    ul = torch.from_numpy(u[take]).cuda(); Al = torch.empty_like(ul)
    slab.apply_dev(ul.data_ptr(), Al.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    ref = Au[take]
    err = float(np.linalg.norm(Al.cpu().numpy() - ref) / np.linalg.norm(ref))
    slab.close()
    return err


def secondary_measurements(d4est_b200, peak_gbs):
    """BASELINE configs 3 and 4 on ONE GPU (not the headline metric): the curved mixed-hp cubed sphere with hanging
    faces (apply + mortar path) and the multigrid-preconditioned flexible CG solve on it, on the hierarchy captured
    from the reference's own coarsening.  The tables are made on this box by the unmodified reference + the product
    glue (oracle/_ref/d4b_mg_harness --capture-only: mesh, reference apply for the right-hand side, hierarchy walk;
    no reference solve -- a reference V-cycle on this mesh takes minutes)."""
    import ctypes as C
    import numpy as np
    import torch
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import fixture, solver_oracle as so  # noqa: E401  (table reader + parameter names only; nothing is computed by the oracle)
    from d4est_b200 import capi
    harness = os.path.join(ROOT, "oracle", "_ref", "d4b_mg_harness")
    inp = "oracle/inputs/cubed_sphere_hp_l2_mg.input"
    if not os.path.exists(harness):
        return {"unavailable": "oracle/_ref/d4b_mg_harness not built"}
    path = "/tmp/d4b_bench_hp_mg.fix"
    t0 = time.time()
    r = subprocess.run([harness, inp, "random_hp", "3", path, "--capture-only"], cwd=ROOT, stdout=subprocess.PIPE,
                       stderr=subprocess.DEVNULL, text=True)
    if r.returncode != 0:
        return {"unavailable": "d4b_mg_harness failed"}
    meta = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    t_tables = time.time() - t0
    fx = fixture.load(path)
    os.remove(path)
    L = d4est_b200.lib()
    nl = int(fx["mg_meta"][0])
    levels = [d4est_b200.MeshLevel(fixture.sub(fx, f"L{l}.")) for l in range(nl)]
    top = levels[-1]; n = top.local_nodes
    workload = (f"cubed_sphere 13 trees, min_level 2, region degrees 2/3/2, deg_quad_inc 1, 3 random-hp steps: "
                f"{top.n_local} elements, {n} DoF, p = {int(fixture.sub(fx, f'L{nl - 1}.')['deg'][:top.n_local].min())}.."
                f"{int(fixture.sub(fx, f'L{nl - 1}.')['deg'][:top.n_local].max())}, hanging faces")
    # --- config 3: apply on the curved hp mesh
    ref = fx["mg_rhs"]     # = the reference's own A u* (d4est_laplacian_with_opt_apply_aij run by the harness)
    parity = float(np.linalg.norm(top.apply_host(fx["mg_ustar"]) - ref) / np.linalg.norm(ref))
    u = torch.rand(n, dtype=torch.float64, device="cuda"); Au = torch.empty_like(u)
    st = torch.cuda.current_stream().cuda_stream
    for _ in range(5):
        top.apply_dev(u.data_ptr(), Au.data_ptr(), st)
    torch.cuda.synchronize()
    reps = 100
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = L.d4b_kernel_launches()
    e0.record()
    for _ in range(reps):
        top.apply_dev(u.data_ptr(), Au.data_ptr(), st)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    split = top.kernel_split
    hp = {"workload": workload, "ms_per_apply": ms, "gdof_per_s": n / ms / 1e6,
          "roofline_frac": top.algorithmic_bytes / (ms * 1e-3) / 1e9 / peak_gbs,
          "algorithmic_GBps": top.algorithmic_bytes / ms / 1e6, "launches_per_apply": int(L.d4b_kernel_launches() - l0) // reps,
          "elements_regular_hp_generic": split, "rel_l2_vs_reference_Au": parity,
          "reference_cpu_apply_seconds_1core": meta.get("seconds_reference_apply_top_level")}
    # --- config 4: MG-preconditioned flexible CG on the captured hierarchy
    trs = [capi.Transfer(fixture.sub(fx, f"T{l}.")) for l in range(nl - 1)]
    par = so.mg_params(fx["mg_params"])
    kw = {k: par[k] for k in ("vcycle_imax", "vcycle_rtol", "vcycle_atol", "cheby_imax", "cheby_eigs_cg_imax",
                              "cheby_eigs_lmax_lmin_ratio", "cheby_eigs_max_multiplier", "eigs_reuse_fromdownvcycle",
                              "eigs_reuse_fromlastvcycle", "use_new_cg_eigs", "bottom_kind", "bottom_imax",
                              "bottom_eigs_cg_imax", "bottom_eigs_lmax_lmin_ratio", "bottom_eigs_max_multiplier",
                              "bottom_atol", "bottom_rtol")}
    mg = capi.Multigrid(levels, trs, **kw)
    rhs = np.ascontiguousarray(fx["mg_rhs"]); it, res = C.c_int(), C.c_double()
    out = {}
    # The flexible-CG update: the reference's in-house driver (d4est_solver_fcg_improved.c) loses the previous search
    # direction through an aliased d4est_linalg_vec_axpyeqz call (see d4b_solver.cu); that arithmetic is reproduced
    # iteration for iteration on the 300-element mesh (tests/test_mg_reference_gpu.py) but stagnates on this one, so the
    # timed solve uses the update its source comments describe (option fcg_textbook_update = what PETSc's KSPFCG does)
    # and the reference-exact variant is reported beside it after a bounded number of iterations.
    capi.check(L.d4b_set_option(b"fcg_textbook_update", 1))
    try:
        for rep in range(2):
            uh = np.zeros(n)
            l0 = L.d4b_kernel_launches(); t0 = time.perf_counter()
            capi.check(L.d4b_mg_fcg_solve_host(mg.handle, uh.ctypes.data_as(C.c_void_p), rhs.ctypes.data_as(C.c_void_p), 200,
                                               par["fcg_atol"], par["fcg_rtol"], C.byref(it), C.byref(res)))
            out = {"seconds": time.perf_counter() - t0, "iterations": it.value, "launches": int(L.d4b_kernel_launches() - l0),
                   "fcg_update": "textbook"}
    finally:
        capi.check(L.d4b_set_option(b"fcg_textbook_update", 0))
    true_res = float(np.linalg.norm(rhs - top.apply_host(uh)) / np.linalg.norm(rhs))
    ux = np.zeros(n); t0 = time.perf_counter()
    capi.check(L.d4b_mg_fcg_solve_host(mg.handle, ux.ctypes.data_as(C.c_void_p), rhs.ctypes.data_as(C.c_void_p), 40,
                                       par["fcg_atol"], par["fcg_rtol"], C.byref(it), C.byref(res)))
    exact = {"iterations": it.value, "seconds": time.perf_counter() - t0,
             "true_relative_residual": float(np.linalg.norm(rhs - top.apply_host(ux)) / np.linalg.norm(rhs)), "imax": 40}
    nodes = [levels[l].local_nodes for l in range(nl)]
    # operator applies of one V-cycle per level (CDS options: cg_eigs 15 + Chebyshev 15, down and up; bottom once):
    per_smooth = (1 + par["cheby_eigs_cg_imax"]) + (par["cheby_imax"] + 1)
    fine_equiv = sum((2 if l > 0 else 1) * per_smooth * nodes[l] / nodes[-1] for l in range(nl))
    t_ref_apply = meta.get("seconds_reference_apply_top_level") or 0.
    mgo = {"workload": workload + f"; {nl} levels {[lv.n_local for lv in levels]} elements; Chebyshev smoother 15 + cg_eigs 15, "
           "Chebyshev bottom solver, flexible CG to rtol 1e-10 (reference-exact arithmetic)",
           **out, "true_relative_residual": true_res, "h2d_d2h": "u and rhs up once, u down once",
           "reference_exact_fcg_update": exact,
           "reference_cpu_seconds_estimate": (out["iterations"] + 1) * (1 + fine_equiv) * t_ref_apply, "reference_cores": 1,
           "reference_estimate_how": (f"(iterations + 1) x (1 + {fine_equiv:.1f} fine-level-equivalent applies per V-cycle) x "
                                      f"{t_ref_apply:.2f} s, the reference's own apply time on this mesh on one core of this box; "
                                      "the reference solve itself was run on the 300-element mesh of tests/test_mg_reference_gpu.py "
                                      "(43 iterations both, see oracle/_ref/fixtures/mg_cubed_sphere_hp1.json)")}
    return {"curved_hp": hp, "mg_fcg": mgo, "tables_seconds": t_tables}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--n", type=int, default=100, help="brick elements per direction (per GPU)")
    ap.add_argument("--deg", type=int, default=3)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-secondary", action="store_true", help="skip the curved-hp apply and MG-FCG solve (N=1 only)")
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
    if os.environ.get("D4B_BENCH_RANK_TIMES"):
        print(f"[bench] rank {rank}: {total_ms / args.steps:.4f} ms per step, median {per[len(per) // 2]:.4f}", file=sys.stderr, flush=True)
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
    capi.check(L.d4b_host_free_pinned(hp)); capi.check(L.d4b_host_free_pinned(hq))

    # one smoother call through host buffers (d4b_cheby_smooth_host: the call the glue installs behind
    # d4est_solver_multigrid_smoother_t.smooth): 2 uploads + 2 downloads around cg_eigs (1 + 15 applies) and the
    # Chebyshev sweep (15 + 1 applies)
    smooth = None
    if world == 1:
        hb = [ctypes.c_void_p() for _ in range(3)]
        for b in hb:
            capi.check(L.d4b_host_alloc_pinned(ndof * 8, ctypes.byref(b)))
        hv = [np.ctypeslib.as_array(ctypes.cast(b, ctypes.POINTER(ctypes.c_double)), shape=(ndof,)) for b in hb]
        hv[1][:] = u.cpu().numpy()
        eig = ctypes.c_double(0.)
        ts = []
        for rep in range(3):
            hv[0][:] = 0.
            torch.cuda.synchronize(); t0 = time.perf_counter()
            capi.check(L.d4b_cheby_smooth_host(lvl.handle, hb[0], hb[1], hb[2], 15, 15, 30., 1., 0, 1, ctypes.byref(eig)))
            ts.append(time.perf_counter() - t0)
        applies = 32
        smooth = {"value": ndof * applies / min(ts) / 1e9, "unit": UNIT, "applies_per_call": applies, "seconds_per_call": min(ts),
                  "h2d_bytes_per_call": 2 * ndof * 8, "d2h_bytes_per_call": 2 * ndof * 8,
                  "what": "d4b_cheby_smooth_host = one d4est smoother call (cg_eigs 15 + Chebyshev 15) on host vectors"}
        for b in hb:
            capi.check(L.d4b_host_free_pinned(b))
    hcheck = None
    if world > 1:
        err = halo_check(d4est_b200, rank, world, args.deg)
        et = torch.tensor([err], dtype=torch.float64, device="cuda")
        dist.all_reduce(et, op=dist.ReduceOp.MAX)
        hcheck = {"max_rel": float(et.item()), "ok": bool(float(et.item()) < 1e-12),
                  "what": "slab apply through the NCCL halo vs the same elements of a whole-brick apply, max over ranks"}

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
                        "steps": e2e_steps, "max_abs_diff_vs_device_path": chk, "per_smoother_call": smooth},
                "gpu_launches": launches, "clocks": clocks}
        if hcheck is not None:
            line["halo_check"] = hcheck
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(args)
        if world == 1 and not args.no_secondary:
            lvl.close()
            try:
                line["secondary"] = secondary_measurements(d4est_b200, peak)
            except Exception as e:   # the headline line must not depend on the secondary run
                line["secondary"] = {"unavailable": f"{type(e).__name__}: {e}"}
        print(json.dumps(line))
    if world > 1:
        lvl.close()
        capi.check(L.d4b_comm_finalize())
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
