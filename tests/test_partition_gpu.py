"""N-rank layout on the GPU through the C ABI: rank-local tables + ghost values == serial apply."""
import os
import subprocess
import sys

import numpy as np
import pytest

import fixture
import d4est_b200
from d4est_b200 import partition as P

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True, params=["two_pass", "single_pass"])
def hp_mode(request):
    """Levels with neighbour ranks take the two-pass hp kernels only from 4 096 elements up by default; these small
    rank-local levels are run both ways: two-pass for every element whose neighbours are local (the elements along
    the rank boundary stay single-pass and feed traces to their two-pass neighbours), and single-pass throughout."""
    L = d4est_b200.lib()
    d4est_b200.capi.check(L.d4b_set_option(b"two_pass_min_elements", 0 if request.param == "two_pass" else 1 << 30))
    os.environ["D4B_TP_MIN"] = "0" if request.param == "two_pass" else str(1 << 30)   # for the NCCL worker processes
    yield request.param
    d4est_b200.capi.check(L.d4b_set_option(b"two_pass_min_elements", -1))
    del os.environ["D4B_TP_MIN"]


def _fix(name, golden_dir, ref_fixture_dir):
    for d in (golden_dir, ref_fixture_dir):
        p = os.path.join(d, name + ".fix")
        if os.path.exists(p):
            return fixture.load(p)
    pytest.skip(f"fixture {name} not built")


@pytest.mark.parametrize("name,nranks", [("brick_p3_l2", 2), ("brick_p3_l3", 3), ("cubed_sphere_hp1", 4),
                                         ("cubed_sphere_hp2", 3)])
def test_n_rank_equals_reference(name, nranks, golden_dir, ref_fixture_dir):
    """Every rank's Au (ghost values supplied through d4b_set_ghost_host, so the element kernels read
    neighbours from u_ghost and the regular-element kernel folds with its own a+) equals the reference's Au."""
    fx = _fix(name, golden_dir, ref_fixture_dir)
    ref, u = fx["Au_ref_with_opt"], fx["u"]
    for me in P.partition(fx, nranks):
        lvl = d4est_b200.MeshLevel(me)
        assert lvl.ghost_nodes == P.ghost_vector(fx, me, u).size
        lvl.set_ghost(P.ghost_vector(fx, me, u))
        Au = lvl.apply_host(P.local_vector(fx, me, u))
        a, b = me["node_range"]
        err = np.linalg.norm(Au - ref[a:b]) / np.linalg.norm(ref[a:b])
        lvl.close()
        assert err < 1e-12, (name, nranks, err)


def test_nccl_world_size_2(golden_dir):
    """Two ranks, two GPUs: NCCL halo (pack kernel + ncclSend/Recv) inside d4b_apply_host."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29519", os.path.join(ROOT, "tests", "mp_halo_worker.py"),
           os.path.join(golden_dir, "cubed_sphere_hp1.fix"), "gpu"]
    r = subprocess.run(cmd, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:]
    print(r.stdout[-1500:])


def test_nccl_world_size_2_krylov_reductions(golden_dir):
    """Two ranks, two GPUs: CG and flexible CG on the partitioned brick (halo over ncclSend/Recv inside every apply,
    dots over ncclAllReduce on device scalars) reproduce the 1-rank iterates and the 1-rank iteration count."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29521", os.path.join(ROOT, "tests", "mp_halo_worker.py"),
           os.path.join(golden_dir, "brick_p3_l2.fix"), "gpu", "krylov"]
    r = subprocess.run(cmd, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:]
    print(r.stdout[-1500:])


def _morton_order(nx, ny, nz):
    """Element coordinates in the generator's order (Morton key over (ix, iy, iz), d4b_brick.cu morton3)."""
    ix, iy, iz = np.meshgrid(np.arange(nx), np.arange(ny), np.arange(nz), indexing="ij")
    ix, iy, iz = ix.reshape(-1), iy.reshape(-1), iz.reshape(-1)
    key = np.zeros(ix.size, np.int64)
    for b in range(10):
        key |= ((ix >> b) & 1) << (3 * b) | ((iy >> b) & 1) << (3 * b + 1) | ((iz >> b) & 1) << (3 * b + 2)
    o = np.argsort(key, kind="stable")
    return ix[o], iy[o], iz[o]


@pytest.mark.parametrize("n,nranks,deg", [((4, 4, 4), 2, 3), ((6, 4, 2), 3, 3), ((4, 4, 2), 2, 2), ((24, 24, 16), 2, 3)])
def test_brick_slabs_equal_whole_brick(n, nranks, deg):
    """d4b_brick_create_slab: the slabs of nx x ny x (nz*nranks), fed with the right ghost planes, reproduce the
    apply of the whole brick (exercises ghost neighbours in the regular-element kernels)."""
    nx, ny, nz = n
    N3 = (deg + 1) ** 3
    box = (0., 1., 0., 1., 0., float(nranks))
    whole = d4est_b200.MeshLevel.brick((nx, ny, nz * nranks), deg=deg, box=box)
    rng = np.random.default_rng(7)
    u = rng.uniform(-1, 1, whole.local_nodes)
    Au = whole.apply_host(u)
    wx, wy, wz = _morton_order(nx, ny, nz * nranks)
    gid = {(int(a), int(b), int(c)): e for e, (a, b, c) in enumerate(zip(wx, wy, wz))}
    sx, sy, sz = _morton_order(nx, ny, nz)
    for r in range(nranks):
        slab = d4est_b200.MeshLevel.brick((nx, ny, nz), deg=deg, box=box, rank=r, nranks=nranks)
        el = [gid[(int(a), int(b), int(c) + r * nz)] for a, b, c in zip(sx, sy, sz)]
        ul = np.concatenate([u[e * N3:(e + 1) * N3] for e in el])
        planes = ([r * nz - 1] if r > 0 else []) + ([(r + 1) * nz] if r < nranks - 1 else [])
        ug = np.concatenate([u[gid[(i, j, z)] * N3:(gid[(i, j, z)] + 1) * N3] for z in planes
                             for j in range(ny) for i in range(nx)]) if planes else np.zeros(0)
        assert slab.ghost_nodes == ug.size
        slab.set_ghost(ug)
        Al = slab.apply_host(ul)
        ref = np.concatenate([Au[e * N3:(e + 1) * N3] for e in el])
        assert np.linalg.norm(Al - ref) <= 1e-13 * np.linalg.norm(ref), (r, np.linalg.norm(Al - ref) / np.linalg.norm(ref))
        if slab.n_local >= 8192:
            # pinned caller buffers: the host pipeline with the late launch for the elements next to a ghost
            import ctypes
            from d4est_b200 import capi
            L = capi.lib()
            hp, hq = ctypes.c_void_p(), ctypes.c_void_p()
            capi.check(L.d4b_host_alloc_pinned(ul.size * 8, ctypes.byref(hp)))
            capi.check(L.d4b_host_alloc_pinned(ul.size * 8, ctypes.byref(hq)))
            try:
                hu = np.ctypeslib.as_array(ctypes.cast(hp, ctypes.POINTER(ctypes.c_double)), shape=(ul.size,))
                hA = np.ctypeslib.as_array(ctypes.cast(hq, ctypes.POINTER(ctypes.c_double)), shape=(ul.size,))
                hu[:] = ul
                hA[:] = np.nan
                slab.apply_host(hu, out=hA)
                assert np.linalg.norm(hA - ref) <= 1e-13 * np.linalg.norm(ref)
            finally:
                capi.check(L.d4b_host_free_pinned(hp)); capi.check(L.d4b_host_free_pinned(hq))
        slab.close()
    whole.close()
