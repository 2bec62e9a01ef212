"""Host-side logic of the N>1 path on CPU: rank-local tables, halo lists, gloo exchange."""
import os
import subprocess
import sys

import numpy as np
import pytest

import fixture
import dg_oracle
from d4est_b200 import partition as P

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def brick(golden_dir):
    return fixture.load(os.path.join(golden_dir, "brick_p3_l2.fix"))


@pytest.mark.parametrize("nranks", [2, 3])
def test_rank_local_apply_equals_serial(brick, nranks):
    """oracle(rank-local tables, halo of u) == oracle(serial tables) on every rank's node range."""
    fx = brick
    ref = dg_oracle.apply(fx, fx["u"])
    seen = 0
    for me in P.partition(fx, nranks):
        a, b = me["node_range"]
        Au = dg_oracle.apply(me, P.local_vector(fx, me, fx["u"]), P.ghost_vector(fx, me, fx["u"]))
        assert np.linalg.norm(Au - ref[a:b]) <= 1e-13 * np.linalg.norm(ref[a:b])
        seen += b - a
    assert seen == fx["u"].size


def test_rank_local_twin_trace_apply_equals_reference(golden_dir):
    """The two-pass formulation on rank-local tables of the curved hp mesh (3 ranks): records whose neighbour is a
    ghost have no local twin and take the neighbour side directly from the halo (on the device: the single-pass
    kernel), every other interior record takes it from the twin's trace; the result is the reference's Au."""
    fx = fixture.load(os.path.join(golden_dir, "cubed_sphere_hp1.fix"))
    ref = fx["Au_ref_with_opt"]
    n_twin = n_ghost_side = 0
    for me in P.partition(fx, 3):
        a, b = me["node_range"]
        nl = int(me["hdr"][0])
        tw = dg_oracle.twin_records(me)
        nbr = me["hm"]["nbr"][:int(me["hm_begin"][nl])]
        assert ((tw >= 0) == ((nbr >= 0) & (nbr < nl))).all()
        n_twin += int((tw >= 0).sum()); n_ghost_side += int((nbr >= nl).sum())
        Au = dg_oracle.apply(me, P.local_vector(fx, me, fx["u"]), P.ghost_vector(fx, me, fx["u"]), twin_traces=True)
        assert np.linalg.norm(Au - ref[a:b]) <= 1e-12 * np.linalg.norm(ref[a:b])
    assert n_twin > 0 and n_ghost_side > 0


def test_halo_lists_are_consistent(golden_dir):
    """What rank r sends to p is exactly, and in order, what p holds as ghosts owned by r (curved hp mesh)."""
    fx = fixture.load(os.path.join(golden_dir, "cubed_sphere_hp1.fix"))
    parts = P.partition(fx, 4)
    for r, me in enumerate(parts):
        assert list(me["peer_rank"]) == sorted(set(me["peer_rank"])) and r not in me["peer_rank"]
        for p, peer in enumerate(me["peer_rank"]):
            other = parts[int(peer)]
            q = list(other["peer_rank"]).index(r)
            sent = [me["owned"][0] + int(e) for e in me["send_elem"][me["send_begin"][p]:me["send_begin"][p + 1]]]
            held = list(other["ghost_gid"][other["recv_begin"][q]:other["recv_begin"][q + 1]])
            assert sent == held and len(sent) > 0
        nl = int(me["hdr"][0])
        assert np.all(me["hm"]["nbr"] < nl + int(me["hdr"][1]))
        u = fx["u"]
        # packed mirrors land in the peer's ghost slice
        for p, peer in enumerate(me["peer_rank"]):
            other = parts[int(peer)]; q = list(other["peer_rank"]).index(r)
            a, b = P.recv_slice(other, q)
            assert np.array_equal(P.pack_mirrors(me, P.local_vector(fx, me, u), p), P.ghost_vector(fx, other, u)[a:b])


def test_gloo_world_size_2(golden_dir):
    """Two processes, gloo: halo exchange + rank-local oracle apply == serial oracle apply."""
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", OMP_NUM_THREADS="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29517", os.path.join(ROOT, "tests", "mp_halo_worker.py"),
           os.path.join(golden_dir, "brick_p3_l2.fix"), "cpu"]
    r = subprocess.run(cmd, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:]
    assert "rank 0/2" in r.stdout and "rank 1/2" in r.stdout
