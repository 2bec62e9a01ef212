"""Host-side logic of the N-rank multigrid (d4est_b200.partition.partition_hierarchy) on the hierarchy captured from
the reference's own coarsening: rank-local transfer tables + rank-local vectors reproduce the serial restriction /
prolongation (they stay rank-local because families are never split), levels tile the serial level, halo lists are
consistent, and the rank-local oracle apply with ghost values equals the serial one on every level.  CPU only."""
import os
import numpy as np
import pytest
import fixture
import dg_oracle
import solver_oracle as so
from d4est_b200 import partition as P

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FIX = os.path.join(ROOT, "oracle", "_ref", "fixtures", "mg_cubed_sphere_hp1.fix")


@pytest.mark.parametrize("nranks", [2, 3])
def test_partitioned_hierarchy_reproduces_serial_transfers_and_applies(nranks):
    if not os.path.exists(FIX):
        pytest.skip("multigrid fixture not built")
    fx = fixture.load(FIX)
    nl = int(fx["mg_meta"][0])
    levels = [fixture.sub(fx, f"L{l}.") for l in range(nl)]
    tabs = [fixture.sub(fx, f"T{l}.") for l in range(nl - 1)]
    ranks = P.partition_hierarchy(levels, tabs, nranks, base=0)
    # levels tile the serial ones
    for li in range(nl):
        assert sum(int(ranks[r][0][li]["hdr"][0]) for r in range(nranks)) == int(levels[li]["hdr"][0])
        assert ranks[0][0][li]["node_range"][0] == 0
        for r in range(1, nranks):
            assert ranks[r][0][li]["node_range"][0] == ranks[r - 1][0][li]["node_range"][1]
    # transfers are rank-local and equal the serial ones on the rank's node ranges
    for l in range(nl - 1):
        T = tabs[l]
        ser_r = so.transfer(T, T["probe_fine"], T["probe_coarse"].size, True)
        ser_p = so.transfer(T, T["probe_coarse"], T["probe_fine"].size, False)
        for r in range(nranks):
            lv, tr = ranks[r]
            (c0, c1), (f0, f1) = lv[l]["node_range"], lv[l + 1]["node_range"]
            rr = so.transfer(tr[l], T["probe_fine"][f0:f1], c1 - c0, True)
            assert np.linalg.norm(rr - ser_r[c0:c1]) <= 1e-14 * np.linalg.norm(ser_r)
            pp = so.transfer(tr[l], T["probe_coarse"][c0:c1], f1 - f0, False)
            assert np.linalg.norm(pp - ser_p[f0:f1]) <= 1e-14 * np.linalg.norm(ser_p)
    # the operator of a coarse level, rank by rank with ghost values, equals the serial apply
    l = 1
    u = np.sin(0.7 * np.arange(levels[l]["node_off"][int(levels[l]["hdr"][0])]))
    ser = dg_oracle.apply(levels[l], u)
    for r in range(nranks):
        me = ranks[r][0][l]
        a, b = me["node_range"]
        loc = dg_oracle.apply(me, P.local_vector(levels[l], me, u), P.ghost_vector(levels[l], me, u))
        assert np.linalg.norm(loc - ser[a:b]) <= 1e-13 * np.linalg.norm(ser)
