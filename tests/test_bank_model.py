"""tools/fast_bank_sim.py (the shared-memory bank-conflict model that chose the paddings of the line-task kernel,
d4est_b200/csrc/d4b_fast_pads.inc): the committed paddings are what the model says they are.  No GPU needed."""
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import fast_bank_sim as sim  # noqa: E402


def _pads():
    txt = open(os.path.join(ROOT, "d4est_b200", "csrc", "d4b_fast_pads.inc")).read()
    for m in re.finditer(r"^FAST_PAD\(([^)]*)\) /\* modelled wavefronts (\d+) -> (\d+) \(ideal (\d+)\) \*/", txt, re.M):
        v = [int(x) for x in m.group(1).split(",")]
        yield v, int(m.group(2)), int(m.group(3)), int(m.group(4))


def test_committed_paddings_reproduce_the_modelled_counts():
    rows = list(_pads())
    assert len(rows) >= 10
    for v, before, after, ideal in rows:
        N, Nq = v[0], v[1]
        t0, i0 = sim.simulate(sim.Layout(N, Nq))
        t1, i1 = sim.simulate(sim.Layout(N, Nq, **sim.current_pads(N, Nq)))
        assert (t0, i0) == (before, ideal), (N, Nq)
        assert (t1, i1) == (after, ideal), (N, Nq)
        assert after < before


def test_padded_layouts_keep_the_resident_ctas():
    """a padding must not cost a resident CTA: shared memory allows at least min(4, what the default strides allow) CTAs
    per SM (4 = the register limit of the 128-register instances)"""
    for v, _, _, _ in _pads():
        L0 = sim.Layout(v[0], v[1])
        L1 = sim.Layout(v[0], v[1], **sim.current_pads(v[0], v[1]))
        ctas = lambda L: (227 * 1024) // (L.PER * L.EPB * 8)
        assert ctas(L1) >= min(4, ctas(L0)), (v[0], v[1], ctas(L0), ctas(L1))
