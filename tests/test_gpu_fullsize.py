"""GPU parity at BASELINE.json's full C2 size: exact equality with the C oracle on the whole 10 M-row, 8-level build
(records, av bits, every row's final node; the oracle's histogram formulation needs well under a second per level
on the host cores), plus the size-independent properties: conservation of rows through every split, linearity of
the histogram over row shards, idempotence of the level loop, agreement of the final node ids with the recorded tree."""
import os
import numpy as np
import pytest

from avenir_b200 import synth
from avenir_b200._lib import AVT_NODE_NONE, CRIT_ENTROPY, CRIT_GINI
from avenir_b200.engine import Engine, binary_split_table, grow_levels
from tests import helpers

pytestmark = pytest.mark.gpu

N_FULL = 10_000_000   # config C2


@pytest.fixture(scope="module")
def c2():
    cols, cls, n_bins, names = synth.call_hangup_encoded(N_FULL)
    return cols, cls, n_bins, binary_split_table(n_bins, [True, True, True, False])


def test_c2_full_size_properties(c2):
    cols, cls, n_bins, table = c2
    eng = Engine(0)
    eng.load_dataset(cols, cls, n_bins, 2)
    eng.set_splits(table)
    eng.forest_init(1)
    # histogram: every attribute's bins partition the rows; class totals agree with the data
    h = eng.histogram([1], fetch=True)[0]
    off = np.concatenate([[0], np.cumsum(n_bins)])
    for a in range(4):
        assert int(h[off[a]:off[a + 1]].sum()) == N_FULL
        assert np.array_equal(h[off[a]:off[a + 1]].sum(axis=0), np.bincount(cls, minlength=2))
        assert np.array_equal(h[off[a]:off[a + 1], 0] + h[off[a]:off[a + 1], 1], np.bincount(cols[a], minlength=n_bins[a]))
    # level loop, 8 expansions (the bench step)
    (rec, nrec), frontier = eng.grow(8, CRIT_ENTROPY, [1], max_records=2048)
    recs = [rec[i] for i in range(nrec)]
    pops = {(0, 0): N_FULL}
    for r in recs:
        kids = [(r.child[s], r.child_pop[s]) for s in range(8) if r.child[s] >= 0]
        assert sum(p for _, p in kids) == pops[(r.level, r.node)]          # rows are conserved by every split
        assert all(p > 0 for _, p in kids) and r.min_av >= 0.0
        for c, p in kids:
            pops[(r.level + 1, c)] = p
    leaves = {c: p for (lvl, c), p in pops.items() if lvl == 8}
    assert len(leaves) == frontier[0] and sum(leaves.values()) == N_FULL
    nid = eng.node_ids(0)
    assert np.array_equal(np.bincount(nid, minlength=frontier[0]), np.array([leaves[c] for c in range(frontier[0])]))
    assert (nid != AVT_NODE_NONE).all()
    # entropy never increases from a node to the weighted average of its children (concavity)
    stat = {(0, 0): None}
    for r in recs:
        for s in range(8):
            if r.child[s] >= 0:
                stat[(r.level + 1, r.child[s])] = r.child_stat[s]
    for r in recs:
        parent = stat.get((r.level, r.node))
        if parent is not None:
            assert r.min_av <= parent + 1e-12
    # idempotence: the same build again gives the same tree and the same row assignment
    eng.forest_init(1)
    (rec2, nrec2), frontier2 = eng.grow(8, CRIT_ENTROPY, [1], max_records=2048)
    assert nrec2 == nrec and frontier2 == frontier
    for i in range(nrec):
        assert (rec2[i].split, list(rec2[i].child_pop), rec2[i].min_av) == (rec[i].split, list(rec[i].child_pop), rec[i].min_av)
    assert np.array_equal(eng.node_ids(0), nid)


def test_c2_histogram_is_linear_over_row_shards(c2):
    cols, cls, n_bins, table = c2
    rng = np.random.default_rng(3)
    nid = rng.integers(0, 37, size=N_FULL).astype(np.uint16)
    whole = Engine(0)
    whole.load_dataset(cols, cls, n_bins, 2)
    whole.forest_init(1, None, [nid])
    h = whole.histogram([37], fetch=True)
    whole.close()
    parts = np.zeros_like(h)
    for lo, hi in ((0, 3_333_333), (3_333_333, 7_000_001), (7_000_001, N_FULL)):
        e = Engine(0)
        e.load_dataset([c[lo:hi] for c in cols], cls[lo:hi], n_bins, 2)
        e.forest_init(1, None, [nid[lo:hi]])
        parts += e.histogram([37], fetch=True)
        e.close()
    assert np.array_equal(parts, h)


@pytest.mark.parametrize("criterion,n", [(CRIT_ENTROPY, N_FULL), (CRIT_GINI, N_FULL), (CRIT_GINI, 1_000_000)])
def test_c2_full_build_equals_oracle(c2, criterion, n):
    """The bench workload itself (C2: 10 M rows, 8 expansions), bit for bit against the oracle."""
    cols, cls, n_bins, table = c2
    g, o = Engine(0), helpers.OracleEngine(threads=os.cpu_count() or 8)
    for e in (g, o):
        e.load_dataset([c[:n] for c in cols], cls[:n], n_bins, 2)
        e.forest_init(1)
        e.set_splits(table)
    (rec, nrec), frontier = g.grow(8, criterion, [1], max_records=2048)
    ho, fo = grow_levels(o, table, 8, criterion)
    want = [x for level in ho for x in level]
    assert frontier == fo and nrec == len(want)
    for i, (t, nd, s, pops, av, ties) in enumerate(want):
        assert (rec[i].node, rec[i].split, rec[i].n_ties) == (nd, s, ties)
        assert [p for p in rec[i].child_pop if p] == pops and helpers.bits(rec[i].min_av) == helpers.bits(av)
    assert np.array_equal(g.node_ids(0), o.node_ids(0))
