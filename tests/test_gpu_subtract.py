"""Sibling subtraction (AVT_SUBTRACT=1): the level loop counts only the smaller children of every split and
derives the most populous one as parent - siblings.  Exact in integers, so every record and node id must
still equal the oracle's.  Off by default (measured neutral in round 1); kept green here."""
import numpy as np
import pytest

from avenir_b200 import synth
from avenir_b200._lib import CRIT_ENTROPY, CRIT_GINI
from avenir_b200.engine import Engine, binary_split_table, grow_levels
from tests import helpers

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("shape", ["call_hangup", "forest"])
def test_subtraction_gives_the_same_trees(shape, monkeypatch):
    monkeypatch.setenv("AVT_SUBTRACT", "1")
    rng = np.random.default_rng(9)
    if shape == "call_hangup":
        n, T, crit = 250_003, 1, CRIT_ENTROPY
        cols, cls, n_bins, _ = synth.call_hangup_encoded(n)
        cat, C, weights = [True, True, True, False], 2, None
    else:
        n, T, crit = 120_000, 3, CRIT_GINI
        n_bins = [16] * 10 + [4] * 4
        cat, C = [False] * 10 + [True] * 4, 4
        cols, cls = synth.binned_dataset(n, n_bins, C, seed=12, signal_attrs=5)
        weights = [rng.poisson(1.0, n).clip(0, 255).astype(np.uint8) for _ in range(T)]
    table = binary_split_table(n_bins, cat)
    g, o = Engine(0), helpers.OracleEngine(threads=8)
    for e in (g, o):
        e.load_dataset(cols, cls, n_bins, C)
        e.forest_init(T, weights, None)
        e.set_splits(table)
    (rec, nrec), frontier = g.grow(7, crit, [1] * T, max_records=1 << 15)
    ho, fo = grow_levels(o, table, 7, crit, n_trees=T)
    want = [(lv, t, nd, s, pops, av, ties) for lv, level in enumerate(ho) for (t, nd, s, pops, av, ties) in level]
    assert frontier == fo and nrec == len(want)
    for i, (lv, t, nd, s, pops, av, ties) in enumerate(want):
        r = rec[i]
        assert (r.level, r.tree, r.node, r.split, r.n_ties) == (lv, t, nd, s, ties), i
        assert [p for p in r.child_pop if p] == pops and helpers.bits(r.min_av) == helpers.bits(av)
    for t in range(T):
        assert np.array_equal(g.node_ids(t), o.node_ids(t))
