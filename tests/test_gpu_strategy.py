"""Candidate-attribute strategies inside avt_grow_levels_ex (getSplitAttributes, tree/DecisionTreeBuilder.java:365-381;
tree/SplitManager.java:149-166 `all`, :195-232 `notUsedYet`) against the oracle-backed level loop that applies the
same per-node candidate sets on the host: records, av bits and final node ids."""
import numpy as np
import pytest

from avenir_b200 import _lib, synth
from avenir_b200._lib import CRIT_ENTROPY, CRIT_GINI
from avenir_b200.engine import Engine, binary_split_table, grow_levels
from tests import helpers

pytestmark = pytest.mark.gpu


def compare(g, o, table, levels, criterion, T, strategy, root_mask=None, class_order=None):
    (rec, nrec), frontier = g.grow(levels, criterion, [1] * T, class_order=class_order, max_records=1 << 14, strategy=strategy,
                                   root_cand_mask=root_mask)
    ho, fo = grow_levels(o, table, levels, criterion, n_trees=T, class_order=class_order, strategy=strategy, root_cand_mask=root_mask)
    want = [(lv, t, nd, s, pops, av, ties) for lv, level in enumerate(ho) for (t, nd, s, pops, av, ties) in level]
    assert frontier == fo and nrec == len(want)
    for i, (lv, t, nd, s, pops, av, ties) in enumerate(want):
        r = rec[i]
        assert (r.level, r.tree, r.node, r.split, r.n_ties) == (lv, t, nd, s, ties), i
        assert [p for p in r.child_pop if p] == pops and helpers.bits(r.min_av) == helpers.bits(av)
    for t in range(T):
        assert np.array_equal(g.node_ids(t), o.node_ids(t))
    return want, frontier


def test_not_used_yet_on_call_hangup_stops_after_four_levels():
    """detr.properties' default strategy on the tutorial schema: 4 attributes, so no node can be expanded a fifth time
    (SURVEY 8d, C2 note); the item-parallel scan kernel path."""
    n = 120_011
    cols, cls, n_bins, _ = synth.call_hangup_encoded(n)
    table = binary_split_table(n_bins, [True, True, True, False])
    g, o = Engine(0), helpers.OracleEngine(threads=4)
    for e in (g, o):
        e.load_dataset(cols, cls, n_bins, 2)
        e.forest_init(2, [None, np.random.default_rng(3).poisson(1.0, n).clip(0, 255).astype(np.uint8)], None)
        e.set_splits(table)
    want, frontier = compare(g, o, table, 6, CRIT_ENTROPY, 2, _lib.ATTRS_NOT_USED_YET)
    assert max(lv for lv, *_ in want) == 3 and frontier == [0, 0]
    # along every recorded path no attribute repeats
    g.close()


@pytest.mark.parametrize("strategy", [_lib.ATTRS_NOT_USED_YET, _lib.ATTRS_ALL])
def test_strategies_on_a_forest_schema_with_a_custom_base_set(strategy):
    """rafo.properties shape (general scan kernel, forest histogram kernel), 40 attributes = two mask words, with a
    dtb.custom.base.attributes-style root set that leaves out some attributes."""
    n, T = 80_003, 3
    n_bins = [8] * 30 + [4] * 10
    cat = [False] * 30 + [True] * 10
    cols, cls = synth.binned_dataset(n, n_bins, 3, seed=77, signal_attrs=5)
    table = binary_split_table(n_bins, cat)
    root = np.zeros((T, 2), dtype=np.uint32)
    root[:, 0] = 0xFFFFFFFF & ~(1 << 1)      # attribute 1 (a signal attribute) is not offered
    root[:, 1] = 0x0F                        # attributes 32-35 only
    g, o = Engine(0), helpers.OracleEngine(threads=8)
    for e in (g, o):
        e.load_dataset(cols, cls, n_bins, 3)
        e.forest_init(T, None, None)
        e.set_splits(table)
    want, _ = compare(g, o, table, 5, CRIT_GINI, T, strategy, root_mask=root, class_order=[1, 0, 2])
    used = {table.attr[s] for (_, _, _, s, _, _, _) in want}
    assert 1 not in used and not any(36 <= a < 40 for a in used)
    g.close()
