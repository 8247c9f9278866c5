"""GPU parity on the shapes of BASELINE.json's configs C3, C4, C5, at sizes the C oracle
finishes in seconds.  (C1/C2 = call_hangup: tests/test_gpu_tool.py and tests/test_gpu_fullsize.py.)"""
import numpy as np
import pytest

from avenir_b200 import synth
from avenir_b200._lib import CRIT_ENTROPY, CRIT_GINI
from avenir_b200.engine import Engine, SplitTable, binary_split_table, grow_levels
from oracle.javaemu import JavaRandom
from tests import helpers

pytestmark = pytest.mark.gpu


def bagging_weights(n, batch, seed):
    """Multiplicities of explore/BaggingSampler.java:101-122 (per batch of `batch` rows, that many uniform
    draws) from a seeded java.util.Random stream -- the same array feeds the engine and the oracle."""
    rng = JavaRandom(seed)
    w = np.zeros(n, dtype=np.int64)
    for s in range(0, n, batch):
        m = min(batch, n - s)
        draws = np.fromiter((int(rng.next_double() * m) for _ in range(m)), dtype=np.int64, count=m)
        w[s:s + m] += np.bincount(draws, minlength=m)
    assert w.sum() == n and w.max() <= 255
    return w.astype(np.uint8)


def compare_level_loops(g, o, table, levels, criterion, n_trees, class_order=None):
    (rec, nrec), frontier = g.grow(levels, criterion, [1] * n_trees, class_order=class_order, max_records=1 << 16)
    ho, fo = grow_levels(o, table, levels, criterion, n_trees=n_trees, class_order=class_order)
    want = [(lv, t, nd, s, pops, av, ties) for lv, level in enumerate(ho) for (t, nd, s, pops, av, ties) in level]
    assert frontier == fo and nrec == len(want)
    for i, (lv, t, nd, s, pops, av, ties) in enumerate(want):
        r = rec[i]
        assert (r.level, r.tree, r.node, r.split, r.n_ties) == (lv, t, nd, s, ties), i
        assert [p for p in r.child_pop if p] == pops and helpers.bits(r.min_av) == helpers.bits(av)
    for t in range(n_trees):
        assert np.array_equal(g.node_ids(t), o.node_ids(t))
    return frontier


def test_c3_forest_shape():
    """C3: bagged forest, 32 attributes (24 int with 16 bins + 8 categorical with 4 levels -> 416 candidate
    splits), 4 classes, gini; parity subset T = 5 trees, 300 k rows, 6 levels, BaggingSampler multiplicities."""
    n, T = 300_000, 5
    n_bins = [16] * 24 + [4] * 8
    cat = [False] * 24 + [True] * 8
    cols, cls = synth.binned_dataset(n, n_bins, 4, seed=33, signal_attrs=6)
    table = binary_split_table(n_bins, cat)
    assert len(table) == 416 and sum(n_bins) == 416
    weights = [bagging_weights(n, 10_000, 100 + t) for t in range(T)]
    g, o = Engine(0), helpers.OracleEngine(threads=8)
    for e in (g, o):
        e.load_dataset(cols, cls, n_bins, 4)
        e.forest_init(T, weights, None)
        e.set_splits(table)
    frontier = compare_level_loops(g, o, table, 6, CRIT_GINI, T, class_order=[0, 1, 2, 3])
    assert all(1 < f <= 64 for f in frontier)
    st = g.stats()
    assert st.last_hist_variant in (0, 2)     # shared-memory paths: by rows (attribute groups) or the forest kernel


def test_c5_wide_shape():
    """C5: 512 int attributes x 64 bins (63 splits each = 32 256 candidates), gini, deep tree: the frontier
    outgrows shared memory: hist_rot_kernel at the shallow levels, then rows in node-contiguous order (hist_sorted_kernel);
    the same loop with that kernel switched off falls back to hist_smem_kernel / warp-aggregated global atomics."""
    n = 60_000
    n_bins = [64] * 512
    cols, cls = synth.binned_dataset(n, n_bins, 2, seed=55, signal_attrs=16)
    table = binary_split_table(n_bins, [False] * 512)
    assert len(table) == 512 * 63
    g, o = Engine(0), helpers.OracleEngine(threads=8)
    for e in (g, o):
        e.load_dataset(cols, cls, n_bins, 2)
        e.forest_init(1)
        e.set_splits(table)
    frontier = compare_level_loops(g, o, table, 10, CRIT_GINI, 1)
    assert frontier[0] > 400                  # beyond (200 KB / 64 bins / 2 classes / 4 B) nodes
    assert g.stats().last_hist_variant == 4


def test_c5_wide_shape_global_fallback(monkeypatch):
    monkeypatch.setenv("AVT_HIST_SORTED", "0")
    n = 40_000
    n_bins = [64] * 512
    cols, cls = synth.binned_dataset(n, n_bins, 2, seed=56, signal_attrs=16)
    table = binary_split_table(n_bins, [False] * 512)
    g, o = Engine(0), helpers.OracleEngine(threads=8)
    for e in (g, o):
        e.load_dataset(cols, cls, n_bins, 2)
        e.forest_init(1)
        e.set_splits(table)
    frontier = compare_level_loops(g, o, table, 10, CRIT_GINI, 1)
    assert frontier[0] > 400 and g.stats().last_hist_variant == 1


@pytest.mark.parametrize("criterion", [CRIT_GINI, CRIT_ENTROPY])
def test_c4_cpg_high_cardinality_shape(criterion):
    """C4: ClassPartitionGenerator sweep, high-cardinality categoricals (1 000 levels, uint16 codes) with explicit
    candidate lists (prefix splits in cardinality order + one-vs-rest; exhaustive set partitions are impossible
    at this cardinality) and int attributes with bucketWidth grids and maxSplit 3."""
    n = 100_000
    n_cat, n_int = 6, 4
    n_bins = [1000] * n_cat + [20] * n_int
    cols, cls = synth.binned_dataset(n, n_bins, 2, seed=44, signal_attrs=3)
    # Zipf-like skew on the categorical codes
    rng = np.random.default_rng(4)
    for a in range(n_cat):
        cols[a] = np.minimum((rng.zipf(1.1, n) - 1) % 1000, 999).astype(np.uint16)
    table = SplitTable()
    for a in range(n_cat):
        for k in range(0, 999, 9):
            table.add_lut(a, [0 if b <= k else 1 for b in range(1000)])
        for k in range(0, 1000, 13):
            table.add_lut(a, [0 if b == k else 1 for b in range(1000)])
    for a in range(n_cat, n_cat + n_int):
        for i in range(19):
            table.add_numeric(a, [i])
            for j in range(i + 1, 19):
                table.add_numeric(a, [i, j])
    g, o = Engine(0), helpers.OracleEngine(threads=8)
    for e in (g, o):
        e.load_dataset(cols, cls, n_bins, 2)
        e.forest_init(1)
        e.set_splits(table)
        e.histogram([1])
    sg, so = g.cpg_sweep(0, criterion, [1, 0]), o.cpg_sweep(0, criterion, [1, 0])
    for s in range(len(table)):
        assert helpers.bits(sg[s].stat) == helpers.bits(so[s].stat), s
        assert helpers.bits(sg[s].split_info) == helpers.bits(so[s].split_info), s
    # gain ratio as the reducer prints it (explore/ClassPartitionGenerator.java:533-538) is a function of these two
    parent = 0.9
    with np.errstate(divide="ignore", invalid="ignore"):
        ratio_g = (parent - np.array([sg[s].stat for s in range(len(table))])) / np.array([sg[s].split_info for s in range(len(table))])
        ratio_o = (parent - np.array([so[s].stat for s in range(len(table))])) / np.array([so[s].split_info for s in range(len(table))])
    assert np.array_equal(ratio_g.view(np.uint64), ratio_o.view(np.uint64))
