"""hist_forest_kernel (avt_forest.cuh; lane = attribute, conflict-free [bin][attribute] counters, rows that are not
counted skipped by whole warps) against the C oracle: raw counts on ragged shapes, and whole level loops with the
fused repartition and sibling subtraction -- records, av bits and every row's final node."""
import os

import numpy as np
import pytest

from avenir_b200 import synth
from avenir_b200._lib import AVT_NODE_NONE, CRIT_ENTROPY, CRIT_GINI
from avenir_b200.engine import Engine, binary_split_table
from tests import helpers
from tests.test_gpu_configs import bagging_weights, compare_level_loops

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n,n_bins,n_classes,n_trees,n_nodes", [
    (100_003, [16] * 24 + [4] * 8, 4, 3, 5),      # C3 schema, ragged row count, several nodes
    (33, [7] * 16, 2, 1, 1),                      # two tiles, the second with one row
    (50_000, [5, 9, 2] * 7 + [30], 3, 2, 3),      # 22 attributes: lanes without an attribute
    (70_001, [12] * 40, 2, 2, 2),                 # 40 attributes: two attribute chunks
    (20_000, [200] * 16, 2, 1, 1),                # many bins per attribute
])
def test_forest_histogram_equals_oracle(n, n_bins, n_classes, n_trees, n_nodes):
    cols, cls = synth.binned_dataset(n, n_bins, n_classes, seed=5, signal_attrs=3)
    rng = np.random.default_rng(n)
    nid = []
    for t in range(n_trees):
        x = rng.integers(0, n_nodes, size=n).astype(np.uint16)
        x[rng.random(n) < 0.07] = AVT_NODE_NONE
        nid.append(x)
    weights = [None if t == 1 else rng.poisson(1.0, n).clip(0, 255).astype(np.uint8) for t in range(n_trees)]
    weights[0][:7] = [255, 0, 3, 0, 0, 1, 200]
    g, o = Engine(0), helpers.OracleEngine(threads=4)
    for e in (g, o):
        e.load_dataset(cols, cls, n_bins, n_classes)
        e.forest_init(n_trees, weights, nid)
    g.set_hist_variant(2)
    hg = g.histogram([n_nodes] * n_trees, fetch=True)
    assert g.stats().last_hist_variant == 2
    assert np.array_equal(hg, o.histogram([n_nodes] * n_trees, fetch=True))
    # unweighted forest: the other template instance
    for e in (g, o):
        e.forest_init(n_trees, None, nid)
    assert np.array_equal(g.histogram([n_nodes] * n_trees, fetch=True), o.histogram([n_nodes] * n_trees, fetch=True))
    g.close()


@pytest.mark.parametrize("subtract", ["auto", "0"])
@pytest.mark.parametrize("weighted", [True, False])
def test_forest_level_loop_equals_oracle(subtract, weighted, monkeypatch):
    if subtract != "auto":
        monkeypatch.setenv("AVT_SUBTRACT", subtract)
    n, T = 200_017, 4
    n_bins = [16] * 12 + [4] * 6 + [9] * 2
    cat = [False] * 12 + [True] * 6 + [False] * 2
    cols, cls = synth.binned_dataset(n, n_bins, 3, seed=91, signal_attrs=5)
    table = binary_split_table(n_bins, cat)
    weights = [bagging_weights(n, 10_000, 7 + t) if t != 2 else None for t in range(T)] if weighted else None
    g, o = Engine(0), helpers.OracleEngine(threads=8)
    for e in (g, o):
        e.load_dataset(cols, cls, n_bins, 3)
        e.forest_init(T, weights, None)
        e.set_splits(table)
    if subtract == "auto":
        g.set_hist_variant(2)   # every level through the forest kernel (with subtraction the deepest level's counted nodes fit)
    frontier = compare_level_loops(g, o, table, 6, CRIT_GINI, T, class_order=[2, 0, 1])
    assert max(frontier) > 8
    g.close()


def test_forest_loop_two_attribute_chunks_entropy():
    n, T = 60_000, 2
    n_bins = [8] * 40
    cols, cls = synth.binned_dataset(n, n_bins, 2, seed=17, signal_attrs=4)
    table = binary_split_table(n_bins, [False] * 40)
    g, o = Engine(0), helpers.OracleEngine(threads=8)
    for e in (g, o):
        e.load_dataset(cols, cls, n_bins, 2)
        e.forest_init(T, None, None)
        e.set_splits(table)
    g.set_hist_variant(2)       # two trees: the automatic choice would be hist_smem_kernel
    compare_level_loops(g, o, table, 4, CRIT_ENTROPY, T)
    assert g.stats().last_hist_variant == 2
    g.close()
