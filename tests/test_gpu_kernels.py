"""GPU parity, kernel level: every kernel of libavenir_tree.so, called through the C-ABI, against the
C oracle on the same seeded inputs.  Counts, chosen splits, node ids: bit-exact.  gini: bit-exact.
entropy / av: bit-exact as well (same fdlibm log, no FMA) -- the 1e-9 relative budget of north_star is
asserted explicitly where the comparison is against the literal oracle."""
import numpy as np
import pytest

from avenir_b200 import synth
from avenir_b200._lib import AVT_NODE_NONE, CRIT_CLASS_CONF, CRIT_ENTROPY, CRIT_GINI, CRIT_HELLINGER
from avenir_b200.engine import Engine, SplitTable, binary_split_table, grow_levels
from tests import helpers

pytestmark = pytest.mark.gpu


def make_pair(cols, cls, n_bins, n_classes, table, n_trees=1, weights=None, node_ids=None):
    g = Engine(0)
    o = helpers.OracleEngine()
    for e in (g, o):
        e.load_dataset(cols, cls, n_bins, n_classes)
        e.forest_init(n_trees, weights, node_ids)
        e.set_splits(table)
    return g, o


def random_node_ids(rng, n, n_nodes, p_none=0.1):
    nid = rng.integers(0, n_nodes, size=n).astype(np.uint16)
    nid[rng.random(n) < p_none] = AVT_NODE_NONE
    return nid


@pytest.mark.parametrize("n", [1, 15, 16, 17, 8191, 8193, 100003])
@pytest.mark.parametrize("variant", [0, 1, 3])
def test_histogram_ragged_sizes(n, variant):
    rng = np.random.default_rng(n)
    n_bins = [2, 4, 2, 10]
    cols, cls = synth.binned_dataset(n, n_bins, 2, seed=n)
    table = binary_split_table(n_bins, [True, True, True, False])
    nid = random_node_ids(rng, n, 5)
    g, o = make_pair(cols, cls, n_bins, 2, table, 1, None, [nid])
    g.set_hist_variant(variant)
    hg = g.histogram([5], fetch=True)
    ho = o.histogram([5], fetch=True)
    assert g.stats().last_hist_variant == variant
    assert np.array_equal(hg, ho)
    assert hg.sum() == 4 * int((nid != AVT_NODE_NONE).sum())


@pytest.mark.parametrize("variant", [0, 1])
def test_histogram_forest_weights_wide_codes(variant):
    """3 trees, bootstrap multiplicities (one tree unweighted), a uint16 column (300 bins), 3 classes."""
    n = 50021
    rng = np.random.default_rng(3)
    n_bins = [7, 300, 16, 3]
    cols, cls = synth.binned_dataset(n, n_bins, 3, seed=11)
    table = binary_split_table(n_bins, [False, False, False, True])
    weights = [rng.poisson(1.0, n).clip(0, 255).astype(np.uint8), None, rng.poisson(1.0, n).clip(0, 255).astype(np.uint8)]
    nids = [random_node_ids(rng, n, k) for k in (3, 1, 6)]
    g, o = make_pair(cols, cls, n_bins, 3, table, 3, weights, nids)
    g.set_hist_variant(variant)
    hg = g.histogram([3, 1, 6], fetch=True)
    ho = o.histogram([3, 1, 6], fetch=True)
    assert np.array_equal(hg, ho)


@pytest.mark.parametrize("n_bins,n_classes", [([2, 4, 2, 10], 2), ([3, 5, 7], 3), ([200], 4), ([2, 2, 2, 2], 2), ([16, 9], 5)])
def test_joint_code_histogram_forest_and_weights(n_bins, n_classes):
    """hist_joint_kernel (one counter space per node indexed by the packed byte, marginals at flush): several trees, one of them
    unweighted, many nodes, rows outside the frontier, schemas whose highest field is not a power of two."""
    n = 70_019
    rng = np.random.default_rng(len(n_bins) * 100 + n_classes)
    cols, cls = synth.binned_dataset(n, n_bins, n_classes, seed=21)
    table = binary_split_table(n_bins, [False] * len(n_bins))
    weights = [rng.poisson(1.0, n).clip(0, 255).astype(np.uint8), None, rng.poisson(1.0, n).clip(0, 255).astype(np.uint8)]
    nids = [random_node_ids(rng, n, k) for k in (7, 1, 20)]
    g, o = make_pair(cols, cls, n_bins, n_classes, table, 3, weights, nids)
    g.set_hist_variant(3)
    hg = g.histogram([7, 1, 20], fetch=True)
    assert g.stats().last_hist_variant == 3
    assert np.array_equal(hg, o.histogram([7, 1, 20], fetch=True))
    g.close()


@pytest.mark.parametrize("weights", [False, True])
@pytest.mark.parametrize("n,nodes", [(1, 1), (33, 5), (70_001, 700), (200_000, 3000)])
def test_histogram_sorted_rows_variant(n, nodes, weights):
    """hist_sorted_kernel (variant 4): one tree, 40 attributes of mixed widths (two 32-attribute chunks), rows outside the
    frontier, empty nodes, frontiers far beyond a block's shared memory, multiplicities incl. 0."""
    rng = np.random.default_rng(n + nodes)
    n_bins = [16] * 20 + [64] * 12 + [3] * 8
    cols, cls = synth.binned_dataset(n, n_bins, 3, seed=9)
    table = binary_split_table(n_bins, [False] * 32 + [True] * 8)
    nid = random_node_ids(rng, n, nodes, 0.2)
    nid[nid == 3] = 4 if nodes > 4 else nid[nid == 3]        # node 3 stays empty when the frontier is large enough
    w = [rng.poisson(1.0, n).clip(0, 255).astype(np.uint8)] if weights else None
    g, o = make_pair(cols, cls, n_bins, 3, table, 1, w, [nid])
    g.set_hist_variant(4)
    hg = g.histogram([nodes], fetch=True)
    assert g.stats().last_hist_variant == 4
    assert np.array_equal(hg, o.histogram([nodes], fetch=True))
    g.close()


def test_pair_class_joint_counts_by_using_one_attributes_bins_as_nodes():
    """The sufficient statistic of the pair-wise count consumers of explore/ (MutualInformation's feature-pair and
    feature-pair-class distributions, explore/MutualInformation.java:181-214) from the existing entry points: repartition
    the root by attribute i with the identity table (node = bin of i), then one level histogram is the joint count
    (bin of i, bin of j, class) of i with EVERY other attribute j."""
    n = 120_000
    n_bins = [7, 300, 12, 5, 33]          # one uint16 column among them
    cols, cls = synth.binned_dataset(n, n_bins, 3, seed=17, signal_attrs=3)
    g = Engine(0)
    g.load_dataset(cols, cls, n_bins, 3)
    off = np.concatenate([[0], np.cumsum(n_bins)])
    for i in (0, 2, 4):
        g.forest_init(1)
        g.histogram([1])                                    # the root frontier (the repartition applies to the last counted one)
        lut = np.full((1, max(n_bins)), AVT_NODE_NONE, dtype=np.uint16)
        lut[0, :n_bins[i]] = np.arange(n_bins[i])
        g.repartition([i], lut)
        h = g.histogram([n_bins[i]], fetch=True)            # [n_bins[i]][total bins][3]
        for j in range(len(n_bins)):
            want = np.zeros((n_bins[i], n_bins[j], 3), dtype=np.uint32)
            np.add.at(want, (cols[i].astype(np.int64), cols[j].astype(np.int64), cls.astype(np.int64)), 1)
            assert np.array_equal(h[:, off[j]:off[j + 1], :], want), (i, j)
    g.close()


def test_histogram_many_nodes_uses_fewer_replicas_then_global():
    n = 200000
    rng = np.random.default_rng(5)
    n_bins = [64] * 6
    cols, cls = synth.binned_dataset(n, n_bins, 2, seed=5)
    table = binary_split_table(n_bins, [False] * 6)
    for nodes, want_variant in ((1, 0), (40, 0), (300, 0), (2000, 1)):
        nid = random_node_ids(rng, n, nodes, 0.0)
        g, o = make_pair(cols, cls, n_bins, 2, table, 1, None, [nid])
        hg = g.histogram([nodes], fetch=True)
        ho = o.histogram([nodes], fetch=True)
        st = g.stats()
        assert st.last_hist_variant == want_variant, (nodes, st.last_hist_variant)
        assert np.array_equal(hg, ho)
        g.close()


@pytest.mark.parametrize("criterion", [CRIT_GINI, CRIT_ENTROPY])
@pytest.mark.parametrize("n_classes", [2, 4])
def test_scan_eval_bit_exact(criterion, n_classes):
    n = 30000
    rng = np.random.default_rng(criterion * 10 + n_classes)
    n_bins = [2, 4, 2, 10, 40, 5]
    cat = [True, True, True, False, False, True]
    cols, cls = synth.binned_dataset(n, n_bins, n_classes, seed=77)
    table = binary_split_table(n_bins, cat)
    nodes = 9
    nid = random_node_ids(rng, n, nodes)
    nid[nid == 7] = 3                                   # node 7 empty
    order = list(rng.permutation(n_classes))
    mask = rng.integers(1, 2 ** 6, size=(nodes, 1)).astype(np.uint32)
    mask[2] = 0                                          # node 2 has no candidate attribute
    g, o = make_pair(cols, cls, n_bins, n_classes, table, 1, None, [nid])
    g.histogram([nodes]); o.histogram([nodes])
    bg = g.scan(criterion, mask, order)
    bo = o.scan(criterion, mask, order)
    for k in range(nodes):
        assert bg[k].population == bo[k].population
        assert bg[k].best_split == bo[k].best_split and bg[k].n_ties == bo[k].n_ties, k
        assert helpers.bits(bg[k].node_stat) == helpers.bits(bo[k].node_stat)
        if bo[k].best_split >= 0:
            assert helpers.bits(bg[k].min_av) == helpers.bits(bo[k].min_av)
        ag, ao = g.fetch_av(k), o.fetch_av(k)
        assert np.array_equal(np.isnan(ag), np.isnan(ao))
        assert np.array_equal(ag[~np.isnan(ag)].view(np.uint64), ao[~np.isnan(ao)].view(np.uint64))
        assert np.array_equal(g.fetch_hist(k), o.fetch_hist(k))
    assert bg[2].best_split == -1 and bg[7].population == 0
    live = [k for k in range(nodes) if bo[k].best_split >= 0]
    eg = g.eval(criterion, live, [bo[k].best_split for k in live], order)
    eo = o.eval(criterion, live, [bo[k].best_split for k in live], order)
    for i in range(len(live)):
        assert helpers.bits(eg[i].av) == helpers.bits(eo[i].av)
        for sg in range(2):
            assert helpers.bits(eg[i].seg_stat[sg]) == helpers.bits(eo[i].seg_stat[sg])
            assert list(eg[i].seg_count[sg])[:n_classes] == list(eo[i].seg_count[sg])[:n_classes]
    assert np.array_equal(g.class_counts(), o.class_counts())


def test_scan_high_cardinality_lut_path():
    """1000-level categorical attribute (uint16 codes, B*C above the staging buffer): prefix + one-vs-rest
    candidate lists as segment tables (config C4's shape, reduced)."""
    n = 60000
    n_bins = [1000, 21]
    cols, cls = synth.binned_dataset(n, n_bins, 2, seed=9, signal_attrs=1)
    table = SplitTable()
    for k in range(0, 999, 37):
        table.add_lut(0, [0 if b <= k else 1 for b in range(1000)])
    for k in range(0, 1000, 91):
        table.add_lut(0, [0 if b == k else 1 for b in range(1000)])
    for k in range(20):
        table.add_numeric(1, [k])
    g, o = make_pair(cols, cls, n_bins, 2, table)
    g.histogram([1]); o.histogram([1])
    bg, bo = g.scan(CRIT_ENTROPY), o.scan(CRIT_ENTROPY)
    assert bg[0].best_split == bo[0].best_split and helpers.bits(bg[0].min_av) == helpers.bits(bo[0].min_av)
    assert np.array_equal(g.fetch_av(0).view(np.uint64), o.fetch_av(0).view(np.uint64))


@pytest.mark.parametrize("criterion", [CRIT_GINI, CRIT_ENTROPY, CRIT_HELLINGER, CRIT_CLASS_CONF])
def test_cpg_sweep_bit_exact(criterion):
    n = 40000
    n_bins = [5, 12, 2]
    cols, cls = synth.binned_dataset(n, n_bins, 2, seed=21)
    table = SplitTable()
    for lut in ([0, 1, 1, 1, 1], [0, 0, 1, 2, 2], [0, 1, 2, 0, 1], [0, 0, 0, 0, 0]):
        table.add_lut(0, lut, n_seg=3 if max(lut) == 2 else 2)
    for i in range(11):
        table.add_numeric(1, [i])
        for j in range(i + 1, 11):
            table.add_numeric(1, [i, j])
    table.add_lut(2, [0, 1])
    g, o = make_pair(cols, cls, n_bins, 2, table)
    g.histogram([1]); o.histogram([1])
    sg, so = g.cpg_sweep(0, criterion, [1, 0]), o.cpg_sweep(0, criterion, [1, 0])
    for s in range(len(table)):
        assert helpers.bits(sg[s].stat) == helpers.bits(so[s].stat), s
        assert helpers.bits(sg[s].split_info) == helpers.bits(so[s].split_info), s
        for j in range(3):
            assert list(sg[s].seg_count[j])[:2] == list(so[s].seg_count[j])[:2]


def test_repartition_and_level_loop_equal_oracle():
    """The per-level driver loop end to end (histogram -> scan -> eval -> repartition) for 6 levels, 2 trees
    with different bootstrap weights: every level's chosen splits, child populations, av bits and the final
    node id of every row equal the oracle's."""
    n = 120007
    rng = np.random.default_rng(8)
    cols, cls, n_bins, _ = synth.call_hangup_encoded(n)
    table = binary_split_table(n_bins, [True, True, True, False])
    weights = [None, rng.poisson(1.0, n).clip(0, 255).astype(np.uint8)]
    g, o = make_pair(cols, cls, n_bins, 2, table, 2, weights, None)
    hg, fg = grow_levels(g, table, 6, CRIT_ENTROPY, n_trees=2)
    ho, fo = grow_levels(o, table, 6, CRIT_ENTROPY, n_trees=2)
    assert fg == fo and len(hg) == len(ho) == 6
    for lg, lo in zip(hg, ho):
        assert len(lg) == len(lo)
        for a, b in zip(lg, lo):
            assert a[:4] == b[:4] and helpers.bits(a[4]) == helpers.bits(b[4]) and a[5] == b[5]
    for t in range(2):
        assert np.array_equal(g.node_ids(t), o.node_ids(t))
    st = g.stats()
    assert st.hist_launches == 6 and st.repart_launches == 6 and st.hist_ms > 0


def test_api_misuse_is_reported():
    from avenir_b200.engine import EngineError
    g = Engine(0)
    with pytest.raises(EngineError):
        g.forest_init(1)                                 # no data set
    cols, cls = synth.binned_dataset(100, [4], 2)
    g.load_dataset(cols, cls, [4], 2)
    with pytest.raises(EngineError):
        g.histogram([1])                                 # no forest
    g.forest_init(1)
    g.histogram([1])
    with pytest.raises(EngineError):
        g.scan(CRIT_GINI)                                # no split table
    t = SplitTable()
    t.add_numeric(0, [9])
    with pytest.raises(EngineError):
        g.set_splits(t)                                  # bound outside the attribute's bins


@pytest.mark.parametrize("mode", ["fast-scan-lookback", "general-scan-unpacked", "fast-scan-ticket", "joint-code-histogram"])
@pytest.mark.parametrize("criterion", [CRIT_GINI, CRIT_ENTROPY])
def test_in_library_level_loop_equals_oracle(mode, criterion, monkeypatch):
    """avt_grow_levels (device-driven: fused repartition+histogram, scan, on-device planning) against the oracle's
    level loop: every record and every final node id, through each form of the scan / planning kernels."""
    monkeypatch.setenv("AVT_SCAN_FAST", "0" if mode == "general-scan-unpacked" else "1")   # item-parallel scan vs the general kernel
    monkeypatch.setenv("AVT_HIST_PACK", "0" if mode == "general-scan-unpacked" else "1")   # one byte per attribute instead of the bit-packed copy
    monkeypatch.setenv("AVT_LOOKBACK", "0" if mode == "fast-scan-ticket" else "1")         # the last scan block plans the level instead of look-back
    n = 150001
    rng = np.random.default_rng(12)
    cols, cls, n_bins, _ = synth.call_hangup_encoded(n)
    table = binary_split_table(n_bins, [True, True, True, False])
    weights = [None, rng.poisson(1.0, n).clip(0, 255).astype(np.uint8), None]
    g, o = make_pair(cols, cls, n_bins, 2, table, 3, weights, None)
    if mode == "joint-code-histogram":
        g.set_hist_variant(3)      # hist_joint_kernel at every level, its repartition table written by the scan kernel
    (rec, nrec), frontier = g.grow(7, criterion, [1, 1, 1], max_records=4096)
    ho, fo = grow_levels(o, table, 7, criterion, n_trees=3)
    want = [(lvl, t, nd, s, pops, av, ties) for lvl, level in enumerate(ho) for (t, nd, s, pops, av, ties) in level]
    assert frontier == fo and nrec == len(want)
    for i, (lvl, t, nd, s, pops, av, ties) in enumerate(want):
        r = rec[i]
        assert (r.level, r.tree, r.node, r.split, r.n_ties) == (lvl, t, nd, s, ties), i
        assert [p for p in r.child_pop if p] == pops
        assert helpers.bits(r.min_av) == helpers.bits(av)
    for t in range(3):
        assert np.array_equal(g.node_ids(t), o.node_ids(t))


def multi_segment_table():
    """4 attributes; numeric splits with 2 and 3 segments, segment tables with 2, 3 and 4 segments (one of
    them with a segment no bin maps to, one with bins outside every segment)."""
    n_bins = [7, 5, 6, 3]
    t = SplitTable()
    for b in range(6):
        t.add_numeric(0, [b])
    for lo in range(5):
        for hi in range(lo + 1, 6):
            t.add_numeric(0, [lo, hi])
    t.add_lut(1, [0, 1, 0, 1, 1])
    t.add_lut(1, [0, 1, 2, 1, 0])
    t.add_lut(1, [2, 1, 1, 2, 1], n_seg=3)        # segment 0 is empty by construction
    t.add_lut(1, [0, 1, 2, 3, 3])
    for b in range(5):
        t.add_numeric(2, [b])
    t.add_numeric(2, [0, 2, 4])
    t.add_lut(3, [0, 1, 1])
    t.add_lut(3, [0, 1, 2])
    t.add_lut(3, [0, 1, 5], n_seg=2)              # bin 2 belongs to no segment: its rows leave the tree
    return n_bins, t


@pytest.mark.parametrize("criterion", [CRIT_GINI, CRIT_ENTROPY])
def test_item_parallel_scan_equals_general_scan_and_oracle(criterion, monkeypatch):
    """The latency-optimised scan + planning (avt_scan_fast.cuh) against the general kernels, field by field and
    bit by bit, and both against the oracle's level loop; multi-segment splits, 3 classes, 2 trees, weights."""
    n = 60013
    n_bins, table = multi_segment_table()
    cols, cls = synth.binned_dataset(n, n_bins, 3, seed=21)
    rng = np.random.default_rng(5)
    weights = [None, rng.poisson(1.0, n).clip(0, 255).astype(np.uint8)]
    runs = {}
    for fast in ("1", "ticket", "0"):
        monkeypatch.setenv("AVT_SCAN_FAST", "0" if fast == "0" else "1")
        monkeypatch.setenv("AVT_LOOKBACK", "0" if fast == "ticket" else "1")
        g, o = make_pair(cols, cls, n_bins, 3, table, 2, weights, None)
        (rec, nrec), frontier = g.grow(5, criterion, [1, 1], class_order=[1, 2, 0], max_records=8192)
        runs[fast] = ([(r.level, r.tree, r.node, r.split, r.n_ties, list(r.child), list(r.child_pop),
                        [helpers.bits(x) for x in r.child_stat], helpers.bits(r.min_av)) for r in rec[:nrec]],
                      frontier, [g.node_ids(t).copy() for t in range(2)])
    assert runs["1"][0] == runs["0"][0] and runs["1"][1] == runs["0"][1]
    assert runs["ticket"][0] == runs["0"][0] and runs["ticket"][1] == runs["0"][1]
    ho, fo = grow_levels(o, table, 5, criterion, n_trees=2, class_order=[1, 2, 0])
    want = [(lvl, t, nd, s, pops, av, ties) for lvl, level in enumerate(ho) for (t, nd, s, pops, av, ties) in level]
    got, frontier, nids = runs["1"]
    assert frontier == fo and len(got) == len(want) and len(want) > 20
    for r, (lvl, t, nd, s, pops, av, ties) in zip(got, want):
        assert r[:5] == (lvl, t, nd, s, ties)
        assert [p for p in r[6] if p] == pops and r[8] == helpers.bits(av)
    assert any(sum(1 for p in r[6] if p) >= 3 for r in got)              # three-way expansions do occur
    for t in range(2):
        assert np.array_equal(nids[t], runs["0"][2][t]) and np.array_equal(nids[t], o.node_ids(t))


def test_level_loop_with_wide_codes_and_many_attrs():
    """uint16 column + more than 4 attributes (generic gather path of the fused repartition), 3 classes."""
    n = 90001
    n_bins = [6, 300, 9, 4, 12, 3]
    cols, cls = synth.binned_dataset(n, n_bins, 3, seed=4)
    table = binary_split_table(n_bins, [False, False, False, True, False, True])
    g, o = make_pair(cols, cls, n_bins, 3, table)
    (rec, nrec), frontier = g.grow(5, CRIT_GINI, [1], class_order=[2, 0, 1], max_records=1024)
    ho, fo = grow_levels(o, table, 5, CRIT_GINI, class_order=[2, 0, 1])
    assert frontier == fo
    want = [x for level in ho for x in level]
    assert nrec == len(want)
    for i, (t, nd, s, pops, av, ties) in enumerate(want):
        assert (rec[i].node, rec[i].split) == (nd, s) and helpers.bits(rec[i].min_av) == helpers.bits(av)
    assert np.array_equal(g.node_ids(0), o.node_ids(0))
