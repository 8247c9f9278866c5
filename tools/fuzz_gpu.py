"""Developer tool: randomised differential run of the in-library level loop (avt_grow_levels / _ex) against the C oracle.

Random shapes around the kernels' switch-over points: 1..70 attributes (one to three 32-attribute chunks, with and without the
rotated copy), 2..300 bins (uint8 and uint16 columns), 2..6 classes, 1..6 trees with and without multiplicities, 1..9 levels,
gini / entropy, `all` / `notUsedYet`, ragged row counts.  Prints one line per case; exits non-zero on the first mismatch.
    python tools/fuzz_gpu.py [seconds] [seed]
"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from avenir_b200 import _lib, synth
from avenir_b200._lib import CRIT_ENTROPY, CRIT_GINI
from avenir_b200.engine import Engine, binary_split_table, grow_levels
from tests import helpers


def one_case(rng, k):
    shape = rng.integers(0, 5)
    if shape == 0:      # narrow (packed codes, rows/smem and joint paths)
        A = int(rng.integers(1, 5)); n_bins = [int(rng.integers(2, 17)) for _ in range(A)]
    elif shape == 1:    # one chunk of uint8 attributes (forest kernel)
        A = int(rng.integers(16, 33)); n_bins = [int(rng.integers(2, 40)) for _ in range(A)]
    elif shape == 2:    # several chunks
        A = int(rng.integers(33, 71)); n_bins = [int(rng.integers(2, 65)) for _ in range(A)]
    elif shape == 3:    # a wide column among narrow ones (uint16 codes: no rotated copy)
        A = int(rng.integers(3, 20)); n_bins = [int(rng.integers(2, 30)) for _ in range(A)]; n_bins[int(rng.integers(A))] = int(rng.integers(257, 301))
    else:               # few attributes with many bins
        A = int(rng.integers(5, 16)); n_bins = [int(rng.integers(50, 257)) for _ in range(A)]
    C = int(rng.integers(2, 7))
    T = int(rng.integers(1, 7)) if shape in (1, 2) else int(rng.integers(1, 3))
    n = int(rng.integers(1, 60_000)) if rng.random() < 0.8 else int(rng.integers(1, 40))
    levels = int(rng.integers(1, 10))
    crit = CRIT_GINI if rng.random() < 0.6 else CRIT_ENTROPY
    strategy = _lib.ATTRS_ALL if rng.random() < 0.6 else _lib.ATTRS_NOT_USED_YET
    cat = [bool(rng.random() < 0.3) and b <= 6 for b in n_bins]
    cols, cls = synth.binned_dataset(n, n_bins, C, seed=int(rng.integers(1 << 30)), signal_attrs=min(A, int(rng.integers(1, 9))))
    table = binary_split_table(n_bins, cat)
    weights = None
    if rng.random() < 0.5:
        weights = [rng.poisson(1.0, n).clip(0, 255).astype(np.uint8) if rng.random() < 0.8 else None for _ in range(T)]
        if all(w is None for w in weights):
            weights = None
    order = list(rng.permutation(C))
    g, o = Engine(0), helpers.OracleEngine(threads=8)
    for e in (g, o):
        e.load_dataset(cols, cls, n_bins, C)
        e.forest_init(T, weights, None)
        e.set_splits(table)
    (rec, nrec), frontier = g.grow(levels, crit, [1] * T, class_order=order, max_records=1 << 16, strategy=strategy)
    ho, fo = grow_levels(o, table, levels, crit, n_trees=T, class_order=order, strategy=strategy)
    want = [(lv, t, nd, s, pops, av, ties) for lv, level in enumerate(ho) for (t, nd, s, pops, av, ties) in level]
    tag = f"case {k}: shape {shape} A={A} bins<= {max(n_bins)} C={C} T={T} n={n} levels={levels} crit={crit} strategy={strategy} w={'y' if weights else 'n'}"
    assert frontier == fo and nrec == len(want), (tag, frontier, fo, nrec, len(want))
    for i, (lv, t, nd, s, pops, av, ties) in enumerate(want):
        r = rec[i]
        assert (r.level, r.tree, r.node, r.split, r.n_ties) == (lv, t, nd, s, ties), (tag, i)
        assert [p for p in r.child_pop if p] == pops and helpers.bits(r.min_av) == helpers.bits(av), (tag, i)
    for t in range(T):
        assert np.array_equal(g.node_ids(t), o.node_ids(t)), (tag, t)
    v = g.stats().last_hist_variant
    g.close()
    return tag + f" variant {v} frontier {max(frontier)} ok"


def main():
    seconds = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
    seed = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    rng = np.random.default_rng(seed)
    t0, k = time.time(), 0
    variants = {}
    while time.time() - t0 < seconds:
        line = one_case(rng, k)
        v = line.split(" variant ")[1].split()[0]
        variants[v] = variants.get(v, 0) + 1
        if k < 5 or k % 25 == 0:
            print(line, flush=True)
        k += 1
    print(f"{k} cases equal the oracle; last histogram variant of each case: {variants}")


if __name__ == "__main__":
    main()
