"""Developer probe: whole-loop device time of avt_grow_levels on C3- / C5-shaped workloads."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from avenir_b200 import synth
from avenir_b200._lib import CRIT_GINI
from avenir_b200.engine import Engine, binary_split_table

def run(name, n, n_bins, cat, n_classes, T, levels, weights):
    cols, cls = synth.binned_dataset(n, n_bins, n_classes, seed=7, signal_attrs=6)
    table = binary_split_table(n_bins, cat)
    eng = Engine(0)
    eng.load_dataset(cols, cls, n_bins, n_classes)
    eng.set_splits(table)
    rng = np.random.default_rng(1)
    w = [rng.poisson(1.0, n).clip(0, 255).astype(np.uint8) for _ in range(T)] if weights else None
    for rep in range(3):
        eng.forest_init(T, w, None)
        eng.reset_stats()
        eng.timer_begin()
        _, fr = eng.grow(levels, CRIT_GINI, [1] * T)
        ms = eng.timer_end()
        st = eng.stats()
    gbs = st.hist_alg_bytes / (st.hist_ms * 1e-3) / 1e9
    print(f"{name}: N={n} attrs={len(n_bins)} trees={T} levels={levels} subtract={os.environ.get('AVT_SUBTRACT','1')}: total {ms:.3f} ms, hist {st.hist_ms:.3f} ms "
          f"({gbs:.0f} GB/s on algorithmic bytes), scan+plan {st.scan_ms:.3f} ms, repart {st.repart_ms:.3f} ms, frontier {max(fr)}, "
          f"rows*levels*trees/s = {n*levels*T/(ms*1e-3)/1e9:.1f} G")
    eng.close()

if __name__ == "__main__":
    which = sys.argv[1]
    if which == "c3":
        run("C3-shape", int(os.environ.get("C3_ROWS", 10_000_000)), [16] * 24 + [4] * 8, [False] * 24 + [True] * 8, 4, int(os.environ.get("C3_TREES", 4)), 6, True)
    else:
        run("C5-shape", int(os.environ.get("C5_ROWS", 2_000_000)), [64] * 512, [False] * 512, 2, 1, 8, False)
