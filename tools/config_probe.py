"""Developer probe: per-level device times of avt_grow_levels on C3- and C5-shaped workloads (reduced N, T)."""
import os, sys, time
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
    for rep in range(2):
        eng.forest_init(T, w, None)
        fr = [1] * T
        rows = []
        for lvl in range(levels):
            eng.reset_stats()
            eng.timer_begin()
            _, fr = eng.grow(1, CRIT_GINI, fr)
            ms = eng.timer_end()
            st = eng.stats()
            gbs = st.hist_alg_bytes / (st.hist_ms * 1e-3) / 1e9 if st.hist_ms else 0
            rows.append((lvl, max(fr), round(ms * 1e3), round(st.hist_ms * 1e3), round(st.scan_ms * 1e3), round(st.repart_ms * 1e3),
                         st.last_hist_variant, st.last_hist_replicas, st.last_hist_grid, st.last_hist_smem, round(gbs)))
    print(f"== {name}: N={n} attrs={len(n_bins)} bins={sum(n_bins)} splits={len(table)} classes={n_classes} trees={T}")
    print("level, max_frontier, total_us, hist_us, scan_us, repart_us, variant, R, grid, smem, hist_alg_GBps")
    for r in rows: print(r)
    eng.close()

if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    if which in ("c3", "all"):
        run("C3-shape", int(os.environ.get("C3_ROWS", 10_000_000)), [16] * 24 + [4] * 8, [False] * 24 + [True] * 8, 4, int(os.environ.get("C3_TREES", 4)), 6, True)
    if which in ("c5", "all"):
        run("C5-shape", int(os.environ.get("C5_ROWS", 2_000_000)), [64] * 512, [False] * 512, 2, 1, 8, False)
