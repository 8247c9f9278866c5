"""Developer probe for hist_rot_kernel: level-0 histogram time on a C3-shaped data set with (a) every (row, tree) pair counted
(no multiplicities), (b) Poisson(1) multiplicities (37 % of the pairs update the phantom slot), (c) 90 % of the pairs dead.
The counter-update count is the same in all three; a difference is the cost of where the updates land."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from avenir_b200 import synth
from avenir_b200.engine import Engine

n, T = int(os.environ.get("ROWS", 10_000_000)), int(os.environ.get("TREES", 48))
n_bins = [16] * 24 + [4] * 8
cols, cls = synth.binned_dataset_device(n, n_bins, 4, seed=7, signal_attrs=6, device="cuda")
torch.cuda.synchronize()
eng = Engine(0)
eng.load_dataset_device([c.data_ptr() for c in cols], cls.data_ptr(), n, n_bins, 4)
rng = np.random.default_rng(1)
for name, lam in (("all pairs live", None), ("Poisson(1) multiplicities", 1.0), ("90 % dead", 0.105)):
    w = None if lam is None else [rng.poisson(lam, n).clip(0, 255).astype(np.uint8) for _ in range(T)]
    eng.forest_init(T, w, None)
    for rep in range(3):
        eng.reset_stats()
        eng.histogram([1] * T)
        eng.synchronize()
        st = eng.stats()
    print(f"{name}: level-0 histogram {st.hist_ms:.3f} ms for {n} rows x {T} trees (variant {st.last_hist_variant}, grid {st.last_hist_grid}, G {st.last_hist_replicas})")
eng.close()
