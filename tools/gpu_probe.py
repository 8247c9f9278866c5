"""Developer probe (not part of the product): per-call wall/device timings on the C2 workload."""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from avenir_b200 import synth
from avenir_b200._lib import CRIT_ENTROPY
from avenir_b200.engine import Engine, binary_split_table

N = int(os.environ.get("ROWS", 10_000_000)); L = 8
cols, cls, n_bins, _ = synth.call_hangup_encoded(N)
table = binary_split_table(n_bins, [True, True, True, False])
pinned = [torch.from_numpy(c).pin_memory() for c in cols]; pc = torch.from_numpy(cls).pin_memory()
cols_p = [p.numpy() for p in pinned]; cls_p = pc.numpy()
eng = Engine(0)
def tm(f, *a):
    t = time.perf_counter(); r = f(*a); eng.synchronize(); return (time.perf_counter() - t) * 1e3, r
for rep in range(3):
    t1, _ = tm(eng.load_dataset, cols_p, cls_p, n_bins, 2)
    t2, _ = tm(eng.set_splits, table)
    t3, _ = tm(eng.forest_init, 1)
    t4, _ = tm(eng.grow, L, CRIT_ENTROPY, [1])
    t5, _ = tm(eng.node_ids, 0)
    print(f"e2e rep{rep}: load {t1:.2f} splits {t2:.2f} forest {t3:.2f} grow {t4:.2f} nid {t5:.2f} ms")
# per-level device times of the fused loop: run grow level by level
for rep in range(2):
    eng.forest_init(1)
    fr = [1]; out = []
    for lvl in range(L):
        eng.reset_stats()
        eng.timer_begin()
        _, fr = eng.grow(1, CRIT_ENTROPY, fr)
        ms = eng.timer_end()
        st = eng.stats()
        out.append((lvl, fr[0], round(ms * 1e3), round(st.hist_ms * 1e3), round(st.scan_ms * 1e3), round(st.eval_ms * 1e3), round(st.repart_ms * 1e3), st.last_hist_replicas, st.last_hist_grid, st.last_hist_smem))
    print("level, frontier, total_us, hist_us, scan_us, plan_us, repart_us, R, grid, smem")
    for o in out: print(o)
