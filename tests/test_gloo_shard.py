"""N > 1 host logic on CPU (gloo, world_size 2): rows shard across ranks, per-level histograms are summed
with an all-reduce, every rank then makes the same decisions on identical counts.  The kernel layer is the
C oracle here; tests/test_gpu_multi.py repeats it with the real library over NCCL when 2 GPUs exist."""
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from avenir_b200 import synth
from avenir_b200._lib import CRIT_ENTROPY
from avenir_b200.engine import binary_split_table, grow_levels
from avenir_b200.sharding import shard_range, share_unique_id
from tests import helpers


class AllReduceOracle(helpers.OracleEngine):
    """Oracle kernels + the one exchange step of the path: sum of the histograms over the ranks."""

    def histogram(self, nodes_per_tree, fetch=False):
        super().histogram(nodes_per_tree, fetch=False)
        t = torch.from_numpy(self.H.astype(np.int64))
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        self.H[...] = t.numpy().astype(np.uint32)
        return self.H.copy() if fetch else None


def _worker(rank, world, port, n, levels, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        uid = share_unique_id(dist, lambda: bytes(range(128)))
        assert uid == bytes(range(128))
        lo, hi = shard_range(n, rank, world)
        cols, cls, n_bins, _ = synth.call_hangup_encoded(n)
        table = binary_split_table(n_bins, [True, True, True, False])
        eng = AllReduceOracle()
        eng.load_dataset([c[lo:hi] for c in cols], cls[lo:hi], n_bins, 2)
        eng.forest_init(1)
        eng.set_splits(table)
        hist, frontier = grow_levels(eng, table, levels, CRIT_ENTROPY)
        flat = [(lv, t, nd, s, tuple(p), helpers.bits(av), ties) for lv, level in enumerate(hist) for (t, nd, s, p, av, ties) in level]
        np.save(os.path.join(out_dir, f"nid{rank}.npy"), eng.node_ids(0))
        torch.save({"records": flat, "frontier": frontier, "range": (lo, hi)}, os.path.join(out_dir, f"rank{rank}.pt"))
    finally:
        dist.destroy_process_group()


def test_shard_range_partitions_rows():
    for n in (0, 1, 7, 100, 10_000_001):
        for w in (1, 2, 3, 8):
            ranges = [shard_range(n, r, w) for r in range(w)]
            assert ranges[0][0] == 0 and ranges[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(ranges, ranges[1:]))
            sizes = [hi - lo for lo, hi in ranges]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)


def test_two_ranks_build_the_same_tree_as_one(tmp_path):
    n, levels, world = 40_001, 5, 2
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(world, port, n, levels, str(tmp_path)), nprocs=world, join=True)
    cols, cls, n_bins, _ = synth.call_hangup_encoded(n)
    table = binary_split_table(n_bins, [True, True, True, False])
    one = helpers.OracleEngine()
    one.load_dataset(cols, cls, n_bins, 2)
    one.forest_init(1)
    one.set_splits(table)
    hist, frontier = grow_levels(one, table, levels, CRIT_ENTROPY)
    want = [(lv, t, nd, s, tuple(p), helpers.bits(av), ties) for lv, level in enumerate(hist) for (t, nd, s, p, av, ties) in level]
    nid = one.node_ids(0)
    for r in range(world):
        got = torch.load(os.path.join(tmp_path, f"rank{r}.pt"), weights_only=False)
        assert got["records"] == want and got["frontier"] == frontier      # bit-identical decisions on every rank
        lo, hi = got["range"]
        assert np.array_equal(np.load(os.path.join(tmp_path, f"nid{r}.npy")), nid[lo:hi])
