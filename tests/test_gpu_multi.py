"""Multi-GPU parity (needs >= 2 GPUs; skipped otherwise): rows sharded over 2 ranks, histograms summed by
NCCL, or exchanged through NVLink peer memory inside the scan kernel (push and pull form), against the single-GPU
run on all rows -- shard-count invariance, bit for bit."""
import os

import numpy as np
import pytest
import torch

from avenir_b200 import synth
from avenir_b200._lib import CRIT_ENTROPY
from avenir_b200.engine import binary_split_table
from avenir_b200.sharding import shard_range

pytestmark = pytest.mark.gpu


def _worker(rank, world, port, n, levels, out_dir, peer):
    import torch.distributed as dist
    from avenir_b200.engine import Engine, comm_unique_id
    from avenir_b200.sharding import connect_peers, share_unique_id
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    if peer == "pull":
        os.environ["AVT_PEER_PUSH"] = "0"     # flags + pull out of the ranks' buffers (what large levels and the general scan kernel use)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        uid = share_unique_id(dist, comm_unique_id, device="cuda")
        lo, hi = shard_range(n, rank, world)
        cols, cls, n_bins, _ = synth.call_hangup_encoded(n)
        eng = Engine(rank)
        eng.comm_init(world, rank, uid)
        if peer:
            connect_peers(dist, eng)
        eng.load_dataset([c[lo:hi] for c in cols], cls[lo:hi], n_bins, 2)
        eng.set_splits(binary_split_table(n_bins, [True, True, True, False]))
        eng.forest_init(1)
        (rec, nrec), frontier = eng.grow(levels, CRIT_ENTROPY, [1], max_records=4096)
        flat = [(rec[i].level, rec[i].node, rec[i].split, tuple(rec[i].child_pop), rec[i].min_av.hex()) for i in range(nrec)]
        if peer:      # a second build right away: exchange buffers are reused across calls
            eng.forest_init(1)
            (rec2, nrec2), frontier2 = eng.grow(levels, CRIT_ENTROPY, [1], max_records=4096)
            assert frontier2 == frontier and nrec2 == nrec and all(rec2[i].split == rec[i].split for i in range(nrec))
            assert eng.stats().comm_calls > 0
        dist.barrier()      # (ranks synchronise before tearing their engines down)
        torch.save({"records": flat, "frontier": frontier, "range": (lo, hi), "nid": eng.node_ids(0)}, os.path.join(out_dir, f"rank{rank}.pt"))
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("peer", [False, True, "pull"], ids=["nccl", "nvlink-peer-push", "nvlink-peer-pull"])
def test_two_gpus_equal_one(tmp_path, peer):
    import torch.multiprocessing as mp
    from avenir_b200.engine import Engine
    n, levels = 1_000_003, 6
    mp.spawn(_worker, args=(2, 29400 + os.getpid() % 500 + [False, True, "pull"].index(peer), n, levels, str(tmp_path), peer), nprocs=2, join=True)
    cols, cls, n_bins, _ = synth.call_hangup_encoded(n)
    eng = Engine(0)
    eng.load_dataset(cols, cls, n_bins, 2)
    eng.set_splits(binary_split_table(n_bins, [True, True, True, False]))
    eng.forest_init(1)
    (rec, nrec), frontier = eng.grow(levels, CRIT_ENTROPY, [1], max_records=4096)
    want = [(rec[i].level, rec[i].node, rec[i].split, tuple(rec[i].child_pop), rec[i].min_av.hex()) for i in range(nrec)]
    nid = eng.node_ids(0)
    for r in range(2):
        got = torch.load(os.path.join(tmp_path, f"rank{r}.pt"), weights_only=False)
        assert got["records"] == want and got["frontier"] == frontier
        lo, hi = got["range"]
        assert np.array_equal(got["nid"], nid[lo:hi])
