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


def _c3_worker(rank, world, port, n, T, levels, out_dir):
    import torch.distributed as dist
    from avenir_b200 import _lib
    from avenir_b200._lib import CRIT_GINI
    from avenir_b200.engine import Engine, comm_unique_id
    from avenir_b200.sharding import connect_peers, share_unique_id
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        uid = share_unique_id(dist, comm_unique_id, device="cuda")
        lo, hi = shard_range(n, rank, world)
        n_bins = [16] * 24 + [4] * 8
        cols, cls = synth.binned_dataset(hi - lo, n_bins, 4, seed=33, start=lo, signal_attrs=6)
        eng = Engine(rank)
        eng.comm_init(world, rank, uid)
        connect_peers(dist, eng)
        eng.load_dataset(cols, cls, n_bins, 4)
        eng.set_splits(binary_split_table(n_bins, [False] * 24 + [True] * 8))
        eng.forest_init_sampled(T, _lib.SAMPLE_WITH_REPLACE, 100, 10_000, n, lo)
        (rec, nrec), frontier = eng.grow(levels, CRIT_GINI, [1] * T, class_order=[0, 1, 2, 3], max_records=1 << 14, strategy=_lib.ATTRS_NOT_USED_YET)
        flat = [(rec[i].level, rec[i].tree, rec[i].node, rec[i].split, tuple(rec[i].child_pop), rec[i].min_av.hex()) for i in range(nrec)]
        dist.barrier()
        torch.save({"records": flat, "frontier": frontier, "range": (lo, hi), "nid": [eng.node_ids(t) for t in range(T)]}, os.path.join(out_dir, f"rank{rank}.pt"))
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_forest_two_gpus_equal_one(tmp_path):
    """The C3 shape (forest kernel, sibling subtraction, device-side bootstrap per shard, notUsedYet masks, peer exchange for
    the small levels and NCCL for the large ones) on 2 ranks against 1: records and every row's final node."""
    import torch.multiprocessing as mp
    from avenir_b200 import _lib
    from avenir_b200._lib import CRIT_GINI
    from avenir_b200.engine import Engine
    n, T, levels = 400_003, 6, 6
    mp.spawn(_c3_worker, args=(2, 29950 + os.getpid() % 40, n, T, levels, str(tmp_path)), nprocs=2, join=True)
    n_bins = [16] * 24 + [4] * 8
    cols, cls = synth.binned_dataset(n, n_bins, 4, seed=33, signal_attrs=6)
    eng = Engine(0)
    eng.load_dataset(cols, cls, n_bins, 4)
    eng.set_splits(binary_split_table(n_bins, [False] * 24 + [True] * 8))
    eng.forest_init_sampled(T, _lib.SAMPLE_WITH_REPLACE, 100, 10_000, n, 0)
    (rec, nrec), frontier = eng.grow(levels, CRIT_GINI, [1] * T, class_order=[0, 1, 2, 3], max_records=1 << 14, strategy=_lib.ATTRS_NOT_USED_YET)
    want = [(rec[i].level, rec[i].tree, rec[i].node, rec[i].split, tuple(rec[i].child_pop), rec[i].min_av.hex()) for i in range(nrec)]
    for r in range(2):
        got = torch.load(os.path.join(tmp_path, f"rank{r}.pt"), weights_only=False)
        assert got["records"] == want and got["frontier"] == frontier
        lo, hi = got["range"]
        for t in range(T):
            assert np.array_equal(got["nid"][t], eng.node_ids(t)[lo:hi])


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_decision_tree_builder_tool_on_two_gpus(tmp_path):
    """The reference-named tool with its rows sharded over two GPUs (avt.num.ranks / avt.rank, one process per GPU, NCCL
    rendezvous through a file): the decision-path lists equal the single-GPU run's and the ranks' part files hold every
    row exactly once."""
    import json
    import subprocess
    import sys
    from tests import helpers
    d = str(tmp_path)
    helpers.write_schema(os.path.join(d, "call_hangup.json"), "call_hangup.json")
    lines = synth.call_hangup_csv(20_001)
    open(os.path.join(d, "in.txt"), "w").write("\n".join(lines) + "\n")

    def props(name, extra):
        conf = {"field.delim.regex": ",", "field.delim.out": ",", "dtb.split.algorithm": "entropy", "dtb.split.attribute.selection.strategy": "notUsedYet",
                "dtb.sub.sampling.strategy": "withReplace", "dtb.sub.sampling.buffer.size": "5000", "avt.random.seed": "3",
                "dtb.path.stopping.strategy": "maxDepth", "dtb.max.depth.limit": "3",
                "dtb.feature.schema.file.path": os.path.join(d, "call_hangup.json"), "dtb.decision.file.path.out": os.path.join(d, name + ".json")}
        conf.update(extra)
        p = os.path.join(d, name + ".properties")
        open(p, "w").write("".join(f"{k}={v}\n" for k, v in conf.items()))
        return p

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    one = props("one", {})
    subprocess.run([sys.executable, "-m", "avenir_b200", "org.avenir.tree.DecisionTreeBuilder", f"-Dconf.path={one}", "-Davt.dtb.num.levels=4",
                    os.path.join(d, "in.txt"), os.path.join(d, "out_one")], check=True, cwd=root)
    procs = []
    for r in range(2):
        p = props(f"two_r{r}", {"avt.num.ranks": "2", "avt.rank": str(r), "avt.device": str(r), "avt.comm.id.file": os.path.join(d, "nccl.id"),
                                "dtb.decision.file.path.out": os.path.join(d, "two.json")})
        procs.append(subprocess.Popen([sys.executable, "-m", "avenir_b200", "org.avenir.tree.DecisionTreeBuilder", f"-Dconf.path={p}", "-Davt.dtb.num.levels=4",
                                       os.path.join(d, "in.txt"), os.path.join(d, "out_two")], cwd=root))
    assert [p.wait(timeout=300) for p in procs] == [0, 0]
    for k in range(4):
        assert json.load(open(os.path.join(d, f"two.json.level{k}"))) == json.load(open(os.path.join(d, f"one.json.level{k}")))
    whole = sorted(open(os.path.join(d, "out_one", "part-r-00000")).read().splitlines())
    parts = sorted(open(os.path.join(d, "out_two", "part-r-00000")).read().splitlines() + open(os.path.join(d, "out_two", "part-r-00001")).read().splitlines())
    assert parts == whole
