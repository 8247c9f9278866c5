#!/usr/bin/env python
"""bench.py -- split-evaluation throughput of the tree-building hot path (BASELINE.json metric).

    python bench.py [--workload c3|c2|c5|c4|c1] [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--rows R] [--trees T] [--levels L] [--no-verify] [--no-cpu-baseline]

Headline workload (default, the configuration BASELINE.json's metric and north_star's targets are quoted on): C3 --
the bagged forest of resource/rafo.properties: 100 trees over 100 M synthetic rows x 32 attributes (24 int attributes
with 16 bins + 8 categorical with 4 levels = 416 bins, 416 binary candidate splits), 4 classes, gini, 6 expansions below
the root, bootstrap multiplicities = BaggingSampler semantics (batches of 10 000 rows, java.util.Random(100 + t) stream,
generated on the device).  STRONG scaling: the 100 M rows are sharded over the ranks (contiguous ranges), per-level
histograms are summed over NVLink (peer memory inside the scan kernel when they fit the exchange buffers, else one NCCL
all-reduce).  Columns are generated in device memory (counter-based splitmix64, avenir_b200/synth.py) -- nothing of this
size is built on the host.  A "step" grows the whole forest once: 6 x (fused repartition + histogram -> gini scan/arg-min
+ children statistics + level planning), all trees of a level in one launch.  `--workload c2` is the round-1 line (call_hangup tree,
10 M rows per GPU, entropy, 8 expansions, weak scaling); `--workload c5|c4|c1` are the other BASELINE.json configs as one-GPU lines
(bench_extra.py: wide 512-attribute schema, ClassPartitionGenerator sweep, the tutorial through the reference-named tool).

  verify  (default on) BEFORE the timed region the parity subset of SURVEY 8(d) -- 5 trees over the first 1 M rows, same knobs,
          sharded over the same ranks through the same exchange -- is grown and compared with the C oracle on every rank:
          records (level, tree, node, split, ties, child populations, av bits) and every row's final node.  A mismatch aborts.
  value   rows/sec = total rows * levels * trees / step time, inputs resident in HBM; step time = CUDA events the
          library records on its own stream around everything avt_grow_levels enqueues, result copies included
          (avt_stats.last_grow_ms; max over ranks);  L2 is flushed between steps and the inputs are far larger than L2.
  e2e     same metric through the C-ABI from HOST buffers: avt_dataset_load (H2D from pinned memory) +
          avt_forest_init_sampled + avt_grow_levels + read-back of the forest's records, wall clock
  roofline  dominant kernel = the histogram pass: algorithmic bytes / its CUDA-event time
  cpu_baseline  the C oracle (oracle/hist_oracle.c, "port": the reference is JVM code and cannot run on
          this box), bounded sample of the same workload, all host cores and one thread (Hadoop local mode is serial)

`--impl reference` times that CPU port as the reference arm (rank 0 only under torchrun).
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "split_eval_rows_per_sec"
UNIT = "rows/s"
ROWS_DEFAULT = 10_000_000      # C2: rows per GPU
LEVELS_DEFAULT = 8             # C2
CPU_SAMPLE_ROWS = 10_000_000   # C2: the CPU arm runs the full shard: 8 levels over 10 M rows on all host cores
# C3 (rafo.properties forest), SURVEY 8(d)
C3_ROWS, C3_TREES, C3_LEVELS, C3_CLASSES = 100_000_000, 100, 6, 4
C3_BINS = [16] * 24 + [4] * 8
C3_CAT = [False] * 24 + [True] * 8
C3_BATCH, C3_SEED, C3_SIGNAL = 10_000, 100, 6          # bas.batch.size, java.util.Random(seed + t), attributes behind the class rule
C3_VERIFY_ROWS, C3_VERIFY_TREES = 1_000_000, 5         # parity subset
C3_CPU_ROWS, C3_CPU_TREES, C3_CPU_ROWS_1T = 10_000_000, 5, 2_000_000   # bounded CPU samples (all cores / one thread)


def measured_peak():
    try:
        d = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """SM clock + throttle reasons DURING the timed region, through NVML (the library behind nvidia-smi's
    clocks.sm / clocks_event_reasons.* query; in-process so that sampling does not stall CUDA calls)."""
    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}

    def __init__(self, gpu_index: int):
        super().__init__(daemon=True)
        self.samples = []
        self.stop_flag = False
        self.handle = None
        self.nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            # NVML indexes physical GPUs; honour CUDA_VISIBLE_DEVICES when it lists indices
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = gpu_index
            if vis:
                parts = [p for p in vis.split(",") if p.strip().isdigit()]
                if gpu_index < len(parts):
                    idx = int(parts[gpu_index])
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(idx)
            self.nvml = pynvml
        except Exception:
            self.nvml = None

    def run(self):
        if self.nvml is None:
            return
        n = self.nvml
        while not self.stop_flag:
            try:
                sm = n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)
                mx = n.nvmlDeviceGetMaxClockInfo(self.handle, n.NVML_CLOCK_SM)
                rs = n.nvmlDeviceGetCurrentClocksEventReasons(self.handle) if hasattr(n, "nvmlDeviceGetCurrentClocksEventReasons") \
                    else n.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle)
                self.samples.append((sm, mx, rs))
            except Exception:
                pass
            time.sleep(0.002)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0, "source": "nvml unavailable"}
        sm = [s[0] for s in self.samples]
        reasons = set()
        for s in self.samples:
            for bit, name in self.REASONS.items():
                if s[2] & bit:
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(s[1] for s in self.samples)), "reasons": sorted(reasons),
                "samples": len(sm), "source": "nvml"}


def cpu_port_rows_per_sec(rows: int, levels: int, threads: int, repeats: int = 1, data=None):
    """The reference's CPU path restated (oracle, histogram formulation) on `threads` host threads."""
    from avenir_b200 import synth
    from avenir_b200.engine import binary_split_table, grow_levels
    from avenir_b200._lib import CRIT_ENTROPY
    from tests.helpers import OracleEngine
    cols, cls, n_bins = data if data is not None else synth.call_hangup_encoded(rows)[:3]
    table = binary_split_table(n_bins, [True, True, True, False])
    o = OracleEngine(threads=threads)
    o.load_dataset(cols, cls, n_bins, 2)
    o.set_splits(table)
    best = None
    for _ in range(repeats):
        o.forest_init(1)
        t0 = time.perf_counter()
        grow_levels(o, table, levels, CRIT_ENTROPY, record=False)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return rows * levels / best, best


def run_reference_c2(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    threads = os.cpu_count() or 1
    rows = min(args.rows, CPU_SAMPLE_ROWS)
    for _ in range(args.warmup and 1):
        cpu_port_rows_per_sec(min(rows, 200_000), args.levels, threads)
    vals = []
    t_all = time.perf_counter()
    from avenir_b200 import synth
    data = synth.call_hangup_encoded(rows)[:3]
    for _ in range(max(1, args.steps)):
        v, _dt = cpu_port_rows_per_sec(rows, args.levels, threads, data=data)
        vals.append(v)
        if time.perf_counter() - t_all > 150:
            break
    value = float(np.mean(vals))
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": len(vals),
            "warmup": args.warmup, "ms_per_step": 1e3 * rows * args.levels / value, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u32 counts / f64 impurity", "data": "synthetic",
            "config": {"workload": "C2 call_hangup tree, entropy, depth 8 (CPU port of the reference path; the JVM reference cannot run on this box)",
                       "rows": rows, "levels": args.levels},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port",
                             "sample": f"{rows} rows x {args.levels} levels of the C2 workload per step"},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))
    return 0


def run_c2(args):
    if args.rows is None:
        args.rows = ROWS_DEFAULT
    if args.levels is None:
        args.levels = LEVELS_DEFAULT
    if args.impl == "reference":
        return run_reference_c2(args)

    import torch
    import torch.distributed as dist
    from avenir_b200 import build, synth
    from avenir_b200._lib import CRIT_ENTROPY
    from avenir_b200.engine import Engine, binary_split_table, comm_unique_id

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the split-evaluation path has no CPU fallback (use --impl reference for the CPU arm)")
    build.build()
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    N, L = args.rows, args.levels
    cols, cls, n_bins, _names = synth.call_hangup_encoded(N, start=rank * N)
    table = binary_split_table(n_bins, [True, True, True, False])
    # pinned host staging for the e2e leg
    pinned = [torch.from_numpy(c).pin_memory() for c in cols]
    pinned_cls = torch.from_numpy(cls).pin_memory()
    cols_p = [p.numpy() for p in pinned]
    cls_p = pinned_cls.numpy()

    eng = Engine(local)
    if world > 1:
        from avenir_b200.sharding import connect_peers, share_unique_id
        eng.comm_init(world, rank, share_unique_id(dist, comm_unique_id, device="cuda"))   # NCCL: fallback for oversized levels
        if os.environ.get("AVT_BENCH_NCCL") != "1":
            connect_peers(dist, eng)       # histograms are summed out of peer memory inside the scan kernel
    eng.load_dataset(cols_p, cls_p, n_bins, 2)
    eng.set_splits(table)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step_resident():
        eng.forest_init(1)
        flush.zero_()
        torch.cuda.synchronize()
        eng.timer_begin()
        _, frontier = eng.grow(L, CRIT_ENTROPY, [1])
        api_ms = eng.timer_end()               # events around the Python call: includes ctypes marshalling on both sides
        return eng.stats().last_grow_ms, api_ms, frontier   # the library's own events around the call's device work

    for _ in range(max(3, args.warmup)):
        step_resident()
    # ---- timed region: K steps, per-launch event timing off (only the bracket around each step)
    eng.set_profiling(False)
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    dev_ms = api_ms_sum = 0.0
    t0 = time.perf_counter()
    for _ in range(args.steps):
        ms, api_ms, frontier = step_resident()
        dev_ms += ms
        api_ms_sum += api_ms
    barrier()
    wall = time.perf_counter() - t0
    # ---- same steps again with per-launch CUDA events on: kernel breakdown + roofline of the dominant kernel
    eng.set_profiling(True)
    eng.reset_stats()
    prof_steps = max(3, min(args.steps, 10))
    for _ in range(prof_steps):
        step_resident()
    barrier()
    st = eng.stats()
    sampler.stop_flag = True
    sampler.join()

    # ---- e2e: host buffers -> tree, through the C-ABI
    nid_pinned = torch.empty(N, dtype=torch.uint16).pin_memory().numpy()

    def step_e2e():
        t = time.perf_counter()
        eng.load_dataset(cols_p, cls_p, n_bins, 2)          # H2D of the step's inputs (pinned host buffers)
        eng.set_splits(table)
        eng.forest_init(1)
        (rec, nrec), _ = eng.grow(L, CRIT_ENTROPY, [1], max_records=1024)   # D2H of the tree (one record per expanded node)
        nid = eng.node_ids(0, out=nid_pinned)                # D2H of every row's leaf
        return time.perf_counter() - t, nid

    step_e2e()
    barrier()
    e2e_s = 0.0
    e2e_steps = max(1, min(args.steps, 5))
    for _ in range(e2e_steps):
        dt, nid = step_e2e()
        e2e_s += dt
    barrier()

    t_dev = torch.tensor([dev_ms / args.steps, e2e_s / e2e_steps * 1e3], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t_dev, op=dist.ReduceOp.MAX)
    ms_step, e2e_ms = float(t_dev[0]), float(t_dev[1])

    if rank == 0:
        peak, peak_src = measured_peak()
        hist_ms_per_launch = st.hist_ms / max(1, st.hist_launches)
        alg_bytes_per_launch = st.hist_alg_bytes / max(1, st.hist_launches)
        achieved = alg_bytes_per_launch / (hist_ms_per_launch * 1e-3) / 1e9 if hist_ms_per_launch > 0 else 0.0
        traffic = None
        try:
            traffic = json.load(open(os.path.join(ROOT, "profiles", "hist_traffic.json"))).get("dram_bytes_per_launch")
        except Exception:
            pass
        line = {
            "metric": METRIC, "value": world * N * L / (ms_step * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u32 counts / f64 impurity", "data": "synthetic",
            "config": {"workload": "C2: call_hangup decision tree (call_hangup.json schema), entropy, attribute strategy all, 8 expansions",
                       "rows_per_gpu": N, "levels": L, "attrs": 4, "bins": 18, "splits": 18, "classes": 2, "trees": 1,
                       "l2": "flushed between steps (256 MiB memset); the 70 MB working set is L2-resident within a step",
                       "final_frontier_nodes": int(frontier[0]), "parallelism": (f"rows sharded over {world} ranks, per-level histograms exchanged through NVLink peer memory inside the scan kernel (no NCCL call on the path)"
                                       if world > 1 and os.environ.get("AVT_BENCH_NCCL") != "1" else
                                       (f"rows sharded over {world} ranks, NCCL all-reduce of uint32 histograms per level" if world > 1 else "1 rank (no exchange)"))},
            "kernel_ms_per_step": {"histogram": st.hist_ms / prof_steps, "scan_plan": st.scan_ms / prof_steps, "eval": st.eval_ms / prof_steps,
                                   "repartition": st.repart_ms / prof_steps, "allreduce": st.comm_ms / prof_steps,
                                   "note": f"CUDA events per launch, from {prof_steps} extra profiled steps"},
            "wall_ms_per_step": wall * 1e3 / args.steps, "api_ms_per_step": api_ms_sum / args.steps,
            "gpu_launches": int((st.hist_launches + st.scan_launches + st.eval_launches + st.repart_launches) * args.steps // prof_steps),
            "e2e": {"value": world * N * L / (e2e_ms * 1e-3), "unit": UNIT, "ms_per_step": e2e_ms,
                    "h2d_bytes_per_step": int(sum(c.nbytes for c in cols_p) + cls_p.nbytes), "d2h_bytes_per_step": int(nid.nbytes)},
            "roofline": {"bound": "hbm", "kernel": {0: "hist_smem_kernel", 1: "hist_global_kernel", 2: "hist_rot_kernel", 3: "hist_joint_kernel", 4: "hist_sorted_kernel"}.get(st.last_hist_variant, "?"),
                         "achieved": achieved, "peak": peak, "peak_source": peak_src, "unit": "GB/s", "frac": achieved / peak,
                         "frac_of_spec_8000": achieved / 8000.0, "traffic": traffic,
                         "alg_bytes_per_launch": alg_bytes_per_launch, "ms_per_launch": hist_ms_per_launch,
                         "launch": {"grid": st.last_hist_grid, "block": st.last_hist_block, "smem": st.last_hist_smem, "replicas": st.last_hist_replicas}},
            "clocks": sampler.summary(),
        }
        if not args.no_cpu_baseline and world == 1:
            threads = os.cpu_count() or 1
            n_cpu = min(N, CPU_SAMPLE_ROWS)
            v, dt = cpu_port_rows_per_sec(n_cpu, L, threads, repeats=3, data=([c[:n_cpu] for c in cols], cls[:n_cpu], n_bins))
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
                                    "sample": f"{n_cpu} rows x {L} levels of the same workload, best of 3 runs of {dt:.2f} s on {threads} threads"}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


# ======================================================================================================================
# C3: the rafo.properties forest
# ======================================================================================================================
C3_WORKLOAD = ("C3: random forest (rafo.properties): {T} bagged trees over {N} synthetic rows x 32 attributes (24 int x 16 bins + "
               "8 categorical x 4 levels, 416 binary candidate splits), 4 classes, gini, {L} expansions, BaggingSampler multiplicities "
               "(batch 10000, java.util.Random(100 + t))")


def c3_table():
    from avenir_b200.engine import binary_split_table
    return binary_split_table(C3_BINS, C3_CAT)


def c3_oracle_setup(rows: int, trees: int, threads: int, total_rows=None):
    """C oracle loaded with the first `rows` rows of the C3 data and `trees` bootstrap samples."""
    from avenir_b200 import synth
    from tests.helpers import OracleEngine, oracle_bagging
    cols, cls = synth.binned_dataset(rows, C3_BINS, C3_CLASSES, seed=synth.DEFAULT_SEED, start=0, signal_attrs=C3_SIGNAL)
    total = rows if total_rows is None else total_rows
    weights = [oracle_bagging(C3_SEED + t, C3_BATCH, total, 0, rows) for t in range(trees)]
    o = OracleEngine(threads=threads)
    o.load_dataset(cols, cls, C3_BINS, C3_CLASSES)
    o.set_splits(c3_table())
    return o, weights


def c3_oracle_forest(rows: int, trees: int, levels: int, threads: int, total_rows=None, want_state: bool = False, setup=None):
    """The reference's CPU path restated (C oracle, histogram formulation) on the first `rows` rows of the C3 data with
    `trees` bagged trees: returns (seconds of the level loop, history, frontier, oracle engine)."""
    from avenir_b200._lib import CRIT_GINI
    from avenir_b200.engine import grow_levels
    o, weights = setup if setup is not None else c3_oracle_setup(rows, trees, threads, total_rows)
    o.forest_init(trees, weights, None)
    t0 = time.perf_counter()
    hist, frontier = grow_levels(o, c3_table(), levels, CRIT_GINI, n_trees=trees, class_order=list(range(C3_CLASSES)), record=want_state)
    return time.perf_counter() - t0, hist, frontier, o


def c3_verify(eng, dist, world: int, rank: int, levels: int):
    """In-run parity gate: the T = 5 / 1 M-row subset of SURVEY 8(d), sharded over the same ranks and summed through the same
    exchange as the timed workload, against the C oracle on the whole subset.  Every rank checks the records (they are
    replicated) and its own rows' final nodes.  Raises on the first difference."""
    import torch
    from avenir_b200 import _lib, synth
    from avenir_b200._lib import CRIT_GINI
    from avenir_b200.sharding import shard_range
    from tests.helpers import bits
    rows, T = C3_VERIFY_ROWS, C3_VERIFY_TREES
    lo, hi = shard_range(rows, rank, world)
    cols, cls = synth.binned_dataset(hi - lo, C3_BINS, C3_CLASSES, seed=synth.DEFAULT_SEED, start=lo, signal_attrs=C3_SIGNAL)
    eng.load_dataset(cols, cls, C3_BINS, C3_CLASSES)
    eng.set_splits(c3_table())
    eng.forest_init_sampled(T, _lib.SAMPLE_WITH_REPLACE, C3_SEED, C3_BATCH, rows, lo)
    (rec, nrec), frontier = eng.grow(levels, CRIT_GINI, [1] * T, class_order=list(range(C3_CLASSES)), max_records=1 << 14)
    threads = max(1, (os.cpu_count() or 1) // world)
    _, hist, fo, o = c3_oracle_forest(rows, T, levels, threads, want_state=True)
    want = [(lv, t, nd, sp, pops, av, ties) for lv, level in enumerate(hist) for (t, nd, sp, pops, av, ties) in level]
    if frontier != fo or nrec != len(want):
        raise SystemExit(f"bench.py verify: rank {rank}: frontier {frontier} / {nrec} records, oracle {fo} / {len(want)}")
    for i, (lv, t, nd, sp, pops, av, ties) in enumerate(want):
        r = rec[i]
        if (r.level, r.tree, r.node, r.split, r.n_ties) != (lv, t, nd, sp, ties) or [p for p in r.child_pop if p] != pops or bits(r.min_av) != bits(av):
            raise SystemExit(f"bench.py verify: rank {rank}: record {i} differs from the oracle: "
                             f"{(r.level, r.tree, r.node, r.split, r.n_ties, list(r.child_pop), r.min_av)} vs {(lv, t, nd, sp, ties, pops, av)}")
    for t in range(T):
        if not np.array_equal(eng.node_ids(t), o.node_ids(t)[lo:hi]):
            raise SystemExit(f"bench.py verify: rank {rank}: final nodes of tree {t} differ from the oracle")
    return {"rows": rows, "trees": T, "levels": levels, "records": nrec, "ranks": world, "result": "records and final nodes equal the C oracle on every rank"}


def run_reference_c3(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    threads = os.cpu_count() or 1
    L = args.levels or C3_LEVELS
    rows, T = min(args.rows or C3_ROWS, C3_CPU_ROWS), min(args.trees or C3_TREES, C3_CPU_TREES)
    if args.warmup:
        c3_oracle_forest(min(rows, 200_000), T, L, threads)
    vals = []
    setup = c3_oracle_setup(rows, T, threads)
    t_all = time.perf_counter()
    for _ in range(max(1, args.steps)):
        dt, _, _, _ = c3_oracle_forest(rows, T, L, threads, setup=setup)
        vals.append(rows * L * T / dt)
        if time.perf_counter() - t_all > 150:
            break
    value = float(np.mean(vals))
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": len(vals),
            "warmup": args.warmup, "ms_per_step": 1e3 * rows * L * T / value, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "u32 counts / f64 impurity", "data": "synthetic",
            "config": {"workload": C3_WORKLOAD.format(T=args.trees or C3_TREES, N=args.rows or C3_ROWS, L=L) +
                       " -- CPU port of the reference path (the JVM reference cannot run on this box), bounded sample per step",
                       "sample_rows": rows, "sample_trees": T, "levels": L},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port",
                             "sample": f"first {rows} rows x {T} trees x {L} levels of the C3 workload per step"},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))
    return 0


def run_c3(args):
    if args.impl == "reference":
        return run_reference_c3(args)
    import torch
    import torch.distributed as dist
    from avenir_b200 import _lib, build, synth
    from avenir_b200._lib import CRIT_GINI
    from avenir_b200.engine import Engine, comm_unique_id
    from avenir_b200.sharding import shard_range

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the split-evaluation path has no CPU fallback (use --impl reference for the CPU arm)")
    build.build()
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    N_total, T, L = args.rows or C3_ROWS, args.trees or C3_TREES, args.levels or C3_LEVELS
    lo, hi = shard_range(N_total, rank, world)
    n_local = hi - lo
    order = list(range(C3_CLASSES))

    eng = Engine(local)
    if world > 1:
        from avenir_b200.sharding import connect_peers, share_unique_id
        eng.comm_init(world, rank, share_unique_id(dist, comm_unique_id, device="cuda"))   # levels beyond the exchange buffers: NCCL all-reduce
        connect_peers(dist, eng)                                                            # small levels: summed out of peer memory in the scan kernel
    verify = None
    if not args.no_verify:
        verify = c3_verify(eng, dist, world, rank, L)

    # ---- the workload: columns generated in device memory, handed to the library as device buffers
    cols, cls = synth.binned_dataset_device(n_local, C3_BINS, C3_CLASSES, seed=synth.DEFAULT_SEED, start=lo, signal_attrs=C3_SIGNAL, device="cuda")
    torch.cuda.synchronize()
    eng.load_dataset_device([c.data_ptr() for c in cols], cls.data_ptr(), n_local, C3_BINS, C3_CLASSES)
    eng.set_splits(c3_table())
    eng.forest_init_sampled(T, _lib.SAMPLE_WITH_REPLACE, C3_SEED, C3_BATCH, N_total, lo)
    sample_ms = eng.stats().sample_ms
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step_resident(max_records=0):
        eng.forest_reset()
        flush.zero_()
        torch.cuda.synchronize()
        eng.timer_begin()
        (rec, nrec), frontier = eng.grow(L, CRIT_GINI, [1] * T, class_order=order, max_records=max_records)
        api_ms = eng.timer_end()
        return eng.stats().last_grow_ms, api_ms, frontier, nrec

    for _ in range(max(3, args.warmup)):
        step_resident()
    eng.set_profiling(False)
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    dev_ms = api_ms_sum = 0.0
    t0 = time.perf_counter()
    for _ in range(args.steps):
        ms, api_ms, frontier, _ = step_resident()
        dev_ms += ms
        api_ms_sum += api_ms
    barrier()
    wall = time.perf_counter() - t0
    # ---- the same step with per-launch CUDA events on: kernel breakdown + roofline of the dominant kernel
    eng.set_profiling(True)
    eng.reset_stats()
    prof_steps = 2
    for _ in range(prof_steps):
        step_resident()
    barrier()
    st = eng.stats()
    sampler.stop_flag = True
    sampler.join()

    # ---- e2e: host buffers -> forest records, through the C-ABI
    pinned = [torch.empty(n_local, dtype=torch.uint8).pin_memory() for _ in cols]
    for p_, c in zip(pinned, cols):
        p_.copy_(c)
    pinned_cls = torch.empty(n_local, dtype=torch.uint8).pin_memory()
    pinned_cls.copy_(cls)
    torch.cuda.synchronize()
    del cols, cls
    torch.cuda.empty_cache()
    cols_p = [p_.numpy() for p_ in pinned]
    cls_p = pinned_cls.numpy()
    table = c3_table()
    max_rec = T * (2 ** L)

    def step_e2e():
        t = time.perf_counter()
        eng.load_dataset(cols_p, cls_p, C3_BINS, C3_CLASSES)                          # H2D of the step's inputs (pinned host buffers)
        eng.set_splits(table)
        eng.forest_init_sampled(T, _lib.SAMPLE_WITH_REPLACE, C3_SEED, C3_BATCH, N_total, lo)
        (rec, nrec), fr = eng.grow(L, CRIT_GINI, [1] * T, class_order=order, max_records=max_rec)   # D2H: one record per expanded node = the forest
        return time.perf_counter() - t, nrec

    step_e2e()
    barrier()
    e2e_s = 0.0
    e2e_steps = max(1, min(args.steps, 3))
    nrec = 0
    for _ in range(e2e_steps):
        dt, nrec = step_e2e()
        e2e_s += dt
    barrier()

    t_dev = torch.tensor([dev_ms / args.steps, e2e_s / e2e_steps * 1e3], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t_dev, op=dist.ReduceOp.MAX)
    ms_step, e2e_ms = float(t_dev[0]), float(t_dev[1])

    if rank == 0:
        peak, peak_src = measured_peak()
        hist_ms_per_launch = st.hist_ms / max(1, st.hist_launches)
        alg_bytes_per_launch = st.hist_alg_bytes / max(1, st.hist_launches)
        achieved = alg_bytes_per_launch / (hist_ms_per_launch * 1e-3) / 1e9 if hist_ms_per_launch > 0 else 0.0
        traffic = traffic_src = None
        try:
            tj = json.load(open(os.path.join(ROOT, "profiles", "hist_traffic_c3.json")))
            traffic, traffic_src = tj.get("dram_bytes_per_launch"), tj.get("source")
        except Exception:
            pass
        # counter updates the pass issues: the kernel is bound by shared-memory atomic throughput, not by HBM
        line = {
            "metric": METRIC, "value": N_total * L * T / (ms_step * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "u32 counts / f64 impurity", "data": "synthetic",
            "config": {"workload": C3_WORKLOAD.format(T=T, N=N_total, L=L), "rows_total": N_total, "rows_per_gpu": n_local, "trees": T, "levels": L,
                       "attrs": 32, "bins": 416, "splits": 416, "classes": 4, "strategy": "all",
                       "l2": "flushed between steps (256 MiB memset); per-GPU inputs (>= 6 GB of node ids, multiplicities and codes) are far larger than L2",
                       "final_frontier_nodes_max": int(max(frontier)), "records": int(nrec),
                       "parallelism": (f"rows sharded over {world} ranks; per-level uint32 histograms summed through NVLink peer memory inside the scan kernel while they fit the "
                                       f"exchange buffers, else one NCCL all-reduce per level" if world > 1 else "1 rank (no exchange)")},
            "verify": verify if verify is not None else "skipped (--no-verify)",
            "kernel_ms_per_step": {"histogram": st.hist_ms / prof_steps, "scan_plan": st.scan_ms / prof_steps, "eval": st.eval_ms / prof_steps,
                                   "repartition": st.repart_ms / prof_steps, "allreduce": st.comm_ms / prof_steps,
                                   "note": f"CUDA events per launch, from {prof_steps} extra profiled steps"},
            "setup_ms": {"bootstrap_multiplicities_on_device": sample_ms},
            "wall_ms_per_step": wall * 1e3 / args.steps, "api_ms_per_step": api_ms_sum / args.steps,
            "gpu_launches": int((st.hist_launches + st.scan_launches + st.eval_launches + st.repart_launches) * args.steps // prof_steps),
            "e2e": {"value": N_total * L * T / (e2e_ms * 1e-3), "unit": UNIT, "ms_per_step": e2e_ms,
                    "h2d_bytes_per_step": int(sum(c.nbytes for c in cols_p) + cls_p.nbytes),
                    "d2h_bytes_per_step": int(nrec * C.sizeof(_lib.GrowRecord))},
            "roofline": {"bound": "hbm", "kernel": {0: "hist_smem_kernel", 1: "hist_global_kernel", 2: "hist_rot_kernel", 3: "hist_joint_kernel", 4: "hist_sorted_kernel"}.get(st.last_hist_variant, "?"),
                         "achieved": achieved, "peak": peak, "peak_source": peak_src, "unit": "GB/s", "frac": achieved / peak,
                         "frac_of_spec_8000": achieved / 8000.0, "traffic": traffic, "traffic_source": traffic_src,
                         "alg_bytes_per_launch": alg_bytes_per_launch, "ms_per_launch": hist_ms_per_launch,
                         "launch": {"grid": st.last_hist_grid, "block": st.last_hist_block, "smem": st.last_hist_smem, "replicas": st.last_hist_replicas},
                         # the unit that actually bounds this pass: one shared-memory counter update per (row, tree, attribute)
                         "shared_atomics": {"updates_per_launch": n_local * T * 32, "achieved_T_per_s": n_local * T * 32 / (hist_ms_per_launch * 1e-3) / 1e12 if hist_ms_per_launch > 0 else 0.0,
                                            "peak_T_per_s": 6.2, "peak_source": "red.shared.add.u32, conflict-free, 32 warps per SM: tools/microbench/atoms_bench.cu, profiles/atoms_bench_r02.txt",
                                            "frac": n_local * T * 32 / (hist_ms_per_launch * 1e-3) / 6.2e12 if hist_ms_per_launch > 0 else 0.0}},
            "clocks": sampler.summary(),
        }
        if not args.no_cpu_baseline and world == 1:
            threads = os.cpu_count() or 1
            rows_c, T_c = min(N_total, C3_CPU_ROWS), min(T, C3_CPU_TREES)
            dt, _, _, _ = c3_oracle_forest(rows_c, T_c, L, threads)
            rows_1 = min(N_total, C3_CPU_ROWS_1T)
            dt1, _, _, _ = c3_oracle_forest(rows_1, T_c, L, 1)
            line["cpu_baseline"] = {"value": rows_c * L * T_c / dt, "unit": UNIT, "cores": threads, "kind": "port",
                                    "sample": f"first {rows_c} rows x {T_c} trees x {L} levels of the same workload, {dt:.2f} s on {threads} threads",
                                    "single_thread": {"value": rows_1 * L * T_c / dt1, "cores": 1,
                                                      "sample": f"first {rows_1} rows x {T_c} trees x {L} levels, {dt1:.2f} s (Hadoop local mode runs its tasks serially)"}}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0



def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c3", choices=["c3", "c2", "c5", "c4", "c1"])
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--rows", type=int, default=None, help="c3: total rows (default 100 M); c2: rows per GPU (default 10 M)")
    ap.add_argument("--trees", type=int, default=None, help="c3: bagged trees (default 100)")
    ap.add_argument("--levels", type=int, default=None)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-verify", action="store_true", help="skip the in-run parity gate (c3)")
    args = ap.parse_args()
    if args.workload == "c2":
        return run_c2(args)
    if args.workload in ("c5", "c4", "c1"):
        # secondary lines (one GPU): bench_extra.py
        if int(os.environ.get("WORLD_SIZE", "1")) > 1 or args.impl == "reference":
            raise SystemExit(f"bench.py --workload {args.workload} is a one-GPU line of our arm only; the scaling runs and the reference arm use c3 (default) or c2")
        import bench_extra
        line = getattr(bench_extra, "run_" + args.workload)(args, ClockSampler, measured_peak)
        print(json.dumps(line))
        return 0
    return run_c3(args)


if __name__ == "__main__":
    sys.exit(main())
