#!/usr/bin/env python
"""bench.py -- split-evaluation throughput of the tree-building hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--rows R] [--levels L]

Workload at every N: config C2 of BASELINE.json -- the call_hangup decision tree (resource/call_hangup.json
schema: 4 feature attributes, 18 bins, 18 binary candidate splits, 2 classes), 10 M synthetic rows PER
GPU (weak scaling: rows shard across ranks, the per-level histograms are exchanged through NVLink peer memory
inside the scan kernel; AVT_BENCH_NCCL=1 uses one NCCL all-reduce per level instead),
entropy, attribute strategy `all`, 8 expansions below the root.  A "step" builds the whole tree once:
8 x (histogram -> fused entropy scan/arg-min -> children statistics -> repartition).

  value   rows/sec = ranks * rows * levels / step time, inputs resident in HBM; step time = CUDA events the
          library records on its own stream around everything avt_grow_levels enqueues, result copies included
          (avt_stats.last_grow_ms; max over ranks);  L2 is flushed between steps.  api_ms_per_step is the same
          step timed by events recorded from Python around the call (adds ctypes marshalling on both sides)
  e2e     same metric through the C-ABI from HOST buffers: avt_dataset_load (H2D from pinned memory) +
          avt_forest_init + avt_grow_levels + read-back of the node ids, wall clock
  roofline  dominant kernel = the histogram pass: algorithmic bytes / its CUDA-event time
  cpu_baseline  the C oracle (oracle/hist_oracle.c, "port": the reference is JVM code and cannot run on
          this box) on all host cores, bounded sample of the same workload

`--impl reference` times that CPU port as the reference arm (rank 0 only under torchrun).
"""
from __future__ import annotations

import argparse
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
ROWS_DEFAULT = 10_000_000
LEVELS_DEFAULT = 8
CPU_SAMPLE_ROWS = 10_000_000   # the CPU arm runs the full C2 shard: 8 levels over 10 M rows on all host cores


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


def run_reference(args):
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--rows", type=int, default=ROWS_DEFAULT)
    ap.add_argument("--levels", type=int, default=LEVELS_DEFAULT)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

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
            "roofline": {"bound": "hbm", "kernel": "hist_smem_kernel" if st.last_hist_variant == 0 else "hist_global_kernel",
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


if __name__ == "__main__":
    sys.exit(main())
