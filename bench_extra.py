"""Secondary bench.py workloads (one GPU each): the BASELINE.json configs that are not the headline line.

  c5  wide synthetic: 20 M rows x 512 int attributes x 64 bins, gini, 12 expansions, strategy `all`
  c4  ClassPartitionGenerator sweep: 10 M rows x 50 attributes (40 categorical with 1 000 Zipf-distributed levels,
      10 int attributes with 20 buckets and up to 3 segments), explicit candidate lists for the categoricals
      (999 prefix splits + 1 000 one-vs-rest), gini and entropy
  c1  the call_hangup tutorial, 100 K rows, detr.properties knobs, through the reference-named tool (CSV file ->
      DecisionPathList JSON); CPU arm = the literal, string-keyed restatement (the MapReduce formulation's cost model)

Every function returns the JSON line's dictionary; bench.py prints it.  All compute goes through libavenir_tree.so; the
oracle is used as the checker (before the timed region) and as the timed CPU baseline only.
"""
from __future__ import annotations

import ctypes as C
import json
import os
import tempfile
import time

import numpy as np

METRIC = "split_eval_rows_per_sec"
UNIT = "rows/s"
ROOT = os.path.dirname(os.path.abspath(__file__))


def _kernel_name(v):
    return {0: "hist_smem_kernel", 1: "hist_global_kernel", 2: "hist_rot_kernel", 3: "hist_joint_kernel", 4: "hist_sorted_kernel"}.get(v, "?")


def _roofline(st, peak, peak_src, traffic_file):
    ms = st.hist_ms / max(1, st.hist_launches)
    alg = st.hist_alg_bytes / max(1, st.hist_launches)
    achieved = alg / (ms * 1e-3) / 1e9 if ms > 0 else 0.0
    traffic = src = None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", traffic_file)))
        traffic, src = tj.get("dram_bytes_per_launch"), tj.get("source")
    except Exception:
        pass
    return {"bound": "hbm", "kernel": _kernel_name(st.last_hist_variant), "achieved": achieved, "peak": peak, "peak_source": peak_src, "unit": "GB/s",
            "frac": achieved / peak, "frac_of_spec_8000": achieved / 8000.0, "traffic": traffic, "traffic_source": src,
            "alg_bytes_per_launch": alg, "ms_per_launch": ms,
            "launch": {"grid": st.last_hist_grid, "block": st.last_hist_block, "smem": st.last_hist_smem, "replicas": st.last_hist_replicas}}


# ----------------------------------------------------------------------------------------------------------------------
# C5
# ----------------------------------------------------------------------------------------------------------------------
C5_ROWS, C5_ATTRS, C5_BINS, C5_LEVELS, C5_SIGNAL = 20_000_000, 512, 64, 12, 16
C5_VERIFY_ROWS, C5_VERIFY_LEVELS = 200_000, 8
C5_CPU_ROWS, C5_CPU_ROWS_1T = 1_000_000, 200_000


def _c5_oracle(rows, levels, threads, want_state=False):
    from avenir_b200 import synth
    from avenir_b200._lib import CRIT_GINI
    from avenir_b200.engine import binary_split_table, grow_levels
    from tests.helpers import OracleEngine
    n_bins = [C5_BINS] * C5_ATTRS
    cols, cls = synth.binned_dataset(rows, n_bins, 2, seed=synth.DEFAULT_SEED, signal_attrs=C5_SIGNAL)
    table = binary_split_table(n_bins, [False] * C5_ATTRS)
    o = OracleEngine(threads=threads)
    o.load_dataset(cols, cls, n_bins, 2)
    o.set_splits(table)
    o.forest_init(1)
    t0 = time.perf_counter()
    hist, frontier = grow_levels(o, table, levels, CRIT_GINI, record=want_state)
    return time.perf_counter() - t0, hist, frontier, o, (cols, cls, n_bins, table)


def run_c5(args, ClockSampler, measured_peak):
    import torch
    from avenir_b200 import build, synth
    from avenir_b200._lib import CRIT_GINI
    from avenir_b200.engine import Engine, binary_split_table
    from tests.helpers import bits
    build.build()
    N, L = args.rows or C5_ROWS, args.levels or C5_LEVELS
    n_bins = [C5_BINS] * C5_ATTRS
    table = binary_split_table(n_bins, [False] * C5_ATTRS)
    eng = Engine(0)
    verify = None
    if not args.no_verify:
        rows_v, L_v = min(N, C5_VERIFY_ROWS), min(L, C5_VERIFY_LEVELS)
        _, hist, fo, o, (cols, cls, _, _) = _c5_oracle(rows_v, L_v, os.cpu_count() or 1, want_state=True)
        eng.load_dataset(cols, cls, n_bins, 2)
        eng.set_splits(table)
        eng.forest_init(1)
        (rec, nrec), frontier = eng.grow(L_v, CRIT_GINI, [1], max_records=1 << 14)
        want = [x for level in hist for x in level]
        if frontier != fo or nrec != len(want):
            raise SystemExit(f"bench.py verify (c5): frontier {frontier} / {nrec} records, oracle {fo} / {len(want)}")
        for i, (t, nd, sp, pops, av, ties) in enumerate(want):
            r = rec[i]
            if (r.node, r.split, r.n_ties) != (nd, sp, ties) or [p for p in r.child_pop if p] != pops or bits(r.min_av) != bits(av):
                raise SystemExit(f"bench.py verify (c5): record {i} differs from the oracle")
        if not np.array_equal(eng.node_ids(0), o.node_ids(0)):
            raise SystemExit("bench.py verify (c5): final nodes differ from the oracle")
        verify = {"rows": rows_v, "levels": L_v, "records": nrec, "result": "records and final nodes equal the C oracle"}
    cols, cls = synth.binned_dataset_device(N, n_bins, 2, seed=synth.DEFAULT_SEED, signal_attrs=C5_SIGNAL, device="cuda")
    torch.cuda.synchronize()
    eng.load_dataset_device([c.data_ptr() for c in cols], cls.data_ptr(), N, n_bins, 2)
    eng.set_splits(table)
    eng.forest_init(1)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")

    def step():
        eng.forest_reset()
        flush.zero_()
        torch.cuda.synchronize()
        _, frontier = eng.grow(L, CRIT_GINI, [1])
        return eng.stats().last_grow_ms, frontier

    for _ in range(max(3, args.warmup)):
        step()
    eng.set_profiling(False)
    sampler = ClockSampler(0)
    sampler.start()
    dev_ms = 0.0
    t0 = time.perf_counter()
    for _ in range(args.steps):
        ms, frontier = step()
        dev_ms += ms
    wall = time.perf_counter() - t0
    eng.set_profiling(True)
    eng.reset_stats()
    step()
    st = eng.stats()
    sampler.stop_flag = True
    sampler.join()
    # e2e from pinned host buffers: H2D of the columns + level loop + the records back
    pinned = [torch.empty(N, dtype=torch.uint8).pin_memory() for _ in cols]
    for p_, c in zip(pinned, cols):
        p_.copy_(c)
    pcls = torch.empty(N, dtype=torch.uint8).pin_memory()
    pcls.copy_(cls)
    torch.cuda.synchronize()
    del cols, cls
    torch.cuda.empty_cache()
    cols_p, cls_p = [p_.numpy() for p_ in pinned], pcls.numpy()

    def step_e2e():
        t = time.perf_counter()
        eng.load_dataset(cols_p, cls_p, n_bins, 2)
        eng.set_splits(table)
        eng.forest_init(1)
        (rec, nrec), _ = eng.grow(L, CRIT_GINI, [1], max_records=1 << 14)
        return time.perf_counter() - t, nrec

    step_e2e()
    e2e_s, nrec = 0.0, 0
    e2e_steps = max(1, min(args.steps, 2))
    for _ in range(e2e_steps):
        dt, nrec = step_e2e()
        e2e_s += dt
    ms_step, e2e_ms = dev_ms / args.steps, e2e_s / e2e_steps * 1e3
    peak, peak_src = measured_peak()
    from avenir_b200 import _lib
    line = {"metric": METRIC, "value": N * L / (ms_step * 1e-3), "unit": UNIT, "n_gpus": 1, "steps": args.steps, "warmup": max(3, args.warmup),
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u32 counts / f64 impurity", "data": "synthetic",
            "config": {"workload": f"C5: wide synthetic, {N} rows x 512 int attributes x 64 bins, 2 classes, gini, strategy all, {L} expansions (histogram-memory stress)",
                       "rows": N, "attrs": C5_ATTRS, "bins": C5_ATTRS * C5_BINS, "splits": len(table), "levels": L, "trees": 1,
                       "l2": "flushed between steps; the 10 GB of codes are far larger than L2", "final_frontier_nodes": int(frontier[0])},
            "verify": verify if verify is not None else "skipped (--no-verify)",
            "kernel_ms_per_step": {"histogram": st.hist_ms, "scan_plan": st.scan_ms, "repartition": st.repart_ms, "note": "CUDA events per launch, one extra profiled step"},
            "wall_ms_per_step": wall * 1e3 / args.steps,
            "gpu_launches": int((st.hist_launches + st.scan_launches + st.repart_launches) * args.steps),
            "e2e": {"value": N * L / (e2e_ms * 1e-3), "unit": UNIT, "ms_per_step": e2e_ms, "h2d_bytes_per_step": int(sum(c.nbytes for c in cols_p) + cls_p.nbytes),
                    "d2h_bytes_per_step": int(nrec * C.sizeof(_lib.GrowRecord))},
            "roofline": _roofline(st, peak, peak_src, "hist_traffic_c5.json"), "clocks": sampler.summary()}
    if not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        rows_c, L_c = min(N, C5_CPU_ROWS), min(L, 8)
        dt, *_ = _c5_oracle(rows_c, L_c, threads)
        rows_1 = min(N, C5_CPU_ROWS_1T)
        dt1, *_ = _c5_oracle(rows_1, L_c, 1)
        line["cpu_baseline"] = {"value": rows_c * L_c / dt, "unit": UNIT, "cores": threads, "kind": "port",
                                "sample": f"first {rows_c} rows x {L_c} levels of the same workload, {dt:.2f} s on {threads} threads",
                                "single_thread": {"value": rows_1 * L_c / dt1, "cores": 1, "sample": f"first {rows_1} rows x {L_c} levels, {dt1:.2f} s"}}
    return line


# ----------------------------------------------------------------------------------------------------------------------
# C4
# ----------------------------------------------------------------------------------------------------------------------
C4_ROWS, C4_CAT, C4_LEVELS_PER_CAT, C4_INT, C4_INT_BUCKETS = 10_000_000, 40, 1000, 10, 20
C4_VERIFY_ROWS, C4_CPU_ROWS = 300_000, 2_000_000


def _c4_cdf():
    p = 1.0 / np.arange(1, C4_LEVELS_PER_CAT + 1, dtype=np.float64) ** 1.1
    return np.cumsum(p / p.sum())


def c4_table():
    """Candidate splits of the C4 sweep: per categorical attribute 999 prefix splits (levels 0..k | rest, in
    `cardinality` order) + 1 000 one-vs-rest splits, as level -> segment tables (exhaustive set partitions of 1 000 levels are
    2^999 splits -- the reference's own generator cannot enumerate them); per int attribute every 1- and 2-point split
    of the bucket grid (maxSplit 3; explore/ClassPartitionGenerator.java:273-316 order: [p0], [p0,p1], ..., [p1], ...)."""
    from avenir_b200._lib import SPLIT_LUT
    from avenir_b200.engine import SplitTable
    t = SplitTable()
    K = C4_LEVELS_PER_CAT
    b = np.arange(K)
    prefix = (b[None, :] > np.arange(K - 1)[:, None]).astype(np.uint8)        # split k: levels 0..k -> segment 0, the rest -> 1
    one_vs_rest = (b[None, :] != np.arange(K)[:, None]).astype(np.uint8)      # split k: level k -> segment 0, the rest -> 1
    per_attr = np.concatenate([prefix, one_vs_rest]).reshape(-1)              # (2K - 1) tables of K entries
    n_per = 2 * K - 1
    t.lut = np.tile(per_attr, C4_CAT)
    for a in range(C4_CAT):
        t.attr.extend([a] * n_per)
        t.kind.extend([SPLIT_LUT] * n_per)
        t.n_seg.extend([2] * n_per)
        t.off.extend(range(a * n_per * K, (a + 1) * n_per * K, K))
    B = C4_INT_BUCKETS
    for a in range(C4_CAT, C4_CAT + C4_INT):
        for p0 in range(B - 1):
            t.add_numeric(a, [p0])
            for p1 in range(p0 + 1, B - 1):
                t.add_numeric(a, [p0, p1])
    return t


def c4_dataset(n, start=0, device=None):
    """Counter-based C4 columns: Zipf(1.1) categorical levels by inverse CDF, uniform int buckets, class from two categorical
    head/tail indicators and one int attribute with 20 % label noise.  numpy (device None) and torch twins are bit-identical."""
    from avenir_b200 import synth
    cdf = _c4_cdf()
    n_bins = [C4_LEVELS_PER_CAT] * C4_CAT + [C4_INT_BUCKETS] * C4_INT
    if device is None:
        rows = np.arange(start, start + n, dtype=np.uint64)
        cols, score = [], np.zeros(n, dtype=np.int64)
        for a in range(C4_CAT):
            code = np.minimum(np.searchsorted(cdf, synth.uniform(synth.DEFAULT_SEED, rows, 64 + a), side="right"), C4_LEVELS_PER_CAT - 1)
            cols.append(code.astype(np.uint16))
            if a < 2:
                score += (code < 3).astype(np.int64)
        for a in range(C4_INT):
            code = np.minimum((synth.uniform(synth.DEFAULT_SEED, rows, 128 + a) * C4_INT_BUCKETS).astype(np.int64), C4_INT_BUCKETS - 1)
            cols.append(code.astype(np.uint8))
            if a == 0:
                score += (code >= 10).astype(np.int64)
        cls = score % 2
        flip = synth.uniform(synth.DEFAULT_SEED, rows, 7) < 0.2
        cls = np.where(flip, 1 - cls, cls).astype(np.uint8)
        return cols, cls, n_bins
    import torch
    M64 = (1 << 64) - 1

    def s64(v):
        v &= M64
        return v - (1 << 64) if v >= (1 << 63) else v

    def lsr(z, k):
        return (z >> k) & ((1 << (64 - k)) - 1)

    def u01(rows, draw):
        x = rows * 1024 + s64(((synth.DEFAULT_SEED << 40) & M64) + draw) + s64(0x9E3779B97F4A7C15)
        z = (x ^ lsr(x, 30)) * s64(0xBF58476D1CE4E5B9)
        z = (z ^ lsr(z, 27)) * s64(0x94D049BB133111EB)
        z = z ^ lsr(z, 31)
        return lsr(z, 11).to(torch.float64) * (1.0 / 9007199254740992.0)

    cdf_t = torch.from_numpy(cdf).to(device)
    cols = [torch.empty(n, dtype=torch.int16 if a < C4_CAT else torch.uint8, device=device) for a in range(C4_CAT + C4_INT)]
    cls = torch.empty(n, dtype=torch.uint8, device=device)
    chunk = 4_000_000
    for s in range(0, n, chunk):
        m = min(chunk, n - s)
        rows = torch.arange(start + s, start + s + m, dtype=torch.int64, device=device)
        score = torch.zeros(m, dtype=torch.int64, device=device)
        for a in range(C4_CAT):
            code = torch.clamp(torch.searchsorted(cdf_t, u01(rows, 64 + a), right=True), max=C4_LEVELS_PER_CAT - 1)
            cols[a][s:s + m] = code.to(torch.int16)
            if a < 2:
                score += (code < 3).to(torch.int64)
        for a in range(C4_INT):
            code = torch.clamp((u01(rows, 128 + a) * C4_INT_BUCKETS).to(torch.int64), max=C4_INT_BUCKETS - 1)
            cols[C4_CAT + a][s:s + m] = code.to(torch.uint8)
            if a == 0:
                score += (code >= 10).to(torch.int64)
        c = score % 2
        flip = u01(rows, 7) < 0.2
        cls[s:s + m] = torch.where(flip, 1 - c, c).to(torch.uint8)
    return cols, cls, n_bins


def _c4_oracle(rows, threads, criteria):
    from tests.helpers import OracleEngine
    cols, cls, n_bins = c4_dataset(rows)
    o = OracleEngine(threads=threads)
    o.load_dataset(cols, cls, n_bins, 2)
    o.set_splits(c4_table())
    o.forest_init(1)
    t0 = time.perf_counter()
    h = o.histogram([1], fetch=True)
    out = [o.cpg_sweep(0, c) for c in criteria]
    return time.perf_counter() - t0, h, out, (cols, cls, n_bins)


def run_c4(args, ClockSampler, measured_peak):
    import torch
    from avenir_b200 import _lib, build
    from avenir_b200._lib import CRIT_ENTROPY, CRIT_GINI
    from avenir_b200.engine import Engine
    from tests.helpers import bits
    build.build()
    N = args.rows or C4_ROWS
    table = c4_table()
    criteria = [CRIT_GINI, CRIT_ENTROPY]
    eng = Engine(0)
    verify = None
    if not args.no_verify:
        rows_v = min(N, C4_VERIFY_ROWS)
        _, h_o, stats_o, (cols, cls, n_bins) = _c4_oracle(rows_v, os.cpu_count() or 1, criteria)
        eng.load_dataset(cols, cls, n_bins, 2)
        eng.set_splits(table)
        eng.forest_init(1)
        if not np.array_equal(eng.histogram([1], fetch=True), h_o):
            raise SystemExit("bench.py verify (c4): root histogram differs from the oracle")
        for c, so in zip(criteria, stats_o):
            sg = eng.cpg_sweep(0, c)
            for i in range(len(table)):
                if bits(sg[i].stat) != bits(so[i].stat) or bits(sg[i].split_info) != bits(so[i].split_info):
                    raise SystemExit(f"bench.py verify (c4): split {i} criterion {c}: stat {sg[i].stat} vs oracle {so[i].stat}")
        verify = {"rows": rows_v, "splits": len(table), "criteria": ["giniIndex", "entropy"], "result": "counts equal, stat and split_info bit-equal to the C oracle"}
    cols, cls, n_bins = c4_dataset(N, device="cuda")
    torch.cuda.synchronize()
    eng.load_dataset_device([c.data_ptr() for c in cols], cls.data_ptr(), N, n_bins, 2)
    eng.set_splits(table)
    eng.forest_init(1)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")

    # result arrays (528 bytes per split and criterion) in page-locked memory, reused by every step
    from avenir_b200.engine import CpgStat
    res_pinned = [torch.empty(C.sizeof(CpgStat) * len(table), dtype=torch.uint8).pin_memory() for _ in criteria]
    res = [(CpgStat * len(table)).from_address(p_.data_ptr()) for p_ in res_pinned]

    def step():
        flush.zero_()
        torch.cuda.synchronize()
        eng.timer_begin()
        eng.histogram([1])
        out = [eng.cpg_sweep(0, c, out=r_) for c, r_ in zip(criteria, res)]
        return eng.timer_end(), out

    for _ in range(max(3, args.warmup)):
        step()
    sampler = ClockSampler(0)
    sampler.start()
    eng.reset_stats()
    dev_ms = 0.0
    t0 = time.perf_counter()
    for _ in range(args.steps):
        ms, out = step()
        dev_ms += ms
    wall = time.perf_counter() - t0
    st = eng.stats()
    sampler.stop_flag = True
    sampler.join()
    pinned = [torch.empty(N, dtype=c.dtype).pin_memory() for c in cols]
    for p_, c in zip(pinned, cols):
        p_.copy_(c)
    pcls = torch.empty(N, dtype=torch.uint8).pin_memory()
    pcls.copy_(cls)
    torch.cuda.synchronize()
    del cols, cls
    torch.cuda.empty_cache()
    cols_p = [p_.numpy().view(np.uint16) if p_.dtype == torch.int16 else p_.numpy() for p_ in pinned]
    cls_p = pcls.numpy()

    def step_e2e():
        t = time.perf_counter()
        eng.load_dataset(cols_p, cls_p, n_bins, 2)
        eng.set_splits(table)
        eng.forest_init(1)
        eng.histogram([1])
        out = [eng.cpg_sweep(0, c) for c in criteria]
        return time.perf_counter() - t

    step_e2e()
    e2e_steps = max(1, min(args.steps, 3))
    e2e_ms = sum(step_e2e() for _ in range(e2e_steps)) / e2e_steps * 1e3
    ms_step = dev_ms / args.steps
    peak, peak_src = measured_peak()
    line = {"metric": METRIC, "value": N / (ms_step * 1e-3), "unit": UNIT, "n_gpus": 1, "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": ms_step,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u32 counts / f64 statistics", "data": "synthetic",
            "config": {"workload": f"C4: ClassPartitionGenerator sweep, {N} rows x 50 attributes (40 categorical x 1000 Zipf levels, 10 int x 20 buckets), "
                                   f"{len(table)} candidate splits (prefix + one-vs-rest lists, 1- and 2-point int splits), giniIndex and entropy, one node (all rows)",
                       "rows": N, "attrs": 50, "bins": sum(n_bins), "splits": len(table), "levels": 1, "l2": "flushed between steps"},
            "verify": verify if verify is not None else "skipped (--no-verify)",
            "kernel_ms_per_step": {"histogram": st.hist_ms / args.steps, "cpg_sweeps": st.scan_ms / args.steps,
                                   "note": "CUDA events per launch (one warp per split); the rest of ms_per_step is the copy of the results (528 bytes per split and criterion) into page-locked host arrays"},
            "wall_ms_per_step": wall * 1e3 / args.steps, "gpu_launches": int(st.hist_launches + 2 * args.steps),
            "e2e": {"value": N / (e2e_ms * 1e-3), "unit": UNIT, "ms_per_step": e2e_ms, "h2d_bytes_per_step": int(sum(c.nbytes for c in cols_p) + cls_p.nbytes),
                    "d2h_bytes_per_step": int(2 * len(table) * C.sizeof(_lib.CpgStat))},
            "roofline": _roofline(st, peak, peak_src, "hist_traffic_c4.json"), "clocks": sampler.summary()}
    if not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        rows_c = min(N, C4_CPU_ROWS)
        dt, *_ = _c4_oracle(rows_c, threads, criteria)
        dt1, *_ = _c4_oracle(min(rows_c, 500_000), 1, criteria)
        line["cpu_baseline"] = {"value": rows_c / dt, "unit": UNIT, "cores": threads, "kind": "port",
                                "sample": f"first {rows_c} rows of the same sweep, {dt:.2f} s on {threads} threads",
                                "single_thread": {"value": min(rows_c, 500_000) / dt1, "cores": 1, "sample": f"{min(rows_c, 500_000)} rows, {dt1:.2f} s"}}
    return line


# ----------------------------------------------------------------------------------------------------------------------
# C1
# ----------------------------------------------------------------------------------------------------------------------
C1_ROWS, C1_LITERAL_ROWS, C1_TOOL_ROWS = 100_000, 40_000, 2_000_000
C1_CONF = {"field.delim.regex": ",", "field.delim.out": ",", "dtb.split.algorithm": "giniIndex", "dtb.split.attribute.selection.strategy": "notUsedYet",
           "dtb.sub.sampling.strategy": "none", "dtb.path.stopping.strategy": "maxDepth", "dtb.max.depth.limit": "2"}   # resource/detr.properties


def run_c1(args, ClockSampler, measured_peak):
    import torch
    from avenir_b200 import _lib, build, synth
    from avenir_b200._lib import CRIT_GINI
    from avenir_b200.engine import Engine, binary_split_table
    from oracle import literal
    from tests import helpers
    build.build()
    N = args.rows or C1_ROWS
    L = 3   # the tutorial's three jobs: root pass + two expansions (resource/call_hangup_dec_tree_tutorial.txt)
    lib = _lib.load()
    tmp = tempfile.mkdtemp(prefix="avt_c1_")
    schema = helpers.write_schema(os.path.join(tmp, "call_hangup.json"), "call_hangup.json")
    lines = synth.call_hangup_csv(N)
    data = os.path.join(tmp, "in.txt")
    open(data, "w").write("\n".join(lines) + "\n")
    conf = dict(C1_CONF, **{"dtb.feature.schema.file.path": os.path.join(tmp, "call_hangup.json"), "dtb.decision.file.path.out": os.path.join(tmp, "dec.json")})
    pp = os.path.join(tmp, "detr.properties")
    open(pp, "w").write("".join(f"{k}={v}\n" for k, v in conf.items()))
    tool_args = [f"-Dconf.path={pp}", f"-Davt.dtb.num.levels={L}", data, os.path.join(tmp, "out")]

    def tool():
        t = time.perf_counter()
        rc, err = helpers.tool_run(lib, "org.avenir.tree.DecisionTreeBuilder", tool_args)
        if rc != 0:
            raise SystemExit("bench.py (c1): DecisionTreeBuilder failed: " + err)
        return time.perf_counter() - t

    # checker + timed CPU arm: the literal restatement on a bounded prefix, compared with the tool on the same prefix
    verify = None
    n_lit = min(N, C1_LITERAL_ROWS)
    t0 = time.perf_counter()
    want = literal.dtb_build(lines[:n_lit], literal.load_properties(pp), literal.FeatureSchema(schema), L - 1, seed=0)
    lit_s = time.perf_counter() - t0
    if not args.no_verify:
        d2 = os.path.join(tmp, "prefix.txt")
        open(d2, "w").write("\n".join(lines[:n_lit]) + "\n")
        rc, err = helpers.tool_run(lib, "org.avenir.tree.DecisionTreeBuilder", [tool_args[0], tool_args[1], d2, os.path.join(tmp, "out_prefix")])
        got = [json.load(open(os.path.join(tmp, f"dec.json.level{k}"))) for k in range(len(want))]
        if rc != 0 or got != want:
            raise SystemExit("bench.py verify (c1): the tool's DecisionPathList files differ from the literal restatement")
        verify = {"rows": n_lit, "jobs": L, "result": "every level's DecisionPathList JSON equals the literal restatement"}
    for _ in range(max(3, args.warmup)):
        tool()
    sampler = ClockSampler(0)
    sampler.start()
    e2e_s = sum(tool() for _ in range(args.steps)) / args.steps
    # the same tool on a larger file, with the rows encoded on the device (default) and by the C++ host encoder (avt.encode=host)
    e2e_tool = None
    n_big = 0 if args.rows else C1_TOOL_ROWS
    if n_big:
        big = os.path.join(tmp, "big.txt")
        open(big, "w").write("\n".join(synth.call_hangup_csv(n_big)) + "\n")
        e2e_tool = {"rows": n_big, "jobs": L, "csv_bytes": os.path.getsize(big)}
        for mode in ("device", "host"):
            a2 = [tool_args[0], tool_args[1], f"-Davt.encode={mode}", big, os.path.join(tmp, "out_big_" + mode)]
            best = None
            for _ in range(3):
                t = time.perf_counter()
                rc, err = helpers.tool_run(lib, "org.avenir.tree.DecisionTreeBuilder", a2)
                dt = time.perf_counter() - t
                if rc != 0:
                    raise SystemExit("bench.py (c1): DecisionTreeBuilder failed on the large file: " + err)
                best = dt if best is None else min(best, dt)
            e2e_tool[mode + "_encode_ms"] = best * 1e3
            e2e_tool[mode + "_encode_rows_per_s"] = n_big * L / best
        e2e_tool["note"] = ("avt_tool_run from the CSV file to the DecisionPathList files and the re-keyed part files (best of 3); reading the file, the class-value "
                            "pass and writing the part files are host work in both modes")
    # resident form: encoded columns in HBM, the level loop with the same knobs
    cols, cls, n_bins, _ = synth.call_hangup_encoded(N)
    table = binary_split_table(n_bins, [True, True, True, False])
    eng = Engine(0)
    eng.load_dataset(cols, cls, n_bins, 2)
    eng.set_splits(table)
    eng.forest_init(1)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")

    def step():
        eng.forest_reset()
        flush.zero_()
        torch.cuda.synchronize()
        eng.grow(L - 1, CRIT_GINI, [1], strategy=_lib.ATTRS_NOT_USED_YET)
        return eng.stats().last_grow_ms

    for _ in range(3):
        step()
    eng.reset_stats()
    ms_step = sum(step() for _ in range(args.steps)) / args.steps
    st = eng.stats()
    sampler.stop_flag = True
    sampler.join()
    peak, peak_src = measured_peak()
    line = {"metric": METRIC, "value": N * (L - 1) / (ms_step * 1e-3), "unit": UNIT, "n_gpus": 1, "steps": args.steps, "warmup": max(3, args.warmup),
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u32 counts / f64 impurity", "data": "synthetic",
            "config": {"workload": f"C1: call_hangup tutorial (call_hangup.json, detr.properties: giniIndex, notUsedYet, maxDepth 2), {N} rows, root pass + 2 expansions; "
                                   "value = the two expansions on resident columns, e2e = the reference-named tool from the CSV file to the three DecisionPathList files",
                       "rows": N, "levels": L - 1, "jobs": L, "l2": "flushed between steps"},
            "verify": verify if verify is not None else "skipped (--no-verify)",
            "gpu_launches": int(st.hist_launches + st.scan_launches + st.repart_launches),
            "e2e": {"value": N * L / e2e_s, "unit": UNIT, "ms_per_step": e2e_s * 1e3, "h2d_bytes_per_step": int(os.path.getsize(data)),
                    "d2h_bytes_per_step": 0, "note": "avt_tool_run: reads the CSV file, hands the text to the device encoder, three level passes, writes JSON and part-r files; rows x 3 jobs / wall time"},
            "tree_build_ms": {"tool_csv_to_json": e2e_s * 1e3, "level_loop_resident": ms_step},
            "e2e_tool": e2e_tool,
            "roofline": _roofline(st, peak, peak_src, "hist_traffic.json"), "clocks": sampler.summary(),
            "cpu_baseline": {"value": n_lit * L / lit_s, "unit": UNIT, "cores": 1, "kind": "port",
                             "sample": f"literal (string-keyed, one record per row x split x predicate) restatement of the three jobs on the first {n_lit} rows: {lit_s:.1f} s, "
                                       "one thread -- the MapReduce formulation's cost model; Hadoop local mode itself: unavailable on this box"}}
    return line
