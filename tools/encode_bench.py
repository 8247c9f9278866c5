#!/usr/bin/env python
"""Measurement of the on-device CSV encode: call_hangup CSV -> columnar bin codes.

    python tools/encode_bench.py [--rows R] [--steps K]

Prints one JSON line: rows/s and text GB/s of avt_dataset_load_csv from a pinned host buffer (wall clock around the
C-ABI call = host->device copy of the text + line index + encode + installation), the device-only phases from the
library's events, the HBM roofline of the encode kernel (algorithmic bytes = text bytes read + codes written), and the
CPU baseline: the C restatement (oracle/hist_oracle.c:ora_encode_csv, "port") on all host cores."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=2_000_000)
    ap.add_argument("--steps", type=int, default=5)
    args = ap.parse_args()
    import torch
    from avenir_b200 import synth
    from avenir_b200.engine import Engine
    from tests import helpers
    attrs = [(1, "categorical", list(synth.CUST_TYPES)), (3, "categorical", list(synth.ISSUE_CARD)),
             (4, "categorical", list(synth.TOD)), (5, "int", [int(p) for p in synth.HOLD_POINTS])]
    text = ("\n".join(synth.call_hangup_csv(args.rows)) + "\n").encode()
    pinned = torch.empty(len(text), dtype=torch.uint8).pin_memory()
    pinned.numpy()[:] = np.frombuffer(text, dtype=np.uint8)
    buf = (pinned.data_ptr(), len(text))
    import ctypes as C
    host_text = (C.c_char * buf[1]).from_address(buf[0])
    g = Engine(0)
    for _ in range(2):
        g.load_csv(host_text, ",", attrs, 6, ["T", "F"])
    wall, h2d, idx, encode = [], [], [], []
    for _ in range(args.steps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        rows, _ = g.load_csv(host_text, ",", attrs, 6, ["T", "F"])
        wall.append(time.perf_counter() - t0)
        st = g.stats()
        h2d.append(st.last_csv_h2d_ms)
        idx.append(st.last_csv_index_ms)
        encode.append(st.last_csv_encode_ms)
    w = float(np.median(wall))
    enc_ms = float(np.median(encode))
    alg_bytes = len(text) + rows * 5 + rows * 8     # text read, 4 codes + class written, line offsets read
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6560.6)) if isinstance(peaks, dict) else 6560.6
    threads = os.cpu_count() or 1
    best = None
    for _ in range(3):
        t0 = time.perf_counter()
        helpers.oracle_encode_csv(text, ",", attrs, 6, ["T", "F"], threads=threads)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    t0 = time.perf_counter()
    helpers.oracle_encode_csv(text, ",", attrs, 6, ["T", "F"], threads=1)
    single = time.perf_counter() - t0
    line = {"metric": "csv_encode_rows_per_sec", "unit": "rows/s", "value": rows / w, "text_gb_per_s": len(text) / w / 1e9,
            "rows": rows, "text_bytes": len(text), "steps": args.steps, "wall_ms": w * 1e3,
            "device_ms": {"h2d_text": float(np.median(h2d)), "line_index": float(np.median(idx)), "encode": enc_ms},
            "roofline": {"bound": "hbm", "kernel": "csv_encode_kernel", "achieved": alg_bytes / (enc_ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                         "frac": alg_bytes / (enc_ms * 1e-3) / 1e9 / peak, "alg_bytes_per_launch": alg_bytes, "traffic": None},
            "cpu_baseline": {"value": rows / best, "unit": "rows/s", "cores": threads, "kind": "port",
                             "single_thread_rows_per_s": rows / single, "sample": f"the same {rows} rows, best of 3"}}
    print(json.dumps(line))


if __name__ == "__main__":
    main()
