"""Thin numpy-facing wrapper over the kernel-level C-ABI (include/avenir_tree.h) plus the level loop
used by bench.py and the GPU parity tests.  All compute happens in libavenir_tree.so on the GPU;
nothing here has a CPU fallback."""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence

import numpy as np

from . import _lib
from ._lib import (AVT_MAX_CLASSES, AVT_MAX_SEG, AVT_NODE_NONE, CRIT_ENTROPY, CRIT_GINI, SPLIT_LUT, SPLIT_NUMERIC, SPLIT_RANGES,
                   CpgStat, NodeBest, Split, SplitEval, Stats)


class EngineError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"[{code}] {msg}")
        self.code = code


class CsvRowError(EngineError):
    """avt_dataset_load_csv met a row the reference would throw on (reason = _lib.CSV_BAD_*)."""

    def __init__(self, row: int, reason: int, msg: str):
        super().__init__(-7, msg)
        self.row, self.reason = row, reason


def _ptr(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class SplitTable:
    """Candidate splits in the C-ABI's flat form.  Splits are grouped by non-decreasing attribute."""

    def __init__(self):
        self.attr: List[int] = []
        self.kind: List[int] = []
        self.n_seg: List[int] = []
        self.off: List[int] = []
        self.bounds: List[int] = []
        self.lut: List[int] = []

    def add_numeric(self, attr: int, upper_bins: Sequence[int]):
        self.attr.append(attr); self.kind.append(SPLIT_NUMERIC); self.n_seg.append(len(upper_bins) + 1)
        self.off.append(len(self.bounds)); self.bounds.extend(int(b) for b in upper_bins)

    def add_lut(self, attr: int, seg_of_bin: Sequence[int], n_seg: Optional[int] = None):
        self.attr.append(attr); self.kind.append(SPLIT_LUT)
        self.n_seg.append(max(2, (max(seg_of_bin) + 1) if n_seg is None else n_seg))
        self.off.append(len(self.lut)); self.lut.extend(int(s) for s in seg_of_bin)

    def add_ranges(self, attr: int, ranges: Sequence[Sequence[int]]):
        """Segments as inclusive bin ranges (lo, hi) that may overlap (the reference's multi-way numeric predicates)."""
        self.attr.append(attr); self.kind.append(SPLIT_RANGES); self.n_seg.append(len(ranges))
        self.off.append(len(self.bounds))
        for lo, hi in ranges:
            self.bounds.extend([int(lo), int(hi)])

    def __len__(self):
        return len(self.attr)

    def segment_of_bin(self, s: int, b: int) -> int:
        if self.kind[s] == SPLIT_NUMERIC:
            j = 0
            while j < self.n_seg[s] - 1 and b > self.bounds[self.off[s] + j]:
                j += 1
            return j
        return self.lut[self.off[s] + b]


def categorical_binary_partitions(n_levels: int) -> List[List[int]]:
    """2-group set partitions of levels 0..n-1 in the reference's enumeration order
    (tree/SplitManager.java:452-547 with numGroups = 2), as segment-of-level tables."""
    if n_levels < 2:
        return []
    splits: List[List[List[int]]] = [[[0], [1]]]
    if n_levels > 2:
        splits.append([[0, 1]])
        for idx in range(2, n_levels):
            new = []
            for sp in splits:
                if len(sp) == 2:
                    new.append([sp[0] + [idx], list(sp[1])])
                    new.append([list(sp[0]), sp[1] + [idx]])
                else:
                    new.append([list(sp[0]), [idx]])
            if idx < n_levels - 1:
                new.append([list(range(idx + 1))])
            splits = new
    out = []
    for sp in splits:
        lut = [0] * n_levels
        for g, grp in enumerate(sp):
            for lv in grp:
                lut[lv] = g
        out.append(lut)
    return out


def binary_split_table(n_bins: Sequence[int], categorical: Sequence[bool]) -> SplitTable:
    """DecisionTreeBuilder's binary candidate splits: every grid threshold of a numeric attribute,
    every 2-group partition of a categorical one (maxSplit = 2)."""
    t = SplitTable()
    for a, (b, cat) in enumerate(zip(n_bins, categorical)):
        if cat:
            for lut in categorical_binary_partitions(b):
                t.add_lut(a, lut)
        else:
            for k in range(b - 1):
                t.add_numeric(a, [k])
    return t


class Engine:
    def __init__(self, device: int = 0):
        self.lib = _lib.load()
        h = C.c_void_p()
        rc = self.lib.avt_create(device, C.byref(h))
        if rc != 0:
            raise EngineError(rc, self.lib.avt_last_error(None).decode())
        self.h = h
        self.n_rows = 0
        self.n_bins: List[int] = []
        self.n_classes = 0
        self.n_trees = 0
        self.n_splits = 0
        self.nf = 0
        self._keep = []

    def close(self):
        if getattr(self, "h", None):
            self.lib.avt_destroy(self.h)
            self.h = None

    __del__ = close

    def _check(self, rc: int):
        if rc != 0:
            raise EngineError(rc, self.lib.avt_last_error(self.h).decode())

    # ---- data
    def load_dataset(self, cols: Sequence[np.ndarray], cls: np.ndarray, n_bins: Sequence[int], n_classes: int):
        n = len(cls)
        fixed = []
        for a, c in enumerate(cols):
            want = np.uint16 if n_bins[a] > 256 else np.uint8
            if c.dtype != want or not c.flags.c_contiguous:
                c = np.ascontiguousarray(c, dtype=want)
            if len(c) != n:
                raise ValueError("column length mismatch")
            fixed.append(c)
        cls = np.ascontiguousarray(cls, dtype=np.uint8)
        nb = (C.c_int32 * len(n_bins))(*[int(b) for b in n_bins])
        ptrs = (C.c_void_p * len(fixed))(*[c.ctypes.data for c in fixed])
        self._check(self.lib.avt_dataset_load(self.h, n, len(fixed), nb, ptrs, _ptr(cls), int(n_classes)))
        self.n_rows, self.n_bins, self.n_classes = n, [int(b) for b in n_bins], int(n_classes)
        self.total_bins = sum(self.n_bins)

    def load_dataset_device(self, col_ptrs: Sequence[int], cls_ptr: int, n_rows: int, n_bins: Sequence[int], n_classes: int):
        """avt_dataset_load from buffers that already sit in this engine's device memory (raw device addresses, e.g.
        torch_tensor.data_ptr(); uint8 / uint16 codes as in load_dataset).  The library copies them."""
        nb = (C.c_int32 * len(n_bins))(*[int(b) for b in n_bins])
        ptrs = (C.c_void_p * len(col_ptrs))(*[int(p) for p in col_ptrs])
        self._check(self.lib.avt_dataset_load(self.h, int(n_rows), len(col_ptrs), nb, ptrs, C.c_void_p(int(cls_ptr)), int(n_classes)))
        self.n_rows, self.n_bins, self.n_classes = int(n_rows), [int(b) for b in n_bins], int(n_classes)
        self.total_bins = sum(self.n_bins)

    def load_csv(self, text: bytes, delim: str, attrs: Sequence[tuple], class_ordinal: int, class_values: Sequence[str],
                 skip_fields: int = 0):
        """avt_dataset_load_csv: CSV text -> data set, encoded on the device.  `attrs`: (ordinal, "int", [points]) |
        (ordinal, "double", [points]) | (ordinal, "categorical", [values]), ordinals increasing.  Returns
        (n_rows, first row of every class value or -1).  A row the reference would throw on raises CsvRowError.
        class_ordinal = -1 (rows without a class field): class_values is ignored, every row gets class code 0."""
        if class_ordinal < 0:
            class_values = [""]
        spec = (_lib.CsvAttr * len(attrs))()
        ints, dbls, cat_bytes, cat_off, n_bins = [], [], b"", [], []
        for i, (ordinal, kind, vals) in enumerate(attrs):
            if kind == "int":
                spec[i] = _lib.CsvAttr(ordinal, _lib.CSV_INT, len(vals), len(ints))
                ints.extend(int(v) for v in vals)
                n_bins.append(len(vals) + 1)
            elif kind == "double":
                spec[i] = _lib.CsvAttr(ordinal, _lib.CSV_DOUBLE, len(vals), len(dbls))
                dbls.extend(float(v) for v in vals)
                n_bins.append(len(vals) + 1)
            else:
                spec[i] = _lib.CsvAttr(ordinal, _lib.CSV_CATEGORICAL, len(vals), len(cat_off))
                for v in vals:
                    cat_off.append(len(cat_bytes))
                    cat_bytes += v.encode()
                n_bins.append(len(vals))
        cat_off.append(len(cat_bytes))
        cls_bytes, cls_off = b"", []
        for v in class_values:
            cls_off.append(len(cls_bytes))
            cls_bytes += v.encode()
        cls_off.append(len(cls_bytes))
        ip = np.asarray(ints if ints else [0], dtype=np.int32)
        dp = np.asarray(dbls if dbls else [0.0], dtype=np.float64)
        co = np.asarray(cat_off, dtype=np.int32)
        ko = np.asarray(cls_off, dtype=np.int32)
        first = np.zeros(len(class_values), dtype=np.int64)
        n_rows, bad_row, bad_reason = C.c_int64(0), C.c_int64(-1), C.c_int32(0)
        rc = self.lib.avt_dataset_load_csv(self.h, text, len(text), delim.encode(), skip_fields, len(attrs), spec, _ptr(ip), _ptr(dp),
                                           cat_bytes if cat_bytes else b"\0", _ptr(co), class_ordinal, len(class_values), cls_bytes, _ptr(ko),
                                           C.byref(n_rows), _ptr(first), C.byref(bad_row), C.byref(bad_reason))
        if rc != 0 and bad_row.value >= 0:
            raise CsvRowError(int(bad_row.value), int(bad_reason.value), self.lib.avt_last_error(self.h).decode())
        self._check(rc)
        self.n_rows, self.n_bins, self.n_classes = int(n_rows.value), n_bins, len(class_values)
        self.total_bins = sum(n_bins)
        return self.n_rows, [int(x) for x in first]

    def rows_partition(self, attr: int, segment_of_bin: Sequence[int], n_segments: int):
        """avt_rows_partition: the rows of the text last given to load_csv, scattered on the device into `n_segments` byte
        strings by segment_of_bin[bin code of `attr`] (AVT_NODE_NONE drops the row); returns ([bytes per segment], [rows])."""
        lut = np.asarray(segment_of_bin, dtype=np.uint16)
        seg_bytes = np.zeros(n_segments, dtype=np.int64)
        seg_rows = np.zeros(n_segments, dtype=np.int64)
        self._check(self.lib.avt_rows_partition(self.h, attr, _ptr(lut), n_segments, None, 0, _ptr(seg_bytes), _ptr(seg_rows)))
        total = int(seg_bytes.sum())
        out = np.empty(max(total, 1), dtype=np.uint8)
        self._check(self.lib.avt_rows_partition(self.h, attr, _ptr(lut), n_segments, _ptr(out), total, _ptr(seg_bytes), _ptr(seg_rows)))
        parts, at = [], 0
        for b in seg_bytes:
            parts.append(out[at:at + int(b)].tobytes())
            at += int(b)
        return parts, [int(r) for r in seg_rows]

    def fetch_column(self, attr: int) -> np.ndarray:
        """The installed data set's codes of one attribute (-1: class codes)."""
        dt = np.uint8 if attr < 0 or self.n_bins[attr] <= 256 else np.uint16
        out = np.empty(self.n_rows, dtype=dt)
        self._check(self.lib.avt_dataset_fetch_column(self.h, attr, _ptr(out)))
        return out

    def forest_init(self, n_trees: int = 1, weights: Optional[Sequence[Optional[np.ndarray]]] = None,
                    node_ids: Optional[Sequence[Optional[np.ndarray]]] = None):
        def arr(seq, dt):
            if seq is None:
                return None, []
            keep = [None if x is None else np.ascontiguousarray(x, dtype=dt) for x in seq]
            return (C.c_void_p * n_trees)(*[None if x is None else x.ctypes.data for x in keep]), keep
        wp, k1 = arr(weights, np.uint8)
        npp, k2 = arr(node_ids, np.uint16)
        self._check(self.lib.avt_forest_init(self.h, n_trees, wp, npp))
        self.n_trees = n_trees

    def forest_init_sampled(self, n_trees: int, mode: int, seed: int, param: int, total_rows: Optional[int] = None, row0: int = 0):
        """Trees whose first-pass sub-sampling is generated on the device (avt_forest_init_sampled): mode
        _lib.SAMPLE_WITH_REPLACE (param = bas.batch.size) or _lib.SAMPLE_WITHOUT_REPLACE (param = rate in percent)."""
        total = self.n_rows if total_rows is None else int(total_rows)
        self._check(self.lib.avt_forest_init_sampled(self.h, n_trees, mode, seed, param, total, row0))
        self.n_trees = n_trees

    def forest_reset(self):
        self._check(self.lib.avt_forest_reset(self.h))

    def weights(self, tree: int) -> np.ndarray:
        out = np.empty(self.n_rows, dtype=np.uint8)
        self._check(self.lib.avt_forest_weights(self.h, tree, _ptr(out)))
        return out

    def set_splits(self, t: SplitTable):
        n = len(t)
        arr = (Split * max(n, 1))()
        for i in range(n):
            arr[i] = Split(t.attr[i], t.kind[i], t.n_seg[i], t.off[i])
        b = np.asarray(t.bounds if t.bounds else [0], dtype=np.int32)
        l = np.asarray(t.lut if len(t.lut) else [0], dtype=np.uint8)
        self._check(self.lib.avt_split_table_set(self.h, n, arr, b.ctypes.data_as(C.POINTER(C.c_int32)), len(t.bounds), _ptr(l), len(t.lut)))
        self.n_splits = n

    # ---- level
    def histogram(self, nodes_per_tree: Sequence[int], fetch: bool = False) -> Optional[np.ndarray]:
        npt = (C.c_int32 * len(nodes_per_tree))(*[int(x) for x in nodes_per_tree])
        self.nf = int(sum(nodes_per_tree))
        out = np.empty((self.nf, self.total_bins, self.n_classes), dtype=np.uint32) if fetch else None
        self._check(self.lib.avt_level_histogram(self.h, npt, _ptr(out)))
        return out

    def scan(self, criterion: int, cand_mask: Optional[np.ndarray] = None, class_order: Optional[Sequence[int]] = None):
        out = (NodeBest * self.nf)()
        m = None if cand_mask is None else np.ascontiguousarray(cand_mask, dtype=np.uint32)
        o = None if class_order is None else np.ascontiguousarray(class_order, dtype=np.uint8)
        self._check(self.lib.avt_level_scan(self.h, criterion, _ptr(m), _ptr(o), out))
        return out

    def fetch_av(self, node: int) -> np.ndarray:
        out = np.empty(self.n_splits, dtype=np.float64)
        self._check(self.lib.avt_level_fetch_av(self.h, node, _ptr(out)))
        return out

    def fetch_hist(self, node: int) -> np.ndarray:
        out = np.empty((self.total_bins, self.n_classes), dtype=np.uint32)
        self._check(self.lib.avt_level_fetch_hist(self.h, node, _ptr(out)))
        return out

    def eval(self, criterion: int, nodes: Sequence[int], splits: Sequence[int], class_order: Optional[Sequence[int]] = None):
        n = len(nodes)
        out = (SplitEval * max(n, 1))()
        na = (C.c_int32 * max(n, 1))(*[int(x) for x in nodes])
        sa = (C.c_int32 * max(n, 1))(*[int(x) for x in splits])
        o = None if class_order is None else np.ascontiguousarray(class_order, dtype=np.uint8)
        self._check(self.lib.avt_level_eval(self.h, criterion, _ptr(o), n, na, sa, out))
        return out

    def repartition(self, node_attr: Sequence[int], child_lut: np.ndarray):
        na = (C.c_int32 * len(node_attr))(*[int(x) for x in node_attr])
        lut = np.ascontiguousarray(child_lut, dtype=np.uint16)
        self._check(self.lib.avt_level_repartition(self.h, na, _ptr(lut), lut.shape[1]))

    def grow(self, n_levels: int, criterion: int, nodes_per_tree: Sequence[int], class_order: Optional[Sequence[int]] = None,
             max_records: int = 0, strategy: int = _lib.ATTRS_ALL, root_cand_mask: Optional[np.ndarray] = None):
        """avt_grow_levels_ex: the level loop inside the library, candidate attributes per `strategy`
        (_lib.ATTRS_ALL | _lib.ATTRS_NOT_USED_YET).  Returns (records, final nodes_per_tree)."""
        npt = (C.c_int32 * len(nodes_per_tree))(*[int(x) for x in nodes_per_tree])
        rec = (_lib.GrowRecord * max_records)() if max_records else None
        n = C.c_int64(0)
        o = None if class_order is None else np.ascontiguousarray(class_order, dtype=np.uint8)
        m = None if root_cand_mask is None else np.ascontiguousarray(root_cand_mask, dtype=np.uint32)
        self._check(self.lib.avt_grow_levels_ex(self.h, n_levels, criterion, _ptr(o), strategy, _ptr(m), npt, rec, max_records, C.byref(n)))
        return (rec, int(n.value)), list(npt)

    def cpg_sweep(self, node: int, criterion: int, class_order: Optional[Sequence[int]] = None, out=None):
        """avt_cpg_sweep; `out` = a (CpgStat * n_splits) array to reuse (e.g. CpgStat-array.from_address of page-locked memory:
        a fresh 528-byte-per-split array costs more to allocate and fault in than the sweep takes)."""
        if out is None:
            out = (CpgStat * max(self.n_splits, 1))()
        o = None if class_order is None else np.ascontiguousarray(class_order, dtype=np.uint8)
        self._check(self.lib.avt_cpg_sweep(self.h, node, criterion, _ptr(o), out))
        return out

    def node_ids(self, tree: int = 0, out: Optional[np.ndarray] = None) -> np.ndarray:
        if out is None:
            out = np.empty(self.n_rows, dtype=np.uint16)
        self._check(self.lib.avt_forest_node_ids(self.h, tree, _ptr(out)))
        return out

    def class_counts(self) -> np.ndarray:
        out = np.empty((self.nf, self.n_classes), dtype=np.uint32)
        self._check(self.lib.avt_level_class_counts(self.h, _ptr(out)))
        return out

    # ---- misc
    def timer_begin(self):
        self._check(self.lib.avt_timer_begin(self.h))

    def timer_end(self) -> float:
        ms = C.c_float()
        self._check(self.lib.avt_timer_end(self.h, C.byref(ms)))
        return float(ms.value)

    def synchronize(self):
        self._check(self.lib.avt_synchronize(self.h))

    def stats(self) -> Stats:
        s = Stats()
        self._check(self.lib.avt_get_stats(self.h, C.byref(s)))
        return s

    def reset_stats(self):
        self._check(self.lib.avt_reset_stats(self.h))

    def set_profiling(self, on: bool):
        self._check(self.lib.avt_set_profiling(self.h, 1 if on else 0))

    def set_hist_variant(self, v: int):
        self._check(self.lib.avt_set_hist_variant(self.h, v))

    def peer_export(self) -> bytes:
        buf = C.create_string_buffer(64)
        self._check(self.lib.avt_peer_export(self.h, buf))
        return buf.raw

    def peer_import(self, n_ranks: int, rank: int, handles: Sequence[bytes]):
        blob = C.create_string_buffer(b"".join(handles), 64 * n_ranks)
        self._check(self.lib.avt_peer_import(self.h, n_ranks, rank, blob))

    def comm_init(self, n_ranks: int, rank: int, unique_id: bytes):
        buf = C.create_string_buffer(unique_id, 128)
        self._check(self.lib.avt_comm_init_rank(self.h, n_ranks, rank, buf))


def comm_unique_id() -> bytes:
    lib = _lib.load()
    buf = C.create_string_buffer(128)
    rc = lib.avt_comm_unique_id(buf)
    if rc != 0:
        raise EngineError(rc, lib.avt_last_error(None).decode())
    return buf.raw


def grow_levels(eng, table: SplitTable, n_levels: int, criterion: int, n_trees: int = 1,
                class_order: Optional[Sequence[int]] = None, record: bool = True, strategy: int = _lib.ATTRS_ALL,
                root_cand_mask: Optional[np.ndarray] = None):
    """The per-level driver loop (what resource/detr.sh:34-54 does by hand, one Hadoop job per level):
    histogram -> scan -> children statistics -> repartition, `n_levels` times, for all trees at once.
    Ties resolve to the lowest table index (the tool layer, not this loop, replays the reference's
    HashMap order).  `eng` only needs the Engine methods used here.  Returns per level a list of
    (tree, node, split, [child populations], min_av, n_ties)."""
    lut_w = max(eng.n_bins)
    frontier = [1] * n_trees
    history = []
    n_attrs = len(eng.n_bins)
    mw = (n_attrs + 31) // 32
    masks = None   # candidate attribute sets of the frontier nodes (getSplitAttributes, tree/DecisionTreeBuilder.java:365-381)
    if strategy != _lib.ATTRS_ALL or root_cand_mask is not None:
        full = np.array([(1 << min(32, n_attrs - 32 * w)) - 1 for w in range(mw)], dtype=np.uint32)
        masks = np.tile(full, (n_trees, 1)) if root_cand_mask is None else (np.asarray(root_cand_mask, dtype=np.uint32).reshape(n_trees, mw) & full)
    for _ in range(n_levels):
        nf = sum(frontier)
        if nf == 0:
            break
        eng.histogram(frontier)
        best = eng.scan(criterion, masks, class_order)
        next_masks = []
        nodes = [g for g in range(nf) if best[g].population > 0 and best[g].best_split >= 0]
        splits = [best[g].best_split for g in nodes]
        evals = eng.eval(criterion, nodes, splits, class_order)
        node_attr = [-1] * nf
        child_lut = np.full((nf, lut_w), AVT_NODE_NONE, dtype=np.uint16)
        new_frontier = [0] * n_trees
        level = []
        base = np.concatenate([[0], np.cumsum(frontier)])
        k = 0
        for t in range(n_trees):
            nxt = 0
            for g in range(base[t], base[t + 1]):
                if k >= len(nodes) or nodes[k] != g:
                    continue
                s = splits[k]
                ev = evals[k]
                k += 1
                a = table.attr[s]
                node_attr[g] = a
                child_of_seg = {}
                pops = []
                for sg in range(table.n_seg[s]):
                    n = sum(ev.seg_count[sg][c] for c in range(eng.n_classes))
                    if n:
                        child_of_seg[sg] = nxt
                        nxt += 1
                        pops.append(n)
                        if masks is not None:
                            m = masks[g].copy()
                            if strategy == _lib.ATTRS_NOT_USED_YET:
                                m[a >> 5] &= np.uint32(~(1 << (a & 31)) & 0xFFFFFFFF)
                            next_masks.append(m)
                for b in range(eng.n_bins[a]):
                    sg = table.segment_of_bin(s, b)
                    if sg in child_of_seg:
                        child_lut[g, b] = child_of_seg[sg]
                if record:
                    level.append((t, g - base[t], s, pops, best[g].min_av, best[g].n_ties))
            new_frontier[t] = nxt
        eng.repartition(node_attr, child_lut)
        if masks is not None:
            masks = np.array(next_masks, dtype=np.uint32).reshape(len(next_masks), mw)
        frontier = new_frontier
        history.append(level)
    return history, frontier
