"""Test infrastructure: loaders for the C oracle (oracle/hist_oracle.c), the CPU link-time mock of the
kernel layer (tests/mock/mock_backend.cpp) and an oracle-backed stand-in for avenir_b200.engine.Engine."""
from __future__ import annotations

import ctypes as C
import json
import os
import subprocess
from typing import Optional, Sequence

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BUILD = os.path.join(ROOT, "tests", "_build")
SCHEMA_DIR = os.path.join(ROOT, "tests", "golden")

from avenir_b200._lib import CpgStat, NodeBest, Split, SplitEval  # struct layouts only  # noqa: E402


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.exists(d) and os.path.getmtime(d) > t for d in deps)


_oracle = None


def oracle_lib() -> C.CDLL:
    global _oracle
    if _oracle is None:
        so = os.path.join(ROOT, "oracle", "libhist_oracle.so")
        if _stale(so, [os.path.join(ROOT, "oracle", "hist_oracle.c")]):
            subprocess.run(["make", "-C", os.path.join(ROOT, "oracle")], check=True, capture_output=True)
        lib = C.CDLL(so)
        lib.ora_strict_log.restype = C.c_double
        lib.ora_strict_log.argtypes = [C.c_double]
        lib.ora_impurity.restype = C.c_double
        lib.ora_impurity.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
        _oracle = lib
    return _oracle


_mock = None


def hostmock_lib() -> C.CDLL:
    """Product host layer (avenir_b200/csrc/host) linked against the CPU mock of the kernel layer."""
    global _mock
    if _mock is None:
        os.makedirs(BUILD, exist_ok=True)
        so = os.path.join(BUILD, "libavt_hostmock.so")
        host = [os.path.join(ROOT, "avenir_b200", "csrc", "host", f) for f in ("dtb.cpp", "cpg.cpp", "mop.cpp", "crc.cpp", "dap.cpp", "cce.cpp", "cca.cpp", "tool.cpp")]
        hdrs = [os.path.join(ROOT, "avenir_b200", "csrc", "host", f) for f in ("model.hpp", "javaemu.hpp", "json.hpp", "tool.hpp")]
        mock = os.path.join(ROOT, "tests", "mock", "mock_backend.cpp")
        ora = os.path.join(ROOT, "oracle", "hist_oracle.c")
        if _stale(so, host + hdrs + [mock, ora, os.path.join(ROOT, "include", "avenir_tree.h")]):
            obj = os.path.join(BUILD, "hist_oracle.o")
            subprocess.run(["gcc", "-O2", "-ffp-contract=off", "-fPIC", "-c", ora, "-o", obj], check=True)
            subprocess.run(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-fPIC", "-w", "-shared", "-Wl,-Bsymbolic", "-o", so] + host + [mock, obj, "-lpthread"], check=True)
        lib = C.CDLL(so)
        lib.avt_tool_run.restype = C.c_int
        lib.avt_tool_run.argtypes = [C.c_char_p, C.c_int, C.POINTER(C.c_char_p), C.c_char_p, C.c_size_t]
        _mock = lib
    return _mock


def tool_run(lib, tool: str, args: Sequence[str]):
    arr = (C.c_char_p * len(args))(*[a.encode() for a in args])
    err = C.create_string_buffer(4096)
    rc = lib.avt_tool_run(tool.encode(), len(args), arr, err, len(err))
    return rc, err.value.decode(errors="replace")


def write_schema(path: str, name: str):
    """Writes one of the reference's shipped tree schemas (tests/golden/schemas.json) to `path`."""
    d = json.load(open(os.path.join(SCHEMA_DIR, "schemas.json")))[name]
    json.dump(d, open(path, "w"))
    return d


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class OracleEngine:
    """Same call surface as avenir_b200.engine.Engine, answered by the C oracle on the CPU."""

    def __init__(self, threads: int = 1):
        self.lib = oracle_lib()
        self.threads = threads

    def load_dataset(self, cols, cls, n_bins, n_classes):
        self.cols = [np.ascontiguousarray(c, dtype=np.uint16 if b > 256 else np.uint8) for c, b in zip(cols, n_bins)]
        self.cls = np.ascontiguousarray(cls, dtype=np.uint8)
        self.n_rows = len(self.cls)
        self.n_bins = [int(b) for b in n_bins]
        self.nb_arr = np.asarray(self.n_bins, dtype=np.int32)
        self.n_classes = int(n_classes)
        self.total_bins = sum(self.n_bins)
        self.colptrs = (C.c_void_p * len(self.cols))(*[c.ctypes.data for c in self.cols])

    def forest_init(self, n_trees=1, weights=None, node_ids=None):
        self.n_trees = n_trees
        self.nid = [np.zeros(self.n_rows, dtype=np.uint16) if node_ids is None or node_ids[t] is None
                    else np.array(node_ids[t], dtype=np.uint16) for t in range(n_trees)]
        self.w = [None if weights is None or weights[t] is None else np.ascontiguousarray(weights[t], dtype=np.uint8)
                  for t in range(n_trees)]

    def set_splits(self, t):
        self.table = t
        n = len(t)
        self.splits = (Split * max(n, 1))()
        for i in range(n):
            self.splits[i] = Split(t.attr[i], t.kind[i], t.n_seg[i], t.off[i])
        self.bounds = np.asarray(t.bounds if t.bounds else [0], dtype=np.int32)
        self.lut = np.asarray(t.lut if len(t.lut) else [0], dtype=np.uint8)
        self.n_splits = n

    def histogram(self, nodes_per_tree, fetch=False):
        self.npt = np.asarray(nodes_per_tree, dtype=np.int32)
        self.nf = int(self.npt.sum())
        self.H = np.zeros((self.nf, self.total_bins, self.n_classes), dtype=np.uint32)
        nidp = (C.c_void_p * self.n_trees)(*[x.ctypes.data for x in self.nid])
        wp = (C.c_void_p * self.n_trees)(*[None if x is None else x.ctypes.data for x in self.w])
        self.lib.ora_histogram(C.c_int64(self.n_rows), len(self.cols), _p(self.nb_arr), self.colptrs, _p(self.cls), self.n_classes,
                               self.n_trees, nidp, wp, _p(self.npt), _p(self.H), self.threads)
        return self.H.copy() if fetch else None

    def _order(self, class_order):
        return np.asarray(range(self.n_classes) if class_order is None else class_order, dtype=np.uint8)

    def scan(self, criterion, cand_mask=None, class_order=None):
        out = (NodeBest * self.nf)()
        self.av = np.zeros((self.nf, self.n_splits), dtype=np.float64)
        m = None if cand_mask is None else np.ascontiguousarray(cand_mask, dtype=np.uint32)
        o = self._order(class_order)
        self.lib.ora_scan(_p(self.H), self.nf, len(self.cols), _p(self.nb_arr), self.n_classes, self.n_splits, self.splits,
                          _p(self.bounds), _p(self.lut), _p(m), _p(o), int(criterion == 1), _p(self.av), out)
        return out

    def fetch_av(self, node):
        return self.av[node].copy()

    def fetch_hist(self, node):
        return self.H[node].copy()

    def eval(self, criterion, nodes, splits, class_order=None):
        n = len(nodes)
        out = (SplitEval * max(n, 1))()
        na = np.asarray(nodes if n else [0], dtype=np.int32)
        sa = np.asarray(splits if n else [0], dtype=np.int32)
        o = self._order(class_order)
        self.lib.ora_eval(_p(self.H), len(self.cols), _p(self.nb_arr), self.n_classes, self.splits, _p(self.bounds), _p(self.lut), _p(o),
                          int(criterion == 1), n, _p(na), _p(sa), out)
        return out

    def repartition(self, node_attr, child_lut):
        na = np.asarray(node_attr, dtype=np.int32)
        lut = np.ascontiguousarray(child_lut, dtype=np.uint16)
        nidp = (C.c_void_p * self.n_trees)(*[x.ctypes.data for x in self.nid])
        self.lib.ora_repartition(C.c_int64(self.n_rows), _p(self.nb_arr), self.colptrs, self.n_trees, nidp, _p(self.npt), _p(na), _p(lut),
                                 lut.shape[1])

    def cpg_sweep(self, node, criterion, class_order=None):
        out = (CpgStat * max(self.n_splits, 1))()
        o = self._order(class_order)
        self.lib.ora_cpg(_p(self.H), node, len(self.cols), _p(self.nb_arr), self.n_classes, self.n_splits, self.splits, _p(self.bounds),
                         _p(self.lut), _p(o), criterion, out)
        return out

    def node_ids(self, tree=0):
        return self.nid[tree].copy()

    def class_counts(self):
        return self.H[:, :self.n_bins[0], :].sum(axis=1).astype(np.uint32)


def oracle_bagging(seed: int, batch: int, total: int, row0: int, n: int) -> np.ndarray:
    """oracle/hist_oracle.c:ora_bagging_weights -- multiplicities of rows [row0, row0 + n) of a `total`-row file."""
    lib = oracle_lib()
    lib.ora_bagging_weights.restype = C.c_int
    lib.ora_bagging_weights.argtypes = [C.c_int64, C.c_int32, C.c_int64, C.c_int64, C.c_int64, C.c_void_p]
    out = np.empty(n, dtype=np.uint8)
    if lib.ora_bagging_weights(seed, batch, total, row0, n, out.ctypes.data_as(C.c_void_p)) != 0:
        raise ValueError("bootstrap multiplicity above 255")
    return out


def unordered_model(level_json: dict):
    """A DecisionPathList as an order-free value: the reference-held sample (resource/dec_tree_rules.json) lists its
    paths in an order the current code's HashMap logic does not reproduce, so list order is NOT pinned by anything the
    reference ships; paths of binary splits are disjoint, so prediction does not depend on it.  Model equality that
    a reader of the reference can rely on is therefore this comparison; the exact-order one is between our two restatements."""
    paths = sorted(json.dumps(p, sort_keys=True) for p in level_json.get("decisionPaths", []))
    rest = {k: v for k, v in level_json.items() if k != "decisionPaths"}
    return json.dumps(rest, sort_keys=True), paths


def bits(x: float) -> int:
    import struct
    return struct.unpack("<Q", struct.pack("<d", x))[0]


class _OraCsvAttr(C.Structure):
    _fields_ = [("ordinal", C.c_int32), ("kind", C.c_int32), ("n", C.c_int32), ("off", C.c_int32)]


def oracle_encode_csv(text: bytes, delim: str, attrs, class_ordinal: int, class_values, skip_fields: int = 0, threads: int = 1,
                      count_only: bool = False):
    """oracle/hist_oracle.c:ora_encode_csv.  Returns (columns as int32 arrays, class codes) or the first bad row (int)."""
    import numpy as np
    lib = oracle_lib()
    lib.ora_encode_csv.restype = C.c_int64
    spec = (_OraCsvAttr * len(attrs))()
    ints, dbls, cb, co = [], [], b"", []
    for i, (ordinal, kind, vals) in enumerate(attrs):
        if kind == "int":
            spec[i] = _OraCsvAttr(ordinal, 0, len(vals), len(ints)); ints.extend(int(v) for v in vals)
        elif kind == "double":
            spec[i] = _OraCsvAttr(ordinal, 1, len(vals), len(dbls)); dbls.extend(float(v) for v in vals)
        else:
            spec[i] = _OraCsvAttr(ordinal, 2, len(vals), len(co))
            for v in vals:
                co.append(len(cb)); cb += v.encode()
    co.append(len(cb))
    kb, ko = b"", []
    for v in class_values:
        ko.append(len(kb)); kb += v.encode()
    ko.append(len(kb))
    ip = np.asarray(ints or [0], dtype=np.int32); dp = np.asarray(dbls or [0.0], dtype=np.float64)
    coa = np.asarray(co, dtype=np.int32); koa = np.asarray(ko, dtype=np.int32)
    args = [C.c_char_p(text), C.c_int64(len(text)), C.c_char(delim.encode()), C.c_int(skip_fields), C.c_int(len(attrs)), spec,
            C.c_void_p(ip.ctypes.data), C.c_void_p(dp.ctypes.data), C.c_char_p(cb or b"\0"), C.c_void_p(coa.ctypes.data),
            C.c_int(class_ordinal), C.c_int(len(class_values)), C.c_char_p(kb), C.c_void_p(koa.ctypes.data)]
    n = lib.ora_encode_csv(*args, None, None, C.c_int(threads))
    if count_only:
        return int(n)
    cols = [np.zeros(n, dtype=np.int32) for _ in attrs]
    cls = np.zeros(n, dtype=np.uint8)
    ptrs = (C.c_void_p * len(cols))(*[c.ctypes.data for c in cols])
    rc = lib.ora_encode_csv(*args, ptrs, C.c_void_p(cls.ctypes.data), C.c_int(threads))
    if rc < 0:
        return int(-rc - 1)
    return cols, cls
