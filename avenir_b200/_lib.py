"""ctypes binding of libavenir_tree.so (include/avenir_tree.h).  Fails loudly when the library is
missing or cannot be loaded: there is no Python/CPU fallback for any entry point."""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libavenir_tree.so")

AVT_NODE_NONE = 0xFFFF
AVT_MAX_CLASSES = 16
AVT_MAX_SEG = 8
CRIT_GINI, CRIT_ENTROPY, CRIT_HELLINGER, CRIT_CLASS_CONF = 0, 1, 2, 3
SAMPLE_WITH_REPLACE, SAMPLE_WITHOUT_REPLACE = 1, 2
ATTRS_ALL, ATTRS_NOT_USED_YET = 0, 1
SPLIT_NUMERIC, SPLIT_LUT, SPLIT_RANGES = 0, 1, 2


class Split(C.Structure):
    _fields_ = [("attr", C.c_int32), ("kind", C.c_int32), ("n_seg", C.c_int32), ("off", C.c_int32)]


class CsvAttr(C.Structure):
    _fields_ = [("ordinal", C.c_int32), ("kind", C.c_int32), ("n", C.c_int32), ("off", C.c_int32)]


CSV_INT, CSV_DOUBLE, CSV_CATEGORICAL = 0, 1, 2
CSV_BAD_SHORT_ROW, CSV_BAD_INT, CSV_BAD_DOUBLE, CSV_BAD_CATEGORY, CSV_BAD_CLASS = 1, 2, 3, 4, 5


class NodeBest(C.Structure):
    _fields_ = [("min_av", C.c_double), ("node_stat", C.c_double), ("population", C.c_int64),
                ("best_split", C.c_int32), ("n_ties", C.c_int32)]


class SplitEval(C.Structure):
    _fields_ = [("av", C.c_double), ("seg_stat", C.c_double * AVT_MAX_SEG),
                ("seg_count", (C.c_uint32 * AVT_MAX_CLASSES) * AVT_MAX_SEG)]


class CpgStat(C.Structure):
    _fields_ = [("stat", C.c_double), ("split_info", C.c_double),
                ("seg_count", (C.c_uint32 * AVT_MAX_CLASSES) * AVT_MAX_SEG)]


class GrowRecord(C.Structure):
    _fields_ = [("level", C.c_int32), ("tree", C.c_int32), ("node", C.c_int32), ("split", C.c_int32), ("n_ties", C.c_int32),
                ("child", C.c_int32 * AVT_MAX_SEG), ("child_pop", C.c_int64 * AVT_MAX_SEG), ("child_stat", C.c_double * AVT_MAX_SEG),
                ("min_av", C.c_double)]


class Stats(C.Structure):
    _fields_ = [("hist_ms", C.c_double), ("scan_ms", C.c_double), ("eval_ms", C.c_double), ("repart_ms", C.c_double),
                ("comm_ms", C.c_double), ("h2d_ms", C.c_double),
                ("hist_launches", C.c_int64), ("scan_launches", C.c_int64), ("eval_launches", C.c_int64),
                ("repart_launches", C.c_int64), ("comm_calls", C.c_int64),
                ("hist_alg_bytes", C.c_int64), ("repart_alg_bytes", C.c_int64),
                ("last_hist_variant", C.c_int32), ("last_hist_replicas", C.c_int32),
                ("last_hist_grid", C.c_int32), ("last_hist_block", C.c_int32), ("last_hist_smem", C.c_int32),
                ("reserved0", C.c_int32), ("last_csv_h2d_ms", C.c_double), ("last_csv_index_ms", C.c_double),
                ("last_csv_encode_ms", C.c_double), ("last_grow_ms", C.c_double), ("sample_ms", C.c_double)]


# every symbol include/avenir_tree.h declares: name -> (restype, argtypes)
_P = C.c_void_p
SIGNATURES = {
    "avt_create": (C.c_int, [C.c_int, C.POINTER(_P)]),
    "avt_destroy": (None, [_P]),
    "avt_last_error": (C.c_char_p, [_P]),
    "avt_dataset_load": (C.c_int, [_P, C.c_int64, C.c_int32, C.POINTER(C.c_int32), C.POINTER(_P), _P, C.c_int32]),
    "avt_dataset_load_csv": (C.c_int, [_P, _P, C.c_int64, C.c_char, C.c_int32, C.c_int32, C.POINTER(CsvAttr), _P, _P, _P, _P,
                                      C.c_int32, C.c_int32, _P, _P, C.POINTER(C.c_int64), _P, C.POINTER(C.c_int64), C.POINTER(C.c_int32)]),
    "avt_dataset_fetch_column": (C.c_int, [_P, C.c_int32, _P]),
    "avt_rows_partition": (C.c_int, [_P, C.c_int32, _P, C.c_int32, _P, C.c_int64, _P, _P]),
    "avt_forest_init": (C.c_int, [_P, C.c_int32, C.POINTER(_P), C.POINTER(_P)]),
    "avt_forest_init_sampled": (C.c_int, [_P, C.c_int32, C.c_int32, C.c_int64, C.c_int32, C.c_int64, C.c_int64]),
    "avt_forest_weights": (C.c_int, [_P, C.c_int32, _P]),
    "avt_forest_reset": (C.c_int, [_P]),
    "avt_split_table_set": (C.c_int, [_P, C.c_int32, C.POINTER(Split), C.POINTER(C.c_int32), C.c_int32, _P, C.c_int64]),
    "avt_level_histogram": (C.c_int, [_P, C.POINTER(C.c_int32), _P]),
    "avt_level_scan": (C.c_int, [_P, C.c_int32, _P, _P, C.POINTER(NodeBest)]),
    "avt_level_fetch_av": (C.c_int, [_P, C.c_int32, _P]),
    "avt_level_fetch_hist": (C.c_int, [_P, C.c_int32, _P]),
    "avt_level_eval": (C.c_int, [_P, C.c_int32, _P, C.c_int32, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(SplitEval)]),
    "avt_level_repartition": (C.c_int, [_P, C.POINTER(C.c_int32), _P, C.c_int32]),
    "avt_grow_levels": (C.c_int, [_P, C.c_int32, C.c_int32, _P, C.POINTER(C.c_int32), C.POINTER(GrowRecord), C.c_int64, C.POINTER(C.c_int64)]),
    "avt_grow_levels_ex": (C.c_int, [_P, C.c_int32, C.c_int32, _P, C.c_int32, _P, C.POINTER(C.c_int32), C.POINTER(GrowRecord), C.c_int64,
                                     C.POINTER(C.c_int64)]),
    "avt_cpg_sweep": (C.c_int, [_P, C.c_int32, C.c_int32, _P, C.POINTER(CpgStat)]),
    "avt_predict": (C.c_int, [_P, C.c_int32, C.POINTER(C.c_int32), _P, _P, _P, C.c_int32, _P, _P]),
    "avt_forest_node_ids": (C.c_int, [_P, C.c_int32, _P]),
    "avt_level_class_counts": (C.c_int, [_P, _P]),
    "avt_timer_begin": (C.c_int, [_P]),
    "avt_timer_end": (C.c_int, [_P, C.POINTER(C.c_float)]),
    "avt_synchronize": (C.c_int, [_P]),
    "avt_get_stats": (C.c_int, [_P, C.POINTER(Stats)]),
    "avt_reset_stats": (C.c_int, [_P]),
    "avt_set_profiling": (C.c_int, [_P, C.c_int32]),
    "avt_set_hist_variant": (C.c_int, [_P, C.c_int32]),
    "avt_comm_unique_id": (C.c_int, [_P]),
    "avt_comm_init_rank": (C.c_int, [_P, C.c_int32, C.c_int32, _P]),
    "avt_peer_export": (C.c_int, [_P, _P]),
    "avt_peer_import": (C.c_int, [_P, C.c_int32, C.c_int32, _P]),
    "avt_histogram_static": (C.c_int, [C.c_int, C.c_int64, C.c_int32, C.POINTER(C.c_int32), C.POINTER(_P), _P, C.c_int32, _P, C.c_int32, _P]),
    "avt_tool_run": (C.c_int, [C.c_char_p, C.c_int, C.POINTER(C.c_char_p), C.c_char_p, C.c_size_t]),
}

_lib = None


def load(path: str | None = None) -> C.CDLL:
    """Loads the shared library and binds every declared symbol (raises if one is missing)."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or LIB_PATH
    if not os.path.exists(p):
        raise RuntimeError(f"{p} not found: build it with `python -m avenir_b200.build` (needs nvcc); "
                           "avenir_b200 has no CPU fallback")
    lib = C.CDLL(p)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the export is missing
        fn.restype = res
        fn.argtypes = args
    if path is None:
        _lib = lib
    return lib
