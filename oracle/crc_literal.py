"""
ORACLE / TEST INFRASTRUCTURE ONLY -- never imported by the product (avenir_b200/).

Literal restatement of org.avenir.explore.CramerCorrelation, row by row and matrix by matrix:
  explore/CramerCorrelation.java:104-141   mapper set-up: one ContingencyMatrix per (source, destination) pair with
                                           src != dst, pairs in source-major order; the destination field list is
                                           built while the FIRST source attribute is handled
  explore/CramerCorrelation.java:157-183   map: items = row.split(delim); for every source field, for every
                                           destination field: matrix of attrPairs[k++] gets increment(srcIndex, dstIndex)
  explore/CramerCorrelation.java:146-152   cleanup: every pair's matrix is emitted
  explore/CramerCorrelation.java:215-235   reduce (keys sorted): sum the matrices, `srcName,dstName,cramerIndex`
  util/ContingencyMatrix.java:87-124       cramerIndex
and of org.avenir.explore.HeterogeneityReductionCorrelation (the same mapper under `cac.*` keys,
explore/CategoricalCorrelation.java:66-150; concentration / uncertainty coefficient, util/ContingencyMatrix.java:142-187).
PARITY UNPINNED (no JVM, no reference tests for these jobs).  Math.log10 is restated with the platform's log10.  chombo's FeatureField.cardinalityIndex is taken to be
List.indexOf (-1 when absent => ArrayIndexOutOfBoundsException in TabularData.increment); Tuple keys are taken to
sort element by element.
"""
from __future__ import annotations

from typing import Dict, List, Sequence, Tuple

from .javaemu import double_to_string
from .literal import FeatureSchema, java_split


class ArrayIndexOutOfBounds(Exception):
    """The reference's mapper would die with an ArrayIndexOutOfBoundsException / IndexOutOfBoundsException here."""


def cramer_index(table: List[List[int]]) -> float:
    """ContingencyMatrix.cramerIndex (util/ContingencyMatrix.java:87-124)."""
    n_row, n_col = len(table), len(table[0])
    row_sum = []
    for i in range(n_row):
        s = 0
        for j in range(n_col):
            s += table[i][j]
        row_sum.append(1 if s == 0 else s)
    col_sum = []
    for j in range(n_col):
        s = 0
        for i in range(n_row):
            s += table[i][j]
        col_sum.append(1 if s == 0 else s)
    pearson = 0.0
    for i in range(n_row):
        for j in range(n_col):
            pearson += (float(table[i][j]) * table[i][j]) / (float(row_sum[i]) * col_sum[j])
    pearson -= 1.0
    smaller = n_row if n_row < n_col else n_col
    return pearson / (smaller - 1)


def _aggregates(table: List[List[int]]):
    """ContingencyMatrix.getAggregates + rowSumAsDouble / colSumAsDouble (util/ContingencyMatrix.java:62-85, 126-140)."""
    n_row, n_col = len(table), len(table[0])
    total = 0
    row_sum = []
    for i in range(n_row):
        s = 0
        for j in range(n_col):
            s += table[i][j]
            total += table[i][j]
        row_sum.append(1 if s == 0 else s)
    col_sum = []
    for j in range(n_col):
        s = 0
        for i in range(n_row):
            s += table[i][j]
        col_sum.append(1 if s == 0 else s)
    div = lambda a, b: a / b if b else (float("nan") if a == 0 else float("inf"))      # Java double division by zero
    return [div(float(r), total) for r in row_sum], [div(float(c), total) for c in col_sum], total


def concentration_coeff(table: List[List[int]]) -> float:
    """ContingencyMatrix.concentrationCoeff (util/ContingencyMatrix.java:142-164)."""
    row, col, total = _aggregates(table)
    sum_one = 0.0
    for i in range(len(table)):
        sq = 0.0
        for j in range(len(table[0])):
            elem = float(table[i][j]) / total if total else float("nan")
            sq += elem * elem
        sum_one += sq / row[i]
    sum_two = 0.0
    for j in range(len(table[0])):
        sum_two += col[j] * col[j]
    return (sum_one - sum_two) / (1.0 - sum_two)


def _log10(x: float) -> float:
    """Math.log10 (the platform's log10; 0 -> -Infinity, negative/NaN -> NaN)."""
    import math
    if x != x or x < 0:
        return float("nan")
    return float("-inf") if x == 0 else math.log10(x)


def uncertainty_coeff(table: List[List[int]]) -> float:
    """ContingencyMatrix.uncertaintyCoeff (util/ContingencyMatrix.java:166-187): an empty cell contributes
    0 * log10(0) = NaN, exactly as in the reference."""
    row, col, total = _aggregates(table)
    mul = lambda a, b: float("nan") if (a == 0 and b in (float("inf"), float("-inf"))) or (b == 0 and a in (float("inf"), float("-inf"))) else a * b
    sum_one = 0.0
    for i in range(len(table)):
        for j in range(len(table[0])):
            elem = float(table[i][j]) / total if total else float("nan")
            sum_one += mul(elem, _log10(elem * col[j] / row[i]))
    sum_two = 0.0
    for j in range(len(table[0])):
        sum_two += mul(col[j], _log10(col[j]))
    return sum_one / sum_two if sum_two else (float("nan") if sum_one == 0 or sum_one != sum_one else float("inf") * (1 if sum_one > 0 else -1))


def cramer_correlation(lines: Sequence[str], conf: Dict[str, str], schema: FeatureSchema) -> List[str]:
    """explore/CramerCorrelation.java with one mapper and one reducer: output lines of part-r-00000."""
    return correlation_job(lines, conf, schema, "crc.source.attributes", "crc.dest.attributes", cramer_index)


def heterogeneity_reduction_correlation(lines: Sequence[str], conf: Dict[str, str], schema: FeatureSchema) -> List[str]:
    """explore/HeterogeneityReductionCorrelation.java: CategoricalCorrelation's mapper (explore/CategoricalCorrelation.java:66-150,
    same bookkeeping, `cac.*` keys) + a reducer printing the concentration or the uncertainty coefficient (:71-88)."""
    gini = conf.get("hrc.heterogeneity.algorithm", "gini") == "gini"
    return correlation_job(lines, conf, schema, "cac.first.set.attributes", "cac.second.set..attributes", concentration_coeff if gini else uncertainty_coeff)


def correlation_job(lines: Sequence[str], conf: Dict[str, str], schema: FeatureSchema, source_key: str, dest_key: str, stat) -> List[str]:
    delim_regex = conf.get("field.delim.regex", ",")
    delim_out = conf.get("field.delim.out", ",")
    source = [int(x) for x in conf[source_key].split(",")]
    dest = [int(x) for x in conf[dest_key].split(",")]
    # ---- mapper set-up
    matrices: Dict[Tuple[int, int], List[List[int]]] = {}
    src_fields, dst_fields, pairs = [], [], []
    dst_initialized = False
    for s in source:
        sf = schema.find_field_by_ordinal(s)
        src_fields.append(sf)
        for d in dest:
            if s != d:
                df = schema.find_field_by_ordinal(d)
                matrices[(s, d)] = [[0] * len(df.cardinality) for _ in sf.cardinality]
                if not dst_initialized:
                    dst_fields.append(df)
                pairs.append((s, d))
        dst_initialized = True
    # ---- map
    for ln in lines:
        items = java_split(ln, delim_regex, False)
        k = 0
        for sf in src_fields:
            if sf.ordinal >= len(items):
                raise ArrayIndexOutOfBounds(ln)
            v = items[sf.ordinal]
            si = sf.cardinality.index(v) if v in sf.cardinality else -1
            for df in dst_fields:
                if df.ordinal >= len(items):
                    raise ArrayIndexOutOfBounds(ln)
                w = items[df.ordinal]
                di = df.cardinality.index(w) if w in df.cardinality else -1
                if k >= len(pairs):
                    raise ArrayIndexOutOfBounds("attrPairs.get(%d)" % k)
                m = matrices[pairs[k]]
                k += 1
                if si < 0 or di < 0 or si >= len(m) or di >= len(m[0]):
                    raise ArrayIndexOutOfBounds(ln)
                m[si][di] += 1
    # ---- reduce: keys in sorted order
    out = []
    for (s, d) in sorted(set(pairs)):
        sf, df = schema.find_field_by_ordinal(s), schema.find_field_by_ordinal(d)
        out.append(sf.name + delim_out + df.name + delim_out + double_to_string(stat(matrices[(s, d)])))
    return out
