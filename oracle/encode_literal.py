"""
ORACLE / TEST INFRASTRUCTURE ONLY -- never imported by the product (avenir_b200/).

Literal restatement of what the reference does to one text row before it can evaluate a predicate, in the
histogram form the engine uses ("on-device encode"):
  tree/DecisionTreeBuilder.java:268        items = value.toString().split(fieldDelimRegex)   (java_split: trailing
                                           empty strings are dropped)
  tree/DecisionTreeBuilder.java:309-320    per candidate attribute: Integer.parseInt / Double.parseDouble / the raw
                                           string, in attribute order; items[ordinal] beyond the row throws
  tree/SplitManager.java:256-330, 741-814  numeric predicates over the point grid  => bin = number of points < value
  tree/SplitManager.java:897-950           categorical `in` predicates             => bin = index in the cardinality list
  tree/DecisionTreeBuilder.java:749        class value = items[classOrdinal]
PARITY UNPINNED (no JVM here).  Double.parseDouble is restated with Python's float() on Java's decimal grammar
(both are correctly rounded); Java's type suffixes (1.0f, 2d) and hex literals are outside the restatement.
"""
from __future__ import annotations

import bisect
import re
from typing import List, Optional, Sequence, Tuple

from .literal import java_split

BAD_SHORT_ROW, BAD_INT, BAD_DOUBLE, BAD_CATEGORY, BAD_CLASS = 1, 2, 3, 4, 5

_INT = re.compile(r"^[+-]?[0-9]+$")
_DBL = re.compile(r"^[+-]?(NaN|Infinity|(([0-9]+\.?[0-9]*)|(\.[0-9]+))([eE][+-]?[0-9]+)?)$")


def parse_java_int(s: str) -> Optional[int]:
    """Integer.parseInt: optional sign, ASCII digits, int32 range; no blanks."""
    if not _INT.match(s):
        return None
    v = int(s)
    return v if -2147483648 <= v <= 2147483647 else None


def parse_java_double(s: str) -> Optional[float]:
    """Double.parseDouble: String.trim() (everything <= U+0020), then Java's FloatingPointLiteral."""
    t = s.strip("".join(chr(c) for c in range(33)))
    if not _DBL.match(t):
        return None
    return float(t.replace("Infinity", "inf").replace("NaN", "nan"))


def encode_row(line: str, delim: str, attrs: Sequence[tuple], class_ordinal: int, class_values: Sequence[str],
               skip_fields: int = 0) -> Tuple[Optional[List[int]], Optional[int], int]:
    """One row -> ([bin per attribute], class code, 0) or (None, None, reason).  attrs as in Engine.load_csv."""
    if line.endswith("\r"):
        line = line[:-1]
    items = java_split(line, re.escape(delim), False)
    bins: List[int] = []
    for ordinal, kind, vals in attrs:
        if ordinal + skip_fields >= len(items):
            return None, None, BAD_SHORT_ROW
        raw = items[ordinal + skip_fields]
        if kind == "int":
            v = parse_java_int(raw)
            if v is None:
                return None, None, BAD_INT
            bins.append(bisect.bisect_left(list(vals), v))
        elif kind == "double":
            d = parse_java_double(raw)
            if d is None:
                return None, None, BAD_DOUBLE
            bins.append(sum(1 for p in vals if p < d))
        else:
            if raw not in vals:
                return None, None, BAD_CATEGORY
            bins.append(list(vals).index(raw))
    if class_ordinal < 0:      # rows without a class field (tree/DataPartitioner.java's input): one class, code 0
        return bins, 0, 0
    if class_ordinal + skip_fields >= len(items):
        return None, None, BAD_SHORT_ROW
    c = items[class_ordinal + skip_fields]
    if c not in class_values:
        return None, None, BAD_CLASS
    return bins, list(class_values).index(c), 0


def encode_text(text: str, delim: str, attrs: Sequence[tuple], class_ordinal: int, class_values: Sequence[str], skip_fields: int = 0):
    """Whole text -> (columns, class codes, first row per class value) or the first offending (row, reason)."""
    lines = text.split("\n")
    if lines and lines[-1] == "":
        lines.pop()
    cols = [[] for _ in attrs]
    cls: List[int] = []
    first = [-1] * len(class_values)
    for r, ln in enumerate(lines):
        bins, c, reason = encode_row(ln, delim, attrs, class_ordinal, class_values, skip_fields)
        if reason:
            return None, None, None, (r, reason)
        for a, b in enumerate(bins):
            cols[a].append(b)
        cls.append(c)
        if first[c] < 0:
            first[c] = r
    return cols, cls, first, None
