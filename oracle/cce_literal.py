"""
ORACLE / TEST INFRASTRUCTURE ONLY -- never imported by the product (avenir_b200/).

Literal restatement of org.avenir.explore.CategoricalContinuousEncoding, row by row:
  explore/CategoricalContinuousEncoding.java:98-148   mapper: items = row.split(delim, -1); per listed ordinal the key (ord, items[ord])
                                                      with (1, 0) for a positive row, (0, 1) otherwise; for weight of evidence also (-1, "*")
  explore/CategoricalContinuousEncoding.java:155-175  combiner: sums
  explore/CategoricalContinuousEncoding.java:243-256  reducer: one ClassAttributeCounter per key
  explore/CategoricalContinuousEncoding.java:198-236  cleanup: supervised ratio (pos * scale) / total in int arithmetic, or
                                                      (int)(Math.log((pos / allPos) / (max(neg, 1) / allNeg)) * scale); `ord,value,code`
PARITY UNPINNED (no JVM, no reference tests or fixtures for this job).  The reducer iterates a HashMap keyed by chombo Tuples whose
hashCode is not in the reference's tree: the ORDER of the output lines is unknown, this restatement returns them in reducer key
order (ordinal, value) and callers compare as sets.  Math.log = the platform's log.
"""
from __future__ import annotations

import math
from typing import Dict, List, Sequence

from .literal import java_split


class JobFailure(Exception):
    pass


def _i32(v: int) -> int:
    v &= 0xFFFFFFFF
    return v - (1 << 32) if v >= (1 << 31) else v


def _jint(x: float) -> int:
    """(int) of a double in Java: NaN -> 0, saturating, truncation toward zero."""
    if x != x:
        return 0
    if x >= 2147483647.0:
        return 2147483647
    if x <= -2147483648.0:
        return -2147483648
    return int(x)


def _java_int_div(a: int, b: int) -> int:
    if b == 0:
        raise JobFailure("ArithmeticException: / by zero")
    q = abs(a) // abs(b)
    return _i32(q if (a >= 0) == (b >= 0) else -q)


def categorical_continuous_encoding(lines: Sequence[str], conf: Dict[str, str]) -> List[str]:
    delim = conf.get("field.delim.regex", ",")
    delim_out = conf.get("field.delim", ",")
    for key, what in (("coe.cat.attribute.ordinals", "missing categorical attribute ordinals"), ("coe.encoding.strategy", "missing encoding strategy"),
                      ("coe.class.attr.ordinal", "missing class atrribute ordinal"), ("coe.pos.class.attr.value", "missing positive class attribute value"),
                      ("coe.output.scale", "missing output scale")):
        if not conf.get(key):
            raise JobFailure("IllegalStateException: " + what)
    ordinals = [int(x) for x in conf["coe.cat.attribute.ordinals"].split(",")]
    woe = conf["coe.encoding.strategy"] == "weightOfEvidence"
    class_ord = int(conf["coe.class.attr.ordinal"])
    pos_value = conf["coe.pos.class.attr.value"]
    scale = int(conf["coe.output.scale"])
    counters: Dict[tuple, List[int]] = {}
    for line in lines:
        items = java_split(line, delim, True)
        if class_ord >= len(items):
            raise JobFailure(f"ArrayIndexOutOfBoundsException: {class_ord}")
        positive = items[class_ord] == pos_value
        keys = []
        for o in ordinals:
            if o >= len(items):
                raise JobFailure(f"ArrayIndexOutOfBoundsException: {o}")
            keys.append((o, items[o]))
        if woe:
            keys.append((-1, "*"))
        for k in keys:
            c = counters.setdefault(k, [0, 0])
            c[0 if positive else 1] += 1
    all_counter = counters.pop((-1, "*"), None) if woe else None
    out = []
    for (o, v) in sorted(counters):
        pos, neg = counters[(o, v)]
        if woe:
            def ddiv(a, b):   # Java double division
                if b == 0:
                    return float("nan") if a == 0 else math.copysign(float("inf"), a)
                return a / b
            w = ddiv(float(pos), float(all_counter[0]))
            w = ddiv(w, ddiv(float(1 if neg == 0 else neg), float(all_counter[1]))) if not math.isnan(w) else w
            if math.isnan(w):
                lw = w
            elif w == 0:
                lw = float("-inf")
            elif w < 0:
                lw = float("nan")
            elif math.isinf(w):
                lw = w
            else:
                lw = math.log(w)
            value = _jint(lw * scale)   # (inf * 0 is NaN -> 0)
        else:
            value = _java_int_div(_i32(pos * scale), _i32(pos + neg))
        out.append(f"{o}{delim_out}{v}{delim_out}{value}")
    return out
