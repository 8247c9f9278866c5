"""
ORACLE / TEST INFRASTRUCTURE ONLY -- never imported by the product (avenir_b200/).

Literal restatement of org.avenir.explore.CategoricalClassAffinity, record by record:
  explore/CategoricalClassAffinity.java:119-137   mapper: items = line.split(delim, -1); key (featureOrd, 0 if the class value is the
                                                  positive one else 1), value (classValue, featureValue, distribution)
  explore/CategoricalClassAffinity.java:72, 80    Utility.setTuplePairSecondarySorting: one reduce call per featureOrd, values in key
                                                  order (0 before 1)
  explore/CategoricalClassAffinity.java:186-232   reduce: two HashMap<String, Double> that are FIELDS of the reducer; map 0 cleared at
                                                  the group's first value, map 1 at every later change of the class value
  explore/CategoricalClassAffinity.java:236-261   findAffinityScore over posDistr.keySet(); negDistr.get(...) unboxed (null -> NPE)
  explore/CategoricalClassAffinity.java:213-230   per strategy: "algorithm: <name>", Collections.sort (stable, descending), one record
                                                  holding the group's lines joined by "\\n" (an empty list dies in substring(0, -1))
PARITY UNPINNED (no JVM, no reference tests or fixtures for this job); chombo's secondary sorting is taken to group by the first
key element and to keep the arrival order of equal keys; Math.log = the platform's log.
"""
from __future__ import annotations

import math
from typing import Dict, List, Sequence

from .encode_literal import parse_java_double, parse_java_int
from .javaemu import JavaHashMap, double_to_string
from .literal import java_split


class JobFailure(Exception):
    pass


def categorical_class_affinity(lines: Sequence[str], conf: Dict[str, str]) -> List[str]:
    delim = conf.get("field.delim.regex", ",")
    delim_out = conf.get("field.delim", ",")
    pos_value = conf.get("cca.pos.class.attr.value")
    if not pos_value:
        raise JobFailure("IllegalStateException: missing positive class attribute value")
    if not conf.get("cca.affinity.strategy"):
        raise JobFailure("IllegalStateException: missing affinity strategy list")
    strategies = conf["cca.affinity.strategy"].split(",")
    records = []
    for line in lines:
        items = java_split(line, delim, True)
        ord_ = parse_java_int(items[0])
        if ord_ is None:
            raise JobFailure(f'NumberFormatException: For input string: "{items[0]}"')
        if len(items) < 4:
            raise JobFailure(f"ArrayIndexOutOfBoundsException: {len(items)}")
        distr = parse_java_double(items[3])
        if distr is None:
            raise JobFailure(f'NumberFormatException: For input string: "{items[3]}"')
        records.append(((ord_, 0 if items[1] == pos_value else 1), (items[1], items[2], distr)))
    records.sort(key=lambda r: r[0])   # stable: arrival order inside a key
    maps = [JavaHashMap(), JavaHashMap()]   # reducer fields
    out: List[str] = []
    i = 0
    while i < len(records):
        j = i
        while j < len(records) and records[j][0][0] == records[i][0][0]:
            j += 1
        class_val, cur = None, None
        for _, (cv, fv, d) in records[i:j]:
            if class_val is None:
                class_val, cur = cv, maps[0]
                cur.clear()
            elif cv != class_val:
                class_val, cur = cv, maps[1]
                cur.clear()
            cur.put(fv, d)
        for strategy in strategies:
            out.append("algorithm: " + strategy)
            scores = []
            for fv in maps[0].keys():
                p = maps[0][fv]
                n = maps[1].get(fv)
                if n is None:
                    raise JobFailure("NullPointerException")

                def div(a, b):
                    if b == 0:
                        return float("nan") if a == 0 or a != a else math.copysign(float("inf"), a) * (1 if math.copysign(1, b) > 0 else -1)
                    return a / b

                if strategy == "oddsRatio":
                    scores.append((fv, div(div(p, 1 - p), div(n, 1 - n))))
                elif strategy == "distrDiff":
                    scores.append((fv, p - n))
                elif strategy == "minRisk":
                    scores.append((fv, p * (1 - n)))
                elif strategy == "klDiff":
                    r = div(p, n)
                    lg = float("nan") if (r != r or r < 0) else (float("-inf") if r == 0 else (r if math.isinf(r) else math.log(r)))
                    scores.append((fv, p * lg))
            if not scores:
                raise JobFailure("StringIndexOutOfBoundsException: String index out of range: -1")
            # Collections.sort with the comparator: right < -> 1, right > -> -1, else 0 (stable)
            import functools
            scores.sort(key=functools.cmp_to_key(lambda a, b: 1 if a[1] < b[1] else (-1 if a[1] > b[1] else 0)))
            for fv, sc in scores:
                out.append(f"{records[i][0][0]}{delim_out}{fv}{delim_out}{double_to_string(sc)}")
        i = j
    return out
