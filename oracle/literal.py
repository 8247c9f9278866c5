"""
ORACLE / TEST INFRASTRUCTURE ONLY -- never imported by the product (avenir_b200/).

Literal CPU restatement of the reference's tree-building path, kept in the reference's own
formulation (string keys, one emitted record per row x candidate split x true predicate, hash-map
counting) so that results -- including every HashMap-order-dependent choice -- can be derived the
way a JDK-8 JVM running the reference would derive them.  Pure-Python loops: small inputs only.

PARITY UNPINNED: the reference has no tests and cannot run here (no JVM; chombo/hoidla jars
absent).  What pins this file: the golden vectors of tests/golden/ (gini values and
JSON shape of resource/dec_tree_rules.json, split enumerations) and agreement with
the independently written histogram formulation (oracle/hist_oracle.c).

Every function cites the reference lines it follows (paths relative to
/root/reference/src/main/java/org/avenir/).
Assumed semantics of absent third-party code (chombo 1.0 / hoidla 1.0) are those of SURVEY App. D.
"""
from __future__ import annotations

import json
import math
import re
from typing import Dict, List, Optional, Sequence, Tuple

from .javaemu import (JavaHashMap, JavaRandom, double_to_string, fdlibm_log)

ROOT_PATH = "$root"      # tree/DecisionTreeBuilder.java:63
CHILD_PATH = "$child"    # :64
SPLIT_DELIM = ":"        # :66


# ----------------------------------------------------------------------------------------------
# Java arithmetic helpers
# ----------------------------------------------------------------------------------------------
def jdiv(a: float, b: float) -> float:
    """Java double division (IEEE: x/0 = +-Infinity, 0/0 = NaN)."""
    try:
        return a / b
    except ZeroDivisionError:
        if a != a or a == 0.0:
            return float("nan")
        neg = (a < 0) != (math.copysign(1.0, b) < 0)
        return float("-inf") if neg else float("inf")


def jlog(x: float) -> float:
    """Math.log restated as StrictMath.log (fdlibm), SURVEY App. E."""
    return fdlibm_log(x)


def jmul(a: float, b: float) -> float:
    return a * b  # Python floats follow IEEE for inf/nan products


def java_split(s: str, regex: str, keep_trailing: bool) -> List[str]:
    """String.split(regex[, -1]).  Without the -1 limit Java drops trailing empty strings."""
    parts = re.split(regex, s)
    if not keep_trailing:
        while len(parts) > 1 and parts[-1] == "":
            parts.pop()
        if len(parts) == 1 and parts[0] == "" and s != "":
            parts = []
    return parts


def java_int_cast(x: float) -> int:
    """(int) of a double: truncation toward zero (saturation not needed at our ranges)."""
    return int(x)


# ----------------------------------------------------------------------------------------------
# configuration + schema (chombo stand-ins)
# ----------------------------------------------------------------------------------------------
def load_properties(path: str) -> Dict[str, str]:
    """java.util.Properties subset: key=value, '#'/'!' comments, later keys override."""
    conf: Dict[str, str] = {}
    with open(path) as f:
        for line in f:
            s = line.strip()
            if not s or s[0] in "#!":
                continue
            m = re.match(r"([^=:\s]+)\s*[=:]?\s*(.*)$", s)
            if m:
                conf[m.group(1)] = m.group(2)
    return conf


class Field:
    def __init__(self, d: dict):
        self.name = d.get("name")
        self.ordinal = int(d["ordinal"])
        self.is_id = bool(d.get("id", False))
        self.feature = bool(d.get("feature", False))
        self.data_type = d.get("dataType")
        self.min = float(d.get("min", 0))
        self.max = float(d.get("max", 0))
        self.split_scan_interval = float(d.get("splitScanInterval", 0))
        self.max_split = int(d.get("maxSplit", 0))
        self.cardinality = list(d.get("cardinality") or [])
        self.bucket_width = int(d.get("bucketWidth", 0))

    def is_integer(self):
        return self.data_type == "int"

    def is_double(self):
        return self.data_type == "double"

    def is_categorical(self):
        return self.data_type == "categorical"


class FeatureSchema:
    def __init__(self, d: dict):
        self.fields = [Field(x) for x in d["fields"]]

    @staticmethod
    def load(path: str) -> "FeatureSchema":
        with open(path) as f:
            return FeatureSchema(json.load(f))

    def find_field_by_ordinal(self, o: int) -> Field:
        for f in self.fields:
            if f.ordinal == o:
                return f
        raise KeyError(o)

    def feature_field_ordinals(self) -> List[int]:
        return [f.ordinal for f in self.fields if f.feature]

    def find_class_attr_field(self) -> Field:
        for f in self.fields:
            if not f.is_id and not f.feature:
                return f
        raise KeyError("class attribute")


# ----------------------------------------------------------------------------------------------
# tree/SplitManager.java
# ----------------------------------------------------------------------------------------------
class Predicate:
    """IntPredicate / DoublePredicate / CategoricalPredicate (tree/SplitManager.java:741-950)."""

    def __init__(self, attr: int, op: str, kind: str, value, other=None):
        self.attr, self.op, self.kind, self.value, self.other = attr, op, kind, value, other
        self._s: Optional[str] = None

    def evaluate(self, operand) -> bool:
        if self.kind == "cat":
            if self.op != "in":
                raise ValueError("Illegal categorical attribute operator")
            return operand in self.value                      # :928-937
        v, o = self.value, self.other
        if self.op == "ge":                                   # :771-796 / :851-874
            r = operand >= v
            if o is not None:
                r = r and operand < o
        elif self.op == "gt":
            r = operand > v
            if o is not None:
                r = r and operand <= o
        elif self.op == "le":
            r = operand <= v
            if o is not None:
                r = r and operand > o
        elif self.op == "lt":
            r = operand < v
            if o is not None:
                r = r and operand >= o
        else:
            raise ValueError("Illegal attribute operator")
        return r

    def __str__(self) -> str:
        if self._s is None:
            if self.kind == "cat":
                self._s = f"{self.attr} {self.op} " + ":".join(self.value)      # :942-948
            elif self.kind == "int":
                self._s = f"{self.attr} {self.op} {self.value}"                 # :805-812
                if self.other is not None:
                    self._s += f" {self.other}"
            else:
                self._s = f"{self.attr} {self.op} {double_to_string(self.value)}"  # :880-889
                if self.other is not None:
                    self._s += f" {double_to_string(self.other)}"
        return self._s


class SplitManager:
    def __init__(self, schema: FeatureSchema, pred_delim: str):
        self.schema = schema
        self.pred_delim = pred_delim
        self.custom_base: Optional[List[int]] = None
        self._all_split_predicates: Dict[int, List[List[Predicate]]] = {}
        self._remaining: Dict[str, List[int]] = {}

    # :141-147
    def base_attribute_ordinals(self) -> List[int]:
        return self.custom_base if self.custom_base is not None else self.schema.feature_field_ordinals()

    def get_all_attributes(self) -> List[int]:                                   # :152-157
        return list(self.base_attribute_ordinals())

    def _current_attributes(self, path: str) -> List[int]:                       # :195-208
        attrs = []
        for sp in java_split(path, self.pred_delim, False):
            if sp != ROOT_PATH:
                attrs.append(int(re.split(r"\s+", sp)[0]))
        return attrs

    def get_remaining_attributes(self, path: Optional[str]) -> List[int]:        # :213-232
        if path is None:
            return self.get_all_attributes()
        c = self._remaining.get(path)
        if c is None:
            cur = [] if path == ROOT_PATH else self._current_attributes(path)
            c = [o for o in self.base_attribute_ordinals() if o not in cur]
            self._remaining[path] = c
        return c

    # ---- numeric enumeration (:256-330, :337-399)
    def int_split_points(self, field: Field) -> List[List[int]]:
        out: List[List[int]] = []
        mn, mx, itv = field.min, field.max, field.split_scan_interval
        if java_int_cast(jdiv(mx - mn, itv)) == 0:                                # :298-301
            itv = (mx - mn) / 2

        def rec(splits: Optional[List[int]]):
            if splits is None:
                sp = java_int_cast(mn + itv)                                     # :305
                while sp < mx:
                    ns = [sp]
                    out.append(ns)
                    rec(ns)
                    sp = java_int_cast(sp + itv)     # `int split; split += double` (App. B-8)
            else:
                ln = len(splits)
                if ln < field.max_split - 1:                                     # :314
                    sp = java_int_cast(splits[-1] + itv)                         # :316
                    while sp < mx:
                        ns = splits + [sp]
                        out.append(ns)
                        rec(ns)
                        sp = java_int_cast(sp + itv)
        rec(None)
        return out

    def double_split_points(self, field: Field) -> List[List[float]]:
        out: List[List[float]] = []
        mn, mx, itv = field.min, field.max, field.split_scan_interval
        if java_int_cast(jdiv(mx - mn, itv)) == 0:
            itv = (mx - mn) / 2

        def rec(splits: Optional[List[float]]):
            if splits is None:
                sp = mn + itv                                                    # :374
                while sp < mx:
                    ns = [sp]
                    out.append(ns)
                    rec(ns)
                    sp = sp + itv
            else:
                if len(splits) < field.max_split - 1:                            # :383
                    sp = splits[-1] + itv                                        # :385
                    while sp < mx:
                        ns = splits + [sp]
                        out.append(ns)
                        rec(ns)
                        sp = sp + itv
        rec(None)
        return out

    @staticmethod
    def _numeric_predicates(attr: int, pts: Sequence, kind: str) -> List[Predicate]:
        """createIntAttrPredicates / createDoubleAttrPredicates (:635-694).  For doubles the 4-arg
        constructor drops otherBound (:840-843)."""
        preds: List[Predicate] = []
        n = len(pts)
        if n == 1:
            preds.append(Predicate(attr, "le", kind, pts[0]))
            preds.append(Predicate(attr, "gt", kind, pts[0]))
        else:
            for i in range(n):
                if i == n - 1:
                    preds.append(Predicate(attr, "le", kind, pts[i]))
                    preds.append(Predicate(attr, "gt", kind, pts[i]))
                elif i == 0:
                    preds.append(Predicate(attr, "le", kind, pts[i]))
                else:
                    other = pts[i - 1] if kind == "int" else None
                    preds.append(Predicate(attr, "le", kind, pts[i], other))
        return preds

    # ---- categorical enumeration (:405-569)
    def categorical_partitions(self, cardinality: List[str], max_split: int) -> List[List[List[str]]]:
        total: List[List[List[str]]] = []
        for gr in range(2, max_split + 1):                                       # :414
            sl: List[List[List[str]]] = []
            self._create_cat_partitions(sl, cardinality, 0, gr, guard=True)
            total.extend(sl)
        return total

    def _create_cat_partitions(self, split_list, cardinality, idx, num_groups, guard):
        """createCategoricalPartitions (:452-515).  guard=False gives the ClassPartitionGenerator
        twin (explore/ClassPartitionGenerator.java:322-390), which lacks the size>numGroups test."""
        if idx == 0:
            full = [[cardinality[i]] for i in range(num_groups)]                 # :557-568
            if guard:
                split_list.append(full)
                if len(cardinality) > num_groups:                                # :462
                    split_list.extend(self._create_partial(cardinality, num_groups - 1, num_groups, guard))
                    self._create_cat_partitions(split_list, cardinality, idx + num_groups, num_groups, guard)
            else:
                partial = self._create_partial(cardinality, num_groups - 1, num_groups, guard)
                split_list.append(full)
                split_list.extend(partial)
                self._create_cat_partitions(split_list, cardinality, idx + num_groups, num_groups, guard)
        elif idx < len(cardinality):
            new_list = []
            el = cardinality[idx]
            for sp in split_list:
                if len(sp) == num_groups:                                        # :477
                    for i in range(num_groups):
                        nsp = [list(g) + ([el] if j == i else []) for j, g in enumerate(sp)]
                        new_list.append(nsp)
                else:                                                            # :493
                    nsp = [list(g) for g in sp]
                    nsp.append([el])
                    new_list.append(nsp)
            if idx < len(cardinality) - 1:                                       # :507
                new_list.extend(self._create_partial(cardinality, idx, num_groups, guard))
            split_list[:] = new_list
            self._create_cat_partitions(split_list, cardinality, idx + 1, num_groups, guard)

    def _create_partial(self, cardinality, idx, num_groups, guard):
        """createPartialCategoricalSplit (:524-547)."""
        out = []
        if num_groups == 2:
            out.append([[cardinality[i] for i in range(idx + 1)]])
        else:
            self._create_cat_partitions(out, cardinality[: idx + 1], 0, num_groups - 1, guard)
        return out

    def split_predicates(self, attr: int) -> List[List[Predicate]]:
        """create{Int,Double,Categorical}AttrSplitPredicates, memoised (:256-283,337-353,405-445)."""
        sp = self._all_split_predicates.get(attr)
        if sp is None:
            f = self.schema.find_field_by_ordinal(attr)
            if f.is_integer():
                sp = [self._numeric_predicates(attr, p, "int") for p in self.int_split_points(f)]
            elif f.is_double():
                sp = [self._numeric_predicates(attr, p, "dbl") for p in self.double_split_points(f)]
            elif f.is_categorical():
                sp = [[Predicate(attr, "in", "cat", g) for g in groups]
                      for groups in self.categorical_partitions(f.cardinality, f.max_split)]
            else:
                raise TypeError("NullPointerException: attribute type has no split predicates")
            self._all_split_predicates[attr] = sp
        return sp


# ----------------------------------------------------------------------------------------------
# util/InfoContentStat.java
# ----------------------------------------------------------------------------------------------
class InfoContentStat:
    def __init__(self):
        self.class_val_count = JavaHashMap()
        self.class_val_pr = JavaHashMap()
        self.total_count = 0
        self.stat = 0.0

    def count_class_val(self, cv: str, count: int):                              # :58-64
        if self.class_val_count.get(cv) is None:
            self.class_val_count.put(cv, 0)
        self.class_val_count.put(cv, self.class_val_count.get(cv) + count)

    def process_stat(self, is_entropy: bool) -> float:                           # :71-101
        self.stat = 0.0
        self.total_count = 0
        for k in self.class_val_count.keys():
            self.total_count += self.class_val_count.get(k)
        if is_entropy:
            log2 = jlog(2.0)
            for k in self.class_val_count.keys():
                pr = jdiv(float(self.class_val_count.get(k)), float(self.total_count))
                self.stat -= jdiv(jmul(pr, jlog(pr)), log2)
                self.class_val_pr.put(k, pr)
        else:
            pr_sq = 0.0
            for k in self.class_val_count.keys():
                pr = jdiv(float(self.class_val_count.get(k)), float(self.total_count))
                pr_sq += pr * pr
                self.class_val_pr.put(k, pr)
            self.stat = 1.0 - pr_sq
        return self.stat


# ----------------------------------------------------------------------------------------------
# tree/DecisionPathList.java  (kept as plain dicts in the Jackson bean shape)
# ----------------------------------------------------------------------------------------------
def _pred_bean(attribute=0, predicate_str=None, operator=None, value_int=0, value_dbl=0.0,
               categorical_values=None, other_int=None, other_dbl=None) -> dict:
    return {"attribute": attribute, "predicateStr": predicate_str, "operator": operator,
            "valueInt": value_int, "valueDbl": value_dbl, "categoricalValues": categorical_values,
            "otherBoundInt": other_int, "otherBoundDbl": other_dbl}


def deserialize_predicate(pred_str: str, field: Field) -> dict:
    """DecisionPathPredicate.create{Int,Double,Categorical}Predicate (:402-447)."""
    items = re.split(r"\s+", pred_str)
    if field.is_integer():
        return _pred_bean(int(items[0]), pred_str, items[1], value_int=int(items[2]),
                          other_int=int(items[3]) if len(items) == 4 else None)
    if field.is_double():
        return _pred_bean(int(items[0]), pred_str, items[1], value_dbl=float(items[2]),
                          other_dbl=float(items[3]) if len(items) == 4 else None)
    if field.is_categorical():
        return _pred_bean(int(items[0]), pred_str, items[1], categorical_values=items[2].split(":"))
    raise ValueError("invalid data type for predicates")


def create_predicates(path: str, schema: FeatureSchema, pred_delim: str) -> List[dict]:
    """DecisionPathPredicate.createPredicates (:454-472)."""
    if path == ROOT_PATH:
        return [_pred_bean(predicate_str=ROOT_PATH)]
    out = []
    for item in java_split(path, pred_delim, False):
        if item == ROOT_PATH:
            out.append(_pred_bean(predicate_str=ROOT_PATH))
        else:
            attr = int(re.split(r"\s+", item)[0])
            out.append(deserialize_predicate(item, schema.find_field_by_ordinal(attr)))
    return out


def find_majority_class(class_val_pr: JavaHashMap) -> Tuple[str, float]:
    """DecisionPath.findMajorityClass (:175-192)."""
    maj = -1.0
    out = ""
    for cv in class_val_pr.keys():
        if class_val_pr.get(cv) > maj:
            maj = class_val_pr.get(cv)
            out = cv
    mn = 1.0 - maj
    if mn > 0:
        conf = maj / mn
        conf = 100.0 if conf > 100 else conf
    else:
        conf = 100.0
    return out, conf


def strip_split_id(preds: List[str]) -> List[str]:
    """DecisionPathList.stripSplitId (:97-107); BasicUtils.splitOnFirstOccurence assumed App. D."""
    out = []
    for p in preds:
        if p == ROOT_PATH:
            out.append(p)
        else:
            i = p.find(SPLIT_DELIM)
            if i < 0:
                raise ValueError("delimiter not found in " + p)
            out.append(p[i + 1:])
    return out


def find_decision_path(dec_path_list: dict, pred_strings: List[str]) -> Optional[dict]:
    """findDecisionPath(String[]) + isMatchedByPredicates (:71-81, :198-214): prefix match."""
    for dp in dec_path_list["decisionPaths"]:
        preds = dp.get("predicates")
        if preds is None:
            if pred_strings[0] == ROOT_PATH:
                return dp
            continue
        ok = True
        for i, p in enumerate(preds):
            if i >= len(pred_strings):
                raise IndexError("ArrayIndexOutOfBoundsException")
            if p["predicateStr"] != pred_strings[i]:
                ok = False
                break
        if ok:
            return dp
    return None


def find_decision_path_by_string(dec_path_list: dict, s: str) -> Optional[dict]:
    """findDecisionPath(String) + isMatchedByPredicateString (:86-95, :220-228).  For non-root
    entries the reference compares against Object.toString() of the predicates, never equal."""
    for dp in dec_path_list["decisionPaths"]:
        if dp.get("predicates") is None and s == ROOT_PATH:
            return dp
    return None


# ----------------------------------------------------------------------------------------------
# tree/DecisionTreeBuilder.java
# ----------------------------------------------------------------------------------------------
class DecisionTreeBuilderLiteral:
    """One MapReduce job = one tree level (tree/DecisionTreeBuilder.java:70-95)."""

    def __init__(self, conf: Dict[str, str], schema: FeatureSchema, dec_path_in: Optional[dict],
                 rng: Optional[JavaRandom] = None):
        g = conf.get
        self.conf = conf
        self.schema = schema
        self.dec_path_list = dec_path_in
        self.tree_available = dec_path_in is not None                             # :149-154
        self.field_delim_regex = g("field.delim.regex", ",")                      # :143
        self.field_delim_out = g("field.delim.out", ",")                          # :433
        self.dec_path_delim = g("dtb.dec.path.delim", ";")                        # :157
        self.split_manager = SplitManager(schema, self.dec_path_delim)
        cb = g("dtb.custom.base.attributes")
        if cb:
            self.split_manager.custom_base = [int(x) for x in cb.split(",")]      # :160-166
        self.attr_strategy = g("dtb.split.attribute.selection.strategy", "notUsedYet")
        self.class_field = schema.find_class_attr_field()
        self.sampling = g("dtb.sub.sampling.strategy", "withReplace")             # :179
        self.sampling_rate = int(g("dtb.sub.sampling.rate", "0") or 0)
        self.sampling_buffer_size = int(g("dtb.sub.sampling.buffer.size", "10000"))
        self.algorithm = g("dtb.split.algorithm", "giniIndex")                    # :434
        self.total_population = int(g("dtb.total.population", "-1"))              # :439
        self.split_sel = g("dtb.split.select.strategy", "best")                   # :442
        self.top_split_count = int(g("dtb.top.split.count", "3"))
        self.stop_strategy = g("dtb.path.stopping.strategy", "minInfoGain")       # :448
        self.max_depth = int(g("dtb.max.depth.limit") or -1)
        self.min_info_gain = float(g("dtb.min.info.gain.limit") or -1)
        self.min_population = int(g("dtb.min.population.limit") or -1)
        if self.stop_strategy not in ("maxDepth", "minInfoGain", "minPopulation"):
            raise ValueError("invalid stopping strategy " + self.stop_strategy)
        self.rng = rng or JavaRandom(0)
        self.valid_dec_paths: Dict[str, bool] = {}
        self.emitted: List[Tuple[Tuple[str, ...], str]] = []

    # ---------------- mapper
    def _root_map(self, record: str):                                            # :248-258
        self.emitted.append(((ROOT_PATH,), record))

    def map_all(self, lines: Sequence[str]):
        """BuilderMapper.map + cleanup (:196-244).  `withReplace` follows the INTENDED semantics
        (buffer of B rows -> B uniform draws; = explore/BaggingSampler.java:101-122), not the
        shipped code, whose shadowed samplingBufferSize makes it emit one row."""
        if not self.tree_available:
            if self.sampling == "withReplace":
                B = self.sampling_buffer_size
                for s in range(0, len(lines), B):
                    buf = lines[s:s + B]
                    n = len(buf)
                    for _ in range(n):
                        sel = java_int_cast(self.rng.next_double() * n)
                        sel = n - 1 if sel == n else sel
                        self._root_map(buf[sel])
            elif self.sampling == "withoutReplace":
                for ln in lines:
                    if java_int_cast(self.rng.next_double() * 100) < self.sampling_rate:   # :232-235
                        self._root_map(ln)
            else:
                for ln in lines:
                    self._root_map(ln)
        else:
            for ln in lines:
                self._path_map(ln)

    def _split_attributes(self, cur_path: Optional[str]) -> List[int]:           # :365-381
        if self.attr_strategy == "all":
            return self.split_manager.get_all_attributes()
        if self.attr_strategy == "notUsedYet":
            return self.split_manager.get_remaining_attributes(cur_path)
        if self.attr_strategy in ("randomAll", "randomNotUsedYet"):
            raise TypeError("NullPointerException: splitAttrs is never assigned (App. B-4)")
        raise ValueError("invalid splitting attribute selection strategy")

    def _path_map(self, record: str):                                            # :266-359
        items = java_split(record, self.field_delim_regex, True)
        cur = items[0]
        predicates = strip_split_id(java_split(cur, self.dec_path_delim, False))
        cur_path = self.dec_path_delim.join(predicates)
        status = self.valid_dec_paths.get(cur_path)
        if status is None:
            status = find_decision_path(self.dec_path_list, predicates) is not None
            self.valid_dec_paths[cur_path] = status
        if not status:
            return
        split_id = 0
        for attr in self._split_attributes(cur_path):
            if attr > len(items) - 2:
                raise IndexError("attrbite index out of bound attr:%d" % attr)
            field = self.schema.find_field_by_ordinal(attr)
            if field.is_integer():
                val = int(items[attr + 1])
            elif field.is_double():
                val = float(items[attr + 1])
            elif field.is_categorical():
                val = items[attr + 1]
            else:
                raise TypeError("NullPointerException")
            for preds in self.split_manager.split_predicates(attr):
                split_id += 1
                matched = False
                for p in preds:
                    if p.evaluate(val):
                        matched = True
                        key = tuple(java_split(items[0], self.dec_path_delim, False)) + \
                            (str(split_id) + SPLIT_DELIM + str(p),)
                        pos = record.index(self.field_delim_regex)
                        self.emitted.append((key, record[pos + len(self.field_delim_regex):]))
                if not matched:
                    raise RuntimeError("no matching predicate for attribute: %d" % attr)

    # ---------------- reducer
    def reduce_all(self) -> Tuple[dict, List[str]]:
        """Sort/shuffle (chombo Tuple.compareTo assumed element-wise, App. D) then
        BuilderReducer.reduce (:730-767) and cleanup (:466-472)."""
        class_ord = self.class_field.ordinal
        dec_paths = JavaHashMap()
        out_lines: List[str] = []
        order = sorted(range(len(self.emitted)), key=lambda i: self.emitted[i][0])  # stable
        prev_key = None
        stats = None
        for i in order:
            key, value = self.emitted[i]
            if key != prev_key:
                prev_key = key
                dec_path = self.dec_path_delim.join(key)
                if len(key) > 1:
                    parent = self.dec_path_delim.join(key[:-1])
                    child = key[-1]
                else:
                    parent = key[0]
                    child = CHILD_PATH
                cand = dec_paths.get(parent)
                if cand is None:
                    cand = JavaHashMap()
                    dec_paths.put(parent, cand)
                stats = cand.get(child)
                if stats is None:
                    stats = InfoContentStat()
                    cand.put(child, stats)
            cv = java_split(value, self.field_delim_out, False)[class_ord]
            stats.count_class_val(cv, 1)
            out_lines.append(dec_path + self.field_delim_out + value)
        if self.tree_available:
            return self._expand_tree(dec_paths), out_lines
        return self._generate_root(dec_paths), out_lines

    def _pr_dict(self, m: JavaHashMap) -> dict:
        return {k: v for k, v in m.items()}   # dict keeps the HashMap iteration order

    def _finish(self, paths: Optional[List[dict]]) -> dict:                      # :713-725
        if self.total_population > 0:
            if paths is None:
                raise TypeError("NullPointerException in withTotalPopulation")
            for p in paths:
                p["support"] = jdiv(float(p["population"]), float(self.total_population))
        return {"decisionPaths": paths}

    def _generate_root(self, dec_paths: JavaHashMap) -> dict:                    # :478-493
        st = dec_paths.get(ROOT_PATH).get(CHILD_PATH)
        st.process_stat(self.algorithm == "entropy")
        path = {"predicates": None, "population": st.total_count, "infoContent": st.stat,
                "stopped": False, "classValPr": self._pr_dict(st.class_val_pr),
                "outputClassVal": "", "confidence": 0.0, "support": 0.0}
        return self._finish([path])

    def _should_stop(self, st: InfoContentStat, parent_stat: float, depth: int) -> bool:
        """tree/DecisionPathStoppingStrategy.java:57-70."""
        if self.stop_strategy == "minPopulation":
            return st.total_count < self.min_population
        if self.stop_strategy == "minInfoGain":
            return (parent_stat - st.stat) < self.min_info_gain
        return depth >= self.max_depth

    def _expand_tree(self, dec_paths: JavaHashMap) -> dict:                      # :499-616
        is_entropy = self.algorithm == "entropy"
        # infoContentBySplit (:636-668)
        by_split = JavaHashMap()
        for parent in dec_paths.keys():
            filt = self.dec_path_delim.join(strip_split_id(java_split(parent, self.dec_path_delim, False)))
            sic = by_split.get(filt)
            if sic is None:
                sic = JavaHashMap()
                by_split.put(filt, sic)
            child_stats = dec_paths.get(parent)
            for pred in child_stats.keys():
                i = pred.index(SPLIT_DELIM)
                sid, predicate = pred[:i], pred[i + 1:]
                pic = sic.get(sid)
                if pic is None:
                    pic = JavaHashMap()
                    sic.put(sid, pic)
                pic.put(predicate, child_stats.get(pred))
        new_paths: Optional[List[dict]] = None
        for parent in by_split.keys():
            if parent == ROOT_PATH:                                              # :692-705
                pdp = find_decision_path_by_string(self.dec_path_list, ROOT_PATH)
            else:
                pdp = find_decision_path(self.dec_path_list, java_split(parent, self.dec_path_delim, False))
            if pdp is None:
                raise RuntimeError("parent decision path not found: " + parent)
            parent_stat = pdp["infoContent"]
            min_info = 1000000.0
            selected, selected_attr = None, -1
            ranked: Dict[float, Tuple[str, int]] = {}
            sic = by_split.get(parent)
            for sid in sic.keys():                                               # :527
                pic = sic.get(sid)
                weighted = 0.0
                total = 0
                attr = -1
                for predicate in pic.keys():                                     # :537
                    attr = int(re.split(r"\s+", predicate)[0])
                    st = pic.get(predicate)
                    weighted += st.process_stat(is_entropy) * st.total_count
                    total += st.total_count
                av = jdiv(weighted, float(total))
                if self.split_sel == "best":
                    if av < min_info:                                            # :551
                        min_info, selected, selected_attr = av, sid, attr
                elif self.split_sel == "randomAmongTop":
                    ranked[av] = (sid, attr)                                     # TreeMap.put: equal keys overwrite
                else:
                    raise RuntimeError("ivalid split slection strategy")
            if self.split_sel == "randomAmongTop":                               # :622-631
                top = [ranked[k] for k in sorted(ranked.keys())][: self.top_split_count]
                selected, selected_attr = top[java_int_cast(self.rng.next_double() * len(top))]
            pic = sic.get(selected)
            parent_preds = create_predicates(parent, self.schema, self.dec_path_delim)
            field = self.schema.find_field_by_ordinal(selected_attr)
            for pred_str in pic.keys():                                          # :586
                pb = deserialize_predicate(pred_str, field)
                st = pic.get(pred_str)
                stop = self._should_stop(st, parent_stat, len(parent_preds) + 1)
                out_cv, conf = find_majority_class(st.class_val_pr)
                path = {"predicates": parent_preds + [pb], "population": st.total_count,
                        "infoContent": st.stat, "stopped": stop,
                        "classValPr": self._pr_dict(st.class_val_pr),
                        "outputClassVal": out_cv, "confidence": conf, "support": 0.0}
                if new_paths is None:
                    new_paths = []
                new_paths.append(path)
        return self._finish(new_paths)

    def run(self, lines: Sequence[str]) -> Tuple[dict, List[str]]:
        self.emitted = []
        self.map_all(lines)
        return self.reduce_all()


def dtb_build(lines: Sequence[str], conf: Dict[str, str], schema: FeatureSchema, expansions: int,
              seed: int = 0, selected_only: bool = True) -> List[dict]:
    """The hand-run level loop of resource/detr.sh:34-54: root pass, then `expansions` passes,
    each fed the previous pass's output file and decision-path list.  Returns the path list of
    every pass.  selected_only drops, between passes, the records the next mapper would reject
    anyway (tree/DecisionTreeBuilder.java:279-292) -- same result, less memory."""
    rng = JavaRandom(seed)
    out = []
    job = DecisionTreeBuilderLiteral(conf, schema, None, rng)
    dpl, data = job.run(lines)
    out.append(dpl)
    for _ in range(expansions):
        if dpl["decisionPaths"] is None:
            break
        job = DecisionTreeBuilderLiteral(conf, schema, dpl, rng)
        new_dpl, new_data = job.run(data)
        out.append(new_dpl)
        if selected_only and new_dpl["decisionPaths"] is not None:
            keep = set()
            for p in new_dpl["decisionPaths"]:
                keep.add(job.dec_path_delim.join(x["predicateStr"] for x in p["predicates"]))
            filt = []
            for ln in new_data:
                head = ln.split(job.field_delim_out, 1)[0]
                sp = job.dec_path_delim.join(strip_split_id(head.split(job.dec_path_delim)))
                if sp in keep:
                    filt.append(ln)
            new_data = filt
        dpl, data = new_dpl, new_data
    return out


# ----------------------------------------------------------------------------------------------
# explore/BaggingSampler.java
# ----------------------------------------------------------------------------------------------
def bagging_sample_indices(n_rows: int, batch_size: int, rng: JavaRandom) -> List[int]:
    """BaggingMapper.map/cleanup/emitBatch (explore/BaggingSampler.java:95-122): per batch of
    `bas.batch.size` rows emit that many uniform draws; tail batch likewise.  Returns the drawn
    row indices in emission order."""
    out = []
    for s in range(0, n_rows, batch_size):
        n = min(batch_size, n_rows - s)
        for _ in range(n):
            out.append(s + java_int_cast(rng.next_double() * n))
    return out


# ----------------------------------------------------------------------------------------------
# explore/ClassPartitionGenerator.java + util/AttributeSplitHandler.java + util/AttributeSplitStat.java
# ----------------------------------------------------------------------------------------------
def java_list_to_string(items: Sequence[str]) -> str:
    return "[" + ", ".join(items) + "]"


class CpgSplit:
    """AttributeSplitHandler.IntegerSplit / CategoricalSplit (util/AttributeSplitHandler.java:135-233)."""

    def __init__(self, points: Optional[List[int]] = None, groups: Optional[List[List[str]]] = None):
        self.points, self.groups = points, groups
        if points is not None:
            self.key = ";".join(str(p) for p in points)                          # :44
        else:
            self.key = ":".join(java_list_to_string(g) for g in groups)          # :208-215

    def segment_index(self, value: str) -> int:
        if self.points is not None:                                              # :148-155
            v = int(value)
            i = 0
            while i < len(self.points) and v > self.points[i]:
                i += 1
            return i
        for i, g in enumerate(self.groups):                                      # :192-206
            if value in g:
                return i
        raise ValueError("split segment not found for " + value)


def cpg_create_partitions(schema: FeatureSchema, split_attrs: Sequence[int],
                          max_cat_groups: int) -> Dict[int, List[CpgSplit]]:
    """PartitionGeneratorMapper.createPartitions (explore/ClassPartitionGenerator.java:235-272)."""
    sm = SplitManager(schema, ";")
    out: Dict[int, List[CpgSplit]] = {}
    for a in split_attrs:
        f = schema.find_field_by_ordinal(a)
        if f.is_integer():
            mn = java_int_cast(f.min + 0.01)                                     # :282-284
            mx = java_int_cast(f.max + 0.01)
            bw = f.bucket_width
            lst: List[List[int]] = []

            def rec(splits):
                if splits is None:
                    sp = mn + bw
                    while sp < mx:
                        ns = [sp]
                        lst.append(ns)
                        rec(ns)
                        sp += bw
                elif len(splits) < f.max_split - 1:                              # :296
                    sp = splits[-1] + bw
                    while sp < mx:
                        ns = splits + [sp]
                        lst.append(ns)
                        rec(ns)
                        sp += bw
            if bw <= 0 and mn < mx:
                raise RuntimeError("bucketWidth 0: the reference loops forever (:287)")
            rec(None)
            out[a] = [CpgSplit(points=p) for p in lst]
        elif f.is_categorical():
            if f.max_split > max_cat_groups:                                     # :250-254
                raise ValueError("more than %d split groups not allwed for categorical attr" % max_cat_groups)
            final = []
            for gr in range(2, f.max_split + 1):
                sl = []
                sm._create_cat_partitions(sl, f.cardinality, 0, gr, guard=False)
                final.extend(sl)
            out[a] = [CpgSplit(groups=g) for g in final]
        # double attrs: no branch in the reference -> no split list -> NPE at next() (a18)
    return out


class SplitStatSegment:
    """util/AttributeSplitStat.java:346-472."""

    def __init__(self, idx: int):
        self.idx = idx
        self.class_val_count = JavaHashMap()
        self.class_val_pr = JavaHashMap()
        self.class_val_conf = JavaHashMap()
        self.total_count = 0

    def count_class_val(self, cv, count):
        if self.class_val_count.get(cv) is None:
            self.class_val_count.put(cv, 0)
        self.class_val_count.put(cv, self.class_val_count.get(cv) + count)

    def process_stat(self, algorithm: str) -> float:                             # :379-411
        stat = 0.0
        self.total_count = 0
        for k in self.class_val_count.keys():
            self.total_count += self.class_val_count.get(k)
        if algorithm == "entropy":
            log2 = jlog(2.0)
            for k in self.class_val_count.keys():
                pr = jdiv(float(self.class_val_count.get(k)), float(self.total_count))
                stat -= jdiv(jmul(pr, jlog(pr)), log2)
                self.class_val_pr.put(k, pr)
        elif algorithm == "giniIndex":
            pr_sq = 0.0
            for k in self.class_val_count.keys():
                pr = jdiv(float(self.class_val_count.get(k)), float(self.total_count))
                pr_sq += pr * pr
                self.class_val_pr.put(k, pr)
            stat = 1.0 - pr_sq
        return stat

    def get_total_count(self) -> int:                                            # :413-420
        if self.total_count == 0:
            for k in self.class_val_count.keys():
                self.total_count += self.class_val_count.get(k)
        return self.total_count

    def count_for(self, cv) -> int:
        c = self.class_val_count.get(cv)
        return 0 if c is None else c

    def class_confidence_ratio(self) -> float:                                   # :452-471
        tot = 0.0
        for cv in self.class_val_conf.keys():
            tot += self.class_val_conf.get(cv)
        ratio = JavaHashMap()
        for cv in self.class_val_conf.keys():
            ratio.put(cv, jdiv(self.class_val_conf.get(cv), tot))
        ent = 0.0
        log2 = jlog(2.0)
        for k in ratio.keys():
            ccr = ratio.get(k)
            lg = jlog(ccr) if ccr == ccr else float("nan")
            ent -= jdiv(jmul(ccr, lg), log2)
        return ent


class SplitStat:
    def __init__(self, key: str):
        self.key = key
        self.segments = JavaHashMap()

    def count_class_val(self, seg: int, cv: str, count: int):                    # :118-127
        s = self.segments.get(seg)
        if s is None:
            s = SplitStatSegment(seg)
            self.segments.put(seg, s)
        s.count_class_val(cv, count)

    def info_content(self) -> float:                                             # :153-170
        total = 0
        for si in self.segments.keys():
            total += self.segments.get(si).get_total_count()
        stat = 0.0
        log2 = jlog(2.0)
        for si in self.segments.keys():
            pr = jdiv(float(self.segments.get(si).get_total_count()), float(total))
            stat -= jdiv(jmul(pr, jlog(pr)), log2)
        return stat

    def process_info_content(self, algorithm: str) -> float:                     # :191-218
        stat_arr, count_arr, total = [], [], 0
        for si in self.segments.keys():
            seg = self.segments.get(si)
            stat_arr.append(seg.process_stat(algorithm))
            c = seg.get_total_count()
            count_arr.append(c)
            total += c
        s = 0.0
        for a, c in zip(stat_arr, count_arr):
            s += a * c
        return jdiv(s, float(total))

    def process_hellinger(self, class_values: List[str]) -> float:               # :240-282
        if len(class_values) != 2:
            raise ValueError("Hellinger distance algorithm is only valid for binary valued class attributes")
        cvc = [0, 0]
        for i in range(2):
            for si in self.segments.keys():
                cvc[i] += self.segments.get(si).count_for(class_values[i])
        sm = 0.0
        for si in self.segments.keys():
            seg = self.segments.get(si)
            v0 = jdiv(float(seg.count_for(class_values[0])), float(cvc[0]))
            v0 = math.sqrt(v0) if v0 == v0 else v0
            v1 = jdiv(float(seg.count_for(class_values[1])), float(cvc[1]))
            v1 = math.sqrt(v1) if v1 == v1 else v1
            sm += (v0 - v1) * (v0 - v1)
        return math.sqrt(sm) if sm == sm else sm

    def process_class_conf(self, class_values: List[str]) -> float:              # :298-336
        cvc = JavaHashMap()
        for cv in class_values:
            for si in self.segments.keys():
                c = self.segments.get(si).count_for(cv)
                if cvc.get(cv) is None:
                    cvc.put(cv, 0)
                cvc.put(cv, cvc.get(cv) + c)
        for cv in class_values:
            for si in self.segments.keys():
                seg = self.segments.get(si)
                seg.class_val_conf.put(cv, jdiv(float(seg.count_for(cv)), float(cvc.get(cv))))
        total, sm = 0, 0.0
        for si in self.segments.keys():
            seg = self.segments.get(si)
            ccr = seg.class_confidence_ratio()
            c = seg.get_total_count()
            sm += ccr * c
            total += c
        return jdiv(sm, float(total))


def cpg_run(lines: Sequence[str], conf: Dict[str, str], schema: FeatureSchema) -> List[str]:
    """ClassPartitionGenerator mapper + combiner + reducer (explore/ClassPartitionGenerator.java:
    200-230, 450-463, 514-592), single reducer.  Returns the output lines."""
    g = conf.get
    delim_regex = g("field.delim.regex", ",")
    delim_out = g("field.delim.out", ",")
    algorithm = g("cpg.split.algorithm", "giniIndex")
    at_root_mapper = g("cpg.at.root", "false") == "true"
    strategy = g("cpg.split.attribute.selection.strategy", "userSpecified")
    attrs_str = g("cpg.split.attributes")
    max_cat = int(g("cpg.max.cat.attr.split.groups", "3"))
    class_field = schema.find_class_attr_field()
    if at_root_mapper:
        split_attrs: List[int] = []
    elif strategy == "userSpecified":
        split_attrs = [int(x) for x in attrs_str.split(",")]
    elif strategy == "all":
        split_attrs = schema.feature_field_ordinals()
    else:
        raise NotImplementedError("cpg strategy " + strategy + " (TODO/NPE/unseeded in the reference)")
    parts = {} if at_root_mapper else cpg_create_partitions(schema, split_attrs, max_cat)
    counts: Dict[tuple, int] = {}
    for ln in lines:
        items = java_split(ln, delim_regex, False)
        cv = items[class_field.ordinal]
        if at_root_mapper:
            k = (-1, "null", -1, cv)
            counts[k] = counts.get(k, 0) + 1
        else:
            for a in split_attrs:
                if a not in parts:
                    raise TypeError("NullPointerException: no split list for attribute %d" % a)
                for sp in parts[a]:
                    k = (a, sp.key, sp.segment_index(items[a]), cv)
                    counts[k] = counts.get(k, 0) + 1
    # reducer set-up (:488-511)
    at_root = attrs_str is None
    output_prob = g("cpg.output.split.prob", "false") == "true"
    parent_info = float(g("cpg.parent.info"))
    out: List[str] = []
    if at_root:
        root = InfoContentStat()
        for k in sorted(counts.keys()):
            root.count_class_val(k[3], counts[k])
        return [double_to_string(root.process_stat(algorithm == "entropy"))]
    attr_ordinals = [int(x) for x in attrs_str.split(",")]
    split_stats: Dict[int, JavaHashMap] = {a: JavaHashMap() for a in attr_ordinals}
    class_values: Dict[int, JavaHashMap] = {a: JavaHashMap() for a in attr_ordinals}
    for k in sorted(counts.keys()):          # Tuple order: int, String, int, String element-wise
        a, key, seg, cv = k
        ss = split_stats[a].get(key)
        if ss is None:
            ss = SplitStat(key)
            split_stats[a].put(key, ss)
        ss.count_class_val(seg, cv, counts[k])
        class_values[a].put(cv, True)
    for a in attr_ordinals:
        stats = JavaHashMap()
        cvs = list(class_values[a].keys())
        for key in split_stats[a].keys():                                        # util/AttributeSplitStat.java:94-100
            ss = split_stats[a].get(key)
            if algorithm in ("entropy", "giniIndex"):
                v = ss.process_info_content(algorithm)
            elif algorithm == "hellingerDistance":
                v = ss.process_hellinger(cvs)
            else:
                v = ss.process_class_conf(cvs)
            stats.put(key, v)
        for key in stats.keys():                                                 # :520
            stat = stats.get(key)
            if algorithm in ("entropy", "giniIndex"):
                gain = parent_info - stat
                ratio = jdiv(gain, split_stats[a].get(key).info_content())
                s = f"{a}{delim_out}{key}{delim_out}{double_to_string(ratio)}"
                if output_prob:                                                  # :539-542, :557-566
                    ss = split_stats[a].get(key)
                    cp = JavaHashMap()
                    for si in ss.segments.keys():
                        cp.put(si, ss.segments.get(si).class_val_pr)
                    frag = []
                    for si in cp.keys():
                        for cv in cp.get(si).keys():
                            frag.append(f"{si}{delim_out}{cv}{delim_out}{double_to_string(cp.get(si).get(cv))}")
                    s += delim_out + delim_out.join(frag)
                out.append(s)
            else:
                out.append(f"{a}{delim_out}{key}{delim_out}{double_to_string(stat)}")
    return out
