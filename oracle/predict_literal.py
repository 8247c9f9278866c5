"""
ORACLE / TEST INFRASTRUCTURE ONLY -- never imported by the product (avenir_b200/).

Literal restatement of the reference's prediction path, row by row and predicate by predicate:
  model/ModelPredictor.java:84-254            (mop.* knobs, output modes, error file)
  tree/DecisionTreeModel.java:56-144          (first decision path whose predicates all hold)
  tree/DecisionPathList.java:356-366          (DecisionPath.doPrediction)
  model/ProbabilisticPredictiveModel.java:41-59, model/PredictiveModel.java:84-96 (cost threshold, error counts)
  model/EnsemblePredictiveModel.java:69-110   (weighted vote, stable sort on count)
PARITY UNPINNED (no JVM, no reference tests).  BasicUtils.formatDouble (chombo, absent) is assumed "%.3f".
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

from .javaemu import JavaHashMap
from .literal import FeatureSchema, java_split


class NullPointer(Exception):
    """The reference would die with a NullPointerException here."""


def _evaluate(pred: dict, field, items: List[str]) -> bool:
    """DecisionTreeModel.evaluate (tree/DecisionTreeModel.java:101-141)."""
    op = pred["operator"]
    ordn = field.ordinal
    if field.is_integer():
        operand = int(items[ordn])
        val = pred["valueInt"]
        if op == "le":
            return operand <= val
        if op == "gt":
            r = operand > val
            if pred.get("otherBoundInt") is not None:
                r = r and operand <= pred["otherBoundInt"]
            return r
        raise ValueError("invalid operator type for int attribute")
    if field.is_double():
        operand = float(items[ordn])
        val = pred["valueDbl"]
        if op == "le":
            return operand <= val
        if op == "gt":
            r = operand > val
            if pred.get("otherBoundInt") is not None:      # sic: the int bound (:124-127)
                r = r and operand <= pred["otherBoundInt"]
            return r
        raise ValueError("invalid operator type for double attribute")
    if field.is_categorical():
        if op == "in":
            return items[ordn] in pred["categoricalValues"]
        raise ValueError("invalid operator type for categorical attribute")
    return False


def tree_predict_class_prob(model: dict, schema: FeatureSchema, items: List[str]) -> Tuple[Optional[str], float]:
    """DecisionTreeModel.predictClassProb (:56-93) + DecisionPath.doPrediction."""
    cache: Dict[str, bool] = {}
    matched = None
    for path in model["decisionPaths"]:
        preds = path["predicates"]
        if preds is None:
            raise NullPointer("predicates == null")
        ev = True
        for p in preds[1:]:
            key = p["predicateStr"]                    # DecisionPathPredicate.equals/hashCode use predicateStr
            if key in cache:
                ev = ev and cache[key]
            else:
                v = _evaluate(p, schema.find_field_by_ordinal(p["attribute"]), items)
                ev = ev and v
                cache[key] = v
            if not ev:
                break
        if ev:
            matched = path
            break
    if matched is None:
        raise NullPointer("no decision path matched")
    pred_cl, max_pr = None, 0.0
    for cl, pr in matched["classValPr"].items():       # Jackson keeps file order (LinkedHashMap)
        if pr > max_pr:
            pred_cl, max_pr = cl, pr
    return pred_cl, max_pr


def model_predictor(lines: Sequence[str], conf: Dict[str, str], schema: FeatureSchema, models: List[dict]):
    """Returns (output lines, error-file text or None)."""
    g = conf.get
    delim_regex = g("field.delim.regex", ",")
    delim_out = g("field.delim.out", ",")
    error_counting = g("mop.error.counting.enabled", "false") == "true"
    cost_based = g("mop.cost.based.prediction.enabled", "false") == "true"
    class_values = [s.strip() for s in g("mop.class.attr.values", "").split(",")] if (error_counting or cost_based) else None
    costs = [float(x) for x in g("mop.miss.class.costs").split(",")] if cost_based else None
    class_ord = int(g("mop.class.attr.ord")) if error_counting else -1
    mode = g("mop.output.mode", "withRecord")
    ensemble = len(models) > 1
    weights = [1.0] * len(models)
    if ensemble and g("mop.ensemble.memeber.weights"):
        weights = [float(x) for x in g("mop.ensemble.memeber.weights").split(",")]
    if ensemble and len(models) % 2 == 0:
        raise ValueError("neem odd number of models in ensemble")
    total = errors = fpos = fneg = 0

    def member_predict(model, items):
        cl, pr = tree_predict_class_prob(model, schema, items)
        if cost_based:
            thr = costs[0] / (costs[0] + costs[1])
            if cl is None:
                raise NullPointer("predClassProb.getLeft() == null")
            pos_prob = pr if cl == class_values[0] else 1.0 - pr
            return class_values[0] if pos_prob > thr else class_values[1]
        return cl

    out = []
    for ln in lines:
        items = java_split(ln, delim_regex, False)
        if not ensemble:
            pred = member_predict(models[0], items)
        else:
            votes = JavaHashMap()
            for m, w in zip(models, weights):
                cl = member_predict(m, items)
                key = "null" if cl is None else cl
                cur = votes.get(key)
                votes.put(key, w if cur is None else cur + w)
            ranked = sorted(list(votes.items()), key=lambda kv: -kv[1])    # stable, like Collections.sort
            pred = ranked[0][0]
            if error_counting:
                raise NullPointer("EnsemblePredictiveModel.countError: items is never set")
        if error_counting:
            total += 1
            if items[class_ord] != pred:
                if pred == class_values[0]:
                    fpos += 1
                else:
                    fneg += 1
                errors += 1
        ps = "null" if pred is None else pred
        if mode == "withRecord":
            out.append(ln + delim_out + ps)
        elif mode == "withKId":
            out.append(items[int(g("mop.rec.id.ordinal"))] + delim_out + ps)
        elif mode == "withActualClassAttr":
            out.append(items[int(g("mop.rec.class.attr.ordinal"))] + delim_out + ps)
        else:
            raise ValueError("invalid output mode")
    err_txt = None
    if error_counting:
        err_txt = "total error: %.3f\nfalse positive error: %.3f\nfalse negative error: %.3f" % (
            errors / total, fpos / total, fneg / total)
    return out, err_txt
