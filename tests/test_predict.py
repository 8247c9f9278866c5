"""ModelPredictor (batch prediction / ensemble vote) through the product against the literal
restatement (oracle/predict_literal.py).  Each case runs twice: on the CPU with the kernel layer mocked at link
time (host logic: mask construction, encoding, output), and -- marked gpu -- through the real library, where
avt_predict's kernel does the matching and the vote."""
import json
import os

import pytest

from avenir_b200 import synth
from oracle import literal
from oracle.predict_literal import model_predictor
from tests import helpers

DTB = "org.avenir.tree.DecisionTreeBuilder"
MOP = "org.avenir.model.ModelPredictor"


def _props(path, d):
    with open(path, "w") as f:
        for k, v in d.items():
            f.write(f"{k}={v}\n")


@pytest.fixture(params=["mock", pytest.param("gpu", marks=pytest.mark.gpu)])
def lib(request):
    if request.param == "mock":
        return helpers.hostmock_lib()
    from avenir_b200 import _lib
    return _lib.load()


def train(lib, d, lines, levels, extra, name):
    """Trains one tree with the product's DecisionTreeBuilder and returns the path of its last level's list."""
    os.makedirs(d, exist_ok=True)
    sp = os.path.join(d, "call_hangup.json")
    schema = helpers.write_schema(sp, "call_hangup.json")
    data = os.path.join(d, f"train_{name}.txt")
    open(data, "w").write("\n".join(lines) + "\n")
    conf = {"dtb.feature.schema.file.path": sp, "dtb.decision.file.path.out": os.path.join(d, f"{name}.json"),
            "dtb.split.attribute.selection.strategy": "all", "dtb.sub.sampling.strategy": "none",
            "dtb.path.stopping.strategy": "maxDepth", "dtb.max.depth.limit": "3"}
    conf.update(extra)
    pp = os.path.join(d, f"{name}.properties")
    _props(pp, conf)
    rc, err = helpers.tool_run(lib, DTB, [f"-Dconf.path={pp}", f"-Davt.dtb.num.levels={levels + 1}", data, os.path.join(d, f"out_{name}")])
    assert rc == 0, err
    return sp, schema, f"{name}.json"


def run_both(lib, d, schema_path, schema, model_names, rows, conf):
    conf = dict(conf, **{"mop.feature.schema.file.path": schema_path, "mop.model.dir.path": d,
                         "mop.model.file.names": ",".join(model_names), "mop.classifier.type": "decTreeClassifier"})
    pp = os.path.join(d, "mop.properties")
    _props(pp, conf)
    data = os.path.join(d, "score.txt")
    open(data, "w").write("\n".join(rows) + "\n")
    rc, err = helpers.tool_run(lib, MOP, [f"-Dconf.path={pp}", data, os.path.join(d, "scored")])
    models = [json.load(open(os.path.join(d, m))) for m in model_names]
    want, want_err = model_predictor(rows, literal.load_properties(pp), literal.FeatureSchema(schema), models)
    assert rc == 0, err
    got = open(os.path.join(d, "scored", "part-r-00000")).read().splitlines()
    return got, want, want_err


def test_single_tree_with_error_counting(lib, tmp_path):
    d = str(tmp_path)
    sp, schema, m = train(lib, d, synth.call_hangup_csv(2500), 3, {"dtb.split.algorithm": "entropy"}, "tree")
    rows = synth.call_hangup_csv(1800, start=50_000)
    conf = {"mop.error.counting.enabled": "true", "mop.class.attr.ord": "6", "mop.class.attr.values": "T,F",
            "map.error.rate.file.path": os.path.join(d, "error.txt")}
    got, want, want_err = run_both(lib, d, sp, schema, [m], rows, conf)
    assert got == want and len(got) == 1800
    assert open(os.path.join(d, "error.txt")).read() == want_err
    assert {ln.rsplit(",", 1)[1] for ln in got} == {"T", "F"}          # the tree does discriminate
    assert float(want_err.split("\n")[0].split(": ")[1]) < 0.35


def test_cost_based_prediction_and_output_modes(lib, tmp_path):
    d = str(tmp_path)
    sp, schema, m = train(lib, d, synth.call_hangup_csv(2000), 4, {}, "tree")
    rows = synth.call_hangup_csv(900, start=90_000)
    base = {"mop.cost.based.prediction.enabled": "true", "mop.class.attr.values": "T,F", "mop.miss.class.costs": "1.0,4.0",
            "map.error.rate.file.path": os.path.join(d, "error.txt")}
    plain, plain_want, _ = run_both(lib, d, sp, schema, [m], rows, {})
    got, want, _ = run_both(lib, d, sp, schema, [m], rows, base)
    assert got == want and plain == plain_want
    assert sum(a != b for a, b in zip(got, plain)) > 0                  # the cost threshold (0.2) moves predictions towards T
    got, want, _ = run_both(lib, d, sp, schema, [m], rows, dict(base, **{"mop.output.mode": "withKId", "mop.rec.id.ordinal": "0"}))
    assert got == want and got[0].count(",") == 1
    got, want, _ = run_both(lib, d, sp, schema, [m], rows, {"mop.output.mode": "withActualClassAttr", "mop.rec.class.attr.ordinal": "6"})
    assert got == want


def test_forest_vote(lib, tmp_path):
    """Three bagged trees (the 'forest' at inference time, model/EnsemblePredictiveModel.java), unequal weights."""
    d = str(tmp_path)
    lines = synth.call_hangup_csv(1500)
    names = []
    for t in range(3):
        sp, schema, m = train(lib, d, lines, 3, {"dtb.sub.sampling.strategy": "withReplace", "dtb.sub.sampling.buffer.size": "400",
                                                 "avt.random.seed": str(11 + t), "dtb.split.algorithm": "giniIndex" if t else "entropy"}, f"tree{t}")
        names.append(m)
    rows = synth.call_hangup_csv(1200, start=70_000)
    got, want, _ = run_both(lib, d, sp, schema, names, rows, {})
    assert got == want
    got_w, want_w, _ = run_both(lib, d, sp, schema, names, rows, {"mop.ensemble.memeber.weights": "0.2,0.5,0.45"})
    assert got_w == want_w
    singles = [run_both(lib, d, sp, schema, [n], rows, {})[0] for n in names]
    assert any(len({s[i] for s in singles}) > 1 for i in range(len(rows)))   # the members do disagree somewhere


def test_reference_failures(lib, tmp_path):
    d = str(tmp_path)
    sp, schema, m = train(lib, d, synth.call_hangup_csv(600), 2, {}, "tree")
    rows = synth.call_hangup_csv(20, start=1000)
    data = os.path.join(d, "score.txt")
    base = {"mop.feature.schema.file.path": sp, "mop.model.dir.path": d, "mop.classifier.type": "decTreeClassifier"}

    def run(extra, lines=rows):
        _props(os.path.join(d, "f.properties"), dict(base, **extra))
        open(data, "w").write("\n".join(lines) + "\n")
        return helpers.tool_run(lib, MOP, [f"-Dconf.path={d}/f.properties", data, os.path.join(d, "scored")])

    rc, err = run({"mop.model.file.names": f"{m},{m}"})
    assert rc == 1 and "odd number of models" in err
    rc, err = run({"mop.model.file.names": m, "mop.classifier.type": "svm"})
    assert rc == 1 and "invalid classifier type" in err
    rc, err = run({"mop.model.file.names": m, "mop.output.mode": "bogus"})
    assert rc == 1 and "invalid output mode" in err
    bad = [rows[0].replace("internet", "fax").replace("cable", "fax").replace("billing", "fax").replace("other", "fax")]
    rc, err = run({"mop.model.file.names": m}, bad)
    model = json.load(open(os.path.join(d, m)))
    uses_issue = any(p["attribute"] == 3 for path in model["decisionPaths"] for p in path["predicates"][1:])
    if uses_issue:          # an issue value no path admits: decPathMatched stays null
        assert rc == 1 and "NullPointerException" in err
    rc, err = run({"mop.model.file.names": f"{m},{m},{m}", "mop.error.counting.enabled": "true", "mop.class.attr.ord": "6",
                   "mop.class.attr.values": "T,F", "map.error.rate.file.path": os.path.join(d, "e.txt")})
    assert rc == 1 and "NullPointerException" in err
