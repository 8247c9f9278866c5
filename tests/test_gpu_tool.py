"""GPU parity, tool level: the reference's entry points (DecisionTreeBuilder / ClassPartitionGenerator,
.properties + schema JSON + CSV in, DecisionPathList JSON / split lines out) through the real
libavenir_tree.so, against the literal restatement of the reference."""
import json
import os

import pytest

from avenir_b200 import synth
from tests import helpers
from tests.test_host_vs_literal import BASE, cpg_rows, cpg_schema, run_cpg_both, run_dtb_both

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lib():
    from avenir_b200 import _lib
    return _lib.load()


@pytest.mark.parametrize("n,levels,strategy,alg", [(3000, 2, "notUsedYet", "giniIndex"),     # config C1's knobs (detr.properties)
                                                   (2000, 5, "all", "entropy"), (300, 7, "all", "giniIndex")])
def test_call_hangup_tree_equals_reference_restatement(lib, tmp_path, n, levels, strategy, alg):
    schema = json.load(open(os.path.join(helpers.SCHEMA_DIR, "schemas.json")))["call_hangup.json"]
    conf = dict(BASE, **{"dtb.split.attribute.selection.strategy": strategy, "dtb.split.algorithm": alg})
    got, want = run_dtb_both(lib, str(tmp_path), synth.call_hangup_csv(n), schema, conf, levels)
    assert len(got) == levels + 1
    for g, w in zip(got, want):
        # structure, counts, chosen splits: exact;  gain values: north_star allows 1e-9 relative
        assert len(g["decisionPaths"]) == len(w["decisionPaths"])
        for pg, pw in zip(g["decisionPaths"], w["decisionPaths"]):
            assert pg["predicates"] == pw["predicates"] and pg["population"] == pw["population"] and pg["stopped"] == pw["stopped"]
            assert abs(pg["infoContent"] - pw["infoContent"]) <= 1e-9 * max(1e-300, abs(pw["infoContent"]))
            assert pg["classValPr"] == pw["classValPr"] and pg["outputClassVal"] == pw["outputClassVal"]
    assert got == want      # and in fact bit-identical


def test_forest_batch_equals_trees_built_one_at_a_time(lib, tmp_path):
    """avt.forest.num.trees=3 (one launch per level for all trees) gives, per tree, the tree a single-tree
    run with that tree's seed gives."""
    schema = json.load(open(os.path.join(helpers.SCHEMA_DIR, "schemas.json")))["call_hangup.json"]
    lines = synth.call_hangup_csv(1200)
    conf = dict(BASE, **{"dtb.sub.sampling.strategy": "withReplace", "dtb.sub.sampling.buffer.size": "500",
                         "dtb.split.attribute.selection.strategy": "all"})
    d = tmp_path / "forest"
    os.makedirs(d)
    json.dump(schema, open(d / "schema.json", "w"))
    (d / "in.txt").write_text("\n".join(lines) + "\n")
    base = dict(conf, **{"dtb.feature.schema.file.path": str(d / "schema.json"), "dtb.decision.file.path.out": str(d / "dec.json")})
    with open(d / "p.properties", "w") as f:
        for k, v in base.items():
            f.write(f"{k}={v}\n")
    rc, err = helpers.tool_run(lib, "DecisionTreeBuilder", [f"-Dconf.path={d / 'p.properties'}", "-Davt.dtb.num.levels=4", "-Davt.forest.num.trees=3",
                                                            "-Davt.random.seed=5", str(d / "in.txt"), str(d / "out")])
    assert rc == 0, err
    for t in range(3):
        got, want = run_dtb_both(lib, str(tmp_path / f"single{t}"), lines, schema, dict(conf, **{"avt.random.seed": str(5 + t)}), 3)
        assert got == want
        assert json.load(open(d / f"dec.json.tree{t}")) == want[-1]


@pytest.mark.parametrize("alg", ["giniIndex", "entropy", "hellingerDistance", "classConfidenceRatio"])
def test_cpg_lines_equal_reference_restatement(lib, tmp_path, alg):
    conf = {"cpg.split.attributes": "2,1,3", "cpg.split.algorithm": alg, "cpg.parent.info": "0.97", "cpg.output.split.prob": "true"}
    got, want = run_cpg_both(lib, str(tmp_path), cpg_rows(600), cpg_schema(), conf)
    assert got == want
