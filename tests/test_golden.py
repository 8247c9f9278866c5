"""The oracle against every result-pinning artefact the reference holds for the tree path
(SURVEY.md section 4 / App. C): the four gini values + probabilities of resource/dec_tree_rules.json,
derived weighted averages, and the split enumerations of the shipped schemas."""
import json
import os

import numpy as np
import pytest

from oracle import literal
from oracle.javaemu import fdlibm_log
from tests import helpers

GOLD = os.path.join(os.path.dirname(__file__), "golden")
KAT = json.load(open(os.path.join(GOLD, "dec_tree_rules_kat.json")))
SCHEMAS = json.load(open(os.path.join(GOLD, "schemas.json")))


def _stat_from_counts(counts, entropy=False):
    st = literal.InfoContentStat()
    for k, v in counts.items():
        st.count_class_val(k, v)
    return st, st.process_stat(entropy)


def test_sample_model_gini_and_probabilities_bit_exact():
    assert [p["infoContent"] for p in KAT["paths"]] == [0.39594381212292684, 0.4284892561983471, 0.49870726802642906, 0.31589923649731366]
    for p in KAT["paths"]:
        st, g = _stat_from_counts(p["classCounts"])
        assert g == p["infoContent"]
        assert [[k, v] for k, v in st.class_val_pr.items()] == p["classValPr"]     # values and T-before-F order
        assert st.total_count == p["population"]


def test_sample_model_gini_c_oracle(oracle):
    order = np.array([0, 1], dtype=np.uint8)
    for p in KAT["paths"]:
        cnt = np.array([p["classCounts"]["T"], p["classCounts"]["F"]], dtype=np.uint32)
        assert oracle.ora_impurity(cnt.ctypes.data, 2, order.ctypes.data, 0) == p["infoContent"]


def test_weighted_averages_appendix_c():
    g = {p["predicateStrs"][-1]: (p["infoContent"], p["population"]) for p in KAT["paths"]}
    a = (g["5 le 420"][0] * 331 + g["5 gt 420"][0] * 550) / 881
    b = (g["5 gt 420"][0] * 550 + g["5 le 420"][0] * 331) / 881
    assert a == b == 0.41626162624492585
    assert (g["5 le 300"][0] * 59 + g["5 gt 300"][0] * 2060) / 2119 == 0.3209892194422017
    assert _stat_from_counts({"T": 469, "F": 412})[1] == 0.4979070064071758
    assert _stat_from_counts({"T": 1683, "F": 436})[1] == 0.3268426232377859
    assert _stat_from_counts({"T": 2152, "F": 848})[1] == 0.40553244444444436


def test_entropy_appendix_c(oracle):
    for counts, want in (({"T": 90, "F": 241}, 0.8441825466830741), ({"T": 379, "F": 171}, 0.8942200620259633)):
        _, h = _stat_from_counts(counts, entropy=True)
        assert abs(h - want) < 1e-12
        cnt = np.array([counts["T"], counts["F"]], dtype=np.uint32)
        order = np.array([0, 1], dtype=np.uint8)
        assert oracle.ora_impurity(cnt.ctypes.data, 2, order.ctypes.data, 1) == h    # C and Python restatements agree bit for bit
    assert oracle.ora_strict_log(0.2719033232628399) == fdlibm_log(0.2719033232628399)


def _sm(name):
    return literal.SplitManager(literal.FeatureSchema(SCHEMAS[name]), ";")


def test_split_enumeration_call_hangup():
    sm = _sm("call_hangup.json")
    hold = [[str(p) for p in sp] for sp in sm.split_predicates(5)]
    assert hold == [[f"5 le {k}", f"5 gt {k}"] for k in range(60, 600, 60)]
    issue = ["|".join(str(p) for p in sp) for sp in sm.split_predicates(3)]
    assert issue == ["3 in internet:billing:other|3 in cable", "3 in internet:billing|3 in cable:other",
                     "3 in internet:other|3 in cable:billing", "3 in internet|3 in cable:billing:other",
                     "3 in internet:cable:other|3 in billing", "3 in internet:cable|3 in billing:other",
                     "3 in internet:cable:billing|3 in other"]
    assert sum(len(sm.split_predicates(a)) for a in (1, 3, 4, 5)) == 18
    # the sample model's predicates are among the enumerated ones
    enumerated = {str(p) for a in (1, 3, 4, 5) for sp in sm.split_predicates(a) for p in sp}
    for p in KAT["paths"]:
        for s in p["predicateStrs"][1:]:
            if s.split()[2] in ("420", "300") or s.startswith("3 in"):
                assert s in enumerated


def test_split_enumeration_other_schemas():
    sm = _sm("teleComChurn.json")
    schema = literal.FeatureSchema(SCHEMAS["teleComChurn.json"])
    by_name = {f.name: f for f in schema.fields}
    assert [p[0] for p in sm.int_split_points(by_name["minUsed"])] == list(range(200, 2200, 200))
    assert [p[0] for p in sm.int_split_points(by_name["csCall"])] == list(range(2, 14, 2))
    assert [p[0] for p in sm.int_split_points(by_name["csEmail"])] == list(range(4, 24, 4))
    f = literal.Field({"ordinal": 1, "dataType": "int", "min": 0, "max": 300, "splitScanInterval": 60, "maxSplit": 3})
    assert sm.int_split_points(f) == [[60], [60, 120], [60, 180], [60, 240], [120], [120, 180], [120, 240], [180], [180, 240], [240]]
    # truncating compound assignment (App. B-8)
    f2 = literal.Field({"ordinal": 1, "dataType": "int", "min": 0, "max": 10, "splitScanInterval": 2.5})
    assert [p[0] for p in sm.int_split_points(f2)] == [2, 4, 6, 8]
    assert len(sm.categorical_partitions(list("abcd"), 2)) == 7
    assert len(sm.categorical_partitions(list("abcde"), 2)) == 15
    assert len(sm.categorical_partitions(list("abcdefghij"), 2)) == 511
    assert len(sm.categorical_partitions(list("abcd"), 3)) == 13


def test_cpg_enumeration():
    schema = literal.FeatureSchema(SCHEMAS["patient.json"])
    age = [f for f in schema.fields if f.name == "age"][0]
    parts = literal.cpg_create_partitions(schema, [age.ordinal], 3)[age.ordinal]
    assert len(parts) == 66
    assert [p.key for p in parts[:3]] == ["25", "25;30", "25;35"] and parts[-1].key == "75"
    es = literal.FeatureSchema(SCHEMAS["emailCampaign.json"])
    camp = [f for f in es.fields if f.is_categorical() and len(f.cardinality) == 9][0]
    sm = literal.SplitManager(es, ";")
    assert len(sm.categorical_partitions(camp.cardinality, 2)) == 255
    assert len(sm.categorical_partitions(camp.cardinality, 3)) == 3280


def test_model_json_shape_matches_sample(tmp_path, hostmock):
    """The product's DecisionPathList writer emits the sample model's fields (plus the three fields the
    current bean adds, tree/DecisionPathList.java:127-129)."""
    from avenir_b200 import synth
    schema = tmp_path / "call_hangup.json"
    helpers.write_schema(str(schema), "call_hangup.json")
    data = tmp_path / "in.txt"
    data.write_text("\n".join(synth.call_hangup_csv(400)) + "\n")
    props = tmp_path / "detr.properties"
    props.write_text(f"dtb.feature.schema.file.path={schema}\ndtb.decision.file.path.out={tmp_path}/dec.json\n"
                     "dtb.sub.sampling.strategy=none\ndtb.path.stopping.strategy=maxDepth\ndtb.max.depth.limit=2\n")
    rc, err = helpers.tool_run(hostmock, "DecisionTreeBuilder", [f"-Dconf.path={props}", "-Davt.dtb.num.levels=3", str(data), str(tmp_path / "out")])
    assert rc == 0, err
    model = json.load(open(tmp_path / "dec.json"))
    sample = KAT["paths"][0]
    for p in model["decisionPaths"]:
        assert set(sample["pathFields"]) <= set(p.keys())
        assert set(p.keys()) - set(sample["pathFields"]) == {"outputClassVal", "confidence", "support"}
        assert sorted(p["predicates"][-1].keys()) == sample["predicateFields"]
        assert p["predicates"][0] == {"attribute": 0, "predicateStr": "$root", "operator": None, "valueInt": 0, "valueDbl": 0.0,
                                      "categoricalValues": None, "otherBoundInt": None, "otherBoundDbl": None}
