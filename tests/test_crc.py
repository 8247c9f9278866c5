"""CramerCorrelation (a count consumer of explore/ on the histogram kernel) through the product against
the literal restatement (oracle/crc_literal.py).  Each case runs on the CPU with the kernel layer mocked at link time
(host logic: pair bookkeeping, matrix extraction, cramerIndex, Double.toString, output order) and -- marked gpu --
through the real library (device CSV encode + root histogram)."""
import json
import os

import numpy as np
import pytest

from avenir_b200 import synth
from oracle import crc_literal, literal
from tests import helpers

CRC = "org.avenir.explore.CramerCorrelation"


@pytest.fixture(params=["mock", pytest.param("gpu", marks=pytest.mark.gpu)])
def lib(request):
    if request.param == "mock":
        return helpers.hostmock_lib()
    from avenir_b200 import _lib
    return _lib.load()


def run_both(lib, d, schema: dict, lines, conf):
    os.makedirs(d, exist_ok=True)
    sp = os.path.join(d, "schema.json")
    json.dump(schema, open(sp, "w"))
    conf = dict(conf, **{"crc.feature.schema.file.path": sp})
    pp = os.path.join(d, "crc.properties")
    with open(pp, "w") as f:
        for k, v in conf.items():
            f.write(f"{k}={v}\n")
    data = os.path.join(d, "in.txt")
    open(data, "w").write("\n".join(lines) + "\n")
    rc, err = helpers.tool_run(lib, CRC, [f"-Dconf.path={pp}", data, os.path.join(d, "out")])
    got = open(os.path.join(d, "out", "part-r-00000")).read().splitlines() if rc == 0 else None
    return rc, err, got, literal.load_properties(pp), literal.FeatureSchema(schema)


def hangup_schema():
    d = json.load(open(os.path.join(helpers.SCHEMA_DIR, "schemas.json")))["call_hangup.json"]
    d = json.loads(json.dumps(d))
    for f in d["fields"]:
        if f["name"] == "hungup":
            f["cardinality"] = ["T", "F"]
    return d


def test_call_hangup_pairs(lib, tmp_path):
    schema = hangup_schema()
    lines = synth.call_hangup_csv(6000)
    rc, err, got, conf, fs = run_both(lib, str(tmp_path), schema, lines, {"crc.source.attributes": "3,1", "crc.dest.attributes": "6,4"})
    assert rc == 0, err
    want = crc_literal.cramer_correlation(lines, conf, fs)
    assert got == want and len(got) == 4
    assert [ln.split(",")[:2] for ln in got] == [["customer type", "time of day"], ["customer type", "hungup"], ["issue", "time of day"], ["issue", "hungup"]]
    vals = {tuple(ln.split(",")[:2]): float(ln.split(",")[2]) for ln in got}
    assert all(-1e-9 < v < 1.0 for v in vals.values())


def test_wide_source_attribute_sparse_cells_and_output_delimiter(lib, tmp_path):
    """A 300-value source attribute (uint16 codes), values that never occur (empty rows count as 1 in the
    denominators), a 16-value destination, field.delim.out."""
    rng = np.random.default_rng(8)
    big = ["v%03d" % i for i in range(300)]
    dest = ["d%02d" % i for i in range(16)]
    schema = {"fields": [{"name": "id", "ordinal": 0, "id": True, "dataType": "string"},
                         {"name": "big", "ordinal": 1, "dataType": "categorical", "feature": True, "cardinality": big},
                         {"name": "small", "ordinal": 2, "dataType": "categorical", "feature": True, "cardinality": ["x", "y", "z"]},
                         {"name": "dest", "ordinal": 4, "dataType": "categorical", "cardinality": dest}]}
    lines = []
    for r in range(5000):
        b = int(rng.integers(0, 250))                        # values 250..299 never occur
        lines.append("%d,%s,%s,pad,%s" % (r, big[b], "xyz"[int(rng.integers(3))], dest[(b + int(rng.integers(0, 3))) % 16]))
    rc, err, got, conf, fs = run_both(lib, str(tmp_path), schema, lines,
                                      {"crc.source.attributes": "1,2", "crc.dest.attributes": "4", "field.delim.out": ";"})
    assert rc == 0, err
    assert got == crc_literal.cramer_correlation(lines, conf, fs) and got[0].startswith("big;dest;")
    assert float(got[0].split(";")[2]) > 10 * abs(float(got[1].split(";")[2]))      # `big` determines `dest`, `small` does not


def test_reference_failures(lib, tmp_path):
    schema = hangup_schema()
    lines = synth.call_hangup_csv(50)
    base = {"crc.source.attributes": "1,3", "crc.dest.attributes": "4"}
    bad = list(lines)
    bad[17] = bad[17].replace("AM", "noon").replace("PM", "noon")
    rc, err, _, conf, fs = run_both(lib, str(tmp_path / "a"), schema, bad, base)
    assert rc == 1 and "ArrayIndexOutOfBoundsException" in err and "row 17" in err
    with pytest.raises(crc_literal.ArrayIndexOutOfBounds):
        crc_literal.cramer_correlation(bad, conf, fs)
    short = list(lines)
    short[3] = ",".join(short[3].split(",")[:4])
    rc, err, _, conf, fs = run_both(lib, str(tmp_path / "b"), schema, short, base)
    assert rc == 1 and "ArrayIndexOutOfBoundsException" in err
    rc, err, _, _, _ = run_both(lib, str(tmp_path / "c"), schema, lines, {"crc.source.attributes": "1,4", "crc.dest.attributes": "4,6"})
    assert rc == 1 and "unsupported" in err and "overlap" in err
    rc, err, _, _, _ = run_both(lib, str(tmp_path / "d"), schema, lines, {"crc.source.attributes": "1"})
    assert rc == 1 and "crc.dest.attributes" in err
    rc, err, _, _, _ = run_both(lib, str(tmp_path / "e"), schema, lines, {"crc.source.attributes": "1", "crc.dest.attributes": "9"})
    assert rc == 1 and "NullPointerException" in err


HRC = "org.avenir.explore.HeterogeneityReductionCorrelation"


def run_hrc(lib, d, schema, lines, conf):
    os.makedirs(d, exist_ok=True)
    sp = os.path.join(d, "schema.json")
    json.dump(schema, open(sp, "w"))
    conf = dict(conf, **{"cac.feature.schema.file.path": sp})
    pp = os.path.join(d, "hrc.properties")
    with open(pp, "w") as f:
        for k, v in conf.items():
            f.write(f"{k}={v}\n")
    data = os.path.join(d, "in.txt")
    open(data, "w").write("\n".join(lines) + "\n")
    rc, err = helpers.tool_run(lib, HRC, [f"-Dconf.path={pp}", data, os.path.join(d, "out")])
    got = open(os.path.join(d, "out", "part-r-00000")).read().splitlines() if rc == 0 else None
    return rc, err, got, literal.load_properties(pp), literal.FeatureSchema(schema)


def test_heterogeneity_reduction_correlation(lib, tmp_path):
    """Same contingency matrices, concentration ("gini") and uncertainty coefficient; the reference's key with two dots."""
    schema = hangup_schema()
    lines = synth.call_hangup_csv(4000)
    keys = {"cac.first.set.attributes": "1,3", "cac.second.set..attributes": "4,6"}
    rc, err, got, conf, fs = run_hrc(lib, str(tmp_path / "g"), schema, lines, keys)
    assert rc == 0, err
    assert got == crc_literal.heterogeneity_reduction_correlation(lines, conf, fs) and len(got) == 4
    assert all(0.0 <= float(ln.rsplit(",", 1)[1]) < 1.0 for ln in got)
    rc, err, got, conf, fs = run_hrc(lib, str(tmp_path / "u"), schema, lines, dict(keys, **{"hrc.heterogeneity.algorithm": "uncertainty"}))
    assert rc == 0, err
    want = crc_literal.heterogeneity_reduction_correlation(lines, conf, fs)
    assert got == want
    # residence callers never have issue "cable"/... combinations missing?  whichever cells are empty give NaN, the dense pairs a number
    assert any(ln.endswith("NaN") for ln in got) or all(float(ln.rsplit(",", 1)[1]) == float(ln.rsplit(",", 1)[1]) for ln in got)
    rc, err, _, _, _ = run_hrc(lib, str(tmp_path / "m"), schema, lines, {"cac.first.set.attributes": "1"})
    assert rc == 1 and "cac.second.set..attributes" in err


def test_uncertainty_coefficient_with_an_empty_cell_is_nan(lib, tmp_path):
    schema = {"fields": [{"name": "a", "ordinal": 0, "dataType": "categorical", "feature": True, "cardinality": ["p", "q"]},
                         {"name": "b", "ordinal": 1, "dataType": "categorical", "cardinality": ["x", "y"]}]}
    dense = ["p,x", "p,y", "q,x", "q,y", "q,y"]
    rc, err, got, conf, fs = run_hrc(lib, str(tmp_path / "d"), schema, dense, {"cac.first.set.attributes": "0", "cac.second.set..attributes": "1",
                                                                             "hrc.heterogeneity.algorithm": "uncertainty"})
    assert rc == 0, err
    assert got == crc_literal.heterogeneity_reduction_correlation(dense, conf, fs) and not got[0].endswith("NaN")
    sparse = ["p,x", "p,x", "q,y"]
    rc, err, got, conf, fs = run_hrc(lib, str(tmp_path / "s"), schema, sparse, {"cac.first.set.attributes": "0", "cac.second.set..attributes": "1",
                                                                              "hrc.heterogeneity.algorithm": "uncertainty"})
    assert rc == 0, err
    assert got == crc_literal.heterogeneity_reduction_correlation(sparse, conf, fs) == ["a,b,NaN"]
