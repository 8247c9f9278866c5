"""CategoricalContinuousEncoding (a count consumer of explore/ on the histogram kernel) through the product
against the literal restatement (oracle/cce_literal.py), on the CPU with the kernel layer mocked at link time and -- marked gpu --
through the real library (device CSV encode + root histogram).  The order of the output lines is unpinned in the reference
(a HashMap keyed by chombo Tuples): lines are compared as sets."""
import os

import numpy as np
import pytest

from oracle import cce_literal
from tests import helpers

CCE = "org.avenir.explore.CategoricalContinuousEncoding"


@pytest.fixture(params=["mock", pytest.param("gpu", marks=pytest.mark.gpu)])
def lib(request):
    if request.param == "mock":
        return helpers.hostmock_lib()
    from avenir_b200 import _lib
    return _lib.load()


def rows(n, seed, n_big=400, classes=("Y", "N", "M")):
    rng = np.random.default_rng(seed)
    out = []
    for i in range(n):
        a = int(rng.zipf(1.3)) % n_big
        b = "ABCDE"[int(rng.integers(5))]
        y = classes[0] if rng.random() < 0.2 + 0.5 * (a % 3 == 0) else classes[1 + int(rng.integers(len(classes) - 1))]
        out.append(f"{i},c{a:04d},{b},{int(rng.integers(100))},{y}")
    return out


def run_both(lib, d, lines, conf):
    os.makedirs(d, exist_ok=True)
    pp = os.path.join(d, "coe.properties")
    with open(pp, "w") as f:
        for k, v in conf.items():
            f.write(f"{k}={v}\n")
    data = os.path.join(d, "in.txt")
    open(data, "w").write("\n".join(lines) + "\n")
    rc, err = helpers.tool_run(lib, CCE, [f"-Dconf.path={pp}", data, os.path.join(d, "out")])
    got = open(os.path.join(d, "out", "part-r-00000")).read().splitlines() if rc == 0 else None
    return rc, err, got


BASE = {"coe.cat.attribute.ordinals": "1,2", "coe.class.attr.ordinal": "4", "coe.pos.class.attr.value": "Y", "coe.output.scale": "1000"}


@pytest.mark.parametrize("strategy", ["supervisedRatio", "weightOfEvidence"])
def test_codes_equal_literal(lib, tmp_path, strategy):
    lines = rows(20_000, 3)
    conf = dict(BASE, **{"coe.encoding.strategy": strategy})
    rc, err, got = run_both(lib, str(tmp_path), lines, conf)
    assert rc == 0, err
    want = cce_literal.categorical_continuous_encoding(lines, conf)
    assert sorted(got) == sorted(want) and len(got) == len(set(got)) > 100
    assert any(ln.startswith("2,") for ln in got)


def test_high_cardinality_values_never_positive_and_other_delimiters(lib, tmp_path):
    """More than 256 values (uint16 codes), values whose rows are never positive (supervised ratio 0, weight of evidence
    log(0) -> Integer.MIN_VALUE), values whose rows are always positive (neg counted as 1), ';' in, '|' out."""
    lines = [ln.replace(",", ";") for ln in rows(30_000, 5, n_big=3000)]
    for strategy in ("supervisedRatio", "weightOfEvidence"):
        conf = dict(BASE, **{"coe.encoding.strategy": strategy, "field.delim.regex": ";", "field.delim": "|", "coe.cat.attribute.ordinals": "1"})
        rc, err, got = run_both(lib, str(tmp_path / strategy), lines, conf)
        assert rc == 0, err
        want = cce_literal.categorical_continuous_encoding(lines, conf)
        assert sorted(got) == sorted(want) and len(got) > 300
        if strategy == "weightOfEvidence":
            assert any(ln.endswith("|-2147483648") for ln in got)


def test_positive_value_absent_int_wraparound_and_duplicate_ordinal(lib, tmp_path):
    lines = rows(5_000, 7)
    # no row is positive: supervised ratio 0 everywhere, weight of evidence 0/0 = NaN -> 0
    for strategy in ("supervisedRatio", "weightOfEvidence"):
        conf = dict(BASE, **{"coe.encoding.strategy": strategy, "coe.pos.class.attr.value": "absent"})
        rc, err, got = run_both(lib, str(tmp_path / ("a" + strategy)), lines, conf)
        assert rc == 0, err
        assert sorted(got) == sorted(cce_literal.categorical_continuous_encoding(lines, conf))
    # pos * scale wraps in int arithmetic; an ordinal listed twice is counted twice
    conf = dict(BASE, **{"coe.encoding.strategy": "supervisedRatio", "coe.output.scale": "2000000000", "coe.cat.attribute.ordinals": "2,2"})
    rc, err, got = run_both(lib, str(tmp_path / "wrap"), lines, conf)
    assert rc == 0, err
    assert sorted(got) == sorted(cce_literal.categorical_continuous_encoding(lines, conf))


@pytest.mark.parametrize("drop,what", [("coe.cat.attribute.ordinals", "missing categorical attribute ordinals"), ("coe.encoding.strategy", "missing encoding strategy"),
                                       ("coe.class.attr.ordinal", "missing class atrribute ordinal"), ("coe.pos.class.attr.value", "missing positive class attribute value"),
                                       ("coe.output.scale", "missing output scale")])
def test_missing_parameters_and_short_rows(lib, tmp_path, drop, what):
    conf = dict(BASE, **{"coe.encoding.strategy": "supervisedRatio"})
    del conf[drop]
    rc, err, got = run_both(lib, str(tmp_path), rows(50, 1), conf)
    assert rc == 1 and what in err
    with pytest.raises(cce_literal.JobFailure):
        cce_literal.categorical_continuous_encoding(rows(50, 1), conf)
    full = dict(BASE, **{"coe.encoding.strategy": "supervisedRatio"})
    rc, err, got = run_both(lib, str(tmp_path / "short"), rows(10, 1) + ["7,c0001"], full)
    assert rc == 1 and "ArrayIndexOutOfBoundsException" in err
