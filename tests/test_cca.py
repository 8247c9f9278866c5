"""CategoricalClassAffinity (SURVEY 8f row 4 names it; host code only -- its input is another job's distribution file) through the
product against the literal restatement (oracle/cca_literal.py): exact output, line by line, incl. the HashMap order of equal
scores, the reducer's maps living across feature groups, and the reference's failures."""
import os

import numpy as np
import pytest

from oracle import cca_literal
from tests import helpers

CCA = "org.avenir.explore.CategoricalClassAffinity"


@pytest.fixture(params=["mock", pytest.param("gpu", marks=pytest.mark.gpu)])
def lib(request):
    if request.param == "mock":
        return helpers.hostmock_lib()
    from avenir_b200 import _lib
    return _lib.load()


def distribution_lines(seed, features=(1, 3, 4), classes=("Y", "N"), n_values=9, shuffle=True):
    rng = np.random.default_rng(seed)
    lines = []
    for f in features:
        vals = [f"v{f}_{k}" for k in range(int(rng.integers(2, n_values)))]
        for c in classes:
            p = rng.random(len(vals)) + 0.05
            p = p / p.sum()
            for v, x in zip(vals, p):
                lines.append(f"{f},{c},{v},{float(x)!r}")
    if shuffle:
        rng.shuffle(lines)
    return [str(x) for x in lines]


def run_both(lib, d, lines, conf):
    os.makedirs(d, exist_ok=True)
    pp = os.path.join(d, "cca.properties")
    with open(pp, "w") as f:
        for k, v in conf.items():
            f.write(f"{k}={v}\n")
    data = os.path.join(d, "in.txt")
    open(data, "w").write("\n".join(lines) + "\n")
    rc, err = helpers.tool_run(lib, CCA, [f"-Dconf.path={pp}", data, os.path.join(d, "out")])
    got = open(os.path.join(d, "out", "part-r-00000")).read().splitlines() if rc == 0 else None
    return rc, err, got


ALL = "oddsRatio,distrDiff,minRisk,klDiff"


@pytest.mark.parametrize("seed", range(6))
def test_scores_equal_literal(lib, tmp_path, seed):
    lines = distribution_lines(seed, n_values=20)
    conf = {"cca.pos.class.attr.value": "Y", "cca.affinity.strategy": ALL}
    rc, err, got = run_both(lib, str(tmp_path), lines, conf)
    assert rc == 0, err
    want = cca_literal.categorical_class_affinity(lines, conf)
    assert got == want and got[0] == "algorithm: oddsRatio" and sum(1 for ln in got if ln.startswith("algorithm")) == 12


def test_equal_scores_stale_negative_map_and_several_other_classes(lib, tmp_path):
    """Equal probabilities (ties keep the HashMap order of the positive map); feature 5 has only positive lines and is scored against
    feature 3's map; with two non-positive class values only the last run survives; other delimiters."""
    names = ("alpha", "beta", "gamma", "delta", "epsilon", "zeta", "eta", "theta", "iota", "kappa", "lambda", "mu", "nu")
    lines = [f"3;Y;{v};0.25" for v in names] + [f"5;Y;{v};0.5" for v in names] + [f"3;N;{v};0.125" for v in names] + [f"3;M;{v};0.0625" for v in names]
    conf = {"cca.pos.class.attr.value": "Y", "cca.affinity.strategy": "distrDiff,oddsRatio", "field.delim.regex": ";", "field.delim": "|"}
    rc, err, got = run_both(lib, str(tmp_path), lines, conf)
    assert rc == 0, err
    assert got == cca_literal.categorical_class_affinity(lines, conf)
    assert "5|alpha|0.4375" in got and "3|alpha|0.1875" in got      # 0.5 - 0.0625 (feature 3's last class), 0.25 - 0.0625


@pytest.mark.parametrize("lines,conf,what", [
    (["1,Y,a,0.5", "1,N,b,0.5"], {"cca.pos.class.attr.value": "Y", "cca.affinity.strategy": "distrDiff"}, "NullPointerException"),
    (["1,Y,a,0.5", "1,N,a,0.5"], {"cca.pos.class.attr.value": "Y", "cca.affinity.strategy": "nothingKnown"}, "StringIndexOutOfBoundsException"),
    (["1,Y,a"], {"cca.pos.class.attr.value": "Y", "cca.affinity.strategy": "distrDiff"}, "ArrayIndexOutOfBoundsException: 3"),
    (["x,Y,a,0.5"], {"cca.pos.class.attr.value": "Y", "cca.affinity.strategy": "distrDiff"}, "NumberFormatException"),
    (["1,Y,a,zero"], {"cca.pos.class.attr.value": "Y", "cca.affinity.strategy": "distrDiff"}, "NumberFormatException"),
    (["1,Y,a,0.5"], {"cca.affinity.strategy": "distrDiff"}, "missing positive class attribute value"),
    (["1,Y,a,0.5"], {"cca.pos.class.attr.value": "Y"}, "missing affinity strategy list"),
])
def test_reference_failures(lib, tmp_path, lines, conf, what):
    rc, err, got = run_both(lib, str(tmp_path), lines, conf)
    assert rc == 1 and what in err, err
    with pytest.raises(cca_literal.JobFailure) as ei:
        cca_literal.categorical_class_affinity(lines, conf)
    assert what.split(":")[0] in str(ei.value)
