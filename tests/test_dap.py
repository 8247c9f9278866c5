"""DataPartitioner (the older node-wise pipeline SplitGenerator -> DataPartitioner) through the product
against the literal restatement (oracle/dap_literal.py).  Each case runs on the CPU with the kernel layer mocked at link
time (host logic: candidate lines, sort, key parsing, reducer layout, files) and -- marked gpu -- through the real library
(device CSV encode + device row scatter, avt_rows_partition).  The kernel-level call is also checked on its own."""
import json
import os

import numpy as np
import pytest

from avenir_b200 import synth
from oracle import dap_literal
from tests import helpers

DAP = "org.avenir.tree.DataPartitioner"
SPG = "org.avenir.tree.SplitGenerator"


@pytest.fixture(params=["mock", pytest.param("gpu", marks=pytest.mark.gpu)])
def lib(request):
    if request.param == "mock":
        return helpers.hostmock_lib()
    from avenir_b200 import _lib
    return _lib.load()


def hangup_schema():
    d = json.load(open(os.path.join(helpers.SCHEMA_DIR, "schemas.json")))["call_hangup.json"]
    return json.loads(json.dumps(d))


def layout(d, schema, lines, split_lines, conf=None, split_path=""):
    """The project directory DataPartitioner expects: <base>/split=root/data[/<split path>]/<rows>, sibling splits/part-r-00000."""
    base = os.path.join(d, "proj")
    node = os.path.join(base, "split=root", "data", split_path) if split_path else os.path.join(base, "split=root", "data")
    os.makedirs(node, exist_ok=True)
    open(os.path.join(node, "rows.txt"), "w").write("".join(ln + "\n" for ln in lines))
    sdir = os.path.join(os.path.dirname(node), "splits")
    os.makedirs(sdir, exist_ok=True)
    open(os.path.join(sdir, "part-r-00000"), "w").write("".join(ln + "\n" for ln in split_lines))
    sp = os.path.join(d, "schema.json")
    json.dump(schema, open(sp, "w"))
    c = {"dap.project.base.path": base, "dap.feature.schema.file.path": sp}
    if split_path:
        c["dap.split.path"] = split_path
    c.update(conf or {})
    pp = os.path.join(d, "dap.properties")
    with open(pp, "w") as f:
        for k, v in c.items():
            f.write(f"{k}={v}\n")
    return pp, c, node


def written(node, rel_files):
    return {rel: open(os.path.join(node, rel)).read() for rel in rel_files}


def run_and_compare(lib, d, schema, lines, split_lines, conf=None, split_path="", seed=None):
    pp, c, node = layout(d, schema, lines, split_lines, conf, split_path)
    args = [f"-Dconf.path={pp}"] + ([f"-Davt.random.seed={seed}"] if seed is not None else [])
    rc, err = helpers.tool_run(lib, DAP, args)
    assert rc == 0, err
    want = dap_literal.dap_run(c, seed)
    assert written(node, want.keys()) == want
    return want


def test_categorical_split_best_line(lib, tmp_path):
    """Three candidate lines; the best statistic is not the first line, so the output directory carries its line index;
    three value groups = three reducers = three segment directories."""
    lines = synth.call_hangup_csv(5000)
    split_lines = ["3;[cable, internet]:[billing, other];0.01", "3;[cable]:[internet, billing]:[other];0.07", "1;[business]:[residence];0.03"]
    want = run_and_compare(lib, str(tmp_path), hangup_schema(), lines, split_lines)
    assert sorted(want) == ["split=1/segment=%d/data/partition.txt" % i for i in range(3)]
    assert sum(v.count("\n") for v in want.values()) == 5000 and all(v for v in want.values())


def test_integer_split_has_one_reducer(lib, tmp_path):
    """App. B-12: an integer split's key has no ':' -> one reducer; both sides land in segment=0, rows <= point first."""
    lines = synth.call_hangup_csv(3000)
    want = run_and_compare(lib, str(tmp_path), hangup_schema(), lines, ["5;500;0.2"])
    assert list(want) == ["split=0/segment=0/data/partition.txt"]
    rows = want["split=0/segment=0/data/partition.txt"].splitlines()
    hold = [int(r.split(",")[5]) for r in rows]
    k = sum(1 for h in hold if h <= 500)
    assert 0 < k < 3000 and all(h <= 500 for h in hold[:k]) and all(h > 500 for h in hold[k:])


def test_colon_joined_points_and_key_mod_reducers(lib, tmp_path):
    """A key '450:560' gives two reducers for three keys: key 2 joins key 0 in reducer 0, behind it."""
    lines = synth.call_hangup_csv(4000)
    want = run_and_compare(lib, str(tmp_path), hangup_schema(), lines, ["5;450:560;0.5", "5;500;0.9"])
    # the second line has the larger statistic (0.9): it wins, one reducer
    assert list(want) == ["split=1/segment=0/data/partition.txt"]
    want = run_and_compare(lib, str(tmp_path / "b"), hangup_schema(), lines, ["5;450:560;0.5", "5;500;0.1"])
    assert sorted(want) == ["split=0/segment=0/data/partition.txt", "split=0/segment=1/data/partition.txt"]
    seg0 = [int(r.split(",")[5]) for r in want["split=0/segment=0/data/partition.txt"].splitlines()]
    assert any(h <= 450 for h in seg0) and any(h > 560 for h in seg0) and not any(450 < h <= 560 for h in seg0)
    k = sum(1 for h in seg0 if h <= 450)
    assert all(h <= 450 for h in seg0[:k])


def test_nested_node_and_stable_ties_and_random_from_top(lib, tmp_path):
    lines = [ln for ln in synth.call_hangup_csv(2000) if ln.split(",")[3] in ("cable", "internet")]
    split_lines = ["1;[business]:[residence];0.25", "4;[AM]:[PM];0.25", "3;[cable]:[internet];0.1", "6;[T]:[F];0.05"]
    want = run_and_compare(lib, str(tmp_path), hangup_schema(), lines, split_lines, split_path="split=1/segment=0/data")
    assert sorted(want)[0].startswith("split=0/")            # equal statistics keep their file order
    for seed in (1, 2, 5):
        w = run_and_compare(lib, str(tmp_path / f"s{seed}"), hangup_schema(), lines, split_lines, split_path="split=1/segment=0/data",
                            conf={"dap.split.selection.strategy": "randomFromTop", "dap.num.top.splits": "3"}, seed=seed)
        assert len(w) == 2


def test_crlf_rows_other_delimiter_and_a_value_listed_twice(lib, tmp_path):
    schema = {"fields": [{"name": "id", "ordinal": 0, "id": True, "dataType": "string"},
                         {"name": "colour", "ordinal": 2, "dataType": "categorical", "feature": True, "cardinality": ["r", "g", "b", "k"]},
                         {"name": "y", "ordinal": 3, "dataType": "categorical", "cardinality": ["0", "1"]}]}
    rng = np.random.default_rng(4)
    lines = ["%d;x;%s;%d" % (i, "rgbk"[int(rng.integers(4))], int(rng.integers(2))) for i in range(777)]
    pp, c, node = layout(str(tmp_path), schema, lines, ["2;[r, g]:[g, b, k];0.3"], conf={"field.delim.regex": ";"})
    # rewrite the rows with CRLF terminators and no terminator after the last row
    open(os.path.join(node, "rows.txt"), "wb").write("\r\n".join(lines).encode())
    rc, err = helpers.tool_run(lib, DAP, [f"-Dconf.path={pp}"])
    assert rc == 0, err
    want = dap_literal.dap_run(c)
    got = written(node, want.keys())
    assert got == want and "\r" not in "".join(got.values())
    assert all(r.split(";")[2] in "rg" for r in got["split=0/segment=0/data/partition.txt"].splitlines())   # 'g' stays with the first group


def test_split_generator_output_feeds_data_partitioner(lib, tmp_path):
    """The pipeline the two jobs were written for: SplitGenerator writes <node>/../splits with ';' as the output delimiter,
    DataPartitioner picks its best line and partitions the node's rows."""
    schema = hangup_schema()
    lines = synth.call_hangup_csv(4000)
    pp, c, node = layout(str(tmp_path), schema, lines, [])
    os.remove(os.path.join(os.path.dirname(node), "splits", "part-r-00000"))
    cp = os.path.join(str(tmp_path), "spg.properties")
    with open(cp, "w") as f:
        f.write(f"project.base.path={c['dap.project.base.path']}\ncpg.feature.schema.file.path={c['dap.feature.schema.file.path']}\n"
                "cpg.split.attributes=1,3\ncpg.split.algorithm=giniIndex\ncpg.parent.info=0.4\nfield.delim.out=;\n")
    rc, err = helpers.tool_run(lib, SPG, [f"-Dconf.path={cp}"])
    assert rc == 0, err
    rc, err = helpers.tool_run(lib, DAP, [f"-Dconf.path={pp}"])
    assert rc == 0, err
    want = dap_literal.dap_run(c)
    # ClassPartitionGenerator also emits the one-group "split" of every categorical attribute, whose gain ratio is x / 0 =
    # Infinity: it sorts first, so the reference pipeline's "best" split at the root has a single segment with every row in it
    assert written(node, want.keys()) == want and list(want) == ["split=0/segment=0/data/partition.txt"]
    assert want["split=0/segment=0/data/partition.txt"].count("\n") == 4000


@pytest.mark.parametrize("split_lines,rows,what", [
    ([], None, "ArrayIndexOutOfBoundsException: 0"),
    (["3;[cable]:[internet]", "1;[business]:[residence];0.1"], None, "ArrayIndexOutOfBoundsException: 2"),
    (["3;[cable]:[internet];abc", "1;[business]:[residence];0.1"], None, "NumberFormatException"),
    (["3;[cable]:[internet];0.5"], None, "split segment not found for billing"),
    (["5;30x;0.5"], None, "NumberFormatException"),
    (["0;10;0.5"], None, "NullPointerException"),
    (["9;10;0.5"], None, "NullPointerException"),
    (["5;300;0.5"], ["1,business,12,cable,AM,3.5,T"], 'For input string: "3.5"'),
    (["3;[cable]:[internet];0.5"], ["1,business"], "ArrayIndexOutOfBoundsException: 3"),
])
def test_reference_failures(lib, tmp_path, split_lines, rows, what):
    lines = rows if rows is not None else synth.call_hangup_csv(300)
    pp, c, node = layout(str(tmp_path), hangup_schema(), lines, split_lines)
    rc, err = helpers.tool_run(lib, DAP, [f"-Dconf.path={pp}"])
    assert rc == 1 and what in err, err
    with pytest.raises(dap_literal.JobFailure) as ei:
        dap_literal.dap_run(c)
    assert what.split(":")[0].split(" ")[0] in str(ei.value) or what in str(ei.value)


def test_unseeded_random_and_missing_base_path(lib, tmp_path):
    pp, c, node = layout(str(tmp_path), hangup_schema(), synth.call_hangup_csv(100), ["3;[cable]:[internet, billing, other];0.5"],
                         conf={"dap.split.selection.strategy": "randomFromTop"})
    rc, err = helpers.tool_run(lib, DAP, [f"-Dconf.path={pp}"])
    assert rc == 1 and "unsupported" in err and "avt.random.seed" in err
    empty = os.path.join(str(tmp_path), "e.properties")
    open(empty, "w").write("debug.on=false\n")
    rc, err = helpers.tool_run(lib, DAP, [f"-Dconf.path={empty}"])
    assert rc == 1 and "base path not defined" in err


@pytest.mark.gpu
def test_rows_partition_kernel_level():
    """avt_rows_partition against a numpy/python restatement: 200 k rows, uint16 codes (300 values), dropped bins, 7 segments."""
    from avenir_b200._lib import AVT_NODE_NONE
    from avenir_b200.engine import Engine
    rng = np.random.default_rng(11)
    vals = ["v%03d" % i for i in range(300)]
    n = 200_000
    code = rng.integers(0, 300, size=n)
    amount = rng.integers(-500, 500, size=n)
    lines = ["%d,%s,%d,tail%d" % (i, vals[code[i]], amount[i], i % 7) for i in range(n)]
    text = ("\n".join(lines) + "\n").encode()
    eng = Engine(0)
    eng.load_csv(text, ",", [(1, "categorical", vals), (2, "int", [-100, 0, 250])], -1, [])
    seg_of_bin = [(b * 7) // 300 if b % 11 else AVT_NODE_NONE for b in range(300)]
    parts, rows = eng.rows_partition(0, seg_of_bin, 7)
    for s in range(7):
        want = [lines[i] for i in range(n) if seg_of_bin[code[i]] == s]
        assert rows[s] == len(want) and parts[s].decode() == "".join(w + "\n" for w in want)
    parts, rows = eng.rows_partition(1, [0, 1, 2, 3], 4)
    bins = np.searchsorted(np.array([-100, 0, 250]), amount, side="left")
    assert rows == [int((bins == s).sum()) for s in range(4)] and sum(len(p) for p in parts) == len(text)
    assert parts[3].decode().splitlines() == [lines[i] for i in range(n) if amount[i] > 250]
    # a data set that did not come from text has no rows to move
    eng.load_dataset([np.zeros(10, np.uint8)], np.zeros(10, np.uint8), [2], 2)
    with pytest.raises(Exception):
        eng.rows_partition(0, [0, 1], 2)
    eng.close()
