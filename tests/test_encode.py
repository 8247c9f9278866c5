"""On-device CSV encode: avt_dataset_load_csv against the literal restatement of the reference's
row parsing (oracle/encode_literal.py).  CPU: the restatement against known answers and against the synthetic
generator's own codes; GPU: the kernels through the C-ABI, bit for bit, incl. every rejection the reference makes."""
import numpy as np
import pytest

from avenir_b200 import synth
from oracle import encode_literal as enc

HANGUP_ATTRS = [(1, "categorical", list(synth.CUST_TYPES)), (3, "categorical", list(synth.ISSUE_CARD)),
                (4, "categorical", list(synth.TOD)), (5, "int", [int(p) for p in synth.HOLD_POINTS])]


def test_restatement_known_answers():
    assert enc.parse_java_int("+12") == 12 and enc.parse_java_int("-2147483648") == -2147483648
    assert enc.parse_java_int("2147483648") is None and enc.parse_java_int(" 1") is None and enc.parse_java_int("") is None
    assert enc.parse_java_int("1.0") is None and enc.parse_java_int("-") is None
    assert enc.parse_java_double(" 1.5e2\t") == 150.0 and enc.parse_java_double(".5") == 0.5 and enc.parse_java_double("5.") == 5.0
    assert enc.parse_java_double("1e") is None and enc.parse_java_double("") is None and enc.parse_java_double("0x10") is None
    attrs = [(0, "int", [10, 20]), (2, "double", [0.5]), (3, "categorical", ["a", "b"])]
    assert enc.encode_row("10,x,0.5,b,Y", ",", attrs, 4, ["N", "Y"]) == ([0, 0, 1], 1, 0)      # bin = number of points < value
    assert enc.encode_row("11,x,0.51,a,N", ",", attrs, 4, ["N", "Y"]) == ([1, 1, 0], 0, 0)
    assert enc.encode_row("21,x,7,a,N,,,", ",", attrs, 4, ["N", "Y"])[0] == [2, 1, 0]           # trailing empties are dropped ...
    assert enc.encode_row("21,x,7,a,,,", ",", attrs, 4, ["N", "Y"])[2] == enc.BAD_SHORT_ROW       # ... so an empty class field is a short row
    assert enc.encode_row("21,x,,a,N", ",", attrs, 4, ["N", "Y"])[2] == enc.BAD_DOUBLE            # an empty field in the middle is parsed
    assert enc.encode_row("2x,x,7,a,N", ",", attrs, 4, ["N", "Y"])[2] == enc.BAD_INT
    assert enc.encode_row("2,x,7,c,N", ",", attrs, 4, ["N", "Y"])[2] == enc.BAD_CATEGORY
    assert enc.encode_row("2,x,7,a,M", ",", attrs, 4, ["N", "Y"])[2] == enc.BAD_CLASS
    assert enc.encode_row("", ",", attrs, 4, ["N", "Y"])[2] == enc.BAD_INT                        # "".split(",") is [""]
    assert enc.encode_row("p;q,10,x,0.5,b,Y", ",", attrs, 4, ["N", "Y"], skip_fields=1) == ([0, 0, 1], 1, 0)


def test_restatement_equals_generator_codes():
    n = 3000
    lines = synth.call_hangup_csv(n)
    cols, cls, first, bad = enc.encode_text("\n".join(lines) + "\n", ",", HANGUP_ATTRS, 6, ["T", "F"])
    want_cols, want_cls, n_bins, names = synth.call_hangup_encoded(n)
    assert bad is None
    for a in range(4):
        assert cols[a] == want_cols[a].tolist()
    assert [["T", "F"][c] for c in cls] == [names[c] for c in want_cls.tolist()]
    assert sorted(range(2), key=lambda k: first[k]) == [["T", "F"].index(x) for x in names]      # first-seen order


def random_schema_text(rng, n, tagged=False, crlf=False):
    pts_i = sorted(rng.choice(np.arange(-500, 500), size=300, replace=False).tolist())     # 301 bins: uint16 column
    pts_d = [-1.5, 0.0, 0.25, 1e3, 2.5e7]
    cats = ["red", "re", "reddish", "", "blue green", "Zz"]
    attrs = [(0, "int", pts_i), (1, "double", pts_d), (3, "categorical", cats), (6, "int", [0])]
    classes = ["yes", "no", "maybe"]
    dbl_forms = ["%d", "%.3f", "%.6e", " %.2f ", "%.17g", "1e%d", "%d.", ".%d"]
    lines = []
    for r in range(n):
        d = rng.normal(0, 1000)
        form = dbl_forms[rng.integers(len(dbl_forms))]
        ds = form % (int(rng.integers(-30, 30)) if form == "1e%d" else (abs(int(d)) if form in ("%d.", ".%d") else (int(d) if form == "%d" else d)))
        f = [str(int(rng.integers(-600, 600))), ds, "junk", cats[rng.integers(len(cats))], "x", classes[rng.integers(3)], str(int(rng.integers(-2, 3)))]
        if rng.random() < 0.1:
            f[0] = "+" + f[0].lstrip("-")
        ln = ",".join(f)
        if tagged:
            ln = "$root;1 le 5," + ln
        lines.append(ln)
    return attrs, classes, ("\r\n" if crlf else "\n").join(lines), lines


@pytest.mark.gpu
@pytest.mark.parametrize("n,tagged,crlf,newline_at_end", [(1, False, False, False), (257, False, False, True), (5003, True, False, True),
                                                          (4099, False, True, False), (60001, False, False, True)])
def test_device_encode_equals_restatement(n, tagged, crlf, newline_at_end):
    from avenir_b200.engine import Engine
    rng = np.random.default_rng(n)
    attrs, classes, text, lines = random_schema_text(rng, n, tagged, crlf)
    if newline_at_end:
        text += "\r\n" if crlf else "\n"
    cols, cls, first, bad = enc.encode_text("\n".join(lines) + "\n", ",", attrs, 5, classes, skip_fields=int(tagged))
    assert bad is None
    g = Engine(0)
    n_rows, got_first = g.load_csv(text.encode(), ",", attrs, 5, classes, skip_fields=int(tagged))
    assert n_rows == n and got_first == first
    for a in range(len(attrs)):
        got = g.fetch_column(a)
        assert got.dtype == (np.uint16 if a == 0 else np.uint8)
        assert np.array_equal(got, np.asarray(cols[a]))
    assert np.array_equal(g.fetch_column(-1), np.asarray(cls, dtype=np.uint8))


@pytest.mark.gpu
def test_device_encode_rejections_match_the_reference_order():
    from avenir_b200._lib import CSV_BAD_CATEGORY, CSV_BAD_CLASS, CSV_BAD_DOUBLE, CSV_BAD_INT, CSV_BAD_SHORT_ROW
    from avenir_b200.engine import CsvRowError, Engine
    attrs = [(0, "int", [10, 20]), (2, "double", [0.5]), (3, "categorical", ["a", "b"])]
    good = ["11,x,0.51,a,N"] * 40
    g = Engine(0)
    cases = [("2x,x,7,a,N", CSV_BAD_INT), ("2147483648,x,7,a,N", CSV_BAD_INT), ("2,x,7..1,a,N", CSV_BAD_DOUBLE), ("2,x,,a,N", CSV_BAD_DOUBLE),
             ("2,x,7,c,N", CSV_BAD_CATEGORY), ("2,x,7,a,M", CSV_BAD_CLASS), ("2,x,7,a,,,", CSV_BAD_SHORT_ROW), ("2,x", CSV_BAD_SHORT_ROW),
             ("", CSV_BAD_INT), ("2x,x", CSV_BAD_INT), ("2,x,7,c", CSV_BAD_CATEGORY), ("2,x,y7,c,M", CSV_BAD_DOUBLE)]
    for i, (row, reason) in enumerate(cases):
        assert enc.encode_row(row, ",", attrs, 4, ["N", "Y"])[2] == reason, row
        lines = good[:7 + i] + [row] + good + [cases[(i + 1) % len(cases)][0]]          # the FIRST offending row is reported
        with pytest.raises(CsvRowError) as ei:
            g.load_csv(("\n".join(lines) + "\n").encode(), ",", attrs, 4, ["N", "Y"])
        assert (ei.value.row, ei.value.reason) == (7 + i, reason), row
    n, first = g.load_csv(("\n".join(good) + "\n").encode(), ",", attrs, 4, ["N", "Y"])       # the engine is usable afterwards
    assert n == 40 and first == [0, -1]


@pytest.mark.gpu
def test_csv_path_grows_the_same_tree_as_encoded_columns():
    from avenir_b200._lib import CRIT_ENTROPY
    from avenir_b200.engine import Engine, binary_split_table
    n = 200_003
    text = ("\n".join(synth.call_hangup_csv(n)) + "\n").encode()
    cols, cls, n_bins, names = synth.call_hangup_encoded(n)
    table = binary_split_table(n_bins, [True, True, True, False])
    a, b = Engine(0), Engine(0)
    a.load_dataset(cols, cls, n_bins, 2)
    rows, _ = b.load_csv(text, ",", HANGUP_ATTRS, 6, names)
    assert rows == n
    out = []
    for e in (a, b):
        e.set_splits(table)
        e.forest_init(1)
        (rec, nrec), frontier = e.grow(6, CRIT_ENTROPY, [1], max_records=2048)
        out.append(([(r.level, r.node, r.split, list(r.child_pop), r.min_av.hex()) for r in rec[:nrec]], frontier, e.node_ids(0).copy()))
    assert out[0][0] == out[1][0] and out[0][1] == out[1][1] and np.array_equal(out[0][2], out[1][2])


def test_c_restatement_equals_python_restatement():
    """oracle/hist_oracle.c:ora_encode_csv (the checker at large sizes and the timed CPU baseline) against
    oracle/encode_literal.py, accepted rows and first rejected row."""
    from tests import helpers
    rng = np.random.default_rng(3)
    attrs, classes, text, lines = random_schema_text(rng, 4000)
    cols, cls, first, bad = enc.encode_text(text + "\n", ",", attrs, 5, classes)
    for threads in (1, 5):
        got_cols, got_cls = helpers.oracle_encode_csv((text + "\n").encode(), ",", attrs, 5, classes, threads=threads)
        assert all(np.array_equal(got_cols[a], np.asarray(cols[a])) for a in range(len(attrs))) and np.array_equal(got_cls, np.asarray(cls))
    lines[1234] = lines[1234].replace(",", ",z", 1)        # attribute 1 (double) of row 1234 becomes unparsable
    lines[3000] = "7"
    assert enc.encode_text("\n".join(lines), ",", attrs, 5, classes)[3][0] == 1234
    assert helpers.oracle_encode_csv("\n".join(lines).encode(), ",", attrs, 5, classes, threads=3) == 1234


@pytest.mark.gpu
def test_device_encode_two_million_rows_equals_c_restatement():
    from avenir_b200.engine import Engine
    from tests import helpers
    n = 2_000_000
    text = ("\n".join(synth.call_hangup_csv(n)) + "\n").encode()
    cols, cls = helpers.oracle_encode_csv(text, ",", HANGUP_ATTRS, 6, ["T", "F"], threads=8)
    g = Engine(0)
    rows, first = g.load_csv(text, ",", HANGUP_ATTRS, 6, ["T", "F"])
    assert rows == n == len(cls) and min(first) == 0
    for a in range(4):
        assert np.array_equal(g.fetch_column(a), cols[a].astype(np.uint8))
    assert np.array_equal(g.fetch_column(-1), cls)
