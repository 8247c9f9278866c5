"""
ORACLE / TEST INFRASTRUCTURE ONLY -- never imported by the product (avenir_b200/).

Literal restatement of org.avenir.tree.DataPartitioner, object by object and row by row:
  tree/DataPartitioner.java:137-151   getNodePath: <base>/split=root/data[/<dap.split.path>]
  tree/DataPartitioner.java:160-206   findBestSplitKey: lines of the sibling `splits/part-r-00000` -> Split objects, Arrays.sort
                                      (descending third field), index 0 or (int)(Math.random() * dap.num.top.splits)
  tree/DataPartitioner.java:208-275   Split: line.split(";"), compareTo on Double.parseDouble(items[2]), segment count =
                                      items[1].split(":").length
  tree/DataPartitioner.java:291-337   PartitionerMapper: IntegerSplit / CategoricalSplit .fromString, key = getSegmentIndex
  util/AttributeSplitHandler.java:146-166, 190-236   the two Split classes
  tree/DataPartitioner.java:344-352   PartitionerReducer: values written unchanged
  tree/DataPartitioner.java:114-129   moveOutputToSegmentDir
PARITY UNPINNED (no JVM, the reference has no tests or fixtures for this job).  Assumptions about code that is not in the
reference's tree (chombo): SecondarySort.RawIntKeyTextPartitioner = key mod numReducers; Utility.getFileLines = the file's
lines; Utility.intArrayFromString = split + Integer.parseInt; Utility.getSiblingPath(p, s) = parent(p)/s.  Rows of one key
keep their input order (one mapper, Hadoop's merge keeps the spill order of equal keys in local mode).  The seeded
java.util.Random behind randomFromTop is this repo's addition.
"""
from __future__ import annotations

import functools
import os
from typing import Dict, List, Optional

from .encode_literal import parse_java_double, parse_java_int
from .javaemu import JavaRandom
from .literal import FeatureSchema, java_int_cast, java_split


class JobFailure(Exception):
    """The reference's job would die with this exception."""


class Split:
    """DataPartitioner.Split (tree/DataPartitioner.java:208-275)."""

    def __init__(self, line: str, index: int):
        self.line, self.index = line, index
        self.items = java_split(line, ";", False)

    def _item(self, k: int) -> str:
        if k >= len(self.items):
            raise JobFailure(f"ArrayIndexOutOfBoundsException: {k}")
        return self.items[k]

    def value(self) -> float:
        v = parse_java_double(self._item(2))
        if v is None:
            raise JobFailure(f'NumberFormatException: For input string: "{self._item(2)}"')
        return v

    def compare_to(self, that: "Split") -> int:
        a, b = self.value(), that.value()
        return -1 if a > b else (1 if a < b else 0)   # descending order

    def split_key(self) -> str:
        return self._item(1)

    def attribute_ordinal(self) -> int:
        v = parse_java_int(self._item(0))
        if v is None:
            raise JobFailure(f'NumberFormatException: For input string: "{self._item(0)}"')
        return v

    def segment_count(self) -> int:
        return len(java_split(self._item(1), ":", False))


class IntegerSplit:
    """util/AttributeSplitHandler.java:133-166."""

    def __init__(self, key: str):
        self.key = key
        self.split_points: List[int] = []

    def from_string(self):
        for it in java_split(self.key, ":", False):
            v = parse_java_int(it)
            if v is None:
                raise JobFailure(f'NumberFormatException: For input string: "{it}"')
            self.split_points.append(v)

    def segment_index(self, value: str) -> int:
        v = parse_java_int(value)
        if v is None:
            raise JobFailure(f'NumberFormatException: For input string: "{value}"')
        i = 0
        while i < len(self.split_points) and v > self.split_points[i]:
            i += 1
        return i


class CategoricalSplit:
    """util/AttributeSplitHandler.java:172-236."""

    def __init__(self, key: str):
        self.key = key
        self.split_sets: List[List[str]] = []

    def from_string(self):
        for piece in java_split(self.key, ":", False):
            if len(piece) < 2:
                raise JobFailure(f"StringIndexOutOfBoundsException: String index out of range: {len(piece) - 1}")
            inner = piece[1:len(piece) - 1]
            self.split_sets.append([it.strip(" \t\r\n") for it in java_split(inner, ",", False)])

    def segment_index(self, value: str) -> int:
        for i, gr in enumerate(self.split_sets):
            if value in gr:
                return i
        raise JobFailure(f"IllegalArgumentException: split segment not found for {value}")


def _input_lines(path: str) -> List[str]:
    files = [path]
    if os.path.isdir(path):
        files = sorted(os.path.join(path, n) for n in os.listdir(path)
                       if not n.startswith((".", "_")) and os.path.isfile(os.path.join(path, n)))
    out: List[str] = []
    for f in files:
        data = open(f, "rb").read().decode()
        lines = data.split("\n")
        if lines and lines[-1] == "":
            lines.pop()
        out.extend(ln[:-1] if ln.endswith("\r") else ln for ln in lines)
    return out


def dap_run(conf: Dict[str, str], seed: Optional[int] = None) -> Dict[str, str]:
    """DataPartitioner.run: returns {relative output file -> contents} under the node path, after writing nothing.
    (The caller compares it with the files the product wrote.)"""
    base = conf.get("dap.project.base.path")
    if base is None or base.strip() == "":
        raise JobFailure("IllegalStateException: base path not defined")
    split_path = conf.get("dap.split.path")
    in_path = base + "/split=root/data" if split_path is None or split_path.strip() == "" else base + "/split=root/data/" + split_path
    candidates = os.path.dirname(in_path) + "/splits/part-r-00000"
    splits = [Split(line, i) for i, line in enumerate(_input_lines(candidates))]
    splits.sort(key=functools.cmp_to_key(lambda a, b: a.compare_to(b)))   # stable, like Arrays.sort(Object[])
    index = 0
    if conf.get("dap.split.selection.strategy", "best") == "randomFromTop":
        assert seed is not None
        index = java_int_cast(JavaRandom(seed).next_double() * int(conf.get("dap.num.top.splits", "5")))
    if index >= len(splits):
        raise JobFailure(f"ArrayIndexOutOfBoundsException: {index}")
    split = splits[index]
    attr = split.attribute_ordinal()
    key = split.split_key()
    n_reducers = split.segment_count()
    out_dir = "split=" + str(split.index)
    # mapper set-up
    schema = FeatureSchema.load(conf["dap.feature.schema.file.path"])
    try:
        field = schema.find_field_by_ordinal(attr)
    except KeyError:
        raise JobFailure("NullPointerException") from None
    if field.is_integer():
        sp = IntegerSplit(key)
    elif field.is_categorical():
        sp = CategoricalSplit(key)
    else:
        raise JobFailure("NullPointerException")
    sp.from_string()
    delim = conf.get("field.delim.regex", ",")
    # map + shuffle: reducer = key mod reducers, reducer input sorted by key, rows of one key in input order
    keyed: List[List[tuple]] = [[] for _ in range(n_reducers)]
    for line in _input_lines(in_path):
        items = java_split(line, delim, False)
        if attr >= len(items):
            raise JobFailure(f"ArrayIndexOutOfBoundsException: {attr}")
        seg = sp.segment_index(items[attr])
        keyed[seg % n_reducers].append((seg, line))
    files: Dict[str, str] = {}
    for i in range(n_reducers):
        rows = sorted(keyed[i], key=lambda t: t[0])   # stable
        files[f"{out_dir}/segment={i}/data/partition.txt"] = "".join(r + "\n" for _, r in rows)
    return files
