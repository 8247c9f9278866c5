"""The reference's job entry points, by name and argument convention, on top of avt_tool_run.

    DecisionTreeBuilder().run(["-Dconf.path=detr.properties", inPath, outPath]) -> 0 | 1
    ClassPartitionGenerator().run([...]) -> 0 | 1

mirrors `ToolRunner.run(new DecisionTreeBuilder(), args)` (tree/DecisionTreeBuilder.java:70-95,773-776;
explore/ClassPartitionGenerator.java:64-91,611-614).  All work happens in libavenir_tree.so.
"""
from __future__ import annotations

import ctypes as C
import sys
from typing import List, Sequence

from . import _lib


class ToolFailure(RuntimeError):
    """The job failed (the reference would exit 1 with a stack trace in the task log)."""


class _Tool:
    java_class = ""

    def __init__(self):
        self.last_error = ""

    def run(self, args: Sequence[str]) -> int:
        lib = _lib.load()
        arr = (C.c_char_p * len(args))(*[a.encode() for a in args])
        err = C.create_string_buffer(4096)
        rc = lib.avt_tool_run(self.java_class.encode(), len(args), arr, err, len(err))
        self.last_error = err.value.decode(errors="replace")
        return rc

    def run_checked(self, args: Sequence[str]) -> None:
        if self.run(args) != 0:
            raise ToolFailure(f"{self.java_class}: {self.last_error}")


class DecisionTreeBuilder(_Tool):
    java_class = "org.avenir.tree.DecisionTreeBuilder"


class ClassPartitionGenerator(_Tool):
    java_class = "org.avenir.explore.ClassPartitionGenerator"


class ModelPredictor(_Tool):
    java_class = "org.avenir.model.ModelPredictor"


class CramerCorrelation(_Tool):
    java_class = "org.avenir.explore.CramerCorrelation"


class SplitGenerator(_Tool):
    java_class = "org.avenir.tree.SplitGenerator"


class HeterogeneityReductionCorrelation(_Tool):
    java_class = "org.avenir.explore.HeterogeneityReductionCorrelation"


class CategoricalClassAffinity(_Tool):
    java_class = "org.avenir.explore.CategoricalClassAffinity"


class CategoricalContinuousEncoding(_Tool):
    java_class = "org.avenir.explore.CategoricalContinuousEncoding"


class DataPartitioner(_Tool):
    """tree/DataPartitioner.java:59-105; takes its paths from dap.project.base.path / dap.split.path, not from arguments."""
    java_class = "org.avenir.tree.DataPartitioner"


def main(argv: List[str] | None = None) -> int:
    argv = list(sys.argv[1:] if argv is None else argv)
    if not argv:
        print("usage: python -m avenir_b200 <DecisionTreeBuilder|ClassPartitionGenerator|ModelPredictor|CramerCorrelation> -Dconf.path=<props> <in> <out>", file=sys.stderr)
        return 2
    tool = _Tool()
    tool.java_class = argv[0]
    rc = tool.run(argv[1:])
    if rc:
        print(f"{argv[0]} failed: {tool.last_error}", file=sys.stderr)
    return rc
