"""The C-ABI library loads, exports every symbol include/avenir_tree.h declares, and -- on a box without
a CUDA device -- refuses to run instead of falling back to a CPU path."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "avenir_tree.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(avt_[a-z0-9_]+)\s*\(", text)))


@pytest.fixture(scope="module")
def lib():
    from avenir_b200 import build, _lib
    build.build()
    return _lib.load()


def test_every_declared_symbol_is_exported_and_bound(lib):
    from avenir_b200 import _lib
    syms = declared_symbols()
    assert len(syms) >= 24
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/avenir_tree.h but not exported"
    assert set(syms) == set(_lib.SIGNATURES), "ctypes binding and header disagree"


def test_struct_layouts_match_header(lib):
    from avenir_b200 import _lib
    assert C.sizeof(_lib.Split) == 16
    assert C.sizeof(_lib.NodeBest) == 32
    assert C.sizeof(_lib.SplitEval) == 8 + 8 * 8 + 8 * 16 * 4
    assert C.sizeof(_lib.CpgStat) == 16 + 8 * 16 * 4


def test_no_cpu_fallback_without_device(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    h = C.c_void_p()
    rc = lib.avt_create(0, C.byref(h))
    assert rc == -2 and not h.value
    assert b"no CPU path" in lib.avt_last_error(None)
    from avenir_b200 import DecisionTreeBuilder
    tool = DecisionTreeBuilder()
    assert tool.run(["-Dconf.path=/nonexistent.properties", "in", "out"]) == 1
    assert tool.last_error


def test_product_does_not_reference_the_oracle():
    """Only tests/, bench.py's baseline legs and __graft_entry__.smoke() may touch oracle/."""
    bad = []
    for base, _, files in os.walk(os.path.join(ROOT, "avenir_b200")):
        for f in files:
            if f.endswith((".py", ".cpp", ".hpp", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(base, f), errors="replace").read()
                if re.search(r"(from|import)\s+oracle|hist_oracle|libhist_oracle|oracle/", txt) and "never" not in txt.lower():
                    bad.append(f)
    assert not bad, bad
