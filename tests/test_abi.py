"""CPU: the C-ABI library builds, loads and exports every symbol include/srb200.h declares (no compute calls)."""
import os
import re

from conftest import ROOT


def test_library_exports_every_declared_symbol():
    from srb200 import build, lib
    build.build()
    L = lib.load()
    hdr = open(os.path.join(ROOT, "include", "srb200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(srb_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 15
    assert declared == set(lib.SIGNATURES), declared ^ set(lib.SIGNATURES)
    for name in declared:
        assert getattr(L, name) is not None
    assert L.srb_version() >= 100
    assert L.srb_stats_row(39) == 81


def test_no_oracle_import_in_product():
    pkg = os.path.join(ROOT, "speech-recognition_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dp, f)).read()
                assert "oracle" not in src, f"{f} mentions the oracle: the product must never route through it"
