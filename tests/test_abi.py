"""CPU: the C-ABI library builds, loads and exports every symbol include/srb200.h declares (no compute calls)."""
import os
import re

import numpy as np

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


def test_em_plan_rejects_long_and_unsorted():
    """host-side tile planner of the fused training pass (no CUDA call)"""
    import ctypes
    from srb200 import lib
    L = lib.load()
    tile = L.srb_em_tile_frames()
    off = np.array([0, 10, 10 + tile + 1], np.int32)
    w = np.zeros(2, np.int32)
    tiles = np.empty(12, np.int32)
    nt, tf = ctypes.c_int32(), ctypes.c_int32()
    assert L.srb_em_plan(off.ctypes.data, w.ctypes.data, 2, tiles.ctypes.data, ctypes.byref(nt), ctypes.byref(tf)) == -2
    off = np.array([0, 10, 20, 30], np.int32)
    w = np.array([1, 0, 0], np.int32)
    assert L.srb_em_plan(off.ctypes.data, w.ctypes.data, 3, tiles.ctypes.data, ctypes.byref(nt), ctypes.byref(tf)) == -1
    w = np.array([0, 0, 1], np.int32)
    assert L.srb_em_plan(off.ctypes.data, w.ctypes.data, 3, tiles.ctypes.data, ctypes.byref(nt), ctypes.byref(tf)) == 0
    assert tf.value == 128 and nt.value == 2 and list(tiles[:8]) == [0, 0, 20, 2, 2, 20, 10, (1 << 8) | 1]
