"""CPU-side checks of the drop-in boundary: libstage_gpu.so loads and exports exactly the symbols
include/stage_gpu.h declares (no compute calls: there is no GPU here)."""
import ctypes
import os
import re

import stage_indexorganized_b200 as sg

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "stage_gpu.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(stage_gpu_\w+)\s*\(", src)))


def test_library_is_built_in_tree():
    assert os.path.exists(sg.LIB_PATH), "run __graft_entry__.build()"


def test_every_declared_symbol_is_exported():
    lib = ctypes.CDLL(sg.LIB_PATH)
    names = _declared()
    assert len(names) >= 15
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/stage_gpu.h but not exported"


def test_python_binding_covers_the_header():
    assert sorted(sg.SYMBOLS) == _declared()


def test_strerror_and_route_host_only():
    import numpy as np
    l = sg.lib()
    assert l.stage_gpu_strerror(0) == b"ok"
    assert b"quiesced" in l.stage_gpu_strerror(-3)
    # routing is host-only logic: YCSB int keys order is signed-byte, LSB first (b_tree.h:99-134)
    keys = np.array([0, 1, 255, 256, 128, 32896], np.uint32).view(np.uint8).reshape(-1, 4)
    split = np.array([0], np.uint32).view(np.uint8).reshape(-1, 4)       # shard 0 owns keys <= 0
    owner = sg.route(keys, split, 4)
    # in reference order 128 (0x80 -> -128) and 32896 sort before 0; 1, 255(-1)?, 256 ...
    #   byte0 of 255 is 0xFF = -1 < 0  -> before key 0 ; byte0 of 1 = 1 > 0 -> after
    assert owner.tolist() == [0, 1, 0, 1, 0, 0]
