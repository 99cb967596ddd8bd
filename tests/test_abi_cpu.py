"""CPU checks of the boundary: the C-ABI library loads and exports every symbol include/margin_b200.h declares
(no compute calls here), and refuses to compute without a device."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from margin_b200 import abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "margin_b200.h")).read()
    src = "\n".join(l for l in src.splitlines() if not l.startswith("static inline"))  # header-only helpers
    return sorted(set(re.findall(r"\b(margin_b200_[a-z_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    if not os.path.exists(abi.LIB_PATH):
        import __graft_entry__ as g
        g.build()
    lib = C.CDLL(abi.LIB_PATH)
    names = _declared()
    assert len(names) >= 13
    for n in names:
        assert hasattr(lib, n), n
    assert sorted(abi.EXPORTED_SYMBOLS) == names


def test_struct_layouts_match_header():
    assert C.sizeof(abi.Model) == 8 * (9 + 16 + 4 + 4) + 8 + 8 * 4 * 51 * 51
    assert C.sizeof(abi.Params) == 48
    assert C.sizeof(abi.BatchStruct) == 88 + 32
    assert C.sizeof(abi.ResultStruct) == 8 + 24 + 24 + 32 + 8 + 8 + 8 + 16 + 8 + 24 + 24
    assert abi.PAIR_DTYPE.itemsize == 12


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a device is present")
    lib = abi.load_library()
    h = C.c_void_p()
    assert lib.margin_b200_create(C.byref(h), 0) != 0
    assert b"no usable CUDA device" in lib.margin_b200_last_error()


def test_symbol_and_run_packing_on_the_host():
    sx = np.array([0, 1 | (3 << 8), 4 | (50 << 8), 7 | (2 << 8)], np.uint16)
    b = abi.Batch([(sx, sx[:2], np.array([[0, 0, 2], [1, 1, 2], [2, 3, 2], [3, 4, 2], [4, 5, 4]], np.int32), 0, 0, 0)], compact=True)
    assert b.packed_x.tolist() == [0, 51 + 3, 4 * 51 + 50, 4 * 51 + 2]
    assert b.anchors_runs.tolist() == [(1 << 8) | (1 << 18) | (1 << 28) | 1, (1 << 8) | (2 << 18) | (1 << 28) | 1,
                                       (1 << 8) | (1 << 18) | (2 << 28)]
    assert b.off_r.tolist() == [0, 3]
    over = sx.copy()
    over[0] = 51 << 8
    assert not abi.Batch([(over, sx, np.zeros((0, 3), np.int32), 0, 0, 0)], compact=True).packed
