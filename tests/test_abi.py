"""CPU: the C-ABI library loads and exports every symbol include/orbb200.h declares (no compute calls)."""
import ctypes
import os
import re

import orbb200

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    txt = open(os.path.join(ROOT, "include", "orbb200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(orbb200_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol():
    assert os.path.exists(orbb200.LIB_PATH), "liborbb200.so not built: run __graft_entry__.build()"
    L = ctypes.CDLL(orbb200.LIB_PATH)
    names = _declared()
    assert len(names) >= 20
    for n in names:
        assert hasattr(L, n), "missing export " + n
    assert sorted(orbb200.EXPORTS) == names


def test_version_and_error_strings():
    L = orbb200.lib()
    assert b"sm_100a" in L.orbb200_version()
    assert isinstance(L.orbb200_last_error(), bytes)


def test_keypoint_layout_is_cv_keypoint():
    assert orbb200.KP_DTYPE.itemsize == 28
    assert orbb200.TOP2_DTYPE.itemsize == 16
