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


def test_level_tables_host_logic_equals_oracle():
    """E1 (ORBextractor.cc:441-461, 468-493, 510-533) runs on the host without a GPU: compare with the oracle."""
    import ctypes as C
    import numpy as np
    import ref_cpu
    L = orbb200.lib()
    for nf, sfac, nl in ((1000, 1.2, 8), (2000, 1.2, 8), (1200, 1.2, 8), (8000, 1.2, 12), (500, 1.3, 5), (40, 1.5, 3)):
        t = [np.zeros(nl, np.float32) for _ in range(4)] + [np.zeros(nl, np.int32), np.zeros(16, np.int32)]
        assert L.orbb200_compute_level_tables(nf, sfac, nl, *[a.ctypes.data_as(C.c_void_p) for a in t]) == 0
        tb = ref_cpu.RefExtractor(nf, sfac, nl, 20, 7).tables()
        for a, k in zip(t, ("scale", "inv_scale", "sigma2", "inv_sigma2", "nfeat", "umax")):
            assert np.array_equal(a, tb[k]), (k, nf, sfac, nl)
        assert t[4].sum() == nf
    assert L.orbb200_compute_level_tables(0, 1.2, 8, *[None] * 6) != 0
    assert b"bad parameters" in L.orbb200_last_error()
