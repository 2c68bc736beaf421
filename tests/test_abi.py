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


def test_vocabulary_facade_scores_equal_reference(tmp_path):
    """include/orbb200/ORBVocabulary.h::score -- host arithmetic, no device -- for all six DBoW2 scoring classes against
    the reference's ScoringObject.cpp outputs (tests/golden/vocab_golden.npz, tools/make_vocab_golden.py)."""
    import subprocess
    import numpy as np
    import vocab_cases as vc
    exe = os.path.join(ROOT, "orb-slam2_b200", "facade_vocab.bin")
    assert os.path.exists(exe), "facade_vocab.bin not built: run __graft_entry__.build()"
    for name in ("a", "b"):
        files = []
        for part in (0, 1):
            bow, _ = vc.gold_maps("%s_w0_s0_l4" % name, part)
            p = str(tmp_path / ("%s%d.txt" % (name, part)))
            open(p, "w").write("".join("%d %s\n" % (k, float(v).hex()) for k, v in bow.items()))
            files.append(p)
        for sc in range(6):
            r = subprocess.run([exe, "--score", str(sc)] + files, capture_output=True, text=True, timeout=60)
            assert r.returncode == 0, r.stderr
            got = [float.fromhex(x) for x in r.stdout.split()[1:3]]
            assert got == list(vc.gold()["%s_scores6" % name][sc]), (name, sc, got)
