"""ctypes face of oracle/_ref/libdbow2_ref.so: the reference's OWN DBoW2 code (compiled by oracle/Makefile from
/root/reference/Thirdparty/DBoW2, wrapper oracle/dbow2_ref.cpp).  TEST INFRASTRUCTURE: the pin for oracle/vocab_oracle.py
and the `"kind": "reference"` CPU baseline of the vocabulary transform.  The .so is prebuilt in the container and
travels to the GPU box; nothing here reads /root/reference at run time."""
import ctypes as C
import os

import numpy as np

_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "libdbow2_ref.so")
_lib = None


def available():
    return os.path.exists(_PATH)


def lib():
    global _lib
    if _lib is None:
        l = C.CDLL(_PATH)
        vp, ip, dp, u8 = C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_double), C.c_void_p
        l.dbow2_ref_new.restype = vp
        l.dbow2_ref_new.argtypes = [C.c_int] * 4
        l.dbow2_ref_free.argtypes = [vp]
        l.dbow2_ref_load_text.argtypes = [vp, C.c_char_p]
        l.dbow2_ref_save_text.argtypes = [vp, C.c_char_p]
        l.dbow2_ref_create.argtypes = [vp, u8, ip, C.c_int, C.c_int]
        l.dbow2_ref_info.argtypes = [vp] + [ip] * 6
        l.dbow2_ref_transform_features.argtypes = [vp, u8, C.c_int, C.c_int, ip, dp, ip]
        l.dbow2_ref_transform.argtypes = [vp, u8, C.c_int, C.c_int, ip, ip, dp, ip, ip, ip, ip]
        l.dbow2_ref_score.restype = C.c_double
        l.dbow2_ref_score.argtypes = [vp, C.c_int, ip, dp, C.c_int, ip, dp]
        _lib = l
    return _lib


def _ip(a):
    return a.ctypes.data_as(C.POINTER(C.c_int))


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


class RefVocabulary:
    """DBoW2::TemplatedVocabulary<cv::Mat, FORB> (== ORB_SLAM2::ORBVocabulary, include/ORBVocabulary.h:31-32)."""

    def __init__(self, k=10, L=5, weighting=0, scoring=0):
        self.h = lib().dbow2_ref_new(k, L, weighting, scoring)

    def __del__(self):
        if getattr(self, "h", None):
            lib().dbow2_ref_free(self.h)
            self.h = None

    def load_text(self, path):
        return lib().dbow2_ref_load_text(self.h, path.encode()) == 0

    def save_text(self, path):
        lib().dbow2_ref_save_text(self.h, path.encode())

    def create(self, desc, counts, seed=1):
        desc = np.ascontiguousarray(desc, dtype=np.uint8)
        counts = np.ascontiguousarray(counts, dtype=np.int32)
        assert counts.sum() == len(desc)
        lib().dbow2_ref_create(self.h, desc.ctypes.data, _ip(counts), len(counts), seed)

    def info(self):
        v = [C.c_int() for _ in range(6)]
        lib().dbow2_ref_info(self.h, *[C.byref(x) for x in v])
        return dict(zip(("k", "L", "nodes", "words", "scoring", "weighting"), [x.value for x in v]))

    def transform_features(self, desc, levelsup):
        desc = np.ascontiguousarray(desc, dtype=np.uint8).reshape(-1, 32)
        n = len(desc)
        word, weight, nid = np.zeros(n, np.int32), np.zeros(n, np.float64), np.zeros(n, np.int32)
        lib().dbow2_ref_transform_features(self.h, desc.ctypes.data, n, levelsup, _ip(word), _dp(weight), _ip(nid))
        return word, weight, nid

    def transform(self, desc, levelsup=4):
        desc = np.ascontiguousarray(desc, dtype=np.uint8).reshape(-1, 32)
        n = len(desc)
        bid, bval = np.zeros(n + 1, np.int32), np.zeros(n + 1, np.float64)
        fnode, foff, ffeat = np.zeros(n + 1, np.int32), np.zeros(n + 2, np.int32), np.zeros(n + 1, np.int32)
        nb, nf = C.c_int(), C.c_int()
        lib().dbow2_ref_transform(self.h, desc.ctypes.data, n, levelsup, C.byref(nb), _ip(bid), _dp(bval), C.byref(nf),
                                  _ip(fnode), _ip(foff), _ip(ffeat))
        bow = {int(bid[i]): float(bval[i]) for i in range(nb.value)}
        fv = {int(fnode[j]): [int(x) for x in ffeat[foff[j]:foff[j + 1]]] for j in range(nf.value)}
        return bow, fv

    def score(self, a, b):
        ia, va = np.array(list(a.keys()), np.int32), np.array(list(a.values()), np.float64)
        ib, vb = np.array(list(b.keys()), np.int32), np.array(list(b.values()), np.float64)
        return lib().dbow2_ref_score(self.h, len(ia), _ip(ia), _dp(va), len(ib), _ip(ib), _dp(vb))
