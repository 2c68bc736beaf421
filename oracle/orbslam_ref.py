"""ctypes face of oracle/_ref/liborbslam_ref.so: the reference's OWN ORBextractor / ORBmatcher / Frame / KeyFrame / MapPoint
code (compiled UNMODIFIED by oracle/Makefile from /root/reference/src, wrapper oracle/orbslam_ref.cpp, OpenCV stand-in
oracle/shim + oracle/cv_shim_impl.cpp).  TEST INFRASTRUCTURE: the pin for oracle/orb_ref.c and for the CUDA path, and the
`"kind": "reference"` CPU baseline.  The .so is prebuilt in the container and travels to the GPU box; nothing here reads
/root/reference at run time."""
import ctypes as C
import os

import numpy as np

_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "liborbslam_ref.so")
KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])
GRID_COLS, GRID_ROWS = 64, 48
_lib = None


def available():
    return os.path.exists(_PATH)


def lib():
    global _lib
    if _lib is None:
        l = C.CDLL(_PATH)
        vp, i, f, d = C.c_void_p, C.c_int, C.c_float, C.c_double
        l.oref_set_monotonic.argtypes = [i]
        l.oref_set_monotonic(1)  # ties of DistributeOctTree's pointer sort resolve by creation order (see orbslam_ref.cpp)
        l.oref_extractor_create.restype = vp
        l.oref_extractor_create.argtypes = [i, f, i, i, i]
        l.oref_extractor_destroy.argtypes = [vp]
        l.oref_extractor_tables.argtypes = [vp] * 7
        l.oref_extract.restype = i
        l.oref_extract.argtypes = [vp, vp, i, i, i, vp, vp, i]
        l.oref_pyramid_level.restype = i
        l.oref_pyramid_level.argtypes = [vp, i, vp, vp, vp]
        l.oref_distribute_octree.restype = i
        l.oref_distribute_octree.argtypes = [vp, vp, vp, vp, i, i, i, i, i, i, i, vp, i]
        l.oref_frame_create.restype = vp
        l.oref_frame_create.argtypes = [vp, vp, i, i, i, f, i, i, i, vp, vp, i, f, f]
        l.oref_frame_destroy.argtypes = [vp]
        l.oref_frame_n.restype = i
        l.oref_frame_n.argtypes = [vp, i]
        l.oref_frame_keys.argtypes = [vp, i, vp, vp]
        l.oref_frame_stereo.argtypes = [vp, vp, vp]
        l.oref_frame_bounds.argtypes = [vp]
        l.oref_frame_grid.argtypes = [vp, vp, vp]
        l.oref_frame_features_in_area.restype = i
        l.oref_frame_features_in_area.argtypes = [vp, f, f, f, i, i, vp, i]
        l.oref_frame_pyramid_level.restype = i
        l.oref_frame_pyramid_level.argtypes = [vp, i, i, vp, vp, vp]
        l.oref_frame_from_keys.restype = vp
        l.oref_frame_from_keys.argtypes = [vp, vp, i, i, i, f, i]
        l.oref_session_create.restype = vp
        l.oref_session_create.argtypes = [i, f, i, i, i, vp, f, f]
        l.oref_session_destroy.argtypes = [vp]
        l.oref_session_frame.restype = i
        l.oref_session_frame.argtypes = [vp, vp, vp, i, i, vp]
        l.oref_distinctive_descriptors.restype = i
        l.oref_distinctive_descriptors.argtypes = [vp, vp, i, vp, vp]
        l.oref_descriptor_distance.restype = i
        l.oref_descriptor_distance.argtypes = [vp, vp]
        l.oref_probe_gemm.argtypes = [vp, i, i, vp, i, i, d, vp, i, i, d, i, vp]
        l.oref_probe_norm.restype = d
        l.oref_probe_norm.argtypes = [vp, i]
        l.oref_probe_norm_l1_diff.restype = d
        l.oref_probe_norm_l1_diff.argtypes = [vp, vp, i, i]
        l.oref_probe_expr.argtypes = [i, vp, vp, vp, f, vp]
        l.oref_probe_undistort.argtypes = [vp, i, vp, vp, i, vp]
        _lib = l
    return _lib


def set_monotonic(on):
    """1 (default): small allocations from a never-reusing arena, so heap-address order == node creation order;
    0: plain malloc, i.e. whatever glibc's reuse pattern gives (ORBextractor.cc:832 sorts by node pointer)."""
    lib().oref_set_monotonic(int(on))


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Extractor:
    """ORB_SLAM2::ORBextractor, the reference's own (src/ORBextractor.cc)."""

    def __init__(self, nfeatures, scale_factor, nlevels, ini_th, min_th):
        self.L = lib()
        self.h = self.L.oref_extractor_create(nfeatures, scale_factor, nlevels, ini_th, min_th)
        self.nlevels, self.nfeatures = nlevels, nfeatures

    def __del__(self):
        try:
            self.L.oref_extractor_destroy(self.h)
        except Exception:
            pass

    def tables(self):
        n = self.nlevels
        sf, isf, s2, is2 = (np.zeros(n, np.float32) for _ in range(4))
        nf, um = np.zeros(n, np.int32), np.zeros(16, np.int32)
        self.L.oref_extractor_tables(self.h, _p(sf), _p(isf), _p(s2), _p(is2), _p(nf), _p(um))
        return dict(scale=sf, inv_scale=isf, sigma2=s2, inv_sigma2=is2, nfeat=nf, umax=um)

    def extract(self, img):
        img = np.ascontiguousarray(img, dtype=np.uint8)
        h, w = img.shape
        cap = 4 * self.nfeatures + 4096
        kps = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        n = self.L.oref_extract(self.h, _p(img), w, h, w, _p(kps), _p(desc), cap)
        assert 0 <= n <= cap
        return kps[:n].copy(), desc[:n].copy()

    def level_image(self, l):
        w, h = C.c_int(), C.c_int()
        if self.L.oref_pyramid_level(self.h, l, C.byref(w), C.byref(h), None) != 0:
            return None
        out = np.zeros((h.value, w.value), np.uint8)
        self.L.oref_pyramid_level(self.h, l, C.byref(w), C.byref(h), _p(out))
        return out

    def distribute(self, cand, minX, maxX, minY, maxY, N, level=0):
        """DistributeOctTree on candidates (structured x/y/resp int arrays); returns indices into `cand`, in result order."""
        x = np.ascontiguousarray(cand["x"], np.int32)
        y = np.ascontiguousarray(cand["y"], np.int32)
        r = np.ascontiguousarray(cand["resp"], np.int32)
        cap = max(len(x), 1) + 16
        out = np.zeros(cap, np.int32)
        n = self.L.oref_distribute_octree(self.h, _p(x), _p(y), _p(r), len(x), minX, maxX, minY, maxY, N, level, _p(out), cap)
        assert n <= cap
        return out[:n].copy()


class RefFrame:
    """ORB_SLAM2::Frame built by the reference's own stereo / monocular constructor (src/Frame.cc)."""

    def __init__(self, imL, imR=None, nfeatures=2000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7,
                 cam=(718.856, 718.856, 607.1928, 185.2157), dist=(0, 0, 0, 0), bf=386.1448, th_depth=35.0, _handle=None):
        self.L = lib()
        if _handle is not None:
            self.h = _handle
            return
        imL = np.ascontiguousarray(imL, np.uint8)
        h, w = imL.shape
        if imR is not None:
            imR = np.ascontiguousarray(imR, np.uint8)
            assert imR.shape == imL.shape
        cam = np.asarray(cam, np.float32)
        dist = np.asarray(dist, np.float32)
        self.h = self.L.oref_frame_create(_p(imL), _p(imR), w, h, nfeatures, scale_factor, nlevels, ini_th, min_th, _p(cam), _p(dist),
                                          len(dist), bf, th_depth)
        self.bounds = np.zeros(6, np.float32)
        self.L.oref_frame_bounds(_p(self.bounds))

    @classmethod
    def from_keys(cls, kps, desc, w, h, scale_factor=1.2, nlevels=8):
        kps = np.ascontiguousarray(kps, KP_DTYPE)
        desc = np.ascontiguousarray(desc, np.uint8)
        hd = lib().oref_frame_from_keys(_p(kps), _p(desc), len(kps), w, h, scale_factor, nlevels)
        f = cls(None, _handle=hd)
        f.bounds = np.zeros(6, np.float32)
        f.L.oref_frame_bounds(_p(f.bounds))
        return f

    def __del__(self):
        try:
            self.L.oref_frame_destroy(self.h)
        except Exception:
            pass

    def n(self, right=False):
        return self.L.oref_frame_n(self.h, int(right))

    def keys(self, which=0):
        """0 = mvKeys + mDescriptors, 1 = mvKeysRight + mDescriptorsRight, 2 = mvKeysUn"""
        n = self.n(which == 1)
        kps = np.zeros(n, KP_DTYPE)
        desc = np.zeros((n, 32), np.uint8)
        self.L.oref_frame_keys(self.h, which, _p(kps), _p(desc))
        return kps, desc

    def stereo(self):
        n = self.n()
        u, d = np.zeros(n, np.float32), np.zeros(n, np.float32)
        self.L.oref_frame_stereo(self.h, _p(u), _p(d))
        return u, d

    def grid(self):
        start = np.zeros(GRID_COLS * GRID_ROWS + 1, np.int32)
        items = np.zeros(max(self.n(), 1), np.int32)
        self.L.oref_frame_grid(self.h, _p(start), _p(items))
        return start, items[:start[-1]]

    def features_in_area(self, x, y, r, min_level=-1, max_level=-1):
        out = np.zeros(max(self.n(), 1), np.int32)
        n = self.L.oref_frame_features_in_area(self.h, x, y, r, min_level, max_level, _p(out), len(out))
        return out[:n].copy()

    def level_image(self, l, right=False):
        w, h = C.c_int(), C.c_int()
        if self.L.oref_frame_pyramid_level(self.h, int(right), l, C.byref(w), C.byref(h), None) != 0:
            return None
        out = np.zeros((h.value, w.value), np.uint8)
        self.L.oref_frame_pyramid_level(self.h, int(right), l, C.byref(w), C.byref(h), _p(out))
        return out


class Session:
    """persistent extractor pair + the reference's stereo / mono Frame constructor per call (timing arm of bench.py)"""

    def __init__(self, nfeatures=2000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7, cam=(718.856, 718.856, 607.1928, 185.2157),
                 bf=386.1448, th_depth=35.0):
        self.L = lib()
        cam = np.asarray(cam, np.float32)
        self.h = self.L.oref_session_create(nfeatures, scale_factor, nlevels, ini_th, min_th, _p(cam), bf, th_depth)

    def __del__(self):
        try:
            self.L.oref_session_destroy(self.h)
        except Exception:
            pass

    def frame(self, imL, imR=None):
        h, w = imL.shape
        ns = C.c_int(0)
        n = self.L.oref_session_frame(self.h, _p(imL), _p(imR), w, h, C.byref(ns))
        return n, ns.value


def distinctive_descriptors(desc, offsets):
    """MapPoint::ComputeDistinctiveDescriptors per group -> (chosen descriptor [G, 32], has [G])"""
    desc = np.ascontiguousarray(desc, np.uint8)
    offsets = np.ascontiguousarray(offsets, np.int32)
    g = len(offsets) - 1
    out = np.zeros((g, 32), np.uint8)
    has = np.zeros(g, np.int32)
    lib().oref_distinctive_descriptors(_p(desc), _p(offsets), g, _p(out), _p(has))
    return out, has


def descriptor_distance(a, b):
    a = np.ascontiguousarray(a, np.uint8)
    b = np.ascontiguousarray(b, np.uint8)
    return lib().oref_descriptor_distance(_p(a), _p(b))


def probe_gemm(A, B, alpha=1.0, Cm=None, beta=0.0, flags=0):
    A = np.ascontiguousarray(A, np.float32)
    B = np.ascontiguousarray(B, np.float32)
    m = A.shape[1] if flags & 1 else A.shape[0]
    n = B.shape[0] if flags & 2 else B.shape[1]
    D = np.zeros((m, n), np.float32)
    if Cm is not None:
        Cm = np.ascontiguousarray(Cm, np.float32)
    lib().oref_probe_gemm(_p(A), A.shape[0], A.shape[1], _p(B), B.shape[0], B.shape[1], alpha, _p(Cm), 0 if Cm is None else Cm.shape[0],
                          0 if Cm is None else Cm.shape[1], beta, flags, _p(D))
    return D


def probe_norm(v):
    v = np.ascontiguousarray(v, np.float32).ravel()
    return lib().oref_probe_norm(_p(v), len(v))


def probe_norm_l1_diff(a, b):
    a = np.ascontiguousarray(a, np.float32)
    b = np.ascontiguousarray(b, np.float32)
    return lib().oref_probe_norm_l1_diff(_p(a), _p(b), a.shape[0], a.shape[1])


def probe_expr(which, R, p, t, s=1.0):
    R = np.ascontiguousarray(R, np.float32)
    p = np.ascontiguousarray(p, np.float32)
    t = np.ascontiguousarray(t, np.float32)
    out = np.zeros((3, 3) if which in (2, 3, 4) else (3, 1), np.float32)
    lib().oref_probe_expr(which, _p(R), _p(p), _p(t), s, _p(out))
    return out


def probe_undistort(pts, cam, dist):
    pts = np.ascontiguousarray(pts, np.float32).reshape(-1, 2)
    cam = np.asarray(cam, np.float32)
    dist = np.asarray(dist, np.float32)
    out = np.zeros_like(pts)
    lib().oref_probe_undistort(_p(pts), len(pts), _p(cam), _p(dist), len(dist), _p(out))
    return out
