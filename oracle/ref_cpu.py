"""ctypes binding of oracle/liborb_ref.so (the plain-C CPU oracle).

TEST INFRASTRUCTURE, NOT PRODUCT: imported only by tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / --impl reference legs.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])
CAND_DTYPE = np.dtype([("x", "<i4"), ("y", "<i4"), ("resp", "<i4")])
TOP2_DTYPE = np.dtype([("best_dist", "<i4"), ("best_idx", "<i4"), ("second_dist", "<i4"), ("second_idx", "<i4")])

_lib = None


def build():
    subprocess.check_call(["make", "-s", "-C", HERE])


def lib():
    global _lib
    if _lib is not None:
        return _lib
    so = os.path.join(HERE, "liborb_ref.so")
    if not os.path.exists(so):
        build()
    try:
        L = C.CDLL(so)
    except OSError:
        build()
        L = C.CDLL(so)
    vp, ip, fp, u8p = C.c_void_p, C.c_int, C.c_float, C.c_void_p
    L.ref_extractor_create.restype = vp
    L.ref_extractor_create.argtypes = [ip, fp, ip, ip, ip]
    L.ref_extractor_destroy.argtypes = [vp]
    L.ref_level_tables.argtypes = [vp] + [vp] * 6
    L.ref_extract.restype = ip
    L.ref_extract.argtypes = [vp, u8p, ip, ip, ip, vp, vp, ip]
    L.ref_level_size.argtypes = [vp, ip, vp, vp]
    L.ref_level_image.restype = vp
    L.ref_level_image.argtypes = [vp, ip]
    L.ref_level_blurred.restype = vp
    L.ref_level_blurred.argtypes = [vp, ip]
    L.ref_level_candidates.restype = ip
    L.ref_level_candidates.argtypes = [vp, ip, vp, ip]
    L.ref_level_nkeypoints.restype = ip
    L.ref_level_nkeypoints.argtypes = [vp, ip]
    L.ref_resize_linear_u8.argtypes = [vp, ip, ip, ip, vp, ip, ip, ip]
    L.ref_gaussian7_u8.argtypes = [vp, ip, ip, ip, vp, ip]
    L.ref_fast9_nms.restype = ip
    L.ref_fast9_nms.argtypes = [vp, ip, ip, ip, ip, vp, ip]
    L.ref_fast_atan2.restype = fp
    L.ref_fast_atan2.argtypes = [fp, fp]
    L.ref_distribute_octree.restype = ip
    L.ref_distribute_octree.argtypes = [vp, ip, ip, ip, ip, ip, ip, vp, ip]
    L.ref_descriptor_distance.restype = ip
    L.ref_descriptor_distance.argtypes = [vp, vp]
    L.ref_bf_top2.argtypes = [vp, ip, vp, ip, vp, ip]
    L.ref_match_candidates.argtypes = [vp, ip, vp, vp, vp, vp]
    L.ref_search_triangulation.argtypes = [vp, vp, vp, ip, vp, vp, vp, vp, vp, vp, fp, fp, vp, vp, ip, vp, vp]
    L.ref_ratio_accept.restype = ip
    L.ref_distinctive_descriptors.argtypes = [vp, vp, ip, vp, vp]
    L.ref_ratio_accept.argtypes = [ip, ip, ip, fp]
    L.ref_assign_grid.argtypes = [vp, ip, fp, fp, fp, fp, ip, ip, vp, vp]
    L.ref_features_in_area.restype = ip
    L.ref_features_in_area.argtypes = [vp, vp, vp, fp, fp, fp, fp, ip, ip, fp, fp, fp, ip, ip, vp, ip]
    L.ref_stereo_match.restype = ip
    L.ref_stereo_match.argtypes = [vp, vp, ip, vp, vp, ip, vp, vp, vp, vp, vp, vp, fp, fp, vp, vp, vp]
    _lib = L
    return L


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class RefExtractor:
    """Mirror of ORB_SLAM2::ORBextractor on the CPU oracle."""

    def __init__(self, nfeatures, scale_factor, nlevels, ini_th, min_th):
        self.L = lib()
        self.h = self.L.ref_extractor_create(nfeatures, scale_factor, nlevels, ini_th, min_th)
        assert self.h
        self.nlevels, self.nfeatures = nlevels, nfeatures

    def __del__(self):
        try:
            self.L.ref_extractor_destroy(self.h)
        except Exception:
            pass

    def tables(self):
        n = self.nlevels
        sf, isf, s2, is2 = (np.zeros(n, np.float32) for _ in range(4))
        nf, um = np.zeros(n, np.int32), np.zeros(16, np.int32)
        self.L.ref_level_tables(self.h, _p(sf), _p(isf), _p(s2), _p(is2), _p(nf), _p(um))
        return dict(scale=sf, inv_scale=isf, sigma2=s2, inv_sigma2=is2, nfeat=nf, umax=um)

    def extract(self, img):
        img = np.ascontiguousarray(img, dtype=np.uint8)
        h, w = img.shape
        cap = self.nfeatures + 8 * self.nlevels + 64
        kps = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        n = self.L.ref_extract(self.h, _p(img), w, h, w, _p(kps), _p(desc), cap)
        assert n >= 0
        return kps[:n].copy(), desc[:n].copy()

    def level_size(self, l):
        w, h = C.c_int(), C.c_int()
        self.L.ref_level_size(self.h, l, C.byref(w), C.byref(h))
        return w.value, h.value

    def _img(self, ptr, l):
        w, h = self.level_size(l)
        if not ptr:
            return None
        return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_uint8)), shape=(h, w)).copy()

    def level_image(self, l):
        return self._img(self.L.ref_level_image(self.h, l), l)

    def level_blurred(self, l):
        return self._img(self.L.ref_level_blurred(self.h, l), l)

    def level_candidates(self, l):
        n = self.L.ref_level_candidates(self.h, l, None, 0)
        out = np.zeros(max(n, 1), CAND_DTYPE)
        self.L.ref_level_candidates(self.h, l, _p(out), n)
        return out[:n]

    def level_nkeypoints(self, l):
        return self.L.ref_level_nkeypoints(self.h, l)


def resize(src, dw, dh):
    src = np.ascontiguousarray(src, np.uint8)
    dst = np.zeros((dh, dw), np.uint8)
    lib().ref_resize_linear_u8(_p(src), src.shape[1], src.shape[0], src.shape[1], _p(dst), dw, dh, dw)
    return dst


def gaussian7(src):
    src = np.ascontiguousarray(src, np.uint8)
    dst = np.zeros_like(src)
    lib().ref_gaussian7_u8(_p(src), src.shape[1], src.shape[0], src.shape[1], _p(dst), src.shape[1])
    return dst


def fast9(roi, th):
    roi = np.ascontiguousarray(roi, np.uint8)
    out = np.zeros(roi.size // 4 + 16, CAND_DTYPE)
    n = lib().ref_fast9_nms(_p(roi), roi.shape[1], roi.shape[1], roi.shape[0], th, _p(out), len(out))
    return out[:n]


def fast_atan2(y, x):
    return np.float32(lib().ref_fast_atan2(float(y), float(x)))


def distribute(cand, minX, maxX, minY, maxY, N):
    cand = np.ascontiguousarray(cand, CAND_DTYPE)
    out = np.zeros(len(cand) + 8, np.int32)
    n = lib().ref_distribute_octree(_p(cand), len(cand), minX, maxX, minY, maxY, N, _p(out), len(out))
    return out[:n]


def descriptor_distance(a, b):
    a, b = np.ascontiguousarray(a, np.uint8), np.ascontiguousarray(b, np.uint8)
    return lib().ref_descriptor_distance(_p(a), _p(b))


def bf_top2(q, db, use_popcnt=False):
    q, db = np.ascontiguousarray(q, np.uint8), np.ascontiguousarray(db, np.uint8)
    out = np.zeros(len(q), TOP2_DTYPE)
    lib().ref_bf_top2(_p(q), len(q), _p(db), len(db), _p(out), int(use_popcnt))
    return out


def match_candidates(q, t, offs, idx):
    q, t = np.ascontiguousarray(q, np.uint8), np.ascontiguousarray(t, np.uint8)
    offs, idx = np.ascontiguousarray(offs, np.int32), np.ascontiguousarray(idx, np.int32)
    out = np.zeros(len(q), TOP2_DTYPE)
    lib().ref_match_candidates(_p(q), len(q), _p(t), _p(offs), _p(idx), _p(out))
    return out


def search_triangulation(q, qxy, qstereo, t, tkp, tstereo, offs, idx, F12, ex, ey, sf, sig2, th_low=50):
    q, t = np.ascontiguousarray(q, np.uint8), np.ascontiguousarray(t, np.uint8)
    qxy = np.ascontiguousarray(qxy, np.float32)
    qstereo, tstereo = np.ascontiguousarray(qstereo, np.uint8), np.ascontiguousarray(tstereo, np.uint8)
    tkp = np.ascontiguousarray(tkp, KP_DTYPE)
    offs, idx = np.ascontiguousarray(offs, np.int32), np.ascontiguousarray(idx, np.int32)
    F12, sf, sig2 = np.ascontiguousarray(F12, np.float32), np.ascontiguousarray(sf, np.float32), np.ascontiguousarray(sig2, np.float32)
    bi, bd = np.zeros(len(q), np.int32), np.zeros(len(q), np.int32)
    lib().ref_search_triangulation(_p(q), _p(qxy), _p(qstereo), len(q), _p(t), _p(tkp), _p(tstereo), _p(offs), _p(idx), _p(F12), ex, ey, _p(sf),
                                   _p(sig2), th_low, _p(bi), _p(bd))
    return bi, bd


def distinctive_descriptors(desc, offsets):
    """MapPoint::ComputeDistinctiveDescriptors for a batch of groups -> (best row per group, its median)"""
    desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    offsets = np.ascontiguousarray(offsets, np.int32)
    g = len(offsets) - 1
    idx = np.zeros(g, np.int32)
    med = np.zeros(g, np.int32)
    lib().ref_distinctive_descriptors(_p(desc), _p(offsets), g, _p(idx), _p(med))
    return idx, med


def assign_grid(kps, minX, minY, invW, invH, cols=64, rows=48):
    kps = np.ascontiguousarray(kps, KP_DTYPE)
    start = np.zeros(cols * rows + 1, np.int32)
    items = np.zeros(max(len(kps), 1), np.int32)
    lib().ref_assign_grid(_p(kps), len(kps), minX, minY, invW, invH, cols, rows, _p(start), _p(items))
    return start, items[:start[-1]]


def features_in_area(kps, start, items, minX, minY, invW, invH, x, y, r, min_level=-1, max_level=-1, cols=64, rows=48):
    kps = np.ascontiguousarray(kps, KP_DTYPE)
    items_c = np.ascontiguousarray(items if len(items) else np.zeros(1, np.int32), np.int32)
    out = np.zeros(max(len(kps), 1), np.int32)
    n = lib().ref_features_in_area(_p(kps), _p(start), _p(items_c), minX, minY, invW, invH, cols, rows, x, y, r,
                                   min_level, max_level, _p(out), len(out))
    return out[:n]


def stereo_match(kL, dL, kR, dR, pyrL, pyrR, sf, isf, bf, b):
    kL, kR = np.ascontiguousarray(kL, KP_DTYPE), np.ascontiguousarray(kR, KP_DTYPE)
    dL, dR = np.ascontiguousarray(dL, np.uint8), np.ascontiguousarray(dR, np.uint8)
    nl = len(pyrL)
    pl = [np.ascontiguousarray(p, np.uint8) for p in pyrL]
    pr = [np.ascontiguousarray(p, np.uint8) for p in pyrR]
    PL = (C.c_void_p * nl)(*[p.ctypes.data for p in pl])
    PR = (C.c_void_p * nl)(*[p.ctypes.data for p in pr])
    lw = np.array([p.shape[1] for p in pl], np.int32)
    lh = np.array([p.shape[0] for p in pl], np.int32)
    sf, isf = np.ascontiguousarray(sf, np.float32), np.ascontiguousarray(isf, np.float32)
    uR, depth = np.zeros(len(kL), np.float32), np.zeros(len(kL), np.float32)
    best = np.zeros(max(len(kL), 1), np.int32)
    n = lib().ref_stereo_match(_p(kL), _p(dL), len(kL), _p(kR), _p(dR), len(kR), PL, PR, _p(lw), _p(lh), _p(sf),
                               _p(isf), bf, b, _p(uR), _p(depth), _p(best))
    return uR, depth, best[:len(kL)], n


def remap_linear(src, mapx, mapy):
    """cv::remap(src, map1, map2, INTER_LINEAR), CV_32FC1 maps, BORDER_CONSTANT 0 (the EuRoC example's rectification)"""
    src = np.ascontiguousarray(src, np.uint8)
    mapx, mapy = np.ascontiguousarray(mapx, np.float32), np.ascontiguousarray(mapy, np.float32)
    dh, dw = mapx.shape
    dst = np.zeros((dh, dw), np.uint8)
    lib().ref_remap_linear_u8(_p(src), src.shape[1], src.shape[0], src.shape[1], _p(mapx), _p(mapy), dw, dh, _p(dst), dw)
    return dst


def init_undistort_rectify_map(K, D, R, P, w, h):
    """cv::initUndistortRectifyMap(K, D = (k1, k2, p1, p2, k3), R, P[:3, :3], (w, h), CV_32F) -> (map1, map2)"""
    K = np.ascontiguousarray(np.asarray(K, np.float64)[:3, :3].reshape(-1))
    R = np.ascontiguousarray(np.asarray(R, np.float64)[:3, :3].reshape(-1))
    P = np.ascontiguousarray(np.asarray(P, np.float64)[:3, :3].reshape(-1))
    d = np.asarray(D, np.float64).reshape(-1)
    D5 = np.zeros(5, np.float64)
    D5[:min(5, len(d))] = d[:5]
    mx, my = np.zeros((h, w), np.float32), np.zeros((h, w), np.float32)
    lib().ref_init_undistort_rectify_map(_p(K), _p(D5), _p(R), _p(P), w, h, _p(mx), _p(my))
    return mx, my
