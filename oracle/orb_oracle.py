"""cv2-backed restatement of ORBextractor::operator() -- the PIN for oracle/orb_ref.c.

TEST INFRASTRUCTURE, NOT PRODUCT.  Imported only from tests/ and tools/.

The reference (/root/reference/src/ORBextractor.cc) delegates resize / FAST /
GaussianBlur / fastAtan2 to OpenCV.  No OpenCV C++ exists on this image, so the
untouched reference cannot be compiled; python cv2 4.13.0 *is* present.  This
module calls cv2 for exactly those four library calls and restates the
reference-own logic (tables :429-534, cell loop :925-1052, quadtree :551-915,
IC_Angle :83-111, computeOrbDescriptor :116-157, assembly :1261-1339) in numpy
float32.  tests/test_oracle_pin.py requires oracle/orb_ref.c (plain C, no cv2)
to agree with it on every output; golden fixtures under tests/golden/ are
generated from it by tools/make_golden.py.
"""
import math
import os
import re

import numpy as np

try:
    import cv2
except Exception:  # pragma: no cover
    cv2 = None

f32 = np.float32
HALF_PATCH = 15
EDGE = 19


def _pattern():
    here = os.path.dirname(os.path.abspath(__file__))
    txt = open(os.path.join(here, "orb_pattern.inc")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    nums = [int(v) for v in re.findall(r"-?\d+", txt)]
    assert len(nums) == 1024
    return np.array(nums, dtype=np.int32).reshape(256, 4)


PATTERN = _pattern()


def cv_round(v):
    return int(np.rint(v))


class Extractor:
    def __init__(self, nfeatures, scale_factor, nlevels, ini_th, min_th):
        self.nfeatures, self.nlevels, self.ini_th, self.min_th = nfeatures, nlevels, ini_th, min_th
        self.scale_factor = float(f32(scale_factor))  # double member initialised from a float arg
        sf = [f32(1.0)]
        for i in range(1, nlevels):
            sf.append(f32(float(sf[-1]) * self.scale_factor))
        self.sf = np.array(sf, dtype=f32)
        self.sig2 = (self.sf * self.sf).astype(f32)
        self.isf = (f32(1.0) / self.sf).astype(f32)
        self.isig2 = (f32(1.0) / self.sig2).astype(f32)
        factor = f32(1.0 / self.scale_factor)
        nd = f32(f32(nfeatures) * (f32(1) - factor)) / (f32(1) - f32(math.pow(float(factor), float(nlevels))))
        nd = f32(nd)
        self.nfeat, s = [], 0
        for _ in range(nlevels - 1):
            self.nfeat.append(cv_round(nd))
            s += self.nfeat[-1]
            nd = f32(nd * factor)
        self.nfeat.append(max(nfeatures - s, 0))
        umax = [0] * (HALF_PATCH + 1)
        vmax = int(math.floor(float(f32(HALF_PATCH) * f32(math.sqrt(2.0)) / f32(2) + f32(1))))
        vmin = int(math.ceil(float(f32(HALF_PATCH) * f32(math.sqrt(2.0)) / f32(2))))
        for v in range(vmax + 1):
            umax[v] = cv_round(math.sqrt(HALF_PATCH * HALF_PATCH - v * v))
        v0 = 0
        for v in range(HALF_PATCH, vmin - 1, -1):
            while umax[v0] == umax[v0 + 1]:
                v0 += 1
            umax[v] = v0
            v0 += 1
        self.umax = umax

    # -- E2 ComputePyramid :1345-1394 ------------------------------------
    def pyramid(self, image):
        h, w = image.shape
        levels = []
        for l in range(self.nlevels):
            scale = self.isf[l]
            sz = (cv_round(f32(w) * scale), cv_round(f32(h) * scale))
            if l == 0:
                levels.append(image.copy())
            else:
                levels.append(cv2.resize(levels[l - 1], sz, interpolation=cv2.INTER_LINEAR))
        return levels

    # -- E3/E4 cell loop :925-1009 ----------------------------------------
    def detect_level(self, img):
        H, W = img.shape
        minB = EDGE - 3
        maxBX, maxBY = W - EDGE + 3, H - EDGE + 3
        width, height = f32(maxBX - minB), f32(maxBY - minB)
        nCols, nRows = int(width / f32(30)), int(height / f32(30))
        out = []
        if nCols < 1 or nRows < 1:
            return out
        wCell = int(math.ceil(float(width / f32(nCols))))
        hCell = int(math.ceil(float(height / f32(nRows))))
        det_ini = cv2.FastFeatureDetector_create(self.ini_th, True)
        det_min = cv2.FastFeatureDetector_create(self.min_th, True)
        for i in range(nRows):
            iniY = minB + i * hCell
            maxY = iniY + hCell + 6
            if iniY >= maxBY - 3:
                continue
            maxY = min(maxY, maxBY)
            for j in range(nCols):
                iniX = minB + j * wCell
                maxX = iniX + wCell + 6
                if iniX >= maxBX - 6:
                    continue
                maxX = min(maxX, maxBX)
                roi = np.ascontiguousarray(img[iniY:maxY, iniX:maxX])
                kps = det_ini.detect(roi)
                if len(kps) == 0:
                    kps = det_min.detect(roi)
                for kp in kps:
                    out.append((int(kp.pt[0]) + j * wCell, int(kp.pt[1]) + i * hCell, int(kp.response)))
        return out

    # -- E5 DistributeOctTree :617-915 ------------------------------------
    @staticmethod
    def distribute(cand, minX, maxX, minY, maxY, N):
        class Node:
            __slots__ = ("x0", "x1", "y0", "y1", "keys", "nomore", "seq")
        seq = [0]

        def new(x0, x1, y0, y1):
            n = Node()
            n.x0, n.x1, n.y0, n.y1, n.keys, n.nomore, n.seq = x0, x1, y0, y1, [], False, seq[0]
            seq[0] += 1
            return n

        def divide(p):
            hx = int(math.ceil(float(f32(p.x1 - p.x0) / f32(2))))
            hy = int(math.ceil(float(f32(p.y1 - p.y0) / f32(2))))
            n1 = new(p.x0, p.x0 + hx, p.y0, p.y0 + hy)
            n2 = new(p.x0 + hx, p.x1, p.y0, p.y0 + hy)
            n3 = new(p.x0, p.x0 + hx, p.y0 + hy, p.y1)
            n4 = new(p.x0 + hx, p.x1, p.y0 + hy, p.y1)
            for k in p.keys:
                x, y, _ = cand[k]
                if x < n1.x1:
                    (n1 if y < n1.y1 else n3).keys.append(k)
                else:
                    (n2 if y < n1.y1 else n4).keys.append(k)
            for n in (n1, n2, n3, n4):
                if len(n.keys) == 1:
                    n.nomore = True
            return n1, n2, n3, n4

        nIni = int(math.floor(float(f32(maxX - minX) / f32(maxY - minY)) + 0.5))  # C round(): half away from zero (positive arg)
        nIni = max(nIni, 1)
        hX = f32(maxX - minX) / f32(nIni)
        nodes = []
        for i in range(nIni):
            nodes.append(new(int(hX * f32(i)), int(hX * f32(i + 1)), 0, maxY - minY))
        for k, (x, y, r) in enumerate(cand):
            nodes[min(int(f32(x) / hX), nIni - 1)].keys.append(k)
        lst = []
        for n in nodes:
            if len(n.keys) == 1:
                n.nomore = True
                lst.append(n)
            elif len(n.keys) > 1:
                lst.append(n)
        finish = False
        while not finish:
            prev = len(lst)
            vs = []
            nexp = 0
            old = list(lst)
            for n in old:
                if n.nomore:
                    continue
                for c in divide(n):
                    if len(c.keys) > 0:
                        lst.insert(0, c)
                        if len(c.keys) > 1:
                            nexp += 1
                            vs.append(c)
                lst.remove(n)
            if len(lst) >= N or len(lst) == prev:
                finish = True
            elif len(lst) + nexp * 3 > N:
                while not finish:
                    prev = len(lst)
                    vprev = sorted(vs, key=lambda n: (len(n.keys), n.seq))
                    vs = []
                    for n in reversed(vprev):
                        for c in divide(n):
                            if len(c.keys) > 0:
                                lst.insert(0, c)
                                if len(c.keys) > 1:
                                    vs.append(c)
                        lst.remove(n)
                        if len(lst) >= N:
                            break
                    if len(lst) >= N or len(lst) == prev:
                        finish = True
        out = []
        for n in lst:
            best = n.keys[0]
            for k in n.keys[1:]:
                if cand[k][2] > cand[best][2]:
                    best = k
            out.append(best)
        return out

    # -- E6 IC_Angle :83-111 ------------------------------------------------
    def ic_angle(self, img, x, y):
        m01 = m10 = 0
        im = img.astype(np.int64)
        for u in range(-HALF_PATCH, HALF_PATCH + 1):
            m10 += u * int(im[y, x + u])
        for v in range(1, HALF_PATCH + 1):
            d = self.umax[v]
            vs = 0
            for u in range(-d, d + 1):
                vp, vm = int(im[y + v, x + u]), int(im[y - v, x + u])
                vs += vp - vm
                m10 += u * (vp + vm)
            m01 += v * vs
        return f32(cv2.fastAtan2(float(f32(m01)), float(f32(m10))))

    # -- E8 computeOrbDescriptor :116-157 ---------------------------------
    @staticmethod
    def descriptor(blur, x, y, angle_deg):
        factor_pi = f32(math.pi / 180.0)
        ang = f32(f32(angle_deg) * factor_pi)
        a, b = f32(math.cos(float(ang))), f32(math.sin(float(ang)))  # libm cosf/sinf are correctly rounded here
        px0, py0, px1, py1 = (PATTERN[:, k].astype(f32) for k in range(4))
        r0 = np.rint(px0 * b + py0 * a).astype(np.int64)
        c0 = np.rint(px0 * a - py0 * b).astype(np.int64)
        r1 = np.rint(px1 * b + py1 * a).astype(np.int64)
        c1 = np.rint(px1 * a - py1 * b).astype(np.int64)
        bits = (blur[y + r0, x + c0] < blur[y + r1, x + c1]).astype(np.uint8)
        return np.packbits(bits.reshape(32, 8), axis=1, bitorder="little").reshape(32)

    # -- E9 operator() :1261-1339 -------------------------------------------
    def extract(self, image, stages=None):
        levels = self.pyramid(image)
        kps, descs = [], []
        per_level = []
        for l, img in enumerate(levels):
            H, W = img.shape
            cand = self.detect_level(img)
            sel = self.distribute(cand, EDGE - 3, W - EDGE + 3, EDGE - 3, H - EDGE + 3, self.nfeat[l])
            size = f32(int(f32(31) * self.sf[l]))
            lk = []
            for k in sel:
                x, y, r = cand[k]
                x += EDGE - 3
                y += EDGE - 3
                lk.append([x, y, size, self.ic_angle(img, x, y), f32(r), l])
            per_level.append((cand, sel, lk))
        blurred = []
        for l, img in enumerate(levels):
            cand, sel, lk = per_level[l]
            if not lk:
                blurred.append(None)
                continue
            blur = cv2.GaussianBlur(img, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
            blurred.append(blur)
            for (x, y, size, ang, r, lev) in lk:
                descs.append(self.descriptor(blur, x, y, ang))
                fx, fy = f32(x), f32(y)
                if l != 0:
                    fx, fy = f32(fx * self.sf[l]), f32(fy * self.sf[l])
                kps.append((fx, fy, size, ang, r, lev, -1))
        if stages is not None:
            stages["levels"] = levels
            stages["blurred"] = blurred
            stages["cand"] = [p[0] for p in per_level]
            stages["sel"] = [p[1] for p in per_level]
        kp_arr = np.zeros(len(kps), dtype=KP_DTYPE)
        for i, k in enumerate(kps):
            kp_arr[i] = k
        d = np.stack(descs) if descs else np.zeros((0, 32), np.uint8)
        return kp_arr, d


KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])
