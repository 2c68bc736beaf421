"""CPU tests that PIN the oracle to the reference's own text.

oracle/_ref/liborbslam_ref.so is the reference's UNMODIFIED src/{ORBextractor,ORBmatcher,Frame,KeyFrame,MapPoint,Map}.cc
compiled by oracle/Makefile against the OpenCV stand-in in oracle/shim (this image has no OpenCV C++).  These tests check
  (1) the stand-in's CV_32F algebra and undistortPoints against the real OpenCV on the image (cv2 4.13),
  (2) the plain-C restatement oracle/orb_ref.c -- the checker every GPU parity test uses -- against that library, live:
      ORBextractor::operator() (keypoints IN ORDER and set-wise, angles, descriptors, pyramid), DistributeOctTree on raw
      candidate lists, the stereo Frame constructor (ComputeStereoMatches, UndistortKeyPoints, AssignFeaturesToGrid,
      GetFeaturesInArea), MapPoint::ComputeDistinctiveDescriptors, DescriptorDistance,
  (3) how often glibc's heap-address order (ORBextractor.cc:832 sorts by node POINTER) changes the selection, against the
      creation-order rule the restatement and the CUDA path implement.
The library is prebuilt in the container and travels to the GPU box; nothing here reads /root/reference.
"""
import os

import numpy as np
import pytest

import orbslam_ref as R
import prologue_oracle as po
import ref_cpu
import synth

pytestmark = pytest.mark.skipif(not R.available(), reason="oracle/_ref/liborbslam_ref.so not built (needs /root/reference)")

EDGE = [
    ((200, 150), (300, 1.2, 6, 20, 7)),      # small image, few cells
    ((333, 517), (500, 1.3, 5, 15, 5)),      # portrait, other scale factor and thresholds
    ((1000, 90), (400, 1.2, 4, 12, 7)),      # panorama: many quadtree roots
    ((641, 479), (5000, 1.2, 8, 20, 7)),     # quota above the candidate count (quadtree runs dry)
    ((640, 480), (40, 1.2, 8, 20, 7)),       # tiny quota (N < initial nodes)
    ((1241, 376), (100, 1.2, 8, 20, 7)),     # wide image, small quota: first split yields more nodes than N
    ((1241, 376), (40, 1.2, 8, 20, 7)),
]


def _same(kr, dr, kc, dc):
    assert len(kr) == len(kc)
    for f in kr.dtype.names:
        assert np.array_equal(kr[f], kc[f]), f
    assert np.array_equal(dr, dc)


def _sorted(k, d):
    o = np.lexsort((k["x"], k["y"], k["octave"]))
    return k[o], d[o]


# ---- (1) the OpenCV stand-in against OpenCV ----------------------------------------------------------------------------
def test_shim_gemm_matches_cv2():
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(5)
    shapes = [((3, 3), (3, 1)), ((3, 3), (3, 3)), ((4, 4), (4, 1)), ((4, 4), (4, 4)), ((3, 4), (4, 1)), ((2, 2), (2, 2)), ((1, 3), (3, 1)),
              ((5, 5), (5, 1)), ((3, 5), (5, 2))]
    n = 0
    for (sa, sb) in shapes:
        for flags in (0, 1, 2, 3):
            for it in range(60):
                A = (rng.standard_normal(sa) * 3).astype(np.float32)
                B = (rng.standard_normal(sb) * 50).astype(np.float32)
                m = A.shape[1] if flags & 1 else A.shape[0]
                ka = A.shape[0] if flags & 1 else A.shape[1]
                kb = B.shape[1] if flags & 2 else B.shape[0]
                nn = B.shape[0] if flags & 2 else B.shape[1]
                if ka != kb:
                    continue
                for (alpha, beta, withc) in ((1.0, 0.0, False), (-1.0, 0.0, False), (1.0, 1.0, True), (0.37, -1.0, True)):
                    Cm = (rng.standard_normal((m, nn)) * 5).astype(np.float32) if withc else None
                    want = cv2.gemm(A, B, alpha, Cm, beta, flags=flags)
                    got = R.probe_gemm(A, B, alpha, Cm, beta, flags)
                    assert np.array_equal(want, got), (sa, sb, flags, alpha, beta)
                    n += 1
    assert n > 2000


def test_shim_expressions_match_cv2():
    """the expression forms ORBmatcher.cc / Frame.cc / KeyFrame.cc write, against the cv2 calls OpenCV's MatExpr folds them into"""
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(6)
    for it in range(400):
        Rm = rng.standard_normal((3, 3)).astype(np.float32)
        p = (rng.standard_normal((3, 1)) * 20).astype(np.float32)
        t = (rng.standard_normal((3, 1)) * 2).astype(np.float32)
        s = np.float32(rng.uniform(0.3, 3.0))
        assert np.array_equal(R.probe_expr(0, Rm, p, t), cv2.gemm(Rm, p, 1, t, 1))                    # R*p + t: one gemm
        assert np.array_equal(R.probe_expr(1, Rm, p, t), cv2.gemm(Rm, t, -1, None, 0, flags=cv2.GEMM_1_T))  # -R.t()*t
        # M/s, s*M and alpha*M.t() evaluate through Mat::convertTo(dst, -1, alpha), which Python does not expose for CV_32F;
        # OpenCV's statement for a CV_32F destination is dst = src * (float)alpha with alpha = 1/s formed once in double
        inv = 1.0 / float(s)
        assert np.array_equal(R.probe_expr(2, Rm, p, t, s), (Rm * np.float32(inv)).astype(np.float32))
        assert np.array_equal(R.probe_expr(3, Rm, p, t, s), (Rm * s).astype(np.float32))               # s*M
        assert np.array_equal(R.probe_expr(4, Rm, p, t, s), (cv2.transpose(Rm) * np.float32(inv)).astype(np.float32))
        assert np.array_equal(R.probe_expr(5, Rm, p, t), cv2.gemm(Rm, t, -1, None, 0))                 # -R*t
        v = (rng.standard_normal(3) * 10).astype(np.float32)
        assert R.probe_norm(v) == cv2.norm(v.reshape(3, 1))
        a = rng.integers(-255, 256, (11, 11)).astype(np.float32)
        b = rng.integers(-255, 256, (11, 11)).astype(np.float32)
        assert R.probe_norm_l1_diff(a, b) == cv2.norm(a, b, cv2.NORM_L1)


def test_shim_undistort_matches_cv2():
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(8)
    for cam, dist in [((517.306408, 516.469215, 318.643040, 255.313989), (0.262383, -0.953104, -0.005358, 0.002628, 1.163314)),
                      ((520.908620, 521.007327, 325.141442, 249.701764), (0.231222, -0.784899, -0.003257, -0.000105, 0.917205)),
                      ((458.654, 457.296, 367.215, 248.375), (-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05))]:
        pts = np.stack([rng.uniform(0, 752, 500), rng.uniform(0, 480, 500)], axis=1).astype(np.float32)
        K = np.array([[cam[0], 0, cam[2]], [0, cam[1], cam[3]], [0, 0, 1]], np.float32)
        want = cv2.undistortPoints(pts.reshape(-1, 1, 2), K, np.asarray(dist, np.float32), None, None, K).reshape(-1, 2)
        got = R.probe_undistort(pts, cam, dist)
        assert np.array_equal(want, got)
        if len(dist) == 5:
            assert np.array_equal(po.undistort_points(pts, cam, dist), got)


# ---- (2) the restatement against the reference's own code ---------------------------------------------------------------
def test_tables_equal_reference():
    for (nf, sfac, nl, ini, mn) in [(1000, 1.2, 8, 20, 7), (2000, 1.2, 8, 20, 7), (1200, 1.2, 8, 20, 7), (8000, 1.2, 12, 20, 7), (500, 1.3, 5, 15, 5),
                                    (40, 1.2, 8, 20, 7), (5000, 1.5, 3, 20, 7)]:
        a = R.Extractor(nf, sfac, nl, ini, mn).tables()
        b = ref_cpu.RefExtractor(nf, sfac, nl, ini, mn).tables()
        for k in a:
            assert np.array_equal(a[k], b[k]), (k, nf, sfac, nl)
        assert a["nfeat"].sum() == nf


@pytest.mark.parametrize("name", ["tum", "kitti", "euroc"])
def test_extract_equals_reference(configs, name):
    w, h, nf, sfac, nl, ini, mn = configs[name]
    for seed in (1000, 31):
        img = synth.textured_frame(w, h, seed)
        er = R.Extractor(nf, sfac, nl, ini, mn)
        kr, dr = er.extract(img)
        ec = ref_cpu.RefExtractor(nf, sfac, nl, ini, mn)
        kc, dc = ec.extract(img)
        assert len(kr) >= nf
        _same(kr, dr, kc, dc)                      # in order: x, y, size, angle, response, octave bit-equal, descriptors byte-equal
        for l in range(nl):
            assert np.array_equal(er.level_image(l), ec.level_image(l)), l


@pytest.mark.parametrize("shape,params", EDGE)
def test_extract_edge_shapes_equal_reference(shape, params):
    w, h = shape
    img = synth.textured_frame(w, h, 77)
    kr, dr = R.Extractor(*params).extract(img)
    kc, dc = ref_cpu.RefExtractor(*params).extract(img)
    _same(kr, dr, kc, dc)


def test_extract_4k_equals_reference():
    """BASELINE config 4 at full size: 3840x2160, 8000 features, 12 levels"""
    img = synth.textured_frame(3840, 2160, 4)
    kr, dr = R.Extractor(8000, 1.2, 12, 20, 7).extract(img)
    kc, dc = ref_cpu.RefExtractor(8000, 1.2, 12, 20, 7).extract(img)
    assert len(kr) >= 8000
    _same(kr, dr, kc, dc)


def test_flat_image_equals_reference():
    img = np.full((480, 640), 90, np.uint8)
    kr, dr = R.Extractor(1000, 1.2, 8, 20, 7).extract(img)
    kc, dc = ref_cpu.RefExtractor(1000, 1.2, 8, 20, 7).extract(img)
    assert len(kr) == 0 and len(kc) == 0


def test_distribute_octree_equals_reference():
    """DistributeOctTree called directly on candidate lists, incl. N below the root count, N above the candidate count,
    duplicate responses (first-max rule) and one-candidate lists.  (Width / height below 0.5 is left out: the reference
    then has zero root nodes and indexes an empty vector, ORBextractor.cc:624-652 -- undefined behaviour, nothing to pin.)"""
    rng = np.random.default_rng(12)
    er = R.Extractor(1000, 1.2, 8, 20, 7)
    cases = 0
    for (W, H, ncand, N) in [(608, 448, 3000, 217), (1209, 344, 6000, 434), (1209, 344, 6000, 2), (1209, 344, 50, 500), (300, 300, 1, 10),
                             (964, 52, 800, 120), (300, 400, 500, 64), (608, 448, 217, 217), (608, 448, 4000, 1), (2000, 100, 3000, 10)]:
        for rep in range(3):
            # distinct pixels in row-major order, the way the cell loop appends them (ORBextractor.cc:997-1004)
            pix = np.sort(rng.choice(W * H, size=min(ncand, W * H), replace=False))
            cand = np.zeros(len(pix), ref_cpu.CAND_DTYPE)
            cand["x"], cand["y"] = pix % W, pix // W
            cand["resp"] = rng.integers(7, 60 if rep else 12, len(pix))  # rep 0: many equal responses
            a = er.distribute(cand, 0, W, 0, H, N)
            b = ref_cpu.distribute(cand, 0, W, 0, H, N)
            assert len(a) == len(b) and len(a) > 0
            assert np.array_equal(a, b), (W, H, ncand, N, rep)
            cases += 1
    assert cases == 30


def test_pointer_tie_break_statistics(configs):
    """ORBextractor.cc:832 sorts (nKeys, node pointer): under plain glibc malloc the order among equal sizes follows the
    heap's reuse pattern.  Report how far that moves the result from the creation-order rule (set-wise), and make sure the
    difference is confined to the tie-break: everything else (count, per-level quota, responses of common points) agrees."""
    w, h, nf, sfac, nl, ini, mn = configs["kitti"]
    img = synth.textured_frame(w, h, 1000)
    kc, dc = ref_cpu.RefExtractor(nf, sfac, nl, ini, mn).extract(img)
    try:
        R.set_monotonic(0)
        km, dm = R.Extractor(nf, sfac, nl, ini, mn).extract(img)
    finally:
        R.set_monotonic(1)
    assert len(km) == len(kc)
    assert np.array_equal(np.bincount(km["octave"], minlength=nl), np.bincount(kc["octave"], minlength=nl))
    sm = set(zip(km["x"].tolist(), km["y"].tolist(), km["octave"].tolist()))
    sc = set(zip(kc["x"].tolist(), kc["y"].tolist(), kc["octave"].tolist()))
    common = len(sm & sc)
    print("pointer tie-break under glibc malloc: %d of %d keypoints in common with the creation-order rule" % (common, len(kc)))
    assert common >= 0.97 * len(kc)
    # a point selected under both rules carries identical angle / response / descriptor
    dmap = {(x, y, o): (a, r, d.tobytes()) for x, y, o, a, r, d in zip(kc["x"].tolist(), kc["y"].tolist(), kc["octave"].tolist(), kc["angle"].tolist(),
                                                                      kc["response"].tolist(), dc)}
    for x, y, o, a, r, d in zip(km["x"].tolist(), km["y"].tolist(), km["octave"].tolist(), km["angle"].tolist(), km["response"].tolist(), dm):
        if (x, y, o) in dmap:
            assert dmap[(x, y, o)] == (a, r, d.tobytes())


def _stereo_inputs(w, h, seed):
    left, right = synth.stereo_pair(w, h, seed)
    return left, right


@pytest.mark.parametrize("name,seed", [("kitti", 1001), ("kitti", 1107), ("euroc", 5)])
def test_stereo_frame_equals_reference(configs, name, seed):
    """Frame::Frame(imLeft, imRight, ...) (Frame.cc:65-125): ExtractORB on two threads, ComputeStereoMatches, AssignFeaturesToGrid"""
    w, h, nf, sfac, nl, ini, mn = configs[name]
    left, right = _stereo_inputs(w, h, seed)
    bf, fx = 386.1448, 718.856
    f = R.RefFrame(left, right, nf, sfac, nl, ini, mn, cam=(fx, fx, 607.1928, 185.2157), dist=(0, 0, 0, 0), bf=bf)
    kL, dL = f.keys(0)
    kR, dR = f.keys(1)
    exL, exR = ref_cpu.RefExtractor(nf, sfac, nl, ini, mn), ref_cpu.RefExtractor(nf, sfac, nl, ini, mn)
    cL, cdL = exL.extract(left)
    cR, cdR = exR.extract(right)
    _same(kL, dL, cL, cdL)
    _same(kR, dR, cR, cdR)
    t = exL.tables()
    pyrL = [exL.level_image(l) for l in range(nl)]
    pyrR = [exR.level_image(l) for l in range(nl)]
    b = np.float32(bf) / np.float32(fx)  # mb = mbf / fx (Frame.cc:117)
    uR, depth, _, nvalid = ref_cpu.stereo_match(cL, cdL, cR, cdR, pyrL, pyrR, t["scale"], t["inv_scale"], np.float32(bf), b)
    ruR, rdepth = f.stereo()
    assert (ruR >= 0).sum() > 100
    assert np.array_equal(ruR, uR) and np.array_equal(rdepth, depth)
    # k1 == 0: mvKeysUn = mvKeys (Frame.cc:473-477); grid and GetFeaturesInArea
    kU, _ = f.keys(2)
    assert np.array_equal(kU, kL)
    minX, maxX, minY, maxY, iw, ih = [float(v) for v in f.bounds]
    assert (minX, maxX, minY, maxY) == (0.0, float(w), 0.0, float(h))
    start, items = ref_cpu.assign_grid(kU, minX, minY, iw, ih)
    rstart, ritems = f.grid()
    assert np.array_equal(start, rstart) and np.array_equal(items, ritems)
    rng = np.random.default_rng(seed)
    nonempty = 0
    for q in range(300):
        x, y = float(rng.uniform(-20, w + 20)), float(rng.uniform(-20, h + 20))
        r = float(rng.choice([3.0, 7.5, 15.0, 40.0, 100.0]))
        lo, hi = [(-1, -1), (0, 2), (2, -1), (-1, 3), (1, 1), (0, 7)][q % 6]
        a = f.features_in_area(x, y, r, lo, hi)
        c = ref_cpu.features_in_area(kU, start, items, minX, minY, iw, ih, x, y, r, lo, hi)
        assert np.array_equal(a, c), (x, y, r, lo, hi)
        nonempty += len(a) > 0
    assert nonempty > 100


def test_mono_frame_undistort_equals_reference(configs):
    """monocular constructor with the TUM1 distortion model: UndistortKeyPoints + ComputeImageBounds (Frame.cc:470-550)"""
    w, h, nf, sfac, nl, ini, mn = configs["tum"]
    img = synth.textured_frame(w, h, 9)
    cam, dist = (517.306408, 516.469215, 318.643040, 255.313989), (0.262383, -0.953104, -0.005358, 0.002628, 1.163314)
    f = R.RefFrame(img, None, nf, sfac, nl, ini, mn, cam=cam, dist=dist, bf=40.0)
    k, d = f.keys(0)
    kc, dc = ref_cpu.RefExtractor(nf, sfac, nl, ini, mn).extract(img)
    _same(k, d, kc, dc)
    kU, _ = f.keys(2)
    un = po.undistort_points(np.stack([k["x"], k["y"]], axis=1), cam, dist)
    assert np.array_equal(kU["x"], un[:, 0]) and np.array_equal(kU["y"], un[:, 1])
    for fld in ("size", "angle", "response", "octave"):
        assert np.array_equal(kU[fld], k[fld])
    corners = po.undistort_points(np.array([[0, 0], [w, 0], [0, h], [w, h]], np.float32), cam, dist)
    minX, maxX = min(corners[0, 0], corners[2, 0]), max(corners[1, 0], corners[3, 0])
    minY, maxY = min(corners[0, 1], corners[1, 1]), max(corners[2, 1], corners[3, 1])
    assert [float(v) for v in f.bounds[:4]] == [float(minX), float(maxX), float(minY), float(maxY)]
    iw, ih = np.float32(64) / (np.float32(maxX) - np.float32(minX)), np.float32(48) / (np.float32(maxY) - np.float32(minY))
    assert float(f.bounds[4]) == float(iw) and float(f.bounds[5]) == float(ih)
    start, items = ref_cpu.assign_grid(kU, float(minX), float(minY), float(iw), float(ih))
    rstart, ritems = f.grid()
    assert np.array_equal(start, rstart) and np.array_equal(items, ritems)


def test_distinctive_descriptors_equal_reference():
    """MapPoint::ComputeDistinctiveDescriptors over real MapPoint / KeyFrame objects (MapPoint.cc:266-340)"""
    import distinct_cases as dcases
    for name in dcases.CASES:
        desc, offs = dcases.case(name)
        out, has = R.distinctive_descriptors(desc, offs)
        idx, med = ref_cpu.distinctive_descriptors(desc, offs)
        sizes = np.diff(offs)
        assert np.array_equal(has, (sizes > 0).astype(np.int32)), name
        for g in range(len(sizes)):
            if sizes[g]:
                assert np.array_equal(out[g], desc[offs[g] + idx[g]]), (name, g)
            else:
                assert idx[g] == -1


def test_descriptor_distance_equals_reference():
    rng = np.random.default_rng(2)
    a = rng.integers(0, 256, (200, 32), dtype=np.uint8)
    b = rng.integers(0, 256, (200, 32), dtype=np.uint8)
    b[:20] = a[:20]
    b[20:40, :4] = a[20:40, :4]
    for i in range(200):
        assert R.descriptor_distance(a[i], b[i]) == ref_cpu.descriptor_distance(a[i], b[i]) == int(np.unpackbits(a[i] ^ b[i]).sum())
