"""GPU parity tests (through the C-ABI) of the path's prologue: colour -> gray fused into the level-0
staging of the extraction (Tracking::GrabImage*, reference Tracking.cc:177-202) and Frame::UndistortKeyPoints (Frame.cc:470-515),
against the oracle and the golden vectors generated from cv2.  Bit-exact."""
import os

import numpy as np
import pytest

import orbb200
import prologue_oracle as po
import ref_cpu
import synth

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = np.load(os.path.join(ROOT, "tests", "golden", "prologue.npz"))


def test_cvt_gray_equals_golden():
    assert np.array_equal(orbb200.cvt_gray(GOLD["rgb"], 1), GOLD["rgb_as_rgb"])
    assert np.array_equal(orbb200.cvt_gray(GOLD["rgb"], 0), GOLD["rgb_as_bgr"])
    assert np.array_equal(orbb200.cvt_gray(GOLD["rgba"], 1), GOLD["rgba_as_rgba"])
    assert np.array_equal(orbb200.cvt_gray(GOLD["rgba"], 0), GOLD["rgba_as_bgra"])


@pytest.mark.parametrize("w,h,ch", [(640, 480, 3), (641, 37, 3), (1241, 376, 4), (17, 9, 3), (16, 1, 4), (333, 240, 3)])
def test_cvt_gray_equals_oracle_all_alignments(w, h, ch):
    rng = np.random.default_rng(w * 7 + h + ch)
    img = rng.integers(0, 256, (h, w, ch), dtype=np.uint8)
    for r_first in (0, 1):
        assert np.array_equal(orbb200.cvt_gray(img, r_first), po.cvt_gray(img, r_first))
    # every byte value of every channel
    ramp = np.zeros((3, 256, ch), np.uint8)
    for c in range(3):
        ramp[c, :, c] = np.arange(256)
    assert np.array_equal(orbb200.cvt_gray(ramp, 1), po.cvt_gray(ramp, 1))


@pytest.mark.parametrize("ch,r_first", [(3, 1), (3, 0), (4, 1)])
def test_extract_color_equals_gray_extraction(ch, r_first):
    """the fused prologue: extraction of a colour frame == extraction of its oracle gray conversion == the CPU oracle"""
    base = synth.textured_frame(640, 480, 1234)
    rng = np.random.default_rng(5)
    img = np.zeros((480, 640, ch), np.uint8)
    for c in range(ch):
        img[..., c] = np.clip(base.astype(np.int32) + rng.integers(-30, 31, base.shape), 0, 255)
    gray = po.cvt_gray(img, r_first)
    ex = orbb200.ORBextractor(1000, 1.2, 8, 20, 7)
    kc, dc = ex.extract_color(img, r_first)
    assert np.array_equal(ex.pyramid_level(0), gray)  # mImGray = level 0 of the pyramid
    kg, dg = ex(gray)
    assert len(kc) == len(kg) and np.array_equal(kc, kg) and np.array_equal(dc, dg)
    kr, dr = ref_cpu.RefExtractor(1000, 1.2, 8, 20, 7).extract(gray)
    for f in ("x", "y", "size", "response", "octave"):
        assert np.array_equal(kc[f], kr[f]), f
    # a batch through the chunked pipeline gives the same
    kb, db, nb = ex.extract_color_batch(np.stack([img, img[::-1].copy(), img]), r_first)
    assert nb[0] == len(kc) and np.array_equal(kb[0, :nb[0]], kc) and np.array_equal(kb[2, :nb[2]], kc)
    assert np.array_equal(db[0, :nb[0]], dc)


@pytest.mark.parametrize("name", ["tum1", "tum2", "wild", "rectified"])
def test_undistort_keypoints_equals_golden(name):
    pts = GOLD["pts"]
    kps = np.zeros(len(pts), orbb200.KP_DTYPE)
    kps["x"], kps["y"] = pts[:, 0], pts[:, 1]
    kps["size"], kps["angle"], kps["response"] = 31.0, 45.5, 77.0
    kps["octave"] = np.arange(len(pts)) % 8
    kps["class_id"] = -1
    out = orbb200.undistort_keypoints(kps, GOLD["cam_" + name], GOLD["dist_" + name])
    ref = GOLD["und_" + name]
    assert np.array_equal(out["x"].view(np.uint32), ref[:, 0].view(np.uint32))
    assert np.array_equal(out["y"].view(np.uint32), ref[:, 1].view(np.uint32))
    for f in ("size", "angle", "response", "octave", "class_id"):
        assert np.array_equal(out[f], kps[f])
    o2 = po.undistort_points(pts, GOLD["cam_" + name], GOLD["dist_" + name])
    assert np.array_equal(out["x"].view(np.uint32), o2[:, 0].view(np.uint32))


def test_undistort_empty_and_single():
    cam, dist = GOLD["cam_tum1"], GOLD["dist_tum1"]
    assert len(orbb200.undistort_keypoints(np.zeros(0, orbb200.KP_DTYPE), cam, dist)) == 0
    kp = np.zeros(1, orbb200.KP_DTYPE)
    kp["x"], kp["y"] = 100.25, 200.5
    o = orbb200.undistort_keypoints(kp, cam, dist)
    r = po.undistort_points(np.array([[100.25, 200.5]], np.float32), cam, dist)
    assert o["x"][0] == r[0, 0] and o["y"][0] == r[0, 1]
