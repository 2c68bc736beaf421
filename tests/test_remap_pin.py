"""CPU: the EuRoC prologue's arithmetic -- cv::remap INTER_LINEAR (CV_32FC1 maps) and cv::initUndistortRectifyMap
(Examples/Stereo/stereo_euroc.cc:96-97, 131-132) -- restated in oracle/orb_ref.c and, for the map set-up, in the product's host-only
orbb200_init_undistort_rectify_map, pinned to the OpenCV on the image (cv2)."""
import numpy as np
import pytest

import orbb200
import ref_cpu

cv2 = pytest.importorskip("cv2")

# Examples/Stereo/EuRoC.yaml (LEFT.* / RIGHT.*)
EUROC = {
    "left": dict(K=[[458.654, 0, 367.215], [0, 457.296, 248.375], [0, 0, 1]], D=[-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05, 0.0],
                 R=[[0.999966347530033, -0.001422739138722922, 0.008079580483432283], [0.001365741834644127, 0.9999741760894847, 0.007055629199258132],
                    [-0.008089410156878961, -0.007044357138835809, 0.9999424675829176]],
                 P=[[435.2046959714599, 0, 367.4517211914062, 0], [0, 435.2046959714599, 252.2008514404297, 0], [0, 0, 1, 0]]),
    "right": dict(K=[[457.587, 0, 379.999], [0, 456.134, 255.238], [0, 0, 1]], D=[-0.28368365, 0.07451284, -0.00010473, -3.555907e-05, 0.0],
                  R=[[0.9999633526194376, -0.003625811871560086, 0.007755443660172947], [0.003680398547259526, 0.9999684752771629, -0.007035845251224894],
                     [-0.007729688520722713, 0.007064130529506649, 0.999945173484644]],
                  P=[[435.2046959714599, 0, 367.4517211914062, -47.90639384423901], [0, 435.2046959714599, 252.2008514404297, 0], [0, 0, 1, 0]]),
}
W, H = 752, 480


def test_remap_oracle_equals_cv2_on_random_maps():
    rng = np.random.default_rng(2)
    for t in range(16):
        sh, sw = int(rng.integers(8, 260)), int(rng.integers(8, 380))
        dh, dw = int(rng.integers(8, 260)), int(rng.integers(8, 380))
        img = rng.integers(0, 256, (sh, sw), dtype=np.uint8)
        spread = float(rng.choice([1.0, 30.0, 3000.0]))
        mx = (rng.uniform(-0.2, 1.2, (dh, dw)) * sw + rng.normal(0, spread, (dh, dw))).astype(np.float32)
        my = (rng.uniform(-0.2, 1.2, (dh, dw)) * sh + rng.normal(0, spread, (dh, dw))).astype(np.float32)
        if t % 4 == 0:   # whole pixels
            mx, my = np.rint(mx), np.rint(my)
        if t % 4 == 1:   # ties of the 1/32 rounding
            mx = (np.rint(mx * 32) / 32 + 1.0 / 64).astype(np.float32)
        assert np.array_equal(ref_cpu.remap_linear(img, mx, my), cv2.remap(img, mx, my, cv2.INTER_LINEAR)), t


@pytest.mark.parametrize("side", ["left", "right"])
def test_rectify_maps_equal_cv2(side):
    c = EUROC[side]
    K, D, R, P = (np.array(c[k], np.float64) for k in "KDRP")
    m1, m2 = cv2.initUndistortRectifyMap(K, D, R, P[:3, :3], (W, H), cv2.CV_32F)
    for name, (a1, a2) in (("oracle", ref_cpu.init_undistort_rectify_map(K, D, R, P, W, H)),
                           ("product (host-only entry point)", orbb200.init_undistort_rectify_map(K, D, R, P, W, H))):
        for a, b in ((a1, m1), (a2, m2)):
            ulp = np.abs(a.view(np.int32).astype(np.int64) - b.view(np.int32).astype(np.int64))
            # OpenCV vectorises the row loop: allow the last bit of the double to fall the other way on a handful of entries
            assert ulp.max() <= 1 and (ulp != 0).mean() < 1e-4, (name, int(ulp.max()), float((ulp != 0).mean()))
    img = np.random.default_rng(5).integers(0, 256, (H, W), dtype=np.uint8)
    a1, a2 = ref_cpu.init_undistort_rectify_map(K, D, R, P, W, H)
    got, want = ref_cpu.remap_linear(img, a1, a2), cv2.remap(img, m1, m2, cv2.INTER_LINEAR)
    assert (got != want).mean() < 1e-4
