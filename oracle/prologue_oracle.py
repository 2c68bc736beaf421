"""TEST INFRASTRUCTURE ONLY (oracle): numpy restatement of the two OpenCV calls that precede the hot path.

* cvt_gray            -- cv::cvtColor(im, gray, CV_RGB2GRAY / CV_BGR2GRAY / CV_RGBA2GRAY / CV_BGRA2GRAY) as called by
                         Tracking::GrabImage* (reference src/Tracking.cc:177-202, 218-231, 255-266), 8-bit path:
                         (R*9798 + G*19235 + B*3735 + (1 << 14)) >> 15.
* undistort_points    -- cv::undistortPoints(mat, mat, mK, mDistCoef, cv::Mat(), mK) as called by Frame::UndistortKeyPoints /
                         Frame::ComputeImageBounds (reference src/Frame.cc:470-515, 517-550): double arithmetic, 5 iterations.

The arithmetic lives in OpenCV (un-vendored, version unpinned by the reference's CMakeLists.txt:32-38); it is pinned to the
OpenCV on this image (cv2 4.13) by tests/test_oracle_pin.py and by the committed golden vectors tests/golden/prologue.npz
(generator tools/make_prologue_golden.py).  Only tests/ may import this module."""
import numpy as np


def cvt_gray(img, r_first):
    img = np.asarray(img)
    c0 = img[..., 0].astype(np.int64)
    c1 = img[..., 1].astype(np.int64)
    c2 = img[..., 2].astype(np.int64)
    r, b = (c0, c2) if r_first else (c2, c0)
    return ((r * 9798 + c1 * 19235 + b * 3735 + (1 << 14)) >> 15).astype(np.uint8)


def undistort_points(pts, cam, dist):
    """pts [N, 2] float32; cam = (fx, fy, cx, cy) float32; dist = (k1, k2, p1, p2, k3) float32 -> [N, 2] float32"""
    pts = np.asarray(pts, np.float32).reshape(-1, 2)
    fx, fy, cx, cy = [np.float64(np.float32(v)) for v in cam]
    k = np.zeros(12, np.float64)
    k[:5] = np.asarray(dist, np.float32).astype(np.float64)
    if np.float32(dist[0]) == 0.0:  # Frame.cc:473-477
        return pts.copy()
    ifx, ify = 1.0 / fx, 1.0 / fy
    u = pts[:, 0].astype(np.float64)
    v = pts[:, 1].astype(np.float64)
    x = (u - cx) * ifx
    y = (v - cy) * ify
    x0, y0 = x.copy(), y.copy()
    live = np.ones(len(x), bool)  # points whose iteration has not hit icdist < 0
    for _ in range(5):
        r2 = x * x + y * y
        icdist = (1 + ((k[7] * r2 + k[6]) * r2 + k[5]) * r2) / (1 + ((k[4] * r2 + k[1]) * r2 + k[0]) * r2)
        neg = live & (icdist < 0)
        dx = 2 * k[2] * x * y + k[3] * (r2 + 2 * x * x) + k[8] * r2 + k[9] * r2 * r2
        dy = k[2] * (r2 + 2 * y * y) + 2 * k[3] * x * y + k[10] * r2 + k[11] * r2 * r2
        nx = (x0 - dx) * icdist
        ny = (y0 - dy) * icdist
        upd = live & ~neg
        x = np.where(upd, nx, np.where(neg, x0, x))
        y = np.where(upd, ny, np.where(neg, y0, y))
        live = upd
    xx = fx * x + 0.0 * y + cx
    yy = 0.0 * x + fy * y + cy
    ww = 1.0 / (0.0 * x + 0.0 * y + 1.0)
    return np.stack([(xx * ww).astype(np.float32), (yy * ww).astype(np.float32)], axis=1)
