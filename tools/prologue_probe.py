"""Measures the path's prologue on TUM-shaped colour frames (640x480x3, 1000 features): the batched
host call with the cvtColor fused into the level-0 staging against the same batch already gray, OpenCV's cvtColor on the host
cores, and Frame::UndistortKeyPoints (device call through host buffers vs cv2.undistortPoints).  Prints one JSON line."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "orb-slam2_b200"), os.path.join(ROOT, "oracle")):
    sys.path.insert(0, p)
import orbb200  # noqa: E402
import prologue_oracle as po  # noqa: E402
import synth  # noqa: E402


def main():
    nimg = int(os.environ.get("PRO_IMAGES", "128"))
    rng = np.random.default_rng(1)
    bases = [synth.textured_frame(640, 480, 3000 + i) for i in range(8)]
    col = np.zeros((nimg, 480, 640, 3), np.uint8)
    for i in range(nimg):
        b = bases[i % 8].astype(np.int32)
        for c in range(3):
            col[i, :, :, c] = np.clip(b + rng.integers(-20, 21, b.shape), 0, 255)
    gray = np.stack([po.cvt_gray(col[i], 1) for i in range(nimg)])
    ex = orbb200.ORBextractor(1000, 1.2, 8, 20, 7)

    def timed(fn, reps=5):
        fn()
        t0 = time.perf_counter()
        for _ in range(reps):
            r = fn()
        return (time.perf_counter() - t0) / reps * 1e3, r

    ms_col, (kc, dc, nc) = timed(lambda: ex.extract_color_batch(col, 1))
    ms_gray, (kg, dg, ng) = timed(lambda: ex.extract_batch(gray))
    same = bool(np.array_equal(nc, ng) and all(np.array_equal(kc[i, :nc[i]], kg[i, :ng[i]]) for i in range(nimg)))
    out = {"probe": "prologue: cvtColor fused into the extraction + UndistortKeyPoints", "images": nimg, "shape": "640x480x3",
           "extract_color_batch_ms": ms_col, "extract_color_batch_frames_per_s": nimg / (ms_col * 1e-3),
           "extract_batch_gray_ms": ms_gray, "extract_batch_gray_frames_per_s": nimg / (ms_gray * 1e-3),
           "h2d_bytes_color": int(col.nbytes), "h2d_bytes_gray": int(gray.nbytes), "parity_color_vs_gray": same}
    try:
        import cv2
        t0 = time.perf_counter()
        for i in range(nimg):
            cv2.cvtColor(col[i], cv2.COLOR_RGB2GRAY)
        out["cv2_cvtColor_ms_per_image_host"] = (time.perf_counter() - t0) / nimg * 1e3
        cam = np.array([517.306408, 516.469215, 318.643040, 255.313989], np.float32)
        dist = np.array([0.262383, -0.953104, -0.005358, 0.002628, 1.163314], np.float32)
        K = np.array([[cam[0], 0, cam[2]], [0, cam[1], cam[3]], [0, 0, 1]], np.float32)
        k0 = kc[0, :nc[0]].copy()
        pts = np.stack([k0["x"], k0["y"]], axis=1).reshape(-1, 1, 2)
        ms_u, un = timed(lambda: orbb200.undistort_keypoints(k0, cam, dist), 50)
        t0 = time.perf_counter()
        for _ in range(50):
            ref = cv2.undistortPoints(pts, K, dist, None, K)
        out["undistort_keypoints_ms"] = ms_u
        out["cv2_undistortPoints_ms"] = (time.perf_counter() - t0) / 50 * 1e3
        out["undistort_points"] = int(len(k0))
        out["parity_undistort_vs_cv2"] = bool(np.array_equal(un["x"].view(np.uint32), ref.reshape(-1, 2)[:, 0].view(np.uint32)) and
                                              np.array_equal(un["y"].view(np.uint32), ref.reshape(-1, 2)[:, 1].view(np.uint32)))
    except ImportError:
        pass
    print(json.dumps(out))


if __name__ == "__main__":
    main()
