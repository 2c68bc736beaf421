"""Generates tests/golden/prologue.npz from the OpenCV on this image (cv2): gray conversions of small seeded colour images
and undistorted points for the TUM1 / TUM2 camera models (Examples/Monocular/TUM1.yaml, TUM2.yaml) plus a strongly
distorted synthetic model that drives some points through the icdist < 0 branch."""
import os

import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CAMS = {
    "tum1": ((517.306408, 516.469215, 318.643040, 255.313989), (0.262383, -0.953104, -0.005358, 0.002628, 1.163314)),
    "tum2": ((520.908620, 521.007327, 325.141442, 249.701764), (0.231222, -0.784899, -0.003257, -0.000105, 0.917205)),
    "wild": ((300.0, 310.0, 320.0, 240.0), (-0.9, 0.35, 0.01, -0.02, -0.6)),
    "rectified": ((718.856, 718.856, 607.1928, 185.2157), (0.0, 0.0, 0.0, 0.0, 0.0)),
}


def main():
    rng = np.random.default_rng(99)
    out = {}
    rgb = rng.integers(0, 256, (61, 97, 3), dtype=np.uint8)
    rgba = rng.integers(0, 256, (33, 50, 4), dtype=np.uint8)
    out["rgb"], out["rgba"] = rgb, rgba
    out["rgb_as_rgb"] = cv2.cvtColor(rgb, cv2.COLOR_RGB2GRAY)
    out["rgb_as_bgr"] = cv2.cvtColor(rgb, cv2.COLOR_BGR2GRAY)
    out["rgba_as_rgba"] = cv2.cvtColor(rgba, cv2.COLOR_RGBA2GRAY)
    out["rgba_as_bgra"] = cv2.cvtColor(rgba, cv2.COLOR_BGRA2GRAY)
    pts = np.stack([rng.uniform(-20, 660, 4000), rng.uniform(-20, 500, 4000)], axis=1).astype(np.float32)
    pts[:4] = [[0, 0], [640, 0], [0, 480], [640, 480]]  # Frame::ComputeImageBounds corners
    out["pts"] = pts
    for name, (cam, dist) in CAMS.items():
        K = np.array([[cam[0], 0, cam[2]], [0, cam[1], cam[3]], [0, 0, 1]], np.float32)
        D = np.array(dist, np.float32)
        out["cam_" + name] = np.array(cam, np.float32)
        out["dist_" + name] = D
        if D[0] == 0:
            out["und_" + name] = pts.copy()  # Frame.cc:473-477
        else:
            out["und_" + name] = cv2.undistortPoints(pts.reshape(-1, 1, 2), K, D, None, K).reshape(-1, 2)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "prologue.npz"), **out)
    print("cv2", cv2.__version__, {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
