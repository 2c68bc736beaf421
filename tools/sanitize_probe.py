"""Small workload over every kernel family (extraction on four shapes, colour input, stereo, brute force, distinctive
descriptors, undistortion): the input of a compute-sanitizer run where the pool allows one
(compute-sanitizer --tool memcheck python tools/sanitize_probe.py -- closed on this round's pool), and a quick
all-kernels smoke otherwise."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "orb-slam2_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import orbb200  # noqa: E402
import synth  # noqa: E402
import distinct_cases as dc  # noqa: E402

for (w, h, nf, nl) in ((640, 480, 1000, 8), (1241, 376, 2000, 8), (97, 131, 200, 4), (333, 240, 500, 8)):
    ex = orbb200.ORBextractor(nf, 1.2, nl, 20, 7)
    img = synth.textured_frame(w, h, 7 + w)
    k, d = ex(img)
    imgs = np.stack([img, img[::-1].copy(), img[:, ::-1].copy()])
    kb, db, nb = ex.extract_batch(imgs)
    col = np.stack([img, img, img], axis=2)
    kc, dcol = ex.extract_color(col, 1)
    print("extract %dx%d: %d keypoints, batch %s, colour %d" % (w, h, len(k), nb.tolist(), len(kc)))
L, R = synth.stereo_pair(1241, 376, 5)
ex = orbb200.ORBextractor(2000, 1.2, 8, 20, 7)
out = ex.extract_stereo_batch(np.stack([L, R]), 386.1448, 0.5371657)
print("stereo matches", int(out[5][0]))
m = orbb200.ORBmatcher()
q, db, _ = synth.descriptor_db(300, 20000, seed=3)
t = m.bruteforce_top2(q, db)
for name in ("small", "ragged_with_empty", "large"):
    desc, offs = dc.case(name)
    gi, gm = m.distinctive_descriptors(desc, offs)
kp = np.zeros(100, orbb200.KP_DTYPE)
kp["x"] = np.linspace(0, 640, 100)
kp["y"] = np.linspace(0, 480, 100)
u = orbb200.undistort_keypoints(kp, [517.3, 516.5, 318.6, 255.3], [0.2624, -0.9531, -0.0054, 0.0026, 1.1633])
g = orbb200.cvt_gray(np.random.default_rng(0).integers(0, 256, (37, 641, 4), dtype=np.uint8), 0)
print("probe done", int(t["best_idx"].sum()), int(gi.sum()), float(u["x"].sum()), int(g.sum()))
