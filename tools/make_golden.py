#!/usr/bin/env python3
"""Generate tests/golden/extract_*.npz from the REFERENCE's own ORBextractor (oracle/_ref/liborbslam_ref.so: the unmodified
/root/reference/src/ORBextractor.cc compiled by oracle/Makefile; node-pointer ties resolved by creation order, see
oracle/orbslam_ref.cpp).  Run in the dev container.  The fixtures are small (keypoints + descriptors of one frame)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "orb-slam2_b200"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import orbslam_ref  # noqa: E402
import synth  # noqa: E402
from conftest import CONFIGS  # noqa: E402

SEEDS = {"tum": 1000, "kitti": 1001, "euroc": 1002}
for name, (w, h, nf, sf, nl, ini, mn) in CONFIGS.items():
    img = synth.textured_frame(w, h, SEEDS[name])
    kp, d = orbslam_ref.Extractor(nf, sf, nl, ini, mn).extract(img)
    out = os.path.join(ROOT, "tests", "golden", "extract_%s.npz" % name)
    np.savez_compressed(out, seed=SEEDS[name], img_sum=int(img.astype(np.uint64).sum()), kps=kp, desc=d, source="reference:src/ORBextractor.cc")
    print(name, len(kp), os.path.getsize(out))
