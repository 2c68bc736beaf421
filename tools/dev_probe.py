"""Times orbb200_extract_batch_device (+ stereo) on 320 resident KITTI images for several chunk / stream settings."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "orb-slam2_b200"))
import bench  # noqa: E402
import orbb200  # noqa: E402

P = int(os.environ.get("PAIRS", "160"))
B = 2 * P
K = bench.KITTI
w, h = K["w"], K["h"]
dev = torch.device("cuda", 0)
imgs = torch.from_numpy(bench.make_pairs(P)).to(dev)
ex = orbb200.ORBextractor(2000, 1.2, 8, 20, 7)
cap = ex.cap
kps_d = torch.empty((B, cap, 7), dtype=torch.float32, device=dev)
desc_d = torch.empty((B, cap, 32), dtype=torch.uint8, device=dev)
n_d = torch.zeros((B,), dtype=torch.int32, device=dev)
uR_d = torch.empty((P, cap), dtype=torch.float32, device=dev)
dep_d = torch.empty((P, cap), dtype=torch.float32, device=dev)
nv_d = torch.zeros((P,), dtype=torch.int32, device=dev)
stream = torch.cuda.ExternalStream(ex.stream(), device=dev)


def step():
    ex.extract_batch_device(imgs.data_ptr(), B, w, h, w, w * h, kps_d.data_ptr(), desc_d.data_ptr(), cap, n_d.data_ptr())
    ex.stereo_match_batch_device(P, K["bf"], K["bf"] / K["fx"], uR_d.data_ptr(), dep_d.data_ptr(), nv_d.data_ptr())


def run(reps=10):
    for _ in range(3):
        step()
    ex.sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(reps):
        step()
    e1.record(stream)
    ex.sync()
    return e0.elapsed_time(e1) / reps


ref = None
for chunk, ns in [(320, 1), (160, 2), (80, 2), (80, 4), (40, 4), (40, 3), (40, 2), (32, 4), (20, 4), (16, 4), (64, 4), (54, 3)]:
    ex.set_pipeline(32, ns, chunk)
    ms = run()
    cs = (int(n_d.sum().item()), int(desc_d[:, :64].to(torch.int64).sum().item()), int(nv_d.sum().item()))
    if ref is None:
        ref = cs
    print("chunk %3d streams %d: %.3f ms  (%.0f frames/s)%s" % (chunk, ns, ms, B / ms * 1e3, "" if cs == ref else "  CHECKSUM DIFFERS"))
