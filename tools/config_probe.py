"""frames/s of the device-resident extraction for the parity-test configs of BASELINE.json (C1, C3, C4); C2 and C5 are bench.py's."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "orb-slam2_b200"))
import orbb200  # noqa: E402
import synth  # noqa: E402

dev = torch.device("cuda", 0)
for name, w, h, nf, nl, B in [("C1 TUM 640x480", 640, 480, 1000, 8, 320), ("C3 EuRoC 752x480", 752, 480, 1200, 8, 320), ("C4 4K 3840x2160", 3840, 2160, 8000, 12, 16)]:
    base = [synth.textured_frame(w, h, 500 + k) for k in range(4)]
    imgs = torch.from_numpy(np.stack([np.roll(base[b % 4], 7 * (b // 4), axis=0) for b in range(B)])).to(dev)
    ex = orbb200.ORBextractor(nf, 1.2, nl, 20, 7)
    cap = ex.cap
    kps = torch.empty((B, cap, 7), dtype=torch.float32, device=dev)
    desc = torch.empty((B, cap, 32), dtype=torch.uint8, device=dev)
    n = torch.zeros((B,), dtype=torch.int32, device=dev)
    stream = torch.cuda.ExternalStream(ex.stream(), device=dev)

    def step():
        ex.extract_batch_device(imgs.data_ptr(), B, w, h, w, w * h, kps.data_ptr(), desc.data_ptr(), cap, n.data_ptr())

    for _ in range(3):
        step()
    ex.sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(10):
        step()
    e1.record(stream)
    ex.sync()
    ms = e0.elapsed_time(e1) / 10
    print("%s: batch %d, %.3f ms per batch, %.0f frames/s, mean keypoints %.0f" % (name, B, ms, B / ms * 1e3, float(n.float().mean())))
    del ex
