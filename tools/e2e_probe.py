"""Times orbb200_extract_batch (host buffers) for several pipeline settings; prints ms per 320 KITTI images.
Run on the GPU box:  python tools/e2e_probe.py"""
import ctypes as C
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "orb-slam2_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402
import orbb200  # noqa: E402

P = int(os.environ.get("PAIRS", "160"))
B = 2 * P
w, h = bench.KITTI["w"], bench.KITTI["h"]
imgs = torch.from_numpy(bench.make_pairs(P)).pin_memory()
ex = orbb200.ORBextractor(2000, 1.2, 8, 20, 7)
cap = ex.cap
kps_h = torch.empty((B, cap, 7), dtype=torch.float32).pin_memory()
desc_h = torch.empty((B, cap, 32), dtype=torch.uint8).pin_memory()
n_h = torch.zeros((B,), dtype=torch.int32).pin_memory()
ptrs = (C.c_void_p * B)(*[imgs.data_ptr() + b * w * h for b in range(B)])
L = orbb200.lib()


def run(reps=6):
    for _ in range(2):
        orbb200.check(L.orbb200_extract_batch(ex.h, ptrs, B, w, h, w, kps_h.data_ptr(), desc_h.data_ptr(), cap, n_h.data_ptr()))
    t0 = time.perf_counter()
    for _ in range(reps):
        orbb200.check(L.orbb200_extract_batch(ex.h, ptrs, B, w, h, w, kps_h.data_ptr(), desc_h.data_ptr(), cap, n_h.data_ptr()))
    return (time.perf_counter() - t0) / reps * 1e3


# raw copies for scale
d = torch.empty_like(imgs, device="cuda")
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(5):
    d.copy_(imgs, non_blocking=True)
torch.cuda.synchronize()
print("H2D %d MB: %.2f ms" % (imgs.numel() >> 20, (time.perf_counter() - t0) / 5 * 1e3))
dk = torch.empty((B, cap, 60), dtype=torch.uint8, device="cuda")
hk = torch.empty((B, cap, 60), dtype=torch.uint8).pin_memory()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(5):
    hk.copy_(dk, non_blocking=True)
torch.cuda.synchronize()
print("D2H %d MB: %.2f ms" % (dk.numel() >> 20, (time.perf_counter() - t0) / 5 * 1e3))
ref = None
for chunk, ns in [(16, 8), (8, 8), (12, 8), (20, 8), (24, 8), (32, 8), (16, 4), (16, 6), (32, 4), (16, 8)]:
    ex.set_pipeline(chunk, ns)
    ms = run()
    cs = int(n_h.numpy().astype(np.int64).sum()) * 1000003 + int(desc_h.numpy()[:, :100].astype(np.int64).sum())
    if ref is None:
        ref = cs
    print("chunk %3d streams %d: %.2f ms  (%.0f frames/s)%s" % (chunk, ns, ms, B / ms * 1e3, "" if cs == ref else "  CHECKSUM DIFFERS"))
