"""e2e with T host threads, one extractor handle each, each taking 1/T of the step's stereo pairs (ctypes releases the GIL)."""
import ctypes as C
import os
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "orb-slam2_b200"))
import bench  # noqa: E402
import orbb200  # noqa: E402

P = 160
B = 2 * P
w, h = bench.KITTI["w"], bench.KITTI["h"]
imgs = torch.from_numpy(bench.make_pairs(P)).pin_memory()
L = orbb200.lib()
for T in (1, 2, 4):
    exs = [orbb200.ORBextractor(2000, 1.2, 8, 20, 7) for _ in range(T)]
    cap = exs[0].cap
    kps_h = torch.empty((B, cap, 7), dtype=torch.float32).pin_memory()
    desc_h = torch.empty((B, cap, 32), dtype=torch.uint8).pin_memory()
    n_h = torch.zeros((B,), dtype=torch.int32).pin_memory()
    u_h = torch.empty((P, cap), dtype=torch.float32).pin_memory()
    d_h = torch.empty((P, cap), dtype=torch.float32).pin_memory()
    nv_h = torch.zeros((P,), dtype=torch.int32).pin_memory()
    per = P // T

    def work(t):
        p0 = t * per
        np_ = per if t < T - 1 else P - p0
        ptrs = (C.c_void_p * (2 * np_))(*[imgs.data_ptr() + (2 * p0 + b) * w * h for b in range(2 * np_)])
        orbb200.check(L.orbb200_extract_stereo_batch(exs[t].h, ptrs, np_, w, h, w, C.c_void_p(kps_h.data_ptr() + 2 * p0 * cap * 28),
                                                     C.c_void_p(desc_h.data_ptr() + 2 * p0 * cap * 32), cap,
                                                     C.c_void_p(n_h.data_ptr() + 2 * p0 * 4), C.c_float(386.1448), C.c_float(0.5371657),
                                                     C.c_void_p(u_h.data_ptr() + p0 * cap * 4), C.c_void_p(d_h.data_ptr() + p0 * cap * 4),
                                                     C.c_void_p(nv_h.data_ptr() + p0 * 4)))

    def step():
        if T == 1:
            work(0)
            return
        th = [threading.Thread(target=work, args=(t,)) for t in range(T)]
        for x in th:
            x.start()
        for x in th:
            x.join()

    for _ in range(3):
        step()
    t0 = time.perf_counter()
    for _ in range(8):
        step()
    ms = (time.perf_counter() - t0) / 8 * 1e3
    print("threads %d: %.2f ms per 320 images (%.0f frames/s), keypoints %d matches %d" % (T, ms, B / ms * 1e3, int(n_h.sum()), int(nv_h.sum())))
