#!/usr/bin/env python3
"""streamed steps (orbb200_extract_stereo_batch_begin / _wait, two in flight) of 160 KITTI pairs for several pipeline settings"""
import ctypes as C, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "orb-slam2_b200")]
import numpy as np, torch, bench, orbb200
P = 160; B = 2 * P
K = bench.KITTI; w, h = K["w"], K["h"]
imgs = torch.from_numpy(bench.make_pairs(P)).pin_memory()
L = orbb200.lib()
ex = orbb200.ORBextractor(2000, 1.2, 8, 20, 7)
cap = ex.cap
def bufs():
    return (torch.empty((B, cap, 7), dtype=torch.float32).pin_memory(), torch.empty((B, cap, 32), dtype=torch.uint8).pin_memory(), torch.zeros((B,), dtype=torch.int32).pin_memory(),
            torch.empty((P, cap), dtype=torch.float32).pin_memory(), torch.empty((P, cap), dtype=torch.float32).pin_memory(), torch.zeros((P,), dtype=torch.int32).pin_memory())
outs = [bufs(), bufs()]
ptrs = (C.c_void_p * B)(*[imgs.data_ptr() + b * w * h for b in range(B)])
def begin(k):
    o = outs[k & 1]; t = C.c_int(-1)
    orbb200.check(L.orbb200_extract_stereo_batch_begin(ex.h, ptrs, P, w, h, w, o[0].data_ptr(), o[1].data_ptr(), cap, o[2].data_ptr(), K["bf"], K["bf"] / 718.856,
                                                       o[3].data_ptr(), o[4].data_ptr(), o[5].data_ptr(), C.byref(t)))
    return t.value
def run(steps=12):
    prev = -1
    for k in range(3):
        cur = begin(k); orbb200.check(L.orbb200_extract_batch_wait(ex.h, prev)); prev = cur
    orbb200.check(L.orbb200_extract_batch_wait(ex.h, prev)); prev = -1
    t0 = time.perf_counter()
    for k in range(steps):
        cur = begin(k); orbb200.check(L.orbb200_extract_batch_wait(ex.h, prev)); prev = cur
    orbb200.check(L.orbb200_extract_batch_wait(ex.h, prev))
    return (time.perf_counter() - t0) / steps * 1e3
for chunk, ns in [(16, 8), (32, 8), (40, 8), (64, 8), (80, 8), (160, 8), (32, 4), (64, 4), (16, 8)]:
    ex.set_pipeline(chunk, ns)
    ms = run()
    print("chunk %3d streams %d: %.2f ms per step (%.0f frames/s) keypoints %d" % (chunk, ns, ms, B / ms * 1e3, int(outs[0][2].sum())), flush=True)
