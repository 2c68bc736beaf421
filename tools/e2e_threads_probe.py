#!/usr/bin/env python3
"""orbb200_extract_stereo_batch (host buffers) for one step of 160 KITTI pairs: one handle / one host thread against T handles on T host
threads, each taking a contiguous share of the step (synchronous calls; the GPU overlaps one handle's tail with the others' copies)."""
import ctypes as C, os, sys, threading, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "orb-slam2_b200")]
import numpy as np, torch, bench, orbb200
P = 160; B = 2 * P
K = bench.KITTI; w, h = K["w"], K["h"]
imgs = torch.from_numpy(bench.make_pairs(P)).pin_memory()
L = orbb200.lib()
def make(T):
    hs = []
    for t in range(T):
        ex = orbb200.ORBextractor(2000, 1.2, 8, 20, 7)
        cap = ex.cap
        p0, p1 = t * P // T, (t + 1) * P // T
        nb = 2 * (p1 - p0)
        out = dict(ex=ex, cap=cap, p0=p0, np=p1 - p0,
                   kps=torch.empty((nb, cap, 7), dtype=torch.float32).pin_memory(), desc=torch.empty((nb, cap, 32), dtype=torch.uint8).pin_memory(),
                   n=torch.zeros((nb,), dtype=torch.int32).pin_memory(), uR=torch.empty((p1 - p0, cap), dtype=torch.float32).pin_memory(),
                   dep=torch.empty((p1 - p0, cap), dtype=torch.float32).pin_memory(), nv=torch.zeros((p1 - p0,), dtype=torch.int32).pin_memory())
        out["ptrs"] = (C.c_void_p * nb)(*[imgs.data_ptr() + (2 * p0 + b) * w * h for b in range(nb)])
        hs.append(out)
    return hs
def call(o):
    orbb200.check(L.orbb200_extract_stereo_batch(o["ex"].h, o["ptrs"], o["np"], w, h, w, o["kps"].data_ptr(), o["desc"].data_ptr(), o["cap"], o["n"].data_ptr(),
                                                 K["bf"], K["bf"] / 718.856, o["uR"].data_ptr(), o["dep"].data_ptr(), o["nv"].data_ptr()))
def step(hs):
    if len(hs) == 1: return call(hs[0])
    th = [threading.Thread(target=call, args=(o,)) for o in hs]
    for t in th: t.start()
    for t in th: t.join()
for T in (1, 2, 3, 4, 1):
    hs = make(T)
    for _ in range(3): step(hs)
    t0 = time.perf_counter()
    for _ in range(10): step(hs)
    ms = (time.perf_counter() - t0) / 10 * 1e3
    print("threads/handles %d: %.2f ms per step (%.0f frames/s), keypoints %d" % (T, ms, B / ms * 1e3, sum(int(o["n"].sum()) for o in hs)), flush=True)
    del hs
