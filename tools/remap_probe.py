#!/usr/bin/env python3
"""EuRoC rectification (cv::remap INTER_LINEAR, 752x480) on the device: ms per batch of 256 stereo-left images resident in HBM, GB/s
against the algorithmic bytes (1 B read + 1 B written + 8 B of map record per pixel), cv2.remap on one host core beside it."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "orb-slam2_b200"), os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")]
import numpy as np, torch, orbb200, synth
from test_remap_pin import EUROC, W, H
c = EUROC["left"]
K, D, R, P = (np.array(c[k], np.float64) for k in "KDRP")
mx, my = orbb200.init_undistort_rectify_map(K, D, R, P, W, H)
rm = orbb200.Remap(mx, my)
B = 256
dev = torch.device("cuda", 0)
src = torch.randint(0, 256, (B, H, W), dtype=torch.uint8, device=dev)
dst = torch.empty_like(src)
st = torch.cuda.Stream(device=dev)
L = orbb200.lib()
def step(): orbb200.check(L.orbb200_remap_batch_device(rm.h, src.data_ptr(), B, W, H, W, W * H, dst.data_ptr(), W, W * H, st.cuda_stream))
for _ in range(3): step()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
st.synchronize()
with torch.cuda.stream(st):
    e0.record()
    for _ in range(20): step()
    e1.record()
st.synchronize()
ms = e0.elapsed_time(e1) / 20
res = {"images": B, "shape": [W, H], "ms_per_batch": ms, "frames_per_s": B / ms * 1e3, "algorithmic_gbs": B * W * H * 10 / ms / 1e6}
try:
    import cv2
    cv2.setNumThreads(1)
    img = src[0].cpu().numpy()
    t0 = time.perf_counter()
    for _ in range(50): cv2.remap(img, mx, my, cv2.INTER_LINEAR)
    res["cv2_remap_one_thread_ms_per_image"] = (time.perf_counter() - t0) / 50 * 1e3
    res["equal_to_cv2"] = bool(np.array_equal(dst[0].cpu().numpy(), cv2.remap(img, mx, my, cv2.INTER_LINEAR)))
except ImportError:
    pass
print(json.dumps(res))
