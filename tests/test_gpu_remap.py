"""GPU: orbb200_remap_* (cv::remap INTER_LINEAR of the EuRoC example, Examples/Stereo/stereo_euroc.cc:131-132) through the C-ABI
against the oracle and cv2 -- bit-exact (integer arithmetic); host API, device API, batches, maps that leave the image; and the
EuRoC loader + StereoRectifier facade on a synthetic sequence."""
import os
import subprocess

import numpy as np
import pytest

import orbb200
import ref_cpu
import synth
from test_remap_pin import EUROC, W, H

pytestmark = pytest.mark.gpu
cv2 = pytest.importorskip("cv2")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_remap_equals_oracle_and_cv2_random_maps():
    rng = np.random.default_rng(12)
    for t in range(8):
        sh, sw = int(rng.integers(8, 300)), int(rng.integers(8, 500))
        dh, dw = int(rng.integers(8, 300)), int(rng.integers(5, 500))
        imgs = rng.integers(0, 256, (3, sh, sw), dtype=np.uint8)
        spread = float(rng.choice([1.0, 40.0, 1e5]))
        mx = (rng.uniform(-0.2, 1.2, (dh, dw)) * sw + rng.normal(0, spread, (dh, dw))).astype(np.float32)
        my = (rng.uniform(-0.2, 1.2, (dh, dw)) * sh + rng.normal(0, spread, (dh, dw))).astype(np.float32)
        if t % 3 == 0:
            mx = (np.rint(mx * 32) / 32 + 1.0 / 64).astype(np.float32)
        got = orbb200.Remap(mx, my)(imgs)
        for b in range(3):
            assert np.array_equal(got[b], ref_cpu.remap_linear(imgs[b], mx, my)), (t, b)
            assert np.array_equal(got[b], cv2.remap(imgs[b], mx, my, cv2.INTER_LINEAR)), (t, b)


def test_euroc_rectification_host_and_device_api():
    import torch
    c = EUROC["left"]
    K, D, R, P = (np.array(c[k], np.float64) for k in "KDRP")
    mx, my = orbb200.init_undistort_rectify_map(K, D, R, P, W, H)
    imgs = np.stack([synth.textured_frame(W, H, 70 + i) for i in range(5)])
    rm = orbb200.Remap(mx, my)
    got = rm(imgs)
    m1, m2 = cv2.initUndistortRectifyMap(K, D, R, P[:3, :3], (W, H), cv2.CV_32F)
    for b in range(5):
        assert np.array_equal(got[b], ref_cpu.remap_linear(imgs[b], mx, my))
        assert (got[b] != cv2.remap(imgs[b], m1, m2, cv2.INTER_LINEAR)).mean() < 1e-4  # the maps may differ in one float ulp on a few entries
    # device API: strided source / destination blocks, asynchronous on a caller stream, chained in front of the extraction
    dev = torch.device("cuda", 0)
    src = torch.zeros((5, H + 3, W + 16), dtype=torch.uint8, device=dev)
    src[:, :H, :W] = torch.from_numpy(imgs).to(dev)
    dst = torch.full((5, H, W + 8), 7, dtype=torch.uint8, device=dev)
    st = torch.cuda.Stream(device=dev)
    orbb200.check(orbb200.lib().orbb200_remap_batch_device(rm.h, src.data_ptr(), 5, W, H, W + 16, (H + 3) * (W + 16), dst.data_ptr(), W + 8, H * (W + 8),
                                                          st.cuda_stream))
    st.synchronize()
    d = dst.cpu().numpy()
    assert np.array_equal(d[:, :, :W], got) and (d[:, :, W:] == 7).all()
    ex = orbb200.ORBextractor(1200, 1.2, 8, 20, 7)
    k_dev, d_dev = ex(np.ascontiguousarray(d[0, :, :W]))
    k_ref, d_ref = ref_cpu.RefExtractor(1200, 1.2, 8, 20, 7).extract(ref_cpu.remap_linear(imgs[0], mx, my))
    assert np.array_equal(k_dev["x"], k_ref["x"]) and np.array_equal(k_dev["y"], k_ref["y"]) and len(k_dev) > 600


def test_euroc_loader_and_rectifier_facade(tmp_path):
    exe = os.path.join(ROOT, "orb-slam2_b200", "loader_test.bin")
    assert os.path.exists(exe), "loader_test.bin not built: run __graft_entry__.build()"
    ld, rd = tmp_path / "cam0" / "data", tmp_path / "cam1" / "data"
    os.makedirs(ld)
    os.makedirs(rd)
    stamps = [1403636579763555584, 1403636579813555456, 1403636579863555584]
    with open(tmp_path / "times.txt", "w") as f:
        f.write("".join("%d\n" % s for s in stamps) + "\n")
    pairs = []
    for i, s in enumerate(stamps):
        L, R = synth.stereo_pair(W, H, 300 + i)
        cv2.imwrite(str(ld / ("%d.png" % s)), L)
        cv2.imwrite(str(rd / ("%d.png" % s)), R)
        pairs.append((L, R))
    with open(tmp_path / "calib.txt", "w") as f:
        for side in ("left", "right"):
            c = EUROC[side]
            for k in "KDRP":
                f.write(" ".join("%.17g" % v for v in np.array(c[k], np.float64).reshape(-1)) + "\n")
        f.write("%d %d\n" % (W, H))
    out = str(tmp_path / "rect.bin")
    r = subprocess.run([exe, "--euroc", str(ld), str(rd), str(tmp_path / "times.txt"), str(tmp_path / "calib.txt"), out], capture_output=True, text=True,
                       timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    lines = [l.split() for l in r.stdout.splitlines() if l.startswith("euroc ")]
    assert len(lines) == 3
    for l, s in zip(lines, stamps):
        assert abs(float(l[1]) - s / 1e9) < 1e-6 and l[2].endswith("cam0/data/%d.png" % s) and l[3].endswith("cam1/data/%d.png" % s)
    got = np.fromfile(out, np.uint8).reshape(2, H, W)
    for side, im, g in (("left", pairs[0][0], got[0]), ("right", pairs[0][1], got[1])):
        c = EUROC[side]
        K, D, R, P = (np.array(c[k], np.float64) for k in "KDRP")
        m1, m2 = cv2.initUndistortRectifyMap(K, D, R, P[:3, :3], (W, H), cv2.CV_32F)
        assert (g != cv2.remap(im, m1, m2, cv2.INTER_LINEAR)).mean() < 1e-4, side
