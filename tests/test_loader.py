"""include/orbb200/DatasetLoader.h (the KITTI / TUM readers of Examples/Stereo/stereo_kitti.cc:130-160 and
Examples/Monocular/mono_tum.cc:128-160, batched into pinned memory): decoded bytes against OpenCV's imread, list parsing,
batching; on a GPU also extraction of the loaded batches against a direct call."""
import os
import subprocess

import numpy as np
import pytest

import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "orb-slam2_b200", "loader_test.bin")


def _make_sequence(td, npairs, w=311, h=97):
    cv2 = pytest.importorskip("cv2")
    os.makedirs(os.path.join(td, "image_0"))
    os.makedirs(os.path.join(td, "image_1"))
    os.makedirs(os.path.join(td, "rgb"))
    imgs = []
    with open(os.path.join(td, "times.txt"), "w") as f:
        for i in range(npairs):
            f.write("%.6e\n" % (0.1 * i))
            L, R = synth.stereo_pair(max(w, 128), max(h, 128), 40 + i)
            L, R = np.ascontiguousarray(L[:h, :w]), np.ascontiguousarray(R[:h, :w])
            if i % 3 == 2:  # some files as binary PGM under the .png name (the decoder goes by content)
                for name, im in (("image_0", L), ("image_1", R)):
                    with open(os.path.join(td, name, "%06d.png" % i), "wb") as g:
                        g.write(b"P5\n# made by test\n%d %d\n255\n" % (w, h) + im.tobytes())
            else:
                params = [cv2.IMWRITE_PNG_COMPRESSION, 1 + i % 9, cv2.IMWRITE_PNG_STRATEGY, i % 5]  # exercises all five row filters
                cv2.imwrite(os.path.join(td, "image_0", "%06d.png" % i), L, params)
                cv2.imwrite(os.path.join(td, "image_1", "%06d.png" % i), R, params)
            imgs += [L, R]
    rgb = np.stack([imgs[0], imgs[1], imgs[2]], axis=2)  # BGR order for cv2.imwrite
    cv2.imwrite(os.path.join(td, "rgb", "1.000000.png"), rgb)
    with open(os.path.join(td, "rgb.txt"), "w") as f:
        f.write("# color images\n# file: 'x.bag'\n# timestamp filename\n1.000000 rgb/1.000000.png\n2.500000 rgb/2.500000.png\n")
    return imgs, rgb


def test_loader_decodes_like_opencv(tmp_path):
    cv2 = pytest.importorskip("cv2")
    assert os.path.exists(EXE), "loader_test.bin not built: run __graft_entry__.build()"
    td = str(tmp_path / "seq")
    os.makedirs(td)
    imgs, rgb = _make_sequence(td, 7)
    out = str(tmp_path / "out.bin")
    r = subprocess.run([EXE, td, "3", out], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    lines = r.stdout.splitlines()
    assert lines[0] == "pairs 7" and "shape 311 97 batches 3" in lines
    assert [l for l in lines if l.startswith("tum ")] == ["tum 1.000000 rgb/1.000000.png", "tum 2.500000 rgb/2.500000.png"]
    assert [l.split()[2:5] for l in lines if l.startswith("batch")] == [["0", "npairs", "3"], ["3", "npairs", "3"], ["6", "npairs", "1"]]
    got = np.fromfile(out, np.uint8).reshape(14, 97, 311)
    for i in range(14):
        assert np.array_equal(got[i], imgs[i]), i
        path = os.path.join(td, "image_%d" % (i % 2), "%06d.png" % (i // 2))
        if (i // 2) % 3 != 2:
            assert np.array_equal(got[i], cv2.imread(path, cv2.IMREAD_GRAYSCALE))
    g = np.fromfile(out + ".tum0", np.uint8).reshape(97, 311)
    want = cv2.cvtColor(cv2.imread(os.path.join(td, "rgb", "1.000000.png"), cv2.IMREAD_COLOR), cv2.COLOR_BGR2GRAY)
    assert np.array_equal(g, want)


def test_loader_reports_missing_and_bad_files(tmp_path):
    assert os.path.exists(EXE)
    td = str(tmp_path / "seq")
    os.makedirs(os.path.join(td, "image_0"))
    os.makedirs(os.path.join(td, "image_1"))
    open(os.path.join(td, "times.txt"), "w").write("0\n1\n")
    r = subprocess.run([EXE, td, "2", str(tmp_path / "o.bin")], capture_output=True, text=True, timeout=60)
    assert r.returncode == 3 and "cannot read" in r.stdout
    open(os.path.join(td, "image_0", "000000.png"), "wb").write(b"\x89PNG\r\n\x1a\n" + b"\0" * 40)
    r = subprocess.run([EXE, td, "2", str(tmp_path / "o.bin")], capture_output=True, text=True, timeout=60)
    assert r.returncode == 3 and "error" in r.stdout


@pytest.mark.gpu
def test_loaded_batches_extract_like_direct_calls(tmp_path):
    import orbb200
    td = str(tmp_path / "seq")
    os.makedirs(td)
    imgs, _ = _make_sequence(td, 5, w=640, h=200)
    out = str(tmp_path / "out.bin")
    r = subprocess.run([EXE, td, "2", out, "--extract", "600"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    n = {int(l.split()[1]): int(l.split()[2]) for l in r.stdout.splitlines() if l.startswith("n ")}
    nv = {int(l.split()[1]): int(l.split()[2]) for l in r.stdout.splitlines() if l.startswith("nv ")}
    ex = orbb200.ORBextractor(600, 1.2, 8, 20, 7)
    kps, desc, nn, uR, dep, nvv = ex.extract_stereo_batch(np.stack(imgs), 386.1448, 386.1448 / 718.856)
    assert [n[i] for i in range(10)] == nn.tolist() and [nv[i] for i in range(5)] == nvv.tolist()
