"""GPU parity tests (through the C-ABI) of the extraction path against the CPU oracle.
Tolerances (BASELINE.json north_star): keypoint x, y, octave, size, response and ORDER bit-exact; angle
within 1e-4 rad; descriptors bit-exact for >= 99.9 % of keypoints."""
import numpy as np
import pytest

import orbb200
import ref_cpu
import synth

pytestmark = pytest.mark.gpu

ANGLE_TOL_DEG = float(np.degrees(1e-4))


def _compare(kg, dg, kc, dc):
    assert len(kg) == len(kc), (len(kg), len(kc))
    for f in ("x", "y", "size", "response", "octave", "class_id"):
        assert np.array_equal(kg[f], kc[f]), (f, np.flatnonzero(kg[f] != kc[f])[:8])
    da = np.abs(kg["angle"].astype(np.float64) - kc["angle"].astype(np.float64))
    da = np.minimum(da, 360.0 - da)
    assert da.max() <= ANGLE_TOL_DEG, da.max()
    bad = (dg != dc).any(axis=1).sum()
    assert bad <= 0.001 * len(kc), "descriptor mismatch on %d of %d keypoints" % (bad, len(kc))
    return int((kg["angle"] != kc["angle"]).sum()), int(bad)


@pytest.mark.parametrize("name", ["tum", "kitti", "euroc"])
def test_stages_and_outputs_match_oracle(configs, name):
    w, h, nf, sfac, nl, ini, mn = configs[name]
    img = synth.textured_frame(w, h, 1000)
    ref = ref_cpu.RefExtractor(nf, sfac, nl, ini, mn)
    kc, dc = ref.extract(img)
    ex = orbb200.ORBextractor(nf, sfac, nl, ini, mn)
    tb = ref.tables()
    assert np.array_equal(ex.GetScaleFactors(), tb["scale"]) and np.array_equal(ex.GetInverseScaleFactors(), tb["inv_scale"])
    assert np.array_equal(ex.GetScaleSigmaSquares(), tb["sigma2"])
    assert np.array_equal(ex.GetInverseScaleSigmaSquares(), tb["inv_sigma2"])
    assert np.array_equal(ex.GetFeaturesPerLevel(), tb["nfeat"])
    kg, dg = ex(img)
    for l in range(nl):
        assert np.array_equal(ex.pyramid_level(l), ref.level_image(l)), ("pyramid", l)
        c = ref.level_candidates(l)
        g = ex.debug_candidates(l)
        assert len(g) == len(c), ("ncand", l, len(g), len(c))
        assert np.array_equal(g[:, 0], c["x"]) and np.array_equal(g[:, 1], c["y"]) and np.array_equal(g[:, 2], c["resp"]), ("cand", l)
        rb = ref.level_blurred(l)
        if rb is not None:
            assert np.array_equal(ex.debug_blurred_level(l), rb), ("blur", l)
    nang, nbad = _compare(kg, dg, kc, dc)
    print(name, "kps", len(kg), "angle bit-diffs", nang, "descriptor row diffs", nbad)


def test_golden_fixture_kitti(configs):
    import os
    path = os.path.join(os.path.dirname(__file__), "golden", "extract_kitti.npz")
    g = np.load(path)
    w, h, nf, sfac, nl, ini, mn = configs["kitti"]
    img = synth.textured_frame(w, h, int(g["seed"]))
    kg, dg = orbb200.ORBextractor(nf, sfac, nl, ini, mn)(img)
    _compare(kg, dg, g["kps"], g["desc"])


def test_batch_equals_single_and_is_deterministic(configs):
    w, h, nf, sfac, nl, ini, mn = configs["euroc"]
    imgs = np.stack([synth.textured_frame(w, h, 2000 + k) for k in range(5)])
    ex = orbb200.ORBextractor(nf, sfac, nl, ini, mn)
    kps, desc, n = ex.extract_batch(imgs)
    kps2, desc2, n2 = ex.extract_batch(imgs)
    assert np.array_equal(n, n2)
    for b in range(5):
        ks, ds = orbb200.ORBextractor(nf, sfac, nl, ini, mn)(imgs[b])
        assert n[b] == len(ks)
        assert np.array_equal(kps[b, :n[b]], ks) and np.array_equal(desc[b, :n[b]], ds)
        assert np.array_equal(kps2[b, :n[b]], ks) and np.array_equal(desc2[b, :n[b]], ds)
    ref = ref_cpu.RefExtractor(nf, sfac, nl, ini, mn)
    for b in (0, 4):
        kc, dc = ref.extract(imgs[b])
        _compare(kps[b, :n[b]], desc[b, :n[b]], kc, dc)


@pytest.mark.parametrize("shape,params", [
    ((200, 150), (300, 1.2, 6, 20, 7)),      # small image, few cells
    ((333, 517), (500, 1.3, 5, 15, 5)),      # portrait, other scale factor and thresholds
    ((1000, 90), (400, 1.2, 4, 12, 7)),      # panorama: many quadtree roots
    ((641, 479), (5000, 1.2, 8, 20, 7)),     # quota above the candidate count (quadtree runs dry)
    ((640, 480), (40, 1.2, 8, 20, 7)),       # tiny quota (N < initial nodes)
])
def test_edge_shapes(shape, params):
    w, h = shape
    img = synth.textured_frame(w, h, 77)
    ref = ref_cpu.RefExtractor(*params)
    kc, dc = ref.extract(img)
    ex = orbb200.ORBextractor(*params)
    kg, dg = ex(img)
    for l in range(params[2]):
        assert np.array_equal(ex.pyramid_level(l), ref.level_image(l)), ("pyramid", l)
    _compare(kg, dg, kc, dc)


def test_flat_and_empty_images():
    ex = orbb200.ORBextractor(1000, 1.2, 8, 20, 7)
    k, d = ex(np.full((480, 640), 128, np.uint8))
    assert len(k) == 0 and d.shape == (0, 32)
    k, d = ex(np.zeros((0, 0), np.uint8))
    assert len(k) == 0
    # unaligned row stride (ROI of a larger image), as cv::Mat ROIs are
    big = synth.textured_frame(700, 500, 5)
    roi = big[7:487, 13:653]
    kc, dc = ref_cpu.RefExtractor(1000, 1.2, 8, 20, 7).extract(np.ascontiguousarray(roi))
    kg, dg = ex(roi)
    _compare(kg, dg, kc, dc)


def test_large_image_4k_properties():
    """BASELINE config 4 (3840x2160, 8000 feats, 12 levels): parity vs the oracle on the full size."""
    img = synth.textured_frame(3840, 2160, 4000)
    params = (8000, 1.2, 12, 20, 7)
    ex = orbb200.ORBextractor(*params)
    kg, dg = ex(img)
    kc, dc = ref_cpu.RefExtractor(*params).extract(img)
    _compare(kg, dg, kc, dc)
    assert len(kg) >= 8000


@pytest.mark.parametrize("layout", ["packed_odd_pitch", "misaligned_base", "aligned_pitch"])
def test_device_api_equals_host_api(configs, layout):
    """orbb200_extract_batch_device on device-resident images (the bench's `value` path), for aligned and unaligned
    caller layouts (packed 1241-byte rows are staged once into the aligned pyramid slot), == the host-buffer API."""
    import torch
    w, h, nf, sfac, nl, ini, mn = configs["kitti"]
    B = 3
    imgs = np.stack([synth.textured_frame(w, h, 3000 + k) for k in range(B)])
    ex = orbb200.ORBextractor(nf, sfac, nl, ini, mn)
    kps_h, desc_h, n_h = ex.extract_batch(imgs)
    if layout == "aligned_pitch":
        pitch, off = 1248, 0
    elif layout == "packed_odd_pitch":
        pitch, off = w, 0
    else:
        pitch, off = w + 2, 3
    img_stride = pitch * h + (13 if layout == "misaligned_base" else 0)
    if layout == "aligned_pitch":
        img_stride = (img_stride + 255) // 256 * 256
    raw = torch.zeros(off + img_stride * B + 64, dtype=torch.uint8, device="cuda")
    for b in range(B):
        view = raw[off + b * img_stride: off + b * img_stride + pitch * h].view(h, pitch)
        view[:, :w] = torch.from_numpy(imgs[b]).cuda()
    cap = ex.cap
    kps_d = torch.zeros((B, cap, 7), dtype=torch.float32, device="cuda")
    desc_d = torch.zeros((B, cap, 32), dtype=torch.uint8, device="cuda")
    n_d = torch.zeros(B, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    ex2 = orbb200.ORBextractor(nf, sfac, nl, ini, mn)
    ex2.extract_batch_device(raw.data_ptr() + off, B, w, h, pitch, img_stride, kps_d.data_ptr(), desc_d.data_ptr(), cap, n_d.data_ptr())
    ex2.sync()
    n = n_d.cpu().numpy()
    assert np.array_equal(n, n_h)
    kd = kps_d.cpu().numpy().view(np.uint8).reshape(B, cap, 28)
    for b in range(B):
        assert np.array_equal(kd[b, :n[b]], kps_h[b, :n[b]].view(np.uint8).reshape(-1, 28))
        assert np.array_equal(desc_d[b, :n[b]].cpu().numpy(), desc_h[b, :n[b]])
    for l in range(nl):
        assert np.array_equal(ex2.pyramid_level(l, 1), ex.pyramid_level(l, 1)), l


def test_pipeline_settings_do_not_change_results(configs):
    """orbb200_extractor_set_pipeline only reschedules: any chunk size / stream count of the host pipeline, and the
    optional chunking of the device variant (aligned and staged layouts), return byte-identical outputs, and the
    stereo matcher that follows on the handle's stream sees all chunks."""
    import torch
    w, h, nf, sfac, nl, ini, mn = configs["kitti"]
    B = 10
    imgs = np.stack([synth.textured_frame(w, h, 4100 + k) for k in range(B)])
    ex = orbb200.ORBextractor(nf, sfac, nl, ini, mn)
    ex.set_pipeline(B, 1, 0)
    kps0, desc0, n0 = ex.extract_batch(imgs)
    st0 = ex.stereo_match_batch(B // 2, 386.1448, 386.1448 / 718.856)
    for chunk, ns in [(1, 4), (3, 2), (4, 3), (7, 4)]:
        ex.set_pipeline(chunk, ns, 0)
        kps, desc, n = ex.extract_batch(imgs)
        assert np.array_equal(n, n0), (chunk, ns)
        for b in range(B):
            assert np.array_equal(kps[b, :n[b]].view(np.uint8), kps0[b, :n[b]].view(np.uint8)), (chunk, ns, b)
            assert np.array_equal(desc[b, :n[b]], desc0[b, :n[b]]), (chunk, ns, b)
        st = ex.stereo_match_batch(B // 2, 386.1448, 386.1448 / 718.856)
        for a, b_ in zip(st, st0):
            assert np.array_equal(a, b_), (chunk, ns)
    cap = ex.cap
    for pitch in (w, 1248):
        raw = torch.zeros((B, h, pitch), dtype=torch.uint8, device="cuda")
        raw[:, :, :w] = torch.from_numpy(imgs).cuda()
        for chunk, ns in [(3, 4), (4, 2)]:
            kps_d = torch.zeros((B, cap, 7), dtype=torch.float32, device="cuda")
            desc_d = torch.zeros((B, cap, 32), dtype=torch.uint8, device="cuda")
            n_d = torch.zeros(B, dtype=torch.int32, device="cuda")
            torch.cuda.synchronize()
            ex.set_pipeline(32, ns, chunk)
            ex.extract_batch_device(raw.data_ptr(), B, w, h, pitch, pitch * h, kps_d.data_ptr(), desc_d.data_ptr(), cap, n_d.data_ptr())
            uR, dep, nv = ex.stereo_match_batch(B // 2, 386.1448, 386.1448 / 718.856)  # ordered after every chunk
            n = n_d.cpu().numpy()
            assert np.array_equal(n, n0), (pitch, chunk, ns)
            kd = kps_d.cpu().numpy().view(np.uint8).reshape(B, cap, 28)
            for b in range(B):
                assert np.array_equal(kd[b, :n[b]], kps0[b, :n[b]].view(np.uint8).reshape(-1, 28)), (pitch, chunk, ns, b)
                assert np.array_equal(desc_d[b, :n[b]].cpu().numpy(), desc0[b, :n[b]]), (pitch, chunk, ns, b)
            assert np.array_equal(uR, st0[0]) and np.array_equal(nv, st0[2]), (pitch, chunk, ns)


def test_single_image_graph_path_equals_batch_pipeline(configs):
    """orbb200_extract replays a CUDA graph per (shape, capacity); results must equal the batch pipeline's (graph off), also
    when shapes alternate (re-capture), for the colour prologue, and the lazy pyramid slab must equal the per-level copies"""
    rng_shapes = [("tum", 1000), ("kitti", 1001), ("tum", 1003)]
    ex = orbb200.ORBextractor(1000, 1.2, 8, 20, 7)
    for name, seed in rng_shapes:
        w, h = configs[name][0], configs[name][1]
        img = synth.textured_frame(w, h, seed)
        ex.set_graph(1)
        k1, d1 = ex(img)
        k1b, d1b = ex(img)  # second call replays the captured graph
        pyr = ex.pyramid(0)
        for l in range(8):
            assert np.array_equal(pyr[l], ex.pyramid_level(l)), l
        ex.set_graph(0)
        k0, d0 = ex(img)
        assert np.array_equal(k1, k0) and np.array_equal(d1, d0)
        assert np.array_equal(k1b, k0) and np.array_equal(d1b, d0)
    img = synth.textured_frame(640, 480, 5)
    col = np.stack([img, img[::-1].copy(), img[:, ::-1].copy()], axis=2)
    ex.set_graph(1)
    kc1, dc1 = ex.extract_color(col, 1)
    ex.set_graph(0)
    kc0, dc0 = ex.extract_color(col, 1)
    assert np.array_equal(kc1, kc0) and np.array_equal(dc1, dc0) and len(kc1) > 500
    # a strided view (rows not contiguous) through the graph path
    big = np.zeros((480, 700), np.uint8)
    big[:, :640] = img
    ex.set_graph(1)
    ks, ds = ex(big[:, :640])
    kr, dr = ref_cpu.RefExtractor(1000, 1.2, 8, 20, 7).extract(img)
    _compare(ks, ds, kr, dr)


@pytest.mark.parametrize("name", ["tum", "kitti"])
def test_quadtree_formulations_agree(configs, name):
    """DistributeOctTree on the device (src/ORBextractor.cc:617-915): the histogram formulation (default depth), the histogram
    formulation forced to overflow its depth (2, 3: blocks fall back mid-way) and the iterative formulation alone (0) must give the
    oracle's keypoints -- also on a clustered image (texture in one corner only: deep, unbalanced tree) and in a batch."""
    w, h, nf, sfac, nl, ini, mn = configs[name]
    img = synth.textured_frame(w, h, 77)
    clustered = np.full_like(img, 128)
    clustered[h // 2:h // 2 + h // 5, w // 3:w // 3 + w // 6] = img[h // 2:h // 2 + h // 5, w // 3:w // 3 + w // 6]
    clustered[20:60, 20:90] = img[20:60, 20:90]
    ref = ref_cpu.RefExtractor(nf, sfac, nl, ini, mn)
    ex = orbb200.ORBextractor(nf, sfac, nl, ini, mn)
    for im in (img, clustered):
        kc, dc = ref.extract(im)
        assert len(kc) > 50
        for depth in (7, 0, 2, 3, 7):
            ex.set_octree_depth(depth)
            for graph in (1, 0):
                ex.set_graph(graph)
                kg, dg = ex(im)
                _compare(kg, dg, kc, dc)
    ex.set_octree_depth(7)
    batch = np.stack([img, clustered, img])
    kps, desc, n = ex.extract_batch(batch)
    for b, im in enumerate((img, clustered, img)):
        kc, dc = ref.extract(im)
        _compare(kps[b, :n[b]], desc[b, :n[b]], kc, dc)


def test_begin_wait_stream_of_stereo_batches_equals_synchronous_calls(configs):
    """orbb200_extract_stereo_batch_begin / orbb200_extract_batch_wait: a stream of steps with two calls in flight on one handle (the
    copies of step k + 1 under the tail of step k, device output blocks and staging buffers reused) gives, step by step, the results
    of the synchronous call -- also when a call of another size comes in between and when a third begin arrives before a wait."""
    import ctypes as C
    w, h, nf, sfac, nl, ini, mn = configs["kitti"]
    L = orbb200.lib()
    ex = orbb200.ORBextractor(nf, sfac, nl, ini, mn)
    ex.set_pipeline(4, 8)  # several chunks per call at this small size
    cap = ex.cap
    bf, b = 386.1448, 386.1448 / 718.856
    steps = []
    for s in range(5):
        npairs = 7 if s != 3 else 3  # step 3 has another size: the calls in flight are completed first
        imgs = np.ascontiguousarray(np.stack([v for p in range(npairs) for v in synth.stereo_pair(w, h, 100 * s + p)]))
        steps.append((npairs, imgs))

    def buffers(npairs):
        return dict(kps=np.zeros((2 * npairs, cap), orbb200.KP_DTYPE), desc=np.zeros((2 * npairs, cap, 32), np.uint8), n=np.zeros(2 * npairs, np.int32),
                    uR=np.zeros((npairs, cap), np.float32), dep=np.zeros((npairs, cap), np.float32), nv=np.zeros(npairs, np.int32))

    def ptrs(imgs):
        return (C.c_void_p * len(imgs))(*[imgs[i].ctypes.data for i in range(len(imgs))])

    want = []
    for npairs, imgs in steps:
        o = buffers(npairs)
        orbb200.check(L.orbb200_extract_stereo_batch(ex.h, ptrs(imgs), npairs, w, h, w, orbb200._p(o["kps"]), orbb200._p(o["desc"]), cap, orbb200._p(o["n"]),
                                                     bf, b, orbb200._p(o["uR"]), orbb200._p(o["dep"]), orbb200._p(o["nv"])))
        want.append(o)
    got, tickets, keep = [], [], []
    for npairs, imgs in steps:  # begin all five back to back: never more than two in flight, the third begin completes the oldest
        o = buffers(npairs)
        t = C.c_int(-5)
        pp = ptrs(imgs)
        keep.append(pp)
        orbb200.check(L.orbb200_extract_stereo_batch_begin(ex.h, pp, npairs, w, h, w, orbb200._p(o["kps"]), orbb200._p(o["desc"]), cap, orbb200._p(o["n"]),
                                                           bf, b, orbb200._p(o["uR"]), orbb200._p(o["dep"]), orbb200._p(o["nv"]), C.byref(t)))
        assert t.value in (0, 1)
        got.append(o)
        tickets.append(t.value)
    for t in tickets[-2:]:
        orbb200.check(L.orbb200_extract_batch_wait(ex.h, t))
    orbb200.check(L.orbb200_extract_batch_wait(ex.h, tickets[0]))  # already complete: a no-op
    for s, (a, g) in enumerate(zip(want, got)):
        assert np.array_equal(a["n"], g["n"]) and a["n"].min() > 500, s
        assert np.array_equal(a["nv"], g["nv"]) and a["nv"].min() > 100, s
        for i in range(len(a["n"])):
            k = a["n"][i]
            assert np.array_equal(a["kps"][i, :k], g["kps"][i, :k]) and np.array_equal(a["desc"][i, :k], g["desc"][i, :k]), (s, i)
        for p in range(len(a["nv"])):
            k = a["n"][2 * p]
            assert np.array_equal(a["uR"][p, :k], g["uR"][p, :k]) and np.array_equal(a["dep"][p, :k], g["dep"][p, :k]), (s, p)
    # the handle is back in its synchronous state: a single-image call right after
    k1, d1 = ex(steps[0][1][0])
    assert np.array_equal(k1, want[0]["kps"][0, :want[0]["n"][0]])
