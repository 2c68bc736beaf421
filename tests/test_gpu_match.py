"""GPU parity tests (through the C-ABI) of the matcher kernels against the CPU oracle: match indices bit-exact."""
import numpy as np
import pytest

import orbb200
import ref_cpu
import synth

pytestmark = pytest.mark.gpu


def _eq(a, b, fields=("best_dist", "best_idx", "second_dist", "second_idx")):
    for f in fields:
        assert np.array_equal(a[f], b[f]), (f, np.flatnonzero(a[f] != b[f])[:8])


def test_descriptor_distance():
    rng = np.random.default_rng(3)
    a = rng.integers(0, 256, (5000, 32), dtype=np.uint8)
    b = rng.integers(0, 256, (5000, 32), dtype=np.uint8)
    b[10] = a[10]
    b[11] = ~a[11]
    d = orbb200.ORBmatcher().DescriptorDistance(a, b)
    ref = np.unpackbits(a ^ b, axis=1).sum(axis=1)
    assert np.array_equal(d, ref) and d[10] == 0 and d[11] == 256
    assert ref_cpu.descriptor_distance(a[0], b[0]) == d[0]


@pytest.mark.parametrize("nq,ndb", [(1, 1), (3, 5), (37, 127), (64, 128), (513, 1000), (2048, 20000)])
def test_bruteforce_top2_equals_oracle(nq, ndb):
    q, db, _ = synth.descriptor_db(nq, ndb, seed=nq + ndb)
    # duplicates and ties: copies of rows at several positions
    if ndb > 40:
        db[ndb - 1] = db[7]
        db[20] = db[7]
        q[0] = db[7]
    t = orbb200.ORBmatcher().bruteforce_top2(q, db)
    _eq(t, ref_cpu.bf_top2(q, db))


def test_bruteforce_two_streams_from_one_thread():
    """two device-resident searches issued back to back from one host thread on two different caller streams share the
    thread's partial top-2 buffer: the second is ordered behind the first and both results equal the oracle"""
    import torch
    dev = torch.device("cuda", 0)
    L = orbb200.lib()
    q1, db1, _ = synth.descriptor_db(512, 60000, seed=3)
    q2, db2, _ = synth.descriptor_db(512, 90000, seed=4)
    tq1, tdb1 = torch.from_numpy(q1).to(dev), torch.from_numpy(db1).to(dev)
    tq2, tdb2 = torch.from_numpy(q2).to(dev), torch.from_numpy(db2).to(dev)
    o1 = torch.empty((512, 4), dtype=torch.int32, device=dev)
    o2 = torch.empty((512, 4), dtype=torch.int32, device=dev)
    s1, s2 = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
    torch.cuda.synchronize()
    for rep in range(3):
        orbb200.check(L.orbb200_match_bruteforce_top2_device(tq1.data_ptr(), 512, tdb1.data_ptr(), len(db1), o1.data_ptr(), 0, s1.cuda_stream))
        orbb200.check(L.orbb200_match_bruteforce_top2_device(tq2.data_ptr(), 512, tdb2.data_ptr(), len(db2), o2.data_ptr(), 0, s2.cuda_stream))
    torch.cuda.synchronize()
    r1, r2 = ref_cpu.bf_top2(q1, db1), ref_cpu.bf_top2(q2, db2)
    for o, r in ((o1, r1), (o2, r2)):
        g = o.cpu().numpy()
        assert np.array_equal(g[:, 0], r["best_dist"]) and np.array_equal(g[:, 1], r["best_idx"])
        assert np.array_equal(g[:, 2], r["second_dist"]) and np.array_equal(g[:, 3], r["second_idx"])


def test_bruteforce_all_equal_rows_first_index_wins():
    db = np.zeros((3000, 32), np.uint8)
    q = np.zeros((10, 32), np.uint8)
    t = orbb200.ORBmatcher().bruteforce_top2(q, db)
    assert (t["best_idx"] == 0).all() and (t["second_idx"] == 1).all() and (t["best_dist"] == 0).all()
    _eq(t, ref_cpu.bf_top2(q, db))


def test_sharded_merge_equals_flat():
    """config 5 structure: database split by rows, per-shard top-2 with global indices, then merge."""
    nq, ndb, shards = 256, 8192, 8
    q, db, _ = synth.descriptor_db(nq, ndb, seed=5)
    db[5000] = db[100]
    q[3] = db[100]
    m = orbb200.ORBmatcher()
    parts = np.zeros((shards, nq), orbb200.TOP2_DTYPE)
    per = ndb // shards
    for s in range(shards):
        p = m.bruteforce_top2(q, db[s * per:(s + 1) * per])
        for f in ("best_idx", "second_idx"):
            p[f] = np.where(p[f] >= 0, p[f] + s * per, -1)
        parts[s] = p
    merged = m.merge_top2(parts)
    _eq(merged, ref_cpu.bf_top2(q, db))
    acc = m.accept_ratio(merged)
    r = ref_cpu.bf_top2(q, db)
    ref_acc = np.array([ref_cpu.lib().ref_ratio_accept(int(a), int(b), 50, float(np.float32(0.6))) for a, b in
                        zip(r["best_dist"], r["second_dist"])], bool)
    assert np.array_equal(acc, ref_acc) and acc.any() and not acc.all()


@pytest.mark.parametrize("nt", [1, 31, 2048, 8192, 8193, 20000])
def test_window_search_grid_build_paths(nt):
    """k_grid_build orders a frame's keypoints by sorting (cell, index) keys in shared memory up to 8 192 keypoints and by an
    ordered one-warp fill beyond: both must reproduce AssignFeaturesToGrid's insertion order (src/Frame.cc:261-279), also for
    keypoints outside the grid and many keypoints per cell."""
    rng = np.random.default_rng(nt)
    w, h = 752.0, 480.0
    kps = np.zeros(nt, orbb200.KP_DTYPE)
    kps["x"] = rng.uniform(-20, w + 20, nt).astype(np.float32)
    kps["y"] = rng.uniform(-20, h + 20, nt).astype(np.float32)
    kps["x"][: nt // 3] = np.float32(100.0) + rng.uniform(0, 5, nt // 3).astype(np.float32)  # a crowded cell
    kps["y"][: nt // 3] = np.float32(100.0)
    kps["octave"] = rng.integers(0, 8, nt)
    desc = rng.integers(0, 256, (nt, 32), dtype=np.uint8)
    invW, invH = np.float32(64) / np.float32(w), np.float32(48) / np.float32(h)
    nq = 300
    win = np.zeros((nq, 5), np.float32)
    win[:, 0], win[:, 1] = rng.uniform(0, w, nq), rng.uniform(0, h, nq)
    win[:40, 0], win[:40, 1] = 102.0, 100.0
    win[:, 2] = rng.choice([5.0, 20.0, 60.0], nq)
    win[:, 3], win[:, 4] = -1, -1
    qdesc = rng.integers(0, 256, (nq, 32), dtype=np.uint8)
    offs, idx, dist = orbb200.ORBmatcher().window_search(kps, desc, (0.0, 0.0, float(invW), float(invH), 64, 48), win, qdesc)
    start, items = ref_cpu.assign_grid(kps, 0.0, 0.0, invW, invH, 64, 48)
    for q in range(nq):
        ref = ref_cpu.features_in_area(kps, start, items, 0.0, 0.0, invW, invH, win[q, 0], win[q, 1], win[q, 2], -1, -1, 64, 48)
        assert np.array_equal(idx[offs[q]:offs[q + 1]], ref), (q, idx[offs[q]:offs[q + 1]][:8], ref[:8])
        if len(ref):
            assert np.array_equal(dist[offs[q]:offs[q + 1]], np.unpackbits(desc[ref] ^ qdesc[q], axis=1).sum(axis=1))


def test_candidate_distances_equal_oracle_and_reference():
    """orbb200_candidate_distances (k_candidate_distances): the distance step of both SearchByBoW overloads and of
    SearchForTriangulation (src/ORBmatcher.cc:214-244, 626-650, 780-815) on random CSR candidate lists -- empty lists, lists
    longer than a warp, repeated targets -- against the C oracle and the reference's own DescriptorDistance."""
    rng = np.random.default_rng(21)
    q, t, _ = synth.descriptor_db(700, 1900, seed=5)
    sizes = rng.choice([0, 0, 1, 2, 5, 31, 32, 33, 64, 100, 257], size=len(q))
    offs = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    idx = rng.integers(0, len(t), int(offs[-1])).astype(np.int32)
    idx[offs[5]:offs[6]] = idx[offs[5]] if sizes[5] else idx[offs[5]:offs[6]]  # one list of a single repeated target
    m = orbb200.ORBmatcher()
    d = m.candidate_distances(q, t, offs, idx)
    want = np.unpackbits(np.repeat(q, sizes, axis=0) ^ t[idx], axis=1).sum(axis=1).astype(np.uint16)
    assert np.array_equal(d, want)
    import orbslam_ref
    for c in rng.integers(0, len(idx), 300):
        qi = int(np.searchsorted(offs, c, side="right") - 1)
        assert int(d[c]) == ref_cpu.descriptor_distance(q[qi], t[idx[c]])
        if orbslam_ref.available():
            assert int(d[c]) == orbslam_ref.descriptor_distance(q[qi], t[idx[c]])
    # top-2 of the same lists
    _eq(m.match_candidates(q, t, offs, idx), ref_cpu.match_candidates(q, t, offs, idx))
    # all lists empty / no queries
    z = np.zeros(len(q) + 1, np.int32)
    assert len(m.candidate_distances(q, t, z, np.zeros(0, np.int32))) == 0
    # malformed lists are rejected on the host, not read out of bounds on the device
    bad = idx.copy()
    bad[3] = len(t)
    with pytest.raises(Exception):
        m.candidate_distances(q, t, offs, bad)
    bad_offs = offs.copy()
    bad_offs[4], bad_offs[5] = bad_offs[5] + 1, bad_offs[4]
    if bad_offs[4] > bad_offs[5]:
        with pytest.raises(Exception):
            m.candidate_distances(q, t, bad_offs, idx)


def test_search_triangulation_equals_oracle():
    """orbb200_search_triangulation (k_triangulation_search): the whole per-query loop of SearchForTriangulation on the device --
    distances, TH_LOW, epipole gate, CheckDistEpipolarLine, later-candidate-wins ties -- against the C restatement
    (src/ORBmatcher.cc:757-815, 147-164) on keyframe-sized inputs (2 000 queries x ~50 candidates), incl. empty lists"""
    rng = np.random.default_rng(33)
    nq, nt = 2000, 2100
    t = rng.integers(0, 256, (nt, 32), dtype=np.uint8)
    tkp = np.zeros(nt, orbb200.KP_DTYPE)
    tkp["x"], tkp["y"] = rng.uniform(0, 640, nt).astype(np.float32), rng.uniform(0, 480, nt).astype(np.float32)
    tkp["octave"] = rng.integers(0, 8, nt)
    q = np.zeros((nq, 32), np.uint8)
    qxy = np.zeros((nq, 2), np.float32)
    sizes = rng.choice([0, 1, 20, 50, 64, 90], size=nq)
    offs = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    idx = rng.integers(0, nt, int(offs[-1])).astype(np.int32)
    for i in range(nq):  # several candidates of a query are near copies at the SAME distance on the query's epipolar line (a row)
        qxy[i] = (rng.uniform(0, 640), rng.uniform(20, 460))
        q[i] = rng.integers(0, 256, 32, dtype=np.uint8)
        for c in range(offs[i], offs[i + 1]):
            if rng.random() < 0.35:
                j = idx[c]
                bits = np.unpackbits(q[i])
                bits[rng.choice(256, int(rng.choice([3, 3, 10, 30, 60])), replace=False)] ^= 1
                t[j] = np.packbits(bits)
                tkp["y"][j] = qxy[i, 1] + np.float32(rng.normal(0, 1.5))
    qst = (rng.random(nq) < 0.3).astype(np.uint8)
    tst = (rng.random(nt) < 0.3).astype(np.uint8)
    F12 = np.array([[0, 0, 0], [0, 0, -1], [0, 1, 0]], np.float32) + rng.normal(0, 1e-4, (3, 3)).astype(np.float32)
    sf = np.float32(1.2) ** np.arange(8, dtype=np.float32)
    sig2 = (sf * sf).astype(np.float32)
    ex, ey = 300.5, 240.25
    gi, gd = orbb200.ORBmatcher().search_triangulation(q, qxy, qst, t, tkp, tst, offs, idx, F12, ex, ey, sf, sig2)
    ri, rd = ref_cpu.search_triangulation(q, qxy, qst, t, tkp, tst, offs, idx, F12, ex, ey, sf, sig2)
    assert np.array_equal(gi, ri) and np.array_equal(gd, rd)
    assert (ri >= 0).sum() > 80 and (ri[sizes == 0] == -1).all()  # (targets are shared between lists: later queries overwrite planted copies)


def test_match_candidates_equals_oracle():
    rng = np.random.default_rng(9)
    nq, nt = 1500, 1800
    q = rng.integers(0, 256, (nq, 32), dtype=np.uint8)
    t = rng.integers(0, 256, (nt, 32), dtype=np.uint8)
    t[5] = t[900]
    counts = rng.integers(0, 90, nq)
    counts[0] = 0
    counts[1] = 1
    counts[2] = 400
    offs = np.zeros(nq + 1, np.int32)
    offs[1:] = np.cumsum(counts)
    idx = rng.integers(0, nt, offs[-1]).astype(np.int32)
    idx[offs[2]:offs[2] + 3] = [900, 5, 900]  # equal distances: first in list order wins
    g = orbb200.ORBmatcher().match_candidates(q, t, offs, idx)
    _eq(g, ref_cpu.match_candidates(q, t, offs, idx))
    assert g["best_idx"][0] == -1 and g["best_dist"][0] == 256


def test_stereo_match_equals_oracle(configs):
    """BASELINE config 2: KITTI-shape stereo pair, extraction of both images + ComputeStereoMatches."""
    w, h, nf, sfac, nl, ini, mn = configs["kitti"]
    left, right = synth.stereo_pair(w, h, 1001)
    exL, exR = orbb200.ORBextractor(nf, sfac, nl, ini, mn), orbb200.ORBextractor(nf, sfac, nl, ini, mn)
    kL, dL = exL(left)
    kR, dR = exR(right)
    bf, fx = 386.1448, 718.856
    b = bf / fx
    uR, depth, nvalid = orbb200.stereo_match(exL, exR, bf, b)
    pyrL = [exL.pyramid_level(l) for l in range(nl)]
    pyrR = [exR.pyramid_level(l) for l in range(nl)]
    ruR, rdepth, _, rn = ref_cpu.stereo_match(kL, dL, kR, dR, pyrL, pyrR, exL.GetScaleFactors(), exL.GetInverseScaleFactors(), bf, b)
    n = len(kL)
    assert rn > 200, rn
    assert nvalid == rn
    assert np.array_equal(uR[:n], ruR) and np.array_equal(depth[:n], rdepth)


def test_stereo_match_batch_equals_single(configs):
    """batched F1 (pairs interleaved in one extract_batch call) == the two-handle single-pair call."""
    w, h, nf, sfac, nl, ini, mn = configs["kitti"]
    bf, fx = 386.1448, 718.856
    b = bf / fx
    pairs = [synth.stereo_pair(w, h, 1100 + k) for k in range(3)]
    imgs = np.stack([im for p in pairs for im in p])
    ex = orbb200.ORBextractor(nf, sfac, nl, ini, mn)
    kps, desc, n = ex.extract_batch(imgs)
    uR, depth, nv = ex.stereo_match_batch(3, bf, b)
    for p in range(3):
        exL, exR = orbb200.ORBextractor(nf, sfac, nl, ini, mn), orbb200.ORBextractor(nf, sfac, nl, ini, mn)
        kL, _ = exL(pairs[p][0])
        exR(pairs[p][1])
        suR, sdepth, snv = orbb200.stereo_match(exL, exR, bf, b)
        nL = len(kL)
        assert nL == n[2 * p] and snv == nv[p] and snv > 200
        assert np.array_equal(uR[p, :nL], suR[:nL]) and np.array_equal(depth[p, :nL], sdepth[:nL])


def test_window_search_equals_oracle_grid_and_area(configs):
    """F3 (AssignFeaturesToGrid) + F2 (GetFeaturesInArea) + distances on the device == the oracle's restatement of
    src/Frame.cc:261-279, 386-442: same candidates in the same order for every window, same distances."""
    w, h, nf, sfac, nl, ini, mn = configs["tum"]
    img = synth.textured_frame(w, h, 4242)
    kps, desc = orbb200.ORBextractor(nf, sfac, nl, ini, mn)(img)
    rng = np.random.default_rng(17)
    minX, minY, maxX, maxY = 0.0, 0.0, float(w), float(h)
    cols, rows = 64, 48
    invW, invH = np.float32(cols) / np.float32(maxX - minX), np.float32(rows) / np.float32(maxY - minY)
    nq = 700
    src = rng.integers(0, len(kps), nq)
    win = np.zeros((nq, 5), np.float32)
    win[:, 0] = kps["x"][src] + rng.normal(0, 3, nq)
    win[:, 1] = kps["y"][src] + rng.normal(0, 3, nq)
    win[:, 2] = rng.choice([2.5, 4.0, 15.0, 40.0, 100.0], nq) * rng.choice([1.0, 1.2, 1.44], nq)
    lev = kps["octave"][src]
    win[:, 3] = np.where(rng.random(nq) < 0.3, -1, lev - 1)
    win[:, 4] = np.where(win[:, 3] < 0, np.where(rng.random(nq) < 0.5, -1, lev), lev)
    win[0, :2] = (-500, -500)        # window entirely outside the grid
    win[1, :3] = (5, 5, 30)          # clipped at the image corner
    win[2, :3] = (w - 2, h - 2, 50)
    qdesc = desc[src].copy()
    qdesc ^= rng.integers(0, 256, qdesc.shape, dtype=np.uint8) & rng.integers(0, 256, qdesc.shape, dtype=np.uint8) & 0x11
    offs, idx, dist = orbb200.ORBmatcher().window_search(kps, desc, (minX, minY, float(invW), float(invH), cols, rows), win, qdesc, cap=100)
    start, items = ref_cpu.assign_grid(kps, minX, minY, invW, invH, cols, rows)
    total = 0
    for q in range(nq):
        ref = ref_cpu.features_in_area(kps, start, items, minX, minY, invW, invH, win[q, 0], win[q, 1], win[q, 2],
                                       int(win[q, 3]), int(win[q, 4]), cols, rows)
        got = idx[offs[q]:offs[q + 1]]
        assert np.array_equal(got, ref), (q, got[:10], ref[:10])
        d = np.unpackbits(desc[ref] ^ qdesc[q], axis=1).sum(axis=1) if len(ref) else np.zeros(0)
        assert np.array_equal(dist[offs[q]:offs[q + 1]], d), q
        total += len(ref)
    assert offs[0] == 0 and offs[-1] == total and total > 5000 and offs[1] == 0


def test_fused_extract_stereo_equals_two_calls(configs):
    """orbb200_extract_stereo_batch (the stereo Frame constructor as one pipelined call) == orbb200_extract_batch followed by
    orbb200_stereo_match_batch, for several chunkings of the pipeline (odd chunk sizes are rounded to whole pairs)."""
    w, h, nf, sfac, nl, ini, mn = configs["kitti"]
    P = 5
    imgs = np.empty((2 * P, h, w), np.uint8)
    for p in range(P):
        imgs[2 * p], imgs[2 * p + 1] = synth.stereo_pair(w, h, 7000 + p)
    bf, b = 386.1448, 386.1448 / 718.856
    ex = orbb200.ORBextractor(nf, sfac, nl, ini, mn)
    kps0, desc0, n0 = ex.extract_batch(imgs)
    uR0, dep0, nv0 = ex.stereo_match_batch(P, bf, b)
    assert nv0.min() > 100
    for chunk, ns in [(16, 8), (2, 3), (5, 2), (4, 8)]:
        ex2 = orbb200.ORBextractor(nf, sfac, nl, ini, mn)
        ex2.set_pipeline(chunk, ns, 0)
        kps, desc, n, uR, dep, nv = ex2.extract_stereo_batch(imgs, bf, b)
        assert np.array_equal(n, n0) and np.array_equal(nv, nv0), (chunk, ns)
        for i in range(2 * P):
            assert np.array_equal(kps[i, :n[i]].view(np.uint8), kps0[i, :n[i]].view(np.uint8)) and np.array_equal(desc[i, :n[i]], desc0[i, :n[i]])
        for p in range(P):
            nl_ = n[2 * p]
            assert np.array_equal(uR[p, :nl_], uR0[p, :nl_]) and np.array_equal(dep[p, :nl_], dep0[p, :nl_]), (chunk, ns, p)


@pytest.mark.parametrize("name", ["small", "ragged_with_empty", "large", "ties"])
def test_distinctive_descriptors_equals_oracle(name):
    """MapPoint::ComputeDistinctiveDescriptors batched over map points: selected row and its median bit-exact"""
    import distinct_cases as dc
    desc, offs = dc.case(name)
    gi, gm = orbb200.ORBmatcher().distinctive_descriptors(desc, offs)
    ri, rm = ref_cpu.distinctive_descriptors(desc, offs)
    assert np.array_equal(gi, ri), np.flatnonzero(gi != ri)[:8]
    assert np.array_equal(gm, rm)


def test_distinctive_descriptors_keyframe_sized_batch():
    """a keyframe's worth of map points (2000 groups of 2..40 observations): equals the oracle; the selected row's median
    is the minimum over its group (size-independent property)"""
    import distinct_cases as dc
    rng = np.random.default_rng(77)
    sizes = rng.integers(2, 41, 2000).tolist()
    desc, offs = dc.groups(78, sizes)
    gi, gm = orbb200.ORBmatcher().distinctive_descriptors(desc, offs)
    ri, rm = ref_cpu.distinctive_descriptors(desc, offs)
    assert np.array_equal(gi, ri) and np.array_equal(gm, rm)
    assert (gi >= 0).all() and (gi < np.diff(offs)).all()


def test_bruteforce_full_config5_size_properties():
    """BASELINE config 5 at its full size (2048 queries x 1 048 576 rows): size-independent properties -- every planted
    near-duplicate (<= 60 flipped bits; random rows sit near 128 +- 8) is the best match at its planted position, best <=
    second, the flat search equals the fold of 8 row shards searched separately (the N-GPU structure), and a sample of
    queries equals the CPU oracle over the whole database."""
    nq, ndb = 2048, 1 << 20
    q, db, pos = synth.descriptor_db(nq, ndb, seed=7)
    m = orbb200.ORBmatcher()
    t = m.bruteforce_top2(q, db)
    assert np.array_equal(t["best_idx"], pos.astype(np.int32))
    assert (t["best_dist"] <= 60).all() and (t["best_dist"] <= t["second_dist"]).all() and (t["second_dist"] > 60).all()
    assert (t["second_idx"] != t["best_idx"]).all() and (t["second_idx"] >= 0).all()
    d_best = np.unpackbits(q ^ db[t["best_idx"]], axis=1).sum(axis=1)
    d_second = np.unpackbits(q ^ db[t["second_idx"]], axis=1).sum(axis=1)
    assert np.array_equal(d_best, t["best_dist"]) and np.array_equal(d_second, t["second_dist"])
    shards = 8
    per = ndb // shards
    parts = np.zeros((shards, nq), orbb200.TOP2_DTYPE)
    for s in range(shards):
        p = m.bruteforce_top2(q, db[s * per:(s + 1) * per])
        for f in ("best_idx", "second_idx"):
            p[f] = np.where(p[f] >= 0, p[f] + s * per, -1)
        parts[s] = p
    _eq(m.merge_top2(parts), t)
    sample = np.arange(0, nq, 128)
    _eq(t[sample], ref_cpu.bf_top2(q[sample], db, use_popcnt=True))


def test_window_search_device_chained_behind_extraction(configs):
    """the windowed search on the device-resident outputs of orbb200_extract_batch_device (keypoint count
    read on the device, no host round trip) == orbb200_window_search on the downloaded outputs; a too small cand_cap only
    drops the tail and still reports the total."""
    import ctypes as C
    import torch
    w, h, nf, sfac, nl, ini, mn = configs["tum"]
    imgs = np.stack([synth.textured_frame(w, h, 900 + i) for i in range(2)])
    ex = orbb200.ORBextractor(nf, sfac, nl, ini, mn)
    cap = ex.cap
    dev = torch.device("cuda", 0)
    d_img = torch.from_numpy(imgs).to(dev)
    d_kps = torch.zeros((2, cap, 7), dtype=torch.float32, device=dev)
    d_desc = torch.zeros((2, cap, 32), dtype=torch.uint8, device=dev)
    d_n = torch.zeros(2, dtype=torch.int32, device=dev)
    ex.extract_batch_device(d_img.data_ptr(), 2, w, h, w, w * h, d_kps.data_ptr(), d_desc.data_ptr(), cap, d_n.data_ptr())
    st = ex.stream()
    # windows around the keypoints of image 0, searched in image 1 (frame-to-frame tracking shape)
    ex.sync()
    n = d_n.cpu().numpy()
    k0 = d_kps[0, :n[0]].cpu().numpy().view(orbb200.KP_DTYPE).reshape(-1)
    k1 = d_kps[1, :n[1]].cpu().numpy().view(orbb200.KP_DTYPE).reshape(-1)
    q0 = d_desc[0, :n[0]].cpu().numpy()
    t1 = d_desc[1, :n[1]].cpu().numpy()
    nq = int(n[0])
    win = np.zeros((nq, 5), np.float32)
    win[:, 0], win[:, 1] = k0["x"], k0["y"]
    win[:, 2] = 15.0 * np.float32(1.2) ** k0["octave"]
    win[:, 3] = k0["octave"] - 1
    win[:, 4] = k0["octave"] + 1
    cols, rows = 64, 48
    invW, invH = float(np.float32(cols) / np.float32(w)), float(np.float32(rows) / np.float32(h))
    offs, idx, dist = orbb200.ORBmatcher().window_search(k1, t1, (0.0, 0.0, invW, invH, cols, rows), win, q0, cap=100)
    total = int(offs[-1])
    assert total > 1000
    L = orbb200.lib()
    d_win = torch.from_numpy(win).to(dev)
    d_ws = torch.zeros(cols * rows + 1 + cap, dtype=torch.int32, device=dev)
    vp = C.c_void_p
    for ccap in (total + 7, total // 2):
        d_offs = torch.full((nq + 1,), -1, dtype=torch.int32, device=dev)
        d_idx = torch.full((max(ccap, 1),), -1, dtype=torch.int32, device=dev)
        d_dist = torch.zeros(max(ccap, 1), dtype=torch.int16, device=dev)
        torch.cuda.synchronize()
        orbb200.check(L.orbb200_window_search_device(vp(d_kps[1].data_ptr()), vp(d_desc[1].data_ptr()), cap, vp(d_n[1:].data_ptr()),
                                                     0.0, 0.0, invW, invH, cols, rows, vp(d_win.data_ptr()), vp(d_desc[0].data_ptr()), nq,
                                                     vp(d_offs.data_ptr()), vp(d_idx.data_ptr()), vp(d_dist.data_ptr()), ccap,
                                                     vp(d_ws.data_ptr()), vp(st)))
        ex.sync()
        assert np.array_equal(d_offs.cpu().numpy(), offs)
        m = min(ccap, total)
        assert np.array_equal(d_idx.cpu().numpy()[:m], idx[:m])
        assert np.array_equal(d_dist.cpu().numpy().view(np.uint16)[:m], dist[:m])
        if ccap > total:
            assert (d_idx.cpu().numpy()[total:] == -1).all()
