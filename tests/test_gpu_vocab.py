"""GPU parity tests (through the C-ABI) of the DBoW2 vocabulary transform (Frame::ComputeBoW, src/Frame.cc:459-466):
word ids, node ids and feature lists bit-exact, BowVector values bit-exact doubles, against the reference's own DBoW2
outputs (tests/golden/vocab_*, tools/make_vocab_golden.py), the live reference library where it was shipped
(oracle/_ref/libdbow2_ref.so) and the oracle restatement (oracle/vocab_oracle.py) on large synthetic trees."""
import numpy as np
import pytest

import orbb200
import vocab_cases as vc
import vocab_oracle as vo

pytestmark = pytest.mark.gpu


def _from_oracle(voc):
    return orbb200.ORBVocabulary.from_arrays(voc.k, voc.L, voc.scoring, voc.weighting, voc.parent, voc.is_leaf, voc.desc, voc.weight)


@pytest.mark.parametrize("name", ["a", "b"])
def test_transform_equals_reference_golden(name, tmp_path):
    g = vc.gold()
    q = g["%s_queries" % name]
    v = orbb200.ORBVocabulary()
    assert v.loadFromTextFile(vc.vocab_path(name))
    info = v.info()
    assert [info["k"], info["L"], info["nodes"], info["words"]] == list(g["%s_info" % name])
    voc = vo.load_text(vc.vocab_path(name))
    for ls in vc.LEVELSUP:
        tag = "%s_w0_s0_l%d" % (name, ls)
        word, weight, nid = v.transform_features(q, ls)
        assert np.array_equal(word, g[tag + "_word"]) and np.array_equal(weight, g[tag + "_weight"])
        ok = g[tag + "_nid"] >= 0
        assert np.array_equal(nid[ok], g[tag + "_nid"][ok])
        assert np.array_equal(nid, vo.transform_features(voc, q, ls)[2])      # incl. the words above the nid level
    for (wt, sc) in vc.VARIANTS:
        v2 = orbb200.ORBVocabulary()
        path = vc.variant_file(name, wt, sc, tmp_path)
        assert v2.loadFromTextFile(path)
        o2 = vo.load_text(path)
        for ls in vc.LEVELSUP:
            res = v2.compute_bow_batch(q, [hi - lo for lo, hi in vc.PARTS], ls)
            pf = vo.transform_features(o2, q, ls)
            for part, (lo, hi) in enumerate(vc.PARTS):
                bow, fv = res[part]
                gb, gf = vc.gold_maps("%s_w%d_s%d_l%d" % (name, wt, sc, ls), part)
                assert list(bow.items()) == list(gb.items()), (wt, sc, ls, part)
                if vc.nid_defined(name, ls, lo, hi):
                    assert list(fv.items()) == list(gf.items()), (wt, sc, ls, part)
                ob, of = vo.transform(o2, q[lo:hi], ls, per_feature=tuple(a[lo:hi] for a in pf))
                assert list(fv.items()) == list(of.items())


@pytest.mark.parametrize("k,L,p_stop,jit,nq", [(10, 4, 0.0, False, 6000), (10, 5, 0.15, False, 6000), (16, 3, 0.0, True, 6000),
                                               (20, 3, 0.1, True, 6000), (2, 10, 0.05, False, 6000), (3, 1, 0.0, False, 6000),
                                               (10, 5, 0.15, False, 30000), (20, 3, 0.1, True, 30000), (2, 10, 0.05, False, 30000)])
def test_transform_equals_oracle_on_synthetic_trees(k, L, p_stop, jit, nq):
    """nq + 2000 features: below 16384 the lane-group mapping runs, above it the thread-per-feature mapping"""
    rng = np.random.default_rng(k * 100 + L)
    voc = vc.random_tree(rng, k, L, p_stop=p_stop, max_nodes=200000, k_jitter=jit)
    v = _from_oracle(voc)
    assert v.info()["nodes"] == voc.nnodes and v.size() == voc.nwords
    q = np.concatenate([vc.queries_near_nodes(rng, voc, nq), rng.integers(0, 256, (2000, 32), dtype=np.uint8)])
    for ls in (0, 2, L, L + 3):
        word, weight, nid = v.transform_features(q, ls)
        ow, owt, onid = vo.transform_features(voc, q, ls)
        assert np.array_equal(word, ow) and np.array_equal(weight, owt) and np.array_equal(nid, onid), (k, L, ls)


def test_compute_bow_batch_ragged_and_empty_images():
    rng = np.random.default_rng(9)
    voc = vc.random_tree(rng, 10, 4, p_stop=0.1)
    v = _from_oracle(voc)
    counts = [0, 1, 700, 0, 33, 1500, 2, 0]
    q = vc.queries_near_nodes(rng, voc, sum(counts), flip=0.05)
    res = v.compute_bow_batch(q, counts, 4)
    pf = vo.transform_features(voc, q, 4)
    o = 0
    for i, c in enumerate(counts):
        ob, of = vo.transform(voc, q[o:o + c], 4, per_feature=tuple(a[o:o + c] for a in pf))
        assert list(res[i][0].items()) == list(ob.items()) and list(res[i][1].items()) == list(of.items()), i
        if c:
            assert abs(sum(res[i][0].values()) - 1.0) < 1e-12 or not res[i][0]      # L1-normalised
            assert sorted(x for f in res[i][1].values() for x in f) == [j for j in range(c) if pf[1][o + j] > 0]
        o += c
    # one call per image gives the same maps as the batch
    assert v.transform(q[1:701], 4) == res[2]
    assert v.compute_bow_batch(np.zeros((0, 32), np.uint8), [], 4) == []
    w, wt, nid = v.transform_features(np.zeros((0, 32), np.uint8), 2)
    assert len(w) == 0


def test_full_size_tree_properties():
    """ORBvoc.txt's shape (k = 10, L = 6: 1.11 M nodes, 10^6 words) at a KITTI batch's feature count: words are in range,
    every feature's node at level L - levelsup is the ancestor of its word, and the walk is deterministic."""
    rng = np.random.default_rng(2)
    voc = vc.random_tree(rng, 10, 6)
    assert voc.nnodes == 1111111 and voc.nwords == 1000000
    v = _from_oracle(voc)
    n = 200000
    q = vc.queries_near_nodes(rng, voc, n, flip=0.12)
    word, weight, nid = v.transform_features(q, 4)
    word_node = np.flatnonzero(voc.is_leaf > 0)
    leaf = word_node[word]
    assert word.min() >= 0 and word.max() < voc.nwords and np.array_equal(weight, voc.weight[leaf])
    anc = leaf.copy()
    for _ in range(4):
        anc = voc.parent[anc]
    assert np.array_equal(anc, nid)
    w2, _, n0 = v.transform_features(q, 0)
    assert np.array_equal(w2, word) and np.array_equal(n0, leaf)
    sub = slice(0, 3000)
    ow, owt, onid = vo.transform_features(voc, q[sub], 4)
    assert np.array_equal(word[sub], ow) and np.array_equal(nid[sub], onid)


def test_loader_errors_and_empty_vocabulary(tmp_path):
    v = orbb200.ORBVocabulary()
    assert not v.loadFromTextFile(str(tmp_path / "missing.txt"))
    bad = tmp_path / "bad.txt"
    bad.write_text("30 4 0 0\n0 1 " + " ".join(["1"] * 32) + " 0.5\n")      # k > 20 (TemplatedVocabulary.h:1385)
    assert not v.loadFromTextFile(str(bad))
    short = tmp_path / "short.txt"
    short.write_text("3 2 0 0\n0 1 1 2 3 0.5\n0 1 " + " ".join(["1"] * 32) + " 0.5\n")
    assert not v.loadFromTextFile(str(short))
    only_header = tmp_path / "hdr.txt"
    only_header.write_text("10 6 0 0\n")
    assert v.loadFromTextFile(str(only_header)) and v.empty()
    assert v.compute_bow_batch(np.zeros((5, 32), np.uint8), [5], 4) == [({}, {})]      # if(empty()) return;
    with pytest.raises(orbb200.OrbB200Error):
        v.transform_features(np.zeros((5, 32), np.uint8), 0)
    # the final newline does not add a node
    a = orbb200.ORBVocabulary()
    txt = open(vc.vocab_path("a")).read()
    p = tmp_path / "nl.txt"
    p.write_text(txt + "\n")
    assert a.loadFromTextFile(str(p)) and a.info()["nodes"] == int(vc.gold()["a_info"][2])


def test_transform_device_on_extracted_descriptors():
    """the device entry chained behind orbb200_extract_batch_device: descriptors never leave HBM"""
    import torch
    import synth
    rng = np.random.default_rng(4)
    voc = vc.random_tree(rng, 10, 4)
    v = _from_oracle(voc)
    ex = orbb200.ORBextractor(1000, 1.2, 8, 20, 7)
    imgs = np.stack([synth.textured_frame(640, 480, 50 + i) for i in range(3)])
    cap = ex.cap
    kps_h, desc_h, n_h = ex.extract_batch(imgs)
    d_img = torch.from_numpy(imgs).cuda()
    d_kps = torch.zeros(3 * cap * 28, dtype=torch.uint8, device="cuda")
    d_desc = torch.zeros((3, cap, 32), dtype=torch.uint8, device="cuda")
    d_n = torch.zeros(3, dtype=torch.int32, device="cuda")
    ex.extract_batch_device(d_img.data_ptr(), 3, 640, 480, 640, 640 * 480, d_kps.data_ptr(), d_desc.data_ptr(), cap, d_n.data_ptr())
    d_word = torch.zeros(3 * cap, dtype=torch.int32, device="cuda")
    d_nid = torch.zeros(3 * cap, dtype=torch.int32, device="cuda")
    d_w = torch.zeros(3 * cap, dtype=torch.float64, device="cuda")
    v.transform_features_device(d_desc.data_ptr(), 3 * cap, 2, d_word.data_ptr(), d_w.data_ptr(), d_nid.data_ptr(), ex.stream())
    ex.sync()
    n = d_n.cpu().numpy()
    word, nid, w = d_word.cpu().numpy().reshape(3, cap), d_nid.cpu().numpy().reshape(3, cap), d_w.cpu().numpy().reshape(3, cap)
    for i in range(3):
        assert n[i] == n_h[i] and n[i] > 500
        ow, owt, onid = vo.transform_features(voc, desc_h[i][:n[i]], 2)
        assert np.array_equal(word[i, :n[i]], ow) and np.array_equal(nid[i, :n[i]], onid) and np.array_equal(w[i, :n[i]], owt)


def test_frame_front_end_resident_in_hbm():
    """extraction -> ComputeBoW chained on one stream with device pointers only (orbb200_extract_batch_device ->
    orbb200_compute_bow_batch_device): same maps as the host call on the downloaded descriptors"""
    import torch
    import synth
    rng = np.random.default_rng(8)
    voc = vc.random_tree(rng, 10, 4, p_stop=0.05)
    v = _from_oracle(voc)
    ex = orbb200.ORBextractor(1000, 1.2, 8, 20, 7)
    B = 5
    imgs = np.stack([synth.textured_frame(640, 480, 80 + i) for i in range(B)])
    imgs[3] = 0                                                # a frame without keypoints
    cap = ex.cap
    _, desc_h, n_h = ex.extract_batch(imgs)
    d_img = torch.from_numpy(imgs).cuda()
    d_kps = torch.zeros(B * cap * 28, dtype=torch.uint8, device="cuda")
    d_desc = torch.zeros((B, cap, 32), dtype=torch.uint8, device="cuda")
    d_n = torch.zeros(B, dtype=torch.int32, device="cuda")
    d_bid = torch.zeros((B, cap), dtype=torch.int32, device="cuda")
    d_bval = torch.zeros((B, cap), dtype=torch.float64, device="cuda")
    d_fkey = torch.zeros((B, cap), dtype=torch.int64, device="cuda")
    d_nb = torch.zeros(B, dtype=torch.int32, device="cuda")
    d_nk = torch.zeros(B, dtype=torch.int32, device="cuda")
    ex.extract_batch_device(d_img.data_ptr(), B, 640, 480, 640, 640 * 480, d_kps.data_ptr(), d_desc.data_ptr(), cap, d_n.data_ptr())
    v.compute_bow_batch_device(d_desc.data_ptr(), d_n.data_ptr(), B, cap, 4, d_bid.data_ptr(), d_bval.data_ptr(), d_nb.data_ptr(),
                               d_fkey.data_ptr(), d_nk.data_ptr(), ex.stream())
    ex.sync()
    n = d_n.cpu().numpy()
    assert np.array_equal(n, n_h) and n[3] == 0
    bid, bval, fkey = d_bid.cpu().numpy(), d_bval.cpu().numpy(), d_fkey.cpu().numpy().view(np.uint64)
    nb, nk = d_nb.cpu().numpy(), d_nk.cpu().numpy()
    host = v.compute_bow_batch(np.concatenate([desc_h[i][:n[i]] for i in range(B)]), n, 4)
    for i in range(B):
        bow, fv = host[i]
        assert list(bow.keys()) == [int(x) for x in bid[i, :nb[i]]] and list(bow.values()) == [float(x) for x in bval[i, :nb[i]]]
        keys = fkey[i, :nk[i]]
        got = {}
        for kx in keys:
            got.setdefault(int(kx >> np.uint64(32)), []).append(int(kx & np.uint64(0xffffffff)))
        assert list(got.items()) == list(fv.items())
        ob, of = vo.transform(voc, desc_h[i][:n[i]], 4, per_feature=vo.transform_features(voc, desc_h[i][:n[i]], 4)) if n[i] else ({}, {})
        assert list(bow.items()) == list(ob.items()) and list(fv.items()) == list(of.items())
