"""Measures the DBoW2 vocabulary transform (Frame::ComputeBoW) on an ORBvoc.txt-shaped tree (k = 10, L = 6, 10^6 words):
the tree-walk kernel alone (device-resident descriptors, CUDA events), the batched host call
(orbb200_compute_bow_batch: H2D + walk + D2H + map epilogue) and the reference's own DBoW2 on the host cores
(oracle/_ref/libdbow2_ref.so, one thread like Frame::ComputeBoW).  Prints one JSON line."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "orb-slam2_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch  # noqa: E402
import orbb200  # noqa: E402
import vocab_cases as vc  # noqa: E402
import vocab_oracle as vo  # noqa: E402


def main():
    nimg = int(os.environ.get("BOW_IMAGES", "320"))
    per = int(os.environ.get("BOW_FEATURES", "2000"))
    reps = int(os.environ.get("BOW_REPS", "20"))
    rng = np.random.default_rng(2)
    voc = vc.random_tree(rng, 10, 6)
    v = orbb200.ORBVocabulary.from_arrays(voc.k, voc.L, 0, 0, voc.parent, voc.is_leaf, voc.desc, voc.weight)
    n = nimg * per
    q = vc.queries_near_nodes(rng, voc, n, flip=0.12)
    d_q = torch.from_numpy(q).cuda()
    d_word = torch.zeros(n, dtype=torch.int32, device="cuda")
    d_nid = torch.zeros(n, dtype=torch.int32, device="cuda")
    d_w = torch.zeros(n, dtype=torch.float64, device="cuda")
    ts = torch.cuda.Stream()                                   # a real stream: 0 would select the vocabulary's own
    torch.cuda.set_stream(ts)
    st = ts.cuda_stream
    flush = torch.zeros(256 << 20, dtype=torch.uint8, device="cuda")
    times = {"cold": [], "warm": []}
    for mode in ("cold", "warm"):
        for it in range(reps + 3):
            if mode == "cold":
                flush.add_(1)                                  # evict the tree from L2
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            v.transform_features_device(d_q.data_ptr(), n, 4, d_word.data_ptr(), d_w.data_ptr(), d_nid.data_ptr(), st)
            e1.record()
            torch.cuda.synchronize()
            if it >= 3:
                times[mode].append(e0.elapsed_time(e1))
    # walk + fold with everything resident in HBM (orbb200_compute_bow_batch_device; [nimg][cap] layout of the extractor)
    d_n = torch.full((nimg,), per, dtype=torch.int32, device="cuda")
    d_bid = torch.zeros(n, dtype=torch.int32, device="cuda")
    d_bval = torch.zeros(n, dtype=torch.float64, device="cuda")
    d_fkey = torch.zeros(n, dtype=torch.int64, device="cuda")
    d_nb = torch.zeros(nimg, dtype=torch.int32, device="cuda")
    d_nk = torch.zeros(nimg, dtype=torch.int32, device="cuda")
    dev_ms = []
    for it in range(reps + 3):
        flush.add_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        v.compute_bow_batch_device(d_q.data_ptr(), d_n.data_ptr(), nimg, per, 4, d_bid.data_ptr(), d_bval.data_ptr(), d_nb.data_ptr(),
                                   d_fkey.data_ptr(), d_nk.data_ptr(), st)
        e1.record()
        torch.cuda.synchronize()
        if it >= 3:
            dev_ms.append(e0.elapsed_time(e1))
    counts = [per] * nimg
    v.compute_bow_batch_raw(q, counts, 4)
    t0 = time.perf_counter()
    for _ in range(3):
        raw = v.compute_bow_batch_raw(q, counts, 4)
    host_ms = (time.perf_counter() - t0) / 3 * 1e3
    out = {"probe": "vocabulary transform", "tree": "k=10 L=6 synthetic (1111111 nodes, 1000000 words)", "images": nimg,
           "features_per_image": per, "levelsup": 4,
           "kernel_ms_tree_in_L2": float(np.median(times["warm"])), "kernel_ms_L2_flushed": float(np.median(times["cold"])),
           "kernel_features_per_s": n / (float(np.median(times["warm"])) * 1e-3),
           "algorithmic_bytes_per_feature": 32 + 6 * 10 * 32 + 6 * 8 + 16,
           "compute_bow_device_ms_walk_plus_fold": float(np.median(dev_ms)),
           "compute_bow_device_images_per_s": nimg / (float(np.median(dev_ms)) * 1e-3),
           "compute_bow_batch_ms": host_ms, "compute_bow_batch_images_per_s": nimg / (host_ms * 1e-3)}
    out["kernel_GBps_algorithmic"] = out["algorithmic_bytes_per_feature"] * n / (out["kernel_ms_tree_in_L2"] * 1e-3) / 1e9
    # parity spot check against the restatement
    ow, owt, onid = vo.transform_features(voc, q[:4000], 4)
    out["parity_first_4000"] = bool(np.array_equal(d_word[:4000].cpu().numpy(), ow) and np.array_equal(d_nid[:4000].cpu().numpy(), onid))
    import dbow2_ref
    if dbow2_ref.available():
        path = "/tmp/bow_probe_vocab.txt"
        vo.save_text(voc, path, final_newline=False)
        rv = dbow2_ref.RefVocabulary()
        t0 = time.perf_counter()
        assert rv.load_text(path)
        out["reference_load_text_s"] = time.perf_counter() - t0
        m = min(nimg, 16)
        t0 = time.perf_counter()
        for i in range(m):
            bow, fv = rv.transform(q[i * per:(i + 1) * per], 4)
        ref_ms = (time.perf_counter() - t0) / m * 1e3
        out["reference_cpu_ms_per_image_1_thread"] = ref_ms
        out["reference_cpu_images_per_s_1_thread"] = 1e3 / ref_ms
        b0, b1 = raw["bow_off"][m - 1], raw["bow_off"][m]
        out["parity_bow_vs_reference"] = bool(list(bow.keys()) == [int(x) for x in raw["bow_id"][b0:b1]] and
                                              list(bow.values()) == [float(x) for x in raw["bow_val"][b0:b1]])
        t0 = time.perf_counter()
        v2 = orbb200.ORBVocabulary()
        assert v2.loadFromTextFile(path)
        out["load_text_s"] = time.perf_counter() - t0
        os.remove(path)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
