"""Writes tests/golden/vocab_*.txt and tests/golden/vocab_golden.npz with the REFERENCE's own DBoW2 code
(oracle/_ref/libdbow2_ref.so, built by oracle/Makefile from /root/reference/Thirdparty/DBoW2): two small vocabularies
trained by TemplatedVocabulary::create() on synthetic ORB-like descriptors, saved with saveToTextFile, and the
reference's transform() outputs on fixed query sets.  Run in the container (needs /root/reference to have built the .so):
    python tools/make_vocab_golden.py
The final newline of each saved file is stripped (see oracle/vocab_oracle.py: the reference's loader would add a phantom
node for it) and the reference RELOADS the stripped file before producing the outputs, so the 6-digit weights in the
file are the weights the golden outputs were computed with."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import dbow2_ref  # noqa: E402
import vocab_oracle as vo  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
# name: (k, L, training images, descriptors per image, centres, flip, seed)
VOCABS = {"a": (6, 4, 24, 400, 3000, 0.15, 7),      # every branch reaches depth L
          "b": (9, 5, 12, 150, 900, 0.18, 11)}      # few descriptors: short branches, words above the nid level
LEVELSUP = (0, 1, 2, 4, 7)
VARIANTS = ((0, 0), (1, 1), (2, 0), (3, 5), (0, 5))  # (weighting, scoring) written into the file's header line


def flat(bow, fv):
    ids = np.array(list(bow.keys()), np.int32)
    vals = np.array(list(bow.values()), np.float64)
    nodes = np.array(list(fv.keys()), np.int32)
    off = np.cumsum([0] + [len(v) for v in fv.values()]).astype(np.int32)
    feat = np.array([x for v in fv.values() for x in v], np.int32)
    return ids, vals, nodes, off, feat


def main():
    out = {}
    for name, (k, L, nimg, per, ncent, flip, seed) in VOCABS.items():
        rng = np.random.default_rng(seed)
        train = vo.synthetic_descriptors(rng, ncent, nimg * per, flip)
        rv = dbow2_ref.RefVocabulary(k, L, 0, 0)
        rv.create(train, [per] * nimg, seed=seed)
        path = os.path.join(GOLD, "vocab_%s.txt" % name)
        rv.save_text(path)
        txt = open(path).read().rstrip("\n")
        open(path, "w").write(txt)
        q = vo.synthetic_descriptors(rng, ncent, 600, flip + 0.05)
        q[:50] = train[:50]                                   # some exact training descriptors
        out["%s_queries" % name] = q
        lines = txt.split("\n")
        for (wt, sc) in VARIANTS:
            tmp = os.path.join(GOLD, "_tmp_vocab.txt")
            open(tmp, "w").write("\n".join(["%d %d %d %d" % (k, L, sc, wt)] + lines[1:]))
            r2 = dbow2_ref.RefVocabulary()
            assert r2.load_text(tmp)
            os.remove(tmp)
            info = r2.info()
            if (wt, sc) == (0, 0):
                r_l1 = r2
            out["%s_info" % name] = np.array([info["k"], info["L"], info["nodes"], info["words"]], np.int32)
            for ls in LEVELSUP:
                tag = "%s_w%d_s%d_l%d" % (name, wt, sc, ls)
                if (wt, sc) == (0, 0):
                    w, wgt, nid = r2.transform_features(q, ls)
                    out[tag + "_word"], out[tag + "_weight"], out[tag + "_nid"] = w, wgt, nid
                for part, (lo, hi) in enumerate(((0, 300), (300, 600))):
                    ids, vals, nodes, off, feat = flat(*r2.transform(q[lo:hi], ls))
                    for nm, arr in zip(("bid", "bval", "fnode", "foff", "ffeat"), (ids, vals, nodes, off, feat)):
                        out["%s_p%d_%s" % (tag, part, nm)] = arr
            if (wt, sc) in ((0, 0), (1, 1)):
                a = r2.transform(q[:300], 4)[0]
                b = r2.transform(q[300:], 4)[0]
                out["%s_w%d_s%d_score" % (name, wt, sc)] = np.array([r2.score(a, b), r2.score(a, a)], np.float64)
        # the six scoring classes on the same two TF-IDF / L1 vectors (score(a, b) and score(b, a))
        a, b = r_l1.transform(q[:300], 4)[0], r_l1.transform(q[300:], 4)[0]
        sc6 = []
        for sc in range(6):
            tmp = os.path.join(GOLD, "_tmp_vocab.txt")
            open(tmp, "w").write("\n".join(["%d %d %d %d" % (k, L, sc, 0)] + lines[1:]))
            r2 = dbow2_ref.RefVocabulary()
            assert r2.load_text(tmp)
            os.remove(tmp)
            sc6.append([r2.score(a, b), r2.score(b, a)])
        out["%s_scores6" % name] = np.array(sc6, np.float64)
        print(name, info, "undefined nid at levelsup 0:", int((out["%s_w0_s0_l0_nid" % name] == -1).sum()))
    np.savez_compressed(os.path.join(GOLD, "vocab_golden.npz"), **out)
    print("wrote", len(out), "arrays")


if __name__ == "__main__":
    main()
