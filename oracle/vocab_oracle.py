"""CPU restatement of the DBoW2 vocabulary path ORB-SLAM2 uses (Frame::ComputeBoW, src/Frame.cc:459-466) -- TEST INFRASTRUCTURE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this module; the
product path (orb-slam2_b200/, include/) never does.

Parity: PINNED.  tests/test_oracle_pin.py checks every function here against the reference's own DBoW2 code compiled
from /root/reference/Thirdparty/DBoW2 where it lies (oracle/Makefile -> oracle/_ref/libdbow2_ref.so, wrapper
oracle/dbow2_ref.cpp, with oracle/shim standing in for the absent OpenCV C++ headers: cv::Mat is only a 1x32 byte
container on this path), on vocabularies TRAINED by the reference's create() and round-tripped through its
saveToTextFile / loadFromTextFile.  The reference ships no tests or golden vectors for DBoW2 and ORBvoc.txt is not in the
checkout (Vocabulary/ORBvoc.txt.tar.gz is fetched by build.sh), so the pin is reference-code-run-here, not fixtures
from upstream; tests/golden/vocab_* are outputs of that reference run (tools/make_vocab_golden.py).

Restated (file:line under Thirdparty/DBoW2/DBoW2/):
  load_text            TemplatedVocabulary.h:1358-1443  loadFromTextFile
  transform_feature    TemplatedVocabulary.h:1218-1259  transform(feature, word_id, weight, nid, levelsup)
  transform            TemplatedVocabulary.h:1127-1190  transform(features, BowVector, FeatureVector, levelsup)
                       BowVector.cpp:34-84 (addWeight, addIfNotExist, normalize), FeatureVector.cpp:31-45 (addFeature)
  score                ScoringObject.cpp:23-69 (L1), :74-124 (L2) ... via TemplatedVocabulary::score (:1194-1198)
  distance             FORB.cpp:81-99 (bit-count of the xor of two 256-bit strings)

Two places where the reference's behaviour is undefined and this restatement (and the CUDA path) picks a definition:
  * loadFromTextFile reads one more "node" from the empty string after the file's final newline (while(!f.eof()) +
    getline): a phantom extra child of the root whose descriptor is uninitialised memory.  Here blank lines are skipped.
    The pin tests strip the final newline before handing a file to the reference so that both sides hold the same tree.
  * when a word sits ABOVE level L - levelsup (a short branch; HKmeans stops splitting small clusters) the reference
    never writes *nid and ComputeBoW files the feature under an indeterminate node id.  Here such a feature is filed
    under its word's own node id (the deepest node on its path).  The pin on nid therefore covers features whose path
    reaches that level, and the wrapper presets nid = -1 so the others are identifiable.
"""
import numpy as np

TF_IDF, TF, IDF, BINARY = 0, 1, 2, 3
L1_NORM, L2_NORM, CHI_SQUARE, KL, BHATTACHARYYA, DOT_PRODUCT = 0, 1, 2, 3, 4, 5

_POP8 = np.array([bin(i).count("1") for i in range(256)], dtype=np.int32)


def distance(a, b):
    """FORB::distance (FORB.cpp:81-99)."""
    return int(_POP8[np.bitwise_xor(a, b)].sum())


class Vocabulary:
    """The tree as loadFromTextFile leaves it: node 0 is the root, children in file order, words numbered in file order."""

    def __init__(self, k, L, scoring, weighting, parent, is_leaf, desc, weight):
        self.k, self.L, self.scoring, self.weighting = int(k), int(L), int(scoring), int(weighting)
        self.parent = np.asarray(parent, dtype=np.int32)       # [nnodes], parent[0] unused
        self.is_leaf = np.asarray(is_leaf, dtype=np.uint8)     # [nnodes] the file's flag (a node WITHOUT children ends the walk)
        self.desc = np.ascontiguousarray(desc, dtype=np.uint8).reshape(-1, 32)
        self.weight = np.asarray(weight, dtype=np.float64)
        n = len(self.parent)
        self.children = [[] for _ in range(n)]
        for nid in range(1, n):                                # m_nodes[pid].children.push_back(nid), :1411
            self.children[self.parent[nid]].append(nid)
        self.word_id = np.full(n, -1, dtype=np.int32)
        wid = 0
        for nid in range(1, n):                                # :1430-1437
            if self.is_leaf[nid] > 0:
                self.word_id[nid] = wid
                wid += 1
        self.nwords = wid

    @property
    def nnodes(self):
        return len(self.parent)

    def empty(self):
        return self.nwords == 0


def load_text(path):
    """loadFromTextFile (:1358-1443).  First line "k L scoring weighting"; then one node per line:
    "parent isLeaf b0 ... b31 weight" with node ids 1, 2, ... in line order."""
    with open(path, "r") as f:
        head = f.readline().split()
        k, L, n1, n2 = int(head[0]), int(head[1]), int(head[2]), int(head[3])
        if k < 0 or k > 20 or L < 1 or L > 10 or n1 < 0 or n1 > 5 or n2 < 0 or n2 > 3:    # :1385-1389
            raise ValueError("Vocabulary loading failure: This is not a correct text file!")
        parent, leaf, desc, weight = [0], [0], [np.zeros(32, np.uint8)], [0.0]
        for line in f:
            t = line.split()
            if not t:
                continue                                       # see the header: the reference makes a phantom node here
            parent.append(int(t[0]))
            leaf.append(int(t[1]))
            desc.append(np.array([int(x) for x in t[2:34]], dtype=np.int64).astype(np.uint8))   # (unsigned char)n
            weight.append(float(t[34]))
    return Vocabulary(k, L, n1, n2, parent, leaf, np.stack(desc), weight)


def save_text(voc, path, final_newline=True):
    """saveToTextFile's format (:1450-1472): enough digits for the weights to survive the round trip exactly."""
    lines = ["%d %d %d %d" % (voc.k, voc.L, voc.scoring, voc.weighting)]
    for nid in range(1, voc.nnodes):
        lines.append("%d %d %s %s" % (voc.parent[nid], 1 if voc.is_leaf[nid] else 0,
                                      " ".join(str(int(b)) for b in voc.desc[nid]), repr(float(voc.weight[nid]))))
    with open(path, "w") as f:
        f.write("\n".join(lines) + ("\n" if final_newline else ""))


def transform_feature(voc, feature, levelsup=0):
    """transform(feature, word_id, weight, nid, levelsup) (:1218-1259): walk from the root to a word, at every node
    taking the child with the smallest distance, the FIRST one on ties (strict <)."""
    nid_level = voc.L - levelsup
    nid = 0 if nid_level <= 0 else None
    final_id, current_level = 0, 0
    while True:
        current_level += 1
        nodes = voc.children[final_id]
        final_id = nodes[0]
        best_d = distance(feature, voc.desc[final_id])
        for c in nodes[1:]:
            d = distance(feature, voc.desc[c])
            if d < best_d:
                best_d, final_id = d, c
        if current_level == nid_level:
            nid = final_id
        if not voc.children[final_id]:                         # isLeaf() == children.empty()
            break
    if nid is None:
        nid = final_id                                         # indeterminate in the reference; see the header
    return int(voc.word_id[final_id]), float(voc.weight[final_id]), int(nid)


def transform_features(voc, feats, levelsup=0):
    """transform_feature for every row of feats, level-synchronous in numpy (same arithmetic, same tie rule)."""
    feats = np.ascontiguousarray(feats, dtype=np.uint8).reshape(-1, 32)
    n = len(feats)
    maxc = max(len(c) for c in voc.children)
    ctab = np.full((voc.nnodes, maxc), -1, dtype=np.int64)
    for i, c in enumerate(voc.children):
        ctab[i, :len(c)] = c
    ccount = np.array([len(c) for c in voc.children], dtype=np.int64)
    nid_level = voc.L - levelsup
    cur = np.zeros(n, dtype=np.int64)
    nid = np.full(n, -1, dtype=np.int64)
    if nid_level <= 0:
        nid[:] = 0
    active = np.ones(n, dtype=bool)
    level = 0
    while active.any():
        level += 1
        idx = np.nonzero(active)[0]
        ch = ctab[cur[idx]]                                    # [m, maxc]
        valid = ch >= 0
        d = _POP8[np.bitwise_xor(feats[idx][:, None, :], voc.desc[np.where(valid, ch, 0)])].sum(axis=2)
        d = np.where(valid, d, 1 << 20)
        best = np.argmin(d, axis=1)                            # first minimum
        cur[idx] = ch[np.arange(len(idx)), best]
        if level == nid_level:
            nid[idx] = cur[idx]
        active[idx] = ccount[cur[idx]] > 0
    nid = np.where(nid < 0, cur, nid)
    return voc.word_id[cur].astype(np.int32), voc.weight[cur].copy(), nid.astype(np.int32)


def must_normalize(scoring):
    """ScoringObject.h:74-89: (mustNormalize, norm) per scoring class."""
    return {L1_NORM: (True, "L1"), L2_NORM: (True, "L2"), CHI_SQUARE: (True, "L1"), KL: (True, "L1"),
            BHATTACHARYYA: (True, "L1"), DOT_PRODUCT: (False, "L1")}[scoring]


def transform(voc, feats, levelsup=4, per_feature=None):
    """transform(features, v, fv, levelsup) (:1127-1190).  Returns (bow, fv): bow = {word: value} and fv = {node: [feature
    indices]}, both in ascending key order like the std::maps they restate."""
    bow, fv = {}, {}
    if voc.empty():
        return bow, fv
    feats = np.ascontiguousarray(feats, dtype=np.uint8).reshape(-1, 32)
    must, norm = must_normalize(voc.scoring)
    if per_feature is None:
        pf = [transform_feature(voc, f, levelsup) for f in feats]
        per_feature = ([p[0] for p in pf], [p[1] for p in pf], [p[2] for p in pf])
    words, weights, nids = per_feature
    tf = voc.weighting in (TF, TF_IDF)
    for i in range(len(feats)):
        wid, w, nid = int(words[i]), float(weights[i]), int(nids[i])
        if w > 0:                                              # not stopped
            if tf:
                bow[wid] = bow[wid] + w if wid in bow else w   # addWeight, BowVector.cpp:34-46
            elif wid not in bow:
                bow[wid] = w                                   # addIfNotExist, :50-58
            fv.setdefault(nid, []).append(i)                   # addFeature, FeatureVector.cpp:31-45
    bow = dict(sorted(bow.items()))
    fv = dict(sorted(fv.items()))
    if tf and bow and not must:                                # :1163-1169
        nd = float(len(bow))
        for k_ in bow:
            bow[k_] /= nd
    if must:                                                   # BowVector::normalize, :62-84 (sums in key order)
        s = 0.0
        if norm == "L1":
            for v in bow.values():
                s += abs(v)
        else:
            for v in bow.values():
                s += v * v
            s = float(np.sqrt(s))
        if s > 0.0:
            for k_ in bow:
                bow[k_] /= s
    return bow, fv


def score(voc, a, b):
    """TemplatedVocabulary::score (:1194-1198) for the two scorings whose arithmetic is restated here: L1 (ORBvoc.txt's,
    ScoringObject.cpp:23-69) and L2 (:74-124); a, b are {word: value} dicts."""
    common = sorted(set(a) & set(b))
    s = 0.0
    if voc.scoring == L1_NORM:
        for k_ in common:
            s += abs(a[k_] - b[k_]) - abs(a[k_]) - abs(b[k_])
        return -s / 2.0
    if voc.scoring == L2_NORM:
        for k_ in common:
            s += a[k_] * b[k_]
        return 1.0 if s >= 1 else 1.0 - float(np.sqrt(1.0 - s))
    raise NotImplementedError("scoring %d" % voc.scoring)


def synthetic_descriptors(rng, ncenters, n, flip=0.08):
    """ORB-like test descriptors: n noisy copies of ncenters random 256-bit strings (each bit flipped with prob. flip)."""
    centers = rng.integers(0, 256, size=(ncenters, 32), dtype=np.uint8)
    pick = rng.integers(0, ncenters, size=n)
    noise = np.packbits(rng.random((n, 256)) < flip, axis=1)
    return np.bitwise_xor(centers[pick], noise)
