"""Shared by the CPU pin tests and the GPU parity tests of the DBoW2 vocabulary transform: the committed golden
fixtures (tests/golden/vocab_*.txt + vocab_golden.npz, outputs of the reference's own DBoW2 code, see
tools/make_vocab_golden.py) and a generator of large synthetic trees."""
import os

import numpy as np

import vocab_oracle as vo

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
LEVELSUP = (0, 1, 2, 4, 7)
VARIANTS = ((0, 0), (1, 1), (2, 0), (3, 5), (0, 5))   # (weighting, scoring)
PARTS = ((0, 300), (300, 600))
_gold = None


def gold():
    global _gold
    if _gold is None:
        _gold = dict(np.load(os.path.join(GOLD, "vocab_golden.npz")))
    return _gold


def vocab_path(name):
    return os.path.join(GOLD, "vocab_%s.txt" % name)


def variant_file(name, weighting, scoring, tmpdir):
    """the golden tree with another (scoring, weighting) in its header line, as the generator handed it to the reference"""
    lines = open(vocab_path(name)).read().split("\n")
    k, L = lines[0].split()[:2]
    p = os.path.join(str(tmpdir), "vocab_%s_w%d_s%d.txt" % (name, weighting, scoring))
    open(p, "w").write("\n".join(["%s %s %d %d" % (k, L, scoring, weighting)] + lines[1:]))
    return p


def gold_maps(tag, part):
    g = gold()
    pre = "%s_p%d_" % (tag, part)
    bow = dict(zip((int(x) for x in g[pre + "bid"]), (float(x) for x in g[pre + "bval"])))
    off = g[pre + "foff"]
    fv = {int(n): [int(x) for x in g[pre + "ffeat"][off[j]:off[j + 1]]] for j, n in enumerate(g[pre + "fnode"])}
    return bow, fv


def nid_defined(name, ls, lo, hi):
    """False where the reference left *nid unwritten for some feature of the range (its FeatureVector is then garbage)"""
    return bool((gold()["%s_w0_s0_l%d_nid" % (name, ls)][lo:hi] >= 0).all())


def random_tree(rng, k, L, p_stop=0.0, max_nodes=None, weighting=0, scoring=0, k_jitter=False):
    """A vocabulary with random node descriptors in the loader's array form (ids in depth-first file order like
    saveToTextFile writes them is NOT required: parents only have to precede their children).  p_stop makes inner
    nodes stop early (short branches); k_jitter varies the child count per node in [2, k]."""
    parent, leaf, level = [0], [0], [0]
    frontier = [0]
    while frontier:
        nxt = []
        for nid in frontier:
            kids = int(rng.integers(2, k + 1)) if k_jitter else k
            for _ in range(kids):
                parent.append(nid)
                leaf.append(0)
                level.append(level[nid] + 1)
                me = len(parent) - 1
                stop = level[me] >= L or (p_stop > 0 and level[me] >= 1 and rng.random() < p_stop) or \
                    (max_nodes is not None and len(parent) + len(nxt) * k > max_nodes)
                if stop:
                    leaf[me] = 1
                else:
                    nxt.append(me)
        frontier = nxt
    n = len(parent)
    desc = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    weight = np.where(np.array(leaf) > 0, np.round(rng.random(n) * 9.0, 5), 0.0)
    weight[rng.random(n) < 0.03] = 0.0                     # stopped words
    return vo.Vocabulary(k, L, scoring, weighting, parent, leaf, desc, weight)


def queries_near_nodes(rng, voc, n, flip=0.1):
    """descriptors near random WORDS of the tree (so that walks spread over it) with exact ties sprinkled in"""
    words = np.flatnonzero(voc.is_leaf > 0)
    pick = words[rng.integers(0, len(words), n)]
    noise = np.packbits(rng.random((n, 256)) < flip, axis=1)
    q = np.bitwise_xor(voc.desc[pick], noise)
    q[::17] = voc.desc[pick[::17]]
    return q
