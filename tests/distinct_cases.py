"""Seeded groups of observed descriptors for MapPoint::ComputeDistinctiveDescriptors (reference MapPoint.cc:266-340):
shared by the oracle pin test (CPU) and the GPU parity test."""
import numpy as np


def groups(seed, sizes, flips=40):
    """each group = noisy copies of one base descriptor (what the keyframes observing one map point hold);
    returns desc [total, 32] u8 and offsets [G+1] i32"""
    rng = np.random.default_rng(seed)
    rows, offs = [], [0]
    for n in sizes:
        base = rng.integers(0, 256, 32, dtype=np.uint8)
        for _ in range(n):
            bits = np.unpackbits(base)
            k = int(rng.integers(0, flips + 1))
            bits[rng.choice(256, k, replace=False)] ^= 1
            rows.append(np.packbits(bits))
        offs.append(offs[-1] + n)
    desc = np.array(rows, np.uint8).reshape(-1, 32) if rows else np.zeros((0, 32), np.uint8)
    return desc, np.array(offs, np.int32)


def numpy_restatement(desc, offsets):
    """independent statement of the same selection: sort every row of the NxN distance matrix (diagonal 0),
    take index int(0.5*(N-1)), first strict minimum"""
    idx, med = [], []
    for g in range(len(offsets) - 1):
        d = desc[offsets[g]:offsets[g + 1]]
        n = len(d)
        if n == 0:
            idx.append(-1)
            med.append(0)
            continue
        D = np.unpackbits(d[:, None, :] ^ d[None, :, :], axis=2).sum(axis=2)
        m = np.sort(D, axis=1)[:, int(0.5 * (n - 1))]
        idx.append(int(np.argmin(m)))
        med.append(int(m.min()))
    return np.array(idx, np.int32), np.array(med, np.int32)


CASES = {
    "small": (11, [1, 2, 3, 4, 5, 7, 8, 9, 16, 31, 32]),
    "ragged_with_empty": (12, [0, 3, 0, 1, 12, 0]),
    "large": (13, [33, 64, 65, 100, 257, 500]),
    "ties": (14, [6] * 40),
}


def case(name):
    seed, sizes = CASES[name]
    desc, offs = groups(seed, sizes, flips=(2 if name == "ties" else 40))
    if name == "ties":  # exact duplicates: several rows share the least median, the first must win
        for g in range(0, 40, 2):
            o = offs[g]
            desc[o + 1] = desc[o + 3]
            desc[o + 4] = desc[o + 3]
    return desc, offs
