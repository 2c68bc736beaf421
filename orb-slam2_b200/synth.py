"""Seeded synthetic inputs for tests and bench.py: textured grayscale
frames of KITTI / TUM / EuRoC shape, stereo pairs, and descriptor databases.

numpy only; deterministic for a given seed on any machine.
"""
import numpy as np


def _box_blur(a, r):
    """separable box blur of radius r with edge replication (float32 in/out)"""
    if r <= 0:
        return a
    k = 2 * r + 1
    for axis in (0, 1):
        pad = [(0, 0), (0, 0)]
        pad[axis] = (r + 1, r)
        p = np.pad(a, pad, mode="edge")
        c = np.cumsum(p, axis=axis, dtype=np.float64)
        if axis == 0:
            a = ((c[k:, :] - c[:-k, :]) / k).astype(np.float32)
        else:
            a = ((c[:, k:] - c[:, :-k]) / k).astype(np.float32)
    return a


def textured_frame(width, height, seed):
    """u8 frame: 3 octaves of band-limited noise + ~300 random rectangles/discs (strong FAST
    corners) + ~10 % flat regions (so the minThFAST retry path runs) + contrast stretch."""
    rng = np.random.default_rng(seed)
    img = np.zeros((height, width), np.float32)
    for r, wgt in ((1, 0.5), (2, 0.3), (4, 0.2)):
        n = rng.integers(0, 256, (height, width)).astype(np.float32)
        img += wgt * _box_blur(n, r)
    img = (img - img.mean()) * 2.5 + 128.0
    nshapes = max(60, int(300 * (width * height) / (1241.0 * 376.0)))
    yy, xx = np.mgrid[0:height, 0:width]
    for _ in range(nshapes):
        g = float(rng.integers(0, 256))
        cx, cy = int(rng.integers(0, width)), int(rng.integers(0, height))
        s = int(rng.integers(6, 40))
        if rng.random() < 0.6:
            x0, x1 = max(cx - s, 0), min(cx + s, width)
            y0, y1 = max(cy - s // 2, 0), min(cy + s // 2 + 1, height)
            img[y0:y1, x0:x1] = 0.7 * g + 0.3 * img[y0:y1, x0:x1]
        else:
            y0, y1 = max(cy - s, 0), min(cy + s + 1, height)
            x0, x1 = max(cx - s, 0), min(cx + s + 1, width)
            m = (xx[y0:y1, x0:x1] - cx) ** 2 + (yy[y0:y1, x0:x1] - cy) ** 2 <= s * s
            sub = img[y0:y1, x0:x1]
            sub[m] = 0.7 * g + 0.3 * sub[m]
    # flat patches: low-contrast regions where iniThFAST finds nothing
    nflat = max(4, int(0.10 * width * height / (80 * 60)))
    for _ in range(nflat):
        cx, cy = int(rng.integers(0, width)), int(rng.integers(0, height))
        x0, x1 = max(cx - 40, 0), min(cx + 40, width)
        y0, y1 = max(cy - 30, 0), min(cy + 30, height)
        sub = img[y0:y1, x0:x1]
        img[y0:y1, x0:x1] = sub.mean() + 0.12 * (sub - sub.mean())
    lo, hi = np.percentile(img, (0.5, 99.5))
    img = (img - lo) * (255.0 / max(hi - lo, 1.0))
    return np.clip(np.rint(img), 0, 255).astype(np.uint8)


def stereo_pair(width, height, seed):
    """left = textured_frame(seed); right = left shifted by a per-row-constant disparity
    (5..60 px, smooth in y) + N(0,2) noise."""
    left = textured_frame(width, height, seed)
    rng = np.random.default_rng(seed + 500000)
    ctrl = rng.uniform(5, 60, size=8)
    disp = np.interp(np.linspace(0, 7, height), np.arange(8), ctrl)
    disp = np.rint(disp).astype(np.int64)
    right = np.empty_like(left)
    cols = np.arange(width)
    for y in range(height):
        src = np.clip(cols + disp[y], 0, width - 1)
        right[y] = left[y, src]
    noise = rng.normal(0, 2.0, size=right.shape)
    right = np.clip(np.rint(right.astype(np.float32) + noise), 0, 255).astype(np.uint8)
    return left, right


def descriptor_db(nq, ndb, seed=7, max_flips=60):
    """C5 input: uniform random 256-bit rows with nq planted near-duplicates of the queries
    (k in 0..max_flips random bit flips) at seeded positions.  Returns (queries, db, planted_pos)."""
    rng = np.random.default_rng(seed)
    db = rng.integers(0, 256, (ndb, 32), dtype=np.uint8)
    q = rng.integers(0, 256, (nq, 32), dtype=np.uint8)
    pos = rng.choice(ndb, size=min(nq, ndb), replace=False)
    for i, p in enumerate(pos):
        row = np.unpackbits(q[i])
        k = int(rng.integers(0, max_flips + 1))
        flips = rng.choice(256, size=k, replace=False)
        row[flips] ^= 1
        db[p] = np.packbits(row)
    return q, db, pos
