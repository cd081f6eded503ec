"""Seeded synthetic inputs: no images ship with the reference.

Integer-only numpy so the same seed gives the same bytes on every machine: filled rectangles and
discs on a mid-grey plane, a separable binomial blur, small additive noise. Yields ~7 FAST
candidates per kept keypoint at 640x480 and exercises the minThFAST fallback and octree phase 2.
"""
import numpy as np


def _binomial_blur(img):
    k = np.array([1, 4, 6, 4, 1], np.int32)
    a = img.astype(np.int32)
    p = np.pad(a, ((0, 0), (2, 2)), mode="edge")
    a = sum(k[i] * p[:, i:i + a.shape[1]] for i in range(5))
    p = np.pad(a, ((2, 2), (0, 0)), mode="edge")
    a = sum(k[i] * p[i:i + a.shape[0], :] for i in range(5))
    return ((a + 128) >> 8).astype(np.int32)


def frame(w, h, seed, kind="shapes"):
    """u8 (h, w) frame. kind: shapes | lowcontrast | constant | noise."""
    rng = np.random.Generator(np.random.PCG64(seed))
    if kind == "constant":
        return np.full((h, w), 128, np.uint8)
    if kind == "noise":
        return rng.integers(0, 256, (h, w), dtype=np.uint8)
    img = np.full((h, w), 128, np.int32)
    n = max(1, (w * h) // 600)
    kinds = rng.integers(0, 2, n)
    cx = rng.integers(0, w, n)
    cy = rng.integers(0, h, n)
    side = rng.integers(4, 60, n)
    rad = rng.integers(2, 30, n)
    grey = rng.integers(0, 256, n)
    for i in range(n):
        if kinds[i] == 0:
            x0, y0 = max(0, cx[i] - side[i] // 2), max(0, cy[i] - side[i] // 2)
            img[y0:min(h, y0 + side[i]), x0:min(w, x0 + side[i])] = grey[i]
        else:
            r = int(rad[i])
            x0, x1 = max(0, cx[i] - r), min(w, cx[i] + r + 1)
            y0, y1 = max(0, cy[i] - r), min(h, cy[i] + r + 1)
            yy, xx = np.ogrid[y0:y1, x0:x1]
            m = (xx - cx[i]) ** 2 + (yy - cy[i]) ** 2 <= r * r
            img[y0:y1, x0:x1][m] = grey[i]
    img = _binomial_blur(img)
    noise = rng.integers(-2, 3, (h, w)) + rng.integers(-2, 3, (h, w))
    img = img + noise
    if kind == "lowcontrast":
        img = 128 + (img - 128) // 12
    return np.clip(img, 0, 255).astype(np.uint8)


def stereo_pair(w, h, seed):
    """Left frame and a right frame = left shifted by a per-row-band disparity + fresh noise."""
    left = frame(w, h, seed)
    rng = np.random.Generator(np.random.PCG64(seed + 1_000_003))
    right = np.empty_like(left)
    band = 16
    for y0 in range(0, h, band):
        d = int(rng.integers(1, 64))
        rows = left[y0:y0 + band]
        shifted = np.empty_like(rows)
        shifted[:, :w - d] = rows[:, d:]
        shifted[:, w - d:] = rows[:, w - 1:w]
        right[y0:y0 + band] = shifted
    noise = rng.integers(-1, 2, (h, w))
    right = np.clip(right.astype(np.int32) + noise, 0, 255).astype(np.uint8)
    return left, right


def descriptors(n, seed):
    rng = np.random.Generator(np.random.PCG64(seed))
    return rng.integers(0, 256, (n, 32), dtype=np.uint8)


def planted_queries(db, nq, seed, frac=0.5, max_flips=40):
    """nq random queries; a fraction are database rows with U[0,max_flips] flipped bits."""
    rng = np.random.Generator(np.random.PCG64(seed))
    q = rng.integers(0, 256, (nq, 32), dtype=np.uint8)
    nplant = int(nq * frac)
    rows = rng.integers(0, db.shape[0], nplant)
    for i in range(nplant):
        d = db[rows[i]].copy()
        k = int(rng.integers(0, max_flips + 1))
        bits = rng.choice(256, k, replace=False)
        for b in bits:
            d[b >> 3] ^= np.uint8(1 << (b & 7))
        q[i] = d
    return q
