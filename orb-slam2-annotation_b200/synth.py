"""Seeded synthetic inputs shared by the tests, bench.py and the CPU baseline.

No datasets are reachable (no network, and the reference ships no images), so every workload uses
these generators.  numpy only; deterministic for a given (shape, seed).
"""
import numpy as np


def _cubic_upsample_1d(g, n, cell):
    """Catmull-Rom upsample along axis 0 of grid g (len >= n/cell + 3) to n samples."""
    t = np.arange(n, dtype=np.float64) / cell
    i = np.floor(t).astype(np.int64)
    f = (t - i)[:, None]
    p0, p1, p2, p3 = g[i], g[i + 1], g[i + 2], g[i + 3]
    return p1 + 0.5 * f * (p2 - p0 + f * (2 * p0 - 5 * p1 + 4 * p2 - p3 + f * (3 * (p1 - p2) + p3 - p0)))


def synth(w, h, seed):
    """Textured grayscale frame: multi-octave value noise + filled rectangles + mild blur + sensor noise.

    Yields a realistic FAST corner density (a few thousand raw corners per VGA pyramid, a good share of
    30-px cells needing the minTh fallback); uniform noise would give ~10x too many corners.
    """
    rng = np.random.default_rng(seed)
    img = np.zeros((h, w), np.float64)
    for o in range(2, 7):
        cell = 1 << o
        g = rng.random((h // cell + 4, w // cell + 4))
        rows = _cubic_upsample_1d(g, h, cell)                       # (h, gw)
        full = _cubic_upsample_1d(np.ascontiguousarray(rows.T), w, cell).T   # (h, w)
        img += full * cell
    img -= img.min()
    img *= 255.0 / max(img.max(), 1e-9)
    nrect = (w * h) // 3000
    xs = rng.integers(0, w, nrect); ys = rng.integers(0, h, nrect)
    ws = rng.integers(4, 40, nrect); hs = rng.integers(4, 40, nrect)
    gs = rng.integers(0, 256, nrect)
    for x, y, rw, rh, g in zip(xs, ys, ws, hs, gs):
        img[y:y + rh, x:x + rw] = g
    k = np.exp(-0.5 * (np.arange(-1, 2) / 0.7) ** 2); k /= k.sum()
    p = np.pad(img, 1, mode="reflect")
    img = k[0] * p[:-2, 1:-1] + k[1] * p[1:-1, 1:-1] + k[2] * p[2:, 1:-1]
    p = np.pad(img, 1, mode="reflect")
    img = k[0] * p[1:-1, :-2] + k[1] * p[1:-1, 1:-1] + k[2] * p[1:-1, 2:]
    img += rng.integers(-4, 5, (h, w))
    return np.clip(np.rint(img), 0, 255).astype(np.uint8)


def synth_stereo(w, h, seed):
    """Left frame + right frame = left warped by a smooth disparity field (4..60 px, nearest pixel)
    plus independent noise."""
    left = synth(w, h, seed)
    rng = np.random.default_rng(seed + 777_000)
    g = rng.random((h // 64 + 4, w // 64 + 4))
    rows = _cubic_upsample_1d(g, h, 64)
    disp = _cubic_upsample_1d(np.ascontiguousarray(rows.T), w, 64).T
    disp = 4 + (disp - disp.min()) / max(disp.max() - disp.min(), 1e-9) * 56
    xs = np.clip(np.arange(w)[None, :] + np.rint(disp).astype(np.int64), 0, w - 1)
    right = np.take_along_axis(left, xs, axis=1).astype(np.int16)
    right += rng.integers(-3, 4, (h, w)).astype(np.int16)
    return left, np.clip(right, 0, 255).astype(np.uint8)


def synth_descriptor_db(m, seed=42):
    """m x 32 uniform random descriptor bytes (config C5 database)."""
    rng = np.random.default_rng(seed)
    return rng.integers(0, 256, (m, 32), dtype=np.uint8)


def synth_queries(db, nq, seed=43, max_flips=60):
    """Half of the queries are database rows with k~U{0..max_flips} random bit flips, half fresh random.
    If `db` holds exactly nq//2 rows they are used as the planted rows in order."""
    rng = np.random.default_rng(seed)
    m = db.shape[0]
    q = rng.integers(0, 256, (nq, 32), dtype=np.uint8)
    nhit = nq // 2
    rows = np.arange(nhit) if m == nhit else rng.integers(0, m, nhit)
    hit = db[rows].copy()
    bits = np.unpackbits(hit, axis=1)
    for i in range(nhit):
        k = int(rng.integers(0, max_flips + 1))
        if k:
            pos = rng.choice(256, k, replace=False)
            bits[i, pos] ^= 1
    q[:nhit] = np.packbits(bits, axis=1)
    return q


def synth_vocabulary(k=10, L=3, seed=9, stop_fraction=0.05, ragged=False):
    """Seeded DBoW2-style vocabulary tree in loadFromTextFile order: (parent, is_leaf, desc, weight).
    Children are noisy copies of their parent's descriptor (so the descent is meaningful), leaf weights are positive idf-like
    doubles, a few leaves are stopped words (weight 0).  ragged=True makes some inner nodes leaves early and varies the fan-out."""
    rng = np.random.default_rng(seed)
    parent = [0]; leaf = [0]; desc = [np.zeros(32, np.uint8)]; weight = [0.0]; level = [0]
    frontier = [0]
    root_desc = rng.integers(0, 256, 32, dtype=np.uint8)
    desc[0] = root_desc
    while frontier:
        nxt = []
        for p in frontier:
            nch = k if not ragged else int(rng.integers(2, k + 1))
            for _ in range(nch):
                d = desc[p].copy()
                nflip = int(rng.integers(8, 48)) if level[p] == 0 else int(rng.integers(4, 28))
                for b in rng.integers(0, 256, nflip):
                    d[b >> 3] ^= np.uint8(1 << (b & 7))
                nid = len(parent)
                parent.append(p); desc.append(d); level.append(level[p] + 1)
                is_leaf = level[p] + 1 == L or (ragged and level[p] + 1 >= 2 and rng.random() < 0.15)
                leaf.append(1 if is_leaf else 0)
                weight.append(0.0 if not is_leaf else (0.0 if rng.random() < stop_fraction else float(rng.uniform(0.5, 9.5))))
                if not is_leaf:
                    nxt.append(nid)
        frontier = nxt
    return (np.asarray(parent, np.int32), np.asarray(leaf, np.uint8), np.stack(desc).astype(np.uint8), np.asarray(weight, np.float64))


def synth_vocabulary_features(desc, is_leaf, n, seed=10):
    """Query descriptors: noisy copies of random leaf descriptors plus some fresh random rows."""
    rng = np.random.default_rng(seed)
    leaves = np.flatnonzero(is_leaf)
    out = desc[rng.choice(leaves, n)].copy()
    noise = rng.random((n, 32, 8)) < 0.05
    out ^= np.packbits(noise, axis=2).reshape(n, 32)
    fresh = rng.random(n) < 0.1
    out[fresh] = rng.integers(0, 256, (int(fresh.sum()), 32), dtype=np.uint8)
    return out
