"""Seeded synthetic inputs for the ORB front-end.

Images are procedurally textured 8-bit frames: 4-octave value noise, random rectangles with
+-40..120 contrast, 2 % salt noise and a ~15 % low-contrast band (which forces the extractor's
iniThFAST -> minThFAST retry, reference src/ORBextractor.cc:843-850).  Stereo right images are the
left image resampled with a smooth disparity field, so Frame::ComputeStereoMatches
(reference src/Frame.cc:516-704) finds true sub-pixel matches.  numpy only; no GPU, no oracle.
"""
import numpy as np

SEED_BASE = 0x5EED0000


def _value_noise(rng, h, w, cell):
    gh, gw = h // cell + 2, w // cell + 2
    g = rng.random((gh, gw), dtype=np.float32)
    ys = np.arange(h, dtype=np.float32) / cell
    xs = np.arange(w, dtype=np.float32) / cell
    y0 = ys.astype(np.int32); x0 = xs.astype(np.int32)
    fy = (ys - y0)[:, None]; fx = (xs - x0)[None, :]
    a = g[y0][:, x0]; b = g[y0][:, x0 + 1]; c = g[y0 + 1][:, x0]; d = g[y0 + 1][:, x0 + 1]
    return (a * (1 - fx) + b * fx) * (1 - fy) + (c * (1 - fx) + d * fx) * fy


def make_image(w, h, idx=0, n_rect=None):
    """One synthetic grayscale frame (h, w) uint8, deterministic in (w, h, idx)."""
    rng = np.random.default_rng(SEED_BASE + int(idx))
    img = np.zeros((h, w), np.float32)
    amp = 1.0
    for cell in (64, 32, 16, 8):
        img += amp * _value_noise(rng, h, w, cell)
        amp *= 0.5
    img = 40.0 + 150.0 * img / 1.875
    if n_rect is None:
        n_rect = max(200, min(2000, (w * h) // 600))
    x0 = rng.integers(0, w - 4, n_rect); y0 = rng.integers(0, h - 4, n_rect)
    rw = rng.integers(3, max(4, w // 12), n_rect); rh = rng.integers(3, max(4, h // 8), n_rect)
    con = rng.integers(40, 121, n_rect) * rng.choice((-1, 1), n_rect)
    for i in range(n_rect):
        img[y0[i]:y0[i] + rh[i], x0[i]:x0[i] + rw[i]] += con[i]
    # low-contrast band: ~15 % of the columns get their contrast divided by 6
    bx = int(rng.integers(0, max(1, w - int(0.15 * w))))
    bw = int(0.15 * w)
    band = img[:, bx:bx + bw]
    img[:, bx:bx + bw] = 128.0 + (band - 128.0) / 6.0
    salt = rng.random((h, w)) < 0.02
    img[salt] += rng.integers(-60, 61, int(salt.sum()))
    return np.clip(np.rint(img), 0, 255).astype(np.uint8)


def make_stereo_pair(w, h, idx=0, fx=718.856):
    """(left, right) with right(x, y) = left(x + d(x, y), y), d smooth in [2, min(0.25 fx, w/8)]."""
    left = make_image(w, h, idx)
    rng = np.random.default_rng(SEED_BASE + 0x100000 + int(idx))
    dmax = min(0.25 * fx, w / 8.0)
    d = 2.0 + (dmax - 2.0) * _value_noise(rng, h, w, 256)
    xs = np.arange(w, dtype=np.float32)[None, :] + d
    x0 = np.clip(np.floor(xs).astype(np.int32), 0, w - 1); x1 = np.clip(x0 + 1, 0, w - 1)
    f = (xs - np.floor(xs)).astype(np.float32)
    rows = np.arange(h)[:, None]
    L = left.astype(np.float32)
    right = L[rows, x0] * (1 - f) + L[rows, x1] * f
    return left, np.clip(np.rint(right), 0, 255).astype(np.uint8)


def make_descriptor_sets(n_a, n_b, idx=0, match_frac=0.7, flip_p=0.08):
    """Two descriptor sets (n, 32) uint8: match_frac of B are rows of A with Binomial(256, flip_p)
    flipped bits, permuted.  Returns (A, B, perm) with perm[j] = source row in A or -1."""
    rng = np.random.default_rng(SEED_BASE + 0x200000 + int(idx))
    A = rng.integers(0, 256, (n_a, 32), dtype=np.uint8)
    B = rng.integers(0, 256, (n_b, 32), dtype=np.uint8)
    n_m = min(int(match_frac * n_b), n_a)
    src = rng.permutation(n_a)[:n_m]
    dst = rng.permutation(n_b)[:n_m]
    flips = rng.random((n_m, 256)) < flip_p
    B[dst] = A[src] ^ np.packbits(flips, axis=1, bitorder="little")
    perm = np.full(n_b, -1, np.int64)
    perm[dst] = src
    return A, B, perm
