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


# ------------------------------------------------------------------------------------------
# Matcher scenarios: SoA stand-ins for Frame / KeyFrame / MapPoint state.
def _scale_factors(n_levels=8, s=1.2):
    sf = np.ones(n_levels, np.float32)
    for i in range(1, n_levels):
        sf[i] = np.float32(np.float64(sf[i - 1]) * np.float64(np.float32(s)))
    return sf


def make_frame_arrays(n, idx=0, w=1241, h=376, n_levels=8, stereo=True, coarse_desc=False):
    """Random frame: keypoints uniform in the image, octaves by the per-level quota, random angles."""
    rng = np.random.default_rng(SEED_BASE + 0x300000 + int(idx))
    x = rng.uniform(20, w - 20, n).astype(np.float32); y = rng.uniform(20, h - 20, n).astype(np.float32)
    quota = np.array([1.2 ** -l for l in range(n_levels)]); quota /= quota.sum()
    octave = rng.choice(n_levels, n, p=quota).astype(np.int32)
    angle = rng.uniform(0, 360, n).astype(np.float32)
    desc = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    if coarse_desc:  # few distinct bit patterns -> many equal distances (tie-breaking paths)
        desc = np.repeat(rng.integers(0, 2, (n, 4), dtype=np.uint8) * 255, 8, axis=1)
    u_right = None
    if stereo:
        u_right = np.where(rng.random(n) < 0.6, x - rng.uniform(1, 60, n), -1.0).astype(np.float32)
    return dict(x=x, y=y, octave=octave, angle=angle, desc=desc, u_right=u_right, sf=_scale_factors(n_levels),
                bounds=(0.0, 0.0, float(w), float(h)))


def perturb_descriptors(rng, desc, flip_p=0.08):
    flips = rng.random((len(desc), 256)) < flip_p
    return desc ^ np.packbits(flips, axis=1, bitorder="little")


def make_projection_scenario(n_kp=2000, n_mp=2000, idx=0, match_frac=0.7, coarse_desc=False, **kw):
    """Frame + MapPoints whose projections fall near frame keypoints (SearchByProjection(F, MPs))."""
    rng = np.random.default_rng(SEED_BASE + 0x400000 + int(idx))
    f = make_frame_arrays(n_kp, idx, coarse_desc=coarse_desc, **kw)
    src = rng.integers(0, n_kp, n_mp)
    matched = rng.random(n_mp) < match_frac
    px = np.where(matched, f["x"][src] + rng.normal(0, 2, n_mp), rng.uniform(0, f["bounds"][2], n_mp)).astype(np.float32)
    py = np.where(matched, f["y"][src] + rng.normal(0, 2, n_mp), rng.uniform(0, f["bounds"][3], n_mp)).astype(np.float32)
    desc = np.where(matched[:, None], perturb_descriptors(rng, f["desc"][src]), rng.integers(0, 256, (n_mp, 32), dtype=np.uint8)).astype(np.uint8)
    if coarse_desc:
        desc = f["desc"][src]
    level = np.clip(f["octave"][src] + rng.integers(0, 2, n_mp), 0, len(f["sf"]) - 1).astype(np.int32)
    ur = f["u_right"][src] if f["u_right"] is not None else np.full(n_mp, -1.0)
    mps = dict(valid=(rng.random(n_mp) < 0.9), proj_x=px, proj_y=py,
               proj_xr=np.where(ur > 0, ur + rng.normal(0, 2, n_mp), px - 10).astype(np.float32), level=level,
               view_cos=np.where(rng.random(n_mp) < 0.5, 0.9995, 0.95).astype(np.float32), desc=desc,
               blocks=(rng.random(n_mp) < 0.9))
    occupied = rng.random(n_kp) < 0.1
    return f, mps, occupied


def make_lastframe_scenario(n_kp=2000, idx=0, cam=(718.856, 718.856, 607.19, 185.2, 386.1448), **kw):
    """Current frame + last-frame MapPoints given in the current camera frame (SearchByProjection(Cur, Last))."""
    rng = np.random.default_rng(SEED_BASE + 0x500000 + int(idx))
    f = make_frame_arrays(n_kp, idx, **kw)
    fx, fy, cx, cy, mbf = cam
    n = n_kp
    src = rng.permutation(n_kp)
    z = rng.uniform(4, 60, n).astype(np.float32)
    u = f["x"][src] + rng.normal(0, 3, n); v = f["y"][src] + rng.normal(0, 3, n)
    xc = ((u - cx) * z / fx).astype(np.float32); yc = ((v - cy) * z / fy).astype(np.float32)
    z[rng.random(n) < 0.03] *= -1  # behind the camera -> invzc < 0
    dangle = np.where(rng.random(n) < 0.8, rng.normal(0, 5, n), rng.uniform(0, 360, n))
    last = dict(valid=(rng.random(n) < 0.8), xc=xc, yc=yc, zc=z, octave=f["octave"][src].copy(),
                angle=np.mod(f["angle"][src] + dangle, 360).astype(np.float32), desc=perturb_descriptors(rng, f["desc"][src]),
                blocks=(rng.random(n) < 0.85))
    return f, last, cam


def make_bow_scenario(n1=2000, n2=2000, idx=0, n_nodes=120, coarse_desc=False, **kw):
    """Two feature sets with FeatureVectors: matched features mostly share a vocabulary node."""
    rng = np.random.default_rng(SEED_BASE + 0x600000 + int(idx))
    a = make_frame_arrays(n1, 2 * idx, coarse_desc=coarse_desc, **kw)
    b = make_frame_arrays(n2, 2 * idx + 1, coarse_desc=coarse_desc, **kw)
    nm = min(n1, n2) * 7 // 10
    ia = rng.permutation(n1)[:nm]; ib = rng.permutation(n2)[:nm]
    if not coarse_desc:
        b["desc"][ib] = perturb_descriptors(rng, a["desc"][ia], 0.05)
    b["angle"][ib] = np.mod(a["angle"][ia] + np.where(rng.random(nm) < 0.8, rng.normal(20, 4, nm), rng.uniform(0, 360, nm)), 360).astype(np.float32)
    b["x"][ib] = np.clip(a["x"][ia] + rng.normal(0, 15, nm), 20, 1220).astype(np.float32)
    b["y"][ib] = np.clip(a["y"][ia] + rng.normal(0, 1.0, nm), 20, 355).astype(np.float32)
    b["octave"][ib] = a["octave"][ia]
    node_a = rng.integers(0, n_nodes, n1) * 3 + 7
    node_b = rng.integers(0, n_nodes, n2) * 3 + 7
    same = rng.random(nm) < 0.9
    node_b[ib[same]] = node_a[ia[same]]
    node_b[node_b % 5 == 0] += 1      # some nodes exist on one side only (merge-join skips)
    fva = {}; fvb = {}
    for i in range(n1):
        fva.setdefault(int(node_a[i]), []).append(i)
    for i in range(n2):
        fvb.setdefault(int(node_b[i]), []).append(i)
    return a, fva, b, fvb, (rng.random(n1) < 0.6), (rng.random(n2) < 0.3)


def make_projected_points_scenario(n_kp=2000, n_pts=2000, idx=0, th=10.0, level_span=(1, 1), match_frac=0.7, coarse_desc=False, **kw):
    """Frame / KeyFrame + MapPoints already projected into it by the caller (the shared input of the
    relocalisation / loop SearchByProjection variants, Fuse and SearchBySim3): projection, radius
    th * sf[predicted level], accepted octaves [pred - level_span[0], pred + level_span[1]]."""
    rng = np.random.default_rng(SEED_BASE + 0x700000 + int(idx))
    f = make_frame_arrays(n_kp, 100 + idx, coarse_desc=coarse_desc, **kw)
    src = rng.integers(0, n_kp, n_pts)
    matched = rng.random(n_pts) < match_frac
    u = np.where(matched, f["x"][src] + rng.normal(0, 2, n_pts), rng.uniform(0, f["bounds"][2], n_pts)).astype(np.float32)
    v = np.where(matched, f["y"][src] + rng.normal(0, 2, n_pts), rng.uniform(0, f["bounds"][3], n_pts)).astype(np.float32)
    desc = np.where(matched[:, None], perturb_descriptors(rng, f["desc"][src]), rng.integers(0, 256, (n_pts, 32), dtype=np.uint8)).astype(np.uint8)
    if coarse_desc:
        desc = f["desc"][src]
    pred = np.clip(f["octave"][src] + rng.integers(0, 2, n_pts), 0, len(f["sf"]) - 1).astype(np.int32)
    urs = f["u_right"][src] if f["u_right"] is not None else np.full(n_pts, -1.0)
    dangle = np.where(rng.random(n_pts) < 0.8, rng.normal(30, 5, n_pts), rng.uniform(0, 360, n_pts))
    pts = dict(valid=(rng.random(n_pts) < 0.9), u=u, v=v,
               ur=np.where(urs >= 0, urs + rng.normal(0, 1.5, n_pts), u - 10).astype(np.float32),
               radius=(np.float32(th) * f["sf"][pred]).astype(np.float32),
               min_level=(pred - level_span[0]).astype(np.int32), max_level=(pred + level_span[1]).astype(np.int32),
               desc=desc, angle=np.mod(f["angle"][src] + dangle, 360).astype(np.float32))
    occupied = rng.random(n_kp) < 0.15
    return f, pts, occupied


def make_sim3_scenario(n1=2000, n2=2000, idx=0, th=7.5):
    """Two KeyFrames with mutually projected MapPoints (SearchBySim3): for ~70 % of the features the
    MapPoint of KF1 feature i lands on its KF2 partner j and vice versa."""
    rng = np.random.default_rng(SEED_BASE + 0x800000 + int(idx))
    a = make_frame_arrays(n1, 300 + 2 * idx); b = make_frame_arrays(n2, 301 + 2 * idx)
    nm = min(n1, n2) * 7 // 10
    ia = rng.permutation(n1)[:nm]; ib = rng.permutation(n2)[:nm]
    part_a = rng.integers(0, n2, n1); part_a[ia] = ib      # partner of each KF1 feature in KF2 (random for the rest)
    part_b = rng.integers(0, n1, n2); part_b[ib] = ia

    def project(dst, partner, n_src):
        pred = np.clip(dst["octave"][partner] + rng.integers(0, 2, n_src), 0, 7).astype(np.int32)
        return dict(valid=(rng.random(n_src) < 0.85), u=(dst["x"][partner] + rng.normal(0, 2, n_src)).astype(np.float32),
                    v=(dst["y"][partner] + rng.normal(0, 2, n_src)).astype(np.float32), ur=None,
                    radius=(np.float32(th) * dst["sf"][pred]).astype(np.float32), min_level=(pred - 1).astype(np.int32),
                    max_level=pred, desc=perturb_descriptors(rng, dst["desc"][partner], 0.05), angle=None)
    return a, project(b, part_a, n1), b, project(a, part_b, n2)


def make_initialization_scenario(n1=2000, n2=2000, idx=0, coarse_desc=False):
    """Two monocular frames for SearchForInitialization: F2 keypoints are F1 keypoints moved by a small
    flow; vbPrevMatched starts as the F1 positions (reference src/Tracking.cc:650-652)."""
    rng = np.random.default_rng(SEED_BASE + 0x900000 + int(idx))
    a = make_frame_arrays(n1, 200 + 2 * idx, stereo=False, coarse_desc=coarse_desc)
    b = make_frame_arrays(n2, 201 + 2 * idx, stereo=False, coarse_desc=coarse_desc)
    nm = min(n1, n2) * 8 // 10
    ia = rng.permutation(n1)[:nm]; ib = rng.permutation(n2)[:nm]
    if not coarse_desc:
        b["desc"][ib] = perturb_descriptors(rng, a["desc"][ia], 0.04)
    b["x"][ib] = np.clip(a["x"][ia] + rng.normal(8, 20, nm), 20, 1220).astype(np.float32)
    b["y"][ib] = np.clip(a["y"][ia] + rng.normal(0, 20, nm), 20, 355).astype(np.float32)
    b["octave"][ib] = a["octave"][ia]
    b["angle"][ib] = np.mod(a["angle"][ia] + np.where(rng.random(nm) < 0.8, rng.normal(5, 4, nm), rng.uniform(0, 360, nm)), 360).astype(np.float32)
    # several F1 keypoints share one F2 partner (duplicates in F1) -> the displacement rule (:509-513)
    dup = rng.permutation(n1)[: n1 // 10]
    srcs = rng.choice(ia, len(dup))
    a["x"][dup] = a["x"][srcs] + rng.normal(0, 1, len(dup)).astype(np.float32)
    a["y"][dup] = a["y"][srcs] + rng.normal(0, 1, len(dup)).astype(np.float32)
    a["octave"][dup] = a["octave"][srcs]
    if not coarse_desc:
        a["desc"][dup] = perturb_descriptors(rng, a["desc"][srcs], 0.02)
    return a, b, a["x"].copy(), a["y"].copy()


def make_vocabulary(k=10, L=3, seed=0, ragged=True, stop_frac=0.02):
    """A synthetic DBoW2 vocabulary tree in loadFromTextFile's node order (depth-first, as createTree /
    saveToTextFile emit it): each node's descriptor is its parent's with ~1/8 of the bits re-drawn, so
    descents are decisive near the root and tie-prone near the leaves; `ragged` drops children of some
    nodes (k-means clusters can come out empty); a few words get weight 0 (stopped words)."""
    rng = np.random.default_rng(0x5EED0000 + 0xB0000 + seed)
    parent, leaf, desc, weight = [0], [0], [np.zeros(32, np.uint8)], [0.0]

    def grow(pid, pdesc, level):
        nk = k if not ragged or level == 1 else int(rng.integers(max(2, k - 4), k + 1))
        for _ in range(nk):
            mask = np.packbits(rng.random(256) < (0.5 if level == 1 else 0.125))
            d = (pdesc & ~mask) | (rng.integers(0, 256, 32).astype(np.uint8) & mask)
            nid = len(parent)
            is_leaf = level == L
            parent.append(pid); leaf.append(int(is_leaf)); desc.append(d)
            weight.append(0.0 if not is_leaf else (0.0 if rng.random() < stop_frac else float(rng.uniform(0.5, 12.0))))
            if not is_leaf:
                grow(nid, d, level + 1)

    grow(0, np.zeros(32, np.uint8), 1)
    return dict(k=k, L=L, scoring=0, weighting=0, parent=np.array(parent, np.int32), is_leaf=np.array(leaf, np.uint8),
                desc=np.stack(desc), weight=np.array(weight, np.float64))


def write_text_vocabulary(path, voc):
    """ORBvoc.txt layout (TemplatedVocabulary.h:1452-1480 saveToTextFile).  No newline after the last node:
    loadFromTextFile's `while(!f.eof())` loop (:1402) would read one more, empty, line and append a node
    with unspecified fields under the root."""
    lines = ["%d %d %d %d" % (voc["k"], voc["L"], voc["scoring"], voc["weighting"])]
    for i in range(1, len(voc["parent"])):
        lines.append("%d %d %s %.17g" % (voc["parent"][i], voc["is_leaf"][i], " ".join(str(int(b)) for b in voc["desc"][i]), voc["weight"][i]))
    with open(path, "w") as f:
        f.write("\n".join(lines))


def vocabulary_queries(voc, n, seed=0, tie_frac=0.1):
    """Descriptors to transform: noisy copies of random leaves, plus some exact copies of INNER nodes' parents
    (equidistant children: exercises the first-child-wins tie rule)."""
    rng = np.random.default_rng(0x5EED0000 + 0xB1000 + seed)
    leaves = np.nonzero(voc["is_leaf"])[0]
    src = voc["desc"][rng.choice(leaves, n)]
    q = perturb_descriptors(rng, src)
    inner = np.nonzero(voc["is_leaf"] == 0)[0][1:]
    nt = int(n * tie_frac)
    if len(inner) and nt:
        q[:nt] = voc["desc"][rng.choice(inner, nt)]
    q[nt:nt + max(1, n // 50)] = rng.integers(0, 2, (max(1, n // 50), 1)).astype(np.uint8) * 255   # all-zero / all-one rows
    return q


def make_rectify_maps(w, h, seed=0, src_w=None, src_h=None):
    """Float x / y maps shaped like cv::initUndistortRectifyMap(..., CV_32F, M1, M2) output: a small rotation plus
    radial distortion around the principal point, so a band along the borders maps outside the source image."""
    rng = np.random.default_rng(SEED_BASE + 0xEC7000 + seed)
    src_w = w if src_w is None else src_w; src_h = h if src_h is None else src_h
    fx = 0.61 * w; fy = fx; cx = 0.49 * w; cy = 0.52 * h
    k1, k2 = -0.28 + 0.02 * rng.standard_normal(), 0.07 + 0.01 * rng.standard_normal()
    th = 0.01 * rng.standard_normal(3)
    R = np.array([[1, -th[2], th[1]], [th[2], 1, -th[0]], [-th[1], th[0], 1]])
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float64)
    X = (xx - cx) / fx; Y = (yy - cy) / fy
    P = np.stack([X, Y, np.ones_like(X)], -1) @ R.T
    x = P[..., 0] / P[..., 2]; y = P[..., 1] / P[..., 2]
    r2 = x * x + y * y
    d = 1 + k1 * r2 + k2 * r2 * r2
    mx = (fx * x * d + 0.5 * src_w + 1.3).astype(np.float32); my = (fy * y * d + 0.5 * src_h - 0.8).astype(np.float32)
    return mx, my
