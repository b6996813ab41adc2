"""Self-checks of SURVEY.md 8(c) on the GPU: inputs engineered for the places where an almost-right kernel goes wrong.

  - cv::fastAtan2 on every sign / ratio class, (0, 0) included, and on integer moments of the orientation stage's range;
  - resize + 7x7 blur on 1-pixel checkerboards, constants and noise at EVERY level size of every configuration (each level size
    becomes the size of a level-0 image: all tile remainders and edge-reflection cases of the TMA kernels occur);
  - FAST on images full of plateau ties (equal scores side by side in the non-maximum suppression);
  - the quadtree on candidates that reach the sorted phase (src/ORBextractor.cc:692-757) with EQUAL counts at the early break,
    witnessed by the oracle's own counters;
  - a property test (hypothesis) of the claim order of SearchByProjection(Frame, MapPoints): few keypoints, many points, descriptors
    from a tiny pool (distance ties), any mixture of blocking and non-blocking points.
Everything is compared bit for bit with the CPU oracle (whose primitives are pinned to OpenCV 4.13 by tests/golden)."""
import ctypes as C
import importlib
import numpy as np
import pytest
from hypothesis import HealthCheck, given, settings, strategies as st

import orc

PKG = "orb_slam2-chinese-annotation_b200"
pytestmark = pytest.mark.gpu
capi = importlib.import_module(PKG + ".capi")
synth = importlib.import_module(PKG + ".synth")
ext = importlib.import_module(PKG + ".extractor")
mt = importlib.import_module(PKG + ".matcher")
views = importlib.import_module(PKG + ".views")

CONFIGS = [(640, 480, 1000), (1241, 376, 2000), (752, 480, 1200), (3840, 2160, 8000)]   # BASELINE.json configs 0..4 geometries


def _atan2_gpu(y, x):
    y = np.ascontiguousarray(y, np.float32); x = np.ascontiguousarray(x, np.float32)
    out = np.empty(len(y), np.float32)
    capi.check(capi.lib().orb_fast_atan2_device(capi.ptr(y), capi.ptr(x), len(y), 0, capi.ptr(out)))
    return out


def test_fast_atan2_every_class():
    """IC_Angle's fastAtan2((float)m_01, (float)m_10) (src/ORBextractor.cc:112): (0, 0) -> 0, the axes, the diagonals (ax == ay),
    both octant branches, denormal and huge ratios, and 200 k integer moment pairs of the stage's range (|m| <= 749*15*255)."""
    special = [0.0, -0.0, 1.0, -1.0, 2.0, -2.0, 0.5, -0.5, 1e-30, -1e-30, 1e30, -1e30, 2864925.0, -2864925.0, 3.0, -3.0, 1.0000001, 0.99999994]
    ys, xs = np.meshgrid(np.array(special, np.float32), np.array(special, np.float32))
    rng = np.random.default_rng(0)
    m = 749 * 15 * 255
    yi = rng.integers(-m, m + 1, 200000).astype(np.float32); xi = rng.integers(-m, m + 1, 200000).astype(np.float32)
    ysm = rng.integers(-60, 61, 20000).astype(np.float32); xsm = rng.integers(-60, 61, 20000).astype(np.float32)   # many exact ratio ties
    y = np.concatenate([ys.ravel(), yi, ysm]); x = np.concatenate([xs.ravel(), xi, xsm])
    got = _atan2_gpu(y, x)
    want = np.array([orc.fast_atan2(a, b) for a, b in zip(y.tolist(), x.tolist())], np.float32)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    assert _atan2_gpu([0.0], [0.0])[0] == 0.0 and _atan2_gpu([-0.0], [-0.0])[0] == orc.fast_atan2(-0.0, -0.0)
    assert np.all((got >= 0) & (got <= 360.0))


def _level_sizes(w, h, nf):
    o = orc.Extractor(nf, 1.2, 8, 20, 7)
    o.extract(np.zeros((h, w), np.uint8))
    return [o.level(l).shape[::-1] for l in range(8)]


def _patterns(w, h, seed):
    yy, xx = np.mgrid[0:h, 0:w]
    rng = np.random.default_rng(seed)
    return {"checker1": (((xx + yy) & 1) * 255).astype(np.uint8), "checker1_low": (100 + ((xx + yy) & 1) * 7).astype(np.uint8),
            "zero": np.zeros((h, w), np.uint8), "white": np.full((h, w), 255, np.uint8), "const90": np.full((h, w), 90, np.uint8),
            "columns": ((xx & 1) * 255).astype(np.uint8), "rows": ((yy & 1) * 255).astype(np.uint8),
            "noise": rng.integers(0, 256, (h, w), dtype=np.uint8)}


@pytest.mark.parametrize("w,h,nf", CONFIGS)
def test_resize_and_blur_on_checker_constant_noise_at_every_level_size(w, h, nf):
    """Every level size of the configuration as a level-0 image: all 8 pyramid levels (OpenCV 8u INTER_LINEAR, src/ORBextractor.cc:1195)
    and their 7x7 sigma-2 blurs (:1145, BORDER_REFLECT_101) bit-equal to the oracle's primitives."""
    checked = 0
    for li, (lw, lh) in enumerate(_level_sizes(w, h, nf)):
        if lw < 64 or lh < 64:
            continue
        big = lw * lh > 2_000_000
        g = ext.ORBextractor(nf, 1.2, 8, 20, 7, max_width=lw, max_height=lh)
        o = orc.Extractor(nf, 1.2, 8, 20, 7)
        for name, img in _patterns(lw, lh, li).items():
            if big and name not in ("checker1", "const90", "noise"):
                continue
            g(img); o.extract(img)
            for l in range(8):
                lv = o.level(l)
                assert np.array_equal(g.image_pyramid(l), lv), (name, lw, lh, l)
                assert np.array_equal(g.image_pyramid(l, blurred=True), orc.gauss7(lv)), (name, lw, lh, l, "blur")
                checked += 1
    assert checked >= 8 * 3 * 6


@pytest.mark.parametrize("block,levels,seed", [(2, 4, 0), (3, 3, 1), (4, 6, 2), (5, 2, 3)])
def test_fast_on_plateau_ties(block, levels, seed):
    """Block images with few grey levels: neighbouring pixels score EQUAL, and cv::FAST's non-maximum suppression keeps a pixel only if
    its score is strictly greater than all 8 neighbours (ties suppress both).  FAST candidates, quadtree and final keypoints equal
    the oracle's; the image is checked to contain such ties."""
    w, h = 640, 480
    rng = np.random.default_rng(seed)
    small = rng.integers(0, levels, ((h + block - 1) // block, (w + block - 1) // block)) * (255 // max(levels - 1, 1))
    img = np.kron(small, np.ones((block, block), np.int64))[:h, :w].astype(np.uint8)
    sc = orc.fast_score_map(img).astype(np.int32)
    ties = int(((sc[:, 1:] == sc[:, :-1]) & (sc[:, 1:] > 0)).sum() + ((sc[1:, :] == sc[:-1, :]) & (sc[1:, :] > 0)).sum())
    assert ties > 500, ties
    g = ext.ORBextractor(1000, 1.2, 8, 20, 7, max_width=w, max_height=h)
    o = orc.Extractor(1000, 1.2, 8, 20, 7)
    kps, desc = g(img); okps, odesc = o.extract(img)
    for l in range(8):
        gc, oc = g.level_candidates(l), o.level_candidates(l)
        assert gc.shape == oc.shape and np.array_equal(gc, oc), l
    assert len(kps) == len(okps) and np.array_equal(desc, odesc)
    for f in ("x", "y", "response", "octave"):
        assert np.array_equal(kps[f], okps[f]), f
    assert np.array_equal(kps["angle"].view(np.uint32), okps["angle"].view(np.uint32))


def test_quadtree_sorted_phase_with_equal_counts():
    """Periodic images: the candidates form a lattice, so the quadtree's nodes come in groups of EQUAL count.  For several nFeatures
    the sorted phase (src/ORBextractor.cc:692-757) stops in the middle of such a group, where only the (count, pointer) order - the
    canonical 'later-created node first' tie-break of this repository - decides which node is expanded.  The oracle's witness
    counters prove the case occurred; the GPU's per-level selection equals the oracle's."""
    w, h = 768, 384
    rng = np.random.default_rng(7)
    tile = rng.integers(0, 2, (6, 6)) * 200 + 20
    tile = np.kron(tile, np.ones((8, 8), np.int64))                      # a 48x48 tile of 8x8 blocks
    img = np.tile(tile, (h // 48 + 1, w // 48 + 1))[:h, :w].astype(np.uint8)
    total = {"rounds": 0, "equal": 0, "decisive": 0}
    for nf in (300, 500, 700, 1000, 1400, 2000):
        g = ext.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
        o = orc.Extractor(nf, 1.2, 8, 20, 7)
        g(img); o.extract(img)
        n_per_level = o.tables()["n_per_level"]
        for l in range(8):
            gk, ok = g.level_keypoints(l), o.level_keypoints(l)
            assert len(gk) == len(ok) and np.array_equal(gk[:, 0], ok["x"].astype(np.int32)) and np.array_equal(gk[:, 1], ok["y"].astype(np.int32)), (nf, l)
            cands = o.level_candidates(l)
            lw, lh = o.level(l).shape[::-1]
            kept, stt = orc.distribute_stats(cands, 16, lw - 16, 16, lh - 16, int(n_per_level[l]))
            assert len(kept) == len(ok), (nf, l)
            total["rounds"] += int(stt[0]); total["equal"] += int(stt[1]); total["decisive"] += int(stt[2])
    assert total["rounds"] > 5 and total["equal"] > 100 and total["decisive"] >= 3, total


@settings(max_examples=60, deadline=None, suppress_health_check=[HealthCheck.too_slow, HealthCheck.function_scoped_fixture])
@given(n_kp=st.integers(1, 48), n_mp=st.integers(1, 160), pool=st.integers(1, 4), spread=st.integers(2, 40), th=st.sampled_from([1.0, 3.0, 5.0]),
       p_block=st.sampled_from([0.0, 0.3, 0.9, 1.0]), seed=st.integers(0, 2 ** 31 - 1))
def test_claim_order_of_search_by_projection(n_kp, n_mp, pool, spread, th, p_block, seed):
    """SearchByProjection(Frame, MapPoints) (src/ORBmatcher.cc:47-133): points are served in index order; a keypoint holding a point with
    observations is skipped by later points, one holding a point without observations is taken over.  Few keypoints crowded into
    a small window, many points, descriptors from a pool of <= 4 rows (distances tie all the time): the device's optimistic
    group resolution with serial replay must end in the same assignment as the serial loop."""
    rng = np.random.default_rng(seed)
    f = synth.make_frame_arrays(max(n_kp, 1), seed % 1000)
    cx, cy = 300.0, 150.0
    f["x"] = (cx + rng.integers(0, spread, n_kp)).astype(np.float32); f["y"] = (cy + rng.integers(0, spread, n_kp)).astype(np.float32)
    f["octave"] = rng.integers(0, 3, n_kp).astype(np.int32)
    rows = rng.integers(0, 256, (pool, 32), dtype=np.uint8)
    rows[1:] = rows[0] ^ (rng.integers(0, 256, (pool - 1, 32), dtype=np.uint8) & 0x03)      # a few bits apart: below TH_HIGH
    f["desc"] = rows[rng.integers(0, pool, n_kp)]
    f["u_right"] = np.where(rng.random(n_kp) < 0.5, f["x"] - 5, -1).astype(np.float32)
    F = views.Frame(f["x"], f["y"], f["octave"], f["angle"], f["desc"], f["sf"], f["bounds"], f["u_right"])
    mps = dict(valid=rng.random(n_mp) < 0.95, proj_x=(cx + rng.uniform(-2, spread + 2, n_mp)).astype(np.float32),
               proj_y=(cy + rng.uniform(-2, spread + 2, n_mp)).astype(np.float32), proj_xr=(cx - 5 + rng.uniform(-2, spread + 2, n_mp)).astype(np.float32),
               level=rng.integers(0, 3, n_mp).astype(np.int32), view_cos=np.where(rng.random(n_mp) < 0.5, 0.9995, 0.95).astype(np.float32),
               desc=rows[rng.integers(0, pool, n_mp)], blocks=rng.random(n_mp) < p_block)
    occ = rng.random(n_kp) < 0.1
    m = _matcher()
    for ratio in (0.6, 0.8):
        m.mfNNratio = ratio
        n_g, m_g = m.SearchByProjection(F, mps, th, occ)
        n_o, m_o = orc.search_by_projection_mps(F, mps, th, ratio, occ)
        assert n_g == n_o and np.array_equal(m_g, m_o)


_M = []


def _matcher():
    if not _M:
        _M.append(mt.ORBmatcher(0.6, True, max_n=256, max_pairs=1))
    return _M[0]
