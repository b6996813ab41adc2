"""Pins the oracle's OpenCV-primitive restatements (oracle/prim.hpp) against the real cv2 when it
is importable (authoring container), and always against the committed vectors in tests/golden/
(made by tests/golden/make_golden.py from cv2 4.13.0)."""
import importlib
import os
import numpy as np
import pytest
import orc

synth = importlib.import_module("orb_slam2-chinese-annotation_b200.synth")
GOLD = os.path.join(os.path.dirname(__file__), "golden")

try:
    import cv2
    cv2.setNumThreads(1)
except Exception:  # pragma: no cover
    cv2 = None

needs_cv2 = pytest.mark.skipif(cv2 is None, reason="cv2 not importable")

LEVELS_KITTI = [(1241, 376), (1034, 313), (862, 261), (718, 218), (598, 181), (499, 151), (416, 126), (346, 105)]


def _images():
    rng = np.random.default_rng(7)
    yield "synth", synth.make_image(1241, 376, 3)
    yield "noise", rng.integers(0, 256, (376, 1241), dtype=np.uint8)
    yield "const", np.full((120, 160), 77, np.uint8)
    chk = ((np.arange(97)[:, None] + np.arange(131)[None, :]) & 1).astype(np.uint8) * 255
    yield "checker", chk


@needs_cv2
def test_resize_chain_matches_cv2():
    for name, img in _images():
        cur = img
        h0, w0 = img.shape
        for l in range(1, 8):
            s = np.float32(1.0) / np.float32(1.2 ** l)
            dw, dh = int(round(w0 * float(s))), int(round(h0 * float(s)))
            if dw < 8 or dh < 8:
                break
            ref = cv2.resize(cur, (dw, dh), interpolation=cv2.INTER_LINEAR)
            got = orc.resize_linear(cur, dw, dh)
            assert np.array_equal(ref, got), (name, l)
            cur = ref


@needs_cv2
def test_resize_odd_ratios_matches_cv2():
    rng = np.random.default_rng(11)
    for _ in range(20):
        h, w = rng.integers(20, 200, 2)
        img = rng.integers(0, 256, (h, w), dtype=np.uint8)
        dw = int(rng.integers(max(8, w // 2), w + 1)); dh = int(rng.integers(max(8, h // 2), h + 1))
        assert np.array_equal(cv2.resize(img, (dw, dh), interpolation=cv2.INTER_LINEAR), orc.resize_linear(img, dw, dh))


@needs_cv2
def test_gauss7_matches_cv2():
    for name, img in _images():
        ref = cv2.GaussianBlur(img, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
        assert np.array_equal(ref, orc.gauss7(img)), name
    rng = np.random.default_rng(5)
    for (w, h) in [(8, 8), (9, 31), (346, 105), (33, 7)]:
        img = rng.integers(0, 256, (h, w), dtype=np.uint8)
        ref = cv2.GaussianBlur(img, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
        assert np.array_equal(ref, orc.gauss7(img)), (w, h)


@needs_cv2
def test_border_matches_cv2():
    rng = np.random.default_rng(6)
    img = rng.integers(0, 256, (40, 55), dtype=np.uint8)
    ref = cv2.copyMakeBorder(img, 19, 19, 19, 19, cv2.BORDER_REFLECT_101)
    assert np.array_equal(ref, orc.border101(img, 19))


def _cv_fast(img, th):
    fd = cv2.FastFeatureDetector_create(int(th), True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
    return np.array([(int(k.pt[0]), int(k.pt[1]), int(k.response)) for k in fd.detect(img)], np.int32).reshape(-1, 3)


@needs_cv2
def test_fast_matches_cv2_including_order():
    rng = np.random.default_rng(8)
    imgs = [synth.make_image(300, 200, 1), synth.make_image(640, 480, 2),
            rng.integers(0, 256, (64, 64), dtype=np.uint8),
            (rng.integers(0, 4, (90, 120)) * 60).astype(np.uint8)]  # plateaus -> score ties
    for img in imgs:
        for th in (7, 20, 40):
            ref = _cv_fast(img, th)
            got = orc.fast9_16(img, th)
            assert ref.shape == got.shape and np.array_equal(ref, got), (img.shape, th, len(ref), len(got))


@needs_cv2
def test_fast_small_cells_match_cv2():
    img = synth.make_image(400, 300, 4)
    rng = np.random.default_rng(9)
    for _ in range(60):
        w = int(rng.integers(7, 45)); h = int(rng.integers(7, 45))
        x = int(rng.integers(0, 400 - w)); y = int(rng.integers(0, 300 - h))
        view = img[y:y + h, x:x + w]
        for th in (7, 20):
            ref = _cv_fast(np.ascontiguousarray(view), th)
            got = orc.fast9_16(np.ascontiguousarray(view), th)
            assert np.array_equal(ref, got)


def test_score_map_is_threshold_free_fast():
    """corner at threshold t  <=>  score >= t, and the detector response equals the score."""
    img = synth.make_image(200, 150, 5)
    sc = orc.fast_score_map(img)
    for th in (7, 20):
        kp = orc.fast9_16(img, th)
        assert len(kp) > 0
        assert np.all(sc[kp[:, 1], kp[:, 0]] == kp[:, 2])
        assert np.all(kp[:, 2] >= th)


@needs_cv2
def test_fast_atan2_matches_cv2():
    rng = np.random.default_rng(10)
    ys = rng.integers(-60000, 60000, 20000).astype(np.float32)
    xs = rng.integers(-60000, 60000, 20000).astype(np.float32)
    ys[:50] = 0; xs[25:75] = 0
    for y, x in zip(ys, xs):
        assert cv2.fastAtan2(float(y), float(x)) == orc.fast_atan2(y, x), (y, x)


def test_primitives_match_committed_golden():
    g = np.load(os.path.join(GOLD, "primitives.npz"))
    img = synth.make_image(int(g["w"]), int(g["h"]), int(g["idx"]))
    assert np.array_equal(img, g["img"]), "synthetic generator drifted"
    assert np.array_equal(orc.resize_linear(img, int(g["rw"]), int(g["rh"])), g["resized"])
    assert np.array_equal(orc.gauss7(img), g["blurred"])
    assert np.array_equal(orc.border101(img, 19), g["bordered"])
    for th in (7, 20):
        assert np.array_equal(orc.fast9_16(img, th), g["fast%d" % th])
    at = np.array([orc.fast_atan2(y, x) for y, x in g["atan_in"]], np.float32)
    assert np.array_equal(at, g["atan_out"])


def test_remap_and_gray_match_cv2_golden():
    """cv::remap(INTER_LINEAR, float maps, BORDER_CONSTANT) and cv::cvtColor(*2GRAY) restatements vs vectors made by the real
    cv2 4.13 (tests/golden/make_golden_rectify.py), incl. half-step rounding, border, far-outside and NaN coordinates."""
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "rectify.npz"))
    assert np.array_equal(orc.remap_linear_u8(g["img"], g["mx"], g["my"]), g["remapped"])
    assert np.array_equal(orc.gray_u8(g["bgr"], False), g["bgr2gray"]) and np.array_equal(orc.gray_u8(g["bgr"], True), g["rgb2gray"])
    assert np.array_equal(orc.gray_u8(g["bgra"], False), g["bgra2gray"]) and np.array_equal(orc.gray_u8(g["bgra"], True), g["rgba2gray"])
    assert (g["remapped"] == 0).mean() < 0.5


def test_remap_matches_live_cv2():
    cv2 = pytest.importorskip("cv2")
    import importlib
    synth = importlib.import_module("orb_slam2-chinese-annotation_b200.synth")
    for (w, h, seed) in [(752, 480, 0), (640, 480, 1)]:
        img = synth.make_image(w, h, seed)
        mx, my = synth.make_rectify_maps(w, h, seed)
        assert np.array_equal(orc.remap_linear_u8(img, mx, my), cv2.remap(img, mx, my, cv2.INTER_LINEAR))
