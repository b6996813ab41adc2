"""GPU parity of stereo rectification (cv::remap INTER_LINEAR, float maps) and gray conversion (cv::cvtColor) — the steps
in front of the extractor (reference Examples/Stereo/stereo_euroc.cc:136-137, src/Tracking.cc:174-199) — against the oracle
restatements (pinned to cv2 4.13 by tests/test_oracle_primitives.py) and the committed cv2 golden vectors."""
import importlib
import os
import numpy as np
import pytest
import orc

synth = importlib.import_module("orb_slam2-chinese-annotation_b200.synth")
rect = importlib.import_module("orb_slam2-chinese-annotation_b200.rectify")
ext = importlib.import_module("orb_slam2-chinese-annotation_b200.extractor")
pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "rectify.npz")


def test_remap_and_gray_match_cv2_golden():
    g = np.load(GOLD)
    r = rect.Rectifier(g["mx"], g["my"])
    assert np.array_equal(r.remap(g["img"]), g["remapped"])          # incl. NaN / inf / half-step / border coordinates
    assert np.array_equal(r.gray(g["bgr"], False), g["bgr2gray"]) and np.array_equal(r.gray(g["bgr"], True), g["rgb2gray"])
    assert np.array_equal(r.gray(g["bgra"], False), g["bgra2gray"]) and np.array_equal(r.gray(g["bgra"], True), g["rgba2gray"])
    assert r.launch_count == 5


@pytest.mark.parametrize("w,h,sw,sh,B", [(752, 480, 752, 480, 3), (1241, 376, 1241, 376, 2), (601, 333, 640, 360, 2)])
def test_remap_matches_oracle(w, h, sw, sh, B):
    mx, my = synth.make_rectify_maps(w, h, 1, sw, sh)
    imgs = np.stack([synth.make_image(sw, sh, 30 + b) for b in range(B)])
    r = rect.Rectifier(mx, my, src_size=(sw, sh), max_batch=B)
    out = r.remap(imgs)
    for b in range(B):
        assert np.array_equal(out[b], orc.remap_linear_u8(imgs[b], mx, my)), b
    assert 0.0 < (out == 0).mean() < 0.5


def test_rectified_pair_through_extractor():
    """EuRoC flow: raw pair -> remap -> ORBextractor, all results identical to the oracle chain."""
    w, h, nf = 752, 480, 1200
    mx, my = synth.make_rectify_maps(w, h, 2)
    raw = synth.make_image(w, h, 77)
    r = rect.Rectifier(mx, my)
    rectified = r.remap(raw)
    g = ext.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
    k, d = g(rectified)
    o = orc.Extractor(nf, 1.2, 8, 20, 7)
    ok, od = o.extract(orc.remap_linear_u8(raw, mx, my))
    assert len(k) == len(ok) and np.array_equal(k["x"], ok["x"]) and np.array_equal(k["angle"], ok["angle"]) and np.array_equal(d, od)
