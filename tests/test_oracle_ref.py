"""Pins the oracle's restatement of ORBextractor (oracle/orb_oracle.cpp) against the REFERENCE'S OWN
src/ORBextractor.cc compiled verbatim (oracle/_ref, built by `make -C oracle ref` where
/root/reference exists; the .so travels to the GPU box).  Keypoints, angles and descriptors must be
bit-identical.  The reference runs under a monotonic heap (oracle/ref_wrap.cpp), which is what
makes its pointer-compared quadtree ties deterministic."""
import importlib
import numpy as np
import pytest
import orc

synth = importlib.import_module("orb_slam2-chinese-annotation_b200.synth")
pytestmark = pytest.mark.skipif(not orc.ref_available(), reason="oracle/_ref not built (needs /root/reference)")

CASES = [(640, 480, 1000, 0), (1241, 376, 2000, 1), (752, 480, 1200, 2), (333, 257, 500, 3), (200, 120, 300, 4), (1241, 376, 2000, 5)]


@pytest.mark.parametrize("w,h,nf,idx", CASES)
def test_restatement_equals_reference(w, h, nf, idx):
    img = synth.make_image(w, h, idx)
    r = orc.RefExtractor(nf, 1.2, 8, 20, 7); o = orc.Extractor(nf, 1.2, 8, 20, 7)
    rk, rd = r.extract(img); ok, od = o.extract(img)
    assert len(rk) == len(ok) and len(rk) > 100
    for f in ("x", "y", "size", "response", "octave"):
        assert np.array_equal(rk[f], ok[f]), f
    assert np.array_equal(rk["angle"].view(np.uint32), ok["angle"].view(np.uint32))
    assert np.array_equal(rd, od)
    for l in range(8):
        assert np.array_equal(r.level(l), o.level(l))


def test_reference_special_images():
    rng = np.random.default_rng(0)
    imgs = [rng.integers(0, 256, (300, 400), dtype=np.uint8),                       # dense: sorted-phase quadtree with many ties
            np.full((300, 400), 90, np.uint8),                                       # no corners at all
            (128 + (rng.integers(0, 256, (300, 400)) - 128) // 12).astype(np.uint8)]  # minThFAST retry only
    for img in imgs:
        r = orc.RefExtractor(800, 1.2, 8, 20, 7); o = orc.Extractor(800, 1.2, 8, 20, 7)
        rk, rd = r.extract(img); ok, od = o.extract(img)
        assert len(rk) == len(ok)
        assert np.array_equal(rk, ok) and np.array_equal(rd, od)


def test_reference_is_deterministic_under_the_monotonic_heap():
    img = synth.make_image(640, 480, 9)
    r = orc.RefExtractor(1000, 1.2, 8, 20, 7)
    a = r.extract(img); junk = [bytearray(100 * (i % 7 + 1)) for i in range(1000)]; b = r.extract(img)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and junk


SETTINGS = [  # the non-default ORBextractor settings of tests/test_gpu_extractor.py::test_other_extractor_settings
    (640, 480, 700, 1.1, 5, 12, 5),
    (640, 480, 400, 1.5, 3, 30, 10),
    (752, 480, 1500, 1.2, 12, 20, 7),
    (500, 375, 600, 1.3, 6, 20, 20),
    (640, 480, 300, 1.2, 1, 40, 7),
]


@pytest.mark.parametrize("w,h,nf,sf,nl,ini,mn", SETTINGS)
def test_restatement_equals_reference_at_other_settings(w, h, nf, sf, nl, ini, mn):
    """Every reference yaml uses 1.2 / 8 / 20 / 7; the class takes any.  The restatement the GPU tests are checked against must follow
    the verbatim reference there too: scale tables, features per level, pyramid bytes, keypoints, angle bits, descriptors."""
    r = orc.RefExtractor(nf, sf, nl, ini, mn); o = orc.Extractor(nf, sf, nl, ini, mn)
    for idx in (10, 11):
        img = synth.make_image(w, h, idx)
        rk, rd = r.extract(img); ok, od = o.extract(img)
        assert len(rk) == len(ok) and len(rk) > 50
        for f in ("x", "y", "size", "response", "octave"):
            assert np.array_equal(rk[f], ok[f]), f
        assert np.array_equal(rk["angle"].view(np.uint32), ok["angle"].view(np.uint32))
        assert np.array_equal(rd, od)
        for l in range(nl):
            assert np.array_equal(r.level(l), o.level(l))
