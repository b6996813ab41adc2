"""GPU parity of Frame::ComputeStereoMatches (reference src/Frame.cc:516-704): matched set and
SAD-median cut identical, uRight / depth within 1e-5 relative (north_star), in practice bit-equal."""
import importlib
import numpy as np
import pytest
import orc

synth = importlib.import_module("orb_slam2-chinese-annotation_b200.synth")
ext = importlib.import_module("orb_slam2-chinese-annotation_b200.extractor")
stereo = importlib.import_module("orb_slam2-chinese-annotation_b200.stereo")
pytestmark = pytest.mark.gpu

CASES = [  # w, h, nfeatures, mbf, fx   (KITTI00-02.yaml, EuRoC.yaml)
    (1241, 376, 2000, 386.1448, 718.856),
    (752, 480, 1200, 47.90639384423901, 435.2046959714599),
]


@pytest.mark.parametrize("w,h,nf,mbf,fx", CASES)
def test_stereo_matches_oracle(w, h, nf, mbf, fx):
    B = 3
    gl = ext.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=B)
    gr = ext.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=B)
    pairs = [synth.make_stereo_pair(w, h, i, fx=fx) for i in range(B)]
    kl, dl, nl = gl.extract_batch(np.stack([p[0] for p in pairs]))
    kr, dr, nr = gr.extract_batch(np.stack([p[1] for p in pairs]))
    ur, dp, n = stereo.compute_stereo_matches(gl, gr, mbf, fx, 0.0, batch=B)
    ol = orc.Extractor(nf, 1.2, 8, 20, 7); orr = orc.Extractor(nf, 1.2, 8, 20, 7)
    total = 0
    for b in range(B):
        okl, odl = ol.extract(pairs[b][0]); okr, odr = orr.extract(pairs[b][1])
        assert n[b] == len(okl) == nl[b]
        our, odp, osad, kept = orc.stereo_match(ol, orr, okl, odl, okr, odr, mbf, 0.0, fx)
        g_ur, g_dp = ur[b][:n[b]], dp[b][:n[b]]
        assert np.array_equal(g_ur < 0, our < 0), "matched set differs: %d vs %d" % ((g_ur >= 0).sum(), (our >= 0).sum())
        m = our >= 0
        assert np.allclose(g_ur[m], our[m], rtol=1e-5, atol=0)
        assert np.allclose(g_dp[m], odp[m], rtol=1e-5, atol=0)
        assert np.array_equal(g_ur, our) and np.array_equal(g_dp, odp)  # in fact bit-equal
        total += int(m.sum())
    assert total > 50 * B, "synthetic pairs should yield many stereo matches, got %d" % total
