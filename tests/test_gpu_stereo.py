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


def test_stereo_frames_chunked_equals_stepwise():
    """orb_stereo_frames (chunk-pipelined Frame constructor) == extract_batch x2 + compute_stereo_matches,
    on a batch large enough to be cut into several chunks with a ragged tail."""
    w, h, nf, mbf, fx = 752, 480, 1200, 47.90639384423901, 435.2046959714599
    B = 19
    gl = ext.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=B)
    gr = ext.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=B)
    pool = [synth.make_stereo_pair(w, h, i, fx=fx) for i in range(4)]
    il = np.stack([pool[(i * 3) % 4][0] for i in range(B)]); ir = np.stack([pool[(i * 3) % 4][1] for i in range(B)])
    il[7] = 128; ir[7] = 128  # a flat pair: zero keypoints in the middle of a chunk
    kl, dl, nl = gl.extract_batch(il)
    kr, dr, nr = gr.extract_batch(ir)
    ur, dp, n = stereo.compute_stereo_matches(gl, gr, mbf, fx, 0.0, batch=B)
    o = stereo.stereo_frames(gl, gr, il, ir, mbf, fx)
    assert np.array_equal(o["n_l"], nl) and np.array_equal(o["n_r"], nr) and nl[7] == 0
    for b in range(B):
        assert np.array_equal(o["kps_l"][b][:nl[b]], kl[b][:nl[b]]) and np.array_equal(o["desc_l"][b][:nl[b]], dl[b][:nl[b]])
        assert np.array_equal(o["kps_r"][b][:nr[b]], kr[b][:nr[b]]) and np.array_equal(o["desc_r"][b][:nr[b]], dr[b][:nr[b]])
        assert np.array_equal(o["u_right"][b][:nl[b]], ur[b][:nl[b]]) and np.array_equal(o["depth"][b][:nl[b]], dp[b][:nl[b]])
    # and against the oracle on one pair
    ol = orc.Extractor(nf, 1.2, 8, 20, 7); orr = orc.Extractor(nf, 1.2, 8, 20, 7)
    okl, odl = ol.extract(il[18]); okr, odr = orr.extract(ir[18])
    our, odp, _, _ = orc.stereo_match(ol, orr, okl, odl, okr, odr, mbf, 0.0, fx)
    assert np.array_equal(o["u_right"][18][:nl[18]], our) and np.array_equal(o["desc_l"][18][:nl[18]], odl)
