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

CASES = [  # w, h, nfeatures, mbf, fx, (scaleFactor, nlevels, iniThFAST, minThFAST)   (KITTI00-02.yaml, EuRoC.yaml, two non-default pyramids)
    (1241, 376, 2000, 386.1448, 718.856, (1.2, 8, 20, 7)),
    (752, 480, 1200, 47.90639384423901, 435.2046959714599, (1.2, 8, 20, 7)),
    (640, 480, 900, 40.0, 500.0, (1.1, 5, 12, 5)),       # search radius and window scale follow mvScaleFactors / mvInvScaleFactors
    (752, 480, 1000, 47.90639384423901, 435.2046959714599, (1.2, 12, 20, 7)),
]


@pytest.mark.parametrize("w,h,nf,mbf,fx,cfg", CASES)
def test_stereo_matches_oracle(w, h, nf, mbf, fx, cfg):
    B = 3
    gl = ext.ORBextractor(nf, *cfg, max_width=w, max_height=h, max_batch=B)
    gr = ext.ORBextractor(nf, *cfg, max_width=w, max_height=h, max_batch=B)
    pairs = [synth.make_stereo_pair(w, h, i, fx=fx) for i in range(B)]
    kl, dl, nl = gl.extract_batch(np.stack([p[0] for p in pairs]))
    kr, dr, nr = gr.extract_batch(np.stack([p[1] for p in pairs]))
    ur, dp, n = stereo.compute_stereo_matches(gl, gr, mbf, fx, 0.0, batch=B)
    ol = orc.Extractor(nf, *cfg); orr = orc.Extractor(nf, *cfg)
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


def test_async_enqueue_wait_two_batches_in_flight():
    """orb_stereo_frames_enqueue / _wait: two different batches in flight on one handle pair, waited for in order, each equal to
    the synchronous call on the same images (chunk-granular hand-over between consecutive batches)."""
    w, h, nf, mbf, fx = 752, 480, 1200, 47.90639384423901, 435.2046959714599
    B = 19
    gl = ext.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=B)
    gr = ext.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=B)
    pool = [synth.make_stereo_pair(w, h, i, fx=fx) for i in range(5)]
    batches = []
    for k in range(3):
        il = np.stack([pool[(i * 3 + k) % 5][0] for i in range(B)]); ir = np.stack([pool[(i * 3 + k) % 5][1] for i in range(B)])
        batches.append((np.ascontiguousarray(il), np.ascontiguousarray(ir)))   # (synth's right images are views: the stack is not C-ordered)
    with pytest.raises(ValueError):
        stereo.stereo_frames_enqueue(gl, gr, il, ir[:, :, ::-1], mbf, fx, {})
    want = [{kk: v.copy() for kk, v in stereo.stereo_frames(gl, gr, il, ir, mbf, fx).items()} for il, ir in batches]
    cap = gl.kp_cap
    KP = importlib.import_module("orb_slam2-chinese-annotation_b200.capi").KP_DTYPE
    def bufs():
        return dict(kps_l=np.zeros((B, cap), KP), desc_l=np.zeros((B, cap, 32), np.uint8), n_l=np.zeros(B, np.int32),
                    kps_r=np.zeros((B, cap), KP), desc_r=np.zeros((B, cap, 32), np.uint8), n_r=np.zeros(B, np.int32),
                    u_right=np.zeros((B, cap), np.float32), depth=np.zeros((B, cap), np.float32))
    outs = [bufs() for _ in range(3)]
    t0 = stereo.stereo_frames_enqueue(gl, gr, batches[0][0], batches[0][1], mbf, fx, outs[0])
    t1 = stereo.stereo_frames_enqueue(gl, gr, batches[1][0], batches[1][1], mbf, fx, outs[1])
    capi = importlib.import_module("orb_slam2-chinese-annotation_b200.capi")
    with pytest.raises(capi.OrbError):   # a third batch in flight is refused
        stereo.stereo_frames_enqueue(gl, gr, batches[2][0], batches[2][1], mbf, fx, outs[2])
    with pytest.raises(capi.OrbError):   # tickets are waited for in order
        stereo.stereo_frames_wait(gl, gr, t1, outs[1])
    stereo.stereo_frames_wait(gl, gr, t0, outs[0])
    t2 = stereo.stereo_frames_enqueue(gl, gr, batches[2][0], batches[2][1], mbf, fx, outs[2])
    stereo.stereo_frames_wait(gl, gr, t1, outs[1])
    stereo.stereo_frames_wait(gl, gr, t2, outs[2])
    for k in range(3):
        assert np.array_equal(outs[k]["n_l"], want[k]["n_l"]) and np.array_equal(outs[k]["n_r"], want[k]["n_r"])
        for b in range(B):
            nl, nr = want[k]["n_l"][b], want[k]["n_r"][b]
            for f in ("kps_l", "desc_l", "u_right", "depth"):
                assert np.array_equal(outs[k][f][b][:nl], want[k][f][b][:nl]), (k, b, f)
            for f in ("kps_r", "desc_r"):
                assert np.array_equal(outs[k][f][b][:nr], want[k][f][b][:nr]), (k, b, f)
    # the synchronous entry point still works afterwards
    o = stereo.stereo_frames(gl, gr, batches[0][0], batches[0][1], mbf, fx)
    assert np.array_equal(o["n_l"], want[0]["n_l"])


def test_multi_gpu_front_end_equals_single_handle():
    """orb_multi_stereo_frames: every visible GPU takes a contiguous block of the host batch, in sub-batches of max_batch_per_gpu;
    the gathered outputs equal one handle pair's."""
    w, h, nf, mbf, fx = 640, 480, 1000, 40.0, 517.3
    B = 7
    pool = [synth.make_stereo_pair(w, h, i, fx=fx) for i in range(3)]
    il = np.stack([pool[i % 3][0] for i in range(B)]); ir = np.stack([pool[i % 3][1] for i in range(B)])
    gl = ext.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=B)
    gr = ext.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=B)
    want = stereo.stereo_frames(gl, gr, il, ir, mbf, fx)
    m = stereo.MultiGpuFrontEnd(nf, 1.2, 8, 20, 7, w, h, max_batch_per_gpu=2)
    assert m.n_devices >= 1 and m.kp_cap == gl.kp_cap
    got = m.stereo_frames(il, ir, mbf, fx)
    assert np.array_equal(got["n_l"], want["n_l"]) and np.array_equal(got["n_r"], want["n_r"])
    for b in range(B):
        nl = want["n_l"][b]
        for f in ("kps_l", "desc_l", "u_right", "depth"):
            assert np.array_equal(got[f][b][:nl], want[f][b][:nl]), (b, f)


@pytest.mark.parametrize("B", [1, 2])
def test_stereo_frames_from_strided_and_pinned_views(B):
    """orb_stereo_frames on one-chunk calls: pageable dense (pinned mirrors both ways), pageable ROI views (row-wise CPU copy into the
    mirror), pinned sources and pinned outputs (direct copies) all give the bytes of the stepwise path."""
    import ctypes as C
    import torch
    capi = importlib.import_module("orb_slam2-chinese-annotation_b200.capi")
    w, h, nf, mbf, fx = 640, 360, 800, 40.0, 500.0
    gl = ext.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=B)
    gr = ext.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=B)
    pairs = [synth.make_stereo_pair(w, h, 30 + i, fx=fx) for i in range(B)]
    L = np.ascontiguousarray(np.stack([p[0] for p in pairs])); R = np.ascontiguousarray(np.stack([p[1] for p in pairs]))
    kl, dl, nl = gl.extract_batch(L); kr, dr, nr = gr.extract_batch(R)
    ur, dp, n = stereo.compute_stereo_matches(gl, gr, mbf, fx, 0.0, batch=B)
    assert (ur[0][:n[0]] >= 0).sum() > 50
    cap = gl.kp_cap

    def run(vl, vr, pinned_out):
        mk = (lambda s, dt: torch.zeros(s, dtype=torch.uint8).pin_memory().numpy().view(dt).reshape(s[:-1] + (-1,))) if pinned_out else None
        if pinned_out:
            o = dict(kps_l=mk((B, cap * 28), capi.KP_DTYPE), desc_l=mk((B, cap * 32), np.uint8).reshape(B, cap, 32),
                     kps_r=mk((B, cap * 28), capi.KP_DTYPE), desc_r=mk((B, cap * 32), np.uint8).reshape(B, cap, 32),
                     u_right=mk((B, cap * 4), np.float32), depth=mk((B, cap * 4), np.float32))
        else:
            o = dict(kps_l=np.zeros((B, cap), capi.KP_DTYPE), desc_l=np.zeros((B, cap, 32), np.uint8), kps_r=np.zeros((B, cap), capi.KP_DTYPE),
                     desc_r=np.zeros((B, cap, 32), np.uint8), u_right=np.zeros((B, cap), np.float32), depth=np.zeros((B, cap), np.float32))
        o["n_l"] = np.zeros(B, np.int32); o["n_r"] = np.zeros(B, np.int32)
        assert vl.strides == vr.strides
        capi.check(capi.lib().orb_stereo_frames(gl.handle, gr.handle, C.c_void_p(vl.ctypes.data), C.c_void_p(vr.ctypes.data), B, w, h, vl.strides[1],
                                                C.c_size_t(vl.strides[0]), C.c_float(mbf), C.c_float(0.0), C.c_float(fx), capi.ptr(o["kps_l"]),
                                                capi.ptr(o["desc_l"]), capi.ptr(o["n_l"]), capi.ptr(o["kps_r"]), capi.ptr(o["desc_r"]), capi.ptr(o["n_r"]),
                                                capi.ptr(o["u_right"]), capi.ptr(o["depth"]), cap))
        return o

    bigL = np.full((B, h + 6, w + 40), 9, np.uint8); bigR = bigL.copy()
    bigL[:, 3:3 + h, 17:17 + w] = L; bigR[:, 3:3 + h, 17:17 + w] = R
    pinL = torch.from_numpy(L.copy()).pin_memory().numpy(); pinR = torch.from_numpy(R.copy()).pin_memory().numpy()
    cases = [("pageable dense", L, R, False), ("pageable roi", bigL[:, 3:3 + h, 17:17 + w], bigR[:, 3:3 + h, 17:17 + w], False),
             ("pinned in, pageable out", pinL, pinR, False), ("pageable in, pinned out", L, R, True), ("pinned both", pinL, pinR, True)]
    for name, vl, vr, po in cases:
        for rep in range(2):
            o = run(vl, vr, po)
            assert np.array_equal(o["n_l"], nl) and np.array_equal(o["n_r"], nr), name
            for b in range(B):
                a, c = nl[b], nr[b]
                assert np.array_equal(o["desc_l"][b][:a], dl[b][:a]) and np.array_equal(o["desc_r"][b][:c], dr[b][:c]), (name, b)
                assert np.array_equal(o["kps_l"][b][:a], kl[b][:a]) and np.array_equal(o["kps_r"][b][:c], kr[b][:c]), (name, b)
                assert np.array_equal(o["u_right"][b][:a], ur[b][:a]) and np.array_equal(o["depth"][b][:a], dp[b][:a]), (name, b)
