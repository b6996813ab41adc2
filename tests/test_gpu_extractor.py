"""GPU parity: the CUDA extractor (through the C ABI) against the CPU oracle, stage by stage and
end to end.  Bit-exact for pyramid, blur, FAST candidates, quadtree selection, keypoints,
descriptors; the angle is compared bit-exactly too (north_star allows 1e-5 relative)."""
import importlib
import numpy as np
import pytest
import orc

pkg = importlib.import_module("orb_slam2-chinese-annotation_b200")
synth = importlib.import_module("orb_slam2-chinese-annotation_b200.synth")
pytestmark = pytest.mark.gpu

CONFIGS = [  # (w, h, nfeatures)  BASELINE.json configs 1, 2, 3 + small/odd shapes
    (640, 480, 1000),
    (1241, 376, 2000),
    (752, 480, 1200),
    (333, 257, 500),
    (200, 120, 300),
    # small levels make single very wide / tall FAST cells (window / floor(window / 30) up to 59 pixels): the per-warp shared memory
    # of k_fast_cells plus its static mbarriers crosses the 48 KB default here (a refused launch once left stale lists behind)
    (257, 193, 1000),
    (179, 134, 1000),
]


def _gpu_extractor(w, h, nf, batch=1):
    ext = importlib.import_module("orb_slam2-chinese-annotation_b200.extractor")
    return ext.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=batch)


def _compare(img, g, o, nlevels=8):
    kps, desc = g(img)
    okps, odesc = o.extract(img)
    for l in range(nlevels):
        assert np.array_equal(g.image_pyramid(l), o.level(l)), "pyramid level %d" % l
        oc = o.level_candidates(l); gc = g.level_candidates(l)
        assert oc.shape == gc.shape and np.array_equal(oc, gc), "FAST candidates level %d (%d vs %d)" % (l, len(gc), len(oc))
        ok = o.level_keypoints(l); gk = g.level_keypoints(l)
        assert len(ok) == len(gk), "quadtree count level %d: %d vs %d" % (l, len(gk), len(ok))
        assert np.array_equal(gk[:, 0], ok["x"].astype(np.int32)) and np.array_equal(gk[:, 1], ok["y"].astype(np.int32)), "quadtree level %d" % l
        ob = o.blurred_level(l)
        if ob is not None:
            assert np.array_equal(g.image_pyramid(l, blurred=True), ob), "blur level %d" % l
    assert len(kps) == len(okps)
    for f in ("x", "y", "size", "response", "octave"):
        assert np.array_equal(kps[f], okps[f]), f
    assert np.all(kps["class_id"] == -1)
    assert np.array_equal(kps["angle"].view(np.uint32), okps["angle"].view(np.uint32)), "angle bits"
    assert np.array_equal(desc, odesc), "descriptors: %d rows differ" % int(np.any(desc != odesc, axis=1).sum())
    return len(kps)


def test_high_resolution_stress_config():
    """BASELINE.json configs[4]: 3840x2160 monocular, nFeatures=8000 (largest level size a packed key holds)."""
    w, h, nf = 3840, 2160, 8000
    g = _gpu_extractor(w, h, nf, batch=2)
    o = orc.Extractor(nf, 1.2, 8, 20, 7)
    img = synth.make_image(w, h, 0)
    n = _compare(img, g, o)
    assert n >= 8000
    # a batch of two distinct 4K frames equals two single calls
    img2 = synth.make_image(w, h, 1)
    kps, desc, cnt = g.extract_batch(np.stack([img2, img]))
    okps, odesc = o.extract(img2)
    assert cnt[0] == len(okps) and np.array_equal(desc[0][:cnt[0]], odesc)
    assert np.array_equal(kps[0][:cnt[0]]["angle"].view(np.uint32), okps["angle"].view(np.uint32))
    assert cnt[1] == n


@pytest.mark.parametrize("w,h,nf", CONFIGS)
def test_extractor_matches_oracle(w, h, nf):
    g = _gpu_extractor(w, h, nf)
    o = orc.Extractor(nf, 1.2, 8, 20, 7)
    for idx in range(3):
        n = _compare(synth.make_image(w, h, idx), g, o)
        assert n > 0


SETTINGS = [  # (w, h, nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST): every reference yaml uses 1.2 / 8 / 20 / 7; the class takes any
    (640, 480, 700, 1.1, 5, 12, 5),
    (640, 480, 400, 1.5, 3, 30, 10),
    (752, 480, 1500, 1.2, 12, 20, 7),
    (500, 375, 600, 1.3, 6, 20, 20),
    (640, 480, 300, 1.2, 1, 40, 7),
]


@pytest.mark.parametrize("w,h,nf,sf,nl,ini,mn", SETTINGS)
def test_other_extractor_settings(w, h, nf, sf, nl, ini, mn):
    ext = importlib.import_module("orb_slam2-chinese-annotation_b200.extractor")
    g = ext.ORBextractor(nf, sf, nl, ini, mn, max_width=w, max_height=h, max_batch=1)
    o = orc.Extractor(nf, sf, nl, ini, mn)
    tg, to = g._tables(), o.tables()
    for k in to:
        assert np.array_equal(tg[k], to[k]), k
    for idx in range(2):
        assert _compare(synth.make_image(w, h, 10 + idx), g, o, nlevels=nl) > 0


def test_tables_match_oracle():
    g = _gpu_extractor(640, 480, 1000)
    o = orc.Extractor(1000, 1.2, 8, 20, 7)
    tg, to = g._tables(), o.tables()
    for k in to:
        assert np.array_equal(tg[k], to[k]), k
    assert list(tg["umax"]) == [15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3]
    assert list(tg["n_per_level"]) == [217, 181, 151, 126, 105, 87, 73, 60]


def test_noise_and_flat_images():
    w, h = 400, 300
    g = _gpu_extractor(w, h, 800)
    o = orc.Extractor(800, 1.2, 8, 20, 7)
    rng = np.random.default_rng(3)
    _compare(rng.integers(0, 256, (h, w), dtype=np.uint8), g, o)          # dense corners: sorted-phase quadtree
    flat = np.full((h, w), 90, np.uint8)
    kps, desc = g(flat)
    assert len(kps) == 0 and desc.shape == (0, 32)
    sparse = flat.copy(); sparse[100:104, 150:154] = 255; sparse[200:203, 50:53] = 0
    _compare(sparse, g, o)                                                   # few corners: quadtree stops without growth
    low = (128 + (rng.integers(0, 256, (h, w)) - 128) // 12).astype(np.uint8)
    _compare(low, g, o)                                                      # only the minThFAST retry fires


def test_empty_image_and_bad_type():
    g = _gpu_extractor(64, 64, 100)
    kps, desc = g(np.empty((0, 0), np.uint8))
    assert len(kps) == 0 and desc.shape == (0, 32)
    with pytest.raises(ValueError):
        g(np.zeros((64, 64), np.float32))


def test_batch_equals_single():
    w, h, nf, B = 752, 480, 1200, 5
    g = _gpu_extractor(w, h, nf, batch=B)
    imgs = np.stack([synth.make_image(w, h, 10 + i) for i in range(B)])
    kps, desc, counts = g.extract_batch(imgs)
    o = orc.Extractor(nf, 1.2, 8, 20, 7)
    for b in range(B):
        ok, od = o.extract(imgs[b])
        assert counts[b] == len(ok)
        for f in ("x", "y", "angle", "response", "octave"):
            assert np.array_equal(kps[b][:counts[b]][f], ok[f])
        assert np.array_equal(desc[b][:counts[b]], od)


def test_streaming_batches_equal_the_synchronous_call():
    """orb_extract_enqueue / _wait: two batches in flight on one handle (three issued back to back, waited for in order) give exactly
    what orb_extract_batch gives for the same images; the synchronous call is refused while a ticket is out."""
    capi = importlib.import_module("orb_slam2-chinese-annotation_b200.capi")
    w, h, nf, B = 640, 480, 1000, 20
    g = _gpu_extractor(w, h, nf, batch=B)
    batches = [np.stack([synth.make_image(w, h, 100 * r + i) for i in range(B)]) for r in range(3)]
    ref = [tuple(a.copy() for a in g.extract_batch(im)) for im in batches]
    t0, o0 = g.extract_enqueue(batches[0])
    t1, o1 = g.extract_enqueue(batches[1])
    with pytest.raises(capi.OrbError):
        g.extract_enqueue(batches[2])                      # a third one has to wait
    with pytest.raises(capi.OrbError):
        g.extract_batch(batches[2])                        # the synchronous call would reuse the slots of the tickets in flight
    got0 = g.extract_wait(t0, o0)
    t2, o2 = g.extract_enqueue(batches[2])
    got1 = g.extract_wait(t1, o1)
    got2 = g.extract_wait(t2, o2)
    for (rk, rd, rc), (gk, gd, gc) in zip(ref, (got0, got1, got2)):
        assert np.array_equal(rc, gc)
        for b in range(B):
            n = rc[b]
            assert np.array_equal(rk[b][:n], gk[b][:n]) and np.array_equal(rd[b][:n], gd[b][:n])
    kps, desc, counts = g.extract_batch(batches[0])        # and the handle is usable synchronously again
    assert np.array_equal(counts, ref[0][2])


def test_bordered_pyramid_matches_reference_layout():
    w, h = 333, 257
    g = _gpu_extractor(w, h, 500)
    img = synth.make_image(w, h, 2)
    g(img)
    o = orc.Extractor(500, 1.2, 8, 20, 7); o.extract(img)
    for l in (0, 3, 7):
        assert np.array_equal(g.image_pyramid(l, border=19), orc.border101(o.level(l), 19))


def _abi_extract(g, imgs_view, cap):
    """orb_extract_batch straight through the C ABI on a (B, H, W) view with arbitrary row / image strides."""
    import ctypes as C
    capi = importlib.import_module("orb_slam2-chinese-annotation_b200.capi")
    B, H, W = imgs_view.shape
    kps = np.zeros((B, cap), capi.KP_DTYPE); desc = np.zeros((B, cap, 32), np.uint8); cnt = np.zeros(B, np.int32)
    capi.check(capi.lib().orb_extract_batch(g._h, C.c_void_p(imgs_view.ctypes.data), B, W, H, imgs_view.strides[1], C.c_size_t(imgs_view.strides[0]),
                                            capi.ptr(kps), capi.ptr(desc), cap, capi.ptr(cnt)))
    return kps, desc, cnt


@pytest.mark.parametrize("B", [1, 3, 12])
def test_strided_pinned_and_pageable_sources_give_the_same_bytes(B):
    """Small calls from pageable memory go through the pinned mirrors (CPU copy in, asynchronous copies, CPU copy out); row-strided
    views (a cv::Mat ROI), pinned sources and larger batches take the other upload paths.  All must equal the dense call; the output
    rows are `cap` apart with cap != the handle's own capacity."""
    import torch
    w, h, nf = 333, 257, 500
    g = _gpu_extractor(w, h, nf, batch=B)
    dense = np.stack([synth.make_image(w, h, 20 + i) for i in range(B)])
    ref_k, ref_d, ref_n = g.extract_batch(dense)
    o = orc.Extractor(nf, 1.2, 8, 20, 7)
    ok0, od0 = o.extract(dense[0])
    assert ref_n[0] == len(ok0) and np.array_equal(ref_d[0][:ref_n[0]], od0)
    cap = g.kp_cap + 37
    # (a) ROI of a larger pageable array: row stride > w, gap between images
    big = np.full((B, h + 9, w + 23), 77, np.uint8)
    big[:, 4:4 + h, 11:11 + w] = dense
    # (b) the same in pinned memory; (c) dense pinned
    pin_big = torch.from_numpy(big.copy()).pin_memory().numpy()
    pin_dense = torch.from_numpy(dense.copy()).pin_memory().numpy()
    for name, view in (("pageable roi", big[:, 4:4 + h, 11:11 + w]), ("pinned roi", pin_big[:, 4:4 + h, 11:11 + w]), ("pinned dense", pin_dense),
                       ("pageable dense", dense)):
        for rep in range(2):   # twice: the mirrors are reused
            k, d, n = _abi_extract(g, view, cap)
            assert np.array_equal(n, ref_n), name
            for b in range(B):
                assert np.array_equal(d[b][:n[b]], ref_d[b][:n[b]]), (name, b)
                assert np.array_equal(k[b][:n[b]], ref_k[b][:n[b]]), (name, b)
    # a capacity below the keypoint count is reported, not overrun
    capi = importlib.import_module("orb_slam2-chinese-annotation_b200.capi")
    with pytest.raises(capi.OrbError) as e:
        _abi_extract(g, dense, int(ref_n.max()) - 5)
    assert e.value.status == capi.ORB_ERR_CAPACITY
    k, d, n = _abi_extract(g, dense, cap)
    assert np.array_equal(n, ref_n) and np.array_equal(d[0][:n[0]], ref_d[0][:n[0]])
