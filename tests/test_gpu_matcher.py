"""GPU parity of the ORBmatcher paths against the CPU oracle: match indices and counts identical,
including tie-heavy descriptor sets, occupied keypoints and non-blocking MapPoints."""
import importlib
import numpy as np
import pytest
import orc

synth = importlib.import_module("orb_slam2-chinese-annotation_b200.synth")
views = importlib.import_module("orb_slam2-chinese-annotation_b200.views")
matcher = importlib.import_module("orb_slam2-chinese-annotation_b200.matcher")
capi = importlib.import_module("orb_slam2-chinese-annotation_b200.capi")
pytestmark = pytest.mark.gpu


def _frame(d):
    return views.Frame(d["x"], d["y"], d["octave"], d["angle"], d["desc"], d["sf"], d["bounds"], d["u_right"])


@pytest.fixture(scope="module")
def gm():
    return matcher.ORBmatcher(0.8, True, max_n=4096, max_pairs=8)


def test_descriptor_distance(gm):
    rng = np.random.default_rng(0)
    a = rng.integers(0, 256, (500, 32), dtype=np.uint8); b = rng.integers(0, 256, (500, 32), dtype=np.uint8)
    b[:5] = a[:5]; b[5] = ~a[5]
    d = gm.DescriptorDistance(a, b)
    assert list(d) == [orc.descriptor_distance(a[i], b[i]) for i in range(500)]
    assert d[0] == 0 and d[5] == 256


@pytest.mark.parametrize("na,nb", [(2000, 2000), (1, 1), (257, 63), (130, 1000)])
def test_bruteforce(gm, na, nb):
    A, B, _ = synth.make_descriptor_sets(na, nb, 1)
    gi, gd, gs = gm.match_bruteforce(A, B)
    oi, od, os_ = orc.match_bruteforce(A, B)
    assert np.array_equal(gi, oi) and np.array_equal(gd, od) and np.array_equal(gs, os_)


def test_bruteforce_batch_and_ties(gm):
    rng = np.random.default_rng(2)
    A = np.repeat(rng.integers(0, 2, (4, 300, 4), dtype=np.uint8) * 255, 8, axis=2)
    B = np.repeat(rng.integers(0, 2, (4, 280, 4), dtype=np.uint8) * 255, 8, axis=2)
    gi, gd, gs = gm.match_bruteforce(A, B)
    for p in range(4):
        oi, od, os_ = orc.match_bruteforce(A[p], B[p])
        assert np.array_equal(gi[p], oi) and np.array_equal(gd[p], od) and np.array_equal(gs[p], os_)


def test_grid_assignment(gm):
    for idx, n in [(0, 2000), (1, 1), (2, 777)]:
        d = synth.make_frame_arrays(n, idx)
        d["x"][: n // 10] = np.float32(d["bounds"][2]) - 0.2   # round() lands on column 64 -> dropped
        f = _frame(d)
        gs, gi = gm.assign_features_to_grid(f)
        os_, oi = orc.assign_features_to_grid(f)
        assert np.array_equal(gs, os_) and np.array_equal(gi, oi)


def test_projection_search_that_outgrows_the_candidate_buffer():
    """A single-pair search queues the candidate fill and the resolve against the buffer's present capacity and learns the lists' size
    with the results; lists that do not fit make it grow the buffer and queue both again.  A handle for 2000 features starts with
    128 k candidate slots; a wide search window needs more."""
    small = matcher.ORBmatcher(0.8, True, max_n=2000, max_pairs=1)
    f, mps, occ = synth.make_projection_scenario(2000, 2000, 9)
    F = _frame(f)
    for th in (3.0, 80.0, 3.0):   # fits, outgrows (twice the work), fits again in the grown buffer
        n_g, m_g = small.SearchByProjection(F, mps, th)
        n_o, m_o = orc.search_by_projection_mps(F, mps, th, small.mfNNratio, None)
        assert n_g == n_o and np.array_equal(m_g, m_o), th
        if th > 10:
            assert capi.lib().orb_matcher_last_candidates(small._h) > 2000 * 64
    small.close()


def test_grid_assignment_between_4096_and_8192_keypoints():
    """8192 sort keys (32 KB of dynamic shared memory) beside the kernel's 24 KB of static tables pass the 48 KB a launch gets
    without the opt-in attribute: the launch was refused for these sizes."""
    big = matcher.ORBmatcher(0.8, True, max_n=8192)
    for idx, n in [(3, 4097), (4, 6000), (5, 8192)]:
        f = _frame(synth.make_frame_arrays(n, idx))
        gs, gi = big.assign_features_to_grid(f)
        os_, oi = orc.assign_features_to_grid(f)
        assert np.array_equal(gs, os_) and np.array_equal(gi, oi)
    f, mps, occ = synth.make_projection_scenario(6000, 6000, 7)
    F = _frame(f)
    n_g, m_g = big.SearchByProjection(F, mps, 3.0)
    n_o, m_o = orc.search_by_projection_mps(F, mps, 3.0, big.mfNNratio, None)
    assert n_g == n_o and np.array_equal(m_g, m_o) and n_o > 100
    big.close()


@pytest.mark.parametrize("idx,th,coarse", [(0, 1.0, False), (1, 3.0, False), (2, 5.0, False), (3, 3.0, True), (4, 1.0, True)])
def test_search_by_projection_mappoints(gm, idx, th, coarse):
    f, mps, occ = synth.make_projection_scenario(2000, 2000, idx, coarse_desc=coarse)
    F = _frame(f)
    for use_occ in (False, True):
        n_g, m_g = gm.SearchByProjection(F, mps, th, occupied=occ if use_occ else None)
        n_o, m_o = orc.search_by_projection_mps(F, mps, th, gm.mfNNratio, occ if use_occ else None)
        assert n_g == n_o and np.array_equal(m_g, m_o), (idx, th, n_g, n_o, int((m_g != m_o).sum()))
    assert n_o > 100


@pytest.mark.parametrize("idx,th,fwd,bwd", [(0, 7.0, False, False), (1, 15.0, False, False), (2, 7.0, True, False), (3, 15.0, False, True)])
def test_search_by_projection_lastframe(gm, idx, th, fwd, bwd):
    f, last, cam = synth.make_lastframe_scenario(2000, idx)
    F = _frame(f)
    for check_ori in (True, False):
        gm.mbCheckOrientation = check_ori
        n_g, m_g = gm.SearchByProjectionLastFrame(F, last, cam, th, fwd, bwd)
        n_o, m_o = orc.search_by_projection_last(F, last, cam, th, fwd, bwd, check_ori)
        assert n_g == n_o and np.array_equal(m_g, m_o), (idx, n_g, n_o, int((m_g != m_o).sum()))
    gm.mbCheckOrientation = True
    assert n_o > 100


@pytest.mark.parametrize("idx,coarse,ratio", [(0, False, 0.7), (1, False, 0.75), (2, True, 0.9), (3, False, 0.6)])
def test_search_by_bow(gm, idx, coarse, ratio):
    a, fva, b, fvb, has_a, _ = synth.make_bow_scenario(2000, 1900, idx, coarse_desc=coarse)
    KF, F = _frame(a), _frame(b)
    va, vb = views.FeatureVector(fva), views.FeatureVector(fvb)
    gm.mfNNratio = ratio
    n_g, m_g = gm.SearchByBoW(KF, va, has_a, F, vb)
    n_o, m_o = orc.search_by_bow(KF, va, has_a, F, vb, ratio, True)
    gm.mfNNratio = 0.8
    assert n_g == n_o and np.array_equal(m_g, m_o), (idx, n_g, n_o)
    if not coarse:
        assert n_o > 50


@pytest.mark.parametrize("idx,only_stereo,coarse", [(0, False, False), (1, True, False), (2, False, True)])
def test_search_for_triangulation(gm, idx, only_stereo, coarse):
    a, fva, b, fvb, has_a, has_b = synth.make_bow_scenario(2000, 2000, 10 + idx, coarse_desc=coarse)
    KF1, KF2 = _frame(a), _frame(b)
    va, vb = views.FeatureVector(fva), views.FeatureVector(fvb)
    # pure x-translation: F12 = [t]_x up to scale -> epipolar lines are horizontal (y2 = y1)
    F12 = np.array([[0, 0, 0], [0, 0, -1], [0, 1, 0]], np.float32)
    sigma2 = (a["sf"] * a["sf"]).astype(np.float32)
    n_g, pairs = gm.SearchForTriangulation(KF1, va, has_a, KF2, vb, has_b, F12, (2000.0, 188.0), sigma2, only_stereo)
    n_o, m_o = orc.search_for_triangulation(KF1, va, has_a, KF2, vb, has_b, F12, (2000.0, 188.0), sigma2, only_stereo, True)
    assert n_g == n_o, (n_g, n_o)
    assert pairs == [(int(i), int(m_o[i])) for i in np.nonzero(m_o >= 0)[0]]
    if not coarse:
        assert n_o > 20


@pytest.mark.parametrize("idx,ratio,coarse", [(0, 0.75, False), (1, 0.6, False), (2, 0.9, True)])
def test_search_by_bow_keyframes(gm, idx, ratio, coarse):
    """SearchByBoW(KF, KF) (reference src/ORBmatcher.cc:581-716, LoopClosing.cc:287)."""
    a, fva, b, fvb, _, _ = synth.make_bow_scenario(2000, 1950, 20 + idx, coarse_desc=coarse)
    rng = np.random.default_rng(idx)
    has_a = rng.random(2000) < 0.7; has_b = rng.random(1950) < 0.7
    KF1, KF2 = _frame(a), _frame(b)
    va, vb = views.FeatureVector(fva), views.FeatureVector(fvb)
    for check_ori in (True, False):
        gm.mfNNratio = ratio; gm.mbCheckOrientation = check_ori
        n_g, m_g = gm.SearchByBoWKeyFrames(KF1, va, has_a, KF2, vb, has_b)
        n_o, m_o = orc.search_by_bow_keyframes(KF1, va, has_a, KF2, vb, has_b, ratio, check_ori)
        assert n_g == n_o and np.array_equal(m_g, m_o), (idx, n_g, n_o)
    gm.mfNNratio = 0.8; gm.mbCheckOrientation = True
    if not coarse:
        assert n_o > 50


@pytest.mark.parametrize("idx,th,span,max_dist,ori,coarse", [
    (0, 10.0, (1, 1), 100, True, False),    # relocalisation: SearchByProjection(F, KF, found, 10, 100), levels pred-1..pred+1
    (1, 3.0, (1, 1), 64, True, False),      # relocalisation refinement (3, 64)
    (2, 10.0, (1, 0), 50, False, False),    # loop closing: SearchByProjection(KF, Scw, points, matched, 10), levels pred-1..pred
    (3, 10.0, (1, 1), 100, True, True),     # tie-heavy descriptors
])
def test_search_by_projection_claiming(gm, idx, th, span, max_dist, ori, coarse):
    f, pts, occ = synth.make_projected_points_scenario(2000, 2000, idx, th=th, level_span=span, coarse_desc=coarse)
    F = _frame(f)
    for use_occ in (True, False):
        n_g, m_g, b_g = gm.SearchByProjectionClaiming(F, pts, occ if use_occ else None, max_dist, check_ori=ori)
        n_o, m_o, b_o = orc.search_by_projection_claiming(F, pts, occ if use_occ else None, max_dist, ori)
        assert n_g == n_o and np.array_equal(m_g, m_o) and np.array_equal(b_g, b_o), (idx, n_g, n_o, int((m_g != m_o).sum()))
    assert n_o > 100


@pytest.mark.parametrize("idx,th,chi2,coarse", [(0, 3.0, True, False), (1, 4.0, False, False), (2, 3.0, True, True), (3, 20.0, True, False)])
def test_fuse_candidates(gm, idx, th, chi2, coarse):
    """Search step of Fuse(KF, MapPoints, th) (chi-square gate) and Fuse(KF, Scw, ...) (reference :903-1207)."""
    f, pts, _ = synth.make_projected_points_scenario(2000, 2500, 10 + idx, th=th, level_span=(1, 0), coarse_desc=coarse)
    KF = _frame(f)
    inv_s2 = (1.0 / (f["sf"] * f["sf"])).astype(np.float32) if chi2 else None
    n_g, i_g, d_g = gm.FuseCandidates(KF, pts, inv_s2)
    n_o, i_o, d_o = orc.search_projected_best(KF, pts, inv_s2, 50)
    assert n_g == n_o and np.array_equal(i_g, i_o) and np.array_equal(d_g, d_o), (idx, n_g, n_o, int((i_g != i_o).sum()))
    assert n_o > 100


@pytest.mark.parametrize("idx", [0, 1])
def test_search_by_sim3(gm, idx):
    a, p12, b, p21 = synth.make_sim3_scenario(1500, 1400, idx)
    KF1, KF2 = _frame(a), _frame(b)
    n_g, m_g = gm.SearchBySim3(KF1, p12, KF2, p21)
    n_o, m_o = orc.search_by_sim3(KF1, p12, KF2, p21)
    assert n_g == n_o and np.array_equal(m_g, m_o), (n_g, n_o)
    assert n_o > 300


@pytest.mark.parametrize("idx,window,ratio,coarse", [(0, 100, 0.9, False), (1, 30, 0.9, False), (2, 100, 0.9, True), (3, 100, 0.6, False)])
def test_search_for_initialization(gm, idx, window, ratio, coarse):
    a, b, px, py = synth.make_initialization_scenario(2000, 1900, idx, coarse_desc=coarse)
    F1, F2 = _frame(a), _frame(b)
    for check_ori in (True, False):
        gm.mfNNratio = ratio; gm.mbCheckOrientation = check_ori
        gx, gy = px.copy(), py.copy()
        n_g, m_g = gm.SearchForInitialization(F1, F2, gx, gy, window)
        n_o, m_o = orc.search_for_initialization(F1, F2, px, py, window, ratio, check_ori)
        assert n_g == n_o and np.array_equal(m_g, m_o), (idx, n_g, n_o, int((m_g != m_o).sum()))
        ok = m_o >= 0
        assert np.array_equal(gx[ok], F2.x[m_o[ok]]) and np.array_equal(gx[~ok], px[~ok])
    gm.mfNNratio = 0.8; gm.mbCheckOrientation = True
    if not coarse:
        assert n_o > 100


def _repeat_points(d, k, n_keep):
    """Every point k times in a row: neighbours in the visiting order compete for exactly the same keypoints, which is the case
    the group-optimistic resolve kernel has to replay serially."""
    out = {}
    for key, v in d.items():
        out[key] = np.repeat(v[:n_keep], k, axis=0) if isinstance(v, np.ndarray) and len(v) >= n_keep else v
    return out


@pytest.mark.parametrize("idx,k", [(0, 2), (1, 3), (3, 4)])
def test_projection_searches_with_competing_neighbours(gm, idx, k):
    """SearchByProjection(F, MapPoints), (Cur, Last) and the claiming variant when consecutive points want the same keypoint."""
    f, mps, occ = synth.make_projection_scenario(2000, 2000, idx, coarse_desc=(idx == 3))
    F = _frame(f)
    mpk = _repeat_points(mps, k, 2000 // k)
    mpk["blocks"] = (np.arange(len(mpk["blocks"])) % 3 != 0)          # mix of closing and non-closing points
    n_g, m_g = gm.SearchByProjection(F, mpk, 3.0, occupied=occ)
    n_o, m_o = orc.search_by_projection_mps(F, mpk, 3.0, gm.mfNNratio, occ)
    assert n_g == n_o and np.array_equal(m_g, m_o), (idx, k, n_g, n_o, int((m_g != m_o).sum()))
    assert n_o > 100
    f2, last, cam = synth.make_lastframe_scenario(2000, idx)
    F2 = _frame(f2)
    lk = _repeat_points(last, k, 2000 // k)
    n_g, m_g = gm.SearchByProjectionLastFrame(F2, lk, cam, 15.0, False, False)
    n_o, m_o = orc.search_by_projection_last(F2, lk, cam, 15.0, False, False, gm.mbCheckOrientation)
    assert n_g == n_o and np.array_equal(m_g, m_o), (idx, k, n_g, n_o)
    f3, pts, occ3 = synth.make_projected_points_scenario(2000, 2000, idx, th=10.0, level_span=(1, 1), coarse_desc=(idx == 3))
    F3 = _frame(f3)
    pk = _repeat_points(pts, k, 2000 // k)
    n_g, m_g, b_g = gm.SearchByProjectionClaiming(F3, pk, occ3, 100, check_ori=True)
    n_o, m_o, b_o = orc.search_by_projection_claiming(F3, pk, occ3, 100, True)
    assert n_g == n_o and np.array_equal(m_g, m_o) and np.array_equal(b_g, b_o), (idx, k, n_g, n_o)


@pytest.mark.parametrize("idx,n_src", [(0, 40), (1, 12), (2, 200)])
def test_projection_search_with_long_claim_chains(gm, idx, n_src):
    """Every MapPoint projects onto one of only n_src keypoints with near-identical descriptors, all points closing: point w's answer
    depends on the claims of MANY earlier points (the parallel fixpoint resolve of the single-frame call needs one round per link of
    the longest chain; the batched one-warp kernel replays serially).  Both must equal the sequential loop of src/ORBmatcher.cc:47-141."""
    f, mps, occ = synth.make_projection_scenario(2000, 2000, 30 + idx, coarse_desc=True)
    F = _frame(f)
    rng = np.random.default_rng(100 + idx)
    src = rng.integers(0, n_src, 2000)                                  # few targets: heavy competition
    mps["proj_x"] = (f["x"][src] + rng.normal(0, 1.0, 2000)).astype(np.float32)
    mps["proj_y"] = (f["y"][src] + rng.normal(0, 1.0, 2000)).astype(np.float32)
    mps["desc"] = f["desc"][src].copy()
    mps["level"] = f["octave"][src].astype(np.int32)
    mps["valid"] = np.ones(2000, bool)
    mps["blocks"] = rng.random(2000) < 0.8
    occ = np.zeros(2000, bool)
    for th in (3.0, 5.0):
        n_g, m_g = gm.SearchByProjection(F, mps, th, occupied=occ)
        n_o, m_o = orc.search_by_projection_mps(F, mps, th, gm.mfNNratio, occ)
        assert n_g == n_o and np.array_equal(m_g, m_o), (idx, th, n_g, n_o, int((m_g != m_o).sum()))
    # the same points through the batched entry point (one resolve warp per pair): identical
    res = gm.SearchByProjectionBatch([F, F], [mps, mps], 3.0, occupied=[occ, occ])
    n_o, m_o = orc.search_by_projection_mps(F, mps, 3.0, gm.mfNNratio, occ)
    for p in range(2):
        assert res[p][0] == n_o and np.array_equal(res[p][1], m_o), (idx, p)
