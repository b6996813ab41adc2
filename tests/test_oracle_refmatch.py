"""Pins the matcher / stereo ORACLE against the REFERENCE ITSELF.

oracle/_ref/ref_match_driver is the reference's own src/ORBmatcher.cc, src/Frame.cc, src/KeyFrame.cc,
src/MapPoint.cc, src/Map.cc, src/KeyFrameDatabase.cc, src/ORBextractor.cc and the vendored DBoW2
compiled verbatim (oracle/Makefile target `refmatch`, against oracle/cvshim) and driven on real
ORB_SLAM2::Frame / KeyFrame / MapPoint objects.  The scenario tests of tests/test_gpu_cpp_dropin.py
(which compare the CUDA drop-in with the oracle) are re-run here with the reference in the drop-in's
place: same scenario files, same expectations, no GPU.  Built only where /root/reference exists;
skipped when the binary is absent."""
import importlib
import os
import subprocess
import numpy as np
import pytest
import cppio
import orc
import worldsim
import test_gpu_cpp_dropin as D

synth = importlib.import_module("orb_slam2-chinese-annotation_b200.synth")
views = importlib.import_module("orb_slam2-chinese-annotation_b200.views")

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
BIN = os.path.join(ROOT, "oracle", "_ref", "ref_match_driver")


@pytest.fixture(scope="module")
def dropin():
    if os.path.exists("/root/reference/src/ORBmatcher.cc"):
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "refmatch"])
    if not os.path.exists(BIN):
        pytest.skip("oracle/_ref/ref_match_driver not built (needs /root/reference)")

    def run(mode, arrays, tmp_path, *extra):
        fin, fout = str(tmp_path / (mode + ".in")), str(tmp_path / (mode + ".out"))
        cppio.write(fin, arrays)
        r = subprocess.run([BIN, mode, fin, fout] + list(extra), capture_output=True, text=True, timeout=600, cwd=str(tmp_path))
        assert r.returncode == 0, r.stderr
        return cppio.read(fout)
    return run


# the drop-in scenario tests, with the reference standing where the CUDA drop-in stands
test_search_by_projection_mappoints = D.test_search_by_projection_mappoints
test_search_by_projection_lastframe = D.test_search_by_projection_lastframe
test_search_by_bow = D.test_search_by_bow
test_search_for_triangulation = D.test_search_for_triangulation
test_search_for_initialization = D.test_search_for_initialization
test_search_by_bow_keyframes = D.test_search_by_bow_keyframes
test_relocalisation_search_by_projection = D.test_relocalisation_search_by_projection
test_loop_search_by_projection = D.test_loop_search_by_projection
test_fuse_sim3 = D.test_fuse_sim3
test_search_by_sim3 = D.test_search_by_sim3
test_orbvocabulary_transform = D.test_orbvocabulary_transform     # ORBVocabulary::loadFromTextFile + transform: the reference's DBoW2
test_orbvocabulary_score = D.test_orbvocabulary_score             # ORBVocabulary::score: the reference's ScoringObject.cpp pins the oracle's merge


REFKP = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4")])


@pytest.mark.parametrize("w,h,nf,mbf,fx,seed", [(1241, 376, 2000, 386.1448, 718.856, 0), (752, 480, 1200, 47.90639384423901, 435.2046959714599, 1),
                                               (640, 480, 1000, 40.0, 517.3, 2)])
def test_compute_stereo_matches(dropin, tmp_path, w, h, nf, mbf, fx, seed):
    """The reference's own Frame::ComputeStereoMatches (src/Frame.cc:516-704, verbatim) on both reference
    extractors' outputs and pyramids == the oracle's stereo port: mvuRight and mvDepth bit-equal."""
    left, right = synth.make_stereo_pair(w, h, seed, fx=fx)
    out = dropin("stereo", {"params": np.array([w, h, nf, 8, 20, 7], np.int32), "scale": D._f32(1.2), "left": left.reshape(-1), "right": right.reshape(-1),
                            "stereo": np.array([mbf, fx], np.float32)}, tmp_path)
    ol = orc.Extractor(nf, 1.2, 8, 20, 7); orr = orc.Extractor(nf, 1.2, 8, 20, 7)
    okl, odl = ol.extract(left); okr, odr = orr.extract(right)
    rkl = out["kl"].view(REFKP); rkr = out["kr"].view(REFKP)
    assert len(rkl) == len(okl) and len(rkr) == len(okr)
    for f in ("x", "y", "angle", "octave"):
        assert np.array_equal(rkl[f], okl[f]) and np.array_equal(rkr[f], okr[f])
    assert np.array_equal(out["dl"].reshape(-1, 32), odl) and np.array_equal(out["dr"].reshape(-1, 32), odr)
    # maxD = mbf / mb with mb = mbf / fx (src/Frame.cc:117,529-531), as computed by the reference in float
    max_d = float(out["max_d"][0])
    assert abs(max_d - np.float32(fx)) <= 2 * np.spacing(np.float32(fx))
    our, odp, _, kept = orc.stereo_match(ol, orr, okl, odl, okr, odr, np.float32(mbf), 0.0, max_d)
    assert np.array_equal(out["u_right"], our) and np.array_equal(out["depth"], odp)
    assert (our >= 0).sum() > 50
    # the C ABI takes max_d = fx: same result on these inputs
    our2, odp2, _, _ = orc.stereo_match(ol, orr, okl, odl, okr, odr, np.float32(mbf), 0.0, np.float32(fx))
    assert np.array_equal(our2, our) and np.array_equal(odp2, odp)


@pytest.mark.parametrize("w,h,nf,mbf,fx,seed,cfg", [(640, 480, 900, 40.0, 500.0, 3, (1.1, 5, 12, 5)),
                                                   (752, 480, 1000, 47.90639384423901, 435.2046959714599, 4, (1.2, 12, 20, 7))])
def test_compute_stereo_matches_at_other_settings(dropin, tmp_path, w, h, nf, mbf, fx, seed, cfg):
    """The same with the non-default pyramids of tests/test_gpu_stereo.py (scale factor 1.1 x 5 levels, 1.2 x 12 levels): the search
    radius and the window scale follow mvScaleFactors / mvInvScaleFactors in the verbatim Frame.cc and in the port alike."""
    sf, nl, ini, mn = cfg
    left, right = synth.make_stereo_pair(w, h, seed, fx=fx)
    left = np.ascontiguousarray(left); right = np.ascontiguousarray(right)
    out = dropin("stereo", {"params": np.array([w, h, nf, nl, ini, mn], np.int32), "scale": D._f32(sf), "left": left.reshape(-1), "right": right.reshape(-1),
                            "stereo": np.array([mbf, fx], np.float32)}, tmp_path)
    ol = orc.Extractor(nf, sf, nl, ini, mn); orr = orc.Extractor(nf, sf, nl, ini, mn)
    okl, odl = ol.extract(left); okr, odr = orr.extract(right)
    rkl = out["kl"].view(REFKP); rkr = out["kr"].view(REFKP)
    assert len(rkl) == len(okl) and len(rkr) == len(okr)
    for f in ("x", "y", "angle", "octave"):
        assert np.array_equal(rkl[f], okl[f]) and np.array_equal(rkr[f], okr[f])
    assert np.array_equal(out["dl"].reshape(-1, 32), odl) and np.array_equal(out["dr"].reshape(-1, 32), odr)
    our, odp, _, _ = orc.stereo_match(ol, orr, okl, odl, okr, odr, np.float32(mbf), 0.0, float(out["max_d"][0]))
    assert np.array_equal(out["u_right"], our) and np.array_equal(out["depth"], odp)
    assert (our >= 0).sum() > 50


@pytest.mark.parametrize("seed", [0, 1])
def test_grid_and_features_in_area(dropin, tmp_path, seed):
    """Frame::AssignFeaturesToGrid / GetFeaturesInArea and KeyFrame::GetFeaturesInArea (verbatim) == oracle grid."""
    f = synth.make_frame_arrays(2000, 700 + seed)
    rng = np.random.default_rng(seed)
    # keypoints on the image border and on exact cell boundaries exercise round() and the dropped last cell
    f["x"][:8] = np.array([0, f["bounds"][2] - 1e-3, f["bounds"][2], 620.5, 9.6953125, 19.390625, 0.5, 1240.9], np.float32)
    f["y"][:8] = np.array([0, f["bounds"][3] - 1e-3, f["bounds"][3], 188.0, 3.9166667, 7.8333335, 375.9, 0.4], np.float32)
    F, blob = D._frame_blob("f_", f)
    Q = 400
    qx = rng.uniform(-30, f["bounds"][2] + 30, Q).astype(np.float32); qy = rng.uniform(-30, f["bounds"][3] + 30, Q).astype(np.float32)
    qr = rng.choice(np.array([2.5, 7.0, 15.0, 40.0, 100.0], np.float32), Q).astype(np.float32)
    qmin = rng.integers(-1, 6, Q).astype(np.int32); qmax = np.where(rng.random(Q) < 0.3, -1, qmin + rng.integers(0, 3, Q)).astype(np.int32)
    blob.update({"qx": qx, "qy": qy, "qr": qr, "qmin": qmin, "qmax": qmax})
    out = dropin("grid", blob, tmp_path)
    cs, ci = orc.assign_features_to_grid(F)
    assert np.array_equal(out["cell_start"], cs) and np.array_equal(out["cell_items"], ci)
    st, it = orc.features_in_area_batch(F, qx, qy, qr, qmin, qmax)
    assert np.array_equal(out["q_start"], st) and np.array_equal(out["q_items"], it)
    st, it = orc.features_in_area_batch(F, qx, qy, qr, np.full(Q, -1), np.full(Q, -1))
    assert np.array_equal(out["k_start"], st) and np.array_equal(out["k_items"], it)
    assert out["q_start"][-1] > 1000


@pytest.mark.parametrize("idx,th", [(0, 3.0), (1, 3.0)])
def test_fuse(dropin, tmp_path, idx, th):
    """Fuse(KF, MapPoints, th) (src/ORBmatcher.cc:903-1054) with the reference's real MapPoint::Replace /
    AddObservation / KeyFrame::ReplaceMapPointMatch (src/MapPoint.cc:97-211) applied by the reference itself:
    the oracle's candidate search + a replay of those map updates gives the same map."""
    f, wd, e_kp, e_nobs, e_bad, in_kf = D._fuse_scenario(idx, 1.0)
    KF, blob = D._world_blob(f, wd, np.zeros(2000, bool), th, False)
    blob.update({"e_kp": e_kp, "e_nobs": e_nobs, "e_bad": views.u8(e_bad), "p_in_kf": in_kf})
    out = dropin("fuse", blob, tmp_path)
    pts = worldsim.gate_keyframe(f, wd, th, sim3=False, skip=wd["null"] | wd["bad"] | (in_kf >= 0), bf=worldsim.CAM[4])
    inv_s2 = (np.float32(1.0) / (f["sf"] * f["sf"]).astype(np.float32)).astype(np.float32)
    _, best, _ = orc.search_projected_best(KF, pts, inv_s2, 50)
    M, E = 2500, 600
    stereo_kp = KF.u_right >= 0 if KF.u_right is not None else np.zeros(2000, bool)
    kf_mp = np.full(2000, -1, np.int64); kf_mp[e_kp] = 100000 + np.arange(E); kf_mp[in_kf[in_kf >= 0]] = np.nonzero(in_kf >= 0)[0]
    p_bad = wd["bad"].copy(); p_idx = in_kf.copy(); p_nobs = wd["nobs"].copy(); p_repl = np.full(M, -1); eb = e_bad.copy(); e_repl = np.full(E, -1)
    n = 0
    for i in range(M):
        if best[i] < 0 or p_bad[i] or p_idx[i] >= 0:
            continue
        slot = best[i]
        cur = kf_mp[slot]
        if cur >= 0:
            cur_bad = eb[cur - 100000] if cur >= 100000 else p_bad[cur]
            cur_nobs = e_nobs[cur - 100000] if cur >= 100000 else p_nobs[cur]
            if not cur_bad:
                if cur_nobs > p_nobs[i]:           # pMP->Replace(pMPinKF): pMP observes nothing yet
                    p_bad[i] = True; p_repl[i] = cur
                elif cur >= 100000:                # pMPinKF->Replace(pMP): scenario 'existing' points carry no observation map
                    eb[cur - 100000] = True; e_repl[cur - 100000] = i
                else:                              # a scenario point sitting in the keyframe hands its slot over
                    p_bad[cur] = True; p_repl[cur] = i
                    s = p_idx[cur]; p_idx[cur] = -1
                    if s >= 0:
                        kf_mp[s] = i; p_idx[i] = s; p_nobs[i] += 2 if stereo_kp[s] else 1
        else:
            p_idx[i] = slot; p_nobs[i] += 2 if stereo_kp[slot] else 1; kf_mp[slot] = i
        n += 1
    assert out["n"][0] == n and n > 100
    assert np.array_equal(out["p_bad"] != 0, p_bad) and np.array_equal(out["p_repl"], p_repl)
    assert np.array_equal(out["e_bad"] != 0, eb) and np.array_equal(out["e_repl"], e_repl)
    assert np.array_equal(out["kf_mp"], kf_mp) and np.array_equal(out["p_idx"], p_idx) and np.array_equal(out["p_nobs"], p_nobs)


def test_shim_algebra_matches_cv2(dropin, tmp_path):
    """oracle/cvshim's cv::Mat products / norms (what the verbatim reference build computes its poses and
    projections with) == the real OpenCV 4.13 (cv2.gemm, cv2.norm) bit for bit, and == the numpy restatement
    tests/worldsim.py uses for the host-side gates."""
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(5)
    n = 4000
    R = rng.standard_normal((n, 3, 3)).astype(np.float32); x = (rng.standard_normal((n, 3)) * 9).astype(np.float32)
    t = rng.standard_normal((n, 3)).astype(np.float32); T4 = rng.standard_normal((n, 4, 4)).astype(np.float32)
    out = dropin("algebra", {"R": R.reshape(-1), "x": x.reshape(-1), "t": t.reshape(-1), "T4": T4.reshape(-1)}, tmp_path)
    rxt = np.stack([cv2.gemm(R[i], x[i].reshape(3, 1), 1.0, t[i].reshape(3, 1), 1.0)[:, 0] for i in range(n)])
    ow = np.stack([cv2.gemm(R[i], t[i].reshape(3, 1), -1.0, None, 0.0, flags=cv2.GEMM_1_T)[:, 0] for i in range(n)])
    rr = np.stack([cv2.gemm(R[i], R[i], 1.0, None, 0.0) for i in range(n)]); tt = np.stack([cv2.gemm(T4[i], T4[i], 1.0, None, 0.0) for i in range(n)])
    nrm = np.array([cv2.norm(x[i].reshape(3, 1)) for i in range(n)]).astype(np.float32)
    assert np.array_equal(out["rxt"].reshape(n, 3), rxt) and np.array_equal(out["ow"].reshape(n, 3), ow)
    assert np.array_equal(out["rr"].reshape(n, 3, 3), rr) and np.array_equal(out["tt"].reshape(n, 4, 4), tt)
    assert np.array_equal(out["norm"], nrm)
    for i in range(0, n, 40):   # the numpy restatement of the gates
        assert np.array_equal(worldsim.mul_row_add(R[i], x[i:i + 1], t[i])[0], rxt[i])
        assert np.array_equal(worldsim.camera_centre(R[i], t[i]), ow[i])
    assert np.array_equal(worldsim.norm3(x), nrm)


bow = importlib.import_module("orb_slam2-chinese-annotation_b200.bow")


@pytest.mark.parametrize("k,L,levelsup,seed,ragged", [(10, 3, 1, 0, True), (10, 4, 2, 1, False), (6, 5, 4, 2, True), (10, 3, 4, 3, True)])
def test_bow_transform_vs_reference_dbow2(dropin, tmp_path, k, L, levelsup, seed, ragged):
    """Frame::ComputeBoW through the reference's own DBoW2 (loadFromTextFile + transform, verbatim) == the oracle's
    tree descent + the host assembly of BowVector / FeatureVector (orb_slam2-chinese-annotation_b200/bow.py), doubles bit-equal."""
    voc = synth.make_vocabulary(k, L, seed, ragged=ragged)
    vf = str(tmp_path / "voc.txt")
    synth.write_text_vocabulary(vf, voc)
    parsed = bow.parse_text_vocabulary(vf)
    for key in ("parent", "is_leaf", "desc", "weight"):
        assert np.array_equal(parsed[key], voc[key]), key
    q = synth.vocabulary_queries(voc, 2000, seed)
    fin, fout = str(tmp_path / "bt.in"), str(tmp_path / "bt.out")
    cppio.write(fin, {"desc": q.reshape(-1), "levelsup": np.array([levelsup], np.int32)})
    r = subprocess.run([BIN, "bowtransform", fin, fout, vf], capture_output=True, text=True, timeout=600, cwd=str(tmp_path))
    assert r.returncode == 0, r.stderr
    out = cppio.read(fout)
    word, node = orc.bow_transform(voc, q, levelsup)
    assert out["info"][0] == voc["is_leaf"].sum() and out["info"][1] == L
    assert np.array_equal(out["word"], word) and np.array_equal(out["node"], node)
    leaves = np.nonzero(voc["is_leaf"])[0]
    assert np.array_equal(out["weight"].view(np.float64), voc["weight"][leaves][word])
    ids, vals, fv = bow.assemble(word, node, voc["weight"][leaves], voc["weighting"], voc["scoring"])
    assert np.array_equal(out["bow_ids"], ids) and np.array_equal(out["bow_vals"].view(np.float64), vals)
    assert np.array_equal(out["fv_ids"], fv.node_ids) and np.array_equal(out["fv_start"], fv.start) and np.array_equal(out["fv_items"], fv.items)
    assert len(ids) > 100 and abs(vals.sum() - 1.0) < 1e-9
    if levelsup < L:
        assert len(fv.node_ids) > 5
    else:
        assert list(fv.node_ids) == [0]          # L - levelsup <= 0: every feature files under the root


@pytest.mark.parametrize("idx,cos_limit", [(0, 0.5), (1, 0.5), (2, 0.8)])
def test_is_in_frustum(dropin, tmp_path, idx, cos_limit):
    """Frame::isInFrustum (src/Frame.cc:303-366, verbatim, built with the reference's -O3 + FMA flags) == the oracle's
    restatement, every float bit-equal (mTrackProjX/Y/XR, mTrackViewCos), and mnTrackScaleLevel == PredictScale(dist)."""
    f, wd = worldsim.frustum_scenario(idx)
    F, blob = D._world_blob(f, wd, np.zeros(2000, bool), 1.0, False)
    blob["cos_limit"] = D._f32(cos_limit)
    out = dropin("frustum", blob, tmp_path)
    R, t, Ow = out["rt_ow"][:9], out["rt_ow"][9:12], out["rt_ow"][12:15]
    assert np.array_equal(Ow, worldsim.camera_centre(wd["pose"][:3, :3], wd["pose"][:3, 3]))
    o = orc.is_in_frustum(R, t, Ow, worldsim.CAM, f["bounds"], cos_limit, wd["xw"], wd["normal"], wd["mind"], wd["maxd"])
    ok = o["in_view"] != 0
    assert np.array_equal(out["in_view"] != 0, ok)
    for k in ("u", "v", "ur", "view_cos"):
        assert np.array_equal(out[k][ok], o[k][ok]), k
    assert np.array_equal(out["level"][ok], worldsim.predict_scale(wd["maxd"][ok], o["dist"][ok], f["sf"]))
    assert 0.3 < ok.mean() < 0.97


def _distinctive_blob(idx):
    kf_desc, start, okf, orow = worldsim.distinctive_scenario(idx)
    f = synth.make_frame_arrays(50, 990)
    F, blob = D._frame_blob("f_", f)
    blob.update({"dims": np.array(kf_desc.shape[:2], np.int32), "kf_desc": kf_desc.reshape(-1), "start": start, "obs_kf": okf, "obs_row": orow})
    return blob, start


@pytest.mark.parametrize("idx", [0, 1])
def test_compute_distinctive_descriptors(dropin, tmp_path, idx):
    """MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:250-326, verbatim) == the oracle on the rows in the
    order the reference gathered them (std::map<KeyFrame*, size_t>: KeyFrame address order)."""
    blob, start = _distinctive_blob(idx)
    out = dropin("distinctive", blob, tmp_path)
    gs, gd = out["g_start"], out["g_desc"].reshape(-1, 32)
    assert np.array_equal(np.diff(gs), np.diff(start))
    best = orc.distinctive_descriptors(gs, gd)
    chosen = out["chosen"].reshape(-1, 32)
    n_obs = np.diff(gs)
    assert np.all(best[n_obs == 0] == -1) and np.all(chosen[n_obs == 0] == 0)
    has = n_obs > 0
    assert np.array_equal(chosen[has], gd[gs[:-1][has] + best[has]])
    assert (best[has] > 0).sum() > 50


def test_soak_reference_vs_oracle_over_seeds(dropin, tmp_path):
    """The scenario tests above on 8 further seeds each (80 scenarios; 200 were run when this was written: 0 disagreements):
    the reference's own matcher and the oracle agree bit for bit on every one."""
    n = 0
    for idx in range(4, 12):
        D.test_search_by_projection_mappoints(dropin, tmp_path, idx, 3.0, idx % 3 == 0)
        D.test_search_by_projection_lastframe(dropin, tmp_path, idx, 7.0 + idx % 9, idx % 2 == 0)
        D.test_search_by_bow(dropin, tmp_path, idx, idx % 2 == 1, 0.7 + 0.01 * (idx % 20))
        D.test_search_for_triangulation(dropin, tmp_path, idx, idx % 2 == 0)
        D.test_search_for_initialization(dropin, tmp_path, idx, 30 + 5 * idx, 0.9, idx % 2 == 0)
        D.test_search_by_bow_keyframes(dropin, tmp_path, idx, 0.6 + 0.01 * idx, idx % 2 == 0)
        D.test_relocalisation_search_by_projection(dropin, tmp_path, idx, 3.0 + idx % 8, 64 + idx, idx % 2 == 0)
        D.test_loop_search_by_projection(dropin, tmp_path, idx, 5 + idx % 6)
        D.test_fuse_sim3(dropin, tmp_path, idx, 4.0)
        D.test_search_by_sim3(dropin, tmp_path, idx, 1.0 + 0.02 * idx)
        n += 10
    assert n == 80
