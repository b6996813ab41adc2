"""The C++ drop-in classes (include/ORBextractor.h, include/ORBmatcher.h: the reference's class
names and signatures over the C ABI) driven the way the reference's callers drive them, by
tests/cpp/test_dropin.cpp, and compared with the CPU oracle on the same scenarios."""
import importlib
import os
import subprocess
import numpy as np
import pytest
import cppio
import orc

synth = importlib.import_module("orb_slam2-chinese-annotation_b200.synth")
views = importlib.import_module("orb_slam2-chinese-annotation_b200.views")
pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
BIN = os.path.join(HERE, "cpp", "build", "test_dropin")

# cv::KeyPoint as laid out by include/orb_b200_cv.h (and by OpenCV): pt.x pt.y size angle response octave class_id
KP = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])


@pytest.fixture(scope="module")
def dropin():
    subprocess.check_call(["make", "-s", "-C", os.path.join(HERE, "cpp")])
    assert os.path.exists(BIN)

    def run(mode, arrays, tmp_path, *extra):
        fin, fout = str(tmp_path / (mode + ".in")), str(tmp_path / (mode + ".out"))
        cppio.write(fin, arrays)
        r = subprocess.run([BIN, mode, fin, fout] + list(extra), capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr
        return cppio.read(fout)
    return run


def _frame_blob(p, d):
    f = views.Frame(d["x"], d["y"], d["octave"], d["angle"], d["desc"], d["sf"], d["bounds"], d["u_right"])
    out = {p + "x": f.x, p + "y": f.y, p + "octave": f.octave, p + "angle": f.angle, p + "desc": f.desc.reshape(-1),
           p + "u_right": f.u_right if f.u_right is not None else np.full(len(f.x), -1.0, np.float32), p + "sf": f.sf,
           p + "grid": np.array([f.bounds[0], f.bounds[1], f.bounds[2], f.bounds[3], f.inv_w, f.inv_h], np.float32)}
    return f, out


def _fv_blob(p, mapping):
    v = views.FeatureVector(mapping)
    return v, {p + "fv_ids": v.node_ids, p + "fv_start": v.start, p + "fv_items": v.items}


def _f32(v):
    return np.array([v], np.float32)


@pytest.mark.parametrize("w,h,nf", [(640, 480, 1000), (1241, 376, 2000)])
def test_extractor_operator(dropin, tmp_path, w, h, nf):
    img = synth.make_image(w, h, 11)
    out = dropin("extract", {"params": np.array([w, h, nf, 8, 20, 7], np.int32), "scale": _f32(1.2), "image": img.reshape(-1)}, tmp_path)
    o = orc.Extractor(nf, 1.2, 8, 20, 7)
    okp, odesc = o.extract(img)
    kps = out["kps"].view(KP)
    assert len(kps) == len(okp)
    for f in ("x", "y", "size", "angle", "response", "octave"):
        assert np.array_equal(kps[f], okp[f]), f
    assert (kps["class_id"] == -1).all()
    assert np.array_equal(out["desc"].reshape(-1, 32), odesc)
    assert np.array_equal(out["sf"], o.tables()["sf"])
    lw, lh = out["l3_dims"]
    assert np.array_equal(out["l3"].reshape(lh, lw), o.level(3))      # mvImagePyramid[3] is readable by callers
    assert out["batch_same"][0] == 1 and out["empty_untouched"][0] == 1


@pytest.mark.parametrize("idx,th,coarse", [(0, 1.0, False), (1, 3.0, False), (3, 3.0, True)])
def test_search_by_projection_mappoints(dropin, tmp_path, idx, th, coarse):
    f, mps, occ = synth.make_projection_scenario(2000, 2000, idx, coarse_desc=coarse)
    F, blob = _frame_blob("f_", f)
    blob.update({"mp_valid": views.u8(mps["valid"]), "mp_blocks": views.u8(mps["blocks"]), "mp_proj_x": mps["proj_x"], "mp_proj_y": mps["proj_y"],
                 "mp_proj_xr": mps["proj_xr"], "mp_level": mps["level"], "mp_view_cos": mps["view_cos"], "mp_desc": mps["desc"].reshape(-1),
                 "occupied": views.u8(occ), "ratio": _f32(0.8), "th": _f32(th)})
    out = dropin("mappoints", blob, tmp_path)
    n_o, m_o = orc.search_by_projection_mps(F, mps, th, np.float32(0.8), occ)
    assert out["n"][0] == n_o and np.array_equal(out["match"], m_o)
    assert n_o > 100


@pytest.mark.parametrize("idx,th,check_ori", [(0, 7.0, True), (1, 15.0, True), (2, 7.0, False)])
def test_search_by_projection_lastframe(dropin, tmp_path, idx, th, check_ori):
    f, last, cam = synth.make_lastframe_scenario(2000, idx)
    F, blob = _frame_blob("f_", f)
    blob.update({"l_valid": views.u8(last["valid"]), "l_blocks": views.u8(last["blocks"]), "l_xc": last["xc"], "l_yc": last["yc"], "l_zc": last["zc"],
                 "l_octave": last["octave"], "l_angle": last["angle"], "l_desc": last["desc"].reshape(-1),
                 "cam": np.array(cam, np.float32), "th": _f32(th), "check_ori": np.array([int(check_ori)], np.int32)})
    out = dropin("lastframe", blob, tmp_path)
    n_o, m_o = orc.search_by_projection_last(F, last, cam, th, False, False, check_ori)
    assert out["n"][0] == n_o and np.array_equal(out["match"], m_o)
    assert n_o > 100


@pytest.mark.parametrize("idx,coarse,ratio", [(0, False, 0.7), (2, True, 0.9)])
def test_search_by_bow(dropin, tmp_path, idx, coarse, ratio):
    a, fva, b, fvb, has_a, has_b = synth.make_bow_scenario(2000, 1900, idx, coarse_desc=coarse)
    KF, blob = _frame_blob("a_", a)
    F, bb = _frame_blob("b_", b); blob.update(bb)
    va, ba = _fv_blob("a_", fva); vb, bb = _fv_blob("b_", fvb); blob.update(ba); blob.update(bb)
    blob.update({"has_a": views.u8(has_a), "has_b": views.u8(has_b), "ratio": _f32(ratio)})
    out = dropin("bow", blob, tmp_path)
    n_o, m_o = orc.search_by_bow(KF, va, has_a, F, vb, np.float32(ratio), True)
    assert out["n"][0] == n_o and np.array_equal(out["match"], m_o)


@pytest.mark.parametrize("idx,only_stereo", [(0, False), (1, True)])
def test_search_for_triangulation(dropin, tmp_path, idx, only_stereo):
    a, fva, b, fvb, has_a, has_b = synth.make_bow_scenario(2000, 2000, 10 + idx)
    KF1, blob = _frame_blob("a_", a)
    KF2, bb = _frame_blob("b_", b); blob.update(bb)
    va, ba = _fv_blob("a_", fva); vb, bb = _fv_blob("b_", fvb); blob.update(ba); blob.update(bb)
    F12 = np.array([[0, 0, 0], [0, 0, -1], [0, 1, 0]], np.float32)
    blob.update({"has_a": views.u8(has_a), "has_b": views.u8(has_b), "ratio": _f32(0.6), "epipole": np.array([2000.0, 188.0], np.float32),
                 "f12": F12.reshape(-1), "only_stereo": np.array([int(only_stereo)], np.int32)})
    out = dropin("tri", blob, tmp_path)
    sigma2 = (a["sf"] * a["sf"]).astype(np.float32)
    n_o, m_o = orc.search_for_triangulation(KF1, va, has_a, KF2, vb, has_b, F12, (2000.0, 188.0), sigma2, only_stereo, True)
    assert out["n"][0] == n_o
    assert out["pairs"].reshape(-1, 2).tolist() == [[int(i), int(m_o[i])] for i in np.nonzero(m_o >= 0)[0]]
    assert n_o > 20


@pytest.mark.parametrize("idx,window,ratio,check_ori", [(0, 100, 0.9, True), (1, 30, 0.9, False)])
def test_search_for_initialization(dropin, tmp_path, idx, window, ratio, check_ori):
    a, b, px, py = synth.make_initialization_scenario(2000, 1900, idx)
    F1, blob = _frame_blob("a_", a)
    F2, bb = _frame_blob("b_", b); blob.update(bb)
    blob.update({"prev_x": px, "prev_y": py, "ratio": _f32(ratio), "check_ori": np.array([int(check_ori)], np.int32),
                 "window": np.array([window], np.int32)})
    out = dropin("init", blob, tmp_path)
    n_o, m_o = orc.search_for_initialization(F1, F2, px, py, window, np.float32(ratio), check_ori)
    assert out["n"][0] == n_o and np.array_equal(out["match"], m_o)
    ok = m_o >= 0      # vbPrevMatched refreshed from the matches (reference :571-574)
    assert np.array_equal(out["prev_x"][ok], F2.x[m_o[ok]]) and np.array_equal(out["prev_x"][~ok], px[~ok])
    assert np.array_equal(out["prev_y"][ok], F2.y[m_o[ok]]) and np.array_equal(out["prev_y"][~ok], py[~ok])
    assert n_o > 100


@pytest.mark.parametrize("idx,ratio,check_ori", [(0, 0.75, True), (1, 0.6, False)])
def test_search_by_bow_keyframes(dropin, tmp_path, idx, ratio, check_ori):
    a, fva, b, fvb, has_a, has_b = synth.make_bow_scenario(2000, 1900, 20 + idx)
    KF1, blob = _frame_blob("a_", a)
    KF2, bb = _frame_blob("b_", b); blob.update(bb)
    va, ba = _fv_blob("a_", fva); vb, bb = _fv_blob("b_", fvb); blob.update(ba); blob.update(bb)
    blob.update({"has_a": views.u8(has_a), "has_b": views.u8(has_b), "ratio": _f32(ratio), "check_ori": np.array([int(check_ori)], np.int32)})
    out = dropin("bowkf", blob, tmp_path)
    n_o, m_o = orc.search_by_bow_keyframes(KF1, va, has_a, KF2, vb, has_b, np.float32(ratio), check_ori)
    assert out["n"][0] == n_o and np.array_equal(out["match"], m_o)
    assert n_o > 50


import worldsim


def _world_blob(f, wd, occ, th, check_ori, max_dist=100):
    F, blob = _frame_blob("f_", f)
    blob.update(worldsim.blob(wd))
    blob.update({"pose": wd["pose"].reshape(-1), "cam": np.array(worldsim.CAM, np.float32), "occupied": views.u8(occ), "th": _f32(th),
                 "check_ori": np.array([int(check_ori)], np.int32), "max_dist": np.array([max_dist], np.int32)})
    return F, blob


@pytest.mark.parametrize("idx,th,max_dist,check_ori", [(0, 10.0, 100, True), (1, 3.0, 64, True), (2, 10.0, 100, False)])
def test_relocalisation_search_by_projection(dropin, tmp_path, idx, th, max_dist, check_ori):
    """Tracking::Relocalization's SearchByProjection(F, KF, sFound, th, ORBdist): host gates in the header + claiming search."""
    f = synth.make_frame_arrays(2000, 500 + idx)
    wd = worldsim.make_world(f, idx, 2000, with_null=True, perturb=synth.perturb_descriptors)
    occ = np.random.default_rng(idx).random(2000) < 0.15
    F, blob = _world_blob(f, wd, occ, th, check_ori, max_dist)
    out = dropin("reloc", blob, tmp_path)
    pts = worldsim.gate_relocalisation(f, wd, th)
    n_o, m_o, _ = orc.search_by_projection_claiming(F, pts, occ, max_dist, check_ori)
    assert out["n"][0] == n_o and np.array_equal(out["match"], m_o)
    assert n_o > 100 and 0.5 < pts["valid"].mean() < 0.95


@pytest.mark.parametrize("idx,th", [(0, 10), (1, 5)])
def test_loop_search_by_projection(dropin, tmp_path, idx, th):
    """LoopClosing::ComputeSim3's SearchByProjection(KF, Scw, points, matched, th)."""
    f = synth.make_frame_arrays(2000, 510 + idx)
    wd = worldsim.make_world(f, 10 + idx, 2000, scale=1.0 + 0.3 * idx, perturb=synth.perturb_descriptors)
    occ = np.random.default_rng(idx).random(2000) < 0.15
    assert wd["found"].sum() <= occ.sum()
    F, blob = _world_blob(f, wd, occ, float(th), False)
    out = dropin("loop", blob, tmp_path)
    pts = worldsim.gate_keyframe(f, wd, th, sim3=True, skip=wd["bad"] | wd["found"])
    n_o, m_o, _ = orc.search_by_projection_claiming(F, pts, occ, 50, False)
    assert out["n"][0] == n_o and np.array_equal(out["match"], m_o)
    assert n_o > 100


def _fuse_scenario(idx, scale):
    f = synth.make_frame_arrays(2000, 520 + idx)
    wd = worldsim.make_world(f, 20 + idx, 2500, scale=scale, with_null=(scale == 1.0), perturb=synth.perturb_descriptors)
    rng = np.random.default_rng(77 + idx)
    perm = rng.permutation(2000)
    e_kp = perm[:600].astype(np.int32)                       # keypoints that already hold a MapPoint
    e_nobs = rng.integers(1, 6, 600).astype(np.int32); e_bad = rng.random(600) < 0.1
    in_kf = np.full(2500, -1, np.int32)                      # candidates already observed in the keyframe
    sel = np.nonzero(rng.random(2500) < 0.05)[0]
    in_kf[sel] = perm[600:600 + len(sel)]
    in_kf[wd["null"]] = -1
    return f, wd, e_kp, e_nobs, e_bad, in_kf


@pytest.mark.parametrize("idx,th", [(0, 3.0), (1, 5.0)])
def test_fuse(dropin, tmp_path, idx, th):
    """Fuse(KF, MapPoints, th): gates + chi-square candidate search on the device, map updates replayed in point order."""
    f, wd, e_kp, e_nobs, e_bad, in_kf = _fuse_scenario(idx, 1.0)
    KF, blob = _world_blob(f, wd, np.zeros(2000, bool), th, False)
    blob.update({"e_kp": e_kp, "e_nobs": e_nobs, "e_bad": views.u8(e_bad), "p_in_kf": in_kf})
    out = dropin("fuse", blob, tmp_path)
    pts = worldsim.gate_keyframe(f, wd, th, sim3=False, skip=wd["null"] | wd["bad"] | (in_kf >= 0), bf=worldsim.CAM[4])
    inv_s2 = (np.float32(1.0) / (f["sf"] * f["sf"]).astype(np.float32)).astype(np.float32)
    _, best, _ = orc.search_projected_best(KF, pts, inv_s2, 50)
    M, E = 2500, 600
    kf_mp = np.full(2000, -1, np.int64); kf_mp[e_kp] = 100000 + np.arange(E); kf_mp[in_kf[in_kf >= 0]] = np.nonzero(in_kf >= 0)[0]
    p_bad = wd["bad"].copy(); p_idx = in_kf.copy(); p_nobs = wd["nobs"].copy(); p_repl = np.full(M, -1); eb = e_bad.copy(); e_repl = np.full(E, -1)
    n = 0
    for i in range(M):
        if best[i] < 0 or p_bad[i] or p_idx[i] >= 0:
            continue
        cur = kf_mp[best[i]]
        if cur >= 0:
            cur_bad = eb[cur - 100000] if cur >= 100000 else p_bad[cur]
            cur_nobs = e_nobs[cur - 100000] if cur >= 100000 else p_nobs[cur]
            if not cur_bad:
                if cur_nobs > p_nobs[i]:
                    p_bad[i] = True; p_repl[i] = cur
                elif cur >= 100000:
                    eb[cur - 100000] = True; e_repl[cur - 100000] = i
                else:
                    p_bad[cur] = True; p_repl[cur] = i
        else:
            p_idx[i] = best[i]; p_nobs[i] += 1; kf_mp[best[i]] = i
        n += 1
    assert out["n"][0] == n and n > 100
    assert np.array_equal(out["kf_mp"], kf_mp) and np.array_equal(out["p_idx"], p_idx)
    assert np.array_equal(out["p_bad"] != 0, p_bad) and np.array_equal(out["p_repl"], p_repl)
    assert np.array_equal(out["e_bad"] != 0, eb) and np.array_equal(out["e_repl"], e_repl)
    assert (p_repl >= 0).sum() > 5 and (e_repl >= 0).sum() > 5 and (p_idx >= 0).sum() > (in_kf >= 0).sum() + 50


@pytest.mark.parametrize("idx,th", [(0, 4.0), (1, 4.0)])
def test_fuse_sim3(dropin, tmp_path, idx, th):
    """Fuse(KF, Scw, points, th, vpReplacePoint) of LoopClosing::SearchAndFuse."""
    f, wd, e_kp, e_nobs, e_bad, in_kf = _fuse_scenario(2 + idx, 1.0 + 0.25 * (idx + 1))
    KF, blob = _world_blob(f, wd, np.zeros(2000, bool), th, False)
    blob.update({"e_kp": e_kp, "e_nobs": e_nobs, "e_bad": views.u8(e_bad), "p_in_kf": in_kf})
    out = dropin("fuse_scw", blob, tmp_path)
    pts = worldsim.gate_keyframe(f, wd, th, sim3=True, skip=wd["bad"] | (in_kf >= 0))
    _, best, _ = orc.search_projected_best(KF, pts, None, 50)
    M, E = 2500, 600
    kf_mp = np.full(2000, -1, np.int64); kf_mp[e_kp] = 100000 + np.arange(E); kf_mp[in_kf[in_kf >= 0]] = np.nonzero(in_kf >= 0)[0]
    p_idx = in_kf.copy(); replace = np.full(M, -1); n = 0
    for i in range(M):
        if best[i] < 0:
            continue
        cur = kf_mp[best[i]]
        if cur >= 0:
            cur_bad = e_bad[cur - 100000] if cur >= 100000 else wd["bad"][cur]
            if not cur_bad:
                replace[i] = cur
        else:
            p_idx[i] = best[i]; kf_mp[best[i]] = i
        n += 1
    assert out["n"][0] == n and n > 100
    assert np.array_equal(out["kf_mp"], kf_mp) and np.array_equal(out["p_idx"], p_idx) and np.array_equal(out["replace"], replace)
    assert (replace >= 0).sum() > 20


@pytest.mark.parametrize("idx,s12", [(0, 1.0), (1, 1.1)])
def test_search_by_sim3(dropin, tmp_path, idx, s12):
    """SearchBySim3(KF1, KF2, matches12, s12, R12, t12, th): both projections gated on the host, mutual agreement on the device."""
    n1, n2, th = 1500, 1400, 7.5
    a = synth.make_frame_arrays(n1, 540 + 2 * idx); b = synth.make_frame_arrays(n2, 541 + 2 * idx)
    rng = np.random.default_rng(900 + idx)
    nm = min(n1, n2) * 7 // 10
    ia = rng.permutation(n1)[:nm]; ib = rng.permutation(n2)[:nm]
    part_a = rng.integers(0, n2, n1); part_a[ia] = ib      # partner of each KF1 feature in KF2
    part_b = rng.integers(0, n1, n2); part_b[ib] = ia
    # MapPoints of KF1's features land on their KF2 partners under T2w; those of KF2 on KF1 partners under T1w
    wd1 = worldsim.make_world(b, 30 + 2 * idx, n1, match_frac=0.9, with_null=True, perturb=synth.perturb_descriptors, src=part_a)
    wd2 = worldsim.make_world(a, 31 + 2 * idx, n2, scale=s12, match_frac=0.9, with_null=True, perturb=synth.perturb_descriptors, src=part_b)
    wd2["maxd"] = (wd2["maxd"] * np.float32(s12)).astype(np.float32); wd2["mind"] = (wd2["mind"] * np.float32(s12)).astype(np.float32)
    R12 = wd2["R"] @ wd1["R"].T
    t12 = s12 * wd2["t"] - s12 * R12 @ wd1["t"]
    R12, t12 = R12.astype(np.float32), t12.astype(np.float32)
    pre = np.where(rng.random(n1) < 0.05, rng.integers(0, n2, n1), -1).astype(np.int32)
    A, blob = _frame_blob("a_", a)
    B, bb = _frame_blob("b_", b); blob.update(bb)
    blob.update(worldsim.blob(wd1, "pa_")); blob.update(worldsim.blob(wd2, "pb_"))
    blob.update({"pose_a": wd2["pose"].reshape(-1), "pose_b": wd1["pose"].reshape(-1), "cam": np.array(worldsim.CAM, np.float32), "pre": pre,
                 "r12": R12.reshape(-1), "t12": t12, "s12": _f32(s12), "th": _f32(th)})
    out = dropin("sim3", blob, tmp_path)
    sR12, sR21, t21 = worldsim.sim3_matrices(s12, R12, t12)
    done1 = pre >= 0
    done2 = np.zeros(n2, bool); done2[pre[done1]] = True
    p1 = worldsim.gate_sim3(b, wd1, wd2["pose"][:3, :3], wd2["pose"][:3, 3], sR21, t21, th, wd1["null"] | done1 | wd1["bad"])
    p2 = worldsim.gate_sim3(a, wd2, wd1["pose"][:3, :3], wd1["pose"][:3, 3], sR12, t12, th, wd2["null"] | done2 | wd2["bad"])
    n_o, m_o = orc.search_by_sim3(A, p1, B, p2)
    assert out["n"][0] == n_o
    assert np.array_equal(out["match"], np.where(m_o >= 0, m_o, pre))
    assert n_o > 200


@pytest.mark.parametrize("k,L,levelsup,seed", [(10, 3, 1, 0), (10, 4, 4, 1), (6, 5, 4, 2)])
def test_orbvocabulary_transform(dropin, tmp_path, k, L, levelsup, seed):
    """include/ORBVocabulary.h: loadFromTextFile + transform(features, BowVector, FeatureVector, levelsup) as Frame::ComputeBoW calls it."""
    bow = importlib.import_module("orb_slam2-chinese-annotation_b200.bow")
    voc = synth.make_vocabulary(k, L, seed)
    vf = str(tmp_path / "voc.txt")
    synth.write_text_vocabulary(vf, voc)
    q = synth.vocabulary_queries(voc, 2000, seed)
    out = dropin("bowtransform", {"desc": q.reshape(-1), "levelsup": np.array([levelsup], np.int32)}, tmp_path, vf)
    word, node = orc.bow_transform(voc, q, levelsup)
    leaves = np.nonzero(voc["is_leaf"])[0]
    ids, vals, fv = bow.assemble(word, node, voc["weight"][leaves], voc["weighting"], voc["scoring"])
    assert out["info"][0] == len(leaves) and out["info"][1] == L
    assert np.array_equal(out["bow_ids"], ids) and np.array_equal(out["bow_vals"].view(np.float64), vals)
    assert np.array_equal(out["fv_ids"], fv.node_ids) and np.array_equal(out["fv_start"], fv.start) and np.array_equal(out["fv_items"], fv.items)
    assert len(ids) > 100


@pytest.mark.parametrize("scoring,weighting,seed", [(0, 0, 0), (1, 0, 1), (2, 1, 2), (3, 0, 3), (4, 2, 4), (5, 0, 5), (0, 3, 6)])
def test_orbvocabulary_score(dropin, tmp_path, scoring, weighting, seed):
    """include/ORBVocabulary.h: score(BowVector, BowVector) as KeyFrameDatabase.cc:137,267 / LoopClosing.cc:145 call it, for the six
    DBoW2 scoring types (ScoringObject.cpp): both orders of two overlapping frames and each frame against itself, doubles bit-equal
    to the oracle's merge; the two BowVectors equal the oracle's assembly."""
    bow = importlib.import_module("orb_slam2-chinese-annotation_b200.bow")
    voc = synth.make_vocabulary(9, 4, seed)
    voc["scoring"], voc["weighting"] = scoring, weighting
    vf = str(tmp_path / "voc.txt")
    synth.write_text_vocabulary(vf, voc)
    qa = synth.vocabulary_queries(voc, 700, seed); qb = synth.vocabulary_queries(voc, 600, seed + 50)
    qb[:250] = qa[100:350]     # the two frames share words
    out = dropin("bowscore", {"desc_a": qa.reshape(-1), "desc_b": qb.reshape(-1)}, tmp_path, vf)
    leaves = np.nonzero(voc["is_leaf"])[0]
    vec = []
    for q in (qa, qb):
        word, node = orc.bow_transform(voc, q, 4)
        ids, vals, _ = bow.assemble(word, node, voc["weight"][leaves], weighting, scoring)
        vec.append((ids, vals))
    # The L1 family (ORBvoc.txt is L1 / TF-IDF) has no a * b + c in it and is bit-equal everywhere.  The L2 norm of a BowVector and the
    # L2 / KL / dot-product sums are "acc += a * b": the reference's own build (-O3 -march=native, CMakeLists.txt:10-11) contracts them
    # into FMAs, a build without FMA does not, so those three are compared to 1e-12 relative and the vector to 1 ulp.
    exact = scoring in (0, 2, 4)
    got_vec = [(out["ids_a"], out["vals_a"].view(np.float64)), (out["ids_b"], out["vals_b"].view(np.float64))]
    for g, v in zip(got_vec, vec):
        assert np.array_equal(g[0], v[0])
        if scoring != 1: assert np.array_equal(g[1], v[1])
        else: assert np.allclose(g[1], v[1], rtol=4e-16, atol=0)
    assert len(np.intersect1d(vec[0][0], vec[1][0])) > 50 and len(np.setdiff1d(vec[0][0], vec[1][0])) > 50
    pairs = [(0, 1), (1, 0), (0, 0), (1, 1)]
    want = np.array([orc.bow_score(*got_vec[i], *got_vec[j], scoring) for i, j in pairs], np.float64)
    got = out["scores"].view(np.float64)
    mine = np.array([bow.score(*got_vec[i], *got_vec[j], scoring) for i, j in pairs], np.float64)   # the Python host mirror
    assert np.array_equal(mine.view(np.uint64), want.view(np.uint64))
    if exact:
        assert np.array_equal(got.view(np.uint64), want.view(np.uint64)), (got, want)
    else:
        assert np.allclose(got, want, rtol=1e-12, atol=1e-15), (got, want)
    if scoring == 0:
        assert 0.0 < got[0] < 1.0 and abs(got[2] - 1.0) < 1e-12


REFKP = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4")])


@pytest.mark.parametrize("w,h,nf,mbf,fx,seed", [(1241, 376, 2000, 386.1448, 718.856, 0), (752, 480, 1200, 47.90639384423901, 435.2046959714599, 1)])
@pytest.mark.parametrize("mode", ["stereo", "stereo1"])
def test_stereo_frame_on_two_threads(dropin, tmp_path, w, h, nf, mbf, fx, seed, mode):
    """The stereo Frame constructor's hot path as the reference runs it (src/Frame.cc:81-95): operator() of the left and right
    drop-in extractor on two threads, joined, then ComputeStereoMatches = orb_stereo_match on the two handles; against the oracle.
    stereo1: the same through ORBextractor::ExtractStereo (one orb_stereo_frames call from one thread)."""
    left, right = synth.make_stereo_pair(w, h, seed, fx=fx)
    out = dropin(mode, {"params": np.array([w, h, nf, 8, 20, 7], np.int32), "scale": _f32(1.2), "left": np.ascontiguousarray(left).reshape(-1),
                            "right": np.ascontiguousarray(right).reshape(-1), "stereo": np.array([mbf, fx], np.float32)}, tmp_path)
    ol = orc.Extractor(nf, 1.2, 8, 20, 7); orr = orc.Extractor(nf, 1.2, 8, 20, 7)
    okl, odl = ol.extract(left); okr, odr = orr.extract(right)
    rkl = out["kl"].view(REFKP); rkr = out["kr"].view(REFKP)
    assert len(rkl) == len(okl) and len(rkr) == len(okr)
    for f in ("x", "y", "angle", "response", "octave", "size"):
        assert np.array_equal(rkl[f], okl[f]) and np.array_equal(rkr[f], okr[f]), f
    assert np.array_equal(out["dl"].reshape(-1, 32), odl) and np.array_equal(out["dr"].reshape(-1, 32), odr)
    our, odp, _, _ = orc.stereo_match(ol, orr, okl, odl, okr, odr, np.float32(mbf), 0.0, float(out["max_d"][0]))
    assert np.array_equal(out["u_right"], our) and np.array_equal(out["depth"], odp)
    assert (our >= 0).sum() > 50


def test_soak_dropin_vs_oracle_over_seeds(dropin, tmp_path):
    """Every matcher scenario on 4 further seeds (40 scenarios) through the CUDA drop-in; tests/test_oracle_refmatch.py runs
    the same seeds (and more) through the reference itself."""
    for idx in range(4, 8):
        test_search_by_projection_mappoints(dropin, tmp_path, idx, 3.0, idx % 3 == 0)
        test_search_by_projection_lastframe(dropin, tmp_path, idx, 7.0 + idx % 9, idx % 2 == 0)
        test_search_by_bow(dropin, tmp_path, idx, idx % 2 == 1, 0.7 + 0.01 * (idx % 20))
        test_search_for_triangulation(dropin, tmp_path, idx, idx % 2 == 0)
        test_search_for_initialization(dropin, tmp_path, idx, 30 + 5 * idx, 0.9, idx % 2 == 0)
        test_search_by_bow_keyframes(dropin, tmp_path, idx, 0.6 + 0.01 * idx, idx % 2 == 0)
        test_relocalisation_search_by_projection(dropin, tmp_path, idx, 3.0 + idx % 8, 64 + idx, idx % 2 == 0)
        test_loop_search_by_projection(dropin, tmp_path, idx, 5 + idx % 6)
        test_fuse_sim3(dropin, tmp_path, idx, 4.0)
        test_search_by_sim3(dropin, tmp_path, idx, 1.0 + 0.02 * idx)
