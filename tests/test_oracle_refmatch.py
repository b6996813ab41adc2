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

    def run(mode, arrays, tmp_path):
        fin, fout = str(tmp_path / (mode + ".in")), str(tmp_path / (mode + ".out"))
        cppio.write(fin, arrays)
        r = subprocess.run([BIN, mode, fin, fout], capture_output=True, text=True, timeout=600, cwd=str(tmp_path))
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
