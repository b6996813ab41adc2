"""The drop-in headers inside the reference's REAL classes.

oracle/_ref/dropin_real_driver (oracle/Makefile target `dropin`) is the reference's own src/Frame.cc, src/KeyFrame.cc,
src/MapPoint.cc, src/Map.cc and src/KeyFrameDatabase.cc, compiled where they lie, with THIS repo's include/ORBextractor.h,
include/ORBmatcher.h and include/ORBVocabulary.h in place of the reference's three headers, patches/*.patch applied to
Frame.cc (ComputeStereoMatches -> orb_stereo_match) and MapPoint.cc (batched descriptor distances), linked with
liborb_b200.so instead of ORBextractor.cc / ORBmatcher.cc / DBoW2's vocabulary.  The scenario driver is the same source file as
oracle/_ref/ref_match_driver (the all-reference build): here every scenario runs through BOTH binaries - the drop-in build on
the GPU, the reference build on the host - and the two output files must be equal byte for byte; the scenario's own
expectations (the oracle's) are then checked on that output as well.  Both binaries are built in the authoring container
(they need /root/reference) and travel to the GPU box prebuilt; skipped when absent."""
import os
import subprocess
import numpy as np
import pytest
import cppio
import test_gpu_cpp_dropin as D
import test_oracle_refmatch as R

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REAL = os.path.join(ROOT, "oracle", "_ref", "dropin_real_driver")
PERPAIR = os.path.join(ROOT, "oracle", "_ref", "dropin_real_driver_perpair")
VERBATIM = os.path.join(ROOT, "oracle", "_ref", "ref_match_driver")

# modes whose output holds doubles that the reference's build computes with contracted FMAs (see test_orbvocabulary_score)
TOLERANT = {"bowscore"}


def _run(binary, mode, fin, fout, cwd, extra):
    r = subprocess.run([binary, mode, fin, fout] + list(extra), capture_output=True, text=True, timeout=600, cwd=cwd)
    assert r.returncode == 0, r.stderr
    with open(fout, "rb") as f:
        return f.read()


@pytest.fixture(scope="module")
def dropin():
    if not (os.path.exists(REAL) and os.path.exists(VERBATIM)):
        pytest.skip("oracle/_ref/dropin_real_driver or ref_match_driver not built (needs /root/reference at build time)")
    stats = {"scenarios": 0}

    def run(mode, arrays, tmp_path, *extra):
        fin = str(tmp_path / (mode + ".in"))
        cppio.write(fin, arrays)
        real = _run(REAL, mode, fin, str(tmp_path / (mode + ".real.out")), str(tmp_path), extra)
        ref = _run(VERBATIM, mode, fin, str(tmp_path / (mode + ".ref.out")), str(tmp_path), extra)
        if mode not in TOLERANT:
            assert real == ref, "drop-in build and reference build wrote different %s outputs" % mode
        stats["scenarios"] += 1
        return cppio.read(str(tmp_path / (mode + ".real.out")))
    run.stats = stats
    return run


# every matcher entry point, on the reference's real Frame / KeyFrame / MapPoint objects
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
# Fuse with the reference's real MapPoint::Replace / AddObservation / KeyFrame::ReplaceMapPointMatch replayed by the drop-in
test_fuse = R.test_fuse
# the stereo Frame: drop-in extractors' operator(), the patched Frame::ComputeStereoMatches
test_compute_stereo_matches = R.test_compute_stereo_matches
# Frame::ComputeBoW on the drop-in vocabulary (loadFromTextFile, transform, single-feature transform), score()
test_bow_transform_vs_reference_dbow2 = R.test_bow_transform_vs_reference_dbow2
test_orbvocabulary_transform = D.test_orbvocabulary_transform
test_orbvocabulary_score = D.test_orbvocabulary_score
# MapPoint::ComputeDistinctiveDescriptors through ORBmatcher::DescriptorDistances (patches/MapPoint_*.patch)
test_compute_distinctive_descriptors = R.test_compute_distinctive_descriptors
# host-only members of the real classes keep working next to the drop-in headers
test_grid_and_features_in_area = R.test_grid_and_features_in_area
test_is_in_frustum = R.test_is_in_frustum


@pytest.mark.parametrize("w,h,nf,mbf,fx,seed", [(1241, 376, 2000, 386.1448, 718.856, 0), (752, 480, 1200, 47.90639384423901, 435.2046959714599, 1)])
def test_stereo_frame_constructor(dropin, tmp_path, w, h, nf, mbf, fx, seed):
    """The reference's stereo Frame CONSTRUCTOR (src/Frame.cc:60-126, unmodified but for the patched ComputeStereoMatches) over
    the drop-in extractors: two ExtractORB threads -> operator(), UndistortKeyPoints, ComputeStereoMatches, bounds, grid.  Its
    keypoints / descriptors / mvuRight / mvDepth equal the all-reference build's "stereo" output on the same pair, and the
    lazily fetched mvImagePyramid equals the oracle's pyramid."""
    import importlib
    import orc
    synth = importlib.import_module("orb_slam2-chinese-annotation_b200.synth")
    left, right = synth.make_stereo_pair(w, h, seed, fx=fx)
    arrays = {"params": np.array([w, h, nf, 8, 20, 7], np.int32), "scale": D._f32(1.2), "left": np.ascontiguousarray(left).reshape(-1),
              "right": np.ascontiguousarray(right).reshape(-1), "stereo": np.array([mbf, fx], np.float32)}
    fin = str(tmp_path / "sc.in")
    cppio.write(fin, arrays)
    _run(REAL, "stereo_ctor", fin, str(tmp_path / "ctor.out"), str(tmp_path), ())
    _run(VERBATIM, "stereo", fin, str(tmp_path / "ref.out"), str(tmp_path), ())
    a = cppio.read(str(tmp_path / "ctor.out")); b = cppio.read(str(tmp_path / "ref.out"))
    for k in ("kl", "dl", "kr", "dr", "u_right", "depth", "max_d"):
        assert np.array_equal(a[k], b[k]), k
    assert (a["depth"] > 0).sum() > 50
    o = orc.Extractor(nf, 1.2, 8, 20, 7); o.extract(left)
    pyr = a["pyr"].reshape(-1, 3)
    for l in range(8):
        m = o.level(l)
        flat = m.reshape(-1).astype(np.uint32)
        pw = np.full(len(flat), 31, np.uint32); pw[0] = 1
        pw = np.multiply.accumulate(pw, dtype=np.uint32)[::-1]          # the driver's rolling checksum: sum = sum * 31 + px (mod 2^32)
        acc = (flat * pw).sum(dtype=np.uint32)
        assert (pyr[l][0], pyr[l][1]) == (m.shape[1], m.shape[0]) and np.uint32(pyr[l][2]) == acc, l


def test_single_pair_descriptor_distance_cost(dropin, tmp_path):
    """ORBmatcher::DescriptorDistance per pair (one upload + launch + synchronise each) against one DescriptorDistances call on the
    same pairs: same distances; the measured costs are printed for INTEGRATION.md 3.4 (the batched route is the supported one).
    Also the unpatched MapPoint.cc (per-pair loop) against the patched one on a ComputeDistinctiveDescriptors scenario."""
    rng = np.random.default_rng(5)
    n = 2000
    a = rng.integers(0, 256, (n, 32), dtype=np.uint8); b = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    fin = str(tmp_path / "dd.in")
    cppio.write(fin, {"a": a.reshape(-1), "b": b.reshape(-1)})
    _run(REAL, "dd_cost", fin, str(tmp_path / "dd.out"), str(tmp_path), ())
    out = cppio.read(str(tmp_path / "dd.out"))
    want = np.unpackbits(a ^ b, axis=1).sum(axis=1).astype(np.int32)
    assert np.array_equal(out["one"], want) and np.array_equal(out["batch"], want)
    per_pair_us, batch_us = float(out["us"][0]), float(out["us"][1])
    print("\nDescriptorDistance: %.1f us per single-pair call; DescriptorDistances: %.1f us for %d pairs (%.3f us per pair)"
          % (per_pair_us, batch_us, n, batch_us / n))
    rec = {"single_pair_call_us": per_pair_us, "batched_call_us": batch_us, "pairs": n}
    if os.path.exists(PERPAIR):
        import time
        blob, _ = R._distinctive_blob(0)
        if blob is not None:
            fin2 = str(tmp_path / "dist.in")
            cppio.write(fin2, blob)
            t0 = time.perf_counter(); p = _run(REAL, "distinctive", fin2, str(tmp_path / "d1.out"), str(tmp_path), ()); t1 = time.perf_counter()
            q = _run(PERPAIR, "distinctive", fin2, str(tmp_path / "d2.out"), str(tmp_path), ()); t2 = time.perf_counter()
            assert p == q
            rec.update({"distinctive_batched_s": t1 - t0, "distinctive_per_pair_s": t2 - t1})
            print("ComputeDistinctiveDescriptors scenario: %.2f s batched (patched MapPoint.cc), %.2f s per pair (unpatched), process wall time" % (t1 - t0, t2 - t1))
    out_dir = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out_dir):
        import json
        with open(os.path.join(out_dir, "descriptor_distance_cost.json"), "w") as fjs:
            json.dump(rec, fjs)
    assert batch_us / n < per_pair_us
