"""GPU parity of the two steps that feed SearchByProjection(Frame, MapPoints): Frame::isInFrustum batched over the
local map (reference src/Frame.cc:303-366) and MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:250-326).
The oracle they are compared with is pinned against the verbatim reference build (tests/test_oracle_refmatch.py)."""
import importlib
import numpy as np
import pytest
import orc
import worldsim

mt = importlib.import_module("orb_slam2-chinese-annotation_b200.matcher")
pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("idx,cos_limit,n", [(0, 0.5, 3000), (1, 0.5, 3000), (2, 0.8, 777), (3, 0.5, 20000)])
def test_is_in_frustum_matches_oracle(idx, cos_limit, n):
    f, wd = worldsim.frustum_scenario(idx, n)
    R, t = wd["pose"][:3, :3], wd["pose"][:3, 3]
    Ow = worldsim.camera_centre(R, t)
    m = mt.ORBmatcher(max_n=max(8192, n))
    g = m.isInFrustum(R, t, Ow, worldsim.CAM, f["bounds"], cos_limit, wd["xw"], wd["normal"], wd["mind"], wd["maxd"])
    o = orc.is_in_frustum(R, t, Ow, worldsim.CAM, f["bounds"], cos_limit, wd["xw"], wd["normal"], wd["mind"], wd["maxd"])
    for k in ("in_view", "u", "v", "ur", "view_cos", "dist"):     # bit-equal, including the values of points that fail a gate
        assert np.array_equal(g[k], o[k]), k
    assert 0.3 < (g["in_view"] != 0).mean() < 0.97
    e = m.isInFrustum(R, t, Ow, worldsim.CAM, f["bounds"], cos_limit, np.zeros((0, 3)), np.zeros((0, 3)), np.zeros(0), np.zeros(0))
    assert len(e["u"]) == 0


@pytest.mark.parametrize("idx,n_kf,n_points", [(0, 24, 400), (1, 24, 400), (2, 150, 300)])
def test_distinctive_descriptors_match_oracle(idx, n_kf, n_points):
    kf_desc, start, okf, orow = worldsim.distinctive_scenario(idx, n_kf=n_kf, n_points=n_points)
    desc = kf_desc[okf, orow]
    m = mt.ORBmatcher(max_n=8192)
    best = m.ComputeDistinctiveDescriptors(start, desc)
    obest = orc.distinctive_descriptors(start, desc)
    assert np.array_equal(best, obest)
    assert (best > 0).sum() > 50 and (np.diff(start) == 0).any() and np.diff(start).max() >= min(n_kf, 100)
