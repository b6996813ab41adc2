"""Batched grid-guided searches (BASELINE.json configs[3]: many frame pairs per call): the results of ONE library call over a ragged
batch of frame pairs must equal the CPU oracle's pair by pair (SearchByProjection(Frame, MapPoints) src/ORBmatcher.cc:47-133 and
SearchByProjection(Cur, Last) :1460-1619)."""
import importlib
import numpy as np
import pytest

import orc

PKG = "orb_slam2-chinese-annotation_b200"
pytestmark = pytest.mark.gpu
synth = importlib.import_module(PKG + ".synth")
mt = importlib.import_module(PKG + ".matcher")
views = importlib.import_module(PKG + ".views")


def _frame(f):
    return views.Frame(f["x"], f["y"], f["octave"], f["angle"], f["desc"], f["sf"], f["bounds"], f["u_right"])


@pytest.mark.parametrize("th", [1.0, 3.0, 5.0])
def test_mappoints_batch_equals_oracle(th):
    sizes = [(2000, 2000), (1500, 1800), (700, 50), (1, 3), (2000, 1200), (64, 2000), (1999, 1999)]
    scen = [synth.make_projection_scenario(n, m, 10 + i, coarse_desc=(i % 3 == 2)) for i, (n, m) in enumerate(sizes)]
    frames = [_frame(s[0]) for s in scen]
    m = mt.ORBmatcher(0.6, True, max_n=2048, max_pairs=len(scen))
    for use_occ in (False, True):
        occ = [s[2] for s in scen] if use_occ else None
        got = m.SearchByProjectionBatch(frames, [s[1] for s in scen], th, occ)
        for p, s in enumerate(scen):
            n_o, m_o = orc.search_by_projection_mps(frames[p], s[1], th, m.mfNNratio, s[2] if use_occ else None)
            assert got[p][0] == n_o, (p, got[p][0], n_o)
            assert np.array_equal(got[p][1], m_o), p


@pytest.mark.parametrize("th,ori", [(7.0, True), (15.0, True), (15.0, False)])
def test_lastframe_batch_equals_oracle(th, ori):
    ns = [2000, 1300, 2000, 17, 900, 2000]
    scen = [synth.make_lastframe_scenario(n, 20 + i) for i, n in enumerate(ns)]
    cam = scen[0][2]
    curs = [_frame(s[0]) for s in scen]
    fwd = [i % 3 == 1 for i in range(len(ns))]
    bwd = [i % 3 == 2 for i in range(len(ns))]
    m = mt.ORBmatcher(0.9, ori, max_n=2048, max_pairs=len(scen))
    got = m.SearchByProjectionLastFrameBatch(curs, [s[1] for s in scen], cam, th, fwd, bwd)
    for p, s in enumerate(scen):
        n_o, m_o = orc.search_by_projection_last(curs[p], s[1], cam, th, fwd[p], bwd[p], ori)
        assert got[p][0] == n_o, (p, got[p][0], n_o)
        assert np.array_equal(got[p][1], m_o), p


def test_batch_equals_single_calls_and_capacity():
    scen = [synth.make_projection_scenario(800, 900, 40 + i) for i in range(4)]
    frames = [_frame(s[0]) for s in scen]
    m = mt.ORBmatcher(0.6, True, max_n=1024, max_pairs=4)
    got = m.SearchByProjectionBatch(frames, [s[1] for s in scen], 3.0, [s[2] for s in scen])
    for p, s in enumerate(scen):
        n1, m1 = m.SearchByProjection(frames[p], s[1], 3.0, s[2])
        assert got[p][0] == n1 and np.array_equal(got[p][1], m1)
    small = mt.ORBmatcher(0.6, True, max_n=1024, max_pairs=2)
    capi = importlib.import_module(PKG + ".capi")
    with pytest.raises(capi.OrbError):
        small.SearchByProjectionBatch(frames, [s[1] for s in scen], 3.0)
