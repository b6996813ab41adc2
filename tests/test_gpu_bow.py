"""GPU parity of the bag-of-words transform (Frame::ComputeBoW -> DBoW2 transform, reference src/Frame.cc:439-449,
Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1128-1283): word and node ids bit-exact against the oracle, which
tests/test_oracle_refmatch.py pins against the reference's own DBoW2."""
import importlib
import numpy as np
import pytest
import orc

synth = importlib.import_module("orb_slam2-chinese-annotation_b200.synth")
bow = importlib.import_module("orb_slam2-chinese-annotation_b200.bow")
pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("k,L,levelsup,seed,ragged,n", [(10, 3, 1, 0, True, 2000), (10, 4, 2, 1, False, 2000), (6, 5, 4, 2, True, 1200),
                                                        (10, 3, 4, 3, True, 777), (20, 3, 1, 4, True, 2000), (10, 6, 4, 5, False, 8000)])
def test_bow_transform_matches_oracle(k, L, levelsup, seed, ragged, n):
    if L == 6:   # the shape of ORBvoc.txt (10^6 words) is too slow to generate in Python: a depth-6 tree with 4 children
        k = 4
    voc = synth.make_vocabulary(k, L, seed, ragged=ragged)
    v = bow.ORBVocabulary(max_features=8192)
    v.load_arrays(**voc)
    assert v.size() == int(voc["is_leaf"].sum())
    q = synth.vocabulary_queries(voc, n, seed)
    word, node = v.descend(q, levelsup)
    oword, onode = orc.bow_transform(voc, q, levelsup)
    assert np.array_equal(word, oword) and np.array_equal(node, onode)
    ids, vals, fv = v.transform(q, levelsup)
    leaves = np.nonzero(voc["is_leaf"])[0]
    oids, ovals, ofv = bow.assemble(oword, onode, voc["weight"][leaves], voc["weighting"], voc["scoring"])
    assert np.array_equal(ids, oids) and np.array_equal(vals, ovals) and np.array_equal(fv.items, ofv.items)
    assert v.launch_count() == 2


def test_bow_transform_edge_cases():
    voc = synth.make_vocabulary(10, 3, 9)
    v = bow.ORBVocabulary(max_features=64)
    v.load_arrays(**voc)
    w, n = v.descend(np.zeros((0, 32), np.uint8))
    assert len(w) == 0 and len(n) == 0
    q = synth.vocabulary_queries(voc, 65, 9)
    capi = importlib.import_module("orb_slam2-chinese-annotation_b200.capi")
    with pytest.raises(capi.OrbError):
        v.descend(q)                                  # more than max_features
    w1, n1 = v.descend(q[:1], 1)
    ow, on = orc.bow_transform(voc, q[:1], 1)
    assert np.array_equal(w1, ow) and np.array_equal(n1, on)


def test_one_vocabulary_from_four_threads():
    """The reference transforms on ONE shared vocabulary from two threads at once (Frame::ComputeBoW on Tracking, KeyFrame::ComputeBoW on
    LocalMapping).  Four threads, different frames each, 40 calls per thread on one handle (ctypes releases the GIL, the library
    serialises the staged call): every result equals the oracle's for that thread's frame."""
    import threading
    voc = synth.make_vocabulary(10, 4, 21)
    v = bow.ORBVocabulary(max_features=4096)
    v.load_arrays(**voc)
    T = 4
    qs = [synth.vocabulary_queries(voc, 1500 + 100 * t, 30 + t) for t in range(T)]
    want = [orc.bow_transform(voc, q, 4) for q in qs]
    bad = []

    def work(t):
        for _ in range(40):
            w, n = v.descend(qs[t], 4)
            if not (np.array_equal(w, want[t][0]) and np.array_equal(n, want[t][1])):
                bad.append(t)
                return

    ts = [threading.Thread(target=work, args=(t,)) for t in range(T)]
    for th in ts:
        th.start()
    for th in ts:
        th.join()
    assert not bad, bad
