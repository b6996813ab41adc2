"""ORBVocabulary mirror: Frame::ComputeBoW / KeyFrame::ComputeBoW on the device.

Reference: src/Frame.cc:439-449, src/KeyFrame.cc:60-71, include/ORBVocabulary.h,
Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h (loadFromTextFile :1362-1450, transform :1128-1283).
The tree descent of every descriptor runs in the CUDA library (csrc/bow.cu through the C ABI);
assembling the std::map-shaped BowVector / FeatureVector from the per-feature (word, node) pairs
follows the reference's order of operations on the host, so the doubles are bit-equal.
"""
import ctypes as C
import numpy as np
from . import capi
from .views import FeatureVector

TF_IDF, TF, IDF, BINARY = range(4)        # DBoW2::WeightingType (BowVector.h)
L1_NORM, L2_NORM, CHI_SQUARE, KL, BHATTACHARYYA, DOT_PRODUCT = range(6)   # DBoW2::ScoringType


def parse_text_vocabulary(path):
    """The text format of loadFromTextFile: 'k L scoring weighting', then per node
    'parent isLeaf b0 .. b31 weight'; node ids are line numbers (root = 0)."""
    with open(path) as f:
        k, L, scoring, weighting = [int(t) for t in f.readline().split()[:4]]
        rows = np.loadtxt(f, dtype=np.float64, ndmin=2)
    n = len(rows) + 1
    parent = np.zeros(n, np.int32); leaf = np.zeros(n, np.uint8); desc = np.zeros((n, 32), np.uint8); weight = np.zeros(n, np.float64)
    parent[1:] = rows[:, 0].astype(np.int32); leaf[1:] = rows[:, 1] > 0
    desc[1:] = rows[:, 2:34].astype(np.uint8); weight[1:] = rows[:, 34]
    return dict(k=k, L=L, scoring=scoring, weighting=weighting, parent=parent, is_leaf=leaf, desc=desc, weight=weight)


class ORBVocabulary:
    def __init__(self, max_features=1 << 16, device=0):
        self.h = None
        self.max_features = int(max_features)
        self.device = int(device)

    def loadFromTextFile(self, path):
        return self.load_arrays(**parse_text_vocabulary(path))

    def load_arrays(self, k, L, scoring, weighting, parent, is_leaf, desc, weight):
        self.close()
        self.k, self.L, self.scoring, self.weighting = int(k), int(L), int(scoring), int(weighting)
        parent = np.ascontiguousarray(parent, np.int32); is_leaf = np.ascontiguousarray(is_leaf, np.uint8)
        desc = np.ascontiguousarray(desc, np.uint8)
        leaves = np.nonzero(is_leaf[1:])[0] + 1
        self.word_weight = np.ascontiguousarray(weight, np.float64)[leaves]   # indexed by word id (leaf order)
        h = C.c_void_p()
        capi.check(capi.lib().orb_vocabulary_create(self.L, len(parent), capi.ptr(parent), capi.ptr(is_leaf), capi.ptr(desc),
                                                    self.max_features, self.device, C.byref(h)))
        self.h = h
        return True

    def close(self):
        if self.h is not None:
            capi.lib().orb_vocabulary_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def size(self):
        nw = C.c_int32()
        capi.check(capi.lib().orb_vocabulary_info(self.h, None, C.byref(nw), None))
        return nw.value

    def launch_count(self):
        return capi.lib().orb_vocabulary_launch_count(self.h)

    def descend(self, desc, levelsup=4):
        """Per descriptor: (word id, node id at level L - levelsup)."""
        desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        n = len(desc)
        word = np.empty(max(n, 1), np.int32); node = np.empty(max(n, 1), np.int32)
        capi.check(capi.lib().orb_bow_transform(self.h, capi.ptr(desc), n, int(levelsup), capi.ptr(word), capi.ptr(node)))
        return word[:n], node[:n]

    def score(self, a, b):
        """score(BowVector, BowVector) under the vocabulary's scoring type; a, b = (ids, values) as transform() returns them."""
        return score(a[0], a[1], b[0], b[1], self.scoring)

    def transform(self, desc, levelsup=4):
        """transform(features, BowVector&, FeatureVector&, levelsup): returns (word ids ascending, values float64, FeatureVector)."""
        word, node = self.descend(desc, levelsup)
        return assemble(word, node, self.word_weight, self.weighting, self.scoring)


def score(ids_a, vals_a, ids_b, vals_b, scoring=L1_NORM):
    """ORBVocabulary::score(BowVector, BowVector) (src/KeyFrameDatabase.cc:137,267, src/LoopClosing.cc:145 ->
    Thirdparty/DBoW2/DBoW2/ScoringObject.cpp:21-313): host-side merge of two id-sorted sparse vectors, like the C++ drop-in
    (include/ORBVocabulary.h); terms are summed in ascending id order."""
    import math
    log_eps = math.log(np.finfo(np.float64).eps)
    term = {L1_NORM: lambda v, w: abs(v - w) - abs(v) - abs(w),
            L2_NORM: lambda v, w: v * w, DOT_PRODUCT: lambda v, w: v * w,
            CHI_SQUARE: lambda v, w: v * w / (v + w) if v + w != 0.0 else 0.0,
            KL: lambda v, w: v * math.log(v / w) if v != 0 and w != 0 else 0.0,
            BHATTACHARYYA: lambda v, w: math.sqrt(v * w)}[scoring]
    a = list(zip((int(k) for k in ids_a), (float(v) for v in vals_a))); b = list(zip((int(k) for k in ids_b), (float(v) for v in vals_b)))
    i = j = 0
    acc = 0.0
    while i < len(a) and j < len(b):
        if a[i][0] == b[j][0]:
            acc += term(a[i][1], b[j][1]); i += 1; j += 1
        elif a[i][0] < b[j][0]:
            if scoring == KL: acc += a[i][1] * (math.log(a[i][1]) - log_eps)   # a word only the first vector holds
            i += 1
        else:
            j += 1
    if scoring == KL:
        for k, v in a[i:]:
            if v != 0: acc += v * (math.log(v) - log_eps)
    return {L1_NORM: -acc / 2.0, L2_NORM: 1.0 if acc >= 1 else 1.0 - math.sqrt(max(1.0 - acc, 0.0)), CHI_SQUARE: 2.0 * acc}.get(scoring, acc)


def assemble(word, node, word_weight, weighting=TF_IDF, scoring=L1_NORM):
    """BowVector / FeatureVector from the per-feature descent, in the reference's order of double operations
    (TemplatedVocabulary.h:1146-1207, BowVector.cpp:26-84)."""
    w = word_weight[word] if len(word) else np.zeros(0)
    keep = w > 0                                             # 'not stopped'
    ids, inv = np.unique(word[keep], return_inverse=True)
    vals = np.zeros(len(ids), np.float64)
    if weighting in (TF_IDF, TF):
        np.add.at(vals, inv, w[keep])                        # v[id] += w, feature order
    else:
        first = np.full(len(ids), -1)
        first[inv[::-1]] = np.nonzero(keep)[0][::-1]
        vals = w[first] if len(ids) else vals                # addIfNotExist
    must = scoring in (L1_NORM, L2_NORM, CHI_SQUARE, KL, BHATTACHARYYA)
    norm_l2 = scoring == L2_NORM
    if len(ids) and not must and weighting in (TF_IDF, TF):
        vals = vals / np.float64(len(ids))
    if must and len(ids):
        acc = np.cumsum(vals * vals if norm_l2 else np.abs(vals))[-1]   # sequential accumulation in key order
        norm = np.sqrt(acc) if norm_l2 else acc
        if norm > 0.0:
            vals = vals / norm
    fv = {}
    for i in np.nonzero(keep)[0]:
        fv.setdefault(int(node[i]), []).append(int(i))
    return ids.astype(np.uint32), vals, FeatureVector(fv)
