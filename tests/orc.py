"""TEST INFRASTRUCTURE — ctypes binding of the CPU oracle (oracle/liborb_oracle.so).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module; the product package never does.
"""
import ctypes as C
import os
import subprocess
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_SO = os.path.join(ROOT, "oracle", "liborb_oracle.so")

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4")])

_u8p = C.POINTER(C.c_uint8)
_i32p = C.POINTER(C.c_int32)
_f32p = C.POINTER(C.c_float)


def build():
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle")])


def _load():
    if not os.path.exists(_SO):
        build()
    lib = C.CDLL(_SO)
    lib.orc_extractor_create.restype = C.c_void_p
    lib.orc_extractor_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
    lib.orc_extractor_destroy.argtypes = [C.c_void_p]
    lib.orc_extractor_tables.argtypes = [C.c_void_p] + [C.c_void_p] * 6
    lib.orc_extract.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int]
    lib.orc_level.restype = C.c_void_p
    lib.orc_level.argtypes = [C.c_void_p, C.c_int, _i32p, _i32p, _i32p]
    lib.orc_blurred_level.restype = C.c_void_p
    lib.orc_blurred_level.argtypes = [C.c_void_p, C.c_int, _i32p, _i32p, _i32p]
    lib.orc_level_candidates.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
    lib.orc_level_keypoints.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
    lib.orc_fast_atan2.restype = C.c_float
    lib.orc_fast_atan2.argtypes = [C.c_float, C.c_float]
    return lib


lib = _load()


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def resize_linear(src, dw, dh):
    src = np.ascontiguousarray(src, np.uint8)
    dst = np.empty((dh, dw), np.uint8)
    lib.orc_resize_linear_u8(_p(src), src.shape[1], src.shape[0], src.strides[0], _p(dst), dw, dh, dw)
    return dst


def gauss7(src):
    src = np.ascontiguousarray(src, np.uint8)
    dst = np.empty_like(src)
    lib.orc_gauss7_u8(_p(src), src.shape[1], src.shape[0], src.strides[0], _p(dst), dst.strides[0])
    return dst


def border101(src, b):
    src = np.ascontiguousarray(src, np.uint8)
    dst = np.empty((src.shape[0] + 2 * b, src.shape[1] + 2 * b), np.uint8)
    lib.orc_border101(_p(src), src.shape[1], src.shape[0], src.strides[0], _p(dst), dst.strides[0], b)
    return dst


def fast9_16(img, th):
    img = np.ascontiguousarray(img, np.uint8)
    cap = img.size // 4 + 16
    out = np.empty((cap, 3), np.int32)
    n = lib.orc_fast9_16(_p(img), img.shape[1], img.shape[0], img.strides[0], th, _p(out), cap)
    return out[:n].copy()


def fast_score_map(img):
    img = np.ascontiguousarray(img, np.uint8)
    out = np.empty(img.shape, np.int32)
    lib.orc_fast_score_map(_p(img), img.shape[1], img.shape[0], img.strides[0], _p(out))
    return out


def fast_atan2(y, x):
    return lib.orc_fast_atan2(float(y), float(x))


def sincosf(x):
    s = C.c_float(); c = C.c_float()
    lib.orc_sincosf(C.c_float(x), C.byref(s), C.byref(c))
    return s.value, c.value


def distribute(cands, min_x, max_x, min_y, max_y, n_features):
    cands = np.ascontiguousarray(cands, np.int32)
    cap = len(cands) + 8
    out = np.empty(cap, np.int32)
    n = lib.orc_distribute(_p(cands), len(cands), min_x, max_x, min_y, max_y, n_features, _p(out), cap)
    return out[:n].copy()


def distribute_stats(cands, min_x, max_x, min_y, max_y, n_features):
    """distribute() plus the witness of the sorted phase: (kept, [rounds, adjacent equal counts, early breaks between equal counts])."""
    cands = np.ascontiguousarray(cands, np.int32)
    cap = len(cands) + 8
    out = np.empty(cap, np.int32); st = np.zeros(3, np.int32)
    lib.orc_distribute_stats.restype = C.c_int
    n = lib.orc_distribute_stats(_p(cands), len(cands), int(min_x), int(max_x), int(min_y), int(max_y), int(n_features), _p(out), cap, _p(st))
    return out[:n].copy(), st


def descriptor_distance(a, b):
    a = np.ascontiguousarray(a, np.uint8); b = np.ascontiguousarray(b, np.uint8)
    return lib.orc_descriptor_distance(_p(a), _p(b))


class Extractor:
    """Oracle restatement of ORB_SLAM2::ORBextractor (reference src/ORBextractor.cc)."""

    def __init__(self, nfeatures=1000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7):
        self.nlevels = nlevels
        self.h = lib.orc_extractor_create(nfeatures, scale_factor, nlevels, ini_th, min_th)
        assert self.h

    def __del__(self):
        if getattr(self, "h", None):
            lib.orc_extractor_destroy(self.h)
            self.h = None

    def tables(self):
        L = self.nlevels
        sf, inv, s2, inv2 = (np.empty(L, np.float32) for _ in range(4))
        npl = np.empty(L, np.int32); umax = np.empty(16, np.int32)
        lib.orc_extractor_tables(self.h, _p(sf), _p(inv), _p(s2), _p(inv2), _p(npl), _p(umax))
        return dict(sf=sf, inv_sf=inv, sigma2=s2, inv_sigma2=inv2, n_per_level=npl, umax=umax)

    def extract(self, img, cap=None):
        img = np.ascontiguousarray(img, np.uint8)
        self._img = img
        cap = cap or 200000
        kps = np.empty(cap, KP_DTYPE); desc = np.empty((cap, 32), np.uint8)
        if img.size == 0:
            return kps[:0], desc[:0]
        n = lib.orc_extract(self.h, _p(img), img.shape[1], img.shape[0], img.strides[0], _p(kps), _p(desc), cap)
        assert n >= 0, "capacity"
        return kps[:n].copy(), desc[:n].copy()

    def _lvl(self, fn, l):
        w = C.c_int32(); h = C.c_int32(); s = C.c_int32()
        p = fn(self.h, l, C.byref(w), C.byref(h), C.byref(s))
        if not p:
            return None
        buf = (C.c_uint8 * (s.value * h.value)).from_address(p)
        return np.frombuffer(buf, np.uint8).reshape(h.value, s.value)[:, :w.value].copy()

    def level(self, l):
        return self._lvl(lib.orc_level, l)

    def blurred_level(self, l):
        return self._lvl(lib.orc_blurred_level, l)

    def level_candidates(self, l):
        cap = 1 << 20
        out = np.empty((cap, 3), np.int32)
        n = lib.orc_level_candidates(self.h, l, _p(out), cap)
        return out[:n].copy()

    def level_keypoints(self, l):
        cap = 1 << 16
        out = np.empty(cap, KP_DTYPE)
        n = lib.orc_level_keypoints(self.h, l, _p(out), cap)
        return out[:n].copy()


def stereo_match(oL, oR, kpL, dL, kpR, dR, mbf, min_d, max_d):
    """Frame::ComputeStereoMatches on two oracle extractors (their last extract() call)."""
    t = oL.tables()
    nL, nR = len(kpL), len(kpR)
    kpL = np.ascontiguousarray(kpL); kpR = np.ascontiguousarray(kpR)
    dL = np.ascontiguousarray(dL); dR = np.ascontiguousarray(dR)
    ur = np.empty(max(nL, 1), np.float32); dp = np.empty(max(nL, 1), np.float32); sad = np.empty(max(nL, 1), np.int32)
    lib.orc_stereo_match.argtypes = [C.c_void_p] * 4 + [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                                                       C.c_float, C.c_float, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]
    kept = lib.orc_stereo_match(oL.h, oR.h, _p(kpL), _p(dL), nL, _p(kpR), _p(dR), nR, _p(t["sf"]), _p(t["inv_sf"]),
                                mbf, min_d, max_d, _p(ur), _p(dp), _p(sad))
    return ur[:nL], dp[:nL], sad[:nL], kept


def match_bruteforce(a, b):
    a = np.ascontiguousarray(a, np.uint8); b = np.ascontiguousarray(b, np.uint8)
    bi = np.empty(len(a), np.int32); bd = np.empty(len(a), np.int32); sd = np.empty(len(a), np.int32)
    lib.orc_match_bruteforce(_p(a), _p(b), len(a), len(b), _p(bi), _p(bd), _p(sd))
    return bi, bd, sd


# ---- matcher oracle (oracle/match_oracle.cpp) over the same plain view structs as the C ABI
import importlib as _il
_views = _il.import_module("orb_slam2-chinese-annotation_b200.views")
_mt = _il.import_module("orb_slam2-chinese-annotation_b200.matcher")   # only for the struct packers _mp_view/_last_view


def assign_features_to_grid(frame):
    cs = np.empty(64 * 48 + 1, np.int32); ci = np.empty(max(frame.N, 1), np.int32); n = C.c_int32()
    g = frame.view().grid
    lib.orc_assign_features_to_grid(_p(frame.x), _p(frame.y), frame.N, C.byref(g), _p(cs), _p(ci), C.byref(n))
    return cs, ci[:n.value].copy()


def features_in_area_batch(frame, qx, qy, qr, qmin, qmax):
    """GetFeaturesInArea for a batch of windows: (start[nq + 1], items)."""
    qx = np.ascontiguousarray(qx, np.float32); qy = np.ascontiguousarray(qy, np.float32); qr = np.ascontiguousarray(qr, np.float32)
    qmin = np.ascontiguousarray(qmin, np.int32); qmax = np.ascontiguousarray(qmax, np.int32)
    nq = len(qx); cap = max(frame.N, 1) * nq
    st = np.empty(nq + 1, np.int32); it = np.empty(cap, np.int32)
    fv = frame.view()
    lib.orc_features_in_area.argtypes = [C.c_void_p, C.c_int] + [C.c_void_p] * 7 + [C.c_int]
    n = lib.orc_features_in_area(C.byref(fv), nq, _p(qx), _p(qy), _p(qr), _p(qmin), _p(qmax), _p(st), _p(it), cap)
    assert n >= 0
    return st, it[:n].copy()


def search_by_projection_mps(F, mappoints, th, nn_ratio, occupied=None):
    mp, keep = _mt._mp_view(mappoints)
    match = np.empty(max(F.N, 1), np.int32)
    occ = None if occupied is None else _views.u8(occupied)
    fv = F.view()
    lib.orc_search_by_projection_mps.argtypes = [C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_void_p, C.c_void_p]
    n = lib.orc_search_by_projection_mps(C.byref(fv), C.byref(mp), th, nn_ratio, None if occ is None else occ.ctypes.data, _p(match))
    return n, match[:F.N]


def search_by_projection_last(Cur, last, cam, th, forward, backward, check_ori, occupied=None):
    lv, keep = _mt._last_view(last)
    match = np.empty(max(Cur.N, 1), np.int32)
    occ = None if occupied is None else _views.u8(occupied)
    cv = Cur.view(); c = _views.Camera(*[float(v) for v in cam])
    lib.orc_search_by_projection_last.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    n = lib.orc_search_by_projection_last(C.byref(cv), C.byref(lv), C.byref(c), th, int(forward), int(backward), int(check_ori),
                                          None if occ is None else occ.ctypes.data, _p(match))
    return n, match[:Cur.N]


def search_by_bow(KF, kfv, kf_has_mp, F, ffv, nn_ratio, check_ori):
    match = np.empty(max(F.N, 1), np.int32)
    hm = _views.u8(kf_has_mp)
    kv = KF.view(); fv = F.view(); a = kfv.view(); b = ffv.view()
    lib.orc_search_by_bow.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_int, C.c_void_p]
    n = lib.orc_search_by_bow(C.byref(kv), C.byref(a), _p(hm), C.byref(fv), C.byref(b), nn_ratio, int(check_ori), _p(match))
    return n, match[:F.N]


def search_for_triangulation(KF1, fv1, has1, KF2, fv2, has2, F12, epipole, sigma2, only_stereo, check_ori):
    match = np.empty(max(KF1.N, 1), np.int32)
    h1 = _views.u8(has1); h2 = _views.u8(has2)
    f12 = np.ascontiguousarray(F12, np.float32).reshape(9); s2 = np.ascontiguousarray(sigma2, np.float32)
    k1 = KF1.view(); k2 = KF2.view(); a = fv1.view(); b = fv2.view()
    lib.orc_search_for_triangulation.argtypes = [C.c_void_p] * 7 + [C.c_float, C.c_float, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
    n = lib.orc_search_for_triangulation(C.byref(k1), C.byref(a), _p(h1), C.byref(k2), C.byref(b), _p(h2), _p(f12), float(epipole[0]),
                                         float(epipole[1]), _p(s2), int(only_stereo), int(check_ori), _p(match))
    return n, match[:KF1.N]


def _opt(a):
    return None if a is None else a.ctypes.data


def search_by_bow_keyframes(KF1, fv1, has1, KF2, fv2, has2, nn_ratio, check_ori):
    match = np.empty(max(KF1.N, 1), np.int32)
    h1 = _views.u8(has1); h2 = _views.u8(has2)
    k1 = KF1.view(); k2 = KF2.view(); a = fv1.view(); b = fv2.view()
    lib.orc_search_by_bow_keyframes.argtypes = [C.c_void_p] * 6 + [C.c_float, C.c_int, C.c_void_p]
    n = lib.orc_search_by_bow_keyframes(C.byref(k1), C.byref(a), _p(h1), C.byref(k2), C.byref(b), _p(h2), nn_ratio, int(check_ori), _p(match))
    return n, match[:KF1.N]


def search_by_projection_claiming(F, points, occupied, max_dist, check_ori):
    pv, keep = _mt._proj_view(points)
    match = np.empty(max(F.N, 1), np.int32); pbest = np.empty(max(pv.n, 1), np.int32)
    occ = None if occupied is None else _views.u8(occupied)
    fv = F.view()
    lib.orc_search_by_projection_claiming.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    n = lib.orc_search_by_projection_claiming(C.byref(fv), C.byref(pv), _opt(occ), int(max_dist), int(check_ori), _p(match), _p(pbest))
    return n, match[:F.N], pbest[:pv.n]


def search_projected_best(KF, points, inv_level_sigma2, max_dist):
    pv, keep = _mt._proj_view(points)
    bi = np.empty(max(pv.n, 1), np.int32); bd = np.empty(max(pv.n, 1), np.int32)
    s2 = None if inv_level_sigma2 is None else np.ascontiguousarray(inv_level_sigma2, np.float32)
    kv = KF.view()
    lib.orc_search_projected_best.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
    n = lib.orc_search_projected_best(C.byref(kv), C.byref(pv), _opt(s2), int(max_dist), _p(bi), _p(bd))
    return n, bi[:pv.n], bd[:pv.n]


def search_by_sim3(KF1, pts1, KF2, pts2):
    p1, k1_ = _mt._proj_view(pts1); p2, k2_ = _mt._proj_view(pts2)
    match = np.empty(max(KF1.N, 1), np.int32)
    k1 = KF1.view(); k2 = KF2.view()
    lib.orc_search_by_sim3.argtypes = [C.c_void_p] * 5
    n = lib.orc_search_by_sim3(C.byref(k1), C.byref(p1), C.byref(k2), C.byref(p2), _p(match))
    return n, match[:KF1.N]


def search_for_initialization(F1, F2, prev_x, prev_y, window, nn_ratio, check_ori):
    match = np.empty(max(F1.N, 1), np.int32)
    f1 = F1.view(); f2 = F2.view()
    px = np.ascontiguousarray(prev_x, np.float32); py = np.ascontiguousarray(prev_y, np.float32)
    lib.orc_search_for_initialization.argtypes = [C.c_void_p] * 4 + [C.c_int, C.c_float, C.c_int, C.c_void_p]
    n = lib.orc_search_for_initialization(C.byref(f1), C.byref(f2), _p(px), _p(py), int(window), nn_ratio, int(check_ori), _p(match))
    return n, match[:F1.N]


# ---- the reference's own ORBextractor.cc compiled verbatim (oracle/_ref; built only where /root/reference exists)
_REF_SO = os.path.join(ROOT, "oracle", "_ref", "libref_orbextractor.so")


def ref_available():
    return os.path.exists(_REF_SO)


class RefExtractor:
    """ORB_SLAM2::ORBextractor from /root/reference/src/ORBextractor.cc (verbatim) behind oracle/ref_wrap.cpp."""
    _lib = None

    def __init__(self, nfeatures=1000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7):
        if RefExtractor._lib is None:
            L = C.CDLL(_REF_SO)
            L.ref_extractor_create.restype = C.c_void_p
            L.ref_extractor_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
            L.ref_extractor_destroy.argtypes = [C.c_void_p]
            L.ref_extract.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int]
            L.ref_level.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, _i32p, _i32p]
            L.ref_levels.argtypes = [C.c_void_p]
            L.ref_level_ptr.restype = C.c_void_p
            L.ref_level_ptr.argtypes = [C.c_void_p, C.c_int, _i32p, _i32p, _i32p]
            RefExtractor._lib = L
        self.h = RefExtractor._lib.ref_extractor_create(nfeatures, scale_factor, nlevels, ini_th, min_th)

    def __del__(self):
        if getattr(self, "h", None):
            RefExtractor._lib.ref_extractor_destroy(self.h)
            self.h = None

    def extract(self, img, cap=200000):
        img = np.ascontiguousarray(img, np.uint8)
        kps = np.empty(cap, KP_DTYPE); desc = np.empty((cap, 32), np.uint8)
        n = RefExtractor._lib.ref_extract(self.h, _p(img), img.shape[1], img.shape[0], img.strides[0], _p(kps), _p(desc), cap)
        assert n >= 0
        return kps[:n].copy(), desc[:n].copy()

    def level(self, l):
        buf = np.empty(1 << 24, np.uint8); w = C.c_int32(); h = C.c_int32()
        assert RefExtractor._lib.ref_level(self.h, l, _p(buf), buf.size, C.byref(w), C.byref(h)) == 0
        return buf[:w.value * h.value].reshape(h.value, w.value).copy()


    def level_views(self):
        """(ptrs, w, h, stride) ctypes arrays of mvImagePyramid for orc_stereo_match_levels."""
        L = RefExtractor._lib
        n = L.ref_levels(self.h)
        ptrs = (C.c_void_p * n)(); w = (C.c_int32 * n)(); h = (C.c_int32 * n)(); s = (C.c_int32 * n)()
        for l in range(n):
            a = C.c_int32(); b = C.c_int32(); c = C.c_int32()
            ptrs[l] = L.ref_level_ptr(self.h, l, C.byref(a), C.byref(b), C.byref(c)); w[l] = a.value; h[l] = b.value; s[l] = c.value
        return n, ptrs, w, h, s


def stereo_match_ref(rL, rR, kpL, dL, kpR, dR, sf, inv_sf, mbf, min_d, max_d):
    """ComputeStereoMatches (oracle port) on the pyramids held by two RefExtractor handles."""
    n, pl, lw, lh, ls = rL.level_views(); _, pr, rw, rh, rs = rR.level_views()
    nL, nR = len(kpL), len(kpR)
    kpL = np.ascontiguousarray(kpL); kpR = np.ascontiguousarray(kpR); dL = np.ascontiguousarray(dL); dR = np.ascontiguousarray(dR)
    ur = np.empty(max(nL, 1), np.float32); dp = np.empty(max(nL, 1), np.float32); sad = np.empty(max(nL, 1), np.int32)
    lib.orc_stereo_match_levels.argtypes = [C.c_int] + [C.c_void_p] * 10 + [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                                                                         C.c_float, C.c_float, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]
    kept = lib.orc_stereo_match_levels(n, pl, lw, lh, ls, pr, rw, rh, rs, _p(kpL), _p(dL), nL, _p(kpR), _p(dR), nR, _p(sf), _p(inv_sf),
                                       mbf, min_d, max_d, _p(ur), _p(dp), _p(sad))
    return ur[:nL], dp[:nL], sad[:nL], kept


def bow_transform(voc, desc, levelsup=4):
    """Vocabulary-tree descent of DBoW2 transform (oracle/bow_oracle.cpp): (word ids, node ids)."""
    desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32); n = len(desc)
    parent = np.ascontiguousarray(voc["parent"], np.int32); leaf = np.ascontiguousarray(voc["is_leaf"], np.uint8); nd = np.ascontiguousarray(voc["desc"], np.uint8)
    word = np.empty(max(n, 1), np.int32); node = np.empty(max(n, 1), np.int32)
    lib.orc_bow_transform.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    lib.orc_bow_transform.restype = None
    lib.orc_bow_transform(int(voc["L"]), len(parent), _p(parent), _p(leaf), _p(nd), _p(desc), n, int(levelsup), _p(word), _p(node))
    return word[:n], node[:n]


def bow_score(ids_a, vals_a, ids_b, vals_b, scoring):
    """ORACLE (pure Python, small inputs): DBoW2 GeneralScoring::score of two BowVectors, the two-iterator merge of
    Thirdparty/DBoW2/DBoW2/ScoringObject.cpp:21-313 (L1 :21-59, L2 :64-106, chi-square :111-150, KL :155-200,
    Bhattacharyya :205-241, dot product :246-282).  scoring: DBoW2::ScoringType 0..5."""
    import math
    log_eps = math.log(2.0 ** -52)
    i = j = 0
    na, nb = len(ids_a), len(ids_b)
    s = 0.0
    while i < na and j < nb:
        vi, wi = float(vals_a[i]), float(vals_b[j])
        if ids_a[i] == ids_b[j]:
            if scoring == 0: s += math.fabs(vi - wi) - math.fabs(vi) - math.fabs(wi)
            elif scoring in (1, 5): s += vi * wi
            elif scoring == 2:
                if vi + wi != 0.0: s += vi * wi / (vi + wi)
            elif scoring == 3:
                if vi != 0 and wi != 0: s += vi * math.log(vi / wi)
            elif scoring == 4: s += math.sqrt(vi * wi)
            i += 1; j += 1
        elif ids_a[i] < ids_b[j]:
            if scoring == 3: s += vi * (math.log(vi) - log_eps)
            i += 1
        else:
            j += 1
    if scoring == 0: return -s / 2.0
    if scoring == 1: return 1.0 if s >= 1 else 1.0 - math.sqrt(1.0 - s)
    if scoring == 2: return 2.0 * s
    if scoring == 3:
        while i < na:
            if vals_a[i] != 0: s += float(vals_a[i]) * (math.log(float(vals_a[i])) - log_eps)
            i += 1
    return s


def is_in_frustum(Rcw, tcw, Ow, cam, bounds, cos_limit, xw, normal, mind, maxd):
    f32 = lambda a: np.ascontiguousarray(a, np.float32)
    Rcw, tcw, Ow, xw, normal, mind, maxd = f32(Rcw).reshape(9), f32(tcw).reshape(3), f32(Ow).reshape(3), f32(xw).reshape(-1, 3), f32(normal).reshape(-1, 3), f32(mind), f32(maxd)
    n = len(xw)
    c = _views.Camera(*[float(v) for v in cam]); g = _views.GridParams(float(bounds[0]), float(bounds[1]), float(bounds[2]), float(bounds[3]), 0.0, 0.0)
    ok = np.zeros(max(n, 1), np.uint8); outs = [np.zeros(max(n, 1), np.float32) for _ in range(5)]
    lib.orc_is_in_frustum.argtypes = [C.c_void_p] * 5 + [C.c_float, C.c_int] + [C.c_void_p] * 10
    lib.orc_is_in_frustum.restype = None
    lib.orc_is_in_frustum(_p(Rcw), _p(tcw), _p(Ow), C.byref(c), C.byref(g), cos_limit, n, _p(xw), _p(normal), _p(mind), _p(maxd), _p(ok), *[_p(o) for o in outs])
    return dict(in_view=ok[:n], u=outs[0][:n], v=outs[1][:n], ur=outs[2][:n], view_cos=outs[3][:n], dist=outs[4][:n])


def distinctive_descriptors(start, desc):
    start = np.ascontiguousarray(start, np.int32); desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    P = len(start) - 1
    best = np.empty(max(P, 1), np.int32)
    lib.orc_distinctive_descriptors.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.orc_distinctive_descriptors.restype = None
    lib.orc_distinctive_descriptors(P, _p(start), _p(desc), _p(best))
    return best[:P]


def remap_linear_u8(src, mapx, mapy):
    src = np.ascontiguousarray(src, np.uint8); mapx = np.ascontiguousarray(mapx, np.float32); mapy = np.ascontiguousarray(mapy, np.float32)
    dh, dw = mapx.shape
    dst = np.empty((dh, dw), np.uint8)
    lib.orc_remap_linear_u8.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int]
    lib.orc_remap_linear_u8.restype = None
    lib.orc_remap_linear_u8(_p(src), src.shape[1], src.shape[0], src.strides[0], _p(mapx), _p(mapy), _p(dst), dw, dh, dw)
    return dst


def gray_u8(src, rgb_order):
    src = np.ascontiguousarray(src, np.uint8)
    h, w, ch = src.shape
    dst = np.empty((h, w), np.uint8)
    lib.orc_gray_u8.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int]
    lib.orc_gray_u8.restype = None
    lib.orc_gray_u8(_p(src), w, h, src.strides[0], ch, int(bool(rgb_order)), _p(dst), w)
    return dst
