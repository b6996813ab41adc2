"""Host-side mirror of ORB_SLAM2::ORBmatcher (reference include/ORBmatcher.h:37-102) over the C
ABI: same constructor (nnratio, checkOri), same constants, same search entry points.  MapPoint /
KeyFrame objects of the reference become flat arrays (views.py); the returned match arrays carry
indices where the reference stores MapPoint pointers."""
import ctypes as C
import numpy as np
from .capi import check, lib, ptr
from .views import (Camera, Frame, FeatureVector, LastFrameView, MapPointView, ProjectedView, GridParams, FrameBatchView,
                    MapPointBatchView, LastFrameBatchView, u8, _addr)


class ORBmatcher:
    TH_LOW = 50
    TH_HIGH = 100
    HISTO_LENGTH = 30

    def __init__(self, nnratio=0.6, checkOri=True, max_n=8192, max_pairs=1, device=0):
        self.mfNNratio = float(nnratio)
        self.mbCheckOrientation = bool(checkOri)
        self._h = C.c_void_p()
        check(lib().orb_matcher_create(int(max_pairs), int(max_n), int(device), C.byref(self._h)))

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            lib().orb_matcher_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close

    @property
    def launch_count(self):
        return lib().orb_matcher_launch_count(self._h)

    # static int DescriptorDistance(const cv::Mat &a, const cv::Mat &b)
    def DescriptorDistance(self, a, b):
        a = np.ascontiguousarray(a, np.uint8).reshape(-1, 32); b = np.ascontiguousarray(b, np.uint8).reshape(-1, 32)
        out = np.empty(len(a), np.int32)
        check(lib().orb_descriptor_distance(self._h, ptr(a), ptr(b), len(a), ptr(out)))
        return int(out[0]) if len(out) == 1 else out

    def match_bruteforce(self, a, b):
        """a: (pairs, n_a, 32) or (n_a, 32); b likewise -> best_idx, best_dist, second_dist."""
        a = np.ascontiguousarray(a, np.uint8); b = np.ascontiguousarray(b, np.uint8)
        single = a.ndim == 2
        if single:
            a = a[None]; b = b[None]
        P, na, _ = a.shape; nb = b.shape[1]
        bi = np.empty((P, na), np.int32); bd = np.empty((P, na), np.int32); sd = np.empty((P, na), np.int32)
        check(lib().orb_match_bruteforce(self._h, ptr(a), ptr(b), P, na, nb, ptr(bi), ptr(bd), ptr(sd)))
        return (bi[0], bd[0], sd[0]) if single else (bi, bd, sd)

    def assign_features_to_grid(self, frame):
        """Frame::AssignFeaturesToGrid -> (cell_start[64*48+1], cell_items)."""
        cs = np.empty(64 * 48 + 1, np.int32); ci = np.empty(max(frame.N, 1), np.int32); n = C.c_int32()
        g = frame.view().grid
        check(lib().orb_assign_features_to_grid(self._h, ptr(frame.x), ptr(frame.y), frame.N, C.byref(g), ptr(cs), ptr(ci), C.byref(n)))
        return cs, ci[:n.value].copy()

    # int SearchByProjection(Frame &F, const std::vector<MapPoint*> &vpMapPoints, const float th=3)
    def SearchByProjection(self, F, mappoints, th=3.0, occupied=None):
        """mappoints: dict with valid, proj_x, proj_y, proj_xr, level, view_cos, desc, blocks."""
        mp = _mp_view(mappoints)
        match = np.empty(max(F.N, 1), np.int32); n = C.c_int32()
        occ = None if occupied is None else u8(occupied)
        fv = F.view()
        check(lib().orb_search_by_projection_mappoints(self._h, C.byref(fv), C.byref(mp[0]), C.c_float(th), C.c_float(self.mfNNratio),
                                                       ptr(occ), ptr(match), C.byref(n)))
        return n.value, match[:F.N]

    # int SearchByProjection(Frame &CurrentFrame, const Frame &LastFrame, const float th, const bool bMono)
    def SearchByProjectionLastFrame(self, Cur, last, cam, th, forward=False, backward=False, occupied=None):
        """last: dict with valid, xc, yc, zc, octave, angle, desc, blocks; cam: (fx, fy, cx, cy, mbf)."""
        lv = _last_view(last)
        match = np.empty(max(Cur.N, 1), np.int32); n = C.c_int32()
        occ = None if occupied is None else u8(occupied)
        cv = Cur.view(); c = Camera(*[float(v) for v in cam])
        check(lib().orb_search_by_projection_lastframe(self._h, C.byref(cv), C.byref(lv[0]), C.byref(c), C.c_float(th), int(forward),
                                                       int(backward), int(self.mbCheckOrientation), ptr(occ), ptr(match), C.byref(n)))
        return n.value, match[:Cur.N]

    # ---- batched forms of the two tracking searches: one call for many frame pairs (BASELINE.json configs[3])
    def SearchByProjectionBatch(self, frames, mappoints, th=3.0, occupied=None):
        """frames: list of Frame (same camera), mappoints: list of dicts as SearchByProjection takes, occupied: list of arrays or None.
        -> list of (nmatches, match) like the single call, from ONE library call."""
        fb = pack_frames(frames)
        qb = pack_points(mappoints, MP_FIELDS)
        occ = None if occupied is None else _pad_stack([u8(o) for o in occupied], fb["stride"], np.uint8)
        match, nm = self.search_mappoints_batch_packed(fb, qb, th, occ)
        return [(int(nm[p]), match[p, :frames[p].N].copy()) for p in range(len(frames))]

    def search_mappoints_batch_packed(self, fb, qb, th, occ=None, match=None, nm=None):
        """Packed (pairs, stride) arrays in, (pairs, stride) matches out; bench.py calls this with pinned buffers."""
        P, S = fb["pairs"], fb["stride"]
        match = np.empty((P, S), np.int32) if match is None else match
        nm = np.empty(P, np.int32) if nm is None else nm
        fv = frame_batch_view(fb)
        mv = MapPointBatchView(P, qb["stride"], _addr(qb["n"]), *[_addr(qb[k]) for k in MP_FIELDS])
        check(lib().orb_search_by_projection_mappoints_batch(self._h, C.byref(fv), C.byref(mv), C.c_float(th), C.c_float(self.mfNNratio),
                                                             ptr(occ), ptr(match), ptr(nm)))
        return match, nm

    def SearchByProjectionLastFrameBatch(self, curs, lasts, cam, th, forward, backward, occupied=None):
        """curs: list of Frame, lasts: list of dicts as SearchByProjectionLastFrame takes, forward / backward: one flag per pair."""
        fb = pack_frames(curs)
        qb = pack_points(lasts, LAST_FIELDS)
        occ = None if occupied is None else _pad_stack([u8(o) for o in occupied], fb["stride"], np.uint8)
        match, nm = self.search_lastframe_batch_packed(fb, qb, cam, th, u8(forward), u8(backward), occ)
        return [(int(nm[p]), match[p, :curs[p].N].copy()) for p in range(len(curs))]

    def search_lastframe_batch_packed(self, fb, qb, cam, th, forward, backward, occ=None, match=None, nm=None):
        P, S = fb["pairs"], fb["stride"]
        match = np.empty((P, S), np.int32) if match is None else match
        nm = np.empty(P, np.int32) if nm is None else nm
        fv = frame_batch_view(fb)
        lv = LastFrameBatchView(P, qb["stride"], _addr(qb["n"]), *[_addr(qb[k]) for k in LAST_FIELDS])
        c = Camera(*[float(v) for v in cam])
        check(lib().orb_search_by_projection_lastframe_batch(self._h, C.byref(fv), C.byref(lv), C.byref(c), C.c_float(th), ptr(forward), ptr(backward),
                                                             int(self.mbCheckOrientation), ptr(occ), ptr(match), ptr(nm)))
        return match, nm

    # int SearchByBoW(KeyFrame *pKF, Frame &F, std::vector<MapPoint*> &vpMapPointMatches)
    def SearchByBoW(self, KF, kf_featvec, kf_has_mappoint, F, f_featvec):
        match = np.empty(max(F.N, 1), np.int32); n = C.c_int32()
        hm = u8(kf_has_mappoint)
        kv = KF.view(); fv = F.view(); a = kf_featvec.view(); b = f_featvec.view()
        check(lib().orb_search_by_bow(self._h, C.byref(kv), C.byref(a), ptr(hm), C.byref(fv), C.byref(b), C.c_float(self.mfNNratio),
                                      int(self.mbCheckOrientation), ptr(match), C.byref(n)))
        return n.value, match[:F.N]

    # int SearchForTriangulation(KeyFrame *pKF1, KeyFrame *pKF2, cv::Mat F12, vector<pair<size_t,size_t>> &vMatchedPairs, bOnlyStereo)
    def SearchForTriangulation(self, KF1, fv1, has_mp1, KF2, fv2, has_mp2, F12, epipole, level_sigma2, bOnlyStereo=False):
        match = np.empty(max(KF1.N, 1), np.int32); n = C.c_int32()
        h1 = u8(has_mp1); h2 = u8(has_mp2)
        f12 = np.ascontiguousarray(F12, np.float32).reshape(9); s2 = np.ascontiguousarray(level_sigma2, np.float32)
        k1 = KF1.view(); k2 = KF2.view(); a = fv1.view(); b = fv2.view()
        check(lib().orb_search_for_triangulation(self._h, C.byref(k1), C.byref(a), ptr(h1), C.byref(k2), C.byref(b), ptr(h2), ptr(f12),
                                                 C.c_float(epipole[0]), C.c_float(epipole[1]), ptr(s2), int(bOnlyStereo),
                                                 int(self.mbCheckOrientation), ptr(match), C.byref(n)))
        m = match[:KF1.N]
        pairs = [(int(i), int(m[i])) for i in np.nonzero(m >= 0)[0]]
        return n.value, pairs


    # int SearchByBoW(KeyFrame *pKF1, KeyFrame *pKF2, std::vector<MapPoint*> &vpMatches12)
    def SearchByBoWKeyFrames(self, KF1, fv1, has_mp1, KF2, fv2, has_mp2):
        match = np.empty(max(KF1.N, 1), np.int32); n = C.c_int32()
        h1 = u8(has_mp1); h2 = u8(has_mp2)
        k1 = KF1.view(); k2 = KF2.view(); a = fv1.view(); b = fv2.view()
        check(lib().orb_search_by_bow_keyframes(self._h, C.byref(k1), C.byref(a), ptr(h1), C.byref(k2), C.byref(b), ptr(h2),
                                                C.c_float(self.mfNNratio), int(self.mbCheckOrientation), ptr(match), C.byref(n)))
        return n.value, match[:KF1.N]

    # int SearchByProjection(Frame &CurrentFrame, KeyFrame *pKF, const set<MapPoint*> &sAlreadyFound, th, ORBdist)
    # int SearchByProjection(KeyFrame *pKF, cv::Mat Scw, const vector<MapPoint*> &vpPoints, vector<MapPoint*> &vpMatched, int th)
    def SearchByProjectionClaiming(self, F, points, occupied, max_dist, check_ori=None):
        """points: dict valid, u, v, ur, radius, min_level, max_level, desc, angle (see orb_projected_view)."""
        pv = _proj_view(points)
        match = np.empty(max(F.N, 1), np.int32); pbest = np.empty(max(pv[0].n, 1), np.int32); n = C.c_int32()
        occ = None if occupied is None else u8(occupied)
        fv = F.view()
        co = self.mbCheckOrientation if check_ori is None else check_ori
        check(lib().orb_search_by_projection_claiming(self._h, C.byref(fv), C.byref(pv[0]), ptr(occ), int(max_dist), int(co), ptr(match),
                                                      ptr(pbest), C.byref(n)))
        return n.value, match[:F.N], pbest[:pv[0].n]

    # search step of int Fuse(KeyFrame *pKF, const vector<MapPoint*> &vpMapPoints, th) and Fuse(pKF, Scw, vpPoints, th, vpReplacePoint)
    def FuseCandidates(self, KF, points, inv_level_sigma2=None, max_dist=None):
        pv = _proj_view(points)
        bi = np.empty(max(pv[0].n, 1), np.int32); bd = np.empty(max(pv[0].n, 1), np.int32); n = C.c_int32()
        s2 = None if inv_level_sigma2 is None else np.ascontiguousarray(inv_level_sigma2, np.float32)
        kv = KF.view()
        check(lib().orb_search_projected_best(self._h, C.byref(kv), C.byref(pv[0]), ptr(s2), int(self.TH_LOW if max_dist is None else max_dist),
                                              ptr(bi), ptr(bd), C.byref(n)))
        return n.value, bi[:pv[0].n], bd[:pv[0].n]

    # int SearchBySim3(KeyFrame *pKF1, KeyFrame *pKF2, vector<MapPoint*> &vpMatches12, s12, R12, t12, th)
    def SearchBySim3(self, KF1, pts1_in_kf2, KF2, pts2_in_kf1):
        p1 = _proj_view(pts1_in_kf2); p2 = _proj_view(pts2_in_kf1)
        match = np.empty(max(KF1.N, 1), np.int32); n = C.c_int32()
        k1 = KF1.view(); k2 = KF2.view()
        check(lib().orb_search_by_sim3(self._h, C.byref(k1), C.byref(p1[0]), C.byref(k2), C.byref(p2[0]), ptr(match), C.byref(n)))
        return n.value, match[:KF1.N]

    # int SearchForInitialization(Frame &F1, Frame &F2, vector<cv::Point2f> &vbPrevMatched, vector<int> &vnMatches12, int windowSize=10)
    def SearchForInitialization(self, F1, F2, prev_x, prev_y, windowSize=10):
        """Returns (nmatches, vnMatches12); prev_x / prev_y (float32 arrays) are updated in place (:571-574)."""
        match = np.empty(max(F1.N, 1), np.int32); n = C.c_int32()
        f1 = F1.view(); f2 = F2.view()
        check(lib().orb_search_for_initialization(self._h, C.byref(f1), C.byref(f2), ptr(prev_x), ptr(prev_y), int(windowSize),
                                                  C.c_float(self.mfNNratio), int(self.mbCheckOrientation), ptr(match), C.byref(n)))
        m = match[:F1.N]
        ok = m >= 0
        prev_x[ok] = F2.x[m[ok]]; prev_y[ok] = F2.y[m[ok]]
        return n.value, m

    # bool Frame::isInFrustum(MapPoint *pMP, float viewingCosLimit), batched over the local map (src/Frame.cc:303-366)
    def isInFrustum(self, Rcw, tcw, Ow, cam, bounds, viewingCosLimit, xw, normal, min_distance, max_distance):
        """-> dict(in_view, u, v, ur, view_cos, dist); mnTrackScaleLevel = PredictScale(dist) stays with the caller."""
        f32 = lambda a: np.ascontiguousarray(a, np.float32)
        Rcw, tcw, Ow, xw, normal = f32(Rcw).reshape(9), f32(tcw).reshape(3), f32(Ow).reshape(3), f32(xw).reshape(-1, 3), f32(normal).reshape(-1, 3)
        mind, maxd = f32(min_distance), f32(max_distance)
        n = len(xw)
        c = Camera(*[float(v) for v in cam]); g = GridParams(float(bounds[0]), float(bounds[1]), float(bounds[2]), float(bounds[3]), 0.0, 0.0)
        ok = np.zeros(max(n, 1), np.uint8); outs = [np.zeros(max(n, 1), np.float32) for _ in range(5)]
        check(lib().orb_is_in_frustum(self._h, ptr(Rcw), ptr(tcw), ptr(Ow), C.byref(c), C.byref(g), C.c_float(viewingCosLimit), n, ptr(xw), ptr(normal),
                                      ptr(mind), ptr(maxd), ptr(ok), *[ptr(o) for o in outs]))
        return dict(in_view=ok[:n], u=outs[0][:n], v=outs[1][:n], ur=outs[2][:n], view_cos=outs[3][:n], dist=outs[4][:n])

    # void MapPoint::ComputeDistinctiveDescriptors(), batched over MapPoints (src/MapPoint.cc:250-326)
    def ComputeDistinctiveDescriptors(self, start, desc):
        """start[P + 1] offsets into desc (rows of each point's observations, reference gather order) -> best row per point."""
        start = np.ascontiguousarray(start, np.int32); desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        P = len(start) - 1
        best = np.empty(max(P, 1), np.int32)
        check(lib().orb_distinctive_descriptors(self._h, P, ptr(start), ptr(desc), ptr(best)))
        return best[:P]


MP_FIELDS = ("valid", "proj_x", "proj_y", "proj_xr", "level", "view_cos", "desc", "blocks")
LAST_FIELDS = ("valid", "xc", "yc", "zc", "octave", "angle", "desc", "blocks")
_DT = dict(valid=np.uint8, blocks=np.uint8, level=np.int32, octave=np.int32, desc=np.uint8)


def _pad_stack(arrays, stride, dtype, width=None):
    out = np.zeros((len(arrays), stride) + (() if width is None else (width,)), dtype)
    for p, a in enumerate(arrays):
        out[p, :len(a)] = a
    return out


def pack_frames(frames):
    """list of Frame -> dict of (pairs, stride[, 32]) arrays for the batched entry points (one camera: grid and scale factors of frames[0])."""
    S = max(max(f.N for f in frames), 1)
    fb = dict(pairs=len(frames), stride=S, n=np.array([f.N for f in frames], np.int32), frame0=frames[0])
    fb["x"] = _pad_stack([f.x for f in frames], S, np.float32); fb["y"] = _pad_stack([f.y for f in frames], S, np.float32)
    fb["octave"] = _pad_stack([f.octave for f in frames], S, np.int32); fb["angle"] = _pad_stack([f.angle for f in frames], S, np.float32)
    fb["u_right"] = None if frames[0].u_right is None else _pad_stack([f.u_right for f in frames], S, np.float32)
    fb["desc"] = _pad_stack([f.desc for f in frames], S, np.uint8, 32)
    return fb


def frame_batch_view(fb):
    f0 = fb["frame0"]
    g = GridParams(f0.bounds[0], f0.bounds[1], f0.bounds[2], f0.bounds[3], f0.inv_w, f0.inv_h)
    return FrameBatchView(fb["pairs"], fb["stride"], _addr(fb["n"]), _addr(fb["x"]), _addr(fb["y"]), _addr(fb["octave"]), _addr(fb["angle"]),
                          _addr(fb["u_right"]), _addr(fb["desc"]), g, _addr(f0.sf), len(f0.sf))


def pack_points(dicts, fields):
    S = max(max(len(d["valid"]) for d in dicts), 1)
    qb = dict(pairs=len(dicts), stride=S, n=np.array([len(d["valid"]) for d in dicts], np.int32))
    for k in fields:
        dt = _DT.get(k, np.float32)
        arrs = [np.asarray(d[k]).astype(dt).reshape((-1, 32) if k == "desc" else (-1,)) for d in dicts]
        qb[k] = _pad_stack(arrs, S, dt, 32 if k == "desc" else None)
    return qb


def _mp_view(d):
    keep = dict(valid=u8(d["valid"]), proj_x=np.ascontiguousarray(d["proj_x"], np.float32), proj_y=np.ascontiguousarray(d["proj_y"], np.float32),
                proj_xr=np.ascontiguousarray(d["proj_xr"], np.float32), level=np.ascontiguousarray(d["level"], np.int32),
                view_cos=np.ascontiguousarray(d["view_cos"], np.float32), desc=np.ascontiguousarray(d["desc"], np.uint8), blocks=u8(d["blocks"]))
    v = MapPointView(len(keep["valid"]), *[_addr(keep[k]) for k in ("valid", "proj_x", "proj_y", "proj_xr", "level", "view_cos", "desc", "blocks")])
    return v, keep


def _last_view(d):
    keep = dict(valid=u8(d["valid"]), xc=np.ascontiguousarray(d["xc"], np.float32), yc=np.ascontiguousarray(d["yc"], np.float32),
                zc=np.ascontiguousarray(d["zc"], np.float32), octave=np.ascontiguousarray(d["octave"], np.int32),
                angle=np.ascontiguousarray(d["angle"], np.float32), desc=np.ascontiguousarray(d["desc"], np.uint8), blocks=u8(d["blocks"]))
    v = LastFrameView(len(keep["valid"]), *[_addr(keep[k]) for k in ("valid", "xc", "yc", "zc", "octave", "angle", "desc", "blocks")])
    return v, keep


def _proj_view(d):
    f32 = lambda a: None if a is None else np.ascontiguousarray(a, np.float32)
    keep = dict(valid=u8(d["valid"]), u=f32(d["u"]), v=f32(d["v"]), ur=f32(d.get("ur")), radius=f32(d["radius"]),
                min_level=np.ascontiguousarray(d["min_level"], np.int32), max_level=np.ascontiguousarray(d["max_level"], np.int32),
                desc=np.ascontiguousarray(d["desc"], np.uint8), angle=f32(d.get("angle")))
    v = ProjectedView(len(keep["valid"]), *[_addr(keep[k]) for k in ("valid", "u", "v", "ur", "radius", "min_level", "max_level", "desc", "angle")])
    return v, keep


GRID_STATIC_MAX_N = 8192   # include/orb_b200.h: ORB_GRID_STATIC_MAX_N


def assign_features_to_grid_static(x, y, n, grid, cell_start, cell_items, n_items, stream=0):
    """orb_assign_features_to_grid_static: Frame::AssignFeaturesToGrid through the allocation-free entry point (no handle; the
    kernel's device buffers are a static __device__ array of the library).  Every array argument is a numpy array (host) or
    an integer address (host or device memory of the current device); n_items likewise.  Calls must not overlap on the GPU."""
    check(lib().orb_assign_features_to_grid_static(ptr(x), ptr(y), int(n), C.byref(grid), ptr(cell_start), ptr(cell_items),
                                                   ptr(n_items), C.c_void_p(stream)))
