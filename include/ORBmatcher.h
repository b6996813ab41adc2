/* Drop-in for the reference's include/ORBmatcher.h over the C ABI of liborb_b200.so: same
 * namespace, class name, constructor, constants and Search* signatures (reference
 * include/ORBmatcher.h:37-102).  Distances, candidate search, ratio tests and the rotation
 * histogram run in sm_100a CUDA kernels; the methods here only gather the reference's objects
 * into the flat views of include/orb_b200.h (taking the same getters, hence the same mutexes, as
 * the reference) and scatter the resulting indices back into the MapPoint* vectors.
 *
 * The methods are templates over the Frame / KeyFrame / MapPoint types so that this header
 * compiles without the rest of ORB-SLAM2: inside the reference tree they instantiate with
 * ORB_SLAM2::Frame, ORB_SLAM2::KeyFrame and ORB_SLAM2::MapPoint (unchanged call sites:
 * matcher.SearchByProjection(mCurrentFrame, mLastFrame, th, bMono) ...), in tests/cpp with small
 * structs carrying the same member names.  No CPU fallback: without a CUDA device the constructor
 * throws.  Not covered yet (SURVEY 8f "next"): SearchForInitialization, SearchBySim3, Fuse,
 * the relocalisation / loop SearchByProjection variants and SearchByBoW(KF, KF).
 */
#ifndef ORBMATCHER_H
#define ORBMATCHER_H

#include <cmath>
#include <map>
#include <set>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "orb_b200.h"
#include "orb_b200_cv.h"

namespace ORB_SLAM2 {

class ORBmatcher {
public:
    ORBmatcher(float nnratio = 0.6, bool checkOri = true) : mfNNratio(nnratio), mbCheckOrientation(checkOri) {}

    // Computes the Hamming distance between two ORB descriptors (src/ORBmatcher.cc:1814-1830)
    static int DescriptorDistance(const cv::Mat& a, const cv::Mat& b) {
        int32_t d = 0;
        Check(orb_descriptor_distance(Handle(), a.ptr<unsigned char>(0), b.ptr<unsigned char>(0), 1, &d));
        return d;
    }
    // n pairs at once: rows of a against rows of b
    static void DescriptorDistances(const cv::Mat& a, const cv::Mat& b, std::vector<int>& dist) {
        const int n = a.rows;
        std::vector<unsigned char> pa((size_t)n * 32), pb((size_t)n * 32);
        for (int i = 0; i < n; ++i) { std::memcpy(&pa[(size_t)i * 32], a.ptr<unsigned char>(i), 32); std::memcpy(&pb[(size_t)i * 32], b.ptr<unsigned char>(i), 32); }
        dist.resize(n);
        if (n) Check(orb_descriptor_distance(Handle(), pa.data(), pb.data(), n, dist.data()));
    }

    // Search matches between Frame keypoints and projected MapPoints. Returns number of matches
    // Used to track the local map (Tracking)                         (src/ORBmatcher.cc:47-133)
    template <class FrameT, class MapPointT>
    int SearchByProjection(FrameT& F, const std::vector<MapPointT*>& vpMapPoints, const float th = 3) {
        FrameArrays fa; GatherFrame(F, fa);
        const int M = (int)vpMapPoints.size();
        std::vector<unsigned char> valid(M, 0), blocks(M, 0), desc((size_t)M * 32, 0);
        std::vector<float> px(M, 0.f), py(M, 0.f), pxr(M, 0.f), vc(M, 1.f);
        std::vector<int32_t> lvl(M, 0);
        for (int i = 0; i < M; ++i) {
            MapPointT* pMP = vpMapPoints[i];
            if (!pMP || !pMP->mbTrackInView) continue;  // :59
            if (pMP->isBad()) continue;                 // :62
            valid[i] = 1;
            px[i] = pMP->mTrackProjX; py[i] = pMP->mTrackProjY; pxr[i] = pMP->mTrackProjXR;
            lvl[i] = pMP->mnTrackScaleLevel; vc[i] = pMP->mTrackViewCos;
            const cv::Mat d = pMP->GetDescriptor();     // clone under mMutexFeatures, as the reference (:84)
            std::memcpy(&desc[(size_t)i * 32], d.template ptr<unsigned char>(0), 32);
            blocks[i] = pMP->Observations() > 0;
        }
        std::vector<unsigned char> occ; GatherOccupied(F, occ);
        orb_mappoint_view mv = {M, valid.data(), px.data(), py.data(), pxr.data(), lvl.data(), vc.data(), desc.data(), blocks.data()};
        std::vector<int32_t> match(fa.n > 0 ? fa.n : 1, -1);
        int32_t nmatches = 0;
        Check(orb_search_by_projection_mappoints(Handle(), &fa.view, &mv, th, mfNNratio, occ.data(), match.data(), &nmatches));
        for (int i = 0; i < fa.n; ++i) if (match[i] >= 0) F.mvpMapPoints[i] = vpMapPoints[match[i]];  // :127
        return nmatches;
    }

    // Project MapPoints tracked in last frame into the current frame and search matches.
    // Used to track from previous frame (Tracking)                   (src/ORBmatcher.cc:1460-1619)
    template <class FrameT>
    int SearchByProjection(FrameT& CurrentFrame, const FrameT& LastFrame, const float th, const bool bMono) {
        FrameArrays fa; GatherFrame(CurrentFrame, fa);
        // :1472-1487 (float matrices; products accumulate in double like cv::gemm on CV_32F)
        float Rcw[9], tcw[3], Rlw[9], tlw[3], twc[3], tlc[3];
        SplitPose(CurrentFrame.mTcw, Rcw, tcw); SplitPose(LastFrame.mTcw, Rlw, tlw);
        for (int r = 0; r < 3; ++r) twc[r] = (float)(-((double)Rcw[0 * 3 + r] * tcw[0] + (double)Rcw[1 * 3 + r] * tcw[1] + (double)Rcw[2 * 3 + r] * tcw[2]));
        for (int r = 0; r < 3; ++r) tlc[r] = MulRowAdd(Rlw + 3 * r, twc, tlw[r]);
        const bool bForward = tlc[2] > CurrentFrame.mb && !bMono;
        const bool bBackward = -tlc[2] > CurrentFrame.mb && !bMono;

        const int M = LastFrame.N;
        std::vector<unsigned char> valid(M, 0), blocks(M, 0), desc((size_t)M * 32, 0);
        std::vector<float> xc(M, 0.f), yc(M, 0.f), zc(M, 0.f), ang(M, 0.f);
        std::vector<int32_t> oct(M, 0);
        for (int i = 0; i < M; ++i) {
            auto* pMP = LastFrame.mvpMapPoints[i];
            if (!pMP || LastFrame.mvbOutlier[i]) continue;  // :1493-1497
            valid[i] = 1;
            const cv::Mat x3Dw = pMP->GetWorldPos();
            const float xw[3] = {x3Dw.template at<float>(0), x3Dw.template at<float>(1), x3Dw.template at<float>(2)};
            xc[i] = MulRowAdd(Rcw + 0, xw, tcw[0]); yc[i] = MulRowAdd(Rcw + 3, xw, tcw[1]); zc[i] = MulRowAdd(Rcw + 6, xw, tcw[2]);  // :1500-1501
            oct[i] = LastFrame.mvKeys[i].octave;                                    // :1520
            ang[i] = LastFrame.mvKeysUn[i].angle;                                   // :1584
            const cv::Mat d = pMP->GetDescriptor();                                 // :1541
            std::memcpy(&desc[(size_t)i * 32], d.template ptr<unsigned char>(0), 32);
            blocks[i] = pMP->Observations() > 0;
        }
        std::vector<unsigned char> occ; GatherOccupied(CurrentFrame, occ);
        orb_lastframe_view lv = {M, valid.data(), xc.data(), yc.data(), zc.data(), oct.data(), ang.data(), desc.data(), blocks.data()};
        orb_camera cam = {CurrentFrame.fx, CurrentFrame.fy, CurrentFrame.cx, CurrentFrame.cy, CurrentFrame.mbf};
        std::vector<int32_t> match(fa.n > 0 ? fa.n : 1, -1);
        int32_t nmatches = 0;
        Check(orb_search_by_projection_lastframe(Handle(), &fa.view, &lv, &cam, th, bForward, bBackward, mbCheckOrientation, occ.data(),
                                                 match.data(), &nmatches));
        // :1577 assigns inside the loop and :1597-1616 resets the culled ones to NULL: net effect = surviving matches
        for (int i = 0; i < fa.n; ++i) if (match[i] >= 0) CurrentFrame.mvpMapPoints[i] = LastFrame.mvpMapPoints[match[i]];
        return nmatches;
    }

    // Search matches between MapPoints in a KeyFrame and ORB in a Frame.
    // Brute force constrained to ORB that belong to the same vocabulary node (at a certain level)
    // Used in Relocalisation and Loop Detection                      (src/ORBmatcher.cc:164-306)
    template <class KeyFrameT, class FrameT, class MapPointT>
    int SearchByBoW(KeyFrameT* pKF, FrameT& F, std::vector<MapPointT*>& vpMapPointMatches) {
        const std::vector<MapPointT*> vpMapPointsKF = pKF->GetMapPointMatches();  // :166
        vpMapPointMatches = std::vector<MapPointT*>(F.N, static_cast<MapPointT*>(NULL));
        FrameArrays ka, fa; GatherFrame(*pKF, ka); GatherFrame(F, fa);
        std::vector<unsigned char> has(ka.n, 0);
        for (int i = 0; i < ka.n; ++i) has[i] = vpMapPointsKF[i] && !vpMapPointsKF[i]->isBad();  // :205-211
        FeatVec kv, fv; Flatten(pKF->mFeatVec, kv); Flatten(F.mFeatVec, fv);
        std::vector<int32_t> match(fa.n > 0 ? fa.n : 1, -1);
        int32_t nmatches = 0;
        Check(orb_search_by_bow(Handle(), &ka.view, &kv.view, has.data(), &fa.view, &fv.view, mfNNratio, mbCheckOrientation, match.data(), &nmatches));
        for (int j = 0; j < fa.n; ++j) if (match[j] >= 0) vpMapPointMatches[j] = vpMapPointsKF[match[j]];  // :245
        return nmatches;
    }

    // Matching to triangulate new MapPoints. Check Epipolar Constraint.   (src/ORBmatcher.cc:718-900)
    template <class KeyFrameT>
    int SearchForTriangulation(KeyFrameT* pKF1, KeyFrameT* pKF2, cv::Mat F12, std::vector<std::pair<size_t, size_t> >& vMatchedPairs,
                               const bool bOnlyStereo) {
        // epipole in the second image (:724-730)
        const cv::Mat Cw = pKF1->GetCameraCenter(), R2w = pKF2->GetRotation(), t2w = pKF2->GetTranslation();
        float R[9], t[3], c[3], C2[3];
        for (int r = 0; r < 3; ++r) { for (int k = 0; k < 3; ++k) R[3 * r + k] = R2w.template at<float>(r, k); t[r] = t2w.template at<float>(r); c[r] = Cw.template at<float>(r); }
        for (int r = 0; r < 3; ++r) C2[r] = MulRowAdd(R + 3 * r, c, t[r]);
        const float invz = 1.0f / C2[2];
        const float ex = pKF2->fx * C2[0] * invz + pKF2->cx;
        const float ey = pKF2->fy * C2[1] * invz + pKF2->cy;

        FrameArrays a1, a2; GatherFrame(*pKF1, a1); GatherFrame(*pKF2, a2);
        std::vector<unsigned char> has1(a1.n, 0), has2(a2.n, 0);
        for (int i = 0; i < a1.n; ++i) has1[i] = pKF1->GetMapPoint(i) != NULL;  // :771-775
        for (int i = 0; i < a2.n; ++i) has2[i] = pKF2->GetMapPoint(i) != NULL;  // :792-796
        FeatVec v1, v2; Flatten(pKF1->mFeatVec, v1); Flatten(pKF2->mFeatVec, v2);
        float f12[9];
        for (int r = 0; r < 3; ++r) for (int k = 0; k < 3; ++k) f12[3 * r + k] = F12.at<float>(r, k);
        std::vector<int32_t> match(a1.n > 0 ? a1.n : 1, -1);
        int32_t nmatches = 0;
        Check(orb_search_for_triangulation(Handle(), &a1.view, &v1.view, has1.data(), &a2.view, &v2.view, has2.data(), f12, ex, ey,
                                           pKF2->mvLevelSigma2.data(), bOnlyStereo, mbCheckOrientation, match.data(), &nmatches));
        vMatchedPairs.clear();
        vMatchedPairs.reserve(nmatches);
        for (int i = 0; i < a1.n; ++i) if (match[i] >= 0) vMatchedPairs.push_back(std::make_pair((size_t)i, (size_t)match[i]));  // :888-897
        return nmatches;
    }

public:
    static const int TH_LOW = 50;
    static const int TH_HIGH = 100;
    static const int HISTO_LENGTH = 30;

    // One matcher workspace per thread: ORBmatcher objects are per-call locals used concurrently from the
    // Tracking, LocalMapping and LoopClosing threads; the workspace outlives them.
    static void SetWorkspace(int max_features, int device) { Slot().max_n = max_features; Slot().device = device; Slot().Reset(); }

protected:
    struct FrameArrays {
        int n;
        std::vector<float> x, y, angle, ur;
        std::vector<int32_t> octave;
        std::vector<unsigned char> desc;
        std::vector<float> sf;
        orb_frame_view view;
    };
    struct FeatVec {
        std::vector<uint32_t> ids, items;
        std::vector<int32_t> start;
        orb_featvec_view view;
    };
    struct WorkspaceSlot {
        orb_matcher* h; int max_n, device;
        WorkspaceSlot() : h(nullptr), max_n(16384), device(0) {}
        ~WorkspaceSlot() { Reset(); }
        void Reset() { if (h) { orb_matcher_destroy(h); h = nullptr; } }
    };
    static WorkspaceSlot& Slot() { static thread_local WorkspaceSlot s; return s; }
    static orb_matcher* Handle() {
        WorkspaceSlot& s = Slot();
        if (!s.h) Check(orb_matcher_create(1, s.max_n, s.device, &s.h));
        return s.h;
    }
    static void Check(int status) {
        if (status != ORB_OK) throw std::runtime_error(std::string("orb_b200: ") + orb_last_error());
    }

    // mvKeysUn / mvuRight / mDescriptors / grid constants / mvScaleFactors of a Frame or KeyFrame
    template <class FrameLike>
    static void GatherFrame(const FrameLike& F, FrameArrays& a) {
        const int n = a.n = F.N;
        a.x.resize(n); a.y.resize(n); a.angle.resize(n); a.octave.resize(n); a.ur.resize(n); a.desc.resize((size_t)n * 32);
        for (int i = 0; i < n; ++i) {
            const cv::KeyPoint& kp = F.mvKeysUn[i];
            a.x[i] = kp.pt.x; a.y[i] = kp.pt.y; a.angle[i] = kp.angle; a.octave[i] = kp.octave;
            a.ur[i] = F.mvuRight[i];
            std::memcpy(&a.desc[(size_t)i * 32], F.mDescriptors.template ptr<unsigned char>(i), 32);
        }
        a.sf.assign(F.mvScaleFactors.begin(), F.mvScaleFactors.end());
        orb_frame_view v;
        v.n = n; v.x = a.x.data(); v.y = a.y.data(); v.octave = a.octave.data(); v.angle = a.angle.data(); v.u_right = a.ur.data();
        v.desc = a.desc.data();
        v.grid.min_x = F.mnMinX; v.grid.min_y = F.mnMinY; v.grid.max_x = F.mnMaxX; v.grid.max_y = F.mnMaxY;
        v.grid.inv_w = F.mfGridElementWidthInv; v.grid.inv_h = F.mfGridElementHeightInv;
        v.scale_factors = a.sf.data(); v.n_levels = (int32_t)a.sf.size();
        a.view = v;
    }
    // keypoints already holding a MapPoint with observations (:90-93, :1550-1553)
    template <class FrameT>
    static void GatherOccupied(const FrameT& F, std::vector<unsigned char>& occ) {
        occ.assign(F.N > 0 ? F.N : 1, 0);
        for (int i = 0; i < F.N; ++i) if (F.mvpMapPoints[i] && F.mvpMapPoints[i]->Observations() > 0) occ[i] = 1;
    }
    // DBoW2::FeatureVector is a std::map<NodeId, std::vector<unsigned int>>: ascending node id
    template <class FeatureVectorT>
    static void Flatten(const FeatureVectorT& fv, FeatVec& out) {
        out.ids.clear(); out.items.clear(); out.start.assign(1, 0);
        for (typename FeatureVectorT::const_iterator it = fv.begin(); it != fv.end(); ++it) {
            out.ids.push_back((uint32_t)it->first);
            for (size_t k = 0; k < it->second.size(); ++k) out.items.push_back((uint32_t)it->second[k]);
            out.start.push_back((int32_t)out.items.size());
        }
        out.view.n_nodes = (int32_t)out.ids.size(); out.view.node_ids = out.ids.data(); out.view.start = out.start.data();
        out.view.items = out.items.data();
    }
    static void SplitPose(const cv::Mat& Tcw, float* R, float* t) {
        for (int r = 0; r < 3; ++r) { for (int k = 0; k < 3; ++k) R[3 * r + k] = Tcw.at<float>(r, k); t[r] = Tcw.at<float>(r, 3); }
    }
    // one row of R * x + t on CV_32F data: cv::gemm accumulates float products in double
    static float MulRowAdd(const float* row, const float* x, float t) {
        return (float)((double)row[0] * x[0] + (double)row[1] * x[1] + (double)row[2] * x[2] + (double)t);
    }

    float mfNNratio;
    bool mbCheckOrientation;
};

}  // namespace ORB_SLAM2

#endif
