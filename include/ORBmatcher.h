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
 * throws.  Every public method of the reference class is covered.  The projection searches of
 * relocalisation, loop closing, Fuse and SearchBySim3 keep the reference's host-side gates
 * (cv::Mat float algebra, MapPoint::PredictScale, IsInImage) in this header, written with the
 * reference's expressions, and replay the map updates (Replace / AddObservation / AddMapPoint) in
 * point order on the host; the candidate search, Hamming distances and claims run on the device.
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
        // :1472-1487 (float matrices; -Rcw.t() * tcw carries a transpose flag: cv::gemm accumulates it in double)
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

    // Matching for the Map Initialization (only used in the monocular case)   (src/ORBmatcher.cc:429-578)
    template <class FrameT>
    int SearchForInitialization(FrameT& F1, FrameT& F2, std::vector<cv::Point2f>& vbPrevMatched, std::vector<int>& vnMatches12, int windowSize = 10) {
        FrameArrays a1, a2; GatherFrame(F1, a1); GatherFrame(F2, a2);
        std::vector<float> px(a1.n > 0 ? a1.n : 1), py(a1.n > 0 ? a1.n : 1);
        for (int i = 0; i < a1.n; ++i) { px[i] = vbPrevMatched[i].x; py[i] = vbPrevMatched[i].y; }
        std::vector<int32_t> match(a1.n > 0 ? a1.n : 1, -1);
        int32_t nmatches = 0;
        Check(orb_search_for_initialization(Handle(), &a1.view, &a2.view, px.data(), py.data(), windowSize, mfNNratio, mbCheckOrientation,
                                            match.data(), &nmatches));
        vnMatches12.assign(match.begin(), match.begin() + a1.n);
        for (int i = 0; i < a1.n; ++i) if (vnMatches12[i] >= 0) vbPrevMatched[i] = F2.mvKeysUn[vnMatches12[i]].pt;  // :571-574
        return nmatches;
    }

    // Loop detection: brute force within vocabulary nodes between two KeyFrames   (src/ORBmatcher.cc:581-716)
    template <class KeyFrameT, class MapPointT>
    int SearchByBoW(KeyFrameT* pKF1, KeyFrameT* pKF2, std::vector<MapPointT*>& vpMatches12) {
        const std::vector<MapPointT*> vpMapPoints1 = pKF1->GetMapPointMatches(), vpMapPoints2 = pKF2->GetMapPointMatches();
        vpMatches12 = std::vector<MapPointT*>(vpMapPoints1.size(), static_cast<MapPointT*>(NULL));
        FrameArrays a1, a2; GatherFrame(*pKF1, a1); GatherFrame(*pKF2, a2);
        std::vector<unsigned char> has1(a1.n > 0 ? a1.n : 1, 0), has2(a2.n > 0 ? a2.n : 1, 0);
        for (int i = 0; i < a1.n; ++i) has1[i] = vpMapPoints1[i] && !vpMapPoints1[i]->isBad();  // :620-624
        for (int i = 0; i < a2.n; ++i) has2[i] = vpMapPoints2[i] && !vpMapPoints2[i]->isBad();  // :639-645
        FeatVec v1, v2; Flatten(pKF1->mFeatVec, v1); Flatten(pKF2->mFeatVec, v2);
        std::vector<int32_t> match(a1.n > 0 ? a1.n : 1, -1);
        int32_t nmatches = 0;
        Check(orb_search_by_bow_keyframes(Handle(), &a1.view, &v1.view, has1.data(), &a2.view, &v2.view, has2.data(), mfNNratio,
                                          mbCheckOrientation, match.data(), &nmatches));
        for (int i = 0; i < a1.n; ++i) if (match[i] >= 0) vpMatches12[i] = vpMapPoints2[match[i]];  // :665
        return nmatches;
    }

    // Project MapPoints seen in KeyFrame into the Frame and search matches.
    // Used in relocalisation (Tracking)                                  (src/ORBmatcher.cc:1622-1759)
    template <class FrameT, class KeyFrameT, class MapPointT>
    int SearchByProjection(FrameT& CurrentFrame, KeyFrameT* pKF, const std::set<MapPointT*>& sAlreadyFound, const float th, const int ORBdist) {
        FrameArrays fa; GatherFrame(CurrentFrame, fa);
        float Rcw[9], tcw[3], Ow[3];
        SplitPose(CurrentFrame.mTcw, Rcw, tcw);
        CameraCentre(Rcw, tcw, Ow);                                              // :1626-1628
        const std::vector<MapPointT*> vpMPs = pKF->GetMapPointMatches();
        const int M = (int)vpMPs.size();
        Projected pv(M);
        for (int i = 0; i < M; ++i) {
            MapPointT* pMP = vpMPs[i];
            if (!pMP || pMP->isBad() || sAlreadyFound.count(pMP)) continue;      // :1639-1643
            const cv::Mat x3Dw = pMP->GetWorldPos();
            const float xw[3] = {x3Dw.template at<float>(0), x3Dw.template at<float>(1), x3Dw.template at<float>(2)};
            const float xc = MulRowAdd(Rcw + 0, xw, tcw[0]), yc = MulRowAdd(Rcw + 3, xw, tcw[1]), zc = MulRowAdd(Rcw + 6, xw, tcw[2]);
            const float invzc = (float)(1.0 / zc);                               // :1650
            const float u = CurrentFrame.fx * xc * invzc + CurrentFrame.cx, v = CurrentFrame.fy * yc * invzc + CurrentFrame.cy;
            if (u < CurrentFrame.mnMinX || u > CurrentFrame.mnMaxX) continue;
            if (v < CurrentFrame.mnMinY || v > CurrentFrame.mnMaxY) continue;
            const float dist3D = Norm3(xw[0] - Ow[0], xw[1] - Ow[1], xw[2] - Ow[2]);   // :1661-1662
            if (dist3D < pMP->GetMinDistanceInvariance() || dist3D > pMP->GetMaxDistanceInvariance()) continue;
            const int lvl = pMP->PredictScale(dist3D, &CurrentFrame);
            pv.Set(i, u, v, 0.f, th * CurrentFrame.mvScaleFactors[lvl], lvl - 1, lvl + 1, pMP->GetDescriptor(), pKF->mvKeysUn[i].angle);
        }
        std::vector<unsigned char> occ(fa.n > 0 ? fa.n : 1, 0);
        for (int i = 0; i < fa.n; ++i) occ[i] = CurrentFrame.mvpMapPoints[i] != NULL;    // :1690
        std::vector<int32_t> match(fa.n > 0 ? fa.n : 1, -1);
        int32_t nmatches = 0;
        Check(orb_search_by_projection_claiming(Handle(), &fa.view, pv.View(true), occ.data(), ORBdist, mbCheckOrientation, match.data(), NULL, &nmatches));
        for (int i = 0; i < fa.n; ++i) if (match[i] >= 0) CurrentFrame.mvpMapPoints[i] = vpMPs[match[i]];
        return nmatches;
    }

    // Project MapPoints using a Similarity Transformation and search matches.
    // Used in loop detection (Loop Closing)                              (src/ORBmatcher.cc:311-427)
    template <class KeyFrameT, class MapPointT>
    int SearchByProjection(KeyFrameT* pKF, cv::Mat Scw, const std::vector<MapPointT*>& vpPoints, std::vector<MapPointT*>& vpMatched, int th) {
        FrameArrays fa; GatherFrame(*pKF, fa);
        const float fx = pKF->fx, fy = pKF->fy, cx = pKF->cx, cy = pKF->cy;
        float Rcw[9], tcw[3], Ow[3];
        DecomposeSim3(Scw, Rcw, tcw);                                            // :320-325
        CameraCentre(Rcw, tcw, Ow);
        std::set<MapPointT*> spAlreadyFound(vpMatched.begin(), vpMatched.end());
        spAlreadyFound.erase(static_cast<MapPointT*>(NULL));
        const int M = (int)vpPoints.size();
        Projected pv(M);
        for (int i = 0; i < M; ++i) {
            MapPointT* pMP = vpPoints[i];
            if (pMP->isBad() || spAlreadyFound.count(pMP)) continue;             // :340
            const cv::Mat p3Dw = pMP->GetWorldPos();
            const float xw[3] = {p3Dw.template at<float>(0), p3Dw.template at<float>(1), p3Dw.template at<float>(2)};
            const float p3Dc[3] = {MulRowAdd(Rcw + 0, xw, tcw[0]), MulRowAdd(Rcw + 3, xw, tcw[1]), MulRowAdd(Rcw + 6, xw, tcw[2])};
            if (p3Dc[2] < 0.0) continue;                                         // :350
            const float invz = 1 / p3Dc[2];
            const float x = p3Dc[0] * invz, y = p3Dc[1] * invz;
            const float u = fx * x + cx, v = fy * y + cy;
            if (!pKF->IsInImage(u, v)) continue;
            const float maxDistance = pMP->GetMaxDistanceInvariance(), minDistance = pMP->GetMinDistanceInvariance();
            const float PO[3] = {xw[0] - Ow[0], xw[1] - Ow[1], xw[2] - Ow[2]};
            const float dist = Norm3(PO[0], PO[1], PO[2]);
            if (dist < minDistance || dist > maxDistance) continue;
            const cv::Mat Pn = pMP->GetNormal();                                 // viewing angle below 60 degrees (:374-377)
            if (Dot3(PO, Pn) < 0.5 * dist) continue;
            const int lvl = pMP->PredictScale(dist, pKF);
            // GetFeaturesInArea(u, v, radius) has no level filter; the level window [lvl-1, lvl] is applied to the candidates (:393-395)
            pv.Set(i, u, v, 0.f, th * pKF->mvScaleFactors[lvl], lvl - 1, lvl, pMP->GetDescriptor(), 0.f);
        }
        std::vector<unsigned char> occ(fa.n > 0 ? fa.n : 1, 0);
        for (int i = 0; i < fa.n; ++i) occ[i] = vpMatched[i] != NULL;            // :391
        std::vector<int32_t> match(fa.n > 0 ? fa.n : 1, -1);
        int32_t nmatches = 0;
        Check(orb_search_by_projection_claiming(Handle(), &fa.view, pv.View(false), occ.data(), TH_LOW, 0, match.data(), NULL, &nmatches));
        for (int i = 0; i < fa.n; ++i) if (match[i] >= 0) vpMatched[i] = vpPoints[match[i]];   // :412
        return nmatches;
    }

    // Project MapPoints into KeyFrame and search for duplicated MapPoints.    (src/ORBmatcher.cc:903-1054)
    template <class KeyFrameT, class MapPointT>
    int Fuse(KeyFrameT* pKF, const std::vector<MapPointT*>& vpMapPoints, const float th = 3.0) {
        FrameArrays fa; GatherFrame(*pKF, fa);
        const cv::Mat mR = pKF->GetRotation(), mt = pKF->GetTranslation(), mO = pKF->GetCameraCenter();
        float Rcw[9], tcw[3], Ow[3];
        for (int r = 0; r < 3; ++r) { for (int k = 0; k < 3; ++k) Rcw[3 * r + k] = mR.template at<float>(r, k); tcw[r] = mt.template at<float>(r); Ow[r] = mO.template at<float>(r); }
        const float fx = pKF->fx, fy = pKF->fy, cx = pKF->cx, cy = pKF->cy, bf = pKF->mbf;
        const int M = (int)vpMapPoints.size();
        Projected pv(M);
        for (int i = 0; i < M; ++i) {
            MapPointT* pMP = vpMapPoints[i];
            if (!pMP) continue;
            if (pMP->isBad() || pMP->IsInKeyFrame(pKF)) continue;                // :927
            float u, v, invz, dist3D;
            if (!ProjectAndGate(pKF, pMP, Rcw, tcw, Ow, fx, fy, cx, cy, u, v, invz, dist3D)) continue;
            const float ur = u - bf * invz;
            const int lvl = pMP->PredictScale(dist3D, pKF);
            pv.Set(i, u, v, ur, th * pKF->mvScaleFactors[lvl], lvl - 1, lvl, pMP->GetDescriptor(), 0.f);
        }
        std::vector<int32_t> best(M > 0 ? M : 1, -1);
        int32_t nfound = 0;
        Check(orb_search_projected_best(Handle(), &fa.view, pv.View(false), pKF->mvInvLevelSigma2.data(), TH_LOW, best.data(), NULL, &nfound));
        // the map updates touch host-side state: replayed in point order (:1042-1070).  A point that an earlier
        // update made bad or put into the keyframe is skipped, as the reference's gate at :927 would.
        int nFused = 0;
        for (int i = 0; i < M; ++i) {
            if (best[i] < 0) continue;
            MapPointT* pMP = vpMapPoints[i];
            if (pMP->isBad() || pMP->IsInKeyFrame(pKF)) continue;
            MapPointT* pMPinKF = pKF->GetMapPoint(best[i]);
            if (pMPinKF) {
                if (!pMPinKF->isBad()) {
                    if (pMPinKF->Observations() > pMP->Observations()) pMP->Replace(pMPinKF);
                    else pMPinKF->Replace(pMP);
                }
            } else {
                pMP->AddObservation(pKF, best[i]);
                pKF->AddMapPoint(pMP, best[i]);
            }
            nFused++;
        }
        return nFused;
    }

    // Project MapPoints into KeyFrame using a given Sim3 and search for duplicated MapPoints.   (src/ORBmatcher.cc:1056-1207)
    template <class KeyFrameT, class MapPointT>
    int Fuse(KeyFrameT* pKF, cv::Mat Scw, const std::vector<MapPointT*>& vpPoints, float th, std::vector<MapPointT*>& vpReplacePoint) {
        FrameArrays fa; GatherFrame(*pKF, fa);
        const float fx = pKF->fx, fy = pKF->fy, cx = pKF->cx, cy = pKF->cy;
        float Rcw[9], tcw[3], Ow[3];
        DecomposeSim3(Scw, Rcw, tcw);
        CameraCentre(Rcw, tcw, Ow);
        const std::set<MapPointT*> spAlreadyFound = pKF->GetMapPoints();
        const int M = (int)vpPoints.size();
        Projected pv(M);
        for (int i = 0; i < M; ++i) {
            MapPointT* pMP = vpPoints[i];
            if (pMP->isBad() || spAlreadyFound.count(pMP)) continue;             // :1087
            float u, v, invz, dist3D;
            if (!ProjectAndGate(pKF, pMP, Rcw, tcw, Ow, fx, fy, cx, cy, u, v, invz, dist3D)) continue;
            const int lvl = pMP->PredictScale(dist3D, pKF);
            pv.Set(i, u, v, 0.f, th * pKF->mvScaleFactors[lvl], lvl - 1, lvl, pMP->GetDescriptor(), 0.f);
        }
        std::vector<int32_t> best(M > 0 ? M : 1, -1);
        int32_t nfound = 0;
        Check(orb_search_projected_best(Handle(), &fa.view, pv.View(false), NULL, TH_LOW, best.data(), NULL, &nfound));
        int nFused = 0;
        for (int i = 0; i < M; ++i) {   // :1184-1202
            if (best[i] < 0) continue;
            MapPointT* pMP = vpPoints[i];
            MapPointT* pMPinKF = pKF->GetMapPoint(best[i]);
            if (pMPinKF) {
                if (!pMPinKF->isBad()) vpReplacePoint[i] = pMPinKF;
            } else {
                pMP->AddObservation(pKF, best[i]);
                pKF->AddMapPoint(pMP, best[i]);
            }
            nFused++;
        }
        return nFused;
    }

    // Search matches between MapPoints seen in KF1 and KF2 transforming by a Sim3 [s12*R12|t12].
    // In the stereo and RGB-D case, s12 = 1                                (src/ORBmatcher.cc:1212-1455)
    template <class KeyFrameT, class MapPointT>
    int SearchBySim3(KeyFrameT* pKF1, KeyFrameT* pKF2, std::vector<MapPointT*>& vpMatches12, const float& s12, const cv::Mat& R12,
                     const cv::Mat& t12, const float th) {
        const float fx = pKF1->fx, fy = pKF1->fy, cx = pKF1->cx, cy = pKF1->cy;    // :1216-1219 (KF1's intrinsics for both directions)
        float R1w[9], t1w[3], R2w[9], t2w[3], sR12[9], sR21[9], t12f[3], t21[3];
        { const cv::Mat R = pKF1->GetRotation(), t = pKF1->GetTranslation(); for (int r = 0; r < 3; ++r) { for (int k = 0; k < 3; ++k) R1w[3 * r + k] = R.template at<float>(r, k); t1w[r] = t.template at<float>(r); } }
        { const cv::Mat R = pKF2->GetRotation(), t = pKF2->GetTranslation(); for (int r = 0; r < 3; ++r) { for (int k = 0; k < 3; ++k) R2w[3 * r + k] = R.template at<float>(r, k); t2w[r] = t.template at<float>(r); } }
        const float inv_s = (float)(1.0 / s12);
        for (int r = 0; r < 3; ++r) {
            for (int k = 0; k < 3; ++k) { sR12[3 * r + k] = s12 * R12.at<float>(r, k); sR21[3 * r + k] = inv_s * R12.at<float>(k, r); }   // :1229-1230
            t12f[r] = t12.at<float>(r);
        }
        for (int r = 0; r < 3; ++r) t21[r] = -MulRow(sR21 + 3 * r, t12f);   // :1231 (-sR21 * t12: plain 3x3 * 3x1 product, float)

        const std::vector<MapPointT*> vpMapPoints1 = pKF1->GetMapPointMatches(), vpMapPoints2 = pKF2->GetMapPointMatches();
        const int N1 = (int)vpMapPoints1.size(), N2 = (int)vpMapPoints2.size();
        std::vector<bool> vbAlreadyMatched1(N1, false), vbAlreadyMatched2(N2, false);
        for (int i = 0; i < N1; ++i) {                                           // :1242-1253
            MapPointT* pMP = vpMatches12[i];
            if (pMP) {
                vbAlreadyMatched1[i] = true;
                const int idx2 = pMP->GetIndexInKeyFrame(pKF2);
                if (idx2 >= 0 && idx2 < N2) vbAlreadyMatched2[idx2] = true;
            }
        }
        FrameArrays a1, a2; GatherFrame(*pKF1, a1); GatherFrame(*pKF2, a2);
        Projected p1(N1), p2(N2);
        for (int dir = 0; dir < 2; ++dir) {
            // dir 0: MapPoints of KF1 into KF2 (:1259-1334); dir 1: MapPoints of KF2 into KF1 (:1337-1410)
            const std::vector<MapPointT*>& src = dir == 0 ? vpMapPoints1 : vpMapPoints2;
            const std::vector<bool>& done = dir == 0 ? vbAlreadyMatched1 : vbAlreadyMatched2;
            const float* Rw = dir == 0 ? R1w : R2w; const float* tw = dir == 0 ? t1w : t2w;
            const float* sR = dir == 0 ? sR21 : sR12; const float* tt = dir == 0 ? t21 : t12f;
            KeyFrameT* dst = dir == 0 ? pKF2 : pKF1;
            Projected& pv = dir == 0 ? p1 : p2;
            for (int i = 0; i < (int)src.size(); ++i) {
                MapPointT* pMP = src[i];
                if (!pMP || done[i]) continue;
                if (pMP->isBad()) continue;
                const cv::Mat p3Dw = pMP->GetWorldPos();
                const float xw[3] = {p3Dw.template at<float>(0), p3Dw.template at<float>(1), p3Dw.template at<float>(2)};
                const float pa[3] = {MulRowAdd(Rw + 0, xw, tw[0]), MulRowAdd(Rw + 3, xw, tw[1]), MulRowAdd(Rw + 6, xw, tw[2])};
                const float pb[3] = {MulRowAdd(sR + 0, pa, tt[0]), MulRowAdd(sR + 3, pa, tt[1]), MulRowAdd(sR + 6, pa, tt[2])};
                if (pb[2] < 0.0) continue;
                const float invz = (float)(1.0 / pb[2]);
                const float x = pb[0] * invz, y = pb[1] * invz;
                const float u = fx * x + cx, v = fy * y + cy;
                if (!dst->IsInImage(u, v)) continue;
                const float dist3D = Norm3(pb[0], pb[1], pb[2]);
                if (dist3D < pMP->GetMinDistanceInvariance() || dist3D > pMP->GetMaxDistanceInvariance()) continue;
                const int lvl = pMP->PredictScale(dist3D, dst);
                pv.Set(i, u, v, 0.f, th * dst->mvScaleFactors[lvl], lvl - 1, lvl, pMP->GetDescriptor(), 0.f);
            }
        }
        std::vector<int32_t> match(N1 > 0 ? N1 : 1, -1);
        int32_t nFound = 0;
        Check(orb_search_by_sim3(Handle(), &a1.view, p1.View(false), &a2.view, p2.View(false), match.data(), &nFound));
        for (int i1 = 0; i1 < N1; ++i1) if (match[i1] >= 0) vpMatches12[i1] = vpMapPoints2[match[i1]];   // :1441-1446
        return nFound;
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
    // flat arrays behind an orb_projected_view
    struct Projected {
        std::vector<unsigned char> valid, desc;
        std::vector<float> u, v, ur, radius, angle;
        std::vector<int32_t> lo, hi;
        orb_projected_view view;
        explicit Projected(int n) : valid(n > 0 ? n : 1, 0), desc((size_t)(n > 0 ? n : 1) * 32, 0), u(valid.size(), 0.f), v(valid.size(), 0.f), ur(valid.size(), 0.f),
                                    radius(valid.size(), 0.f), angle(valid.size(), 0.f), lo(valid.size(), 0), hi(valid.size(), -1) { view.n = n; }
        void Set(int i, float pu, float pv, float pur, float r, int lmin, int lmax, const cv::Mat& d, float ang) {
            valid[i] = 1; u[i] = pu; v[i] = pv; ur[i] = pur; radius[i] = r; lo[i] = lmin; hi[i] = lmax; angle[i] = ang;
            std::memcpy(&desc[(size_t)i * 32], d.ptr<unsigned char>(0), 32);
        }
        const orb_projected_view* View(bool with_angle) {
            view.valid = valid.data(); view.u = u.data(); view.v = v.data(); view.ur = ur.data(); view.radius = radius.data();
            view.min_level = lo.data(); view.max_level = hi.data(); view.desc = desc.data(); view.angle = with_angle ? angle.data() : NULL;
            return &view;
        }
    };
    // Ow = -Rcw.t() * tcw on CV_32F data (gemm accumulates in double)
    static void CameraCentre(const float* R, const float* t, float* Ow) {
        for (int r = 0; r < 3; ++r) Ow[r] = (float)(-((double)R[0 * 3 + r] * t[0] + (double)R[1 * 3 + r] * t[1] + (double)R[2 * 3 + r] * t[2]));
    }
    // Shared gate of Fuse / Fuse(Scw) (:932-972, :1091-1127): positive depth, inside the image, distance inside the
    // scale-invariance region, viewing angle below 60 degrees.
    template <class KeyFrameT, class MapPointT>
    static bool ProjectAndGate(KeyFrameT* pKF, MapPointT* pMP, const float* Rcw, const float* tcw, const float* Ow, float fx, float fy, float cx,
                               float cy, float& u, float& v, float& invz, float& dist3D) {
        const cv::Mat p3Dw = pMP->GetWorldPos();
        const float xw[3] = {p3Dw.template at<float>(0), p3Dw.template at<float>(1), p3Dw.template at<float>(2)};
        const float p3Dc[3] = {MulRowAdd(Rcw + 0, xw, tcw[0]), MulRowAdd(Rcw + 3, xw, tcw[1]), MulRowAdd(Rcw + 6, xw, tcw[2])};
        if (p3Dc[2] < 0.0f) return false;
        invz = 1 / p3Dc[2];
        const float x = p3Dc[0] * invz, y = p3Dc[1] * invz;
        u = fx * x + cx; v = fy * y + cy;
        if (!pKF->IsInImage(u, v)) return false;
        const float maxDistance = pMP->GetMaxDistanceInvariance(), minDistance = pMP->GetMinDistanceInvariance();
        const float PO[3] = {xw[0] - Ow[0], xw[1] - Ow[1], xw[2] - Ow[2]};
        dist3D = Norm3(PO[0], PO[1], PO[2]);
        if (dist3D < minDistance || dist3D > maxDistance) return false;
        const cv::Mat Pn = pMP->GetNormal();
        if (Dot3(PO, Pn) < 0.5 * dist3D) return false;
        return true;
    }
    // Scw -> rotation and translation without the scale (:320-324).  Mat / float is Mat * (1.0 / s) with float elements.
    static void DecomposeSim3(const cv::Mat& Scw, float* R, float* t, float* scale = NULL) {
        float sR[9], st[3];
        SplitPose(Scw, sR, st);
        const float scw = (float)std::sqrt((double)sR[0] * sR[0] + (double)sR[1] * sR[1] + (double)sR[2] * sR[2]);
        const float inv = (float)(1.0 / (double)scw);
        for (int k = 0; k < 9; ++k) R[k] = sR[k] * inv;
        for (int k = 0; k < 3; ++k) t[k] = st[k] * inv;
        if (scale) *scale = scw;
    }
    // a.dot(b) on 3 x 1 CV_32F: accumulated in double
    static double Dot3(const float* a, const cv::Mat& b) {
        return (double)a[0] * b.at<float>(0) + (double)a[1] * b.at<float>(1) + (double)a[2] * b.at<float>(2);
    }
    // (float)cv::norm(3 x 1 CV_32F): sum of squares in double, sqrt in double
    static float Norm3(float a, float b, float c) { return (float)std::sqrt((double)a * a + (double)b * b + (double)c * c); }
    static void SplitPose(const cv::Mat& Tcw, float* R, float* t) {
        for (int r = 0; r < 3; ++r) { for (int k = 0; k < 3; ++k) R[3 * r + k] = Tcw.at<float>(r, k); t[r] = Tcw.at<float>(r, 3); }
    }
    // one row of R * x + t on CV_32F data.  cv::gemm without transpose flags and inner length 2..4 (3x3 * 3x1)
    // takes its unrolled path: float products summed left to right in float, then "+ t" (probed against
    // OpenCV 4.13; products that carry a transpose flag, e.g. -R.t() * t, accumulate in double instead).
    // Each product goes through double (exact) and back, so no compiler can contract it into an FMA.
    static float MulRow(const float* row, const float* x) {
        const float p0 = (float)((double)row[0] * x[0]), p1 = (float)((double)row[1] * x[1]), p2 = (float)((double)row[2] * x[2]);
        return (p0 + p1) + p2;
    }
    static float MulRowAdd(const float* row, const float* x, float t) { return MulRow(row, x) + t; }

    float mfNNratio;
    bool mbCheckOrientation;
};

}  // namespace ORB_SLAM2

#endif
