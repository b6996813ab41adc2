// Drives the C++ drop-in classes (include/ORBextractor.h, include/ORBmatcher.h) the way the
// reference's callers do, on scenarios written by tests/test_gpu_cpp_dropin.py, and writes the
// results back for comparison with the oracle.  Frame / KeyFrame / MapPoint are small structs with
// the reference's member names (include/Frame.h, include/KeyFrame.h, include/MapPoint.h).
//   test_dropin extract <in> <out> | mappoints <in> <out> | lastframe <in> <out> | bow <in> <out> | tri <in> <out>
#include <cstdio>
#include <cmath>
#include <map>
#include <set>
#include <string>
#include <thread>
#include <vector>

#include "ORBextractor.h"
#include "ORBmatcher.h"
#include "ORBVocabulary.h"
#include "blobio.h"

struct Frame;
struct MapPoint {
    bool mbTrackInView; float mTrackProjX, mTrackProjY, mTrackProjXR; int mnTrackScaleLevel; float mTrackViewCos;
    cv::Mat mDescriptor, mWorldPos, mNormal; int nObs; bool mbBad; int id;
    float mfMaxDistance, mfMinDistance;
    std::map<Frame*, size_t> mObservations;
    MapPoint* mpReplaced;
    MapPoint() : mbTrackInView(false), nObs(0), mbBad(false), id(-1), mfMaxDistance(0.f), mfMinDistance(0.f), mpReplaced(NULL) {}
    bool isBad() { return mbBad; }
    cv::Mat GetDescriptor() { return mDescriptor.clone(); }
    cv::Mat GetWorldPos() { return mWorldPos.clone(); }
    cv::Mat GetNormal() { return mNormal.clone(); }
    int Observations() { return nObs; }
    float GetMinDistanceInvariance() { return 0.8f * mfMinDistance; }   // src/MapPoint.cc:404-414
    float GetMaxDistanceInvariance() { return 1.2f * mfMaxDistance; }
    inline int PredictScale(const float& currentDist, Frame* pF);        // src/MapPoint.cc:417-450
    bool IsInKeyFrame(Frame* pKF) { return mObservations.count(pKF) != 0; }
    int GetIndexInKeyFrame(Frame* pKF) { return mObservations.count(pKF) ? (int)mObservations[pKF] : -1; }
    void AddObservation(Frame* pKF, size_t idx) { if (!mObservations.count(pKF)) { mObservations[pKF] = idx; nObs++; } }
    // src/MapPoint.cc:158-211 reduced to what Fuse observes afterwards: the replaced point goes bad
    void Replace(MapPoint* pMP) { if (pMP == this) return; mbBad = true; mpReplaced = pMP; }
};

struct Frame {  // also stands in for KeyFrame
    int N;
    std::vector<cv::KeyPoint> mvKeys, mvKeysUn;
    std::vector<float> mvuRight, mvScaleFactors, mvLevelSigma2, mvInvLevelSigma2;
    cv::Mat mDescriptors, mTcw;
    std::vector<MapPoint*> mvpMapPoints;
    std::vector<bool> mvbOutlier;
    float mnMinX, mnMinY, mnMaxX, mnMaxY, mfGridElementWidthInv, mfGridElementHeightInv;
    float fx, fy, cx, cy, mbf, mb;
    float mfLogScaleFactor; int mnScaleLevels;
    std::map<unsigned, std::vector<unsigned> > mFeatVec;
    cv::Mat mOw;
    std::vector<MapPoint*> GetMapPointMatches() { return mvpMapPoints; }
    std::set<MapPoint*> GetMapPoints() { std::set<MapPoint*> s; for (size_t i = 0; i < mvpMapPoints.size(); ++i) if (mvpMapPoints[i] && !mvpMapPoints[i]->isBad()) s.insert(mvpMapPoints[i]); return s; }
    MapPoint* GetMapPoint(int i) { return mvpMapPoints[i]; }
    void AddMapPoint(MapPoint* pMP, size_t idx) { mvpMapPoints[idx] = pMP; }
    bool IsInImage(const float& x, const float& y) const { return (x >= mnMinX && x < mnMaxX && y >= mnMinY && y < mnMaxY); }
    cv::Mat GetCameraCenter() { return mOw.clone(); }
    cv::Mat GetRotation() { cv::Mat R(3, 3, CV_32F); for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) R.at<float>(r, c) = mTcw.at<float>(r, c); return R; }
    cv::Mat GetTranslation() { cv::Mat t(3, 1, CV_32F); for (int r = 0; r < 3; ++r) t.at<float>(r, 0) = mTcw.at<float>(r, 3); return t; }
};

inline int MapPoint::PredictScale(const float& currentDist, Frame* pF) {
    const float ratio = mfMaxDistance / currentDist;
    int nScale = (int)std::ceil(std::log(ratio) / pF->mfLogScaleFactor);
    if (nScale < 0) nScale = 0;
    else if (nScale >= pF->mnScaleLevels) nScale = pF->mnScaleLevels - 1;
    return nScale;
}

static cv::Mat Identity4() { cv::Mat T = cv::Mat::zeros(4, 4, CV_32F); for (int i = 0; i < 4; ++i) T.at<float>(i, i) = 1.f; return T; }

static void LoadFrame(const Blob& b, const std::string& p, Frame& F) {
    F.N = (int)b.count(p + "x");
    const float* x = b.get<float>(p + "x"); const float* y = b.get<float>(p + "y"); const float* ang = b.get<float>(p + "angle");
    const int* oct = b.get<int>(p + "octave"); const float* ur = b.get<float>(p + "u_right");
    F.mvKeysUn.resize(F.N); F.mvuRight.assign(ur, ur + F.N);
    for (int i = 0; i < F.N; ++i) F.mvKeysUn[i] = cv::KeyPoint(x[i], y[i], 31.f, ang[i], 0.f, oct[i]);
    F.mvKeys = F.mvKeysUn;
    F.mDescriptors.create(F.N, 32, CV_8U);
    if (F.N) std::memcpy(F.mDescriptors.data, b.get<unsigned char>(p + "desc"), (size_t)F.N * 32);
    const float* sf = b.get<float>(p + "sf"); const size_t nl = b.count(p + "sf");
    F.mvScaleFactors.assign(sf, sf + nl);
    F.mvLevelSigma2.resize(nl); for (size_t l = 0; l < nl; ++l) F.mvLevelSigma2[l] = sf[l] * sf[l];
    F.mvInvLevelSigma2.resize(nl); for (size_t l = 0; l < nl; ++l) F.mvInvLevelSigma2[l] = 1.0f / F.mvLevelSigma2[l];
    F.mnScaleLevels = (int)nl; F.mfLogScaleFactor = std::log(nl > 1 ? sf[1] : 1.2f);
    const float* g = b.get<float>(p + "grid");  // min_x, min_y, max_x, max_y, inv_w, inv_h
    F.mnMinX = g[0]; F.mnMinY = g[1]; F.mnMaxX = g[2]; F.mnMaxY = g[3]; F.mfGridElementWidthInv = g[4]; F.mfGridElementHeightInv = g[5];
    F.mvpMapPoints.assign(F.N, (MapPoint*)NULL); F.mvbOutlier.assign(F.N, false);
    F.mTcw = Identity4(); F.mOw = cv::Mat::zeros(3, 1, CV_32F);
    F.fx = F.fy = 1.f; F.cx = F.cy = 0.f; F.mbf = 0.f; F.mb = 0.f;
    if (b.has(p + "fv_ids")) {
        const unsigned* ids = b.get<unsigned>(p + "fv_ids"); const int* st = b.get<int>(p + "fv_start"); const unsigned* it = b.get<unsigned>(p + "fv_items");
        for (size_t k = 0; k < b.count(p + "fv_ids"); ++k) F.mFeatVec[ids[k]] = std::vector<unsigned>(it + st[k], it + st[k + 1]);
    }
}

static cv::Mat Row32(const unsigned char* p);
// MapPoints of a world scenario: position, normal, distance range, descriptor, flags
static void LoadPoints(const Blob& b, const std::string& p, std::vector<MapPoint>& store, std::vector<MapPoint*>& ptrs) {
    const int M = (int)b.count(p + "maxd");
    store.assign(M, MapPoint()); ptrs.assign(M, (MapPoint*)NULL);
    for (int i = 0; i < M; ++i) {
        MapPoint& m = store[i];
        m.id = i;
        m.mWorldPos.create(3, 1, CV_32F); m.mNormal.create(3, 1, CV_32F);
        for (int k = 0; k < 3; ++k) { m.mWorldPos.at<float>(k) = b.get<float>(p + "xw")[3 * i + k]; m.mNormal.at<float>(k) = b.get<float>(p + "normal")[3 * i + k]; }
        m.mfMaxDistance = b.get<float>(p + "maxd")[i]; m.mfMinDistance = b.get<float>(p + "mind")[i];
        m.mDescriptor = Row32(b.get<unsigned char>(p + "desc") + (size_t)i * 32);
        m.mbBad = b.get<unsigned char>(p + "bad")[i] != 0;
        m.nObs = b.get<int>(p + "nobs")[i];
        if (!b.get<unsigned char>(p + "null")[i]) ptrs[i] = &m;
    }
}
static void SetPose(Frame& F, const float* T, const float* cam) {
    F.mTcw.create(4, 4, CV_32F); std::memcpy(F.mTcw.data, T, 64);
    F.fx = cam[0]; F.fy = cam[1]; F.cx = cam[2]; F.cy = cam[3]; F.mbf = cam[4];
    // Ow = -R^T t like KeyFrame::SetPose (src/KeyFrame.cc:67-84), float products accumulated in double
    F.mOw.create(3, 1, CV_32F);
    for (int r = 0; r < 3; ++r)
        F.mOw.at<float>(r) = (float)(-((double)T[0 * 4 + r] * T[3] + (double)T[1 * 4 + r] * T[7] + (double)T[2 * 4 + r] * T[11]));
}

static cv::Mat Row32(const unsigned char* p) { cv::Mat d(1, 32, CV_8U); std::memcpy(d.data, p, 32); return d; }

int main(int argc, char** argv) {
    if (argc < 4) { fprintf(stderr, "usage: test_dropin <mode> <in> <out>\n"); return 2; }
    const std::string mode = argv[1];
    try {
        Blob b; b.load(argv[2]);
        BlobWriter w(argv[3]);
        if (mode == "extract") {
            const int* p = b.get<int>("params");  // w, h, nfeatures, nlevels, iniTh, minTh
            const float sf = b.get<float>("scale")[0];
            cv::Mat img(p[1], p[0], CV_8UC1, const_cast<unsigned char*>(b.get<unsigned char>("image")));
            ORB_SLAM2::ORBextractor ext(p[2], sf, p[3], p[4], p[5]);   // mvImagePyramid: lazy by default, fetched by the read below
            std::vector<cv::KeyPoint> kps; cv::Mat desc;
            ext(img, cv::Mat(), kps, desc);
            // second call on the same instance, as Tracking does frame after frame
            std::vector<cv::KeyPoint> kps2; cv::Mat desc2;
            ext(img, cv::Mat(), kps2, desc2);
            if (kps2.size() != kps.size()) throw std::runtime_error("second call differs");
            w.put("kps", "u1", kps.data(), kps.size() * sizeof(cv::KeyPoint));
            w.put("desc", "u1", desc.data, (size_t)desc.rows * 32);
            std::vector<float> tabs = ext.GetScaleFactors();
            w.put("sf", "f4", tabs.data(), tabs.size());
            const cv::Mat& l3 = ext.mvImagePyramid[3];
            int dims[2] = {l3.cols, l3.rows};
            w.put("l3_dims", "i4", dims, 2);
            w.put("l3", "u1", l3.data, (size_t)l3.rows * l3.cols);
            // batch entry point
            std::vector<cv::Mat> imgs(3, img); std::vector<std::vector<cv::KeyPoint> > bk; std::vector<cv::Mat> bd;
            ext.ExtractBatch(imgs, bk, bd);
            int same = bk.size() == 3 && bk[2].size() == kps.size() && std::memcmp(bd[2].data, desc.data, (size_t)desc.rows * 32) == 0;
            w.put("batch_same", "i4", &same, 1);
            // empty image: outputs untouched (:1095)
            std::vector<cv::KeyPoint> ke(7); cv::Mat de;
            ext(cv::Mat(), cv::Mat(), ke, de);
            int untouched = ke.size() == 7;
            w.put("empty_untouched", "i4", &untouched, 1);
        } else if (mode == "stereo" || mode == "stereo1") {
            // The stereo Frame constructor as the reference runs it (src/Frame.cc:81-95): ExtractORB left / right on two threads, joined,
            // then ComputeStereoMatches (= orb_stereo_match on the two handles, patches/Frame_ComputeStereoMatches.patch).  Output names
            // and layout of oracle/ref_match_driver.cpp's "stereo" mode (the reference's own code on the same input).
            const int* p = b.get<int>("params");
            const float sf = b.get<float>("scale")[0];
            const float mbf = b.get<float>("stereo")[0], fx = b.get<float>("stereo")[1];
            ORB_SLAM2::ORBextractor eL(p[2], sf, p[3], p[4], p[5]), eR(p[2], sf, p[3], p[4], p[5]);
            cv::Mat imL(p[1], p[0], CV_8UC1, const_cast<unsigned char*>(b.get<unsigned char>("left")));
            cv::Mat imR(p[1], p[0], CV_8UC1, const_cast<unsigned char*>(b.get<unsigned char>("right")));
            std::vector<cv::KeyPoint> kl, kr; cv::Mat dl, dr;
            const float mb = mbf / fx, maxd = mbf / mb;   // :117, :529-531
            std::vector<float> uR, depth;
            if (mode == "stereo1") {   // the same front end as ONE call from one thread (ORBextractor::ExtractStereo)
                ORB_SLAM2::ORBextractor::ExtractStereo(eL, eR, imL, imR, mbf, 0.f, maxd, kl, dl, kr, dr, uR, depth);
            } else {
                std::thread tl([&] { eL(imL, cv::Mat(), kl, dl); });
                std::thread tr([&] { eR(imR, cv::Mat(), kr, dr); });
                tl.join(); tr.join();
            }
            const int N = (int)kl.size();
            if (mode == "stereo") { uR.assign(N > 0 ? N : 1, -1.f); depth.assign(N > 0 ? N : 1, -1.f); }
            int32_t n = 0;
            if (mode == "stereo" && N > 0 && orb_stereo_match(eL.Handle(), eR.Handle(), 0, mbf, 0.f, maxd, uR.data(), depth.data(), N, &n) != ORB_OK) throw std::runtime_error(orb_last_error());
            struct RefKp { float x, y, size, angle, response; int32_t octave; };
            std::vector<RefKp> ol(kl.size()), orr(kr.size());
            for (size_t i = 0; i < kl.size(); ++i) { const RefKp q = {kl[i].pt.x, kl[i].pt.y, kl[i].size, kl[i].angle, kl[i].response, kl[i].octave}; ol[i] = q; }
            for (size_t i = 0; i < kr.size(); ++i) { const RefKp q = {kr[i].pt.x, kr[i].pt.y, kr[i].size, kr[i].angle, kr[i].response, kr[i].octave}; orr[i] = q; }
            w.put("kl", "u1", ol.data(), ol.size() * sizeof(RefKp)); w.put("dl", "u1", dl.data, (size_t)dl.rows * 32);
            w.put("kr", "u1", orr.data(), orr.size() * sizeof(RefKp)); w.put("dr", "u1", dr.data, (size_t)dr.rows * 32);
            w.put("u_right", "f4", uR.data(), (size_t)N); w.put("depth", "f4", depth.data(), (size_t)N);
            w.put("max_d", "f4", &maxd, 1);
        } else if (mode == "mappoints") {
            Frame F; LoadFrame(b, "f_", F);
            const int M = (int)b.count("mp_valid");
            std::vector<MapPoint> store(M); std::vector<MapPoint*> mps(M);
            const unsigned char* valid = b.get<unsigned char>("mp_valid"); const unsigned char* blocks = b.get<unsigned char>("mp_blocks");
            for (int i = 0; i < M; ++i) {
                MapPoint& m = store[i];
                m.mbTrackInView = valid[i] != 0; m.mbBad = false; m.id = i;
                m.mTrackProjX = b.get<float>("mp_proj_x")[i]; m.mTrackProjY = b.get<float>("mp_proj_y")[i]; m.mTrackProjXR = b.get<float>("mp_proj_xr")[i];
                m.mnTrackScaleLevel = b.get<int>("mp_level")[i]; m.mTrackViewCos = b.get<float>("mp_view_cos")[i];
                m.mDescriptor = Row32(b.get<unsigned char>("mp_desc") + (size_t)i * 32); m.nObs = blocks[i] ? 1 : 0;
                mps[i] = &m;
            }
            // keypoints that already hold an observed MapPoint (:90-93)
            MapPoint holder; holder.nObs = 3; holder.id = -2;
            const unsigned char* occ = b.get<unsigned char>("occupied");
            for (int i = 0; i < F.N; ++i) if (occ[i]) F.mvpMapPoints[i] = &holder;
            ORB_SLAM2::ORBmatcher matcher(b.get<float>("ratio")[0], true);
            const int n = matcher.SearchByProjection(F, mps, b.get<float>("th")[0]);
            std::vector<int> match(F.N, -1);
            for (int i = 0; i < F.N; ++i) if (F.mvpMapPoints[i] && F.mvpMapPoints[i] != &holder) match[i] = F.mvpMapPoints[i]->id;
            w.put("n", "i4", &n, 1); w.put("match", "i4", match.data(), match.size());
        } else if (mode == "lastframe") {
            Frame Cur, Last; LoadFrame(b, "f_", Cur);
            const int M = (int)b.count("l_valid");
            Last = Cur; Last.N = M; Last.mvKeys.assign(M, cv::KeyPoint()); Last.mvKeysUn.assign(M, cv::KeyPoint());
            Last.mvpMapPoints.assign(M, (MapPoint*)NULL); Last.mvbOutlier.assign(M, false);
            std::vector<MapPoint> store(M);
            for (int i = 0; i < M; ++i) {
                MapPoint& m = store[i];
                m.id = i; m.mbBad = false; m.nObs = b.get<unsigned char>("l_blocks")[i] ? 1 : 0;
                m.mDescriptor = Row32(b.get<unsigned char>("l_desc") + (size_t)i * 32);
                m.mWorldPos.create(3, 1, CV_32F);
                m.mWorldPos.at<float>(0) = b.get<float>("l_xc")[i]; m.mWorldPos.at<float>(1) = b.get<float>("l_yc")[i]; m.mWorldPos.at<float>(2) = b.get<float>("l_zc")[i];
                Last.mvKeys[i].octave = b.get<int>("l_octave")[i]; Last.mvKeysUn[i].angle = b.get<float>("l_angle")[i];
                if (b.get<unsigned char>("l_valid")[i]) Last.mvpMapPoints[i] = &m;
            }
            const float* cam = b.get<float>("cam");
            Cur.fx = cam[0]; Cur.fy = cam[1]; Cur.cx = cam[2]; Cur.cy = cam[3]; Cur.mbf = cam[4]; Cur.mb = 1e9f;  // identity poses: neither forward nor backward
            ORB_SLAM2::ORBmatcher matcher(0.9f, b.get<int>("check_ori")[0] != 0);
            const int n = matcher.SearchByProjection(Cur, Last, b.get<float>("th")[0], false);
            std::vector<int> match(Cur.N, -1);
            for (int i = 0; i < Cur.N; ++i) if (Cur.mvpMapPoints[i]) match[i] = Cur.mvpMapPoints[i]->id;
            w.put("n", "i4", &n, 1); w.put("match", "i4", match.data(), match.size());
        } else if (mode == "bow" || mode == "tri") {
            Frame A, B; LoadFrame(b, "a_", A); LoadFrame(b, "b_", B);
            std::vector<MapPoint> sa(A.N), sb(B.N);
            for (int i = 0; i < A.N; ++i) { sa[i].id = i; sa[i].mbBad = false; if (b.get<unsigned char>("has_a")[i]) A.mvpMapPoints[i] = &sa[i]; }
            for (int i = 0; i < B.N; ++i) { sb[i].id = i; sb[i].mbBad = false; if (b.get<unsigned char>("has_b")[i]) B.mvpMapPoints[i] = &sb[i]; }
            ORB_SLAM2::ORBmatcher matcher(b.get<float>("ratio")[0], true);
            if (mode == "bow") {
                std::vector<MapPoint*> out;
                const int n = matcher.SearchByBoW(&A, B, out);
                std::vector<int> match(B.N, -1);
                for (int j = 0; j < B.N; ++j) if (out[j]) match[j] = out[j]->id;
                w.put("n", "i4", &n, 1); w.put("match", "i4", match.data(), match.size());
            } else {
                // camera centre of KF1 expressed in KF2 = (ex, ey, 1) with unit intrinsics -> epipole (ex, ey)
                const float* ep = b.get<float>("epipole");
                A.mOw.at<float>(0) = ep[0]; A.mOw.at<float>(1) = ep[1]; A.mOw.at<float>(2) = 1.f;
                cv::Mat F12(3, 3, CV_32F); std::memcpy(F12.data, b.get<float>("f12"), 36);
                std::vector<std::pair<size_t, size_t> > pairs;
                const int n = matcher.SearchForTriangulation(&A, &B, F12, pairs, b.get<int>("only_stereo")[0] != 0);
                std::vector<int> flat;
                for (size_t k = 0; k < pairs.size(); ++k) { flat.push_back((int)pairs[k].first); flat.push_back((int)pairs[k].second); }
                w.put("n", "i4", &n, 1); w.put("pairs", "i4", flat.data(), flat.size());
            }
        } else if (mode == "reloc" || mode == "loop") {
            Frame F; LoadFrame(b, "f_", F);
            SetPose(F, b.get<float>("pose"), b.get<float>("cam"));
            std::vector<MapPoint> store; std::vector<MapPoint*> pts; LoadPoints(b, "p_", store, pts);
            const int M = (int)pts.size();
            MapPoint holder; holder.id = -2; holder.nObs = 3;
            const unsigned char* occ = b.get<unsigned char>("occupied"); const unsigned char* found = b.get<unsigned char>("p_found");
            ORB_SLAM2::ORBmatcher matcher(0.75f, b.get<int>("check_ori")[0] != 0);
            int n;
            std::vector<int> match(F.N, -1);
            if (mode == "reloc") {   // Tracking::Relocalization: matcher.SearchByProjection(mCurrentFrame, vpCandidateKFs[i], sFound, 10, 100)
                for (int i = 0; i < F.N; ++i) if (occ[i]) F.mvpMapPoints[i] = &holder;
                Frame K; K.N = M; K.mvpMapPoints = pts; K.mvKeysUn.assign(M, cv::KeyPoint());
                for (int i = 0; i < M; ++i) K.mvKeysUn[i].angle = b.get<float>("p_angle")[i];
                std::set<MapPoint*> sFound;
                for (int i = 0; i < M; ++i) if (found[i] && pts[i]) sFound.insert(pts[i]);
                n = matcher.SearchByProjection(F, &K, sFound, b.get<float>("th")[0], b.get<int>("max_dist")[0]);
                for (int i = 0; i < F.N; ++i) if (F.mvpMapPoints[i] && F.mvpMapPoints[i] != &holder) match[i] = F.mvpMapPoints[i]->id;
            } else {                 // LoopClosing::ComputeSim3: matcher.SearchByProjection(mpCurrentKF, mScw, mvpLoopMapPoints, mvpCurrentMatchedPoints, 10)
                std::vector<MapPoint*> vpMatched(F.N, (MapPoint*)NULL);
                int nf = 0;
                std::vector<int> foundIdx; for (int i = 0; i < M; ++i) if (found[i]) foundIdx.push_back(i);
                for (int i = 0; i < F.N; ++i) if (occ[i]) vpMatched[i] = (nf < (int)foundIdx.size()) ? pts[foundIdx[nf++]] : &holder;
                std::vector<MapPoint*> before = vpMatched;
                cv::Mat Scw(4, 4, CV_32F); std::memcpy(Scw.data, b.get<float>("pose"), 64);
                n = matcher.SearchByProjection(&F, Scw, pts, vpMatched, (int)b.get<float>("th")[0]);
                for (int i = 0; i < F.N; ++i) if (vpMatched[i] && !before[i]) match[i] = vpMatched[i]->id;
            }
            w.put("n", "i4", &n, 1); w.put("match", "i4", match.data(), match.size());
        } else if (mode == "fuse" || mode == "fuse_scw") {
            // LocalMapping::SearchInNeighbors: matcher.Fuse(pKFi, vpMapPointMatches) (src/LocalMapping.cc:584,613);
            // LoopClosing::SearchAndFuse: matcher.Fuse(pKF, cvScw, mvpLoopMapPoints, 4, vpReplacePoints) (src/LoopClosing.cc:652)
            Frame K; LoadFrame(b, "f_", K);
            SetPose(K, b.get<float>("pose"), b.get<float>("cam"));
            std::vector<MapPoint> store; std::vector<MapPoint*> pts; LoadPoints(b, "p_", store, pts);
            const int M = (int)pts.size();
            // MapPoints the keyframe already holds (e_kp[j] = keypoint index of existing point j)
            const int E = (int)b.count("e_kp");
            std::vector<MapPoint> existing(E);
            for (int j = 0; j < E; ++j) {
                existing[j].id = 100000 + j; existing[j].nObs = b.get<int>("e_nobs")[j]; existing[j].mbBad = b.get<unsigned char>("e_bad")[j] != 0;
                K.mvpMapPoints[b.get<int>("e_kp")[j]] = &existing[j];
            }
            // candidate points that are already observed in the keyframe (:927 IsInKeyFrame / :1087 spAlreadyFound)
            const int* inkf = b.get<int>("p_in_kf");
            for (int i = 0; i < M; ++i) if (inkf[i] >= 0 && pts[i]) { pts[i]->mObservations[&K] = (size_t)inkf[i]; K.mvpMapPoints[inkf[i]] = pts[i]; }
            ORB_SLAM2::ORBmatcher matcher;
            int n;
            std::vector<int> replace(M, -1);
            if (mode == "fuse") n = matcher.Fuse(&K, pts, b.get<float>("th")[0]);
            else {
                cv::Mat Scw(4, 4, CV_32F); std::memcpy(Scw.data, b.get<float>("pose"), 64);
                std::vector<MapPoint*> vpReplace(M, (MapPoint*)NULL);
                n = matcher.Fuse(&K, Scw, pts, b.get<float>("th")[0], vpReplace);
                for (int i = 0; i < M; ++i) if (vpReplace[i]) replace[i] = vpReplace[i]->id;
            }
            std::vector<int> kf_mp(K.N, -1), p_bad(M, 0), p_repl(M, -1), p_idx(M, -1), e_bad(E, 0), e_repl(E, -1);
            for (int i = 0; i < K.N; ++i) if (K.mvpMapPoints[i]) kf_mp[i] = K.mvpMapPoints[i]->id;
            for (int i = 0; i < M; ++i) { p_bad[i] = store[i].mbBad; if (store[i].mpReplaced) p_repl[i] = store[i].mpReplaced->id; p_idx[i] = store[i].GetIndexInKeyFrame(&K); }
            for (int j = 0; j < E; ++j) { e_bad[j] = existing[j].mbBad; if (existing[j].mpReplaced) e_repl[j] = existing[j].mpReplaced->id; }
            w.put("n", "i4", &n, 1); w.put("kf_mp", "i4", kf_mp.data(), kf_mp.size()); w.put("replace", "i4", replace.data(), replace.size());
            w.put("p_bad", "i4", p_bad.data(), p_bad.size()); w.put("p_repl", "i4", p_repl.data(), p_repl.size()); w.put("p_idx", "i4", p_idx.data(), p_idx.size());
            w.put("e_bad", "i4", e_bad.data(), e_bad.size()); w.put("e_repl", "i4", e_repl.data(), e_repl.size());
        } else if (mode == "sim3") {
            // LoopClosing::ComputeSim3: matcher.SearchBySim3(mpCurrentKF, pKF, vpMapPointMatches, s, R, t, 7.5) (src/LoopClosing.cc:348)
            Frame A, B; LoadFrame(b, "a_", A); LoadFrame(b, "b_", B);
            SetPose(A, b.get<float>("pose_a"), b.get<float>("cam")); SetPose(B, b.get<float>("pose_b"), b.get<float>("cam"));
            std::vector<MapPoint> sa, sb; std::vector<MapPoint*> pa, pb;
            LoadPoints(b, "pa_", sa, pa); LoadPoints(b, "pb_", sb, pb);
            A.mvpMapPoints = pa; B.mvpMapPoints = pb;
            for (int j = 0; j < B.N; ++j) if (pb[j]) pb[j]->mObservations[&B] = (size_t)j;
            std::vector<MapPoint*> m12(A.N, (MapPoint*)NULL);
            const int* pre = b.get<int>("pre");
            for (int i = 0; i < A.N; ++i) if (pre[i] >= 0) m12[i] = &sb[pre[i]];
            cv::Mat R12(3, 3, CV_32F), t12(3, 1, CV_32F);
            std::memcpy(R12.data, b.get<float>("r12"), 36); std::memcpy(t12.data, b.get<float>("t12"), 12);
            const float s12 = b.get<float>("s12")[0];
            ORB_SLAM2::ORBmatcher matcher(0.75f, true);
            const int n = matcher.SearchBySim3(&A, &B, m12, s12, R12, t12, b.get<float>("th")[0]);
            std::vector<int> match(A.N, -1);
            for (int i = 0; i < A.N; ++i) if (m12[i]) match[i] = m12[i]->id;
            w.put("n", "i4", &n, 1); w.put("match", "i4", match.data(), match.size());
        } else if (mode == "init") {
            Frame A, B; LoadFrame(b, "a_", A); LoadFrame(b, "b_", B);
            std::vector<cv::Point2f> prev(A.N);
            for (int i = 0; i < A.N; ++i) prev[i] = cv::Point2f(b.get<float>("prev_x")[i], b.get<float>("prev_y")[i]);
            ORB_SLAM2::ORBmatcher matcher(b.get<float>("ratio")[0], b.get<int>("check_ori")[0] != 0);
            std::vector<int> m12;
            const int n = matcher.SearchForInitialization(A, B, prev, m12, b.get<int>("window")[0]);
            std::vector<float> px(A.N), py(A.N);
            for (int i = 0; i < A.N; ++i) { px[i] = prev[i].x; py[i] = prev[i].y; }
            w.put("n", "i4", &n, 1); w.put("match", "i4", m12.data(), m12.size());
            w.put("prev_x", "f4", px.data(), px.size()); w.put("prev_y", "f4", py.data(), py.size());
        } else if (mode == "bowkf") {
            Frame A, B; LoadFrame(b, "a_", A); LoadFrame(b, "b_", B);
            std::vector<MapPoint> sa(A.N), sb(B.N);
            for (int i = 0; i < A.N; ++i) { sa[i].id = i; if (b.get<unsigned char>("has_a")[i]) A.mvpMapPoints[i] = &sa[i]; }
            for (int i = 0; i < B.N; ++i) { sb[i].id = i; if (b.get<unsigned char>("has_b")[i]) B.mvpMapPoints[i] = &sb[i]; }
            ORB_SLAM2::ORBmatcher matcher(b.get<float>("ratio")[0], b.get<int>("check_ori")[0] != 0);
            std::vector<MapPoint*> out;
            const int n = matcher.SearchByBoW(&A, &B, out);
            std::vector<int> match(A.N, -1);
            for (int i = 0; i < A.N; ++i) if (out[i]) match[i] = out[i]->id;
            w.put("n", "i4", &n, 1); w.put("match", "i4", match.data(), match.size());
        } else if (mode == "bowtransform") {
            // Frame::ComputeBoW (src/Frame.cc:439-449): mpORBvocabulary->transform(vCurrentDesc, mBowVec, mFeatVec, 4)
            ORB_SLAM2::ORBVocabulary voc;
            if (argc < 5 || !voc.loadFromTextFile(argv[4])) throw std::runtime_error("vocabulary load failed");
            const int N = (int)b.count("desc") / 32;
            cv::Mat descriptors(N, 32, CV_8U);
            std::memcpy(descriptors.data, b.get<unsigned char>("desc"), (size_t)N * 32);
            std::vector<cv::Mat> vCurrentDesc;                       // Converter::toDescriptorVector (src/Converter.cc:27-35)
            for (int j = 0; j < N; ++j) vCurrentDesc.push_back(descriptors.row(j));
            DBoW2::BowVector bv; DBoW2::FeatureVector fv;
            voc.transform(vCurrentDesc, bv, fv, b.get<int>("levelsup")[0]);
            std::vector<unsigned> bw_ids, fv_ids, fv_items; std::vector<double> bw_vals; std::vector<int> fv_start(1, 0);
            for (DBoW2::BowVector::const_iterator it = bv.begin(); it != bv.end(); ++it) { bw_ids.push_back(it->first); bw_vals.push_back(it->second); }
            for (DBoW2::FeatureVector::const_iterator it = fv.begin(); it != fv.end(); ++it) {
                fv_ids.push_back(it->first);
                for (size_t k = 0; k < it->second.size(); ++k) fv_items.push_back(it->second[k]);
                fv_start.push_back((int)fv_items.size());
            }
            w.put("bow_ids", "u4", bw_ids.data(), bw_ids.size()); w.put("bow_vals", "u1", bw_vals.data(), bw_vals.size() * 8);
            w.put("fv_ids", "u4", fv_ids.data(), fv_ids.size()); w.put("fv_start", "i4", fv_start.data(), fv_start.size()); w.put("fv_items", "u4", fv_items.data(), fv_items.size());
            const int info[2] = {(int)voc.size(), voc.getDepthLevels()};
            w.put("info", "i4", info, 2);
        } else if (mode == "bowscore") {
            // KeyFrameDatabase::DetectLoopCandidates / DetectRelocalizationCandidates: mpVoc->score(pKF->mBowVec, pKFi->mBowVec) (src/KeyFrameDatabase.cc:137,267)
            ORB_SLAM2::ORBVocabulary voc;
            if (argc < 5 || !voc.loadFromTextFile(argv[4])) throw std::runtime_error("vocabulary load failed");
            const ORB_SLAM2::ORBVocabulary* mpVoc = &voc;
            DBoW2::BowVector bv[2]; DBoW2::FeatureVector fv[2];
            const char* names[2] = {"desc_a", "desc_b"};
            for (int s = 0; s < 2; ++s) {
                const int n = (int)b.count(names[s]) / 32;
                cv::Mat d(n, 32, CV_8U);
                std::memcpy(d.data, b.get<unsigned char>(names[s]), (size_t)n * 32);
                std::vector<cv::Mat> rows;
                for (int j = 0; j < n; ++j) rows.push_back(d.row(j));
                mpVoc->transform(rows, bv[s], fv[s], 4);
            }
            const float si = mpVoc->score(bv[0], bv[1]);   // the reference narrows to float at the call site
            (void)si;
            const double sc[4] = {mpVoc->score(bv[0], bv[1]), mpVoc->score(bv[1], bv[0]), mpVoc->score(bv[0], bv[0]), mpVoc->score(bv[1], bv[1])};
            w.put("scores", "u1", sc, sizeof(sc));
            for (int s = 0; s < 2; ++s) {
                std::vector<unsigned> ids; std::vector<double> vals;
                for (DBoW2::BowVector::const_iterator it = bv[s].begin(); it != bv[s].end(); ++it) { ids.push_back(it->first); vals.push_back(it->second); }
                w.put(s ? "ids_b" : "ids_a", "u4", ids.data(), ids.size()); w.put(s ? "vals_b" : "vals_a", "u1", vals.data(), vals.size() * 8);
            }
        } else {
            fprintf(stderr, "unknown mode %s\n", mode.c_str());
            return 2;
        }
    } catch (const std::exception& e) {
        fprintf(stderr, "test_dropin: %s\n", e.what());
        return 1;
    }
    return 0;
}
