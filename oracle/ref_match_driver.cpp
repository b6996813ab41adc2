// TEST INFRASTRUCTURE — drives the REFERENCE'S OWN matcher / stereo code on the scenario files of
// tests/test_gpu_cpp_dropin.py (same modes and blob layout as tests/cpp/test_dropin.cpp), so that
// the oracle's expectations for those scenarios are pinned against the reference itself.
//
// Compiled verbatim, where they lie under /root/reference (oracle/Makefile target `refmatch`):
//   src/ORBmatcher.cc  src/Frame.cc  src/KeyFrame.cc  src/MapPoint.cc  src/Map.cc
//   src/KeyFrameDatabase.cc  src/ORBextractor.cc  Thirdparty/DBoW2/DBoW2/*.cpp  Thirdparty/DBoW2/DUtils/*.cpp
// against oracle/cvshim (container types + cv::Mat algebra with OpenCV 4.13's arithmetic) and the
// reference's own headers.  Objects are real ORB_SLAM2::Frame / KeyFrame / MapPoint / Map instances;
// this file only fills their members from the scenario and reads the results back.  Protected
// members (MapPoint::mDescriptor, Frame::AssignFeaturesToGrid ...) are reached by compiling THIS
// translation unit with the access keywords opened; the reference's translation units are untouched.
//   ref_match_driver <mode> <in> <out>
//
// -DORB_B200_DROPIN (oracle/Makefile target `dropin`, binary _ref/dropin_real_driver): the SAME driver over the SAME real
// Frame / KeyFrame / MapPoint / Map / KeyFrameDatabase translation units, but with this repo's include/ORBextractor.h,
// include/ORBmatcher.h and include/ORBVocabulary.h in place of the reference's headers and liborb_b200.so in place of
// ORBextractor.cc / ORBmatcher.cc / DBoW2's vocabulary (Frame.cc / MapPoint.cc with patches/*.patch).  Needs a GPU; its output files
// must equal this driver's byte for byte (tests/test_gpu_dropin_real.py).
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>
#include <list>
#include <map>
#include <mutex>
#include <set>
#include <sstream>
#include <string>
#include <thread>
#include <vector>
#include <unistd.h>
#include <opencv2/core/core.hpp>

#define private public
#define protected public
#include "ORBmatcher.h"
#include "Map.h"
#include "KeyFrameDatabase.h"
#undef private
#undef protected
#include "Converter.h"
#include "../tests/cpp/blobio.h"

using namespace ORB_SLAM2;

struct ref_keypoint { float x, y, size, angle, response; int32_t octave; };
#ifndef ORB_B200_DROPIN
// oracle/ref_wrap.cpp: the reference extractor under the monotonic heap (canonical quadtree tie-break)
extern "C" void* ref_extractor_create(int nfeatures, float scale_factor, int nlevels, int ini_th, int min_th);
extern "C" int ref_extract(void* e, const uint8_t* img, int w, int h, int stride, ref_keypoint* kps, uint8_t* desc, int cap);
#endif

static Map g_map;
static ORBVocabulary g_voc;
static KeyFrameDatabase* g_db = NULL;

static cv::Mat Identity4() { return cv::Mat::eye(4, 4, CV_32F); }
static cv::Mat Row32(const unsigned char* p) { cv::Mat d(1, 32, CV_8U); std::memcpy(d.data, p, 32); return d; }
static cv::Mat Mat44(const float* T) { cv::Mat m(4, 4, CV_32F); std::memcpy(m.data, T, 64); return m; }

// Frame members from a scenario (prefix p): undistorted keypoints, descriptors, scale tables, grid.
// fx.. and the image bounds are static in the reference (include/Frame.h:80-85,173-179): every frame
// of one scenario shares them, as in the reference.
static void LoadFrame(const Blob& b, const std::string& p, Frame& F) {
    F.N = (int)b.count(p + "x");
    const float* x = b.get<float>(p + "x"); const float* y = b.get<float>(p + "y"); const float* ang = b.get<float>(p + "angle");
    const int* oct = b.get<int>(p + "octave"); const float* ur = b.get<float>(p + "u_right");
    F.mvKeysUn.resize(F.N); F.mvuRight.assign(ur, ur + F.N); F.mvDepth.assign(F.N, -1.f);
    for (int i = 0; i < F.N; ++i) F.mvKeysUn[i] = cv::KeyPoint(x[i], y[i], 31.f, ang[i], 0.f, oct[i]);
    F.mvKeys = F.mvKeysUn;
    F.mDescriptors.create(F.N, 32, CV_8U);
    if (F.N) std::memcpy(F.mDescriptors.data, b.get<unsigned char>(p + "desc"), (size_t)F.N * 32);
    const float* sf = b.get<float>(p + "sf"); const size_t nl = b.count(p + "sf");
    F.mvScaleFactors.assign(sf, sf + nl);
    F.mvInvScaleFactors.resize(nl); F.mvLevelSigma2.resize(nl); F.mvInvLevelSigma2.resize(nl);
    for (size_t l = 0; l < nl; ++l) { F.mvInvScaleFactors[l] = 1.0f / sf[l]; F.mvLevelSigma2[l] = sf[l] * sf[l]; F.mvInvLevelSigma2[l] = 1.0f / F.mvLevelSigma2[l]; }
    F.mnScaleLevels = (int)nl; F.mfScaleFactor = nl > 1 ? sf[1] : 1.2f; F.mfLogScaleFactor = std::log(F.mfScaleFactor);
    const float* g = b.get<float>(p + "grid");  // min_x, min_y, max_x, max_y, inv_w, inv_h
    Frame::mnMinX = g[0]; Frame::mnMinY = g[1]; Frame::mnMaxX = g[2]; Frame::mnMaxY = g[3];
    Frame::mfGridElementWidthInv = g[4]; Frame::mfGridElementHeightInv = g[5];
    Frame::mbInitialComputations = false;
    F.mvpMapPoints.assign(F.N, (MapPoint*)NULL); F.mvbOutlier.assign(F.N, false);
    F.mpORBvocabulary = &g_voc; F.mpORBextractorLeft = F.mpORBextractorRight = NULL; F.mpReferenceKF = NULL;
    F.mTimeStamp = 0; F.mnId = Frame::nNextId++;
    F.mK = cv::Mat::eye(3, 3, CV_32F); F.mDistCoef = cv::Mat::zeros(4, 1, CV_32F);
    Frame::fx = Frame::fy = 1.f; Frame::cx = Frame::cy = 0.f; Frame::invfx = Frame::invfy = 1.f;
    F.mbf = 0.f; F.mb = 0.f; F.mThDepth = 0.f;
    for (int i = 0; i < FRAME_GRID_COLS; ++i) for (int j = 0; j < FRAME_GRID_ROWS; ++j) F.mGrid[i][j].clear();
    F.AssignFeaturesToGrid();                  // src/Frame.cc:261-276
    F.SetPose(Identity4());                    // src/Frame.cc:287-300
    if (b.has(p + "fv_ids")) {
        const unsigned* ids = b.get<unsigned>(p + "fv_ids"); const int* st = b.get<int>(p + "fv_start"); const unsigned* it = b.get<unsigned>(p + "fv_items");
        for (size_t k = 0; k < b.count(p + "fv_ids"); ++k) for (int q = st[k]; q < st[k + 1]; ++q) F.mFeatVec.addFeature(ids[k], it[q]);
    }
}
static void SetCamera(Frame& F, const float* cam) {
    Frame::fx = cam[0]; Frame::fy = cam[1]; Frame::cx = cam[2]; Frame::cy = cam[3];
    Frame::invfx = 1.0f / cam[0]; Frame::invfy = 1.0f / cam[1];
    F.mbf = cam[4]; F.mb = F.mbf / Frame::fx;
}

// MapPoints of a world scenario: position, normal, distance range, descriptor, flags.  ids[MapPoint*] = index.
static std::map<MapPoint*, int> g_id;
static MapPoint* NewPoint(const float* xw, KeyFrame* pRef) {
    cv::Mat pos(3, 1, CV_32F);
    for (int k = 0; k < 3; ++k) pos.at<float>(k) = xw ? xw[k] : 0.f;
    return new MapPoint(pos, pRef, &g_map);    // src/MapPoint.cc:32-44
}
static void LoadPoints(const Blob& b, const std::string& p, KeyFrame* pRef, std::vector<MapPoint*>& all, std::vector<MapPoint*>& ptrs, int idBase = 0) {
    const int M = (int)b.count(p + "maxd");
    all.assign(M, (MapPoint*)NULL); ptrs.assign(M, (MapPoint*)NULL);
    for (int i = 0; i < M; ++i) {
        MapPoint* m = NewPoint(b.get<float>(p + "xw") + 3 * i, pRef);
        for (int k = 0; k < 3; ++k) m->mNormalVector.at<float>(k) = b.get<float>(p + "normal")[3 * i + k];
        m->mfMaxDistance = b.get<float>(p + "maxd")[i]; m->mfMinDistance = b.get<float>(p + "mind")[i];
        m->mDescriptor = Row32(b.get<unsigned char>(p + "desc") + (size_t)i * 32);
        m->mbBad = b.get<unsigned char>(p + "bad")[i] != 0;
        m->nObs = b.get<int>(p + "nobs")[i];
        g_id[m] = idBase + i; all[i] = m;
        if (!b.get<unsigned char>(p + "null")[i]) ptrs[i] = m;
    }
}
static int IdOf(MapPoint* m) { return m && g_id.count(m) ? g_id[m] : -1; }

// A KeyFrame whose features are M placeholders carrying only an angle (the KeyFrame side of the
// relocalisation search reads pKF->mvKeysUn[i].angle and GetMapPointMatches(), nothing else).
static KeyFrame* MakeAngleKeyFrame(const Frame& like, int M, const float* angle) {
    Frame S(like);
    S.N = M; S.mvKeys.assign(M, cv::KeyPoint()); S.mvKeysUn.assign(M, cv::KeyPoint());
    for (int i = 0; i < M; ++i) S.mvKeysUn[i].angle = angle[i];
    S.mvuRight.assign(M, -1.f); S.mvDepth.assign(M, -1.f);
    S.mDescriptors = cv::Mat::zeros(M > 0 ? M : 1, 32, CV_8U);
    S.mvpMapPoints.assign(M, (MapPoint*)NULL); S.mvbOutlier.assign(M, false);
    for (int i = 0; i < FRAME_GRID_COLS; ++i) for (int j = 0; j < FRAME_GRID_ROWS; ++j) S.mGrid[i][j].clear();
    return new KeyFrame(S, &g_map, g_db);
}

static void PutMatch(BlobWriter& w, int n, const std::vector<int>& match) { w.put("n", "i4", &n, 1); w.put("match", "i4", match.data(), match.size()); }

static int RunStereo(const Blob& b, BlobWriter& w);
#ifdef ORB_B200_DROPIN
static int RunStereoCtor(const Blob& b, BlobWriter& w);
static int RunDistanceCost(const Blob& b, BlobWriter& w);
#endif
static int RunGrid(const Blob& b, BlobWriter& w);
static int RunAlgebra(const Blob& b, BlobWriter& w);
static int RunFrustum(const Blob& b, BlobWriter& w);
static int RunDistinctive(const Blob& b, BlobWriter& w);
static int RunBowTransform(const Blob& b, BlobWriter& w, const char* vocfile);
static int RunBowScore(const Blob& b, BlobWriter& w, const char* vocfile);

int main(int argc, char** argv) {
    if (argc < 4) { fprintf(stderr, "usage: ref_match_driver <mode> <in> <out>\n"); return 2; }
    const std::string mode = argv[1];
    try {
        Blob b; b.load(argv[2]);
        BlobWriter w(argv[3]);
        g_db = new KeyFrameDatabase(g_voc);
        if (mode == "stereo") return RunStereo(b, w);
#ifdef ORB_B200_DROPIN
        if (mode == "stereo_ctor") return RunStereoCtor(b, w);
        if (mode == "dd_cost") return RunDistanceCost(b, w);
#endif
        if (mode == "grid") return RunGrid(b, w);
        if (mode == "algebra") return RunAlgebra(b, w);
        if (mode == "frustum") return RunFrustum(b, w);
        if (mode == "distinctive") return RunDistinctive(b, w);
        if (mode == "bowtransform") return RunBowTransform(b, w, argc > 4 ? argv[4] : "");
        if (mode == "bowscore") return RunBowScore(b, w, argc > 4 ? argv[4] : "");
        if (mode == "mappoints") {
            Frame F; LoadFrame(b, "f_", F);
            KeyFrame* ref = new KeyFrame(F, &g_map, g_db);
            const int M = (int)b.count("mp_valid");
            std::vector<MapPoint*> mps(M);
            const unsigned char* valid = b.get<unsigned char>("mp_valid"); const unsigned char* blocks = b.get<unsigned char>("mp_blocks");
            for (int i = 0; i < M; ++i) {
                MapPoint* m = NewPoint(NULL, ref); g_id[m] = i;
                m->mbTrackInView = valid[i] != 0;
                m->mTrackProjX = b.get<float>("mp_proj_x")[i]; m->mTrackProjY = b.get<float>("mp_proj_y")[i]; m->mTrackProjXR = b.get<float>("mp_proj_xr")[i];
                m->mnTrackScaleLevel = b.get<int>("mp_level")[i]; m->mTrackViewCos = b.get<float>("mp_view_cos")[i];
                m->mDescriptor = Row32(b.get<unsigned char>("mp_desc") + (size_t)i * 32); m->nObs = blocks[i] ? 1 : 0;
                mps[i] = m;
            }
            MapPoint* holder = NewPoint(NULL, ref); holder->nObs = 3;   // keypoints that already hold an observed MapPoint (:90-93)
            const unsigned char* occ = b.get<unsigned char>("occupied");
            for (int i = 0; i < F.N; ++i) if (occ[i]) F.mvpMapPoints[i] = holder;
            ORBmatcher matcher(b.get<float>("ratio")[0], true);
            const int n = matcher.SearchByProjection(F, mps, b.get<float>("th")[0]);
            std::vector<int> match(F.N, -1);
            for (int i = 0; i < F.N; ++i) if (F.mvpMapPoints[i] && F.mvpMapPoints[i] != holder) match[i] = IdOf(F.mvpMapPoints[i]);
            PutMatch(w, n, match);
        } else if (mode == "lastframe") {
            Frame Cur; LoadFrame(b, "f_", Cur);
            KeyFrame* ref = new KeyFrame(Cur, &g_map, g_db);
            const int M = (int)b.count("l_valid");
            Frame Last(Cur);
            Last.N = M; Last.mvKeys.assign(M, cv::KeyPoint()); Last.mvKeysUn.assign(M, cv::KeyPoint());
            Last.mvpMapPoints.assign(M, (MapPoint*)NULL); Last.mvbOutlier.assign(M, false);
            for (int i = 0; i < M; ++i) {
                const float xw[3] = {b.get<float>("l_xc")[i], b.get<float>("l_yc")[i], b.get<float>("l_zc")[i]};
                MapPoint* m = NewPoint(xw, ref); g_id[m] = i;
                m->nObs = b.get<unsigned char>("l_blocks")[i] ? 1 : 0;
                m->mDescriptor = Row32(b.get<unsigned char>("l_desc") + (size_t)i * 32);
                Last.mvKeys[i].octave = b.get<int>("l_octave")[i]; Last.mvKeysUn[i].angle = b.get<float>("l_angle")[i];
                if (b.get<unsigned char>("l_valid")[i]) Last.mvpMapPoints[i] = m;
            }
            SetCamera(Cur, b.get<float>("cam")); Cur.mb = 1e9f;  // identity poses: neither forward nor backward
            ORBmatcher matcher(0.9f, b.get<int>("check_ori")[0] != 0);
            const int n = matcher.SearchByProjection(Cur, Last, b.get<float>("th")[0], false);
            std::vector<int> match(Cur.N, -1);
            for (int i = 0; i < Cur.N; ++i) if (Cur.mvpMapPoints[i]) match[i] = IdOf(Cur.mvpMapPoints[i]);
            PutMatch(w, n, match);
        } else if (mode == "bow" || mode == "tri" || mode == "bowkf") {
            Frame A, B; LoadFrame(b, "a_", A); LoadFrame(b, "b_", B);
            if (mode == "tri") {   // camera centre of KF1 = (ex, ey, 1), KF2 at the origin with unit intrinsics -> epipole (ex, ey)
                const float* ep = b.get<float>("epipole");
                cv::Mat T = Identity4(); T.at<float>(0, 3) = -ep[0]; T.at<float>(1, 3) = -ep[1]; T.at<float>(2, 3) = -1.f;
                A.SetPose(T);
            }
            KeyFrame* KA = new KeyFrame(A, &g_map, g_db); KeyFrame* KB = new KeyFrame(B, &g_map, g_db);
            std::vector<MapPoint*> sb(B.N, (MapPoint*)NULL);
            for (int i = 0; i < A.N; ++i) if (b.get<unsigned char>("has_a")[i]) { MapPoint* m = NewPoint(NULL, KA); g_id[m] = i; KA->AddMapPoint(m, i); }
            for (int i = 0; i < B.N; ++i) if (b.get<unsigned char>("has_b")[i]) { MapPoint* m = NewPoint(NULL, KB); g_id[m] = i; KB->AddMapPoint(m, i); B.mvpMapPoints[i] = m; }
            const bool ori = b.has("check_ori") ? b.get<int>("check_ori")[0] != 0 : true;
            ORBmatcher matcher(b.get<float>("ratio")[0], ori);
            if (mode == "bow") {
                std::vector<MapPoint*> out;
                const int n = matcher.SearchByBoW(KA, B, out);
                std::vector<int> match(B.N, -1);
                for (int j = 0; j < B.N; ++j) if (out[j]) match[j] = IdOf(out[j]);
                PutMatch(w, n, match);
            } else if (mode == "bowkf") {
                std::vector<MapPoint*> out;
                const int n = matcher.SearchByBoW(KA, KB, out);
                std::vector<int> match(A.N, -1);
                for (int i = 0; i < A.N; ++i) if (out[i]) match[i] = IdOf(out[i]);
                PutMatch(w, n, match);
            } else {
                cv::Mat F12(3, 3, CV_32F); std::memcpy(F12.data, b.get<float>("f12"), 36);
                std::vector<std::pair<size_t, size_t> > pairs;
                const int n = matcher.SearchForTriangulation(KA, KB, F12, pairs, b.get<int>("only_stereo")[0] != 0);
                std::vector<int> flat;
                for (size_t k = 0; k < pairs.size(); ++k) { flat.push_back((int)pairs[k].first); flat.push_back((int)pairs[k].second); }
                w.put("n", "i4", &n, 1); w.put("pairs", "i4", flat.data(), flat.size());
            }
        } else if (mode == "reloc" || mode == "loop") {
            Frame F; LoadFrame(b, "f_", F);
            SetCamera(F, b.get<float>("cam"));
            F.SetPose(Mat44(b.get<float>("pose")));
            const int M = (int)b.count("p_maxd");
            const unsigned char* occ = b.get<unsigned char>("occupied"); const unsigned char* found = b.get<unsigned char>("p_found");
            ORBmatcher matcher(0.75f, b.get<int>("check_ori")[0] != 0);
            std::vector<MapPoint*> all, pts;
            std::vector<int> match(F.N, -1);
            int n;
            if (mode == "reloc") {   // Tracking::Relocalization: matcher.SearchByProjection(mCurrentFrame, vpCandidateKFs[i], sFound, 10, 100)
                KeyFrame* K = MakeAngleKeyFrame(F, M, b.get<float>("p_angle"));
                LoadPoints(b, "p_", K, all, pts);
                MapPoint* holder = NewPoint(NULL, K); holder->nObs = 3;
                for (int i = 0; i < F.N; ++i) if (occ[i]) F.mvpMapPoints[i] = holder;
                for (int i = 0; i < M; ++i) if (pts[i]) K->AddMapPoint(pts[i], i);
                std::set<MapPoint*> sFound;
                for (int i = 0; i < M; ++i) if (found[i] && pts[i]) sFound.insert(pts[i]);
                n = matcher.SearchByProjection(F, K, sFound, b.get<float>("th")[0], b.get<int>("max_dist")[0]);
                for (int i = 0; i < F.N; ++i) if (F.mvpMapPoints[i] && F.mvpMapPoints[i] != holder) match[i] = IdOf(F.mvpMapPoints[i]);
            } else {                 // LoopClosing::ComputeSim3: matcher.SearchByProjection(mpCurrentKF, mScw, mvpLoopMapPoints, mvpCurrentMatchedPoints, 10)
                KeyFrame* K = new KeyFrame(F, &g_map, g_db);
                LoadPoints(b, "p_", K, all, pts);
                MapPoint* holder = NewPoint(NULL, K); holder->nObs = 3;
                std::vector<MapPoint*> vpMatched(F.N, (MapPoint*)NULL);
                int nf = 0;
                std::vector<int> foundIdx; for (int i = 0; i < M; ++i) if (found[i]) foundIdx.push_back(i);
                for (int i = 0; i < F.N; ++i) if (occ[i]) vpMatched[i] = (nf < (int)foundIdx.size()) ? pts[foundIdx[nf++]] : holder;
                std::vector<MapPoint*> before = vpMatched;
                n = matcher.SearchByProjection(K, Mat44(b.get<float>("pose")), pts, vpMatched, (int)b.get<float>("th")[0]);
                for (int i = 0; i < F.N; ++i) if (vpMatched[i] && !before[i]) match[i] = IdOf(vpMatched[i]);
            }
            PutMatch(w, n, match);
        } else if (mode == "fuse" || mode == "fuse_scw") {
            // LocalMapping::SearchInNeighbors: matcher.Fuse(pKFi, vpMapPointMatches) (src/LocalMapping.cc:584,613);
            // LoopClosing::SearchAndFuse: matcher.Fuse(pKF, cvScw, mvpLoopMapPoints, 4, vpReplacePoints) (src/LoopClosing.cc:652)
            Frame F; LoadFrame(b, "f_", F);
            SetCamera(F, b.get<float>("cam"));
            F.SetPose(Mat44(b.get<float>("pose")));
            KeyFrame* K = new KeyFrame(F, &g_map, g_db);
            std::vector<MapPoint*> all, pts; LoadPoints(b, "p_", K, all, pts);
            const int M = (int)pts.size();
            const int E = (int)b.count("e_kp");    // MapPoints the keyframe already holds
            std::vector<MapPoint*> existing(E);
            for (int j = 0; j < E; ++j) {
                MapPoint* m = NewPoint(NULL, K); g_id[m] = 100000 + j;
                m->nObs = b.get<int>("e_nobs")[j]; m->mbBad = b.get<unsigned char>("e_bad")[j] != 0;
                existing[j] = m; K->AddMapPoint(m, b.get<int>("e_kp")[j]);
            }
            const int* inkf = b.get<int>("p_in_kf");   // candidates already observed in the keyframe (:927 / :1087)
            for (int i = 0; i < M; ++i) if (inkf[i] >= 0 && pts[i]) { pts[i]->mObservations[K] = (size_t)inkf[i]; K->AddMapPoint(pts[i], inkf[i]); }
            ORBmatcher matcher;
            int n;
            std::vector<int> replace(M, -1);
            if (mode == "fuse") n = matcher.Fuse(K, pts, b.get<float>("th")[0]);
            else {
                std::vector<MapPoint*> vpReplace(M, (MapPoint*)NULL);
                n = matcher.Fuse(K, Mat44(b.get<float>("pose")), pts, b.get<float>("th")[0], vpReplace);
                for (int i = 0; i < M; ++i) if (vpReplace[i]) replace[i] = IdOf(vpReplace[i]);
            }
            std::vector<int> kf_mp(F.N, -1), p_bad(M, 0), p_repl(M, -1), p_idx(M, -1), p_nobs(M, 0), e_bad(E, 0), e_repl(E, -1);
            for (int i = 0; i < F.N; ++i) kf_mp[i] = IdOf(K->GetMapPoint(i));
            for (int i = 0; i < M; ++i) { p_bad[i] = all[i]->mbBad; p_repl[i] = IdOf(all[i]->mpReplaced); p_idx[i] = all[i]->GetIndexInKeyFrame(K); p_nobs[i] = all[i]->nObs; }
            for (int j = 0; j < E; ++j) { e_bad[j] = existing[j]->mbBad; e_repl[j] = IdOf(existing[j]->mpReplaced); }
            w.put("n", "i4", &n, 1); w.put("kf_mp", "i4", kf_mp.data(), kf_mp.size()); w.put("replace", "i4", replace.data(), replace.size());
            w.put("p_bad", "i4", p_bad.data(), p_bad.size()); w.put("p_repl", "i4", p_repl.data(), p_repl.size()); w.put("p_idx", "i4", p_idx.data(), p_idx.size());
            w.put("p_nobs", "i4", p_nobs.data(), p_nobs.size());
            w.put("e_bad", "i4", e_bad.data(), e_bad.size()); w.put("e_repl", "i4", e_repl.data(), e_repl.size());
        } else if (mode == "sim3") {
            // LoopClosing::ComputeSim3: matcher.SearchBySim3(mpCurrentKF, pKF, vpMapPointMatches, s, R, t, 7.5) (src/LoopClosing.cc:348)
            Frame A, B; LoadFrame(b, "a_", A); LoadFrame(b, "b_", B);
            SetCamera(A, b.get<float>("cam")); SetCamera(B, b.get<float>("cam"));
            A.SetPose(Mat44(b.get<float>("pose_a"))); B.SetPose(Mat44(b.get<float>("pose_b")));
            KeyFrame* KA = new KeyFrame(A, &g_map, g_db); KeyFrame* KB = new KeyFrame(B, &g_map, g_db);
            std::vector<MapPoint*> sa, sb, pa, pb;
            LoadPoints(b, "pa_", KA, sa, pa); LoadPoints(b, "pb_", KB, sb, pb);
            for (int i = 0; i < A.N; ++i) if (pa[i]) KA->AddMapPoint(pa[i], i);
            for (int j = 0; j < B.N; ++j) if (pb[j]) { KB->AddMapPoint(pb[j], j); pb[j]->mObservations[KB] = (size_t)j; }
            std::vector<MapPoint*> m12(A.N, (MapPoint*)NULL);
            const int* pre = b.get<int>("pre");
            for (int i = 0; i < A.N; ++i) if (pre[i] >= 0) m12[i] = sb[pre[i]];
            cv::Mat R12(3, 3, CV_32F), t12(3, 1, CV_32F);
            std::memcpy(R12.data, b.get<float>("r12"), 36); std::memcpy(t12.data, b.get<float>("t12"), 12);
            ORBmatcher matcher(0.75f, true);
            const int n = matcher.SearchBySim3(KA, KB, m12, b.get<float>("s12")[0], R12, t12, b.get<float>("th")[0]);
            std::vector<int> match(A.N, -1);
            for (int i = 0; i < A.N; ++i) if (m12[i]) match[i] = IdOf(m12[i]);
            PutMatch(w, n, match);
        } else if (mode == "init") {
            Frame A, B; LoadFrame(b, "a_", A); LoadFrame(b, "b_", B);
            std::vector<cv::Point2f> prev(A.N);
            for (int i = 0; i < A.N; ++i) prev[i] = cv::Point2f(b.get<float>("prev_x")[i], b.get<float>("prev_y")[i]);
            ORBmatcher matcher(b.get<float>("ratio")[0], b.get<int>("check_ori")[0] != 0);
            std::vector<int> m12;
            const int n = matcher.SearchForInitialization(A, B, prev, m12, b.get<int>("window")[0]);
            std::vector<float> px(A.N), py(A.N);
            for (int i = 0; i < A.N; ++i) { px[i] = prev[i].x; py[i] = prev[i].y; }
            PutMatch(w, n, m12);
            w.put("prev_x", "f4", px.data(), px.size()); w.put("prev_y", "f4", py.data(), py.size());
        } else {
            fprintf(stderr, "unknown mode %s\n", mode.c_str());
            return 2;
        }
    } catch (const std::exception& e) {
        fprintf(stderr, "ref_match_driver: %s\n", e.what());
        return 1;
    }
    return 0;
}

// Frame::ComputeStereoMatches (src/Frame.cc:516-704) on a stereo pair: both images go through the
// reference extractor, the Frame is filled the way its stereo constructor does (:64-126), then the
// reference's own ComputeStereoMatches runs.  (The fork's debugging dump "xiewenzhang.txt", :556,
// lands in the current directory: the test runs this binary inside its tmp_path.)
static int RunStereo(const Blob& b, BlobWriter& w) {
    const int* p = b.get<int>("params");  // w, h, nfeatures, nlevels, iniTh, minTh
    const float sf = b.get<float>("scale")[0];
    const float mbf = b.get<float>("stereo")[0], fx = b.get<float>("stereo")[1];
#ifndef ORB_B200_DROPIN
    ORBextractor* eL = (ORBextractor*)ref_extractor_create(p[2], sf, p[3], p[4], p[5]);
    ORBextractor* eR = (ORBextractor*)ref_extractor_create(p[2], sf, p[3], p[4], p[5]);
    const int cap = p[2] * 4 + 1024;
    std::vector<ref_keypoint> kl(cap), kr(cap); std::vector<uint8_t> dl((size_t)cap * 32), dr((size_t)cap * 32);
    const int nl = ref_extract(eL, b.get<unsigned char>("left"), p[0], p[1], p[0], kl.data(), dl.data(), cap);
    const int nr = ref_extract(eR, b.get<unsigned char>("right"), p[0], p[1], p[0], kr.data(), dr.data(), cap);
    if (nl < 0 || nr < 0) throw std::runtime_error("keypoint capacity");
#else
    // the drop-in class, called the way Frame::ExtractORB calls it (src/Frame.cc:249-258): (*mpORBextractorLeft)(im, cv::Mat(), mvKeys, mDescriptors)
    ORBextractor* eL = new ORBextractor(p[2], sf, p[3], p[4], p[5]);
    ORBextractor* eR = new ORBextractor(p[2], sf, p[3], p[4], p[5]);
    cv::Mat imL(p[1], p[0], CV_8U, (void*)b.get<unsigned char>("left")), imR(p[1], p[0], CV_8U, (void*)b.get<unsigned char>("right"));
    std::vector<cv::KeyPoint> vkl, vkr; cv::Mat mdl, mdr;
    (*eL)(imL, cv::Mat(), vkl, mdl); (*eR)(imR, cv::Mat(), vkr, mdr);
    const int nl = (int)vkl.size(), nr = (int)vkr.size();
    std::vector<ref_keypoint> kl(nl > 0 ? nl : 1), kr(nr > 0 ? nr : 1); std::vector<uint8_t> dl((size_t)(nl > 0 ? nl : 1) * 32), dr((size_t)(nr > 0 ? nr : 1) * 32);
    for (int i = 0; i < nl; ++i) { const cv::KeyPoint& k = vkl[i]; const ref_keypoint q = {k.pt.x, k.pt.y, k.size, k.angle, k.response, k.octave}; kl[i] = q; std::memcpy(&dl[(size_t)i * 32], mdl.ptr(i), 32); }
    for (int i = 0; i < nr; ++i) { const cv::KeyPoint& k = vkr[i]; const ref_keypoint q = {k.pt.x, k.pt.y, k.size, k.angle, k.response, k.octave}; kr[i] = q; std::memcpy(&dr[(size_t)i * 32], mdr.ptr(i), 32); }
#endif
    Frame F;
    F.mpORBextractorLeft = eL; F.mpORBextractorRight = eR;
    F.mnScaleLevels = eL->GetLevels(); F.mfScaleFactor = eL->GetScaleFactor(); F.mfLogScaleFactor = std::log(F.mfScaleFactor);
    F.mvScaleFactors = eL->GetScaleFactors(); F.mvInvScaleFactors = eL->GetInverseScaleFactors();
    F.mvLevelSigma2 = eL->GetScaleSigmaSquares(); F.mvInvLevelSigma2 = eL->GetInverseScaleSigmaSquares();
    F.mvKeys.resize(nl); F.mvKeysRight.resize(nr);
    for (int i = 0; i < nl; ++i) F.mvKeys[i] = cv::KeyPoint(kl[i].x, kl[i].y, kl[i].size, kl[i].angle, kl[i].response, kl[i].octave);
    for (int i = 0; i < nr; ++i) F.mvKeysRight[i] = cv::KeyPoint(kr[i].x, kr[i].y, kr[i].size, kr[i].angle, kr[i].response, kr[i].octave);
    F.mDescriptors.create(nl, 32, CV_8U); F.mDescriptorsRight.create(nr, 32, CV_8U);
    if (nl) std::memcpy(F.mDescriptors.data, dl.data(), (size_t)nl * 32);
    if (nr) std::memcpy(F.mDescriptorsRight.data, dr.data(), (size_t)nr * 32);
    F.N = nl;
    F.mvKeysUn = F.mvKeys;                       // no distortion (src/Frame.cc:454-458)
    Frame::fx = fx; F.mbf = mbf; F.mb = mbf / fx; // :117
    F.ComputeStereoMatches();
    const float maxd = F.mbf / F.mb;              // :529-531 (minZ = mb)
    w.put("kl", "u1", kl.data(), (size_t)nl * sizeof(ref_keypoint)); w.put("dl", "u1", dl.data(), (size_t)nl * 32);
    w.put("kr", "u1", kr.data(), (size_t)nr * sizeof(ref_keypoint)); w.put("dr", "u1", dr.data(), (size_t)nr * 32);
    w.put("u_right", "f4", F.mvuRight.data(), F.mvuRight.size()); w.put("depth", "f4", F.mvDepth.data(), F.mvDepth.size());
    w.put("max_d", "f4", &maxd, 1);
    return 0;
}


#ifdef ORB_B200_DROPIN
// The reference's stereo Frame CONSTRUCTOR itself (src/Frame.cc:60-126) over the drop-in extractors: two ExtractORB threads,
// UndistortKeyPoints, the patched ComputeStereoMatches, image bounds, grid.  Frame::mb is read by ComputeStereoMatches (:529) before
// the constructor assigns it (:117), i.e. upstream reads whatever the storage held; the storage is pre-set to mbf / fx here so that
// the run is defined.  Same output names as the "stereo" mode.
static int RunStereoCtor(const Blob& b, BlobWriter& w) {
    const int* p = b.get<int>("params");
    const float sf = b.get<float>("scale")[0];
    const float mbf = b.get<float>("stereo")[0], fx = b.get<float>("stereo")[1];
    ORBextractor eL(p[2], sf, p[3], p[4], p[5]), eR(p[2], sf, p[3], p[4], p[5]);
    cv::Mat imL(p[1], p[0], CV_8U, (void*)b.get<unsigned char>("left")), imR(p[1], p[0], CV_8U, (void*)b.get<unsigned char>("right"));
    cv::Mat K = cv::Mat::eye(3, 3, CV_32F);
    K.at<float>(0, 0) = fx; K.at<float>(1, 1) = fx; K.at<float>(0, 2) = p[0] * 0.5f; K.at<float>(1, 2) = p[1] * 0.5f;
    cv::Mat dist = cv::Mat::zeros(4, 1, CV_32F);
    Frame::mbInitialComputations = true;
    void* raw = ::operator new(sizeof(Frame));
    std::memset(raw, 0, sizeof(Frame));
    {   // pre-set the one member the constructor reads before writing
        const size_t off = (size_t)((char*)&((Frame*)raw)->mb - (char*)raw);
        const float mb0 = mbf / fx;
        std::memcpy((char*)raw + off, &mb0, sizeof(float));
    }
    Frame* F = new (raw) Frame(imL, imR, 0.0, &eL, &eR, &g_voc, K, dist, mbf, 35.f * mbf / fx);
    const int nl = F->N, nr = (int)F->mvKeysRight.size();
    std::vector<ref_keypoint> kl(nl > 0 ? nl : 1), kr(nr > 0 ? nr : 1);
    for (int i = 0; i < nl; ++i) { const cv::KeyPoint& k = F->mvKeys[i]; const ref_keypoint q = {k.pt.x, k.pt.y, k.size, k.angle, k.response, k.octave}; kl[i] = q; }
    for (int i = 0; i < nr; ++i) { const cv::KeyPoint& k = F->mvKeysRight[i]; const ref_keypoint q = {k.pt.x, k.pt.y, k.size, k.angle, k.response, k.octave}; kr[i] = q; }
    const float maxd = F->mbf / F->mb;
    w.put("kl", "u1", kl.data(), (size_t)nl * sizeof(ref_keypoint)); w.put("dl", "u1", F->mDescriptors.data, (size_t)nl * 32);
    w.put("kr", "u1", kr.data(), (size_t)nr * sizeof(ref_keypoint)); w.put("dr", "u1", F->mDescriptorsRight.data, (size_t)nr * 32);
    w.put("u_right", "f4", F->mvuRight.data(), F->mvuRight.size()); w.put("depth", "f4", F->mvDepth.data(), F->mvDepth.size());
    w.put("max_d", "f4", &maxd, 1);
    // the lazily fetched host pyramid of the drop-in (mvImagePyramid[l]): level sizes and a checksum per level
    std::vector<int> lv;
    for (int l = 0; l < eL.GetLevels(); ++l) {
        const cv::Mat& m = eL.mvImagePyramid[l];
        unsigned sum = 0;
        for (int y = 0; y < m.rows; ++y) for (int x = 0; x < m.cols; ++x) sum = sum * 31u + m.ptr(y)[x];
        lv.push_back(m.cols); lv.push_back(m.rows); lv.push_back((int)sum);
    }
    w.put("pyr", "i4", lv.data(), lv.size());
    return 0;
}

// Cost of ORBmatcher::DescriptorDistance called per pair (upload + launch + synchronise each) against one DescriptorDistances call.
static int RunDistanceCost(const Blob& b, BlobWriter& w) {
    const int n = (int)b.count("a") / 32;
    cv::Mat A(n, 32, CV_8U), B(n, 32, CV_8U);
    std::memcpy(A.data, b.get<unsigned char>("a"), (size_t)n * 32); std::memcpy(B.data, b.get<unsigned char>("b"), (size_t)n * 32);
    std::vector<int> one(n), batch;
    ORBmatcher::DescriptorDistance(A.row(0), B.row(0));   // creates the thread's workspace
    const auto t0 = std::chrono::steady_clock::now();
    for (int i = 0; i < n; ++i) one[i] = ORBmatcher::DescriptorDistance(A.row(i), B.row(i));
    const auto t1 = std::chrono::steady_clock::now();
    ORBmatcher::DescriptorDistances(A, B, batch);
    const auto t2 = std::chrono::steady_clock::now();
    const float us[2] = {(float)(std::chrono::duration<double, std::micro>(t1 - t0).count() / n), (float)std::chrono::duration<double, std::micro>(t2 - t1).count()};
    w.put("one", "i4", one.data(), one.size()); w.put("batch", "i4", batch.data(), batch.size()); w.put("us", "f4", us, 2);
    return 0;
}
#endif

// Frame::AssignFeaturesToGrid / GetFeaturesInArea and KeyFrame::GetFeaturesInArea on query windows
static int RunGrid(const Blob& b, BlobWriter& w) {
    Frame F; LoadFrame(b, "f_", F);
    KeyFrame* K = new KeyFrame(F, &g_map, g_db);
    std::vector<int> start(1, 0), items;
    for (int ix = 0; ix < FRAME_GRID_COLS; ++ix)
        for (int iy = 0; iy < FRAME_GRID_ROWS; ++iy) { for (size_t k = 0; k < F.mGrid[ix][iy].size(); ++k) items.push_back((int)F.mGrid[ix][iy][k]); start.push_back((int)items.size()); }
    w.put("cell_start", "i4", start.data(), start.size()); w.put("cell_items", "i4", items.data(), items.size());
    const int Q = (int)b.count("qx");
    std::vector<int> qs(1, 0), qi, ks(1, 0), ki;
    for (int q = 0; q < Q; ++q) {
        const std::vector<size_t> v = F.GetFeaturesInArea(b.get<float>("qx")[q], b.get<float>("qy")[q], b.get<float>("qr")[q], b.get<int>("qmin")[q], b.get<int>("qmax")[q]);
        for (size_t k = 0; k < v.size(); ++k) qi.push_back((int)v[k]);
        qs.push_back((int)qi.size());
        const std::vector<size_t> u = K->GetFeaturesInArea(b.get<float>("qx")[q], b.get<float>("qy")[q], b.get<float>("qr")[q]);
        for (size_t k = 0; k < u.size(); ++k) ki.push_back((int)u[k]);
        ks.push_back((int)ki.size());
    }
    w.put("q_start", "i4", qs.data(), qs.size()); w.put("q_items", "i4", qi.data(), qi.size());
    w.put("k_start", "i4", ks.data(), ks.size()); w.put("k_items", "i4", ki.data(), ki.size());
    return 0;
}

// The shim's cv::Mat algebra on batches of random matrices, for comparison with the real cv2.gemm / cv2.norm
// (tests/test_oracle_refmatch.py::test_shim_algebra_matches_cv2).
static int RunAlgebra(const Blob& b, BlobWriter& w) {
    const int n = (int)b.count("x") / 3;
    const float* R = b.get<float>("R"); const float* x = b.get<float>("x"); const float* t = b.get<float>("t"); const float* T4 = b.get<float>("T4");
    std::vector<float> rxt(3 * n), ow(3 * n), nrm(n), rr(9 * n), tt(16 * n), dt(n);
    for (int i = 0; i < n; ++i) {
        cv::Mat Rm(3, 3, CV_32F, (void*)(R + 9 * i)), xm(3, 1, CV_32F, (void*)(x + 3 * i)), tm(3, 1, CV_32F, (void*)(t + 3 * i)), Tm(4, 4, CV_32F, (void*)(T4 + 16 * i));
        cv::Mat a = Rm * xm + tm, o = -Rm.t() * tm, r2 = Rm * Rm, t2 = Tm * Tm;
        for (int k = 0; k < 3; ++k) { rxt[3 * i + k] = a.at<float>(k); ow[3 * i + k] = o.at<float>(k); }
        nrm[i] = (float)cv::norm(xm); dt[i] = (float)xm.dot(tm);
        std::memcpy(&rr[9 * i], r2.data, 36); std::memcpy(&tt[16 * i], t2.data, 64);
    }
    w.put("rxt", "f4", rxt.data(), rxt.size()); w.put("ow", "f4", ow.data(), ow.size()); w.put("norm", "f4", nrm.data(), nrm.size());
    w.put("rr", "f4", rr.data(), rr.size()); w.put("tt", "f4", tt.data(), tt.size()); w.put("dot", "f4", dt.data(), dt.size());
    return 0;
}

// Frame::ComputeBoW (src/Frame.cc:439-449): ORBVocabulary::loadFromTextFile + transform(vCurrentDesc, mBowVec, mFeatVec, levelsup),
// all the reference's own DBoW2.  Doubles are written as raw bytes ("u1").
static int RunBowTransform(const Blob& b, BlobWriter& w, const char* vocfile) {
    ORBVocabulary voc;
    if (!voc.loadFromTextFile(vocfile)) throw std::runtime_error("vocabulary load failed");
    Frame F;
    F.mpORBvocabulary = &voc;
    F.N = (int)b.count("desc") / 32;
    F.mDescriptors.create(F.N, 32, CV_8U);
    std::memcpy(F.mDescriptors.data, b.get<unsigned char>("desc"), (size_t)F.N * 32);
    const int levelsup = b.get<int>("levelsup")[0];
    if (levelsup == 4) F.ComputeBoW();            // the reference's call, levelsup hard-wired to 4 (:448)
    else voc.transform(Converter::toDescriptorVector(F.mDescriptors), F.mBowVec, F.mFeatVec, levelsup);
    std::vector<unsigned> bw_ids, fv_ids, fv_items; std::vector<double> bw_vals; std::vector<int> fv_start(1, 0);
    for (DBoW2::BowVector::const_iterator it = F.mBowVec.begin(); it != F.mBowVec.end(); ++it) { bw_ids.push_back(it->first); bw_vals.push_back(it->second); }
    for (DBoW2::FeatureVector::const_iterator it = F.mFeatVec.begin(); it != F.mFeatVec.end(); ++it) {
        fv_ids.push_back(it->first);
        for (size_t k = 0; k < it->second.size(); ++k) fv_items.push_back(it->second[k]);
        fv_start.push_back((int)fv_items.size());
    }
    w.put("bow_ids", "u4", bw_ids.data(), bw_ids.size()); w.put("bow_vals", "u1", bw_vals.data(), bw_vals.size() * 8);
    w.put("fv_ids", "u4", fv_ids.data(), fv_ids.size()); w.put("fv_start", "i4", fv_start.data(), fv_start.size()); w.put("fv_items", "u4", fv_items.data(), fv_items.size());
    // per-feature descent
    std::vector<int> word(F.N), node(F.N); std::vector<double> weight(F.N);
    for (int i = 0; i < F.N; ++i) {
        DBoW2::WordId id; DBoW2::NodeId nid = 0; DBoW2::WordValue wv;
        voc.transform(F.mDescriptors.row(i), id, wv, &nid, levelsup);
        word[i] = (int)id; node[i] = (int)nid; weight[i] = wv;
    }
    w.put("word", "i4", word.data(), word.size()); w.put("node", "i4", node.data(), node.size()); w.put("weight", "u1", weight.data(), weight.size() * 8);
    const int info[2] = {(int)voc.size(), (int)voc.getDepthLevels()};
    w.put("info", "i4", info, 2);
    return 0;
}

// ORBVocabulary::score(BowVector, BowVector) as KeyFrameDatabase::DetectLoopCandidates / DetectRelocalizationCandidates call it
// (src/KeyFrameDatabase.cc:137,267): two descriptor sets transformed by the vocabulary, scored both ways and against themselves.
static int RunBowScore(const Blob& b, BlobWriter& w, const char* vocfile) {
    ORBVocabulary voc;
    if (!voc.loadFromTextFile(vocfile)) throw std::runtime_error("vocabulary load failed");
    DBoW2::BowVector bv[2]; DBoW2::FeatureVector fv[2];
    const char* names[2] = {"desc_a", "desc_b"};
    for (int s = 0; s < 2; ++s) {
        const int n = (int)b.count(names[s]) / 32;
        cv::Mat d(n, 32, CV_8U);
        std::memcpy(d.data, b.get<unsigned char>(names[s]), (size_t)n * 32);
        std::vector<cv::Mat> rows;
        for (int j = 0; j < n; ++j) rows.push_back(d.row(j));
        voc.transform(rows, bv[s], fv[s], 4);
    }
    const double sc[4] = {voc.score(bv[0], bv[1]), voc.score(bv[1], bv[0]), voc.score(bv[0], bv[0]), voc.score(bv[1], bv[1])};
    w.put("scores", "u1", sc, sizeof(sc));
    for (int s = 0; s < 2; ++s) {
        std::vector<unsigned> ids; std::vector<double> vals;
        for (DBoW2::BowVector::const_iterator it = bv[s].begin(); it != bv[s].end(); ++it) { ids.push_back(it->first); vals.push_back(it->second); }
        w.put(s ? "ids_b" : "ids_a", "u4", ids.data(), ids.size()); w.put(s ? "vals_b" : "vals_a", "u1", vals.data(), vals.size() * 8);
    }
    return 0;
}

// Frame::isInFrustum (src/Frame.cc:303-366) per MapPoint of a world scenario, as Tracking::SearchLocalPoints drives it.
static int RunFrustum(const Blob& b, BlobWriter& w) {
    Frame F; LoadFrame(b, "f_", F);
    SetCamera(F, b.get<float>("cam"));
    F.SetPose(Mat44(b.get<float>("pose")));
    KeyFrame* K = new KeyFrame(F, &g_map, g_db);
    std::vector<MapPoint*> all, pts; LoadPoints(b, "p_", K, all, pts);
    const int M = (int)all.size();
    const float lim = b.get<float>("cos_limit")[0];
    std::vector<unsigned char> ok(M); std::vector<float> u(M), v(M), ur(M), vc(M); std::vector<int> lvl(M);
    for (int i = 0; i < M; ++i) {
        MapPoint* m = all[i];
        m->mTrackProjX = m->mTrackProjY = m->mTrackProjXR = m->mTrackViewCos = 0.f; m->mnTrackScaleLevel = -1;
        ok[i] = F.isInFrustum(m, lim) ? 1 : 0;
        if (ok[i] != (m->mbTrackInView ? 1 : 0)) throw std::runtime_error("mbTrackInView differs from the return value");
        u[i] = m->mTrackProjX; v[i] = m->mTrackProjY; ur[i] = m->mTrackProjXR; vc[i] = m->mTrackViewCos; lvl[i] = m->mnTrackScaleLevel;
    }
    cv::Mat Ow = F.GetCameraCenter();
    float pose[15];
    for (int r = 0; r < 3; ++r) { for (int c = 0; c < 3; ++c) pose[3 * r + c] = F.mRcw.at<float>(r, c); pose[9 + r] = F.mtcw.at<float>(r); pose[12 + r] = Ow.at<float>(r); }
    w.put("in_view", "u1", ok.data(), M); w.put("u", "f4", u.data(), M); w.put("v", "f4", v.data(), M); w.put("ur", "f4", ur.data(), M);
    w.put("view_cos", "f4", vc.data(), M); w.put("level", "i4", lvl.data(), M); w.put("rt_ow", "f4", pose, 15);
    return 0;
}

// MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:250-326): KeyFrames holding descriptor rows, MapPoints observing
// (keyframe, row) pairs.  The reference gathers the rows in std::map<KeyFrame*, size_t> order, i.e. by KeyFrame ADDRESS; the rows
// are written back in the order the reference saw them, with the descriptor it chose.
static int RunDistinctive(const Blob& b, BlobWriter& w) {
    const int nkf = b.get<int>("dims")[0], rows = b.get<int>("dims")[1], P = (int)b.count("start") - 1;
    Frame F; LoadFrame(b, "f_", F);
    std::vector<KeyFrame*> kfs(nkf);
    for (int k = 0; k < nkf; ++k) {
        Frame S(F);
        S.N = rows; S.mvKeys.assign(rows, cv::KeyPoint()); S.mvKeysUn.assign(rows, cv::KeyPoint()); S.mvuRight.assign(rows, -1.f); S.mvDepth.assign(rows, -1.f);
        S.mDescriptors.create(rows, 32, CV_8U); std::memcpy(S.mDescriptors.data, b.get<unsigned char>("kf_desc") + (size_t)k * rows * 32, (size_t)rows * 32);
        S.mvpMapPoints.assign(rows, (MapPoint*)NULL); S.mvbOutlier.assign(rows, false);
        for (int i = 0; i < FRAME_GRID_COLS; ++i) for (int j = 0; j < FRAME_GRID_ROWS; ++j) S.mGrid[i][j].clear();
        kfs[k] = new KeyFrame(S, &g_map, g_db);
    }
    const int* start = b.get<int>("start"); const int* okf = b.get<int>("obs_kf"); const int* orow = b.get<int>("obs_row");
    std::vector<unsigned char> gathered, chosen((size_t)P * 32, 0); std::vector<int> gstart(1, 0);
    for (int p = 0; p < P; ++p) {
        MapPoint* m = NewPoint(NULL, kfs[0]);
        for (int q = start[p]; q < start[p + 1]; ++q) m->AddObservation(kfs[okf[q]], (size_t)orow[q]);
        m->ComputeDistinctiveDescriptors();
        const std::map<KeyFrame*, size_t> obs = m->GetObservations();
        for (std::map<KeyFrame*, size_t>::const_iterator it = obs.begin(); it != obs.end(); ++it) {
            const unsigned char* r = it->first->mDescriptors.ptr(it->second);
            gathered.insert(gathered.end(), r, r + 32);
        }
        gstart.push_back((int)gathered.size() / 32);
        cv::Mat d = m->GetDescriptor();
        if (!d.empty()) std::memcpy(&chosen[(size_t)p * 32], d.data, 32);
    }
    w.put("g_start", "i4", gstart.data(), gstart.size()); w.put("g_desc", "u1", gathered.data(), gathered.size()); w.put("chosen", "u1", chosen.data(), chosen.size());
    return 0;
}
