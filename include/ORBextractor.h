/* Drop-in for the reference's include/ORBextractor.h: same namespace, class name, constructor,
 * operator(), getters and public mvImagePyramid member (reference include/ORBextractor.h:45-112),
 * implemented over the C ABI of liborb_b200.so (include/orb_b200.h).  All arithmetic happens in
 * hand-written sm_100a CUDA kernels; there is NO CPU fallback: without a usable CUDA device the
 * constructor throws std::runtime_error.
 *
 * Differences a maintainer should know:
 *   - the handle is sized for images up to (maxWidth x maxHeight), default 4096 x 4096 lazily grown;
 *   - mvImagePyramid is always valid but lazy: the levels of the last operator() call are brought to the host on the
 *     first read after that call (one staged copy, orb_get_pyramid), and not at all when nobody reads them
 *     (Frame::ComputeStereoMatches, its only reader in the reference, is replaced by orb_stereo_match on the
 *     device-resident pyramids).  SetKeepPyramidOnHost(true) downloads eagerly inside operator() instead;
 *   - ExtractBatch() is an additional entry point for many images per call.
 */
#ifndef ORBEXTRACTOR_H
#define ORBEXTRACTOR_H

#include <list>
#include <stdexcept>
#include <string>
#include <vector>

#include "orb_b200.h"
#include "orb_b200_cv.h"

namespace ORB_SLAM2 {

class ORBextractor {
public:
    enum { HARRIS_SCORE = 0, FAST_SCORE = 1 };

    ORBextractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST)
        : mHandle(nullptr), mMaxW(0), mMaxH(0), mMaxBatch(1), mDevice(0), mKeepPyramid(false), nfeatures(nfeatures),
          scaleFactor(scaleFactor), nlevels(nlevels), iniThFAST(iniThFAST), minThFAST(minThFAST) {
        if (orb_device_count() < 1) throw std::runtime_error("ORBextractor (B200): no CUDA device; this build has no CPU path");
        // Tables are produced by the library with the reference's arithmetic (src/ORBextractor.cc:433-464);
        // a 1x1-capacity handle is enough to query them.
        Reserve(64, 64, 1);
        mvScaleFactor.resize(nlevels); mvInvScaleFactor.resize(nlevels); mvLevelSigma2.resize(nlevels); mvInvLevelSigma2.resize(nlevels);
        mnFeaturesPerLevel.resize(nlevels); umax.resize(16);
        Check(orb_extractor_tables(mHandle, mvScaleFactor.data(), mvInvScaleFactor.data(), mvLevelSigma2.data(), mvInvLevelSigma2.data(),
                                   mnFeaturesPerLevel.data(), umax.data()));
        mvImagePyramid.Bind(this);
    }
    ~ORBextractor() { if (mHandle) orb_extractor_destroy(mHandle); }
    ORBextractor(const ORBextractor&) = delete;
    ORBextractor& operator=(const ORBextractor&) = delete;

    // Compute the ORB features and descriptors on an image (reference src/ORBextractor.cc:1091-1169).
    // Mask is ignored, as in the reference.
    void operator()(cv::InputArray _image, cv::InputArray /*mask*/, std::vector<cv::KeyPoint>& _keypoints, cv::OutputArray _descriptors) {
        const cv::Mat image = _image.getMat();
        if (image.empty()) return;  // :1095 (outputs untouched)
        if (image.type() != CV_8UC1) throw std::invalid_argument("ORBextractor: image.type() must be CV_8UC1");  // assert at :1100
        Reserve(image.cols, image.rows, 1);
        const int cap = orb_extractor_max_keypoints(mHandle);
        mKp.resize(cap); mDesc.resize((size_t)cap * 32);
        int32_t n = 0;
        Check(orb_extract(mHandle, image.data, image.cols, image.rows, (int)image.step, mKp.data(), mDesc.data(), cap, &n));
        static_assert(sizeof(cv::KeyPoint) == sizeof(orb_keypoint), "cv::KeyPoint layout");
        _keypoints.resize(n);
        if (n == 0) { _descriptors.release(); }
        else {
            std::memcpy(static_cast<void*>(_keypoints.data()), mKp.data(), (size_t)n * sizeof(orb_keypoint));
            _descriptors.create(n, 32, CV_8U);
            cv::Mat d = _descriptors.getMat();
            for (int i = 0; i < n; ++i) std::memcpy(d.ptr<unsigned char>(i), &mDesc[(size_t)i * 32], 32);
        }
        mvImagePyramid.Invalidate();
        if (mKeepPyramid) DownloadPyramid(0);
    }

    // Many same-sized images in one call: images[b] -> keypoints[b], descriptors[b].
    void ExtractBatch(const std::vector<cv::Mat>& images, std::vector<std::vector<cv::KeyPoint> >& keypoints, std::vector<cv::Mat>& descriptors) {
        const int B = (int)images.size();
        keypoints.assign(B, std::vector<cv::KeyPoint>()); descriptors.assign(B, cv::Mat());
        if (B == 0) return;
        const int w = images[0].cols, h = images[0].rows;
        std::vector<unsigned char> dense((size_t)B * w * h);
        for (int b = 0; b < B; ++b) {
            if (images[b].cols != w || images[b].rows != h || images[b].type() != CV_8UC1) throw std::invalid_argument("ExtractBatch: images must share size and be CV_8UC1");
            for (int y = 0; y < h; ++y) std::memcpy(&dense[((size_t)b * h + y) * w], images[b].ptr<unsigned char>(y), w);
        }
        Reserve(w, h, B);
        const int cap = orb_extractor_max_keypoints(mHandle);
        mKp.resize((size_t)B * cap); mDesc.resize((size_t)B * cap * 32);
        std::vector<int32_t> n(B, 0);
        Check(orb_extract_batch(mHandle, dense.data(), B, w, h, w, (size_t)w * h, mKp.data(), mDesc.data(), cap, n.data()));
        for (int b = 0; b < B; ++b) {
            keypoints[b].resize(n[b]);
            if (n[b] == 0) continue;
            std::memcpy(static_cast<void*>(keypoints[b].data()), &mKp[(size_t)b * cap], (size_t)n[b] * sizeof(orb_keypoint));
            descriptors[b].create(n[b], 32, CV_8U);
            for (int i = 0; i < n[b]; ++i) std::memcpy(descriptors[b].ptr<unsigned char>(i), &mDesc[((size_t)b * cap + i) * 32], 32);
        }
    }

    // The front end of the stereo Frame constructor in ONE call (src/Frame.cc:81-95: ExtractORB on both images from two threads, join,
    // ComputeStereoMatches): both extractions are queued from the calling thread on the two handles' streams, so no thread is created
    // per frame and the two host threads do not take turns at the driver's lock.  Outputs as two operator() calls followed by
    // ComputeStereoMatches: keypoints / descriptors of both sides, mvuRight and mvDepth (-1 = no match).  maxD = mbf / mb (:549).
    static void ExtractStereo(ORBextractor& left, ORBextractor& right, const cv::Mat& imLeft, const cv::Mat& imRight, float mbf, float minD,
                              float maxD, std::vector<cv::KeyPoint>& keysLeft, cv::Mat& descLeft, std::vector<cv::KeyPoint>& keysRight,
                              cv::Mat& descRight, std::vector<float>& uRight, std::vector<float>& depth) {
        if (imLeft.empty() || imRight.empty()) return;
        if (imLeft.type() != CV_8UC1 || imRight.type() != CV_8UC1 || imLeft.cols != imRight.cols || imLeft.rows != imRight.rows ||
            (size_t)imLeft.step != (size_t)imRight.step)
            throw std::invalid_argument("ExtractStereo: two CV_8UC1 images of one size and step");
        left.Reserve(imLeft.cols, imLeft.rows, 1); right.Reserve(imLeft.cols, imLeft.rows, 1);
        const int cap = orb_extractor_max_keypoints(left.mHandle);
        if (orb_extractor_max_keypoints(right.mHandle) != cap) throw std::invalid_argument("ExtractStereo: extractors differ in settings");
        left.mKp.resize(cap); left.mDesc.resize((size_t)cap * 32); right.mKp.resize(cap); right.mDesc.resize((size_t)cap * 32);
        uRight.resize(cap); depth.resize(cap);
        int32_t nl = 0, nr = 0;
        Check(orb_stereo_frames(left.mHandle, right.mHandle, imLeft.data, imRight.data, 1, imLeft.cols, imLeft.rows, (int)imLeft.step,
                                (size_t)imLeft.step * imLeft.rows, mbf, minD, maxD, left.mKp.data(), left.mDesc.data(), &nl, right.mKp.data(),
                                right.mDesc.data(), &nr, uRight.data(), depth.data(), cap));
        left.Deliver(nl, keysLeft, descLeft); right.Deliver(nr, keysRight, descRight);
        uRight.resize(nl); depth.resize(nl);
    }

    int inline GetLevels() { return nlevels; }
    float inline GetScaleFactor() { return (float)scaleFactor; }
    std::vector<float> inline GetScaleFactors() { return mvScaleFactor; }
    std::vector<float> inline GetInverseScaleFactors() { return mvInvScaleFactor; }
    std::vector<float> inline GetScaleSigmaSquares() { return mvLevelSigma2; }
    std::vector<float> inline GetInverseScaleSigmaSquares() { return mvInvLevelSigma2; }

    // Levels of the last operator() call, without the 19-pixel border the reference keeps around them
    // (src/ORBextractor.cc:1181-1203: mvImagePyramid[l] is the ROI inside the bordered buffer).  Reads like the reference's
    // std::vector<cv::Mat> (operator[], size(), iteration, conversion); the host copy is made on the first read after a call.
    class Pyramid {
    public:
        Pyramid() : mOwner(nullptr), mStale(false) {}
        cv::Mat& operator[](size_t l) { Fetch(); return mLevels[l]; }
        cv::Mat& at(size_t l) { Fetch(); return mLevels.at(l); }
        size_t size() const { return mLevels.size(); }
        bool empty() const { return mLevels.empty(); }
        std::vector<cv::Mat>::iterator begin() { Fetch(); return mLevels.begin(); }
        std::vector<cv::Mat>::iterator end() { Fetch(); return mLevels.end(); }
        operator const std::vector<cv::Mat>&() { Fetch(); return mLevels; }
        void resize(size_t n) { mLevels.resize(n); }
        bool OnHost() const { return !mStale; }
    private:
        friend class ORBextractor;
        void Bind(ORBextractor* owner) { mOwner = owner; mLevels.resize(owner->nlevels); }
        void Invalidate() { mStale = true; }
        void Fetch() { if (mStale && mOwner) { mStale = false; mOwner->DownloadPyramid(0); } }
        ORBextractor* mOwner;
        bool mStale;
        std::vector<cv::Mat> mLevels;
    };
    Pyramid mvImagePyramid;

    // ---- additions
    void SetKeepPyramidOnHost(bool on) { mKeepPyramid = on; }
    void SetDevice(int device) { if (device != mDevice) { mDevice = device; Recreate(); } }
    // Capacity for later calls (avoids re-creating the handle when sizes grow).
    void Reserve(int width, int height, int batch) {
        if (mHandle && width <= mMaxW && height <= mMaxH && batch <= mMaxBatch) return;
        if (width > mMaxW) mMaxW = width;
        if (height > mMaxH) mMaxH = height;
        if (batch > mMaxBatch) mMaxBatch = batch;
        Recreate();
    }
    void DownloadPyramid(int image) {
        std::vector<cv::Mat>& lv = mvImagePyramid.mLevels;
        lv.resize(nlevels);
        std::vector<unsigned char*> dst(nlevels);
        std::vector<int32_t> stride(nlevels);
        for (int l = 0; l < nlevels; ++l) {
            int32_t w = 0, h = 0;
            Check(orb_level_size(mHandle, l, &w, &h));
            lv[l].create(h, w, CV_8U);
            dst[l] = lv[l].data; stride[l] = (int32_t)lv[l].step;
        }
        Check(orb_get_pyramid(mHandle, image, dst.data(), stride.data()));
        mvImagePyramid.mStale = false;
    }
    orb_extractor* Handle() { return mHandle; }  // for orb_stereo_match / orb_stereo_frames

protected:
    void Deliver(int n, std::vector<cv::KeyPoint>& keys, cv::Mat& desc) {
        keys.resize(n);
        if (n == 0) desc.release();
        else {
            std::memcpy(static_cast<void*>(keys.data()), mKp.data(), (size_t)n * sizeof(orb_keypoint));
            desc.create(n, 32, CV_8U);
            if (desc.isContinuous()) std::memcpy(desc.data, mDesc.data(), (size_t)n * 32);
            else for (int i = 0; i < n; ++i) std::memcpy(desc.ptr<unsigned char>(i), &mDesc[(size_t)i * 32], 32);
        }
        mvImagePyramid.Invalidate();
        if (mKeepPyramid) DownloadPyramid(0);
    }
    void Recreate() {
        if (mHandle) { orb_extractor_destroy(mHandle); mHandle = nullptr; }
        Check(orb_extractor_create(nfeatures, (float)scaleFactor, nlevels, iniThFAST, minThFAST, mMaxW, mMaxH, mMaxBatch, mDevice, &mHandle));
    }
    static void Check(int status) {
        if (status != ORB_OK) throw std::runtime_error(std::string("orb_b200: ") + orb_last_error());
    }

    orb_extractor* mHandle;
    int mMaxW, mMaxH, mMaxBatch, mDevice;
    bool mKeepPyramid;
    std::vector<orb_keypoint> mKp;
    std::vector<unsigned char> mDesc;

    int nfeatures;
    double scaleFactor;
    int nlevels;
    int iniThFAST;
    int minThFAST;
    std::vector<int> mnFeaturesPerLevel;
    std::vector<int> umax;
    std::vector<float> mvScaleFactor;
    std::vector<float> mvInvScaleFactor;
    std::vector<float> mvLevelSigma2;
    std::vector<float> mvInvLevelSigma2;
};

}  // namespace ORB_SLAM2

#endif
