// TEST INFRASTRUCTURE — C wrapper around the reference's own ORB_SLAM2::ORBextractor
// (/root/reference/src/ORBextractor.cc compiled verbatim, see oracle/Makefile target `ref`).
//
// The reference breaks equal-size ties in DistributeOctTree by comparing ExtractorNode heap
// addresses (sort of pair<int, ExtractorNode*>, src/ORBextractor.cc:703), so its output depends
// on the allocator.  While an extraction runs, this library's operator new hands out strictly
// increasing addresses from a per-thread bump arena (and never reuses freed memory), which makes
// "higher address" == "created later": the canonical tie-break the oracle and the CUDA path use.
#include <sys/mman.h>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>
#include "ORBextractor.h"

namespace {
const size_t kArenaBytes = (size_t)6 << 30;  // virtual, committed lazily
struct Arena { char* base; size_t off; bool active; };
thread_local Arena g_arena = {nullptr, 0, false};

void* arena_alloc(size_t n) {
    Arena& a = g_arena;
    if (!a.base) {
        void* p = mmap(nullptr, kArenaBytes, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
        if (p == MAP_FAILED) std::abort();
        a.base = (char*)p;
    }
    n = (n + 15) & ~(size_t)15;
    if (a.off + n > kArenaBytes) std::abort();
    void* r = a.base + a.off;
    a.off += n;
    return r;
}
inline bool in_arena(void* p) { return g_arena.base && (char*)p >= g_arena.base && (char*)p < g_arena.base + kArenaBytes; }
}  // namespace

void* operator new(size_t n) { if (g_arena.active) return arena_alloc(n); void* p = std::malloc(n ? n : 1); if (!p) throw std::bad_alloc(); return p; }
void* operator new[](size_t n) { return operator new(n); }
void operator delete(void* p) noexcept { if (p && !in_arena(p)) std::free(p); }
void operator delete[](void* p) noexcept { operator delete(p); }
void operator delete(void* p, size_t) noexcept { operator delete(p); }
void operator delete[](void* p, size_t) noexcept { operator delete(p); }

struct ref_keypoint { float x, y, size, angle, response; int32_t octave; };

extern "C" {

void* ref_extractor_create(int nfeatures, float scale_factor, int nlevels, int ini_th, int min_th) {
    return new ORB_SLAM2::ORBextractor(nfeatures, scale_factor, nlevels, ini_th, min_th);
}
void ref_extractor_destroy(void* e) { delete (ORB_SLAM2::ORBextractor*)e; }

// ORBextractor::operator(); returns keypoint count or -1 when cap is too small
int ref_extract(void* ev, const uint8_t* img, int w, int h, int stride, ref_keypoint* kps, uint8_t* desc, int cap) {
    ORB_SLAM2::ORBextractor* e = (ORB_SLAM2::ORBextractor*)ev;
    int n = 0;
    {
        g_arena.off = 0;
        g_arena.active = true;
        {
            cv::Mat image(h, w, CV_8UC1, (void*)img, (size_t)stride), mask, descriptors;
            std::vector<cv::KeyPoint> keypoints;
            (*e)(image, mask, keypoints, descriptors);
            n = (int)keypoints.size();
            if (n <= cap) {
                for (int i = 0; i < n; ++i) {
                    const cv::KeyPoint& k = keypoints[i];
                    kps[i].x = k.pt.x; kps[i].y = k.pt.y; kps[i].size = k.size; kps[i].angle = k.angle; kps[i].response = k.response; kps[i].octave = k.octave;
                    std::memcpy(desc + (size_t)i * 32, descriptors.ptr(i), 32);
                }
            } else n = -1;
        }
        g_arena.active = false;
        if (g_arena.off > ((size_t)256 << 20)) madvise(g_arena.base, g_arena.off, MADV_DONTNEED);
    }
    return n;
}

// mvImagePyramid[level] of the last call (unbordered view): copies into dst (w*h bytes, tight)
int ref_level(void* ev, int level, uint8_t* dst, int cap, int* w, int* h) {
    ORB_SLAM2::ORBextractor* e = (ORB_SLAM2::ORBextractor*)ev;
    if (level < 0 || level >= (int)e->mvImagePyramid.size()) return -1;
    const cv::Mat& m = e->mvImagePyramid[level];
    *w = m.cols; *h = m.rows;
    if (m.cols * m.rows > cap) return -1;
    for (int y = 0; y < m.rows; ++y) std::memcpy(dst + (size_t)y * m.cols, m.ptr(y), (size_t)m.cols);
    return 0;
}
// pointer view of mvImagePyramid[level] (valid until the next ref_extract on this handle)
const uint8_t* ref_level_ptr(void* ev, int level, int* w, int* h, int* stride) {
    ORB_SLAM2::ORBextractor* e = (ORB_SLAM2::ORBextractor*)ev;
    if (level < 0 || level >= (int)e->mvImagePyramid.size()) return nullptr;
    const cv::Mat& m = e->mvImagePyramid[level];
    *w = m.cols; *h = m.rows; *stride = (int)m.step;
    return m.data;
}
int ref_levels(void* ev) { return ((ORB_SLAM2::ORBextractor*)ev)->GetLevels(); }
void ref_scale_factors(void* ev, float* sf) {
    std::vector<float> v = ((ORB_SLAM2::ORBextractor*)ev)->GetScaleFactors();
    for (size_t i = 0; i < v.size(); ++i) sf[i] = v[i];
}

}  // extern "C"
