// Multi-GPU front end in the library (frames are independent, shard by frame index, one host thread and one
// handle pair per GPU, pinned host staging, results gathered by the host; no collective on the data path).  A C++ user of the
// drop-in gets the whole box behind one call; the Python side (shard.py, bench.py under torchrun) does the same per process.
#include "orb_internal.h"
#include <algorithm>
#include <thread>

struct orb_multi {
    int n = 0, max_batch = 0;
    std::vector<int> devices;
    std::vector<orb_extractor*> left, right;
};

extern "C" {

// contiguous blocks: part p of n_parts gets frames [first, first + count) of a batch
void orb_multi_slice(int batch, int n_parts, int part, int* first, int* count) {
    if (n_parts < 1 || part < 0 || part >= n_parts || batch < 0) { if (first) *first = 0; if (count) *count = 0; return; }
    const int base = batch / n_parts, rem = batch % n_parts;
    const int f = part * base + std::min(part, rem);
    if (first) *first = f;
    if (count) *count = base + (part < rem ? 1 : 0);
}

void orb_multi_destroy(orb_multi* m) {
    if (!m) return;
    for (orb_extractor* e : m->left) orb_extractor_destroy(e);
    for (orb_extractor* e : m->right) orb_extractor_destroy(e);
    delete m;
}

int orb_multi_create(int nfeatures, float scale_factor, int nlevels, int ini_th, int min_th, int max_w, int max_h, int max_batch_per_gpu,
                     int n_devices, orb_multi** out) {
    if (!out) return ORB_ERR_BAD_ARG;
    *out = nullptr;
    const int ndev = orb_device_count();
    if (ndev < 1) { orb_set_error("no CUDA device: this library has no CPU fallback"); return ORB_ERR_CUDA; }
    if (n_devices < 0 || n_devices > ndev || max_batch_per_gpu < 1) { orb_set_error("bad device count / batch"); return ORB_ERR_BAD_ARG; }
    orb_multi* m = new orb_multi();
    m->n = n_devices == 0 ? ndev : n_devices;
    m->max_batch = max_batch_per_gpu;
    for (int d = 0; d < m->n; ++d) {
        orb_extractor *l = nullptr, *r = nullptr;
        int st = orb_extractor_create(nfeatures, scale_factor, nlevels, ini_th, min_th, max_w, max_h, max_batch_per_gpu, d, &l);
        if (st == ORB_OK) st = orb_extractor_create(nfeatures, scale_factor, nlevels, ini_th, min_th, max_w, max_h, max_batch_per_gpu, d, &r);
        if (st != ORB_OK) { if (l) orb_extractor_destroy(l); orb_multi_destroy(m); return st; }
        m->devices.push_back(d); m->left.push_back(l); m->right.push_back(r);
    }
    *out = m;
    return ORB_OK;
}

int orb_multi_device_count(const orb_multi* m) { return m ? m->n : 0; }
int orb_multi_max_keypoints(const orb_multi* m) { return m && m->n ? orb_extractor_max_keypoints(m->left[0]) : 0; }

// Frame::Frame(left, right) for a host batch, sharded over the GPUs of the handle: GPU g takes the contiguous block
// orb_multi_slice(batch, n, g) in sub-batches of at most max_batch_per_gpu; outputs land in the caller's [batch][cap] arrays.
int orb_multi_stereo_frames(orb_multi* m, const uint8_t* imgs_l, const uint8_t* imgs_r, int batch, int w, int h, int stride, size_t img_stride,
                            float mbf, float min_d, float max_d, orb_keypoint* kps_l, uint8_t* desc_l, int32_t* n_l, orb_keypoint* kps_r,
                            uint8_t* desc_r, int32_t* n_r, float* u_right, float* depth, int cap) {
    if (!m || batch < 0 || !n_l || !n_r) return ORB_ERR_BAD_ARG;
    if (batch == 0) return ORB_OK;
    std::vector<int> status(m->n, ORB_OK);
    std::vector<std::string> msg(m->n);
    auto work = [&](int g) {
        int first = 0, count = 0;
        orb_multi_slice(batch, m->n, g, &first, &count);
        for (int done = 0; done < count;) {
            const int nb = std::min(m->max_batch, count - done), b0 = first + done;
            const int st = orb_stereo_frames(m->left[g], m->right[g], imgs_l + (size_t)b0 * img_stride, imgs_r + (size_t)b0 * img_stride, nb, w, h, stride,
                                             img_stride, mbf, min_d, max_d, kps_l + (size_t)b0 * cap, desc_l + (size_t)b0 * cap * 32, n_l + b0,
                                             kps_r + (size_t)b0 * cap, desc_r + (size_t)b0 * cap * 32, n_r + b0, u_right + (size_t)b0 * cap,
                                             depth + (size_t)b0 * cap, cap);
            if (st != ORB_OK) { status[g] = st; msg[g] = orb_last_error(); return; }   // the error text is per thread: carry it over
            done += nb;
        }
    };
    std::vector<std::thread> th;
    for (int g = 1; g < m->n; ++g) th.emplace_back(work, g);
    work(0);
    for (std::thread& t : th) t.join();
    for (int g = 0; g < m->n; ++g)
        if (status[g] != ORB_OK) { orb_set_error("GPU " + std::to_string(m->devices[g]) + ": " + msg[g]); return status[g]; }
    return ORB_OK;
}

// ORBextractor::operator() for a host batch of single images, sharded the same way
int orb_multi_extract_batch(orb_multi* m, const uint8_t* imgs, int batch, int w, int h, int stride, size_t img_stride, orb_keypoint* kps,
                            uint8_t* desc, int cap, int32_t* counts) {
    if (!m || batch < 0 || !counts) return ORB_ERR_BAD_ARG;
    if (batch == 0) return ORB_OK;
    std::vector<int> status(m->n, ORB_OK);
    std::vector<std::string> msg(m->n);
    auto work = [&](int g) {
        int first = 0, count = 0;
        orb_multi_slice(batch, m->n, g, &first, &count);
        for (int done = 0; done < count;) {
            const int nb = std::min(m->max_batch, count - done), b0 = first + done;
            const int st = orb_extract_batch(m->left[g], imgs ? imgs + (size_t)b0 * img_stride : nullptr, nb, w, h, stride, img_stride,
                                             kps ? kps + (size_t)b0 * cap : nullptr, desc ? desc + (size_t)b0 * cap * 32 : nullptr, cap, counts + b0);
            if (st != ORB_OK) { status[g] = st; msg[g] = orb_last_error(); return; }
            done += nb;
        }
    };
    std::vector<std::thread> th;
    for (int g = 1; g < m->n; ++g) th.emplace_back(work, g);
    work(0);
    for (std::thread& t : th) t.join();
    for (int g = 0; g < m->n; ++g)
        if (status[g] != ORB_OK) { orb_set_error("GPU " + std::to_string(m->devices[g]) + ": " + msg[g]); return status[g]; }
    return ORB_OK;
}

}  // extern "C"
