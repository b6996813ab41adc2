// The two image steps in front of the extractor, so that raw camera frames can stay on the device:
//   * stereo rectification  cv::remap(im, imRect, M1, M2, cv::INTER_LINEAR)  with CV_32F maps
//     (reference Examples/Stereo/stereo_euroc.cc:97-98 initUndistortRectifyMap(..., CV_32F, M1, M2), :136-137);
//   * cv::cvtColor(im, im, CV_RGB2GRAY | CV_BGR2GRAY | CV_RGBA2GRAY | CV_BGRA2GRAY)  (src/Tracking.cc:174-199).
// Both reproduce OpenCV 4.13's 8-bit fixed point bit for bit (golden vectors made with cv2: tests/golden/rectify.npz).
//
// The maps of a camera never change: at creation they are converted once to the form OpenCV itself
// derives per call (integer source pixel, saturated to short, + 5-bit fractions), stored per
// destination pixel as short2 + u16.  The kernel is a pure gather: 6 B of map, 4 source bytes
// (neighbouring threads share cache lines), 1 B out per pixel; HBM-bound.
#include <algorithm>
#include <climits>
#include <cmath>
#include "orb_internal.h"

struct orb_rectifier {
    int device = 0, dw = 0, dh = 0, sw = 0, sh = 0, max_batch = 0;
    cudaStream_t stream = nullptr;
    short2* d_xy = nullptr;        // [dh * dw] integer source coordinates
    uint16_t* d_a = nullptr;       // [dh * dw] (fy << 5) | fx
    uint8_t* d_src = nullptr;      // staging for host entry points: max_batch * sw * sh * 4 (up to 4 channels)
    uint8_t* d_dst = nullptr;      // max_batch * max(dw * dh, sw * sh)
    long long launches = 0;
};

namespace {

// 4 destination pixels per thread (dw4 = ceil(dw / 4) threads per row).
__global__ void __launch_bounds__(256) k_remap_linear(const short2* __restrict__ xy, const uint16_t* __restrict__ a, int dw, int dh, int dw4,
                                                       const uint8_t* __restrict__ src, int sw, int sh, int sstride, size_t simg,
                                                       uint8_t* __restrict__ dst, int dstride, size_t dimg) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= dw4 * dh) return;
    const int y = t / dw4, x0 = (t - y * dw4) * 4;
    const uint8_t* s = src + blockIdx.y * simg;
    uint8_t* d = dst + blockIdx.y * dimg + (size_t)y * dstride;
    unsigned out = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int x = x0 + k;
        if (x >= dw) break;
        const short2 c = __ldg(xy + (size_t)y * dw + x);
        const int al = __ldg(a + (size_t)y * dw + x);
        const int fx = al & 31, fy = al >> 5, ix = c.x, iy = c.y;
        int v = 0;
        if ((unsigned)ix < (unsigned)(sw - 1) && (unsigned)iy < (unsigned)(sh - 1)) {
            const uint8_t* p = s + (size_t)iy * sstride + ix;
            v = (p[0] * (32 - fx) + p[1] * fx) * (32 - fy) + (p[sstride] * (32 - fx) + p[sstride + 1] * fx) * fy;
        } else if (ix >= -1 && ix < sw && iy >= -1 && iy < sh) {      // touches the border: outside pixels read 0
            const bool x0i = ix >= 0, x1i = ix + 1 < sw, y0i = iy >= 0, y1i = iy + 1 < sh;
            const int p00 = (x0i && y0i) ? s[(size_t)iy * sstride + ix] : 0, p01 = (x1i && y0i) ? s[(size_t)iy * sstride + ix + 1] : 0;
            const int p10 = (x0i && y1i) ? s[(size_t)(iy + 1) * sstride + ix] : 0, p11 = (x1i && y1i) ? s[(size_t)(iy + 1) * sstride + ix + 1] : 0;
            v = (p00 * (32 - fx) + p01 * fx) * (32 - fy) + (p10 * (32 - fx) + p11 * fx) * fy;
        }
        // OpenCV: sum of p * (wy * wx * 32) + 2^14 >> 15  ==  (sum of p * wy * wx + 2^9) >> 10
        out |= (unsigned)((v + 512) >> 10) << (8 * k);
    }
    if (x0 + 3 < dw && (dstride & 3) == 0 && (dimg & 3) == 0) *reinterpret_cast<unsigned*>(d + x0) = out;
    else for (int k = 0; k < 4 && x0 + k < dw; ++k) d[x0 + k] = (uint8_t)(out >> (8 * k));
}

template <int CH>
__global__ void __launch_bounds__(256) k_gray(const uint8_t* __restrict__ src, int w, int h, int sstride, size_t simg, int rgb_order,
                                               uint8_t* __restrict__ dst, int dstride, size_t dimg) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= w) return;
    const uint8_t* p = src + blockIdx.z * simg + (size_t)y * sstride + (size_t)x * CH;
    const int c0 = p[0], g = p[1], c2 = p[2];
    const int r = rgb_order ? c0 : c2, b = rgb_order ? c2 : c0;
    dst[blockIdx.z * dimg + (size_t)y * dstride + x] = (uint8_t)((r * 9798 + g * 19235 + b * 3735 + (1 << 14)) >> 15);
}

}  // namespace

extern "C" {

int orb_rectifier_create(const float* map_x, const float* map_y, int dst_w, int dst_h, int src_w, int src_h, int max_batch, int device,
                         orb_rectifier** out) {
    if (!out) return ORB_ERR_BAD_ARG;
    *out = nullptr;
    if (dst_w < 1 || dst_h < 1 || src_w < 1 || src_h < 1 || max_batch < 1 || ((map_x == nullptr) != (map_y == nullptr))) { orb_set_error("bad rectifier arguments"); return ORB_ERR_BAD_ARG; }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        orb_set_error("no CUDA device: this library has no CPU fallback");
        return ORB_ERR_CUDA;
    }
    if (device < 0 || device >= ndev) { orb_set_error("bad device index"); return ORB_ERR_BAD_ARG; }
    ORB_CUDA(cudaSetDevice(device));
    orb_rectifier* r = new orb_rectifier();
    r->device = device; r->dw = dst_w; r->dh = dst_h; r->sw = src_w; r->sh = src_h; r->max_batch = max_batch;
    const size_t n = (size_t)dst_w * dst_h;
    bool ok = cudaStreamCreateWithFlags(&r->stream, cudaStreamNonBlocking) == cudaSuccess &&
              orb_dev_alloc((void**)&r->d_src, (size_t)max_batch * src_w * src_h * 4) == cudaSuccess &&
              orb_dev_alloc((void**)&r->d_dst, (size_t)max_batch * std::max(n, (size_t)src_w * src_h)) == cudaSuccess;
    if (ok && map_x) {
        // what cv::remap derives from CV_32FC1 maps on every call (imgwarp.cpp, RemapInvoker): sx = cvRound(x * 32) ...
        std::vector<short2> xy(n); std::vector<uint16_t> a(n);
        for (size_t i = 0; i < n; ++i) {
            const float fxs = map_x[i] * 32.f, fys = map_y[i] * 32.f;
            const int sx = (fxs >= -2147483648.f && fxs < 2147483648.f) ? (int)lrintf(fxs) : INT_MIN;   // cvRound: cvtss2si semantics
            const int sy = (fys >= -2147483648.f && fys < 2147483648.f) ? (int)lrintf(fys) : INT_MIN;
            xy[i].x = (short)std::max(-32768, std::min(32767, sx >> 5));
            xy[i].y = (short)std::max(-32768, std::min(32767, sy >> 5));
            a[i] = (uint16_t)(((sy & 31) << 5) | (sx & 31));
        }
        ok = orb_dev_alloc((void**)&r->d_xy, n * sizeof(short2)) == cudaSuccess && orb_dev_alloc((void**)&r->d_a, n * 2) == cudaSuccess &&
             cudaMemcpy(r->d_xy, xy.data(), n * sizeof(short2), cudaMemcpyHostToDevice) == cudaSuccess &&
             cudaMemcpy(r->d_a, a.data(), n * 2, cudaMemcpyHostToDevice) == cudaSuccess;
    }
    if (!ok) {
        orb_set_error(std::string("rectifier allocation: ") + cudaGetErrorString(cudaGetLastError()));
        orb_rectifier_destroy(r);
        return ORB_ERR_CUDA;
    }
    *out = r;
    return ORB_OK;
}

void orb_rectifier_destroy(orb_rectifier* r) {
    if (!r) return;
    cudaSetDevice(r->device);
    if (r->stream) cudaStreamSynchronize(r->stream);
    orb_dev_free(r->d_xy); orb_dev_free(r->d_a); orb_dev_free(r->d_src); orb_dev_free(r->d_dst);
    if (r->stream) cudaStreamDestroy(r->stream);
    delete r;
}
void* orb_rectifier_stream(const orb_rectifier* r) { return r ? (void*)r->stream : nullptr; }
long long orb_rectifier_launch_count(const orb_rectifier* r) { return r ? r->launches : 0; }
int orb_rectifier_sync(orb_rectifier* r) {
    if (!r) return ORB_ERR_BAD_ARG;
    ORB_CUDA(cudaSetDevice(r->device));
    ORB_CUDA(cudaStreamSynchronize(r->stream));
    return ORB_OK;
}

int orb_remap_device(orb_rectifier* r, const uint8_t* d_src, int src_stride, size_t src_img_stride, int batch, uint8_t* d_dst, int dst_stride,
                     size_t dst_img_stride) {
    ORB_NVTX("orb:remap_device");
    if (!r || !d_src || !d_dst || batch < 1 || src_stride < r->sw || dst_stride < r->dw) return ORB_ERR_BAD_ARG;
    if (!r->d_xy) { orb_set_error("rectifier was created without maps"); return ORB_ERR_BAD_ARG; }
    ORB_CUDA(cudaSetDevice(r->device));
    const int dw4 = (r->dw + 3) / 4;
    dim3 grid((dw4 * r->dh + 255) / 256, batch);
    k_remap_linear<<<grid, 256, 0, r->stream>>>(r->d_xy, r->d_a, r->dw, r->dh, dw4, d_src, r->sw, r->sh, src_stride, src_img_stride, d_dst, dst_stride, dst_img_stride);
    r->launches++;
    ORB_CUDA(cudaGetLastError());
    return ORB_OK;
}

int orb_gray_device(orb_rectifier* r, const uint8_t* d_src, int w, int h, int channels, int rgb_order, int src_stride, size_t src_img_stride, int batch,
                    uint8_t* d_dst, int dst_stride, size_t dst_img_stride) {
    if (!r || !d_src || !d_dst || batch < 1 || w < 1 || h < 1 || (channels != 3 && channels != 4) || src_stride < w * channels || dst_stride < w) return ORB_ERR_BAD_ARG;
    ORB_CUDA(cudaSetDevice(r->device));
    dim3 grid((w + 255) / 256, h, batch);
    if (channels == 3) k_gray<3><<<grid, 256, 0, r->stream>>>(d_src, w, h, src_stride, src_img_stride, rgb_order, d_dst, dst_stride, dst_img_stride);
    else k_gray<4><<<grid, 256, 0, r->stream>>>(d_src, w, h, src_stride, src_img_stride, rgb_order, d_dst, dst_stride, dst_img_stride);
    r->launches++;
    ORB_CUDA(cudaGetLastError());
    return ORB_OK;
}

int orb_remap(orb_rectifier* r, const uint8_t* src, int src_stride, int batch, uint8_t* dst, int dst_stride) {
    ORB_NVTX("orb:remap");
    if (!r || !src || !dst || batch < 1 || src_stride < r->sw || dst_stride < r->dw) return ORB_ERR_BAD_ARG;
    if (batch > r->max_batch) { orb_set_error("batch exceeds the rectifier's max_batch"); return ORB_ERR_CAPACITY; }
    ORB_CUDA(cudaSetDevice(r->device));
    const size_t simg = (size_t)r->sw * r->sh, dimg = (size_t)r->dw * r->dh;
    ORB_CUDA(cudaMemcpy2DAsync(r->d_src, r->sw, src, src_stride, r->sw, (size_t)r->sh * batch, cudaMemcpyHostToDevice, r->stream));
    const int st = orb_remap_device(r, r->d_src, r->sw, simg, batch, r->d_dst, r->dw, dimg);
    if (st != ORB_OK) return st;
    ORB_CUDA(cudaMemcpy2DAsync(dst, dst_stride, r->d_dst, r->dw, r->dw, (size_t)r->dh * batch, cudaMemcpyDeviceToHost, r->stream));
    ORB_CUDA(cudaStreamSynchronize(r->stream));
    return ORB_OK;
}

int orb_gray(orb_rectifier* r, const uint8_t* src, int w, int h, int channels, int rgb_order, int src_stride, uint8_t* dst, int dst_stride) {
    if (!r || !src || !dst || w < 1 || h < 1 || (channels != 3 && channels != 4) || src_stride < w * channels || dst_stride < w) return ORB_ERR_BAD_ARG;
    if ((size_t)w * h > (size_t)r->max_batch * r->sw * r->sh) { orb_set_error("image larger than the rectifier's staging area"); return ORB_ERR_CAPACITY; }
    ORB_CUDA(cudaSetDevice(r->device));
    ORB_CUDA(cudaMemcpy2DAsync(r->d_src, (size_t)w * channels, src, src_stride, (size_t)w * channels, h, cudaMemcpyHostToDevice, r->stream));
    const int st = orb_gray_device(r, r->d_src, w, h, channels, rgb_order, w * channels, 0, 1, r->d_dst, w, 0);
    if (st != ORB_OK) return st;
    ORB_CUDA(cudaMemcpy2DAsync(dst, dst_stride, r->d_dst, w, w, h, cudaMemcpyDeviceToHost, r->stream));
    ORB_CUDA(cudaStreamSynchronize(r->stream));
    return ORB_OK;
}

}  // extern "C"
