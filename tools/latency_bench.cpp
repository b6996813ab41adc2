// Per-frame latency through the C++ drop-in classes, timed the way the reference's examples time a frame
// (Examples/Stereo/stereo_kitti.cc:80-97,113-122: steady_clock around the frame, then median / mean).
//
//   stereo: Frame::Frame(left, right) as the reference runs it (src/Frame.cc:81-95): ORBextractor::operator() on the left
//           and right image from two std::threads created for the frame, joined, then ComputeStereoMatches
//           (= orb_stereo_match on the device-resident pyramids).
//   mono:   ORBextractor::operator() on one image (src/Frame.cc:187).
//
//   latency_bench <mode stereo|mono> <images.bin> <w> <h> <pool> <nfeatures> <mbf> <fx> <frames> <warmup> [keep_pyramid 0|1] [threads 0|1|2]
// images.bin: <pool> left images then <pool> right images (mono: <pool> images), w*h bytes each.  Prints one JSON object.
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <thread>
#include <vector>

#include "ORBextractor.h"

int main(int argc, char** argv) {
    if (argc < 11) { fprintf(stderr, "usage: see the header comment\n"); return 2; }
    const std::string mode = argv[1];
    const int w = atoi(argv[3]), h = atoi(argv[4]), pool = atoi(argv[5]), nf = atoi(argv[6]);
    const float mbf = (float)atof(argv[7]), fx = (float)atof(argv[8]);
    const int frames = atoi(argv[9]), warm = atoi(argv[10]);
    const bool keep = argc > 11 && atoi(argv[11]) != 0;
    const bool use_threads = !(argc > 12 && atoi(argv[12]) == 0);
    const bool single_call = argc > 12 && atoi(argv[12]) == 2;   // ORBextractor::ExtractStereo: one orb_stereo_frames call per pair
    const bool stereo = mode == "stereo";
    const size_t px = (size_t)w * h, nimg = (size_t)pool * (stereo ? 2 : 1);
    std::vector<unsigned char> buf(px * nimg);
    FILE* f = fopen(argv[2], "rb");
    if (!f || fread(buf.data(), 1, buf.size(), f) != buf.size()) { fprintf(stderr, "cannot read %s\n", argv[2]); return 2; }
    fclose(f);
    try {
        ORB_SLAM2::ORBextractor left(nf, 1.2f, 8, 20, 7), right(nf, 1.2f, 8, 20, 7);
        left.Reserve(w, h, 1); right.Reserve(w, h, 1);
        left.SetKeepPyramidOnHost(keep); right.SetKeepPyramidOnHost(keep);
        std::vector<cv::KeyPoint> kl, kr;
        cv::Mat dl, dr;
        std::vector<float> u_right, depth;
        std::vector<double> ms;
        long long kp_sum = 0, match_sum = 0;
        for (int i = 0; i < warm + frames; ++i) {
            cv::Mat iml(h, w, CV_8U, buf.data() + px * (size_t)(i % pool), (size_t)w);
            cv::Mat imr(h, w, CV_8U, buf.data() + px * (size_t)((stereo ? pool : 0) + i % pool), (size_t)w);
            const auto t0 = std::chrono::steady_clock::now();
            int nl = 0;
            if (stereo) {
                if (single_call) {
                    ORB_SLAM2::ORBextractor::ExtractStereo(left, right, iml, imr, mbf, 0.f, fx, kl, dl, kr, dr, u_right, depth);
                    nl = (int)kl.size();
                } else {
                if (use_threads) {
                    std::thread tl([&] { left(iml, cv::Mat(), kl, dl); });
                    std::thread tr([&] { right(imr, cv::Mat(), kr, dr); });
                    tl.join(); tr.join();
                } else {
                    left(iml, cv::Mat(), kl, dl);
                    right(imr, cv::Mat(), kr, dr);
                }
                const int cap = (int)kl.size();
                u_right.resize(cap > 0 ? cap : 1); depth.resize(cap > 0 ? cap : 1);
                int32_t n = 0;
                if (orb_stereo_match(left.Handle(), right.Handle(), 0, mbf, 0.f, fx, u_right.data(), depth.data(), cap, &n) != ORB_OK) {
                    fprintf(stderr, "orb_stereo_match: %s\n", orb_last_error());
                    return 1;
                }
                nl = n;
                }
            } else {
                left(iml, cv::Mat(), kl, dl);
                nl = (int)kl.size();
            }
            const auto t1 = std::chrono::steady_clock::now();
            if (i >= warm) {
                ms.push_back(std::chrono::duration<double, std::milli>(t1 - t0).count());
                kp_sum += nl;
                if (stereo) for (int k = 0; k < nl; ++k) match_sum += depth[k] > 0.f;
            }
        }
        std::sort(ms.begin(), ms.end());
        double mean = 0;
        for (double v : ms) mean += v;
        mean /= ms.size();
        printf("{\"mode\": \"%s\", \"frames\": %d, \"warmup\": %d, \"median_ms\": %.4f, \"mean_ms\": %.4f, \"p95_ms\": %.4f, \"min_ms\": %.4f, "
               "\"frames_per_s\": %.1f, \"avg_keypoints_left\": %.1f, \"avg_stereo_matches\": %.1f, \"keep_pyramid_on_host\": %s, "
               "\"extractor_threads\": %d, \"single_call\": %s}\n",
               mode.c_str(), frames, warm, ms[ms.size() / 2], mean, ms[(size_t)(ms.size() * 0.95)], ms[0], 1e3 / mean,
               (double)kp_sum / frames, (double)match_sum / frames, keep ? "true" : "false", stereo && use_threads && !single_call ? 2 : 1, single_call ? "true" : "false");
    } catch (const std::exception& e) {
        fprintf(stderr, "latency_bench: %s\n", e.what());
        return 1;
    }
    return 0;
}
