"""How much of the CPU baseline's time is the SCALAR stand-ins for OpenCV's primitives?

The timed CPU baseline (bench.py cpu_baseline / --impl reference) is the reference's own src/ORBextractor.cc compiled against
oracle/cvshim, whose cv::resize / cv::GaussianBlur / cv::FAST are scalar restatements (bit-exact with OpenCV 4.13, but without its
SIMD kernels).  This tool times, on the same synthetic images and on ONE thread each:
    scalar:  the oracle's resize_linear / gauss7 / fast9_16 over the 8 levels of an image   (what the baseline runs)
    cv2:     cv2.resize / cv2.GaussianBlur / cv2.FastFeatureDetector over the same levels   (what a stock OpenCV build runs)
and rescales a measured baseline:  t_image' = t_image - scalar + cv2.  FAST is timed once per whole level with the initial
threshold on both sides (the reference calls it per 30x30 cell, and a second time with the lower threshold on empty cells; both
sides would pay the same call pattern).  TEST / MEASUREMENT INFRASTRUCTURE: imports the oracle, never the product.

    python tools/cv2_baseline.py [--config kitti|euroc|tum|uhd] [--images N]      -> one JSON object
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))


def primitive_times(images, nfeat, scale=1.2, nlevels=8, ini_th=20, repeats=3):
    """-> dict(scalar_ms, cv2_ms, per-primitive ms) per image, single thread; None if cv2 is not importable."""
    try:
        import cv2
    except ImportError:
        return None
    import orc
    cv2.setNumThreads(1)
    o = orc.Extractor(nfeat, scale, nlevels, ini_th, 7)
    det = cv2.FastFeatureDetector_create(ini_th, True)
    acc = {k: 0.0 for k in ("scalar_resize", "scalar_blur", "scalar_fast", "cv2_resize", "cv2_blur", "cv2_fast")}
    n = 0
    for img in images:
        img = np.ascontiguousarray(img)
        o.extract(img)
        sizes = [o.level(l).shape for l in range(nlevels)]
        for _ in range(repeats):
            lv = [img]
            t0 = time.perf_counter()
            for l in range(1, nlevels):
                lv.append(orc.resize_linear(lv[-1], sizes[l][1], sizes[l][0]))
            t1 = time.perf_counter()
            for m in lv:
                orc.gauss7(m)
            t2 = time.perf_counter()
            for m in lv:
                orc.fast9_16(m, ini_th)
            t3 = time.perf_counter()
            cl = [img]
            for l in range(1, nlevels):
                cl.append(cv2.resize(cl[-1], (sizes[l][1], sizes[l][0]), interpolation=cv2.INTER_LINEAR))
            t4 = time.perf_counter()
            for m in cl:
                cv2.GaussianBlur(m, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101)
            t5 = time.perf_counter()
            for m in cl:
                det.detect(m)
            t6 = time.perf_counter()
            for k, v in zip(acc, (t1 - t0, t2 - t1, t3 - t2, t4 - t3, t5 - t4, t6 - t5)):
                acc[k] += v
            n += 1
    out = {k: 1e3 * v / n for k, v in acc.items()}
    out["scalar_ms"] = out["scalar_resize"] + out["scalar_blur"] + out["scalar_fast"]
    out["cv2_ms"] = out["cv2_resize"] + out["cv2_blur"] + out["cv2_fast"]
    out["cv2_version"] = cv2.__version__
    return out


def adjust(frames_per_s, cores, images_per_frame, prim):
    """A measured all-core baseline (frames/s) rescaled as if the three primitives ran at cv2's single-thread speed."""
    t_image = cores / frames_per_s / images_per_frame * 1e3          # ms of one core per image
    t_adj = max(t_image - prim["scalar_ms"] + prim["cv2_ms"], 1e-6)
    return frames_per_s * t_image / t_adj, t_image, t_adj


def main():
    import importlib
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="kitti")
    ap.add_argument("--images", type=int, default=4)
    a = ap.parse_args()
    synth = importlib.import_module("orb_slam2-chinese-annotation_b200.synth")
    geo = {"kitti": (1241, 376, 2000), "euroc": (752, 480, 1200), "tum": (640, 480, 1000), "uhd": (3840, 2160, 8000)}[a.config]
    imgs = [synth.make_image(geo[0], geo[1], i) for i in range(a.images)]
    p = primitive_times(imgs, geo[2])
    print(json.dumps({"config": a.config, "images": a.images, "per_image_ms_one_thread": p}))


if __name__ == "__main__":
    main()
