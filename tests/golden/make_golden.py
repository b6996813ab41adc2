#!/usr/bin/env python
"""Regenerates tests/golden/primitives.npz from the REAL OpenCV (cv2 4.13.0, the only OpenCV on
the authoring box; the reference's own OpenCV dependency is un-vendored and un-pinned,
/root/reference/CMakeLists.txt:31-37).  Run in the authoring container: python tests/golden/make_golden.py"""
import importlib, os, sys
import numpy as np
import cv2

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
synth = importlib.import_module("orb_slam2-chinese-annotation_b200.synth")
cv2.setNumThreads(1)
assert cv2.__version__.startswith("4.13"), cv2.__version__

w, h, idx = 310, 188, 42
img = synth.make_image(w, h, idx)
rw, rh = int(round(w / 1.2)), int(round(h / 1.2))


def fast(th):
    fd = cv2.FastFeatureDetector_create(th, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
    return np.array([(int(k.pt[0]), int(k.pt[1]), int(k.response)) for k in fd.detect(img)], np.int32).reshape(-1, 3)


rng = np.random.default_rng(1)
atan_in = rng.integers(-60000, 60000, (4096, 2)).astype(np.float32)
atan_in[:16] = [[0, 0], [0, 1], [1, 0], [0, -1], [-1, 0], [1, 1], [-1, 1], [1, -1], [-1, -1], [5, 5], [3, 7], [7, 3],
                [-3, 7], [3, -7], [100, 0], [0, 100]]
atan_out = np.array([cv2.fastAtan2(float(y), float(x)) for y, x in atan_in], np.float32)
np.savez_compressed(os.path.join(os.path.dirname(__file__), "primitives.npz"),
                    w=w, h=h, idx=idx, img=img, rw=rw, rh=rh,
                    resized=cv2.resize(img, (rw, rh), interpolation=cv2.INTER_LINEAR),
                    blurred=cv2.GaussianBlur(img, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101),
                    bordered=cv2.copyMakeBorder(img, 19, 19, 19, 19, cv2.BORDER_REFLECT_101),
                    fast7=fast(7), fast20=fast(20), atan_in=atan_in, atan_out=atan_out, cv2_version=cv2.__version__)
print("wrote primitives.npz")
