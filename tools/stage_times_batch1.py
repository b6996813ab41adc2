#!/usr/bin/env python
"""Stage times of ONE image through the extractor (batch 1, the live-camera case): where the per-frame latency goes.
   python tools/stage_times_batch1.py [kitti|tum]"""
import importlib, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "orb_slam2-chinese-annotation_b200"
synth = importlib.import_module(PKG + ".synth"); ext = importlib.import_module(PKG + ".extractor")
cfg = {"kitti": (1241, 376, 2000), "tum": (640, 480, 1000)}[sys.argv[1] if len(sys.argv) > 1 else "kitti"]
w, h, nf = cfg
img = synth.make_image(w, h, 0)
g = ext.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
for _ in range(5): g(img)
names = ["pyramid", "fast_cells", "quadtree", "blur", "orient_describe"]
g.set_timing(True)
acc = None
for _ in range(20):
    g(img)
    d = g.stage_ms()
    s = np.array([d[n] for n in names])
    acc = s if acc is None else acc + s
g.set_timing(False)
print({n: round(float(v) / 20 * 1000, 1) for n, v in zip(names, acc)}, "us; sum", round(float(acc.sum()) / 20 * 1000, 1), "us")
t0 = time.perf_counter()
for _ in range(200): g(img)
print("host call (numpy in, numpy out): %.1f us per frame" % ((time.perf_counter() - t0) / 200 * 1e6))
