#!/usr/bin/env python
"""Writes a pool of synthetic images for tools/latency_bench: <pool> left images then (stereo) <pool> right images.
   tools/make_images.py <out.bin> <w> <h> <pool> <fx> [mono]"""
import importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
synth = importlib.import_module("orb_slam2-chinese-annotation_b200.synth")
out, w, h, pool, fx = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), float(sys.argv[5])
pairs = [synth.make_stereo_pair(w, h, i, fx=fx) for i in range(pool)]
imgs = [p[0] for p in pairs] + ([] if len(sys.argv) > 6 else [p[1] for p in pairs])
np.stack(imgs).tofile(out)
