#!/usr/bin/env python
"""Regenerates tests/golden/rectify.npz from the REAL OpenCV (cv2 4.13.0): cv2.remap(INTER_LINEAR) with float maps and
cv2.cvtColor to gray, the two steps in front of the extractor (reference Examples/Stereo/stereo_euroc.cc:136-137,
src/Tracking.cc:174-199).  Run in the authoring container: python tests/golden/make_golden_rectify.py"""
import importlib, os, sys
import numpy as np
import cv2

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
synth = importlib.import_module("orb_slam2-chinese-annotation_b200.synth")
cv2.setNumThreads(1)
assert cv2.__version__.startswith("4.13"), cv2.__version__

w, h = 188, 120
img = synth.make_image(w, h, 7)
mx, my = synth.make_rectify_maps(w, h, 3)
# hand-placed edge cases: exact half steps of 1/32 px (round half to even), borders, far outside, NaN
mx[0, :8] = [-1.0, -0.5, -0.015625, 0.015625, w - 1, w - 0.5, w - 1.015625, 1e9]; my[0, :8] = 5.0
my[1, :8] = [-1.0, -0.5, -0.015625, 0.046875, h - 1, h - 0.5, h - 1.015625, -1e9]; mx[1, :8] = 7.25
mx[2, 0] = np.nan; my[2, 1] = np.inf; mx[2, 2] = 40000.0; my[2, 3] = -40000.0
rng = np.random.default_rng(2)
bgr = rng.integers(0, 256, (h, w, 3), dtype=np.uint8); bgra = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
np.savez_compressed(os.path.join(os.path.dirname(__file__), "rectify.npz"), w=w, h=h, img=img, mx=mx, my=my,
                    remapped=cv2.remap(img, mx, my, cv2.INTER_LINEAR), bgr=bgr, bgra=bgra,
                    bgr2gray=cv2.cvtColor(bgr, cv2.COLOR_BGR2GRAY), rgb2gray=cv2.cvtColor(bgr, cv2.COLOR_RGB2GRAY),
                    bgra2gray=cv2.cvtColor(bgra, cv2.COLOR_BGRA2GRAY), rgba2gray=cv2.cvtColor(bgra, cv2.COLOR_RGBA2GRAY), cv2_version=cv2.__version__)
print("wrote rectify.npz")
