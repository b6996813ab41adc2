"""Stereo rectification (cv::remap INTER_LINEAR with the float maps of initUndistortRectifyMap, reference
Examples/Stereo/stereo_euroc.cc:97-98,136-137) and gray conversion (cv::cvtColor, src/Tracking.cc:174-199) on the device,
through the C ABI (csrc/rectify.cu).  No CPU fallback."""
import ctypes as C
import numpy as np
from .capi import check, lib, ptr


class Rectifier:
    def __init__(self, map_x=None, map_y=None, src_size=None, max_batch=1, device=0):
        """map_x / map_y: (dst_h, dst_w) float32 (M1, M2); src_size = (w, h) of the raw image (default: the maps' size)."""
        self._h = C.c_void_p()
        if map_x is None:
            self.map_x = self.map_y = None
            sw, sh = src_size; dw, dh = sw, sh
        else:
            self.map_x = np.ascontiguousarray(map_x, np.float32); self.map_y = np.ascontiguousarray(map_y, np.float32)
            dh, dw = self.map_x.shape
            sw, sh = src_size if src_size is not None else (dw, dh)
        self.dw, self.dh, self.sw, self.sh, self.max_batch = dw, dh, sw, sh, int(max_batch)
        check(lib().orb_rectifier_create(ptr(self.map_x), ptr(self.map_y), dw, dh, sw, sh, self.max_batch, int(device), C.byref(self._h)))

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            lib().orb_rectifier_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close

    @property
    def launch_count(self):
        return lib().orb_rectifier_launch_count(self._h)

    def remap(self, img):
        """img: (h, w) or (batch, h, w) uint8 -> rectified, same leading shape."""
        img = np.ascontiguousarray(img, np.uint8)
        single = img.ndim == 2
        if single:
            img = img[None]
        B = img.shape[0]
        assert img.shape[1:] == (self.sh, self.sw), img.shape
        out = np.empty((B, self.dh, self.dw), np.uint8)
        check(lib().orb_remap(self._h, ptr(img), self.sw, B, ptr(out), self.dw))
        return out[0] if single else out

    def gray(self, img, rgb_order):
        """img: (h, w, 3|4) uint8; rgb_order: channel 0 is R (CV_RGB2GRAY / CV_RGBA2GRAY) else B."""
        img = np.ascontiguousarray(img, np.uint8)
        h, w, ch = img.shape
        out = np.empty((h, w), np.uint8)
        check(lib().orb_gray(self._h, ptr(img), w, h, ch, int(bool(rgb_order)), w * ch, ptr(out), w))
        return out
