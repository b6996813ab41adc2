"""Host-side mirror of ORB_SLAM2::ORBextractor (reference include/ORBextractor.h:45-114) over the
C ABI.  Same constructor arguments, same call semantics: extractor(image, mask) -> (keypoints,
descriptors) with the mask ignored, an empty image giving empty outputs (src/ORBextractor.cc:1095),
and a non-8UC1 image rejected (the reference asserts, :1100)."""
import ctypes as C
import numpy as np
from . import capi
from .capi import KP_DTYPE, check, lib, ptr


class ORBextractor:
    HARRIS_SCORE = 0
    FAST_SCORE = 1

    def __init__(self, nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, max_width=1241, max_height=376,
                 max_batch=1, device=0):
        self._h = C.c_void_p()
        self.nlevels = int(nlevels)
        check(lib().orb_extractor_create(int(nfeatures), C.c_float(scaleFactor), int(nlevels), int(iniThFAST),
                                         int(minThFAST), int(max_width), int(max_height), int(max_batch), int(device),
                                         C.byref(self._h)))
        self.max_batch = int(max_batch)
        self.kp_cap = lib().orb_extractor_max_keypoints(self._h)
        self._last = None

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            lib().orb_extractor_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close

    @property
    def handle(self):
        return self._h

    # ---- getters of the reference class
    def _tables(self):
        L = self.nlevels
        sf, inv, s2, inv2 = (np.empty(L, np.float32) for _ in range(4))
        npl = np.empty(L, np.int32); umax = np.empty(16, np.int32)
        check(lib().orb_extractor_tables(self._h, ptr(sf), ptr(inv), ptr(s2), ptr(inv2), ptr(npl), ptr(umax)))
        return dict(sf=sf, inv_sf=inv, sigma2=s2, inv_sigma2=inv2, n_per_level=npl, umax=umax)

    def GetLevels(self):
        return self.nlevels

    def GetScaleFactor(self):
        return float(self._tables()["sf"][1]) if self.nlevels > 1 else 1.0

    def GetScaleFactors(self):
        return self._tables()["sf"]

    def GetInverseScaleFactors(self):
        return self._tables()["inv_sf"]

    def GetScaleSigmaSquares(self):
        return self._tables()["sigma2"]

    def GetInverseScaleSigmaSquares(self):
        return self._tables()["inv_sigma2"]

    # ---- operator()
    def __call__(self, image, mask=None):
        image = np.asarray(image)
        if image.size == 0:
            return np.empty(0, KP_DTYPE), np.empty((0, 32), np.uint8)
        if image.dtype != np.uint8 or image.ndim != 2:
            raise ValueError("image must be 8UC1 (reference asserts image.type() == CV_8UC1)")
        kps, desc, counts = self.extract_batch(image[None])
        return kps[0][:counts[0]].copy(), desc[0][:counts[0]].copy()

    def extract_batch(self, images):
        """images: (B, H, W) uint8 host array -> (kps (B, cap), desc (B, cap, 32), counts (B,))."""
        images = np.ascontiguousarray(images, np.uint8)
        B, H, W = images.shape
        cap = self.kp_cap
        kps = np.empty((B, cap), KP_DTYPE); desc = np.empty((B, cap, 32), np.uint8); counts = np.zeros(B, np.int32)
        check(lib().orb_extract_batch(self._h, ptr(images), B, W, H, images.strides[1], C.c_size_t(images.strides[0]),
                                      ptr(kps), ptr(desc), cap, ptr(counts)))
        self._last = (B, W, H)
        return kps, desc, counts

    def extract_enqueue(self, images, out=None):
        """Streaming form: queue a (B, H, W) uint8 host batch, return (ticket, out); out = dict(kps, desc, counts) owns the host
        buffers (pass the dict of an earlier, finished call to reuse them).  Keep `images` alive until extract_wait(ticket, out)."""
        assert images.dtype == np.uint8 and images.ndim == 3 and images.flags.c_contiguous
        B, H, W = images.shape
        cap = self.kp_cap
        if out is None:
            out = dict(kps=np.empty((B, cap), KP_DTYPE), desc=np.empty((B, cap, 32), np.uint8), counts=np.zeros(B, np.int32))
        t = C.c_int(0)
        check(lib().orb_extract_enqueue(self._h, ptr(images), B, W, H, images.strides[1], C.c_size_t(images.strides[0]),
                                        ptr(out["kps"]), ptr(out["desc"]), cap, C.byref(t)))
        self._last = (B, W, H)
        return t.value, out

    def extract_wait(self, ticket, out):
        check(lib().orb_extract_wait(self._h, int(ticket), ptr(out["counts"])))
        return out["kps"], out["desc"], out["counts"]

    def extract_device(self, d_ptr, batch, w, h, stride=None, img_stride=None):
        """Images already resident in HBM (raw device pointer); results stay on the device."""
        stride = stride or w
        img_stride = img_stride or stride * h
        check(lib().orb_extract_device(self._h, C.c_void_p(d_ptr), batch, w, h, stride, C.c_size_t(img_stride)))
        self._last = (batch, w, h)

    def download(self, batch):
        cap = self.kp_cap
        kps = np.empty((batch, cap), KP_DTYPE); desc = np.empty((batch, cap, 32), np.uint8); counts = np.zeros(batch, np.int32)
        check(lib().orb_extractor_download(self._h, batch, ptr(kps), ptr(desc), cap, ptr(counts)))
        return kps, desc, counts

    def sync(self):
        check(lib().orb_extractor_sync(self._h))

    @property
    def stream(self):
        return lib().orb_extractor_stream(self._h)

    @property
    def launch_count(self):
        return lib().orb_extractor_launch_count(self._h)

    STAGES = ("pyramid", "fast_cells", "quadtree", "blur", "orient_describe", "stereo_rows", "stereo_match", "stereo_median")

    def set_timing(self, on=True):
        check(lib().orb_extractor_set_timing(self._h, int(on)))

    def stage_ms(self):
        ms = np.zeros(8, np.float32)
        check(lib().orb_extractor_stage_ms(self._h, ptr(ms)))
        return dict(zip(self.STAGES, (float(v) for v in ms)))

    # ---- mvImagePyramid and stage outputs
    def image_pyramid(self, level, image=0, blurred=False, border=0):
        B, W, H = self._last
        buf = np.empty((H + 2 * border, W + 2 * border), np.uint8)
        w = C.c_int(); h = C.c_int()
        check(lib().orb_get_level(self._h, image, level, int(blurred), border, ptr(buf), buf.strides[0], buf.nbytes,
                                  C.byref(w), C.byref(h)))
        return buf[:h.value + 2 * border, :w.value + 2 * border].copy()

    def level_candidates(self, level, image=0):
        cap = 1 << 20
        out = np.empty((cap, 3), np.int32); n = C.c_int32()
        check(lib().orb_get_level_candidates(self._h, image, level, ptr(out), cap, C.byref(n)))
        return out[:n.value].copy()

    def level_keypoints(self, level, image=0):
        cap = self.kp_cap
        out = np.empty((cap, 3), np.int32); n = C.c_int32()
        check(lib().orb_get_level_keypoints(self._h, image, level, ptr(out), cap, C.byref(n)))
        return out[:n.value].copy()
