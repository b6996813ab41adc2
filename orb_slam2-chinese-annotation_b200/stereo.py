"""Frame::ComputeStereoMatches (reference src/Frame.cc:516-704) over the C ABI.

The reference method reads both extractors' pyramids and the frame's keypoints/descriptors; here
the two ORBextractor handles keep all of that device-resident from their last call, so the
function takes the handles plus the camera constants (mbf, and max_d = mbf / mb, which the
reference computes from Frame::mb before it is assigned; pass fx for the upstream intent)."""
import ctypes as C
import numpy as np
from .capi import KP_DTYPE, check, lib, ptr


def compute_stereo_matches(left, right, mbf, max_d, min_d=0.0, batch=1):
    """-> (mvuRight, mvDepth, counts): float32 arrays (batch, cap); -1 marks no match."""
    cap = left.kp_cap
    ur = np.empty((batch, cap), np.float32); dp = np.empty((batch, cap), np.float32); n = np.zeros(batch, np.int32)
    check(lib().orb_stereo_match_batch(left.handle, right.handle, batch, C.c_float(mbf), C.c_float(min_d), C.c_float(max_d),
                                       ptr(ur), ptr(dp), cap, ptr(n)))
    return ur, dp, n


def compute_stereo_matches_device(left, right, mbf, max_d, min_d=0.0, batch=1):
    """Asynchronous; results stay in the left handle's device buffers."""
    check(lib().orb_stereo_match_device(left.handle, right.handle, batch, C.c_float(mbf), C.c_float(min_d), C.c_float(max_d)))


def stereo_frames(left, right, images_l, images_r, mbf, max_d, min_d=0.0, out=None):
    """The stereo Frame constructor (reference src/Frame.cc:60-115) for a batch of host image pairs
    (B, H, W) uint8: ExtractORB left + right, then ComputeStereoMatches, chunk-pipelined on the
    device.  -> dict(kps_l, desc_l, n_l, kps_r, desc_r, n_r, u_right, depth); arrays are
    (B, cap, ...) and valid up to n_l[b] / n_r[b].  `out` may hold preallocated (pinned) arrays."""
    images_l = np.ascontiguousarray(images_l, np.uint8); images_r = np.ascontiguousarray(images_r, np.uint8)
    if images_l.shape != images_r.shape or images_l.ndim != 3:
        raise ValueError("left/right batches must both be (B, H, W)")
    B, H, W = images_l.shape
    cap = left.kp_cap
    o = out or {}
    for k, shape, dt in (("kps_l", (B, cap), KP_DTYPE), ("desc_l", (B, cap, 32), np.uint8), ("n_l", (B,), np.int32),
                         ("kps_r", (B, cap), KP_DTYPE), ("desc_r", (B, cap, 32), np.uint8), ("n_r", (B,), np.int32),
                         ("u_right", (B, cap), np.float32), ("depth", (B, cap), np.float32)):
        if k not in o:
            o[k] = np.empty(shape, dt)
    check(lib().orb_stereo_frames(left.handle, right.handle, ptr(images_l), ptr(images_r), B, W, H, images_l.strides[1],
                                  C.c_size_t(images_l.strides[0]), C.c_float(mbf), C.c_float(min_d), C.c_float(max_d),
                                  ptr(o["kps_l"]), ptr(o["desc_l"]), ptr(o["n_l"]), ptr(o["kps_r"]), ptr(o["desc_r"]), ptr(o["n_r"]),
                                  ptr(o["u_right"]), ptr(o["depth"]), cap))
    left._last = right._last = (B, W, H)
    return o


def stereo_frames_enqueue(left, right, images_l, images_r, mbf, max_d, out, min_d=0.0):
    """Asynchronous stereo_frames: queues the batch and returns a ticket; `out` must hold preallocated (pinned) arrays kps_l, desc_l,
    kps_r, desc_r, u_right, depth and, like the (contiguous) images, stay alive and untouched until stereo_frames_wait(ticket)."""
    for a in (images_l, images_r):
        if a.dtype != np.uint8 or a.ndim != 3 or not a.flags["C_CONTIGUOUS"]:
            raise ValueError("enqueue takes C-contiguous (B, H, W) uint8 batches (they are read after the call returns, so no copy is made here)")
    if images_l.shape != images_r.shape:
        raise ValueError("left/right batches must have one shape")
    B, H, W = images_l.shape
    t = C.c_int(0)
    check(lib().orb_stereo_frames_enqueue(left.handle, right.handle, ptr(images_l), ptr(images_r), B, W, H, images_l.strides[1],
                                          C.c_size_t(images_l.strides[0]), C.c_float(mbf), C.c_float(min_d), C.c_float(max_d),
                                          ptr(out["kps_l"]), ptr(out["desc_l"]), ptr(out["kps_r"]), ptr(out["desc_r"]), ptr(out["u_right"]),
                                          ptr(out["depth"]), left.kp_cap, C.byref(t)))
    left._last = right._last = (B, W, H)
    return t.value


def stereo_frames_wait(left, right, ticket, out):
    """Blocks until the batch of `ticket` is in `out`; fills out["n_l"], out["n_r"]."""
    check(lib().orb_stereo_frames_wait(left.handle, right.handle, int(ticket), ptr(out["n_l"]), ptr(out["n_r"])))
    return out


class MultiGpuFrontEnd:
    """The library's multi-GPU front end (orb_multi_*): one handle pair and one host thread per visible GPU, a host batch is cut
    into contiguous blocks (reference: every frame is independent)."""

    def __init__(self, nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, max_width, max_height, max_batch_per_gpu, n_devices=0):
        self._h = C.c_void_p()
        check(lib().orb_multi_create(int(nfeatures), C.c_float(scaleFactor), int(nlevels), int(iniThFAST), int(minThFAST), int(max_width),
                                     int(max_height), int(max_batch_per_gpu), int(n_devices), C.byref(self._h)))
        lib().orb_multi_device_count.argtypes = [C.c_void_p]; lib().orb_multi_max_keypoints.argtypes = [C.c_void_p]
        self.n_devices = lib().orb_multi_device_count(self._h)
        self.kp_cap = lib().orb_multi_max_keypoints(self._h)

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            lib().orb_multi_destroy.argtypes = [C.c_void_p]
            lib().orb_multi_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close

    def stereo_frames(self, images_l, images_r, mbf, max_d, min_d=0.0, out=None):
        images_l = np.ascontiguousarray(images_l, np.uint8); images_r = np.ascontiguousarray(images_r, np.uint8)
        B, H, W = images_l.shape
        cap = self.kp_cap
        o = out or {}
        for k, shape, dt in (("kps_l", (B, cap), KP_DTYPE), ("desc_l", (B, cap, 32), np.uint8), ("n_l", (B,), np.int32),
                             ("kps_r", (B, cap), KP_DTYPE), ("desc_r", (B, cap, 32), np.uint8), ("n_r", (B,), np.int32),
                             ("u_right", (B, cap), np.float32), ("depth", (B, cap), np.float32)):
            if k not in o:
                o[k] = np.empty(shape, dt)
        check(lib().orb_multi_stereo_frames(self._h, ptr(images_l), ptr(images_r), B, W, H, images_l.strides[1], C.c_size_t(images_l.strides[0]),
                                            C.c_float(mbf), C.c_float(min_d), C.c_float(max_d), ptr(o["kps_l"]), ptr(o["desc_l"]), ptr(o["n_l"]),
                                            ptr(o["kps_r"]), ptr(o["desc_r"]), ptr(o["n_r"]), ptr(o["u_right"]), ptr(o["depth"]), cap))
        return o


def multi_slice(batch, n_parts, part):
    f = C.c_int(0); c = C.c_int(0)
    lib().orb_multi_slice(int(batch), int(n_parts), int(part), C.byref(f), C.byref(c))
    return f.value, c.value
