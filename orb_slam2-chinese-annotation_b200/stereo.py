"""Frame::ComputeStereoMatches (reference src/Frame.cc:516-704) over the C ABI.

The reference method reads both extractors' pyramids and the frame's keypoints/descriptors; here
the two ORBextractor handles keep all of that device-resident from their last call, so the
function takes the handles plus the camera constants (mbf, and max_d = mbf / mb, which the
reference computes from Frame::mb before it is assigned; pass fx for the upstream intent)."""
import ctypes as C
import numpy as np
from .capi import check, lib, ptr


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
