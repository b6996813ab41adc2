"""ctypes binding of the C ABI (include/orb_b200.h) implemented by lib/liborb_b200.so.

Fails loudly when the CUDA library is missing or no device is usable: there is no CPU fallback.
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("ORB_B200_LIB") or os.path.join(_HERE, "lib", "liborb_b200.so")  # override: kernel-variant experiments

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])  # == cv::KeyPoint, 28 bytes

ORB_OK, ORB_ERR_BAD_ARG, ORB_ERR_CAPACITY, ORB_ERR_CUDA, ORB_ERR_INTERNAL = range(5)

EXPORTS = [
    "orb_last_error", "orb_device_count", "orb_extractor_create", "orb_extractor_destroy", "orb_extractor_tables",
    "orb_extractor_max_keypoints", "orb_extract", "orb_extract_batch", "orb_extract_device",
    "orb_extractor_device_results", "orb_extractor_download", "orb_extractor_sync", "orb_extractor_stream",
    "orb_extractor_launch_count", "orb_extractor_set_timing", "orb_extractor_stage_ms", "orb_get_level", "orb_get_pyramid", "orb_debug_guard_check", "orb_fast_atan2_device", "orb_level_size", "orb_get_level_candidates", "orb_get_level_keypoints",
    "orb_stereo_match", "orb_stereo_match_batch", "orb_stereo_match_device", "orb_stereo_device_results", "orb_stereo_frames",
    "orb_multi_create", "orb_multi_destroy", "orb_multi_device_count", "orb_multi_max_keypoints", "orb_multi_slice",
    "orb_multi_stereo_frames", "orb_multi_extract_batch",
    "orb_stereo_frames_enqueue", "orb_stereo_frames_wait", "orb_extract_enqueue", "orb_extract_wait", "orb_extractor_set_copy_only",
    "orb_matcher_create", "orb_matcher_destroy", "orb_matcher_stream", "orb_matcher_launch_count", "orb_matcher_sync",
    "orb_descriptor_distance", "orb_match_bruteforce", "orb_match_bruteforce_device", "orb_assign_features_to_grid",
    "orb_assign_features_to_grid_static",
    "orb_search_by_projection_mappoints", "orb_search_by_projection_lastframe", "orb_search_by_bow",
    "orb_search_by_projection_mappoints_batch", "orb_search_by_projection_lastframe_batch", "orb_matcher_last_candidates",
    "orb_search_for_triangulation", "orb_search_by_bow_keyframes", "orb_search_by_projection_claiming",
    "orb_search_projected_best", "orb_search_by_sim3", "orb_search_for_initialization", "orb_sincosf_sweep_device",
    "orb_vocabulary_create", "orb_vocabulary_destroy", "orb_vocabulary_info", "orb_vocabulary_stream", "orb_vocabulary_launch_count",
    "orb_vocabulary_sync", "orb_bow_transform", "orb_bow_transform_device", "orb_is_in_frustum", "orb_distinctive_descriptors",
    "orb_rectifier_create", "orb_rectifier_destroy", "orb_rectifier_stream", "orb_rectifier_launch_count", "orb_rectifier_sync",
    "orb_remap", "orb_gray", "orb_remap_device", "orb_gray_device",
]


class OrbError(RuntimeError):
    def __init__(self, status, msg):
        super().__init__("orb_b200 status %d: %s" % (status, msg))
        self.status = status


def build(force=False):
    """Compile the CUDA library in-tree (nvcc, sm_100a)."""
    if force:
        subprocess.check_call(["make", "-s", "-C", os.path.join(_HERE, "csrc"), "clean"])
    subprocess.check_call(["make", "-s", "-C", os.path.join(_HERE, "csrc")])


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise OrbError(ORB_ERR_CUDA, "CUDA library %s is not built (run __graft_entry__.build()); no CPU fallback" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        L.orb_last_error.restype = C.c_char_p
        L.orb_extractor_stream.restype = C.c_void_p
        L.orb_matcher_stream.restype = C.c_void_p
        L.orb_extractor_launch_count.restype = C.c_longlong
        L.orb_matcher_launch_count.restype = C.c_longlong
        L.orb_matcher_last_candidates.restype = C.c_longlong
        L.orb_vocabulary_stream.restype = C.c_void_p
        L.orb_rectifier_stream.restype = C.c_void_p
        L.orb_rectifier_launch_count.restype = C.c_longlong
        L.orb_vocabulary_launch_count.restype = C.c_longlong
        for name in ("orb_extractor_stream", "orb_extractor_launch_count", "orb_extractor_max_keypoints",
                     "orb_extractor_destroy", "orb_extractor_sync", "orb_matcher_stream", "orb_matcher_launch_count",
                     "orb_matcher_last_candidates", "orb_matcher_destroy", "orb_matcher_sync", "orb_vocabulary_stream", "orb_vocabulary_launch_count",
                     "orb_vocabulary_destroy", "orb_vocabulary_sync", "orb_rectifier_stream", "orb_rectifier_launch_count",
                     "orb_rectifier_destroy", "orb_rectifier_sync"):
            getattr(L, name).argtypes = [C.c_void_p]
        _lib = L
    return _lib


def check(status):
    if status != ORB_OK:
        raise OrbError(status, lib().orb_last_error().decode())


def ptr(a):
    if a is None:
        return C.c_void_p(0)
    if isinstance(a, int):
        return C.c_void_p(a)
    return a.ctypes.data_as(C.c_void_p)
