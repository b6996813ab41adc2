"""B200-native ORB-SLAM2 front-end hot path (extract + match) behind the reference's API.

The compute path is hand-written sm_100a CUDA in csrc/, exposed through the C ABI declared in
include/orb_b200.h and loaded here with ctypes (see capi.py).  There is no CPU fallback.
"""
