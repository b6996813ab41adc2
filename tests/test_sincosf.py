"""glibc sincosf restatement used by the rBRIEF kernel (csrc/sincosf_glibc.h) vs the host libm, for
EVERY float in [0, 6.2832] (the reference's cos/sin(angle*factorPI), src/ORBextractor.cc:124-125,
compile to one sincosf call).  CPU: the header compiled as C, with and without FMA contraction.
GPU: the same sweep on the device, compared through a 64-bit checksum."""
import ctypes as C
import os
import struct
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LAST = struct.unpack("<I", struct.pack("<f", 6.2832))[0]   # bit pattern of the largest swept float


def _sweep(binary, first=0, last=LAST):
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "tests", "cpp"), "build/" + binary])
    out = subprocess.check_output([os.path.join(ROOT, "tests", "cpp", "build", binary), str(first), str(last)], text=True)
    n, mism, ck_libm, ck_rest = (int(x) for x in out.split())
    return n, mism, ck_libm, ck_rest


@pytest.mark.parametrize("binary", ["sincosf_sweep", "sincosf_sweep_fma"])
def test_restatement_equals_host_libm_exhaustively(binary):
    n, mism, ck_libm, ck_rest = _sweep(binary)
    assert n == LAST + 1 and n > 1_086_000_000
    assert mism == 0 and ck_libm == ck_rest


@pytest.mark.gpu
def test_device_sincosf_equals_host_libm_exhaustively():
    import importlib
    capi = importlib.import_module("orb_slam2-chinese-annotation_b200.capi")
    L = capi.lib()
    ck = C.c_uint64(0)
    L.orb_sincosf_sweep_device.argtypes = [C.c_uint32, C.c_uint32, C.c_int, C.POINTER(C.c_uint64)]
    capi.check(L.orb_sincosf_sweep_device(0, LAST, 0, C.byref(ck)))
    n, mism, ck_libm, _ = _sweep("sincosf_sweep")
    assert mism == 0
    assert ck.value == ck_libm
    # negative zero and a sub-range, as a check of the checksum plumbing itself
    capi.check(L.orb_sincosf_sweep_device(0x3f000000, 0x3f00ffff, 0, C.byref(ck)))
    assert ck.value == _sweep("sincosf_sweep", 0x3f000000, 0x3f00ffff)[2]
