"""Runs one call of every kernel family of liborb_b200.so under ORB_B200_GUARD=1 (4 KB canaries around every device buffer) and
reads the canaries back (orb_debug_guard_check).  Started by tests/test_gpu_guard.py in a fresh process, because the guard mode is
chosen at the library's first allocation.  Prints one JSON line."""
import ctypes as C
import importlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE)); sys.path.insert(0, HERE)
assert os.environ.get("ORB_B200_GUARD") == "1"
PKG = "orb_slam2-chinese-annotation_b200"
capi = importlib.import_module(PKG + ".capi")
synth = importlib.import_module(PKG + ".synth")
ext = importlib.import_module(PKG + ".extractor")
stereo = importlib.import_module(PKG + ".stereo")
mt = importlib.import_module(PKG + ".matcher")
bow = importlib.import_module(PKG + ".bow")
rect = importlib.import_module(PKG + ".rectify")


def check():
    n = C.c_int32(); bad = C.c_int32()
    st = capi.lib().orb_debug_guard_check(C.byref(n), C.byref(bad))
    return st, n.value, bad.value, capi.lib().orb_last_error().decode() if st != 0 else ""


def main():
    import __graft_entry__ as g
    g.smoke()                                               # extractor, stereo, rectify, BoW, brute-force matcher (640x480)
    report = {}
    # batched stereo front end on three geometries, odd sizes included (tile / edge paths of the TMA kernels)
    for (w, h, nf, B) in [(1241, 376, 2000, 9), (752, 480, 1200, 5), (333, 217, 500, 3), (1920, 1080, 3000, 2)]:
        gl = ext.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=B)
        gr = ext.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=B)
        pairs = [synth.make_stereo_pair(w, h, i, fx=500.0) for i in range(2)]
        il = np.ascontiguousarray(np.stack([pairs[i % 2][0] for i in range(B)])); ir = np.ascontiguousarray(np.stack([pairs[i % 2][1] for i in range(B)]))
        o = stereo.stereo_frames(gl, gr, il, ir, 40.0, 500.0)
        report["stereo_%dx%d" % (w, h)] = int(o["n_l"].sum())
        st, n, bad, msg = check()
        assert st == 0 and bad == 0, msg
    # a scale factor that takes the untiled resize path, and many levels
    g2 = ext.ORBextractor(800, 1.5, 6, 20, 7, max_width=640, max_height=480)
    kp, _ = g2(synth.make_image(640, 480, 3)); report["sf1.5"] = len(kp)
    # grid-guided matching: batched searches over ragged pairs, last-frame search with the rotation histogram
    views = importlib.import_module(PKG + ".views")
    fr = lambda f: views.Frame(f["x"], f["y"], f["octave"], f["angle"], f["desc"], f["sf"], f["bounds"], f["u_right"])
    sizes = [(2000, 2000), (700, 50), (1, 3), (1999, 1999)]
    scen = [synth.make_projection_scenario(n, mm, 10 + i) for i, (n, mm) in enumerate(sizes)]
    m = mt.ORBmatcher(0.6, True, max_n=2048, max_pairs=len(scen))
    got = m.SearchByProjectionBatch([fr(s[0]) for s in scen], [s[1] for s in scen], 3.0, [s[2] for s in scen])
    report["mappoint_matches"] = int(sum(g[0] for g in got))
    ls = [synth.make_lastframe_scenario(n, 20 + i) for i, n in enumerate([2000, 17, 900])]
    m2 = mt.ORBmatcher(0.9, True, max_n=2048, max_pairs=3)
    got = m2.SearchByProjectionLastFrameBatch([fr(s[0]) for s in ls], [s[1] for s in ls], ls[0][2], 15.0, [False, True, False], [False, False, True])
    report["lastframe_matches"] = int(sum(g[0] for g in got))
    st, n, bad, msg = check()
    assert st == 0 and bad == 0, msg
    report["buffers_checked"] = n
    # the check itself detects damage: one byte written just in front of a library buffer
    try:
        from cuda.bindings import runtime as rt
        kps = C.c_void_p(); desc = C.c_void_p(); cnt = C.c_void_p(); cap = C.c_int()
        capi.lib().orb_extractor_device_results.argtypes = [C.c_void_p] * 4 + [C.c_void_p]
        capi.check(capi.lib().orb_extractor_device_results(g2.handle, C.byref(kps), C.byref(desc), C.byref(cnt), C.byref(cap)))
        rt.cudaMemset(kps.value - 3, 0, 1)
        rt.cudaDeviceSynchronize()
        st, n, bad, msg = check()
        report["damage_detected"] = bool(st != 0 and bad == 1)
        report["damage_message"] = msg[:160]
    except ImportError:
        report["damage_detected"] = None
    print(json.dumps(report))


if __name__ == "__main__":
    main()
