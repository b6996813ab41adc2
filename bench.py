#!/usr/bin/env python
"""Benchmark of the ORB front-end hot path (BASELINE.json: "ORB extract+match frames/s @1241x376 2000kp").
One step = one batch of KITTI-shape stereo frames through: extract(left), extract(right), ComputeStereoMatches.
A frame is one stereo pair.

  python bench.py --gpus N --steps K --warmup W              our CUDA path (N>1: under torchrun)
  python bench.py --impl reference ...                        the reference's CPU path on host cores

Prints ONE JSON line.  `value` is measured with inputs resident in HBM, `e2e` through the C ABI with host buffers
(pinned H2D and D2H inside the timed region).  Beside the headline the same line carries (untimed for `value`):
  roofline      every kernel against ITS measured roof: HBM copy bandwidth (MEASURED_PEAKS.json) for the pyramid and the
                blur, the measured ALU-pipe issue rate (profiles/MEASURED_INT_PEAKS.json, tools/peaks_int.cu) for FAST,
                quadtree, rBRIEF and stereo; numerators of the INT kernels are instruction counts read by ncu at this
                batch size (profiles/r2_ncu_counts.json, tools/ncu_counts.py)
  sustained     the same step repeated for >= 2 s
  parity_checked  pairs of the timed batch compared bit for bit with the CPU oracle after the timing
  latency       BASELINE configs[1] at batch 1 and configs[0] (one 640x480 frame) through the C++ drop-in classes,
                timed per frame like Examples/Stereo/stereo_kitti.cc:80-122, beside the reference's own CPU latency
  configs       configs[2] (EuRoC 752x480 / 1200, 256 pairs) and configs[4] (3840x2160 / 8000, 128 frames per GPU)
  matching      configs[3] (2000 x 2000 descriptors per pair, 1024 pairs)
"""
import argparse
import ctypes as C
import importlib
import json
import math
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
PKG = "orb_slam2-chinese-annotation_b200"

N_LEVELS, SCALE, INI_TH, MIN_TH = 8, 1.2, 20, 7
METRIC = "ORB extract+match frames/s @1241x376 2000kp"
CFG = {
    # Examples/Stereo/KITTI00-02.yaml:8,25,38-51
    "kitti": dict(w=1241, h=376, nfeat=2000, mbf=386.1448, fx=718.856, stereo=True, pool=16,
                  workload="kitti_stereo_1241x376_2000kp (BASELINE.json configs[1])"),
    # Examples/Stereo/EuRoC.yaml:18-19,25,88-101
    "euroc": dict(w=752, h=480, nfeat=1200, mbf=47.906, fx=435.205, stereo=True, pool=16,
                  workload="euroc_stereo_752x480_1200kp, 256 pairs per GPU (BASELINE.json configs[2])"),
    "uhd": dict(w=3840, h=2160, nfeat=8000, mbf=0.0, fx=2000.0, stereo=False, pool=2,
                workload="uhd_mono_3840x2160_8000kp, 128 frames per GPU = 1024 over 8 GPUs (BASELINE.json configs[4])"),
    # Examples/Monocular/TUM1.yaml:30-43
    "tum": dict(w=640, h=480, nfeat=1000, mbf=40.0, fx=517.3, stereo=False, pool=8,
                workload="tum_mono_640x480_1000kp, one frame (BASELINE.json configs[0])"),
}
CONFIG_KEYS = {"workload": CFG["kitti"]["workload"], "frame": "stereo pair (2 images)", "image": "1241x376 8u", "nfeatures": 2000,
               "levels": N_LEVELS, "scale_factor": SCALE, "fast_thresholds": [INI_TH, MIN_TH]}
CPU_DESC = {
    "reference": "extraction = the reference's own src/ORBextractor.cc compiled verbatim (oracle/_ref), its OpenCV primitives being SCALAR "
                 "restatements pinned to cv2 4.13 (a SIMD OpenCV build is faster: see opencv_simd_adjusted, measured beside it), stereo = line-by-line "
                 "port of Frame::ComputeStereoMatches",
    "port": "CPU oracle: restatement of the reference ORBextractor + ComputeStereoMatches (oracle/_ref not built on this box)",
}


def level_sizes(w, h):
    sf = [np.float32(1.0)]
    for _ in range(1, N_LEVELS):
        sf.append(np.float32(np.float64(sf[-1]) * np.float64(np.float32(SCALE))))
    out = []
    for s in sf:
        inv = np.float32(1.0) / s
        out.append((int(np.rint(np.float32(w) * inv)), int(np.rint(np.float32(h) * inv))))
    return out


def algorithmic_bytes(cfg, n_kp):
    """SURVEY.md 8(d), per image: every level read once + levels 1..7 written + blurred levels written + per-keypoint patches;
    and the same split per kernel."""
    sz = level_sizes(cfg["w"], cfg["h"])
    s = sum(a * b for a, b in sz)
    s1 = s - cfg["w"] * cfg["h"]
    per_kernel = {
        "pyramid": (s - sz[-1][0] * sz[-1][1]) + s1,      # read levels 0..6 once, write levels 1..7
        "fast_cells": s + 4 * 8 * n_kp,                    # read every level once, write candidate keys (about 8 per kept keypoint)
        "quadtree": 3 * 4 * 8 * n_kp,                      # candidate keys: gather + leaf ids (lower bound)
        "blur": 2 * s,                                     # read + write every level
        "orient_describe": n_kp * (749 + 961 + 32 + 28),
    }
    return s + s1 + s + n_kp * (749 + 961 + 32 + 28), per_kernel, s


def synth_pool(cfg):
    synth = importlib.import_module(PKG + ".synth")
    pairs = [synth.make_stereo_pair(cfg["w"], cfg["h"], i, fx=cfg["fx"]) for i in range(cfg["pool"])]
    return np.stack([p[0] for p in pairs]), np.stack([p[1] for p in pairs])


def tile_pool(pool, B, rank):
    reps = (B + len(pool) - 1) // len(pool)
    return np.roll(np.tile(pool, (reps, 1, 1))[:B], rank, axis=0)   # every rank works on its own frames (weak scaling)


class ClockSampler:
    """Samples SM clocks / throttle reasons DURING the timed region: NVML (nvidia_ml_py) polled every few ms from a
    thread, so that even a 70 ms region gets samples on an 8-GPU box; `nvidia-smi -lms` as the fallback."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.sm, self.mx, self.reasons = [], [], [], set()
        self.p = None; self.nv = None; self.run = True
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[index]) if vis and vis.split(",")[index].isdigit() else index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.nv = pynvml
            self.mx.append(float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)))
            self.t = threading.Thread(target=self._poll, daemon=True)
            self.t.start()
            return
        except Exception:
            self.nv = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                       "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.p = None

    def _poll(self):
        nv = self.nv
        get_reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
        bits = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}   # nvml.h nvmlClocksEventReason*
        while self.run:
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                r = int(get_reasons(self.h))
                for k, b in bits.items():
                    if r & b:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.004)

    def _read(self):
        for line in self.p.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.nv is not None:
            self.run = False
            self.t.join(timeout=1)
            return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": max(self.mx) if self.mx else None,
                    "reasons": sorted(self.reasons), "samples": len(self.sm), "source": "nvml"}
        if not self.p:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.p.terminate()
        try:
            self.p.wait(timeout=2)
        except Exception:
            self.p.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for nme, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nme)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "source": "nvidia-smi"}


# ------------------------------------------------------------------------------------------ CPU (oracle = checker / baseline only)
def _orc():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import orc  # TEST INFRASTRUCTURE: used here only as the timed CPU baseline and the post-timing checker
    return orc


def cpu_units_per_s(cfg, n_units, threads, left, right):
    """Times the reference's CPU path (oracle/_ref when built, else the oracle port) on n_units frames (stereo pairs, or
    single images for a mono config) using `threads` host threads (ctypes releases the GIL)."""
    orc = _orc()
    use_ref = orc.ref_available()   # the reference's own ORBextractor.cc compiled verbatim (oracle/_ref)
    nf = cfg["nfeat"]
    tabs = orc.Extractor(nf, SCALE, N_LEVELS, INI_TH, MIN_TH).tables()
    pool = len(left)

    def work(tid, out, lat):
        mk = orc.RefExtractor if use_ref else orc.Extractor
        ol = mk(nf, SCALE, N_LEVELS, INI_TH, MIN_TH); orr = mk(nf, SCALE, N_LEVELS, INI_TH, MIN_TH)
        n = 0
        for i in range(tid, n_units, threads):
            t0 = time.perf_counter()
            kl, dl = ol.extract(left[i % pool])
            if cfg["stereo"]:
                kr, dr = orr.extract(right[i % pool])
                if use_ref:
                    orc.stereo_match_ref(ol, orr, kl, dl, kr, dr, tabs["sf"], tabs["inv_sf"], cfg["mbf"], 0.0, cfg["fx"])
                else:
                    orc.stereo_match(ol, orr, kl, dl, kr, dr, cfg["mbf"], 0.0, cfg["fx"])
            lat.append(time.perf_counter() - t0)
            n += 1
        out[tid] = n

    out = [0] * threads
    lats = [[] for _ in range(threads)]
    ts = [threading.Thread(target=work, args=(t, out, lats[t])) for t in range(threads)]
    t0 = time.perf_counter()
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    dt = time.perf_counter() - t0
    return sum(out) / dt, dt, ("reference" if use_ref else "port"), [x for l in lats for x in l]


def opencv_adjusted(cfg, frames_per_s, cores, left):
    """The CPU baseline rescaled as if cv::resize / GaussianBlur / FAST ran at real OpenCV's (SIMD) single-thread speed instead of the
    scalar stand-ins of oracle/cvshim: both are timed here on one thread over the levels of a few images (tools/cv2_baseline.py).
    None where cv2 is not importable."""
    try:
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import cv2_baseline
        prim = cv2_baseline.primitive_times(left[:3], cfg["nfeat"], SCALE, N_LEVELS, INI_TH, repeats=2)
        if prim is None:
            return None
        adj, t_img, t_adj = cv2_baseline.adjust(frames_per_s, cores, 2 if cfg["stereo"] else 1, prim)
        return {"value": adj, "unit": "frames/s", "ms_per_image_per_core": t_img, "adjusted_ms_per_image_per_core": t_adj,
                "scalar_primitives_ms_per_image": prim["scalar_ms"], "cv2_primitives_ms_per_image": prim["cv2_ms"], "cv2": prim["cv2_version"],
                "how": "value x t / (t - scalar + cv2): resize + GaussianBlur + FAST over the 8 levels, scalar stand-ins vs cv2, 1 thread each"}
    except Exception as e:   # noqa: a measurement aid must not take the bench line down
        return {"error": repr(e)}


def cpu_stereo_latency(cfg, n_pairs, left, right):
    """The reference's own per-frame shape: left and right extraction on two threads (src/Frame.cc:81-84), then
    ComputeStereoMatches; steady clock per frame, median / mean like Examples/Stereo/stereo_kitti.cc:113-122."""
    orc = _orc()
    use_ref = orc.ref_available()
    nf = cfg["nfeat"]
    mk = orc.RefExtractor if use_ref else orc.Extractor
    ol = mk(nf, SCALE, N_LEVELS, INI_TH, MIN_TH); orr = mk(nf, SCALE, N_LEVELS, INI_TH, MIN_TH)
    tabs = orc.Extractor(nf, SCALE, N_LEVELS, INI_TH, MIN_TH).tables()
    ms = []
    for i in range(n_pairs + 2):
        res = {}
        t0 = time.perf_counter()
        if cfg["stereo"]:
            tl = threading.Thread(target=lambda: res.__setitem__("l", ol.extract(left[i % len(left)])))
            tr = threading.Thread(target=lambda: res.__setitem__("r", orr.extract(right[i % len(left)])))
            tl.start(); tr.start(); tl.join(); tr.join()
            kl, dl = res["l"]; kr, dr = res["r"]
            if use_ref:
                orc.stereo_match_ref(ol, orr, kl, dl, kr, dr, tabs["sf"], tabs["inv_sf"], cfg["mbf"], 0.0, cfg["fx"])
            else:
                orc.stereo_match(ol, orr, kl, dl, kr, dr, cfg["mbf"], 0.0, cfg["fx"])
        else:
            ol.extract(left[i % len(left)])
        if i >= 2:
            ms.append((time.perf_counter() - t0) * 1e3)
    return {"median_ms": float(np.median(ms)), "mean_ms": float(np.mean(ms)), "frames": n_pairs, "threads": 2 if cfg["stereo"] else 1,
            "kind": "reference" if use_ref else "port"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cfg = CFG["kitti"]
    left, right = synth_pool(cfg)
    cores = os.cpu_count() or 1
    per_step = 12 * cores   # ~0.1 core-seconds per pair: a step is about 1.5 s of wall clock on all cores
    cpu_units_per_s(cfg, min(per_step, cores), cores, left, right)  # page in
    for _ in range(args.warmup):
        cpu_units_per_s(cfg, per_step, cores, left, right)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        kind = cpu_units_per_s(cfg, per_step, cores, left, right)[2]
    dt = time.perf_counter() - t0
    val = per_step * args.steps / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": CONFIG_KEYS, "pairs_per_step": per_step,
        "cpu_baseline": {"value": val, "unit": "frames/s", "cores": cores, "kind": kind,
                         "sample": "%d stereo pairs per step x %d steps, %d host threads; %s"
                                   % (per_step, args.steps, cores, CPU_DESC[kind]),
                         "opencv_simd_adjusted": opencv_adjusted(cfg, val, cores, left)},
        "e2e": {"value": val, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------ GPU
class Job:
    """One workload (a CFG entry) on this rank's GPU: handles, device-resident inputs, timing helpers."""

    def __init__(self, cfg, B, local, rank, dist):
        import torch
        self.torch = torch
        ext = importlib.import_module(PKG + ".extractor")
        self.stereo_mod = importlib.import_module(PKG + ".stereo")
        self.capi = importlib.import_module(PKG + ".capi")
        self.cfg, self.B, self.local, self.rank, self.dist = cfg, B, local, rank, dist
        self.left_pool, self.right_pool = synth_pool(cfg)
        self.hl = tile_pool(self.left_pool, B, rank)
        self.hr = tile_pool(self.right_pool, B, rank) if cfg["stereo"] else None
        mk = lambda: ext.ORBextractor(cfg["nfeat"], SCALE, N_LEVELS, INI_TH, MIN_TH, max_width=cfg["w"], max_height=cfg["h"],
                                      max_batch=B, device=local)
        self.gl = mk()
        self.gr = mk() if cfg["stereo"] else None
        self.cap = self.gl.kp_cap
        self.dl = torch.from_numpy(self.hl).cuda()
        self.dr = torch.from_numpy(self.hr).cuda() if cfg["stereo"] else None
        torch.cuda.synchronize()
        self.sl = torch.cuda.ExternalStream(self.gl.stream, device=local)

    def close(self):
        self.gl.close()
        if self.gr:
            self.gr.close()
        self.dl = self.dr = None
        self.torch.cuda.empty_cache()

    def step_device(self):
        c = self.cfg
        self.gl.extract_device(self.dl.data_ptr(), self.B, c["w"], c["h"])
        if c["stereo"]:
            self.gr.extract_device(self.dr.data_ptr(), self.B, c["w"], c["h"])
            self.stereo_mod.compute_stereo_matches_device(self.gl, self.gr, c["mbf"], c["fx"], 0.0, batch=self.B)

    def sync_all(self):
        self.gl.sync()
        if self.gr:
            self.gr.sync()
        self.torch.cuda.synchronize()
        if self.dist is not None:
            self.dist.barrier()

    def launches(self):
        return self.gl.launch_count + (self.gr.launch_count if self.gr else 0)

    def max_over_ranks(self, ms):
        t = self.torch.tensor([ms], dtype=self.torch.float64, device="cuda")
        if self.dist is not None:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def time_device(self, steps, warmup):
        """CUDA events on the left handle's stream (the stereo stage makes the left stream wait for the right one, and the next
        step's right extraction waits for the left stream: the left stream covers the whole step chain); max over ranks."""
        torch = self.torch
        for _ in range(warmup):
            self.step_device()
        self.sync_all()
        l0 = self.launches()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(self.sl)
        for _ in range(steps):
            self.step_device()
        e1.record(self.sl)
        self.sync_all()
        return self.max_over_ranks(e0.elapsed_time(e1)), self.launches() - l0

    def stage_times(self, n_prof=3):
        """Per-stage event times, one handle at a time so that the other stream's kernels do not share the SMs."""
        c = self.cfg
        self.gl.set_timing(True)
        if self.gr:
            self.gr.set_timing(True)
        acc = {}
        for _ in range(n_prof):
            self.gl.extract_device(self.dl.data_ptr(), self.B, c["w"], c["h"]); self.gl.sync()
            sr = {}
            if self.gr:
                self.gr.extract_device(self.dr.data_ptr(), self.B, c["w"], c["h"]); self.gr.sync()
                self.stereo_mod.compute_stereo_matches_device(self.gl, self.gr, c["mbf"], c["fx"], 0.0, batch=self.B); self.gl.sync()
                sr = self.gr.stage_ms()
            sl = self.gl.stage_ms()
            for k in sl:
                acc[k] = acc.get(k, 0.0) + sl[k] + (sr.get(k, 0.0) if not k.startswith("stereo") else 0.0)
        self.gl.set_timing(False)
        if self.gr:
            self.gr.set_timing(False)
        return {k: v / n_prof for k, v in acc.items() if c["stereo"] or not k.startswith("stereo")}

    def e2e(self, steps, copy_only=False):
        """The call a user of the reference makes, batched, HOST buffers in and out: Frame::Frame(left, right) = ExtractORB x2 +
        ComputeStereoMatches (src/Frame.cc:60-115) -> orb_stereo_frames_enqueue / _wait, two batches in flight on alternating pinned
        buffer sets (the next batch is queued before the previous one is waited for, as a streaming caller does); a mono config ->
        orb_extract_enqueue / _wait, the same for one handle.  Every step uploads its images from pinned host memory and downloads all of its results."""
        torch, capi, c, B, cap = self.torch, self.capi, self.cfg, self.B, self.cap
        L = capi.lib()
        kp_sz = capi.KP_DTYPE.itemsize
        pl = torch.from_numpy(self.hl).pin_memory()
        sides = 2 if c["stereo"] else 1
        capi.check(L.orb_extractor_set_copy_only(self.gl.handle, int(copy_only)))
        nset = 2
        sets = []
        for _ in range(nset):
            o_k = [torch.empty((B, cap, kp_sz), dtype=torch.uint8).pin_memory() for _ in range(sides)]
            o_d = [torch.empty((B, cap, 32), dtype=torch.uint8).pin_memory() for _ in range(sides)]
            o_n = [np.zeros(B, np.int32) for _ in range(2)]
            o_s = (torch.empty((B, cap), dtype=torch.float32).pin_memory(), torch.empty((B, cap), dtype=torch.float32).pin_memory()) if c["stereo"] else None
            sets.append((o_k, o_d, o_n, o_s))
        if c["stereo"]:
            pr = torch.from_numpy(self.hr).pin_memory()
            tick = C.c_int(0)

            def enqueue(i):
                o_k, o_d, o_n, o_s = sets[i & 1]
                capi.check(L.orb_stereo_frames_enqueue(self.gl.handle, self.gr.handle, C.c_void_p(pl.data_ptr()), C.c_void_p(pr.data_ptr()), B, c["w"], c["h"],
                                                       c["w"], C.c_size_t(c["w"] * c["h"]), C.c_float(c["mbf"]), C.c_float(0.0), C.c_float(c["fx"]),
                                                       C.c_void_p(o_k[0].data_ptr()), C.c_void_p(o_d[0].data_ptr()), C.c_void_p(o_k[1].data_ptr()),
                                                       C.c_void_p(o_d[1].data_ptr()), C.c_void_p(o_s[0].data_ptr()), C.c_void_p(o_s[1].data_ptr()), cap, C.byref(tick)))
                return tick.value

            def wait(i, t):
                o_n = sets[i & 1][2]
                capi.check(L.orb_stereo_frames_wait(self.gl.handle, self.gr.handle, t, capi.ptr(o_n[0]), capi.ptr(o_n[1])))

            def run(n):
                t_prev = enqueue(0)
                for i in range(1, n):
                    t = enqueue(i)
                    wait(i - 1, t_prev)
                    t_prev = t
                wait(n - 1, t_prev)
        else:
            tick = C.c_int(0)

            def enqueue(i):   # orb_extract_enqueue / _wait: the streaming form of orb_extract_batch, two batches in flight
                o_k, o_d, o_n, _ = sets[i & 1]
                capi.check(L.orb_extract_enqueue(self.gl.handle, C.c_void_p(pl.data_ptr()), B, c["w"], c["h"], c["w"], C.c_size_t(c["w"] * c["h"]),
                                                 C.c_void_p(o_k[0].data_ptr()), C.c_void_p(o_d[0].data_ptr()), cap, C.byref(tick)))
                return tick.value

            def run(n):
                t_prev = enqueue(0)
                for i in range(1, n):
                    t = enqueue(i)
                    capi.check(L.orb_extract_wait(self.gl.handle, t_prev, capi.ptr(sets[(i - 1) & 1][2][0])))
                    t_prev = t
                capi.check(L.orb_extract_wait(self.gl.handle, t_prev, capi.ptr(sets[(n - 1) & 1][2][0])))
        run(3)
        self.sync_all()
        t0 = time.perf_counter()
        run(steps)
        self.sync_all()
        ms = self.max_over_ranks((time.perf_counter() - t0) * 1e3)
        capi.check(L.orb_extractor_set_copy_only(self.gl.handle, 0))
        h2d = sides * B * c["w"] * c["h"]
        # slot rows are downloaded whole (cap == the handle's kp_cap): keypoints + descriptors per side, uRight, depth, counts
        d2h = int(B * cap * (sides * (kp_sz + 32) + (8 if c["stereo"] else 0)) + sides * B * 4 + 8)
        self.host_out = sets[(steps - 1) & 1]
        return ms, h2d, d2h


def check_parity(job, picks):
    """Bit-equality of `picks` frames of the batch just timed (outputs of the last host-buffer step) against the CPU oracle."""
    orc = _orc()
    c = job.cfg
    o_k, o_d, o_n, st = job.host_out
    kp_dt = job.capi.KP_DTYPE
    ol = orc.Extractor(c["nfeat"], SCALE, N_LEVELS, INI_TH, MIN_TH); orr = orc.Extractor(c["nfeat"], SCALE, N_LEVELS, INI_TH, MIN_TH)
    ok = 0
    for b in picks:
        okl, odl = ol.extract(job.hl[b])
        n = int(o_n[0][b])
        gk = o_k[0].numpy()[b].view(kp_dt).reshape(-1)[:n]
        good = n == len(okl) and all(np.array_equal(gk[f], okl[f]) for f in ("x", "y", "angle", "response", "octave", "size"))
        good = good and np.array_equal(o_d[0].numpy()[b][:n], odl)
        if c["stereo"]:
            okr, odr = orr.extract(job.hr[b])
            m = int(o_n[1][b])
            good = good and m == len(okr) and np.array_equal(o_d[1].numpy()[b][:m], odr)
            our, odp, _, _ = orc.stereo_match(ol, orr, okl, odl, okr, odr, c["mbf"], 0.0, c["fx"])
            good = good and np.array_equal(st[0].numpy()[b][:n], our) and np.array_equal(st[1].numpy()[b][:n], odp)
        if not good:
            raise SystemExit("bench.py: parity check FAILED on frame %d of the timed batch" % b)
        ok += 1
    return ok


def load_json(path, default):
    try:
        return json.load(open(os.path.join(ROOT, path)))
    except Exception:
        return default


def kernel_counts(counts, names):
    """ncu counts of one stage: one kernel, or the sum over the kernels the stage launches."""
    ks = counts.get("kernels", {})
    if isinstance(names, str):
        return ks.get(names)
    got = [ks[n] for n in names if n in ks]
    if len(got) != len(names):
        return None
    return {f: sum(g[f] for g in got) for f in ("alu_lane_ops_per_unit", "thread_inst_per_unit", "dram_bytes_per_unit")}


def roofline_block(cfg, B, stages, n_kp, clocks):
    """Every stage against its own measured roof.  HBM: algorithmic bytes / event time / measured copy bandwidth.  INT: ALU-pipe
    lane-operations counted by ncu at this batch size / event time / (measured ALU-pipe lanes per clock per SM x SMs x clock)."""
    peaks = load_json("MEASURED_PEAKS.json", {})
    ipk = load_json("profiles/MEASURED_INT_PEAKS.json", {})
    counts = load_json("profiles/r2_ncu_counts.json", {})
    hbm = float(peaks.get("hbm_gbs", 6650.0))
    hbm_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
    sms = int(ipk.get("sms", 148))
    lanes = float(ipk.get("tests", {}).get("lop3", {}).get("lanes_per_clk_per_sm", 64.0))
    sm_mhz = clocks.get("sm_mhz") or 1965.0
    alu_peak = lanes * sms * sm_mhz * 1e6 / 1e9      # G lane-ops/s
    issue_peak = 128.0 * sms * sm_mhz * 1e6 / 1e9    # 4 schedulers x 32 lanes per clock per SM
    _, bytes_k, _ = algorithmic_bytes(cfg, n_kp)
    kernel_of = {"pyramid": "k_resize_tma", "fast_cells": "k_fast_cells", "quadtree": "k_quadtree", "blur": "k_gauss7_tma",
                 "orient_describe": "k_orient_describe", "stereo_match": ("k_stereo_search", "k_stereo_refine"), "stereo_rows": "k_stereo_rows",
                 "stereo_median": "k_stereo_median"}
    bound_of = {"pyramid": "hbm", "blur": "hbm"}
    per = {}
    sides = 2 if cfg["stereo"] else 1
    for k, ms in stages.items():
        if ms <= 0:
            continue
        images = B * (1 if k.startswith("stereo") else sides)     # image (or pair) units behind this stage time
        ent = {"ms_per_step": ms, "bound": bound_of.get(k, "int")}
        if k in bytes_k:
            a = bytes_k[k] * images / (ms * 1e-3) / 1e9
            ent["hbm"] = {"alg_bytes_per_image": int(bytes_k[k]), "achieved_GBs": a, "frac": a / hbm}
        cn = kernel_counts(counts, kernel_of.get(k, ""))
        if cn and counts.get("workload") == cfg["workload"]:
            a = cn["alu_lane_ops_per_unit"] * images / (ms * 1e-3) / 1e9
            t = cn["thread_inst_per_unit"] * images / (ms * 1e-3) / 1e9
            ent["int"] = {"alu_lane_ops_per_unit": cn["alu_lane_ops_per_unit"], "achieved_Gops": a, "frac": a / alu_peak,
                          "issue_frac": t / issue_peak, "dram_bytes_per_unit_ncu": cn.get("dram_bytes_per_unit")}
        per[k] = ent
    dom = max(per, key=lambda k: per[k]["ms_per_step"])
    d = per[dom]
    if d["bound"] == "hbm":
        top = {"bound": "hbm", "kernel": dom, "achieved": d["hbm"]["achieved_GBs"], "peak": hbm, "unit": "GB/s", "frac": d["hbm"]["frac"]}
    elif "int" in d:
        top = {"bound": "int", "kernel": dom, "achieved": d["int"]["achieved_Gops"], "peak": alu_peak,
               "unit": "G lane-ops/s on the ALU pipe (ncu sm__inst_executed_pipe_alu x 32)", "frac": d["int"]["frac"]}
    else:
        top = {"bound": "int", "kernel": dom, "achieved": None, "peak": alu_peak, "unit": "G lane-ops/s on the ALU pipe", "frac": None,
               "note": "profiles/r2_ncu_counts.json has no counts for this workload"}
    cn = (kernel_counts(counts, kernel_of.get(dom, "")) or {}) if counts.get("workload") == cfg["workload"] else {}
    units = B * (1 if dom.startswith("stereo") else 1)   # per launch: one handle, B images
    top["traffic"] = (cn.get("dram_bytes_per_unit") * units) if cn.get("dram_bytes_per_unit") else None
    top["traffic_source"] = counts.get("source")
    top["peak_source"] = {"hbm": hbm_src, "int": "measured: profiles/MEASURED_INT_PEAKS.json (%g lanes/clk/SM x %d SMs) x sampled SM clock %.0f MHz"
                          % (lanes, sms, sm_mhz)}
    top["per_kernel"] = per
    tot, _, _ = algorithmic_bytes(cfg, n_kp)
    top["pipeline_hbm_frac"] = sides * B * tot / (sum(stages.values()) * 1e-3) / 1e9 / hbm
    return top


def run_latency(cfg_name, local):
    """Batch 1 through the C++ drop-in classes (tools/latency_bench.cpp), beside the reference's CPU latency."""
    cfg = CFG[cfg_name]
    exe = os.path.join(ROOT, "tools", "build", "latency_bench")
    if not os.path.exists(exe):
        return {"unavailable": "tools/build/latency_bench not built"}
    left, right = synth_pool(dict(cfg, pool=8))
    path = "/tmp/orb_b200_latency_%s_%d.bin" % (cfg_name, os.getpid())
    (np.concatenate([left.ravel(), right.ravel()]) if cfg["stereo"] else left.ravel()).tofile(path)
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    env = dict(os.environ, CUDA_VISIBLE_DEVICES=vis.split(",")[local] if vis else str(local))
    out = {}
    try:
        base = [exe, "stereo" if cfg["stereo"] else "mono", path, str(cfg["w"]), str(cfg["h"]), "8", str(cfg["nfeat"]), str(cfg["mbf"]), str(cfg["fx"]),
                "200", "20"]
        out["ours"] = json.loads(subprocess.run(base + ["0", "1"], capture_output=True, text=True, env=env, timeout=120).stdout)
        out["ours_with_host_pyramid"] = json.loads(subprocess.run(base + ["1", "1"], capture_output=True, text=True, env=env, timeout=120).stdout)
        if cfg["stereo"]:   # ORBextractor::ExtractStereo: the same front end as one orb_stereo_frames call from one thread
            out["ours_single_call"] = json.loads(subprocess.run(base + ["0", "2"], capture_output=True, text=True, env=env, timeout=120).stdout)
    except Exception as e:   # noqa
        out["error"] = repr(e)
    finally:
        try:
            os.remove(path)
        except OSError:
            pass
    out["reference_cpu"] = cpu_stereo_latency(cfg, 12, left, right)
    out["workload"] = cfg["workload"]
    out["how"] = ("per frame, steady clock, 200 frames after 20 warm-ups (Examples/Stereo/stereo_kitti.cc:80-122); ours = ORB_SLAM2::ORBextractor::operator() "
                  "on two threads + orb_stereo_match, host cv::Mat in, std::vector<cv::KeyPoint> / cv::Mat out; ours_single_call = "
                  "ORB_SLAM2::ORBextractor::ExtractStereo (both extractions and the stereo match queued by one thread, no thread per frame); "
                  "reference = two extraction threads + ComputeStereoMatches on the host CPU")
    return out


def run_extra(name, B, steps, local, rank, world, dist, cores):
    """A further BASELINE.json config: device-resident and host-buffer throughput, HBM fraction, CPU sample."""
    cfg = CFG[name]
    job = Job(cfg, B, local, rank, dist)
    ms, _ = job.time_device(steps, 3)
    val = world * B * steps / (ms * 1e-3)
    stages = job.stage_times(2)
    _, _, counts = job.gl.download(min(B, 8))
    n_kp = float(np.mean(counts))
    e_steps = max(6, steps)   # enough host-buffer steps for two batches in flight to reach their steady state
    e_ms, h2d, d2h = job.e2e(e_steps)
    e_val = world * B * e_steps / (e_ms * 1e-3)
    out = None
    if rank == 0:
        parity = check_parity(job, [0, B - 1] if cfg["w"] < 2000 else [0])
        peaks = load_json("MEASURED_PEAKS.json", {})
        hbm = float(peaks.get("hbm_gbs", 6650.0))
        tot, bytes_k, _ = algorithmic_bytes(cfg, n_kp)
        sides = 2 if cfg["stereo"] else 1
        per = {k: {"ms_per_step": v, **({"hbm_frac": bytes_k[k] * sides * B / (v * 1e-3) / 1e9 / hbm} if k in bytes_k and v > 0 else {})}
               for k, v in stages.items()}
        n_cpu = cores * (2 if cfg["w"] < 2000 else 1)
        cpu_val, cpu_dt, cpu_kind, _ = cpu_units_per_s(cfg, n_cpu, cores, job.left_pool, job.right_pool)
        out = {"workload": cfg["workload"], "unit": "frames/s", "frame": "stereo pair" if cfg["stereo"] else "image",
               "frames_per_step_per_gpu": B, "value": val, "ms_per_step": ms / steps, "steps": steps,
               "e2e": {"value": e_val, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "steps": e_steps},
               "avg_keypoints": n_kp, "stage_ms_per_step": per,
               "pipeline_hbm_frac": sides * B * tot / (sum(stages.values()) * 1e-3) / 1e9 / hbm,
               "alg_bytes_per_image": int(tot), "parity_checked": parity,
               "cpu_baseline": {"value": cpu_val, "unit": "frames/s", "cores": cores, "kind": cpu_kind,
                                "sample": "%d frames on %d host threads in %.1f s" % (n_cpu, cores, cpu_dt)}}
    job.close()
    return out


def run_ours(args):
    import torch
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        # NCCL prints its version banner on stdout when the communicator is created; the driver wants ONE JSON line there
        sys.stdout.flush()
        saved, devnull = os.dup(1), os.open(os.devnull, os.O_WRONLY)
        os.dup2(devnull, 1)
        try:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            os.dup2(saved, 1); os.close(saved); os.close(devnull)

    cfg = CFG["kitti"]
    B = args.batch
    warmup = max(args.warmup, 3)
    cores = os.cpu_count() or 1
    job = Job(cfg, B, local, rank, dist)
    if args.copy_only:   # the host-buffer pipeline with the kernels skipped: the copy ceiling of e2e at this GPU count
        e_steps = max(6, args.steps)
        e_ms, h2d, d2h = job.e2e(e_steps, copy_only=True)
        if rank == 0:
            print(json.dumps({"copy_only": True, "n_gpus": world, "unit": "frames/s", "value": world * B * e_steps / (e_ms * 1e-3), "steps": e_steps,
                              "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "pairs_per_step_per_gpu": B,
                              "per_gpu_GBs": (h2d + d2h) * e_steps / (e_ms * 1e-3) / 1e9, "workload": cfg["workload"]}), flush=True)
        if dist is not None:
            dist.barrier(); dist.destroy_process_group()
        return
    if args.profile_run:   # exactly two device-resident steps and nothing else (tools/ncu_counts.py captures these launches)
        job.step_device(); job.step_device(); job.sync_all()
        print(json.dumps({"profile_run": True, "steps": 2, "pairs_per_step": B, "workload": cfg["workload"]}), flush=True)
        return

    # ---- headline: EXACTLY args.steps steps, device-resident inputs, CUDA events, max over ranks
    job.time_device(0, warmup)
    sampler = ClockSampler(local)
    ms_total, launches = job.time_device(args.steps, 0)
    clocks = sampler.stop()
    value = world * B * args.steps / (ms_total * 1e-3)

    # ---- the same step for >= 2 s (clocks settle, more clock samples)
    s_steps = max(args.steps, int(math.ceil(2000.0 / (ms_total / args.steps))))
    sampler = ClockSampler(local)
    s_ms, _ = job.time_device(s_steps, 0)
    s_clocks = sampler.stop()
    sustained = {"value": world * B * s_steps / (s_ms * 1e-3), "unit": "frames/s", "steps": s_steps, "seconds": s_ms * 1e-3, "clocks": s_clocks}

    stages = job.stage_times(3)
    _, _, counts = job.gl.download(B)
    n_kp = float(np.mean(counts))

    # ---- e2e through the C ABI with pinned host buffers
    e_steps = max(3, args.steps // 2, int(math.ceil(1000.0 / (ms_total / args.steps))))   # >= 1 s of host-buffer steps: the link settles
    e_ms, h2d, d2h = job.e2e(e_steps)
    e_val = world * B * e_steps / (e_ms * 1e-3)
    parity = check_parity(job, [0, 1, B // 2, B - 1]) if rank == 0 else 0
    left_pool, right_pool = job.left_pool, job.right_pool
    job.close()

    # ---- the other BASELINE.json configs (every rank: they are sharded like the headline)
    extras = {}
    if not args.headline_only:
        extras["euroc_752x480_1200"] = run_extra("euroc", 256, 6, local, rank, world, dist, cores)
        extras["uhd_3840x2160_8000"] = run_extra("uhd", 128, 2, local, rank, world, dist, cores)

    if rank == 0:
        matching = None if args.headline_only else bench_matching(local)
        latency = None
        if not args.headline_only:
            latency = {"kitti_stereo_batch1": run_latency("kitti", local), "tum_mono_640x480": run_latency("tum", local)}
        cpu_n = 12 * cores   # ~0.1 core-seconds per pair: about 20 s of CPU work
        cpu_val, cpu_dt, cpu_kind, _ = cpu_units_per_s(cfg, cpu_n, cores, left_pool, right_pool)
        line = {
            "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps, "warmup": warmup,
            "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
            "data": "synthetic", "config": CONFIG_KEYS,
            "batch": {"pairs_per_step_per_gpu": B, "avg_keypoints": n_kp,
                      "l2": "inputs larger than L2: %d MB of images + %d MB of intermediates per step" % (2 * B * cfg["w"] * cfg["h"] >> 20, int(2 * B * 9.5))},
            "e2e": {"value": e_val, "unit": "frames/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "steps": e_steps},
            "gpu_launches": int(launches), "clocks": clocks, "sustained": sustained, "parity_checked": parity,
            "roofline": roofline_block(cfg, B, stages, n_kp, clocks),
            "cpu_baseline": {"value": cpu_val, "unit": "frames/s", "cores": cores, "kind": cpu_kind,
                             "sample": "%d stereo pairs on %d host threads in %.1f s; %s" % (cpu_n, cores, cpu_dt, CPU_DESC[cpu_kind]),
                             "opencv_simd_adjusted": opencv_adjusted(cfg, cpu_val, cores, left_pool)},
            "latency": latency, "configs": extras, "matching": matching,
        }
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


class _Dev:
    """A device tensor where the packed-view helpers expect a numpy array (they only read .ctypes.data)."""
    def __init__(self, t):
        import types
        self.t = t
        self.ctypes = types.SimpleNamespace(data=t.data_ptr())


def bench_matching(local):
    """BASELINE.json configs[3] beside the headline step (rank 0, outside its timed region): ORBmatcher Hamming matching on
    2000 x 2000 descriptor sets, 1024 frame pairs per call:
      bruteforce   best / second best per row, device-resident, CUDA events on the matcher's stream, against the MEASURED POPC rate
      M4           SearchByProjection(Frame, MapPoints) (src/ORBmatcher.cc:47-133; ratio test), th 1 / 3 / 5, ONE batched call
      M5           SearchByProjection(Cur, Last) (:1460-1619; rotation histogram), th 7 / 15, ONE batched call
    M4 / M5 are timed twice: pinned HOST buffers in and out (what a caller of the C ABI sees) and device-resident buffers; the CPU
    figure beside them is the oracle port of the same functions (pinned bit-exact to the reference's ORBmatcher.cc) on all host cores."""
    synth = importlib.import_module(PKG + ".synth")
    mt = importlib.import_module(PKG + ".matcher")
    views = importlib.import_module(PKG + ".views")
    capi = importlib.import_module(PKG + ".capi")
    import torch
    P, N = 1024, 2000
    rng = np.random.default_rng(7)
    U = 32   # unique pairs, tiled to P (the kernel's time does not depend on the content; 131 MB of descriptors > L2)
    a = rng.integers(0, 256, (U, N, 32), dtype=np.uint8)
    b = a[:, rng.permutation(N)] ^ np.packbits(rng.random((U, N, 256)) < 0.08, axis=2)
    a = np.tile(a, (P // U, 1, 1)); b = np.tile(b, (P // U, 1, 1))
    m = mt.ORBmatcher(0.6, True, max_n=2048, max_pairs=P, device=local)
    L = capi.lib()
    st = torch.cuda.ExternalStream(L.orb_matcher_stream(m._h), device=local)
    da = torch.from_numpy(a).cuda(); db = torch.from_numpy(b).cuda()
    bi = torch.empty((P, N), dtype=torch.int32, device="cuda"); bd = torch.empty_like(bi); sd = torch.empty_like(bi)
    torch.cuda.synchronize()

    def run():
        capi.check(L.orb_match_bruteforce_device(m._h, C.c_void_p(da.data_ptr()), C.c_void_p(db.data_ptr()), P, N, N, C.c_void_p(bi.data_ptr()),
                                                 C.c_void_p(bd.data_ptr()), C.c_void_p(sd.data_ptr())))
    for _ in range(2):
        run()
    capi.check(L.orb_matcher_sync(m._h))
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    K = 5
    e0.record(st)
    for _ in range(K):
        run()
    e1.record(st)
    capi.check(L.orb_matcher_sync(m._h))
    bf_ms = e0.elapsed_time(e1) / K
    del da, db, bi, bd, sd
    popc = P * N * N * 8 / (bf_ms * 1e-3)
    ipk = load_json("profiles/MEASURED_INT_PEAKS.json", {})
    popc_lanes = float(ipk.get("tests", {}).get("popc", {}).get("lanes_per_clk_per_sm", 16.0))
    popc_peak = popc_lanes * int(ipk.get("sms", 148)) * 1965e6
    out = {"workload": "2000x2000 descriptors per frame pair, 1024 pairs per call (BASELINE.json configs[3])",
           "bruteforce": {"ms_per_1024_pairs": bf_ms, "pairs_per_s": P / (bf_ms * 1e-3), "popc_Tops": popc / 1e12,
                          "popc_peak_Tops": popc_peak / 1e12, "frac": popc / popc_peak,
                          "peak_source": "measured: profiles/MEASURED_INT_PEAKS.json popc %.2f lanes/clk/SM x SMs x 1965 MHz" % popc_lanes}}

    # ---- grid-guided searches, 1024 pairs per call
    UQ = 16
    cores = os.cpu_count() or 1
    orc = _orc()

    def pin(x):
        t = torch.from_numpy(np.ascontiguousarray(x)).pin_memory()
        return t.numpy(), t

    def tiled(qb, fields):
        o = dict(pairs=P, stride=qb["stride"]); keep = []
        for k in ("n",) + tuple(fields):
            if qb.get(k) is None:
                o[k] = None
                continue
            o[k], t = pin(np.tile(qb[k], (P // UQ,) + (1,) * (qb[k].ndim - 1)))
            keep.append(t)
        return o, keep

    def dev(o, fields):
        d = dict(pairs=o["pairs"], stride=o["stride"])
        for k in ("n",) + tuple(fields):
            d[k] = None if o.get(k) is None else _Dev(torch.from_numpy(o[k]).cuda())
        return d

    def timed(call, reps=3):
        call(); capi.check(L.orb_matcher_sync(m._h))
        t0 = time.perf_counter()
        for _ in range(reps):
            call()
        capi.check(L.orb_matcher_sync(m._h))
        return (time.perf_counter() - t0) / reps * 1e3

    FR = ("x", "y", "octave", "angle", "u_right", "desc")
    res = {}
    # M4
    scen = [synth.make_projection_scenario(N, N, i) for i in range(UQ)]
    frames = [views.Frame(f["x"], f["y"], f["octave"], f["angle"], f["desc"], f["sf"], f["bounds"], f["u_right"]) for f, _, _ in scen]
    fb0 = mt.pack_frames(frames); qb0 = mt.pack_points([s[1] for s in scen], mt.MP_FIELDS)
    fb, k1 = tiled(fb0, FR); fb["frame0"] = frames[0]
    qb, k2 = tiled(qb0, mt.MP_FIELDS)
    occ, k3 = pin(np.tile(mt._pad_stack([mt.u8(s[2]) for s in scen], fb["stride"], np.uint8), (P // UQ, 1)))
    match, k4 = pin(np.empty((P, fb["stride"]), np.int32)); nm, k5 = pin(np.empty(P, np.int32))
    dfb = dev(fb, FR); dfb["frame0"] = frames[0]; dqb = dev(qb, mt.MP_FIELDS)
    docc = torch.from_numpy(occ).cuda(); dmatch = torch.empty((P, fb["stride"]), dtype=torch.int32, device="cuda"); dnm = torch.empty(P, dtype=torch.int32, device="cuda")
    m4 = {}
    for th in (1.0, 3.0, 5.0):
        ms_h = timed(lambda: m.search_mappoints_batch_packed(fb, qb, th, occ, match, nm))
        cands = L.orb_matcher_last_candidates(m._h)
        ms_d = timed(lambda: m.search_mappoints_batch_packed(dfb, dqb, th, docc.data_ptr(), dmatch.data_ptr(), dnm.data_ptr()))
        # post-check of a few pairs of the timed batch against the oracle
        for p in (0, 5, P - 1):
            n_o, m_o = orc.search_by_projection_mps(frames[p % UQ], scen[p % UQ][1], th, m.mfNNratio, scen[p % UQ][2])
            if int(nm[p]) != n_o or not np.array_equal(match[p, :frames[p % UQ].N], m_o):
                raise SystemExit("bench.py: batched SearchByProjection differs from the oracle (pair %d, th %g)" % (p, th))
        m4["th_%g" % th] = {"ms_per_1024_pairs_host_buffers": ms_h, "pairs_per_s_host_buffers": P / (ms_h * 1e-3),
                            "ms_per_1024_pairs_device_resident": ms_d, "pairs_per_s_device_resident": P / (ms_d * 1e-3),
                            "candidates_per_pair": cands / P, "popc_Gops_device_resident": cands * 8 / (ms_d * 1e-3) / 1e9,
                            "matches_pair0": int(nm[0]), "parity_checked": 3}
    # CPU: the same function on all cores (th 3)
    def cpu_rate(fn, n_calls):
        def work(tid, cnt):
            for i in range(tid, n_calls, cores):
                fn(i)
                cnt[tid] += 1
        cnt = [0] * cores
        ts = [threading.Thread(target=work, args=(t, cnt)) for t in range(cores)]
        t0 = time.perf_counter()
        for t in ts:
            t.start()
        for t in ts:
            t.join()
        return sum(cnt) / (time.perf_counter() - t0)
    cpu4 = cpu_rate(lambda i: orc.search_by_projection_mps(frames[i % UQ], scen[i % UQ][1], 3.0, 0.6, scen[i % UQ][2]), 16 * cores)
    res["SearchByProjection_Frame_MapPoints"] = dict(m4, cpu_baseline={"value": cpu4, "unit": "frame pairs/s", "cores": cores, "kind": "port", "th": 3.0,
                                                                       "sample": "%d calls of the oracle port on %d host threads (grid build included)" % (16 * cores, cores)})
    del dfb, dqb, docc, dmatch, dnm, k1, k2, k3
    # M5
    scen5 = [synth.make_lastframe_scenario(N, i) for i in range(UQ)]
    cam = scen5[0][2]
    curs = [views.Frame(f["x"], f["y"], f["octave"], f["angle"], f["desc"], f["sf"], f["bounds"], f["u_right"]) for f, _, _ in scen5]
    fb0 = mt.pack_frames(curs); qb0 = mt.pack_points([s[1] for s in scen5], mt.LAST_FIELDS)
    fb, k1 = tiled(fb0, FR); fb["frame0"] = curs[0]
    qb, k2 = tiled(qb0, mt.LAST_FIELDS)
    fwd, k6 = pin(np.zeros(P, np.uint8)); bwd, k7 = pin(np.zeros(P, np.uint8))
    match, k4 = pin(np.empty((P, fb["stride"]), np.int32))
    dfb = dev(fb, FR); dfb["frame0"] = curs[0]; dqb = dev(qb, mt.LAST_FIELDS)
    dfw = torch.zeros(P, dtype=torch.uint8, device="cuda"); dbw = torch.zeros(P, dtype=torch.uint8, device="cuda")
    dmatch = torch.empty((P, fb["stride"]), dtype=torch.int32, device="cuda"); dnm = torch.empty(P, dtype=torch.int32, device="cuda")
    m5 = {}
    for th in (7.0, 15.0):
        ms_h = timed(lambda: m.search_lastframe_batch_packed(fb, qb, cam, th, fwd, bwd, None, match, nm))
        cands = L.orb_matcher_last_candidates(m._h)
        ms_d = timed(lambda: m.search_lastframe_batch_packed(dfb, dqb, cam, th, dfw.data_ptr(), dbw.data_ptr(), None, dmatch.data_ptr(), dnm.data_ptr()))
        for p in (0, 7, P - 1):
            n_o, m_o = orc.search_by_projection_last(curs[p % UQ], scen5[p % UQ][1], cam, th, False, False, True)
            if int(nm[p]) != n_o or not np.array_equal(match[p, :curs[p % UQ].N], m_o):
                raise SystemExit("bench.py: batched SearchByProjection(Cur, Last) differs from the oracle (pair %d, th %g)" % (p, th))
        m5["th_%g" % th] = {"ms_per_1024_pairs_host_buffers": ms_h, "pairs_per_s_host_buffers": P / (ms_h * 1e-3),
                            "ms_per_1024_pairs_device_resident": ms_d, "pairs_per_s_device_resident": P / (ms_d * 1e-3),
                            "candidates_per_pair": cands / P, "popc_Gops_device_resident": cands * 8 / (ms_d * 1e-3) / 1e9,
                            "matches_pair0": int(nm[0]), "parity_checked": 3}
    cpu5 = cpu_rate(lambda i: orc.search_by_projection_last(curs[i % UQ], scen5[i % UQ][1], cam, 15.0, False, False, True), 16 * cores)
    res["SearchByProjection_Cur_Last"] = dict(m5, cpu_baseline={"value": cpu5, "unit": "frame pairs/s", "cores": cores, "kind": "port", "th": 15.0,
                                                                "sample": "%d calls of the oracle port on %d host threads (grid build included)" % (16 * cores, cores)})
    out["grid_guided"] = res
    # the single-frame call for comparison (one frame per call: launch- and copy-latency bound)
    f, mps, occ1 = scen[0]
    one = mt.ORBmatcher(0.6, True, max_n=8192, max_pairs=1, device=local)
    for _ in range(3):
        one.SearchByProjection(frames[0], mps, 3.0, occ1)
    t0 = time.perf_counter(); reps = 50
    for _ in range(reps):
        n_sp, _ = one.SearchByProjection(frames[0], mps, 3.0, occ1)
    sp_ms = (time.perf_counter() - t0) / reps * 1e3
    out["search_by_projection_single_call"] = {"ms_per_frame": sp_ms, "frames_per_s": 1e3 / sp_ms, "matches": int(n_sp), "th": 3.0}

    # ---- M7 SearchByBoW(KF, F) (src/ORBmatcher.cc:164-306) and M8 SearchForTriangulation (:718-900): one call per keyframe pair as
    # Tracking / LocalMapping make it (host merge-join of the feature vectors + device distances + resolve), beside the oracle port
    def single(call, reps=30):
        for _ in range(3):
            call()
        t0 = time.perf_counter()
        for _ in range(reps):
            r = call()
        return (time.perf_counter() - t0) / reps * 1e3, r
    bow = [synth.make_bow_scenario(N, N - 100, i) for i in range(4)]
    bf = [(views.Frame(a["x"], a["y"], a["octave"], a["angle"], a["desc"], a["sf"], a["bounds"], a["u_right"]),
           views.Frame(b["x"], b["y"], b["octave"], b["angle"], b["desc"], b["sf"], b["bounds"], b["u_right"]),
           views.FeatureVector(fva), views.FeatureVector(fvb), ha, hb, a) for a, fva, b, fvb, ha, hb in bow]
    one.mfNNratio = 0.7
    kf, fr, va, vb, ha, hb, _ = bf[0]
    ms7, (n7, m7) = single(lambda: one.SearchByBoW(kf, va, ha, fr, vb))
    n7o, m7o = orc.search_by_bow(kf, va, ha, fr, vb, 0.7, True)
    if n7 != n7o or not np.array_equal(m7, m7o):
        raise SystemExit("bench.py: SearchByBoW differs from the oracle")
    cpu7 = cpu_rate(lambda i: orc.search_by_bow(bf[i % 4][0], bf[i % 4][2], bf[i % 4][4], bf[i % 4][1], bf[i % 4][3], 0.7, True), 16 * cores)
    out["SearchByBoW_single_call"] = {"ms_per_call": ms7, "calls_per_s": 1e3 / ms7, "matches": int(n7), "parity_checked": 1,
                                      "cpu_baseline": {"value": cpu7, "unit": "calls/s", "cores": cores, "kind": "port",
                                                       "sample": "%d calls of the oracle port on %d host threads" % (16 * cores, cores)}}
    F12 = np.array([[0, 0, 0], [0, 0, -1], [0, 1, 0]], np.float32)   # pure x-translation: horizontal epipolar lines
    sig2 = (bf[0][6]["sf"] * bf[0][6]["sf"]).astype(np.float32)
    hb_t = [np.zeros(len(x[5]), bool) if x[5] is None else x[5] for x in bf]
    ms8, (n8, p8) = single(lambda: one.SearchForTriangulation(kf, va, ha, fr, vb, hb_t[0], F12, (2000.0, 188.0), sig2, False))
    n8o, m8o = orc.search_for_triangulation(kf, va, ha, fr, vb, hb_t[0], F12, (2000.0, 188.0), sig2, False, True)
    if n8 != n8o or p8 != [(int(i), int(m8o[i])) for i in np.nonzero(m8o >= 0)[0]]:
        raise SystemExit("bench.py: SearchForTriangulation differs from the oracle")
    cpu8 = cpu_rate(lambda i: orc.search_for_triangulation(bf[i % 4][0], bf[i % 4][2], bf[i % 4][4], bf[i % 4][1], bf[i % 4][3], hb_t[i % 4], F12,
                                                             (2000.0, 188.0), sig2, False, True), 16 * cores)
    out["SearchForTriangulation_single_call"] = {"ms_per_call": ms8, "calls_per_s": 1e3 / ms8, "matches": int(n8), "parity_checked": 1,
                                                 "cpu_baseline": {"value": cpu8, "unit": "calls/s", "cores": cores, "kind": "port",
                                                                  "sample": "%d calls of the oracle port on %d host threads" % (16 * cores, cores)}}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=256, help="stereo pairs per step per GPU")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--copy-only", action="store_true", help="e2e pipeline with the kernels skipped (copy ceiling), one JSON line")
    ap.add_argument("--profile-run", action="store_true", help="two device-resident steps only, for ncu")
    ap.add_argument("--headline-only", action="store_true", help="skip the extra configs / latency / matching sections (profiling runs)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
