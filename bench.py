#!/usr/bin/env python
"""Benchmark of the ORB front-end hot path (BASELINE.json: "ORB extract+match frames/s @1241x376
2000kp").  One step = one batch of KITTI-shape stereo frames through: extract(left), extract(right),
ComputeStereoMatches.  A frame is one stereo pair.

  python bench.py --gpus N --steps K --warmup W              our CUDA path (N>1: under torchrun)
  python bench.py --impl reference ...                        the reference's CPU path on host cores

Prints ONE JSON line (see the keys below).  `value` is measured with inputs resident in HBM,
`e2e` through the C ABI with host buffers (pinned H2D and D2H inside the timed region).
"""
import argparse
import ctypes as C
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
PKG = "orb_slam2-chinese-annotation_b200"

W, H, NFEAT = 1241, 376, 2000            # Examples/Stereo/KITTI00-02.yaml
MBF, FX = 386.1448, 718.856
N_LEVELS, SCALE, INI_TH, MIN_TH = 8, 1.2, 20, 7
WORKLOAD = "kitti_stereo_1241x376_2000kp (BASELINE.json configs[1])"
METRIC = "ORB extract+match frames/s @1241x376 2000kp"
POOL = 16                                # distinct synthetic stereo pairs, tiled to the batch
CPU_DESC = {
    "reference": "extraction = the reference's own src/ORBextractor.cc compiled verbatim (oracle/_ref, OpenCV primitives = scalar "
                 "restatements pinned to cv2 4.13), stereo = line-by-line port of Frame::ComputeStereoMatches",
    "port": "CPU oracle: restatement of the reference ORBextractor + ComputeStereoMatches (oracle/_ref not built on this box)",
}


# dram__bytes_read.sum + dram__bytes_write.sum per image of the ncu --set full capture named below (batch 32)
NCU_TRAFFIC_SOURCE = "profiles/r1_v6_full_fast_cells.txt"
NCU_TRAFFIC_BYTES_PER_IMAGE = {"fast_cells": 1328000}   # (42.34 MB read + 0.13 MB written) / 32 images; algorithmic: 1.51 MB


def algorithmic_bytes_per_image(n_kp):
    """sum(levels) read + levels 1..7 written + blurred levels written + per-keypoint patches."""
    sizes = [(1241, 376), (1034, 313), (862, 261), (718, 218), (598, 181), (499, 151), (416, 126), (346, 105)]
    s = sum(w * h for w, h in sizes)
    return s + (s - W * H) + s + n_kp * (749 + 961 + 32 + 28)


def synth_pool():
    synth = importlib.import_module(PKG + ".synth")
    pairs = [synth.make_stereo_pair(W, H, i, fx=FX) for i in range(POOL)]
    return np.stack([p[0] for p in pairs]), np.stack([p[1] for p in pairs])


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


# ------------------------------------------------------------------------------------------ CPU
def cpu_pairs_per_s(n_pairs, threads, left, right):
    """Times the CPU oracle (restatement of the reference's ORBextractor + ComputeStereoMatches) on
    n_pairs stereo frames using `threads` host threads (ctypes releases the GIL)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import orc  # TEST INFRASTRUCTURE used here only as the timed CPU baseline

    use_ref = orc.ref_available()   # the reference's own ORBextractor.cc compiled verbatim (oracle/_ref)
    tabs = orc.Extractor(NFEAT, SCALE, N_LEVELS, INI_TH, MIN_TH).tables()

    def work(tid, out):
        if use_ref:
            ol = orc.RefExtractor(NFEAT, SCALE, N_LEVELS, INI_TH, MIN_TH); orr = orc.RefExtractor(NFEAT, SCALE, N_LEVELS, INI_TH, MIN_TH)
        else:
            ol = orc.Extractor(NFEAT, SCALE, N_LEVELS, INI_TH, MIN_TH); orr = orc.Extractor(NFEAT, SCALE, N_LEVELS, INI_TH, MIN_TH)
        n = 0
        for i in range(tid, n_pairs, threads):
            kl, dl = ol.extract(left[i % POOL]); kr, dr = orr.extract(right[i % POOL])
            if use_ref:
                orc.stereo_match_ref(ol, orr, kl, dl, kr, dr, tabs["sf"], tabs["inv_sf"], MBF, 0.0, FX)
            else:
                orc.stereo_match(ol, orr, kl, dl, kr, dr, MBF, 0.0, FX)
            n += 1
        out[tid] = n

    out = [0] * threads
    ts = [threading.Thread(target=work, args=(t, out)) for t in range(threads)]
    t0 = time.perf_counter()
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    dt = time.perf_counter() - t0
    return sum(out) / dt, dt, ("reference" if use_ref else "port")


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    left, right = synth_pool()
    cores = os.cpu_count() or 1
    per_step = 12 * cores   # ~0.1 core-seconds per pair: a step is about 1.5 s of wall clock on all cores
    cpu_pairs_per_s(min(per_step, cores), cores, left, right)  # page in
    for _ in range(args.warmup):
        cpu_pairs_per_s(per_step, cores, left, right)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        kind = cpu_pairs_per_s(per_step, cores, left, right)[2]
    dt = time.perf_counter() - t0
    val = per_step * args.steps / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": WORKLOAD, "frame": "stereo pair", "pairs_per_step": per_step},
        "cpu_baseline": {"value": val, "unit": "frames/s", "cores": cores, "kind": kind,
                         "sample": "%d stereo pairs per step x %d steps, %d host threads; %s"
                                   % (per_step, args.steps, cores, CPU_DESC[kind])},
        "e2e": {"value": val, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------ GPU
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

    ext = importlib.import_module(PKG + ".extractor")
    stereo = importlib.import_module(PKG + ".stereo")
    capi = importlib.import_module(PKG + ".capi")
    B = args.batch
    left, right = synth_pool()
    reps = (B + POOL - 1) // POOL
    # every rank works on its own frames (weak scaling); rotate the pool so ranks differ
    hl = np.roll(np.tile(left, (reps, 1, 1))[:B], rank, axis=0)
    hr = np.roll(np.tile(right, (reps, 1, 1))[:B], rank, axis=0)

    gl = ext.ORBextractor(NFEAT, SCALE, N_LEVELS, INI_TH, MIN_TH, max_width=W, max_height=H, max_batch=B, device=local)
    gr = ext.ORBextractor(NFEAT, SCALE, N_LEVELS, INI_TH, MIN_TH, max_width=W, max_height=H, max_batch=B, device=local)
    cap = gl.kp_cap

    # ---- device-resident inputs for `value`
    dl = torch.from_numpy(hl).cuda(); dr = torch.from_numpy(hr).cuda()
    torch.cuda.synchronize()

    def step_device():
        gl.extract_device(dl.data_ptr(), B, W, H)
        gr.extract_device(dr.data_ptr(), B, W, H)
        stereo.compute_stereo_matches_device(gl, gr, MBF, FX, 0.0, batch=B)

    def sync_all():
        gl.sync(); gr.sync(); torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()

    sl = torch.cuda.ExternalStream(gl.stream, device=local)

    for _ in range(max(args.warmup, 3)):
        step_device()
    sync_all()
    l0 = gl.launch_count + gr.launch_count
    sampler = ClockSampler(local)
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    sync_all()
    # the stereo stage makes the left stream wait for the right stream, and the next step's right
    # extraction waits for the left stream: timing on the left stream covers the whole step chain
    e0.record(sl)
    for _ in range(args.steps):
        step_device()
    e1.record(sl)
    gr.sync()
    sync_all()
    ms_total = e0.elapsed_time(e1)
    clocks = sampler.stop()
    launches = (gl.launch_count + gr.launch_count - l0)
    t = torch.tensor([ms_total], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    value = world * B * args.steps / (ms_total * 1e-3)

    # ---- per-stage event timing (separate, untimed-for-value pass) for the roofline object
    gl.set_timing(True); gr.set_timing(True)
    acc = {}
    n_prof = 3
    for _ in range(n_prof):
        # serialised (one handle at a time) so that the per-stage event times are not inflated by
        # the other stream's kernels sharing the SMs
        gl.extract_device(dl.data_ptr(), B, W, H); gl.sync()
        gr.extract_device(dr.data_ptr(), B, W, H); gr.sync()
        stereo.compute_stereo_matches_device(gl, gr, MBF, FX, 0.0, batch=B); gl.sync()
        sl_ms = gl.stage_ms(); sr_ms = gr.stage_ms()
        for k in sl_ms:
            acc[k] = acc.get(k, 0.0) + sl_ms[k] + (sr_ms[k] if not k.startswith("stereo") else 0.0)
    gl.set_timing(False); gr.set_timing(False)
    stages = {k: v / n_prof for k, v in acc.items()}   # ms per step, both images of the pairs
    _, _, counts = gl.download(B)
    n_kp = float(np.mean(counts))

    # ---- e2e through the C ABI with pinned host buffers
    pl = torch.from_numpy(hl).pin_memory(); pr = torch.from_numpy(hr).pin_memory()
    kp_dt = capi.KP_DTYPE
    o_k = [torch.empty((B, cap, kp_dt.itemsize), dtype=torch.uint8).pin_memory() for _ in range(2)]
    o_d = [torch.empty((B, cap, 32), dtype=torch.uint8).pin_memory() for _ in range(2)]
    o_n = [np.zeros(B, np.int32) for _ in range(3)]
    o_u = torch.empty((B, cap), dtype=torch.float32).pin_memory(); o_z = torch.empty((B, cap), dtype=torch.float32).pin_memory()
    L = capi.lib()

    def step_host():
        # the call a user of the reference makes for a stereo frame: Frame::Frame(left, right, ...) =
        # ExtractORB x2 + ComputeStereoMatches (src/Frame.cc:60-115), batched, host buffers in and out
        capi.check(L.orb_stereo_frames(gl.handle, gr.handle, C.c_void_p(pl.data_ptr()), C.c_void_p(pr.data_ptr()), B, W, H, W,
                                       C.c_size_t(W * H), C.c_float(MBF), C.c_float(0.0), C.c_float(FX),
                                       C.c_void_p(o_k[0].data_ptr()), C.c_void_p(o_d[0].data_ptr()), capi.ptr(o_n[0]),
                                       C.c_void_p(o_k[1].data_ptr()), C.c_void_p(o_d[1].data_ptr()), capi.ptr(o_n[1]),
                                       C.c_void_p(o_u.data_ptr()), C.c_void_p(o_z.data_ptr()), cap))

    for _ in range(3):
        step_host()
    sync_all()
    e_steps = max(3, args.steps // 2)
    t0 = time.perf_counter()
    for _ in range(e_steps):
        step_host()
    sync_all()
    e_ms = (time.perf_counter() - t0) * 1e3
    t = torch.tensor([e_ms], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e_val = world * B * e_steps / (float(t.item()) * 1e-3)
    h2d = 2 * B * W * H
    # slot rows are downloaded whole (cap == the handle's kp_cap here): keypoints + descriptors of both sides, uRight, depth, counts
    d2h = int(B * cap * (2 * (kp_dt.itemsize + 32) + 8) + 2 * B * 4 + 8)

    matching = None
    if rank == 0:
        matching = bench_matching(local)
    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
        ext_stages = {k: v for k, v in stages.items()}
        dom = max(ext_stages, key=lambda k: ext_stages[k])
        # algorithmic bytes of the whole extraction of one image x images per launch,
        # attributed to the pipeline; the dominant kernel's own share is reported next to it
        alg_img = algorithmic_bytes_per_image(n_kp)
        per_kernel_bytes = {
            "pyramid": lambda: (1444097 - W * H) * 2 + W * H,       # read levels 0..6 once, write levels 1..7
            "fast_cells": lambda: 1444097 + 4 * 16000,               # read every level once, write candidate keys
            "quadtree": lambda: 3 * 4 * 16000,                        # gather + partition passes over candidate keys (lower bound)
            "blur": lambda: 2 * 1444097,                              # read + write every level
            "orient_describe": lambda: n_kp * (749 + 961 + 32 + 28),
        }
        dom_bytes = per_kernel_bytes.get(dom, lambda: alg_img)() * B  # per launch (one image side)
        dom_ms_launch = ext_stages[dom] / (1 if dom.startswith("stereo") else 2)
        achieved = dom_bytes / (dom_ms_launch * 1e-3) / 1e9
        pipe_ms = sum(v for k, v in stages.items())
        # the same launch against the INT issue rate: SURVEY 8(d) counts 80 simple integer ops per pyramid pixel for FAST
        sm_mhz = clocks.get("sm_mhz") or 1965.0
        alu_peak = 148 * 128 * sm_mhz * 1e6 / 1e12
        alu_ach = 1444097 * 80 * B / (stages["fast_cells"] / 2 * 1e-3) / 1e12
        traffic = NCU_TRAFFIC_BYTES_PER_IMAGE.get(dom)
        cores = os.cpu_count() or 1
        cpu_n = 12 * cores   # ~0.1 core-seconds per pair: about 20 s of CPU work
        cpu_val, cpu_dt, cpu_kind = cpu_pairs_per_s(cpu_n, cores, left, right)
        line = {
            "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
            "data": "synthetic",
            "config": {"workload": WORKLOAD, "frame": "stereo pair (2 images)", "pairs_per_step_per_gpu": B, "nfeatures": NFEAT,
                       "levels": N_LEVELS, "l2": "inputs larger than L2: %d MB of images + %d MB of intermediates per step"
                       % (2 * B * W * H >> 20, int(2 * B * 9.5)), "avg_keypoints": n_kp},
            "e2e": {"value": e_val, "unit": "frames/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "steps": e_steps},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": (traffic * B if traffic else None), "traffic_source": NCU_TRAFFIC_SOURCE,
                         "peak_source": peak_src,
                         "alu": {"kernel": "fast_cells", "achieved": alu_ach, "peak": alu_peak, "unit": "Top/s (int, algorithmic: 80 per pixel)",
                                 "frac": alu_ach / alu_peak},
                         "pipeline_achieved": 2 * B * alg_img / (pipe_ms * 1e-3) / 1e9,
                         "stage_ms_per_step": stages},
            "matching": matching,
            "cpu_baseline": {"value": cpu_val, "unit": "frames/s", "cores": cores, "kind": cpu_kind,
                             "sample": "%d stereo pairs on %d host threads in %.1f s; %s" % (cpu_n, cores, cpu_dt, CPU_DESC[cpu_kind])},
        }
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def bench_matching(local):
    """BASELINE.json configs[3] beside the headline step (rank 0, outside its timed region): ORBmatcher Hamming matching on
    2000 x 2000 descriptor sets, 1024 frame pairs: brute force (best / second best per row, device-resident, CUDA events on the
    matcher's stream) and grid-guided SearchByProjection(Frame, MapPoints) through the host API (one call per frame)."""
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
    m = mt.ORBmatcher(0.6, True, max_n=8192, max_pairs=1, device=local)
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
    # grid-guided projection search, host API, one frame per call
    f, mps, occ = synth.make_projection_scenario(N, N, 0)
    F = views.Frame(f["x"], f["y"], f["octave"], f["angle"], f["desc"], f["sf"], f["bounds"], f["u_right"])
    for _ in range(3):
        m.SearchByProjection(F, mps, 3.0, occ)
    t0 = time.perf_counter(); reps = 50
    for _ in range(reps):
        n_sp, _ = m.SearchByProjection(F, mps, 3.0, occ)
    sp_ms = (time.perf_counter() - t0) / reps * 1e3
    # the same call from T host threads, one matcher handle (own stream) each: ORB-SLAM2 itself runs its matchers from the
    # Tracking, LocalMapping and LoopClosing threads concurrently (threading)
    T = 8
    ms_ = [mt.ORBmatcher(0.6, True, max_n=8192, max_pairs=1, device=local) for _ in range(T)]
    def worker(mm, out, k):
        for _ in range(3):
            mm.SearchByProjection(F, mps, 3.0, occ)
        t = time.perf_counter()
        for _ in range(reps):
            mm.SearchByProjection(F, mps, 3.0, occ)
        out[k] = time.perf_counter() - t
    outs = [0.0] * T
    ths = [threading.Thread(target=worker, args=(ms_[k], outs, k)) for k in range(T)]
    t0 = time.perf_counter()
    for t in ths:
        t.start()
    for t in ths:
        t.join()
    mt_fps = T * reps / max(outs)
    popc = P * N * N * 8 / (bf_ms * 1e-3)
    return {"workload": "2000x2000 descriptors per frame pair, 1024 pairs (BASELINE.json configs[3])",
            "bruteforce": {"ms_per_1024_pairs": bf_ms, "pairs_per_s": P / (bf_ms * 1e-3), "popc_Tops": popc / 1e12,
                           "popc_peak_Tops": 148 * 16 * 1965e6 / 1e12, "frac": popc / (148 * 16 * 1965e6),
                           "note": "peak = 148 SM x 16 POPC/clk/SM (CUDA C Programming Guide arithmetic-throughput table; not re-measured here) x 1.965 GHz"},
            "search_by_projection_host_api": {"ms_per_frame": sp_ms, "frames_per_s": 1e3 / sp_ms, "matches": int(n_sp), "th": 3.0,
                                              "frames_per_s_8_host_threads": mt_fps}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=256, help="stereo pairs per step per GPU")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
