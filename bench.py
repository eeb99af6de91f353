#!/usr/bin/env python3
"""bench.py — headline benchmark of the ORB front-end hot path on B200 (contract: see the task statement / DESIGN.md).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

metric  : ORB extract frames/s at 640x480 / 1000 kp (BASELINE.json), inputs resident in HBM, device-timed.
A "step": one pass of ORBextractor::operator() over one batch of synthetic frames (per GPU).  The Hamming half of the
metric (brute-force top-2, Gcmp/s) is timed in its own loop and reported in the "hamming" object of the same line.
N > 1   : launched by torch.distributed.run, one rank per GPU; frames shard with no collective (weak scaling), the
          descriptor DB shards by rows and the per-shard top-2 are merged after one NCCL all-gather.
"""
import argparse
import json
import os
import queue
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path[:0] = [os.path.join(ROOT, "orb-slam2-annotation_b200")]

W, H, NFEAT = 640, 480, 1000                 # C1 shape: the configuration the metric is quoted on
FRAMES_PER_STEP = 512                        # per GPU: 157 MB of input per step (> 126 MB L2) + ~2 GB of intermediates
NDISTINCT = 32                               # distinct synthetic frames (tiled to the batch)
NQ, DB_ROWS = 2000, 10_000_000               # C5 shape per GPU (weak scaling: every rank holds a 10 M-row shard)
CAPACITY = 1280                              # per-frame output capacity (1000 requested -> ~1005 returned)


def level_bytes():
    """Algorithmic bytes per frame: B_frame = 3P - P0 - P7 + 60 N_kp; the level kernels' share is 3P - P0 - P7."""
    import numpy as np
    sc = np.float32(1.0); ws, hs = [], []
    for l in range(8):
        inv = np.float32(1.0) / sc
        ws.append(int(np.rint(np.float32(W) * inv))); hs.append(int(np.rint(np.float32(H) * inv)))
        sc = np.float32(np.float64(sc) * np.float64(np.float32(1.2)))
    P = [a * b for a, b in zip(ws, hs)]
    return 3 * sum(P) - P[0] - P[-1], P


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU while the timed region runs (NVML)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.sm, self.reasons, self.max_mhz = index, False, [], set(), None

    def run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            hd = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(hd, nv.NVML_CLOCK_SM)
            names = {nv.nvmlClocksEventReasonHwSlowdown: "hw_slowdown", nv.nvmlClocksEventReasonHwThermalSlowdown: "hw_thermal_slowdown",
                     nv.nvmlClocksEventReasonSwThermalSlowdown: "sw_thermal_slowdown", nv.nvmlClocksEventReasonSwPowerCap: "sw_power_cap"}
            while not self.stop_flag:
                self.sm.append(nv.nvmlDeviceGetClockInfo(hd, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksEventReasons(hd)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
                time.sleep(0.05)
        except Exception as e:      # NVML missing: report what we know
            self.reasons.add(f"nvml_unavailable:{type(e).__name__}")

    def result(self):
        self.stop_flag = True
        self.join(timeout=2)
        return {"sm_mhz": statistics.median(self.sm) if self.sm else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f)["hbm_gbs"], "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic():
    """dram bytes per level-kernel launch sequence from the committed ncu capture, if any (profiles/traffic.json)."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f)
    except Exception:
        return None


def cpu_reference_extract(seconds, nthreads):
    """The reference's own CPU ORBextractor (oracle/_ref when built, else the oracle port) on this box's host cores."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import pyoracle
    from synth import synth
    frames = [synth(W, H, 1 + i) for i in range(8)]
    kind = "reference" if pyoracle.ref_lib() is not None else "port"
    make = (lambda: pyoracle.RefExtractor(NFEAT)) if kind == "reference" else (lambda: pyoracle.OracleExtractor(NFEAT))
    counts = [0] * nthreads
    deadline = [0.0]

    def work(t):
        ex = make()
        ex.extract(frames[0])
        i = 0
        while time.perf_counter() < deadline[0]:
            ex.extract(frames[i % len(frames)])
            i += 1
        counts[t] = i

    deadline[0] = time.perf_counter() + seconds + 0.5
    t0 = time.perf_counter()
    th = [threading.Thread(target=work, args=(t,)) for t in range(nthreads)]
    [t.start() for t in th]; [t.join() for t in th]
    dt = time.perf_counter() - t0
    return sum(counts) / dt, kind, sum(counts), dt


def cpu_cv2_primitives(seconds=3.0):
    """BASELINE.md section 3's other leg: the image primitives of the reference on REAL OpenCV (cv2 4.x, SIMD / IPP, one thread): the 8-level
    INTER_LINEAR pyramid, GaussianBlur 7x7 sigma 2 and FAST-9-16 with NMS at iniTh and minTh on every level (whole level instead of
    per 30-px cell: fewer calls, same pixels).  Octree, IC_Angle and the descriptors are NOT included, so this is an upper bound for a
    reference built on the real library, to set beside the per-thread rate of the timed arm (reference ORBextractor.cpp over restated
    scalar primitives)."""
    try:
        import cv2
    except Exception:
        return None
    from synth import synth
    cv2.setNumThreads(1)
    frames = [synth(W, H, 1 + i) for i in range(4)]
    f_ini = cv2.FastFeatureDetector_create(20, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
    f_min = cv2.FastFeatureDetector_create(7, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)

    def one(img):
        lvl, sc = img, 1.0
        for l in range(8):
            if l > 0:
                sc *= 1.2
                lvl = cv2.resize(lvl, (int(round(W / sc)), int(round(H / sc))), interpolation=cv2.INTER_LINEAR)
            cv2.GaussianBlur(lvl, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101)
            f_ini.detect(lvl[16:-16, 16:-16]); f_min.detect(lvl[16:-16, 16:-16])
    one(frames[0])
    t0 = time.perf_counter(); n = 0
    while time.perf_counter() - t0 < seconds:
        one(frames[n % len(frames)]); n += 1
    dt = time.perf_counter() - t0
    return {"ms_per_frame_1_thread": 1e3 * dt / n, "frames": n, "opencv": cv2.__version__,
            "what": "pyramid + GaussianBlur + whole-level FAST(20) and FAST(7) only (no octree / angles / descriptors): upper bound for the reference on real OpenCV"}


def cpu_reference_match(nthreads, rows=200_000):
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import pyoracle
    from synth import synth_descriptor_db, synth_queries
    db = synth_descriptor_db(rows)
    q = synth_queries(db, NQ)
    t0 = time.perf_counter()
    pyoracle.bf_top2(db, q, nthreads=nthreads)
    dt = time.perf_counter() - t0
    return NQ * rows / dt / 1e9, dt


def run_reference(args, rank, world):
    if rank != 0:
        return
    nthreads = os.cpu_count() or 1
    per_step = 4.0
    t_all = time.perf_counter()
    for _ in range(max(args.warmup, 1)):
        cpu_reference_extract(0.5, nthreads)
    vals = []
    for _ in range(args.steps):
        fps, kind, n, dt = cpu_reference_extract(per_step, nthreads)
        vals.append((fps, n, dt))
    total_frames = sum(v[1] for v in vals); total_t = sum(v[2] for v in vals)
    value = total_frames / total_t
    gcmp, _ = cpu_reference_match(nthreads)
    line = {"impl": "reference", "metric": "orb_extract_frames_per_s_640x480_1000kp", "value": value, "unit": "frames/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total_t / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": f"C1 640x480 1000kp, {total_frames // args.steps} frames/step on {nthreads} host threads (bounded sample)",
                       "hamming_workload": f"{NQ} q x 200000 rows top-2"},
            "cpu_baseline": {"value": value, "unit": "frames/s", "cores": nthreads, "kind": kind,
                             "sample": f"{total_frames} frames in {total_t:.1f} s, reference ORBextractor.cpp over restated OpenCV primitives"},
            "e2e": {"value": value, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "hamming": {"metric": "bf_top2_gcmp_per_s", "value": gcmp, "unit": "Gcmp/s", "cores": nthreads},
            "wall_s": time.perf_counter() - t_all}
    emit(line)


def _pcts(us):
    us = sorted(us)
    return {"p50_us": us[len(us) // 2], "p99_us": us[min(len(us) - 1, int(0.99 * len(us)))], "min_us": us[0], "calls": len(us)}


def measure_latency(local_rank, dev):
    """Config 1 as BASELINE.json words it: ONE synthetic 640x480 frame through the call Frame::ExtractORB makes (Frame.cpp:259-265 ->
    orbx_extract: H2D, kernels, D2H, sync), and the same launch sequence device-resident with plain launches vs one CUDA graph."""
    import ctypes as C
    import numpy as np
    import torch
    import orbb200
    from synth import synth
    L = orbb200.lib()
    img = synth(W, H, 4242)
    ex = orbb200.ORBextractor(NFEAT, 1.2, 8, 20, 7, max_size=(W, H), max_batch=1, device=local_rank)
    kps = np.zeros(CAPACITY, orbb200.KP_DTYPE); desc = np.zeros((CAPACITY, 32), np.uint8); n = C.c_int()
    pin = torch.from_numpy(img.copy()).pin_memory()
    pk = torch.zeros((CAPACITY, 7), dtype=torch.float32).pin_memory(); pd = torch.zeros((CAPACITY, 32), dtype=torch.uint8).pin_memory()

    def call_pageable():
        rc = L.orbx_extract(ex.h, img.ctypes.data_as(C.c_void_p), W, H, W, kps.ctypes.data_as(C.c_void_p), desc.ctypes.data_as(C.c_void_p), CAPACITY, C.byref(n))
        assert rc == 0, rc

    def call_pinned():
        rc = L.orbx_extract(ex.h, C.c_void_p(pin.data_ptr()), W, H, W, C.c_void_p(pk.data_ptr()), C.c_void_p(pd.data_ptr()), CAPACITY, C.byref(n))
        assert rc == 0, rc

    dimg = pin.to(dev)
    dk = torch.zeros((CAPACITY, 7), dtype=torch.float32, device=dev); dd = torch.zeros((CAPACITY, 32), dtype=torch.uint8, device=dev)
    dc = torch.zeros(1, dtype=torch.int32, device=dev)

    def call_device():
        ex.extract_batch_device(dimg.data_ptr(), 1, W, H, W, W * H, dk.data_ptr(), dd.data_ptr(), dc.data_ptr(), CAPACITY)
        ex.sync()

    def timed(fn, calls=300, warm=20):
        for _ in range(warm):
            fn()
        us = []
        for _ in range(calls):
            t0 = time.perf_counter(); fn(); us.append((time.perf_counter() - t0) * 1e6)
        return _pcts(us)

    out = {"workload": "one 640x480 synthetic frame, 1000 kp (BASELINE.json configs[0])",
           "orbx_extract_pageable_host": timed(call_pageable), "keypoints": 0}
    out["keypoints"] = int(n.value)
    out["orbx_extract_pinned_host"] = timed(call_pinned)
    # the monocular Frame constructor (Frame.cpp:171-224, TUM1.yaml calibration): member by member, and as one call
    K4 = (C.c_float * 4)(517.306408, 516.469215, 318.643040, 255.313989)
    D5 = (C.c_float * 5)(0.262383, -0.953104, -0.005358, 0.002628, 1.163314)
    b4 = (C.c_float * 4)()
    assert L.orbm_image_bounds(W, H, K4, D5, 5, b4, local_rank) == 0
    kun = np.zeros(CAPACITY, orbb200.KP_DTYPE)
    vp = C.c_void_p

    def ctor_members():
        call_pageable()
        assert L.orbm_undistort_keypoints(kps.ctypes.data_as(vp), n.value, K4, D5, 5, kun.ctypes.data_as(vp), local_rank) == 0
        fh = vp()
        assert L.orbm_frame_create(C.byref(fh), kun.ctypes.data_as(vp), desc.ctypes.data_as(vp), n.value, b4[0], b4[1], b4[2], b4[3], local_rank) == 0
        L.orbm_frame_destroy(fh)

    def ctor_one_call():
        fh = vp()
        rc = L.orbm_mono_frame_create(ex.h, img.ctypes.data_as(vp), W, H, W, K4, D5, 5, b4[0], b4[1], b4[2], b4[3], C.byref(fh), kps.ctypes.data_as(vp),
                                      kun.ctypes.data_as(vp), desc.ctypes.data_as(vp), C.byref(n), CAPACITY)
        assert rc == 0, rc
        L.orbm_frame_destroy(fh)
    out["mono_frame_constructor_member_calls"] = timed(ctor_members)
    out["mono_frame_constructor_one_call"] = timed(ctor_one_call)
    out["device_resident_plain_launches"] = timed(call_device)
    ref_cnt = int(dc.item()); ref_desc = dd.clone()
    ex.graph_capture(dimg.data_ptr(), 1, W, H, W, W * H, dk.data_ptr(), dd.data_ptr(), dc.data_ptr(), CAPACITY)
    dd.zero_(); dc.zero_()

    def call_graph():
        ex.graph_launch()
        ex.sync()
    out["device_resident_cuda_graph"] = timed(call_graph)
    out["graph_equals_plain"] = bool(int(dc.item()) == ref_cnt and torch.equal(dd, ref_desc))
    # the kernels alone (CUDA events around 50 graph replays, no host in between)
    st = torch.cuda.ExternalStream(ex.stream(), device=dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(50):
        ex.graph_launch()
    e1.record(st)
    ex.sync()
    out["gpu_us_per_frame_back_to_back_graphs"] = e0.elapsed_time(e1) * 1e3 / 50
    return out


def measure_link_floor(world, dev, barrier, max_over_ranks, h2d_bytes, d2h_bytes, frames):
    """What the host can move: every rank copies one step's input bytes H2D and one step's result bytes D2H from / to pinned memory, each
    direction split over two streams (as the two handles of the e2e loop do), both directions and all ranks at the same time; best of three
    passes.  e2e can not beat frames / this time by more than the probe's own imperfection."""
    import torch
    hin = torch.empty(h2d_bytes, dtype=torch.uint8).pin_memory(); din = torch.empty(h2d_bytes, dtype=torch.uint8, device=dev)
    hout = torch.empty(d2h_bytes, dtype=torch.uint8).pin_memory(); dout = torch.empty(d2h_bytes, dtype=torch.uint8, device=dev)
    streams = [torch.cuda.Stream(dev) for _ in range(4)]
    hi, ho = h2d_bytes // 2, d2h_bytes // 2

    def go(n):
        for _ in range(n):
            with torch.cuda.stream(streams[0]):
                din[:hi].copy_(hin[:hi], non_blocking=True)
            with torch.cuda.stream(streams[1]):
                din[hi:].copy_(hin[hi:], non_blocking=True)
            with torch.cuda.stream(streams[2]):
                hout[:ho].copy_(dout[:ho], non_blocking=True)
            with torch.cuda.stream(streams[3]):
                hout[ho:].copy_(dout[ho:], non_blocking=True)
        for st in streams:
            st.synchronize()
    go(2)
    iters, best = 5, None
    for _ in range(3):
        barrier()
        t0 = time.perf_counter()
        go(iters)
        barrier()
        dt = max_over_ranks(time.perf_counter() - t0)
        best = dt if best is None else min(best, dt)
    return {"frames_per_s": world * frames * iters / best, "h2d_GBps_per_gpu": h2d_bytes * iters / best / 1e9, "d2h_GBps_per_gpu": d2h_bytes * iters / best / 1e9,
            "what": f"{world} rank(s) x (H2D {h2d_bytes} B + D2H {d2h_bytes} B) per step from pinned memory, two streams per direction, both directions and all ranks at once, best of 3"}


def measure_configs(rank, world, local_rank, dev, barrier, max_over_ranks):
    """Driver-timed rows for BASELINE.json configs[1..3] (C2 KITTI stereo pair, C3 EuRoC batch, C4 4K batch).  Frames shard over
    ranks with no collective; every row is the whole job (all ranks), device-timed with CUDA events, max over ranks."""
    import ctypes as C
    import numpy as np
    import torch
    import orbb200
    from synth import synth
    L = orbb200.lib()
    rows = {}

    def batch_row(w, h, nf, F, cap, nd, iters, seed):
        distinct = np.stack([synth(w, h, seed + 100 * rank + i) for i in range(nd)])
        frames = torch.from_numpy(np.ascontiguousarray(np.tile(distinct, (F // nd, 1, 1)))).to(dev)
        ex = orbb200.ORBextractor(nf, 1.2, 8, 20, 7, max_size=(w, h), max_batch=F, device=local_rank)
        st = torch.cuda.ExternalStream(ex.stream(), device=dev)
        kps = torch.zeros((F, cap, 7), dtype=torch.float32, device=dev); desc = torch.zeros((F, cap, 32), dtype=torch.uint8, device=dev)
        cnt = torch.zeros(F, dtype=torch.int32, device=dev)
        step = lambda: ex.extract_batch_device(frames.data_ptr(), F, w, h, w, w * h, kps.data_ptr(), desc.data_ptr(), cnt.data_ptr(), cap)
        for _ in range(3):
            step()
        ex.sync(); barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for _ in range(iters):
            step()
        e1.record(st)
        ex.sync(); barrier()
        ms = max_over_ranks(e0.elapsed_time(e1)) / iters
        ex.set_profiling(True)
        for _ in range(2):
            step()
        prof = ex.get_profile(); ex.set_profiling(False)
        sc = np.float32(1.0); P = []
        for l in range(8):
            inv = np.float32(1.0) / sc
            P.append(int(np.rint(np.float32(w) * inv)) * int(np.rint(np.float32(h) * inv)))
            sc = np.float32(np.float64(sc) * np.float64(np.float32(1.2)))
        lvl_bytes = 3 * sum(P) - P[0] - P[-1]
        lvl_ms = prof["level_ms"] / prof["calls"]
        peak, _ = measured_peaks()
        nkp = float(cnt.float().mean().item())
        return {"frames_per_s": world * F / ms * 1e3, "ms_per_step": ms, "frames_per_step_per_gpu": F, "steps": iters, "keypoints_per_frame": nkp,
                "keypoints_per_s": world * F * nkp / ms * 1e3, "distinct_frames": nd,
                "step_share_ms": {k: prof[k] / prof["calls"] for k in ("level_ms", "octree_ms", "desc_ms")},
                "level_kernels_hbm_frac": lvl_bytes * F / lvl_ms / 1e6 / peak}

    rows["C3_euroc_752x480_1200kp_batch"] = batch_row(752, 480, 1200, 512, 1536, 16, 6, 7000)
    rows["C4_4k_3840x2160_8000kp_batch"] = batch_row(3840, 2160, 8000, 64, 9216, 2, 5, 9000)
    rows["C2_kitti_1241x376_2000kp_batch"] = batch_row(1241, 376, 2000, 256, 2560, 8, 6, 8000)
    torch.cuda.empty_cache()

    # ---- C2 pair pipeline through the host-pointer C ABI: extract L, extract R, ComputeStereoMatches, SearchByProjection(last frame) ----
    w, h, nf, cap = 1241, 376, 2000, 2560
    big = synth(w + 80, h + 16, 8100 + rank)
    pairs = []
    for (ox, oy) in ((8, 8), (11, 9)):                      # two consecutive stereo pairs: the camera moved by (3, 1) pixels
        left = big[oy:oy + h, ox + 40:ox + 40 + w].copy()
        right = big[oy:oy + h, ox + 28:ox + 28 + w].copy()  # constant 12-px disparity
        pairs.append((left, right))
    vp = C.c_void_p
    ptr = lambda a: a.ctypes.data_as(vp)
    mbf, bounds = 386.1448, (0.0, 0.0, float(w), float(h))

    class Pipeline:
        """One tracking thread's state: two extractor handles (left / right camera) and its host buffers."""

        def __init__(self, fused=False):
            self.fused = fused
            if fused:                # Frame::Frame (stereo) as one call: both images are one two-frame batch of one handle
                self.exL = orbb200.ORBextractor(nf, 1.2, 8, 20, 7, max_size=(w, h), max_batch=2, device=local_rank)
                self.one_pair = self.one_pair_fused
            else:
                self.exL = orbb200.ORBextractor(nf, 1.2, 8, 20, 7, max_size=(w, h), device=local_rank)
                self.exR = orbb200.ORBextractor(nf, 1.2, 8, 20, 7, max_size=(w, h), device=local_rank)
            self.scale = self.exL.GetScaleFactors()
            self.bufs = [dict(kL=np.zeros(cap, orbb200.KP_DTYPE), dL=np.zeros((cap, 32), np.uint8), kR=np.zeros(cap, orbb200.KP_DTYPE), dR=np.zeros((cap, 32), np.uint8),
                              ur=np.zeros(cap, np.float32), dp=np.zeros(cap, np.float32), mp=np.zeros(cap, np.int32), ob=np.zeros(cap, np.int32)) for _ in range(2)]
            self.nL, self.nR, self.nm = C.c_int(), C.c_int(), C.c_int()
            self.stats = {"stereo": 0, "proj": 0, "kp": 0}
            self.right_job, self.right_done = queue.Queue(), queue.Queue()
            if not fused:
                threading.Thread(target=self._right_worker, daemon=True).start()
            self.one_pair(0, None); n0 = self.nL.value
            self.one_pair(1, None); n1 = self.nL.value
            self.qs = [self.queries_from(self.bufs[1], n1, 3.0, 1.0), self.queries_from(self.bufs[0], n0, -3.0, -1.0)]   # queries for pair 0 come from pair 1 and vice versa
            for i in range(4):
                self.one_pair(i, self.qs[i & 1])
            self.stats.update(stereo=0, proj=0, kp=0)

        def _right_worker(self):
            while True:
                job = self.right_job.get()
                if job is None:
                    return
                right, b = job
                self.right_done.put(L.orbx_extract(self.exR.h, ptr(right), w, h, w, ptr(b["kR"]), ptr(b["dR"]), cap, C.byref(self.nR)))

        def close(self):
            self.right_job.put(None)

        def one_pair(self, i, queries):
            b = self.bufs[i & 1]; left, right = pairs[i & 1]
            nL, nR, nm = self.nL, self.nR, self.nm
            # the reference extracts the two images on two threads (Frame.cpp:84-87: threadLeft / threadRight): so does this pipeline
            self.right_job.put((right, b))
            assert L.orbx_extract(self.exL.h, ptr(left), w, h, w, ptr(b["kL"]), ptr(b["dL"]), cap, C.byref(nL)) == 0
            assert self.right_done.get() == 0
            assert L.orbm_stereo_match(self.exL.h, self.exR.h, ptr(b["kL"]), ptr(b["dL"]), nL.value, ptr(b["kR"]), ptr(b["dR"]), nR.value,
                                       C.c_float(mbf), C.c_float(0.0), ptr(b["ur"]), ptr(b["dp"])) == 0
            fh = vp()
            assert L.orbm_frame_create(C.byref(fh), ptr(b["kL"]), ptr(b["dL"]), nL.value, *[C.c_float(x) for x in bounds], local_rank) == 0
            if queries is not None:
                b["mp"][:nL.value] = -1; b["ob"][:nL.value] = 0
                assert L.orbm_search_by_projection_frame(fh, ptr(b["ur"]), ptr(queries), len(queries), 1, ptr(b["mp"]), ptr(b["ob"]), C.byref(nm)) == 0
                self.stats["proj"] += nm.value
            L.orbm_frame_destroy(fh)
            self.stats["stereo"] += int((b["ur"][:nL.value] > 0).sum()); self.stats["kp"] += nL.value + nR.value

        def one_pair_fused(self, i, queries):
            b = self.bufs[i & 1]; left, right = pairs[i & 1]
            nL, nR, nm = self.nL, self.nR, self.nm
            fh = vp()
            assert L.orbm_stereo_frame_create(self.exL.h, ptr(left), ptr(right), w, h, w, C.c_float(mbf), C.c_float(0.0), *[C.c_float(x) for x in bounds], C.byref(fh),
                                              ptr(b["kL"]), ptr(b["dL"]), C.byref(nL), ptr(b["kR"]), ptr(b["dR"]), C.byref(nR), cap, ptr(b["ur"]), ptr(b["dp"])) == 0
            if queries is not None:
                b["mp"][:nL.value] = -1; b["ob"][:nL.value] = 0
                assert L.orbm_search_by_projection_frame(fh, orbb200.URIGHT_OF_FRAME, ptr(queries), len(queries), 1, ptr(b["mp"]), ptr(b["ob"]), C.byref(nm)) == 0
                self.stats["proj"] += nm.value
            L.orbm_frame_destroy(fh)
            self.stats["stereo"] += int((b["ur"][:nL.value] > 0).sum()); self.stats["kp"] += nL.value + nR.value

        def queries_from(self, b, n, dx, dy):                  # the previous left frame's keypoints projected into the current one
            q = np.zeros(n, orbb200.PROJ_QUERY_DTYPE)
            k = b["kL"][:n]
            q["q"]["u"] = k["x"] + dx; q["q"]["v"] = k["y"] + dy; q["q"]["r"] = 15.0 * self.scale[k["octave"]]
            q["q"]["min_level"] = k["octave"] - 1; q["q"]["max_level"] = k["octave"] + 1; q["q"]["desc"] = b["dL"][:n]
            q["u_right"] = q["q"]["u"] - 12.0; q["max_err_right"] = 15.0 * self.scale[k["octave"]]
            q["id"] = np.arange(n); q["has_obs"] = 1; q["angle"] = k["angle"]
            return q

        def run(self, npairs):
            for i in range(npairs):
                self.one_pair(i, self.qs[i & 1])

    apis = {False: ("per pair, host pointers: orbx_extract (L) and orbx_extract (R) on two threads as in Frame.cpp:84-87, then orbm_stereo_match + orbm_frame_create + "
                    "orbm_search_by_projection_frame + orbm_frame_destroy; wall clock, H2D/D2H inside every call"),
            True: ("per pair, host pointers: orbm_stereo_frame_create (Frame.cpp:61-115 as one call: both images as one batch, stereo matching and grid on the device "
                   "results) + orbm_search_by_projection_frame (mvuRight of the frame) + orbm_frame_destroy; wall clock, H2D/D2H inside every call")}
    for fused, nthreads, npairs in ((False, 1, 40), (False, 4, 40), (True, 1, 40), (True, 4, 40)):
        pipes = [Pipeline(fused) for _ in range(nthreads)]
        barrier()
        t0 = time.perf_counter()
        th = [threading.Thread(target=p_.run, args=(npairs,)) for p_ in pipes]
        [t.start() for t in th]; [t.join() for t in th]
        torch.cuda.synchronize(); barrier()
        dt = max_over_ranks(time.perf_counter() - t0)
        st_ = {k: sum(p_.stats[k] for p_ in pipes) for k in ("stereo", "proj", "kp")}
        tot = nthreads * npairs
        name = ("C2_kitti_stereo_pair_fused_constructor" if fused else "C2_kitti_stereo_pair_pipeline") + ("" if nthreads == 1 else f"_{nthreads}_host_threads")
        rows[name] = {
            "pairs_per_s": world * tot / dt, "frames_per_s": 2 * world * tot / dt, "ms_per_pair_per_thread": dt / npairs * 1e3, "host_threads_per_gpu": nthreads,
            "stereo_matches_per_s": world * st_["stereo"] / dt, "projection_matches_per_s": world * st_["proj"] / dt,
            "matches_per_s": world * (st_["stereo"] + st_["proj"]) / dt, "keypoints_per_frame": st_["kp"] / (2 * tot),
            "stereo_matches_per_pair": st_["stereo"] / tot, "projection_matches_per_pair": st_["proj"] / tot, "api": apis[fused]}
        for p_ in pipes:
            p_.close()
        del pipes
    return rows


def measure_c5_strong(rank, world, local_rank, dev, stream, barrier, max_over_ranks):
    """BASELINE.json configs[4] as stated: 2000 queries x 10 M rows IN TOTAL, rows block-sharded over the ranks (1.25 M per GPU at N = 8),
    per-shard top-2 merged after ONE packed NCCL all-gather (sharding.ShardedDescriptorDB / gather_top2); the merged result is checked."""
    import numpy as np
    import torch
    import torch.distributed as dist
    import orbb200
    import sharding
    S = sharding.ShardedDescriptorDB(DB_ROWS, local_rank)
    nrows = S.end - S.begin
    g = torch.Generator(device=dev); g.manual_seed(4242 + rank)
    shard = torch.randint(0, 256, (nrows, 32), dtype=torch.uint8, device=dev, generator=g)
    S.append_device(shard)
    rng = np.random.default_rng(99)                                   # the same on every rank
    nh = NQ // 2
    planted = np.sort(rng.choice(DB_ROWS, nh, replace=False))
    mine = (planted >= S.begin) & (planted < S.end)
    rows = torch.zeros((nh, 32), dtype=torch.int32, device=dev)
    if mine.any():
        rows[torch.from_numpy(np.flatnonzero(mine)).to(dev)] = shard[torch.from_numpy(planted[mine] - S.begin).to(dev)].to(torch.int32)
    if world > 1:
        dist.all_reduce(rows)                                         # every planted row is owned by exactly one rank
    base = rows.to(torch.uint8).cpu().numpy()
    bits = np.unpackbits(base, axis=1)
    flips = rng.integers(0, 61, nh)
    for i in range(nh):
        if flips[i]:
            bits[i, rng.choice(256, int(flips[i]), replace=False)] ^= 1
    q = rng.integers(0, 256, (NQ, 32), dtype=np.uint8)
    q[:nh] = np.packbits(bits, axis=1)
    qd = torch.from_numpy(q).to(dev)
    del shard
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
    ea, eb, ec = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    d1 = torch.empty(NQ, dtype=torch.int32, device=dev); i1 = torch.empty(NQ, dtype=torch.int64, device=dev); d2 = torch.empty(NQ, dtype=torch.int32, device=dev)

    def step():
        """scan of the local shard | packed all-gather + merge"""
        S.db.top2_device(qd.data_ptr(), NQ, d1.data_ptr(), i1.data_ptr(), d2.data_ptr(), stream.cuda_stream)
        eb.record(stream)
        if world > 1:
            d1a, i1a, d2a = sharding.gather_top2(d1, i1, d2)
            orbb200.merge_top2_device(d1a.data_ptr(), i1a.data_ptr(), d2a.data_ptr(), world, NQ, d1.data_ptr(), i1.data_ptr(), d2.data_ptr(), stream.cuda_stream)
    with torch.cuda.stream(stream):
        for _ in range(3):
            step()
    barrier()
    iters, ms_scan, ms_all = 10, 0.0, 0.0
    l0 = orbb200.launch_count()
    for _ in range(iters):
        with torch.cuda.stream(stream):
            flush.zero_()
            if world > 1:
                dist.barrier()                  # ranks start the step together: the gather then measures the exchange, not the skew
            ea.record(stream)
            step()
            ec.record(stream)
        stream.synchronize()
        ms_scan += ea.elapsed_time(eb); ms_all += ea.elapsed_time(ec)
    barrier()
    launches = orbb200.launch_count() - l0
    ms_all = max_over_ranks(ms_all) / iters; ms_scan = max_over_ranks(ms_scan) / iters
    r1, ri, r2 = d1.cpu().numpy(), i1.cpu().numpy(), d2.cpu().numpy()
    ok = bool(np.array_equal(r1[:nh], flips) and np.array_equal(ri[:nh], planted) and np.all(r1 <= r2) and np.all((ri >= 0) & (ri < DB_ROWS)))
    if not ok:
        raise AssertionError("C5-strong: merged top-2 of the planted queries is wrong")
    return {"metric": "bf_top2_gcmp_per_s", "value": NQ * DB_ROWS / (ms_all * 1e-3) / 1e9, "unit": "Gcmp/s", "scaling": "strong",
            "workload": f"{NQ} queries x {DB_ROWS} rows in total = {nrows} rows per GPU on {world} GPU(s)", "ms_per_step": ms_all,
            "scan_ms": ms_scan, "gather_merge_ms": ms_all - ms_scan, "limiter": "scan" if ms_scan > ms_all - ms_scan else "all-gather + merge",
            "steps": iters, "gpu_launches": int(launches), "queries_per_s": NQ / (ms_all * 1e-3),
            "exchange": "one all_gather_into_tensor of 16 B per query (sharding.gather_top2) + bf_merge_shards_kernel" if world > 1 else "none (one shard)",
            "merge_checked": ok}


def run_ours(args, rank, world, local_rank):
    import numpy as np
    import torch
    import torch.distributed as dist
    import orbb200
    from synth import synth, synth_queries

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    B, K, Wm = FRAMES_PER_STEP, args.steps, args.warmup

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---------------- inputs ----------------
    distinct = np.stack([synth(W, H, 1000 * rank + 1 + i) for i in range(NDISTINCT)])
    host_frames = torch.from_numpy(np.ascontiguousarray(np.tile(distinct, (B // NDISTINCT, 1, 1)))).pin_memory()
    dev_frames = host_frames.to(dev)
    ex = orbb200.ORBextractor(NFEAT, 1.2, 8, 20, 7, max_size=(W, H), max_batch=B, device=local_rank)
    stream = torch.cuda.ExternalStream(ex.stream(), device=dev)
    kps = torch.zeros((B, CAPACITY, 7), dtype=torch.float32, device=dev)
    desc = torch.zeros((B, CAPACITY, 32), dtype=torch.uint8, device=dev)
    cnt = torch.zeros(B, dtype=torch.int32, device=dev)

    def step_extract():
        ex.extract_batch_device(dev_frames.data_ptr(), B, W, H, W, W * H, kps.data_ptr(), desc.data_ptr(), cnt.data_ptr(), CAPACITY)

    # ---------------- extraction: device-resident, device-timed ----------------
    for _ in range(max(Wm, 3)):
        step_extract()
    ex.sync()
    barrier()
    sampler = ClockSampler(local_rank); sampler.start()
    l0 = orbb200.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(K):
        step_extract()
    e1.record(stream)
    ex.sync()
    barrier()
    launches = orbb200.launch_count() - l0
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    clocks = sampler.result()
    value = world * B * K / (ms_total * 1e-3)
    nkp = float(cnt.float().mean().item())

    # ---------------- roofline of the dominant kernel (per-level resize/blur/FAST), live CUDA events ----------------
    ex.set_profiling(True)
    for _ in range(5):
        step_extract()
    prof = ex.get_profile()
    ex.set_profiling(False)
    lvl_bytes, P = level_bytes()
    lvl_ms = prof["level_ms"] / prof["calls"]
    peak, peak_src = measured_peaks()
    achieved = lvl_bytes * B / (lvl_ms * 1e-3) / 1e9
    traffic = ncu_traffic()
    roofline = {"kernel": "level3_kernel (8 launches per batch: TMA-staged tile -> blur + FAST/NMS + next-level resize, fused per pyramid level)", "bound": "hbm",
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "peak_source": peak_src,
                "algorithmic_bytes_per_frame": lvl_bytes, "frames_per_launch_sequence": B, "ms_per_launch_sequence": lvl_ms,
                "traffic": traffic["level_kernel_bytes_per_sequence"] if traffic else None,
                "step_share": {k: prof[k] / prof["calls"] for k in ("level_ms", "octree_ms", "desc_ms")},
                "whole_step": {"algorithmic_bytes_per_frame": lvl_bytes + 60 * nkp,
                               "achieved": (lvl_bytes + 60 * nkp) * B * K / (ms_total * 1e-3) / 1e9 * 1.0,
                               "frac": (lvl_bytes + 60 * nkp) * B * K / (ms_total * 1e-3) / 1e9 / peak}}

    # ---------------- end to end: pinned host frames in, host results out, through the C ABI ----------------
    # Two handles, one host thread each, work on alternate batches (the reference runs its two extractors the same way,
    # Frame.cpp:84-87): while one batch drains its last kernels and result copies, the next batch's frames are already
    # crossing PCIe.  Every call is a complete orbx_extract_batch: H2D of its 512 frames, kernels, D2H of its results.
    import ctypes as C
    L = orbb200.lib()
    NH = 2
    exs = [ex] + [orbb200.ORBextractor(NFEAT, 1.2, 8, 20, 7, max_size=(W, H), max_batch=B, device=local_rank) for _ in range(NH - 1)]
    outs = []
    for _ in range(NH):
        outs.append((torch.zeros((B, CAPACITY, 7), dtype=torch.float32).pin_memory(),
                     torch.zeros((B, CAPACITY, 32), dtype=torch.uint8).pin_memory(), np.zeros(B, np.int32)))

    def step_e2e(i):
        hk, hd, hc = outs[i]
        rc = L.orbx_extract_batch(exs[i].h, C.c_void_p(host_frames.data_ptr()), B, W, H, W, W * H, C.c_void_p(hk.data_ptr()),
                                  C.c_void_p(hd.data_ptr()), hc.ctypes.data_as(C.c_void_p), CAPACITY)
        if rc != 0:
            raise orbb200.OrbError(rc, "orbx_extract_batch")

    def run_e2e(calls_per_handle):
        errs = []

        def work(i):
            try:
                for _ in range(calls_per_handle):
                    step_e2e(i)
            except Exception as e:      # surfaced after join
                errs.append(e)
        th = [threading.Thread(target=work, args=(i,)) for i in range(NH)]
        [t.start() for t in th]; [t.join() for t in th]
        if errs:
            raise errs[0]
    run_e2e(2)
    for i in range(NH):
        assert int(outs[i][2].sum()) == int(cnt.sum().item()), "e2e results differ from the device-resident path"
    barrier()
    t0 = time.perf_counter()
    Ke = max(3, K // 2)                 # calls per handle
    run_e2e(Ke)
    barrier()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    e2e = {"value": world * B * Ke * NH / e2e_s, "unit": "frames/s", "h2d_bytes_per_step": B * W * H,
           "d2h_bytes_per_step": B * CAPACITY * (28 + 32) + 8 * B, "steps": Ke * NH,
           "api": f"orbx_extract_batch (pinned host frames -> host keypoints+descriptors), {NH} handles on {NH} host threads, "
                  "alternate batches in flight"}
    del exs[1:]
    e2e["link_floor"] = measure_link_floor(world, dev, barrier, max_over_ranks, e2e["h2d_bytes_per_step"], e2e["d2h_bytes_per_step"], B)
    e2e["frac_of_link_floor"] = e2e["value"] / e2e["link_floor"]["frames_per_s"]

    # ---------------- Hamming: brute-force top-2, 2000 queries x 10 M rows per GPU ----------------
    g = torch.Generator(device=dev); g.manual_seed(42 + rank)
    db = torch.randint(0, 256, (DB_ROWS, 32), dtype=torch.uint8, device=dev, generator=g)
    D = orbb200.DescriptorDB(DB_ROWS, row_offset=rank * DB_ROWS, device=local_rank)
    D.append_device(db.data_ptr(), DB_ROWS)
    # queries: half are DB rows with k ~ U{0..60} flipped bits, half fresh random rows
    rng = np.random.default_rng(7 + rank)
    planted = torch.from_numpy(rng.integers(0, DB_ROWS, NQ // 2)).to(dev)
    qh = synth_queries(db[planted].cpu().numpy(), NQ, seed=7)
    del db
    qd = torch.from_numpy(qh).to(dev)
    d1 = torch.empty(NQ, dtype=torch.int32, device=dev); i1 = torch.empty(NQ, dtype=torch.int64, device=dev); d2 = torch.empty(NQ, dtype=torch.int32, device=dev)
    if world > 1:
        import sharding
        m1 = torch.empty_like(d1); mi = torch.empty_like(i1); m2 = torch.empty_like(d2)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)

    def step_match():
        with torch.cuda.stream(stream):
            D.top2_device(qd.data_ptr(), NQ, d1.data_ptr(), i1.data_ptr(), d2.data_ptr(), stream.cuda_stream)
            if world > 1:       # the one real exchange step: ONE packed all-gather of the per-shard (d1, idx1, d2) (16 B per query), merge on every rank
                g1, gi, g2 = sharding.gather_top2(d1, i1, d2)
                orbb200.merge_top2_device(g1.data_ptr(), gi.data_ptr(), g2.data_ptr(), world, NQ, m1.data_ptr(), mi.data_ptr(), m2.data_ptr(),
                                          stream.cuda_stream)
    for _ in range(3):
        step_match()
    barrier()
    Km = max(3, min(K, 10))
    lm0 = orbb200.launch_count()
    ms_match = 0.0
    ea, eb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(Km):
        with torch.cuda.stream(stream):
            flush.zero_()                      # L2 flush between timed iterations (the DB itself is 320 MB > L2 as well)
        ea.record(stream)
        step_match()
        eb.record(stream)
        stream.synchronize()
        ms_match += ea.elapsed_time(eb)
    barrier()
    match_launches = orbb200.launch_count() - lm0
    ms_match = max_over_ranks(ms_match)
    gcmp = world * NQ * DB_ROWS * Km / (ms_match * 1e-3) / 1e9
    popc_peak = orbb200.popc_peak(local_rank)
    imma_peak = orbb200.imma_peak(local_rank)
    tc_peak = orbb200.tcgen05_i8_peak(local_rank)
    tops = 512 * gcmp / world / 1e3                     # one 256-bit comparison = 256 int8 multiply-adds on the tensor pipe
    hamming = {"metric": "bf_top2_gcmp_per_s", "value": gcmp, "unit": "Gcmp/s", "ms_per_step": ms_match / Km, "steps": Km,
               "workload": f"{NQ} queries x {DB_ROWS} rows per GPU, top-2 + merge" + (" + NCCL all-gather" if world > 1 else ""),
               "gpu_launches": int(match_launches),
               "kernel": "bf_top2_tcn_kernel: hamming = |q| - sum(+-1 query bits x 0/1 database bits), tcgen05.mma kind::i8 128x256x32, warp-specialised, "
                         "accumulators in TMEM, threshold-filtered top-2 update",
               "roofline": {"bound": "tensor", "achieved": tops, "peak": tc_peak, "unit": "int8 TOP/s", "frac": tops / tc_peak,
                            "peak_source": "measured on this GPU: orbb200_measure_tcgen05_i8_peak = tcgen05.mma kind::i8 128x256x32 back to back from "
                                           "shared-memory operands into TMEM, one CTA per SM, no producers or consumers (nominal dense int8 4500 TOP/s)",
                            "inferred_peak_4x_legacy_imma": 4 * imma_peak, "frac_of_inferred": tops / (4 * imma_peak),
                            "legacy_imma_peak": imma_peak,
                            "popc_equivalent": {"achieved": 8 * gcmp / world, "peak": popc_peak, "unit": "Gpopc/s", "frac": 8 * gcmp / world / popc_peak,
                                                "note": "what the integer-pipe formulation (8 POPC.32 per comparison) would need: the measured "
                                                        "register-only __popc rate caps that design at frac <= 1.6 (carry-save adders, 5 POPC)"}}}

    # the same scan on REAL descriptors: rows = the descriptors this rank just extracted (512 frames x ~1006), repeated to 10 M rows with
    # 0..8 random bit flips per copy (near-duplicates, as a keyframe database holds); queries = descriptors of frames never inserted.
    # The threshold filter of the scan kernels is data dependent: uniform random rows (above) are its best case.
    with torch.cuda.stream(stream):
        valid = torch.arange(CAPACITY, device=dev)[None, :] < cnt[:, None]
        real = desc[valid]                                              # (n_real, 32) uint8
        n_real = real.shape[0]
        reps = (DB_ROWS + n_real - 1) // n_real
        gq = torch.Generator(device=dev); gq.manual_seed(4321 + rank)
        dbr = real.repeat(reps, 1)[:DB_ROWS].contiguous()
        for _ in range(8):                                              # up to 8 flipped bits per row
            byte = torch.randint(0, 32, (DB_ROWS,), device=dev, generator=gq)
            bit = (torch.randint(0, 2, (DB_ROWS,), device=dev, generator=gq) << torch.randint(0, 8, (DB_ROWS,), device=dev, generator=gq)).to(torch.uint8)
            dbr[torch.arange(DB_ROWS, device=dev), byte] ^= bit
        del byte, bit
    stream.synchronize()
    D.close()
    Dr = orbb200.DescriptorDB(DB_ROWS, row_offset=0, device=local_rank)
    Dr.append_device(dbr.data_ptr(), DB_ROWS)
    qex = orbb200.ORBextractor(NFEAT, 1.2, 8, 20, 7, max_size=(W, H), max_batch=2, device=local_rank)
    qk, qdsc = qex(synth(W, H, 777_000 + rank))
    qk2, qdsc2 = qex(synth(W, H, 778_000 + rank))
    qr = torch.from_numpy(np.concatenate([qdsc, qdsc2])[:NQ].copy()).to(dev)
    del qex
    nqr = qr.shape[0]
    ms_real = 0.0
    for i in range(3 + Km):
        with torch.cuda.stream(stream):
            flush.zero_()
        ea.record(stream)
        Dr.top2_device(qr.data_ptr(), nqr, d1.data_ptr(), i1.data_ptr(), d2.data_ptr(), stream.cuda_stream)
        eb.record(stream)
        stream.synchronize()
        if i >= 3:
            ms_real += ea.elapsed_time(eb)
    ms_real = max_over_ranks(ms_real)
    rd1 = d1[:nqr].float()
    hamming["real_descriptors"] = {"value": world * nqr * DB_ROWS * Km / (ms_real * 1e-3) / 1e9, "unit": "Gcmp/s", "ms_per_step": ms_real / Km,
                                   "workload": f"{nqr} ORB descriptors of unseen frames x {DB_ROWS} rows per GPU built from {n_real} extracted descriptors "
                                               "(copies with 0..8 random bit flips)",
                                   "best_distance_mean": float(rd1.mean().item()), "best_distance_p10": float(rd1.quantile(0.1).item()),
                                   "frac_of_uniform_random": (nqr * DB_ROWS * Km / ms_real) / (NQ * DB_ROWS * Km / ms_match)}
    Dr.close()
    del Dr, dbr, flush
    torch.cuda.empty_cache()
    c5 = measure_c5_strong(rank, world, local_rank, dev, stream, barrier, max_over_ranks)
    torch.cuda.empty_cache()
    configs = None if args.no_configs else measure_configs(rank, world, local_rank, dev, barrier, max_over_ranks)
    latency = measure_latency(local_rank, dev) if (rank == 0 and not args.no_configs) else None
    barrier()

    # ---------------- CPU baseline (rank 0, N = 1 only): bounded sample on the host cores ----------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        nthreads = os.cpu_count() or 1
        fps, kind, n, dt = cpu_reference_extract(10.0, nthreads)
        g_cpu, dtm = cpu_reference_match(nthreads)
        cpu = {"value": fps, "unit": "frames/s", "cores": nthreads, "kind": kind,
               "sample": f"{n} frames (8 distinct 640x480 synthetic frames cycled) in {dt:.1f} s; matching {NQ} q x 200000 rows in {dtm:.1f} s",
               "primitives": "the reference's ORBextractor.cpp compiled unmodified over restated scalar OpenCV primitives (oracle/cvshim)",
               "ms_per_frame_per_thread": 1e3 * nthreads / fps, "cv2_primitives_leg": cpu_cv2_primitives(),
               "hamming_gcmp_per_s": g_cpu}

    if rank == 0:
        line = {"metric": "orb_extract_frames_per_s_640x480_1000kp", "value": value, "unit": "frames/s", "n_gpus": world,
                "steps": K, "warmup": max(Wm, 3), "ms_per_step": ms_total / K, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "u8", "data": "synthetic",
                "config": {"workload": f"C1 640x480 TUM1 ORB settings (1000 kp, 8 levels, 1.2, FAST 20/7), {B} frames/GPU/step, "
                                       f"{NDISTINCT} distinct textured frames tiled; + C5-shape Hamming loop (see hamming)",
                           "frames_per_step_per_gpu": B, "keypoints_per_frame": nkp,
                           "l2": "inputs 157 MB/step/GPU > 126 MB L2 and ~2 GB of intermediates rewritten every step",
                           "parallelism": f"frames sharded across {world} GPU(s), no collective; DB row-sharded, one NCCL all-gather"},
                "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu, "hamming": hamming,
                "configs": dict(configs or {}, C5_strong_10M_rows_total=c5), "merge_checked": c5["merge_checked"], "latency": latency}
        emit(line)
    if world > 1:
        dist.destroy_process_group()


_OUT_FD = None


def emit(line):
    """The ONE JSON line goes to the process's original stdout; everything else that libraries print (NCCL's version banner,
    warnings) was redirected to stderr at start-up."""
    data = (json.dumps(line) + "\n").encode()
    os.write(_OUT_FD if _OUT_FD is not None else 1, data)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the C2/C3/C4 rows and the single-frame latency (headline + Hamming only)")
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.gpus > 1 and world == 1:
        # convenience: re-launch under torch.distributed.run (the driver does this itself)
        import subprocess
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}", "--master-addr", "127.0.0.1",
               "--master-port", "29517", os.path.abspath(__file__)] + sys.argv[1:]
        sys.exit(subprocess.call(cmd))
    global _OUT_FD
    sys.stdout.flush()
    _OUT_FD = os.dup(1)
    os.dup2(2, 1)                       # C-level and Python-level prints of libraries -> stderr
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
