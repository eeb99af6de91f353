"""Measurement of the rows either side of the extractor: every kernel family timed device-resident with
CUDA events on its own stream (after warm-up, inputs in HBM), the CPU oracle timed on the same inputs (bounded samples, 1 thread), and for
streaming kernels the algorithmic bytes against the measured HBM peak.  Prints one JSON object; `python tools/bench_rows.py > profiles/...`.
This is a secondary table: bench.py's line stays the headline."""
import ctypes as C
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "orb-slam2-annotation_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests", "golden")]
import numpy as np
import torch
import orbb200
import pyoracle
from synth import synth, synth_vocabulary, synth_vocabulary_features

dev = torch.device("cuda:0")
L = orbb200.lib()
vp = C.c_void_p
rows = {}


def hbm_peak():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        for k in ("hbm_gbs", "hbm_gbps"):
            if k in p:
                return float(p[k]), "MEASURED_PEAKS.json:" + k
        for k, v in p.items():
            if "hbm" in k.lower() and isinstance(v, (int, float)):
                return float(v), "MEASURED_PEAKS.json:" + k
    except Exception:
        pass
    return 6539.5, "fallback (round's measured copy peak)"


HBM, HBM_SRC = hbm_peak()


def gpu_ms(fn, st, iters=20, warm=3):
    for _ in range(warm):
        fn()
    st.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(iters):
        fn()
    e1.record(st)
    st.synchronize()
    return e0.elapsed_time(e1) / iters


def cpu_s(fn, min_s=0.3):
    t0 = time.perf_counter(); n = 0
    while True:
        fn(); n += 1
        t = time.perf_counter() - t0
        if t > min_s:
            return t / n


def tdev(a):
    return torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1).copy()).to(dev)


st = torch.cuda.Stream()
S = st.cuda_stream

# ---- a1-a8 at the other config shapes (C2 KITTI, C3 EuRoC, C4 4K): device-resident batches, step shares by CUDA events ----------
def level_pixels(w, h, nl=8, sf=1.2):
    sc = np.float32(1.0); P = []
    for l in range(nl):
        inv = np.float32(1.0) / sc
        P.append(int(np.rint(np.float32(w) * inv)) * int(np.rint(np.float32(h) * inv)))
        sc = np.float32(np.float64(sc) * np.float64(np.float32(sf)))
    return P


for name, (w, h, nf, F, cap) in {"C2 KITTI 1241x376 / 2000 kp": (1241, 376, 2000, 256, 2560), "C3 EuRoC 752x480 / 1200 kp": (752, 480, 1200, 512, 1536),
                                 "C4 4K 3840x2160 / 8000 kp": (3840, 2160, 8000, 64, 9216)}.items():
    nd = 4 if w > 2000 else 16
    distinct = np.stack([synth(w, h, 5000 + i) for i in range(nd)])
    frames = torch.from_numpy(np.ascontiguousarray(np.tile(distinct, (F // nd, 1, 1)))).to(dev)
    ex = orbb200.ORBextractor(nf, 1.2, 8, 20, 7, max_size=(w, h), max_batch=F)
    est = torch.cuda.ExternalStream(ex.stream(), device=dev)
    kps = torch.zeros((F, cap, 7), dtype=torch.float32, device=dev); desc = torch.zeros((F, cap, 32), dtype=torch.uint8, device=dev)
    cnt = torch.zeros(F, dtype=torch.int32, device=dev)
    step = lambda: ex.extract_batch_device(frames.data_ptr(), F, w, h, w, w * h, kps.data_ptr(), desc.data_ptr(), cnt.data_ptr(), cap)
    ms = gpu_ms(step, est, iters=10)
    ex.sync()
    nkp = float(cnt.float().mean().item())
    ex.set_profiling(True)
    for _ in range(3):
        step()
    prof = ex.get_profile(); ex.set_profiling(False)
    P = level_pixels(w, h)
    lvl_bytes = 3 * sum(P) - P[0] - P[-1]
    lvl_ms = prof["level_ms"] / prof["calls"]
    oe = pyoracle.RefExtractor(nf) if pyoracle.ref_lib() is not None else pyoracle.OracleExtractor(nf)
    tc = cpu_s(lambda: oe.extract(distinct[0]), 1.0)
    rows[f"extract {name} (a1-a8)"] = {"workload": f"{F} frames device-resident, {nd} distinct synthetic frames tiled", "gpu_ms": ms, "frames_per_s": F / ms * 1e3,
                                      "keypoints_per_frame": nkp, "step_share_ms": {k: prof[k] / prof["calls"] for k in ("level_ms", "octree_ms", "desc_ms")},
                                      "level_kernels_algorithmic_GBps": lvl_bytes * F / lvl_ms / 1e6, "level_kernels_hbm_frac": lvl_bytes * F / lvl_ms / 1e6 / HBM,
                                      "cpu_reference_frames_per_s_1_thread": 1 / tc, "parity": "tests/test_gpu_extract.py (same shapes)"}
    del frames, kps, desc, cnt, ex

# ---- f4: cvtColor, 512 BGR VGA frames (bigger than L2) -----------------------------------------------------------------
F, w, h = 512, 640, 480
rng = np.random.default_rng(1)
col = torch.randint(0, 256, (F, h, w, 3), dtype=torch.uint8, device=dev)
gray = torch.empty((F, h, w), dtype=torch.uint8, device=dev)
ms = gpu_ms(lambda: orbb200._check(L.orbx_cvt_gray_device(vp(col.data_ptr()), F, w, h, w * 3, w * h * 3, 3, 0, vp(gray.data_ptr()), w, w * h, vp(S)), "cvt"), st)
one = col[0].cpu().numpy()
tc = cpu_s(lambda: pyoracle.cvt_gray(one, False))
by = F * w * h * 4
rows["cvt_gray_bgr (f4)"] = {"workload": f"{F} x {w}x{h} BGR -> gray", "gpu_ms": ms, "frames_per_s": F / ms * 1e3, "algorithmic_GBps": by / ms / 1e6,
                             "hbm_frac": by / ms / 1e6 / HBM, "cpu_oracle_frames_per_s": 1 / tc, "parity": "tests/test_gpu_extract.py (cv2 golden, all 2^24 colours)"}
del col

# ---- f4: remap (stereo rectification), 256 raw 800x520 -> 752x480 ------------------------------------------------------
import make_golden as GEN
import cv2
F, sw, sh, dw, dh = 256, 752, 480, 752, 480
mx, my = GEN.rectify_maps(cv2, dw, dh)
raw = torch.randint(0, 256, (F, sh, sw), dtype=torch.uint8, device=dev)
rect = torch.empty((F, dh, dw), dtype=torch.uint8, device=dev)
mxc, myc = np.ascontiguousarray(mx, np.float32), np.ascontiguousarray(my, np.float32)
ex = orbb200.ORBextractor(1200, max_size=(dw, dh), max_batch=F)
ex.set_rectification(mxc, myc, (sw, sh))
# orbx_remap_device converts the float maps on every call (host) and synchronises: time the kernel through the handle path instead when available
t0 = time.perf_counter()
for _ in range(5):
    orbb200._check(L.orbx_remap_device(vp(raw.data_ptr()), F, sw, sh, sw, sw * sh, orbb200._p(mxc), orbb200._p(myc), dw, dh, vp(rect.data_ptr()), dw, dw * dh, vp(S)), "remap")
st.synchronize()
ms_call = (time.perf_counter() - t0) / 5 * 1e3
one = raw[0].cpu().numpy()
tc = cpu_s(lambda: pyoracle.remap(one, mxc, myc))
rows["remap_rectify (f4)"] = {"workload": f"{F} x {sw}x{sh} -> {dw}x{dh}, cv::remap INTER_LINEAR", "call_ms_incl_map_upload_and_sync": ms_call,
                              "frames_per_s_call": F / ms_call * 1e3, "cpu_oracle_frames_per_s": 1 / tc,
                              "parity": "tests/test_gpu_extract.py (cv2 golden)"}
del raw, rect, ex

# ---- f2: vocabulary transform, ORBvoc shape (k=10, L=6: 1.11 M nodes) --------------------------------------------------
k, Lv = 10, 6
parent, leaf, vdesc, wt = synth_vocabulary(k, Lv, 9)
voc = orbb200.Vocabulary(k, Lv, parent, leaf, vdesc, wt)
n = 512 * 1000
feats = synth_vocabulary_features(vdesc, leaf, 20000, 3)
fd = tdev(np.tile(feats, (n // len(feats) + 1, 1))[:n])
wid = torch.empty(n, dtype=torch.int32, device=dev); wv = torch.empty(n, dtype=torch.float64, device=dev); nid = torch.empty(n, dtype=torch.int32, device=dev)
ms = gpu_ms(lambda: orbb200._check(L.orbv_transform_device(voc.h, vp(fd.data_ptr()), n, 4, vp(wid.data_ptr()), vp(wv.data_ptr()), vp(nid.data_ptr()), vp(S)), "transform"), st, iters=5)
ov = pyoracle.OracleVocabulary(k, Lv, parent, leaf, vdesc, wt)
tc = cpu_s(lambda: ov.transform(feats, 4), 1.0)
rows["vocab_transform (f2)"] = {"workload": f"{n} descriptors, k={k} L={Lv} ({len(parent)} nodes)", "gpu_ms": ms, "Mdesc_per_s": n / ms / 1e3,
                                "frames_per_s_at_1000": n / 1000 / ms * 1e3, "node_compares_Gcmp_per_s": n * k * Lv / ms / 1e6,
                                "cpu_oracle_Mdesc_per_s": len(feats) / tc / 1e6, "parity": "tests/test_vocab.py"}
del fd, wid, wv, nid

# ---- f3: ComputeDistinctiveDescriptors, 50k map points with 2..20 observations ----------------------------------------
npnt = 50000
sizes = rng.integers(2, 21, npnt); off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
desc = rng.integers(0, 256, (int(off[-1]), 32), dtype=np.uint8)
dd, do = tdev(desc), tdev(off)
br = torch.empty(npnt, dtype=torch.int32, device=dev); bm = torch.empty(npnt, dtype=torch.int32, device=dev)
ms = gpu_ms(lambda: orbb200._check(L.orbm_distinctive_descriptors_device(vp(dd.data_ptr()), vp(do.data_ptr()), npnt, vp(br.data_ptr()), vp(bm.data_ptr()), vp(S)), "dist"), st)
sub = 5000
tc = cpu_s(lambda: pyoracle.distinctive_descriptors(desc[:off[sub]], off[:sub + 1]), 0.5)
pairs = float((sizes.astype(np.int64) ** 2).sum())
rows["distinctive_descriptors (f3)"] = {"workload": f"{npnt} map points x 2..20 observations ({int(off[-1])} rows)", "gpu_ms": ms, "Mpoints_per_s": npnt / ms / 1e3,
                                        "pair_Gcmp_per_s": pairs / ms / 1e6, "cpu_oracle_Mpoints_per_s": sub / tc / 1e6, "parity": "tests/test_gpu_match.py"}

# ---- f1: tracked frame on the device: undistort + grid + SearchByProjection(last frame) ------------------------------
w, h, nf = 752, 480, 1200
K4 = [458.654, 457.296, 367.215, 248.375]; D = [-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05]
big = synth(w + 16, h + 16, 31)
cur = big[8:8 + h, 8:8 + w].copy(); last = big[9:9 + h, 11:11 + w].copy()
oex = pyoracle.OracleExtractor(nf)
kc, dc = oex.extract(cur); kl, dl = oex.extract(last)
scale = oex.tables()["scale"]
klu = pyoracle.undistort_keypoints(kl, K4, D)
bounds = pyoracle.image_bounds(w, h, K4, D)
q = np.zeros(len(kl), orbb200.PROJ_QUERY_DTYPE)
q["q"]["u"] = klu["x"] + 3; q["q"]["v"] = klu["y"] + 1; q["q"]["r"] = 15.0 * scale[kl["octave"]]
q["q"]["min_level"] = kl["octave"] - 1; q["q"]["max_level"] = kl["octave"] + 1; q["q"]["desc"] = dl
q["id"] = np.arange(len(kl)); q["has_obs"] = 1; q["angle"] = kl["angle"]
t_k, t_d, t_q = tdev(kc), tdev(dc), tdev(q)
t_ku = torch.empty_like(t_k)
mp0 = np.full(len(kc), -1, np.int32)
t_mp0 = tdev(mp0); t_mp = torch.empty_like(t_mp0); t_ob = torch.zeros(len(kc), dtype=torch.int32, device=dev); t_sc = torch.zeros(4, dtype=torch.int32, device=dev)
m = orbb200.ORBmatcher(0.9, True)
ms_un = gpu_ms(lambda: orbb200.undistort_keypoints_device(t_k.data_ptr(), len(kc), K4, D, t_ku.data_ptr(), 0, S), st)
with torch.cuda.stream(st):
    Fm = orbb200.FrameMatcher.from_device(t_ku.data_ptr(), t_d.data_ptr(), len(kc), bounds, 0, S)


def track():
    with torch.cuda.stream(st):
        t_mp.copy_(t_mp0, non_blocking=True)
    m.SearchByProjection_device("frame", Fm, t_q.data_ptr(), len(q), t_mp.data_ptr(), t_ob.data_ptr(), t_sc.data_ptr(), t_sc.data_ptr() + 4, stream=S)


ms_tr = gpu_ms(track, st, iters=50)


def grid():
    with torch.cuda.stream(st):
        f = orbb200.FrameMatcher.from_device(t_ku.data_ptr(), t_d.data_ptr(), len(kc), bounds, 0, S)
    return f


t0 = time.perf_counter(); fs = [grid() for _ in range(20)]; st.synchronize(); ms_grid = (time.perf_counter() - t0) / 20 * 1e3
for f in fs:
    f.close()
oq = np.zeros(len(q), pyoracle.QUERY_DTYPE)
oq["u"] = q["q"]["u"]; oq["v"] = q["q"]["v"]; oq["r"] = q["q"]["r"]; oq["minLevel"] = q["q"]["min_level"]; oq["maxLevel"] = q["q"]["max_level"]
oq["desc"] = q["q"]["desc"]; oq["id"] = q["id"]; oq["hasObs"] = q["has_obs"]; oq["angle"] = q["angle"]
kcu = pyoracle.undistort_keypoints(kc, K4, D)
OF = pyoracle.OracleFrame(kcu, dc, bounds)
ob0 = np.zeros(len(kc), np.int32)
tc = cpu_s(lambda: OF.search_last_frame(oq, True, mp0, ob0, None))
tcu = cpu_s(lambda: pyoracle.undistort_keypoints(kc, K4, D))
tcg = cpu_s(lambda: pyoracle.OracleFrame(kcu, dc, bounds))
rows["tracked_frame (f1, a11, a15)"] = {"workload": f"{len(kc)} keypoints, {len(q)} projected points, th=15, {w}x{h} EuRoC calibration",
                                        "undistort_gpu_us": ms_un * 1e3, "grid_create_call_us (incl. 6 cudaMalloc)": ms_grid * 1e3,
                                        "search_resolve_gpu_us (4 launches, no host sync)": ms_tr * 1e3, "matches": int(t_sc.cpu()[0]),
                                        "cpu_oracle_us": {"undistort": tcu * 1e6, "grid": tcg * 1e6, "search": tc * 1e6},
                                        "note": "single-frame latency: launch-bound; frames on different streams overlap", "parity": "tests/test_gpu_match.py"}

# ---- a14: stereo matching C2 --------------------------------------------------------------------------------------------
w, h, nf = 1241, 376, 2000
from synth import synth_stereo
imL, imR = synth_stereo(w, h, 2)
exL = orbb200.ORBextractor(nf, max_size=(w, h)); exR = orbb200.ORBextractor(nf, max_size=(w, h))
kL, dL = exL(imL); kR, dR = exR(imR)
orbb200.stereo_match(exL, exR, kL, dL, kR, dR, 386.1448, 0.537)
t0 = time.perf_counter()
for _ in range(20):
    orbb200.stereo_match(exL, exR, kL, dL, kR, dR, 386.1448, 0.537)
ms_st = (time.perf_counter() - t0) / 20 * 1e3
oL = pyoracle.OracleExtractor(nf); oR = pyoracle.OracleExtractor(nf)
oL.extract(imL); oR.extract(imR)
tc = cpu_s(lambda: pyoracle.stereo_match(oL, oR, 386.1448, 0.537))
rows["stereo_match (a14)"] = {"workload": f"C2 {w}x{h}, {len(kL)} x {len(kR)} keypoints", "call_ms_host_pointers": ms_st, "pairs_per_s_call": 1e3 / ms_st,
                              "cpu_oracle_ms": tc * 1e3, "parity": "tests/test_gpu_match.py::test_stereo_matches"}

print(json.dumps({"hbm_peak_GBps": HBM, "hbm_peak_source": HBM_SRC, "device": torch.cuda.get_device_name(0), "rows": rows}, indent=1))
