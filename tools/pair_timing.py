"""Per-call wall-clock medians of one KITTI stereo pair (1241x376, 2000 kp) through the host-pointer C ABI:
  (a) the reference-shaped sequence  orbx_extract (L) + orbx_extract (R) [one thread, back to back] + orbm_stereo_match + orbm_frame_create +
      orbm_search_by_projection_frame + orbm_frame_destroy, and
  (b) the fused constructor          orbm_stereo_frame_create + orbm_search_by_projection_frame (mvuRight of the frame) + orbm_frame_destroy.
Prints one JSON object (microseconds)."""
import ctypes as C, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "orb-slam2-annotation_b200")]
import numpy as np, torch, orbb200
from synth import synth
w, h, nf = 1241, 376, 2000
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 200
big = synth(w + 80, h + 16, 8100)
pairs = [(big[oy:oy + h, ox + 40:ox + 40 + w].copy(), big[oy:oy + h, ox + 28:ox + 28 + w].copy()) for (ox, oy) in ((8, 8), (11, 9))]
L = orbb200.lib(); vp = C.c_void_p
ptr = lambda a: a.ctypes.data_as(vp)
exL = orbb200.ORBextractor(nf, max_size=(w, h)); exR = orbb200.ORBextractor(nf, max_size=(w, h)); ex2 = orbb200.ORBextractor(nf, max_size=(w, h), max_batch=2)
cap = exL.default_capacity
scale = exL.GetScaleFactors()
b = dict(kL=np.zeros(cap, orbb200.KP_DTYPE), dL=np.zeros((cap, 32), np.uint8), kR=np.zeros(cap, orbb200.KP_DTYPE), dR=np.zeros((cap, 32), np.uint8),
         ur=np.zeros(cap, np.float32), dp=np.zeros(cap, np.float32), mp=np.zeros(cap, np.int32), ob=np.zeros(cap, np.int32))
nL, nR, nm = C.c_int(), C.c_int(), C.c_int()
bounds = [C.c_float(x) for x in (0.0, 0.0, float(w), float(h))]
mbf = C.c_float(386.1448)


def queries(n, dx, dy):
    q = np.zeros(n, orbb200.PROJ_QUERY_DTYPE); k = b["kL"][:n]
    q["q"]["u"] = k["x"] + dx; q["q"]["v"] = k["y"] + dy; q["q"]["r"] = 15.0 * scale[k["octave"]]
    q["q"]["min_level"] = k["octave"] - 1; q["q"]["max_level"] = k["octave"] + 1; q["q"]["desc"] = b["dL"][:n]
    q["u_right"] = q["q"]["u"] - 12.0; q["max_err_right"] = 15.0 * scale[k["octave"]]
    q["id"] = np.arange(n); q["has_obs"] = 1; q["angle"] = k["angle"]
    return q


T = {}
def tick(name, t0):
    t1 = time.perf_counter(); T.setdefault(name, []).append((t1 - t0) * 1e6); return t1


def old_pair(i, q):
    left, right = pairs[i & 1]
    t = t00 = time.perf_counter()
    assert L.orbx_extract(exL.h, ptr(left), w, h, w, ptr(b["kL"]), ptr(b["dL"]), cap, C.byref(nL)) == 0; t = tick("a_extract_left", t)
    assert L.orbx_extract(exR.h, ptr(right), w, h, w, ptr(b["kR"]), ptr(b["dR"]), cap, C.byref(nR)) == 0; t = tick("a_extract_right", t)
    assert L.orbm_stereo_match(exL.h, exR.h, ptr(b["kL"]), ptr(b["dL"]), nL.value, ptr(b["kR"]), ptr(b["dR"]), nR.value, mbf, C.c_float(0.0), ptr(b["ur"]), ptr(b["dp"])) == 0
    t = tick("a_stereo_match", t)
    fh = vp()
    assert L.orbm_frame_create(C.byref(fh), ptr(b["kL"]), ptr(b["dL"]), nL.value, *bounds, 0) == 0; t = tick("a_frame_create", t)
    if q is not None:
        b["mp"][:nL.value] = -1; b["ob"][:nL.value] = 0
        assert L.orbm_search_by_projection_frame(fh, ptr(b["ur"]), ptr(q), len(q), 1, ptr(b["mp"]), ptr(b["ob"]), C.byref(nm)) == 0; t = tick("a_search_frame", t)
    L.orbm_frame_destroy(fh); t = tick("a_frame_destroy", t)
    tick("a_pair", t00)


pinned_pairs = [tuple(torch.from_numpy(x).pin_memory().numpy() for x in pr) for pr in pairs]
use_pinned = False


def new_pair(i, q):
    left, right = (pinned_pairs if use_pinned else pairs)[i & 1]
    t = t00 = time.perf_counter()
    fh = vp()
    assert L.orbm_stereo_frame_create(ex2.h, ptr(left), ptr(right), w, h, w, mbf, C.c_float(0.0), *bounds, C.byref(fh), ptr(b["kL"]), ptr(b["dL"]), C.byref(nL),
                                      ptr(b["kR"]), ptr(b["dR"]), C.byref(nR), cap, ptr(b["ur"]), ptr(b["dp"])) == 0
    t = tick("b_stereo_frame_create", t)
    if q is not None:
        b["mp"][:nL.value] = -1; b["ob"][:nL.value] = 0
        assert L.orbm_search_by_projection_frame(fh, orbb200.URIGHT_OF_FRAME, ptr(q), len(q), 1, ptr(b["mp"]), ptr(b["ob"]), C.byref(nm)) == 0; t = tick("b_search_frame", t)
    L.orbm_frame_destroy(fh); t = tick("b_frame_destroy", t)
    tick("b_pair", t00)


with np.errstate(divide="ignore"):
    old_pair(1, None); q0 = queries(nL.value, 3.0, 1.0)       # queries for pair 0 come from pair 1 and vice versa
    old_pair(0, None); q1 = queries(nL.value, -3.0, -1.0)
    qs = [q0, q1]
    res = {}
    for fn in (old_pair, new_pair):
        for i in range(10):
            fn(i, qs[i & 1])
        T.clear()
        for i in range(reps):
            fn(i, qs[i & 1])
        res.update({k: round(float(np.median(v)), 1) for k, v in T.items()})
        res[fn.__name__ + "_matches"] = int(nm.value)
    use_pinned = True                                         # (c) the fused constructor on page-locked images
    for i in range(10):
        new_pair(i, qs[i & 1])
    T.clear()
    for i in range(reps):
        new_pair(i, qs[i & 1])
    res.update({"c_pinned_images_" + k[2:]: round(float(np.median(v)), 1) for k, v in T.items()})
print(json.dumps(res))
