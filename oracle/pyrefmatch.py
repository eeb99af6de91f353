"""ctypes binding of oracle/_ref/liborb_refmatch.so: the reference's own ORBmatcher / Frame / KeyFrame / MapPoint / DBoW2 code
compiled unmodified (oracle/Makefile `refmatch`, harness oracle/refmatch_capi.cpp).

TEST INFRASTRUCTURE: used by tests/golden/make_golden_match.py to generate the committed fixtures, and by tests that re-check the
fixtures where /root/reference exists.  Nothing on the GPU box needs it (the fixtures travel instead)."""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])
_lib = None
VP = C.c_void_p


def available():
    return os.path.isdir("/root/reference/ORB-SLAM2/src") or os.path.exists(os.path.join(_HERE, "_ref", "liborb_refmatch.so"))


def build(force=False):
    so = os.path.join(_HERE, "_ref", "liborb_refmatch.so")
    if os.path.isdir("/root/reference/ORB-SLAM2/src"):
        srcs = [os.path.join(_HERE, "refmatch_capi.cpp"), os.path.join(_HERE, "cv_prims.hpp"), os.path.join(_HERE, "Makefile")]
        srcs += [os.path.join(dp, f) for d in ("cvshim", "refshim") for dp, _, fs in os.walk(os.path.join(_HERE, d)) for f in fs]
        if force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
            subprocess.run(["make", "-C", _HERE, "refmatch"], check=True, capture_output=True)
    return so


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(build())
        for name in ("rm_vocab_load", "rm_extractor_create", "rm_frame_mono", "rm_frame_stereo", "rm_frame_copy", "rm_map_create",
                     "rm_kf_create", "rm_mappoints_create"):
            getattr(L, name).restype = VP
        L.rm_vocab_load.argtypes = [C.c_char_p]
        L.rm_vocab_size.argtypes = [VP]
        L.rm_vocab_transform.argtypes = [VP, VP, C.c_int, C.c_int, VP, VP, VP]
        L.rm_vocab_bow.argtypes = [VP, VP, C.c_int, C.c_int] + [VP] * 7
        L.rm_extractor_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        L.rm_extractor_level.argtypes = [VP, C.c_int, VP, VP, VP]
        L.rm_frame_mono.argtypes = [VP, C.c_int, C.c_int, VP, VP, VP, VP, C.c_int, C.c_float, C.c_float]
        L.rm_frame_stereo.argtypes = [VP, VP, C.c_int, C.c_int, VP, VP, VP, VP, VP, C.c_int, C.c_float, C.c_float]
        L.rm_frame_copy.argtypes = [VP]
        L.rm_frame_n.argtypes = [VP]
        L.rm_frame_n_right.argtypes = [VP]
        L.rm_frame_get.argtypes = [VP] * 8
        L.rm_frame_statics.argtypes = [VP]
        L.rm_frame_scales.argtypes = [VP] * 7
        L.rm_frame_set_pose.argtypes = [VP, VP]
        L.rm_frame_features_in_area.argtypes = [VP, C.c_float, C.c_float, C.c_float, C.c_int, C.c_int, VP, C.c_int]
        L.rm_frame_grid.argtypes = [VP, VP, VP]
        L.rm_frame_compute_bow.argtypes = [VP] * 8
        L.rm_kf_create.argtypes = [VP, VP]
        L.rm_kf_set_pose.argtypes = [VP, VP]
        L.rm_kf_n.argtypes = [VP]
        L.rm_kf_features_in_area.argtypes = [VP, C.c_float, C.c_float, C.c_float, VP, C.c_int]
        L.rm_kf_compute_bow.argtypes = [VP] * 8
        L.rm_kf_pose.argtypes = [VP] * 4
        L.rm_mappoints_create.argtypes = [VP, VP, C.c_int] + [VP] * 6
        L.rm_mappoints_n.argtypes = [VP]
        L.rm_mappoints_ptrs.argtypes = [VP, VP]
        L.rm_mappoints_set_track.argtypes = [VP] * 7
        L.rm_mappoints_in_frustum.argtypes = [VP, VP, C.c_float] + [VP] * 6
        L.rm_mappoints_set_bad.argtypes = [VP, C.c_int]
        L.rm_mappoints_add_observation.argtypes = [VP, C.c_int, VP, C.c_int]
        L.rm_mappoints_nobs.argtypes = [VP, VP]
        L.rm_mappoints_is_bad.argtypes = [VP, VP]
        L.rm_mappoints_predict_scale.argtypes = [VP] * 5
        L.rm_mappoints_distinctive.argtypes = [VP, C.c_int, VP]
        L.rm_mappoints_observations.argtypes = [VP, C.c_int, VP, VP, C.c_int]
        L.rm_frame_set_mappoints.argtypes = [VP] * 4
        L.rm_frame_get_mappoints.argtypes = [VP] * 3
        L.rm_kf_set_mappoints.argtypes = [VP, VP, VP, C.c_int]
        L.rm_kf_get_mappoints.argtypes = [VP] * 3
        L.rm_trace.argtypes = [C.c_int]
        L.rm_trace_get.argtypes = [VP] * 4
        L.rm_search_local_points.argtypes = [VP, VP, C.c_float, C.c_float]
        L.rm_search_last_frame.argtypes = [VP, VP, C.c_float, C.c_int, C.c_float, C.c_int]
        L.rm_search_reloc.argtypes = [VP, VP, VP, VP, C.c_int, C.c_float, C.c_int, C.c_float, C.c_int]
        L.rm_search_loop.argtypes = [VP, VP, VP, VP, C.c_int, C.c_float]
        L.rm_search_by_bow_kf_frame.argtypes = [VP, VP, VP, VP, C.c_float, C.c_int]
        L.rm_search_by_bow_kf_kf.argtypes = [VP, VP, VP, VP, C.c_float, C.c_int]
        L.rm_search_for_initialization.argtypes = [VP, VP, VP, VP, C.c_int, C.c_float, C.c_int]
        L.rm_search_for_triangulation.argtypes = [VP, VP, VP, C.c_int, VP, C.c_int, C.c_float, C.c_int]
        L.rm_search_by_sim3.argtypes = [VP, VP, VP, VP, C.c_float, VP, VP, C.c_float, C.c_float]
        L.rm_fuse.argtypes = [VP, VP, C.c_float, C.c_float]
        L.rm_fuse_sim3.argtypes = [VP, VP, VP, VP, C.c_float, VP, C.c_float]
        L.rm_descriptor_distance.argtypes = [VP, VP, C.c_int, VP]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(VP) if a is not None else None


def _f32(a):
    return np.ascontiguousarray(a, np.float32)


def reset():
    lib().rm_reset()


class Trace:
    """with Trace() as t: ... ; t.events -> structured array (kind 0 Frame::GetFeaturesInArea, 1 KeyFrame::GetFeaturesInArea,
    2 KeyFrame::GetMapPoint, 3 MapPoint::GetWorldPos, 4 MapPoint::isBad, 5 MapPoint::PredictScale: x = dist, a = level) of the calls ORBmatcher.cpp made, in order."""
    DT = np.dtype([("kind", "<i4"), ("x", "<f4"), ("y", "<f4"), ("r", "<f4"), ("a", "<i4"), ("b", "<i4"), ("obj", "<u8")])

    def __enter__(self):
        lib().rm_trace(1)
        return self

    def __exit__(self, *a):
        L = lib()
        n = L.rm_trace_count()
        kind = np.zeros(n, np.int32); xyr = np.zeros((n, 3), np.float32); ab = np.zeros((n, 2), np.int32); obj = np.zeros(n, np.uint64)
        if n:
            L.rm_trace_get(_p(kind), _p(xyr), _p(ab), _p(obj))
        L.rm_trace(0)
        ev = np.zeros(n, self.DT)
        ev["kind"] = kind; ev["x"] = xyr[:, 0]; ev["y"] = xyr[:, 1]; ev["r"] = xyr[:, 2]; ev["a"] = ab[:, 0]; ev["b"] = ab[:, 1]; ev["obj"] = obj
        self.events = ev


def _bow_call(fn, first, n):
    ids = np.zeros(max(n, 1), np.int32); vals = np.zeros(max(n, 1), np.float64); nb = np.zeros(1, np.int32)
    fnodes = np.zeros(max(n, 1), np.int32); fo = np.zeros(n + 2, np.int32); ff = np.zeros(max(n, 1), np.int32); nf = np.zeros(1, np.int32)
    fn(*first, _p(ids), _p(vals), _p(nb), _p(fnodes), _p(fo), _p(ff), _p(nf))
    return ids[:nb[0]].copy(), vals[:nb[0]].copy(), fnodes[:nf[0]].copy(), fo[:nf[0] + 1].copy(), ff[:fo[nf[0]]].copy()


class Vocabulary:
    """ORBVocabulary loaded with the reference's loadFromTextFile (TemplatedVocabulary.h:1362-1448)."""

    def __init__(self, path):
        self.h = lib().rm_vocab_load(path.encode())
        if not self.h:
            raise RuntimeError("loadFromTextFile failed: " + path)

    def size(self):
        return lib().rm_vocab_size(self.h)

    def transform(self, desc, levelsup=4):
        desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        n = len(desc)
        wid = np.zeros(n, np.int32); w = np.zeros(n, np.float64); nid = np.zeros(n, np.int32)
        lib().rm_vocab_transform(self.h, _p(desc), n, levelsup, _p(wid), _p(w), _p(nid))
        return wid, w, nid

    def bow(self, desc, levelsup=4):
        desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        return _bow_call(lib().rm_vocab_bow, (self.h, _p(desc), len(desc), levelsup), len(desc))


class Extractor:
    def __init__(self, nfeatures=1000, scale=1.2, nlevels=8, ini_th=20, min_th=7):
        self.h = lib().rm_extractor_create(nfeatures, scale, nlevels, ini_th, min_th)
        self.nlevels = nlevels

    def level(self, lv):
        w, h = C.c_int(), C.c_int()
        lib().rm_extractor_level(self.h, lv, None, C.byref(w), C.byref(h))
        out = np.zeros((h.value, w.value), np.uint8)
        lib().rm_extractor_level(self.h, lv, _p(out), C.byref(w), C.byref(h))
        return out


class Frame:
    """ORB_SLAM2::Frame built by the reference's own constructors."""

    def __init__(self, h):
        self.h = h
        L = lib()
        self.N = L.rm_frame_n(h)
        nr = L.rm_frame_n_right(h)
        self.kps = np.zeros(self.N, KP_DTYPE); self.kps_un = np.zeros(self.N, KP_DTYPE); self.desc = np.zeros((self.N, 32), np.uint8)
        self.u_right = np.full(self.N, -1, np.float32); self.depth = np.full(self.N, -1, np.float32)
        self.kps_right = np.zeros(nr, KP_DTYPE); self.desc_right = np.zeros((nr, 32), np.uint8)
        L.rm_frame_get(h, _p(self.kps), _p(self.kps_un), _p(self.desc), _p(self.u_right), _p(self.depth), _p(self.kps_right), _p(self.desc_right))
        sc, isc, s2, is2 = (np.zeros(16, np.float32) for _ in range(4))
        misc = np.zeros(5, np.float32); nl = np.zeros(1, np.int32)
        L.rm_frame_scales(h, _p(sc), _p(isc), _p(s2), _p(is2), _p(misc), _p(nl))
        n = int(nl[0])
        self.nlevels = n
        self.scale, self.inv_scale, self.sigma2, self.inv_sigma2 = sc[:n].copy(), isc[:n].copy(), s2[:n].copy(), is2[:n].copy()
        self.mbf, self.mb, self.th_depth, self.scale_factor, self.log_scale_factor = [np.float32(v) for v in misc]

    @staticmethod
    def mono(img, ex, voc, K4, dist, bf=40.0, th_depth=35.0):
        img = np.ascontiguousarray(img, np.uint8)
        K4 = _f32(K4); dist = _f32(dist)
        return Frame(lib().rm_frame_mono(_p(img), img.shape[1], img.shape[0], ex.h, voc.h if voc else None, _p(K4), _p(dist), len(dist), bf, th_depth))

    @staticmethod
    def stereo(imgL, imgR, exL, exR, voc, K4, dist, bf=40.0, th_depth=35.0):
        imgL = np.ascontiguousarray(imgL, np.uint8); imgR = np.ascontiguousarray(imgR, np.uint8)
        K4 = _f32(K4); dist = _f32(dist)
        return Frame(lib().rm_frame_stereo(_p(imgL), _p(imgR), imgL.shape[1], imgL.shape[0], exL.h, exR.h, voc.h if voc else None,
                                           _p(K4), _p(dist), len(dist), bf, th_depth))

    def copy(self):
        return Frame(lib().rm_frame_copy(self.h))

    @staticmethod
    def statics():
        o = np.zeros(12, np.float32)
        lib().rm_frame_statics(_p(o))
        return dict(zip(("fx", "fy", "cx", "cy", "invfx", "invfy", "min_x", "min_y", "max_x", "max_y", "grid_w_inv", "grid_h_inv"), o))

    def set_pose(self, Tcw):
        T = _f32(Tcw).reshape(4, 4)
        lib().rm_frame_set_pose(self.h, _p(T))

    def features_in_area(self, x, y, r, min_level=-1, max_level=-1):
        out = np.zeros(max(self.N, 1), np.int32)
        n = lib().rm_frame_features_in_area(self.h, x, y, r, min_level, max_level, _p(out), self.N)
        return out[:n].copy()

    def grid(self):
        off = np.zeros(64 * 48 + 1, np.int32); idx = np.zeros(max(self.N, 1), np.int32)
        lib().rm_frame_grid(self.h, _p(off), _p(idx))
        return off, idx[:off[-1]].copy()

    def compute_bow(self):
        return _bow_call(lib().rm_frame_compute_bow, (self.h,), self.N)

    def set_mappoints(self, mps, idx, outlier=None):
        idx = np.ascontiguousarray(idx, np.int32)
        out = None if outlier is None else np.ascontiguousarray(outlier, np.uint8)
        lib().rm_frame_set_mappoints(self.h, mps.h, _p(idx), _p(out))

    def get_mappoints(self, mps):
        idx = np.zeros(self.N, np.int32)
        lib().rm_frame_get_mappoints(self.h, mps.h, _p(idx))
        return idx


class Map:
    def __init__(self):
        self.h = lib().rm_map_create()


class KeyFrame:
    def __init__(self, frame, map_):
        self.h = lib().rm_kf_create(frame.h, map_.h)
        self.N = lib().rm_kf_n(self.h)
        self.frame = frame

    def set_pose(self, Tcw):
        T = _f32(Tcw).reshape(4, 4)
        lib().rm_kf_set_pose(self.h, _p(T))

    def pose(self):
        R = np.zeros((3, 3), np.float32); t = np.zeros(3, np.float32); O = np.zeros(3, np.float32)
        lib().rm_kf_pose(self.h, _p(R), _p(t), _p(O))
        return R, t, O

    def features_in_area(self, x, y, r):
        out = np.zeros(max(self.N, 1), np.int32)
        n = lib().rm_kf_features_in_area(self.h, x, y, r, _p(out), self.N)
        return out[:n].copy()

    def compute_bow(self):
        return _bow_call(lib().rm_kf_compute_bow, (self.h,), self.N)

    def set_mappoints(self, mps, idx, add_observation=True):
        idx = np.ascontiguousarray(idx, np.int32)
        lib().rm_kf_set_mappoints(self.h, mps.h, _p(idx), int(add_observation))

    def get_mappoints(self, mps):
        idx = np.zeros(self.N, np.int32)
        lib().rm_kf_get_mappoints(self.h, mps.h, _p(idx))
        return idx


class MapPoints:
    """A vector<MapPoint*> with the protected state set from arrays."""

    def __init__(self, map_, refkf, pos, desc, normal=None, min_dist=None, max_dist=None, nobs=None):
        pos = _f32(pos).reshape(-1, 3); desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        self.n = len(pos)
        normal = None if normal is None else _f32(normal).reshape(-1, 3)
        min_dist = None if min_dist is None else _f32(min_dist); max_dist = None if max_dist is None else _f32(max_dist)
        nobs = None if nobs is None else np.ascontiguousarray(nobs, np.int32)
        self.h = lib().rm_mappoints_create(map_.h, refkf.h, self.n, _p(pos), _p(desc), _p(normal), _p(min_dist), _p(max_dist), _p(nobs))

    def ptrs(self):
        out = np.zeros(self.n, np.uint64)
        lib().rm_mappoints_ptrs(self.h, _p(out))
        return out

    def set_track(self, in_view, proj_x, proj_y, proj_xr, level, view_cos):
        a = [np.ascontiguousarray(in_view, np.uint8), _f32(proj_x), _f32(proj_y), _f32(proj_xr), np.ascontiguousarray(level, np.int32), _f32(view_cos)]
        lib().rm_mappoints_set_track(self.h, *[_p(x) for x in a])

    def in_frustum(self, frame, viewing_cos_limit=0.5):
        n = self.n
        iv = np.zeros(n, np.uint8); px = np.zeros(n, np.float32); py = np.zeros(n, np.float32); pxr = np.zeros(n, np.float32)
        lv = np.zeros(n, np.int32); vc = np.zeros(n, np.float32)
        lib().rm_mappoints_in_frustum(self.h, frame.h, viewing_cos_limit, _p(iv), _p(px), _p(py), _p(pxr), _p(lv), _p(vc))
        return iv, px, py, pxr, lv, vc

    def set_bad(self, i):
        lib().rm_mappoints_set_bad(self.h, int(i))

    def add_observation(self, i, kf, idx):
        lib().rm_mappoints_add_observation(self.h, int(i), kf.h, int(idx))

    def nobs(self):
        out = np.zeros(self.n, np.int32)
        lib().rm_mappoints_nobs(self.h, _p(out))
        return out

    def is_bad(self):
        out = np.zeros(self.n, np.uint8)
        lib().rm_mappoints_is_bad(self.h, _p(out))
        return out

    def predict_scale(self, dist, frame=None, kf=None):
        dist = _f32(dist)
        out = np.zeros(self.n, np.int32)
        lib().rm_mappoints_predict_scale(self.h, _p(dist), frame.h if frame else None, kf.h if kf else None, _p(out))
        return out

    def distinctive(self, i):
        out = np.zeros(32, np.uint8)
        lib().rm_mappoints_distinctive(self.h, int(i), _p(out))
        return out

    def observations(self, i, cap=256):
        kfs = np.zeros(cap, np.uint64); idx = np.zeros(cap, np.int32)
        n = lib().rm_mappoints_observations(self.h, int(i), _p(kfs), _p(idx), cap)
        return kfs[:n].copy(), idx[:n].copy()


# ---- the searches (return values as the reference returns them) ------------------------------------------------------------------
def search_local_points(frame, mps, th, nnratio):
    return lib().rm_search_local_points(frame.h, mps.h, th, nnratio)


def search_last_frame(cur, last, th, mono, nnratio=0.9, check_ori=True):
    return lib().rm_search_last_frame(cur.h, last.h, th, int(mono), nnratio, int(check_ori))


def search_reloc(cur, kf, mps, already, th, orb_dist, nnratio=0.9, check_ori=True):
    already = np.ascontiguousarray(already, np.int32)
    return lib().rm_search_reloc(cur.h, kf.h, mps.h, _p(already), len(already), th, orb_dist, nnratio, int(check_ori))


def search_loop(kf, Scw, mps, matched, th, nnratio=0.75):
    S = _f32(Scw).reshape(4, 4); m = np.ascontiguousarray(matched, np.int32).copy()
    n = lib().rm_search_loop(kf.h, _p(S), mps.h, _p(m), int(th), nnratio)
    return n, m


def search_by_bow_kf_frame(kf, frame, mps, nnratio=0.7, check_ori=True):
    out = np.zeros(frame.N, np.int32)
    n = lib().rm_search_by_bow_kf_frame(kf.h, frame.h, mps.h, _p(out), nnratio, int(check_ori))
    return n, out


def search_by_bow_kf_kf(kf1, kf2, mps, nnratio=0.75, check_ori=True):
    out = np.zeros(kf1.N, np.int32)
    n = lib().rm_search_by_bow_kf_kf(kf1.h, kf2.h, mps.h, _p(out), nnratio, int(check_ori))
    return n, out


def search_for_initialization(f1, f2, prev_matched, window=100, nnratio=0.9, check_ori=True):
    pm = _f32(prev_matched).reshape(-1, 2).copy(); m12 = np.zeros(f1.N, np.int32)
    n = lib().rm_search_for_initialization(f1.h, f2.h, _p(pm), _p(m12), window, nnratio, int(check_ori))
    return n, m12, pm


def search_for_triangulation(kf1, kf2, F12, only_stereo=False, nnratio=0.6, check_ori=False):
    F = _f32(F12).reshape(3, 3); cap = kf1.N
    pairs = np.zeros((cap, 2), np.int32)
    n = lib().rm_search_for_triangulation(kf1.h, kf2.h, _p(F), int(only_stereo), _p(pairs), cap, nnratio, int(check_ori))
    return n, pairs[:n].copy()


def search_by_sim3(kf1, kf2, mps, matches12, s12, R12, t12, th, nnratio=0.75):
    m = np.ascontiguousarray(matches12, np.int32).copy(); R = _f32(R12).reshape(3, 3); t = _f32(t12).reshape(3)
    n = lib().rm_search_by_sim3(kf1.h, kf2.h, mps.h, _p(m), s12, _p(R), _p(t), th, nnratio)
    return n, m


def fuse(kf, mps, th=3.0, nnratio=0.6):
    return lib().rm_fuse(kf.h, mps.h, th, nnratio)


def fuse_sim3(kf, Scw, mps, kfpoints, th, nnratio=0.8):
    S = _f32(Scw).reshape(4, 4); rep = np.zeros(mps.n, np.int32)
    n = lib().rm_fuse_sim3(kf.h, _p(S), mps.h, kfpoints.h, th, _p(rep), nnratio)
    return n, rep


def descriptor_distance(a, b):
    a = np.ascontiguousarray(a, np.uint8).reshape(-1, 32); b = np.ascontiguousarray(b, np.uint8).reshape(-1, 32)
    out = np.zeros(len(a), np.int32)
    lib().rm_descriptor_distance(_p(a), _p(b), len(a), _p(out))
    return out
