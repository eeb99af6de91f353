"""ctypes binding of the CPU oracle (oracle/liborb_oracle.so) and of oracle/_ref/liborb_ref.so.

TEST INFRASTRUCTURE: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module.  The product (orb-slam2-annotation_b200/) never does.
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])
assert KP_DTYPE.itemsize == 28

QUERY_DTYPE = np.dtype([("u", "<f4"), ("v", "<f4"), ("r", "<f4"), ("minLevel", "<i4"), ("maxLevel", "<i4"),
                        ("uR", "<f4"), ("maxErrR", "<f4"), ("desc", "u1", (32,)), ("id", "<i4"),
                        ("hasObs", "<i4"), ("angle", "<f4")])
assert QUERY_DTYPE.itemsize == 72


def build(force=False):
    so = os.path.join(_HERE, "liborb_oracle.so")
    srcs = [os.path.join(_HERE, f) for f in ("oracle_capi.cpp", "cv_prims.hpp", "orb_extract_oracle.hpp",
                                              "orb_match_oracle.hpp", "dbow_oracle.hpp", "orb_pattern.inc", "Makefile")]
    stale = force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs)
    if stale:
        subprocess.run(["make", "-C", _HERE, "liborb_oracle.so"], check=True, capture_output=True)
    if os.path.isdir("/root/reference/ORB-SLAM2/src"):
        ref = os.path.join(_HERE, "_ref", "liborb_ref.so")
        rsrcs = [os.path.join(_HERE, "ref_capi.cpp")] + [os.path.join(dp, f) for dp, _, fs in os.walk(os.path.join(_HERE, "cvshim")) for f in fs]
        rsrcs = [s for s in rsrcs if os.path.exists(s)]
        if rsrcs and (force or not os.path.exists(ref) or any(os.path.getmtime(s) > os.path.getmtime(ref) for s in rsrcs)):
            subprocess.run(["make", "-C", _HERE, "ref"], check=True, capture_output=True)
        import pyrefmatch                       # the matcher half of the reference (oracle/_ref/liborb_refmatch.so)
        pyrefmatch.build(force)
    return so


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        L = _lib
        L.orc_create.restype = C.c_void_p
        L.orc_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]
        L.orc_destroy.argtypes = [C.c_void_p]
        L.orc_extract.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_long]
        L.orc_get_result.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        L.orc_tables.argtypes = [C.c_void_p] + [C.c_void_p] * 6
        L.orc_level_size.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.orc_level_image.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.orc_level_raw_count.argtypes = [C.c_void_p, C.c_int]
        L.orc_level_raw.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.orc_level_kp_count.argtypes = [C.c_void_p, C.c_int]
        L.orc_level_kps.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.orc_resize.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int]
        L.orc_blur.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.orc_remap.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int]
        L.orc_cvt_gray.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]
        L.orc_fast.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.orc_fast_atan2.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        L.orc_distance_pairs.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.orc_distinctive.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.orc_bf_top2.argtypes = [C.c_void_p, C.c_long, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        L.orc_vocab_create.restype = C.c_void_p
        L.orc_vocab_create.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int]
        L.orc_vocab_destroy.argtypes = [C.c_void_p]
        L.orc_vocab_transform.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_vocab_bow.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int] + [C.c_void_p] * 7
        L.orc_search_reloc.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]
        L.orc_search_loop.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.orc_fuse_best.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.orc_search_by_bow.argtypes = [C.c_void_p] * 6 + [C.c_int] + [C.c_void_p] * 2 + [C.c_int] + [C.c_void_p] * 3 + [C.c_int, C.c_float, C.c_int, C.c_void_p]
        L.orc_search_by_bow_kf.argtypes = ([C.c_void_p] * 3 + [C.c_int] + [C.c_void_p] * 3 + [C.c_int]) * 2 + [C.c_float, C.c_int, C.c_void_p]
        L.orc_search_for_triangulation.argtypes = ([C.c_void_p] * 4 + [C.c_int] + [C.c_void_p] * 3 + [C.c_int]) * 2 + [C.c_void_p, C.c_float, C.c_float, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.orc_frame_create.restype = C.c_void_p
        L.orc_frame_create.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float]
        L.orc_frame_destroy.argtypes = [C.c_void_p]
        L.orc_features_in_area.argtypes = [C.c_void_p, C.c_float, C.c_float, C.c_float, C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.orc_window_search.restype = C.c_long
        L.orc_window_search.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_long]
        L.orc_search_local_points.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_void_p, C.c_void_p]
        L.orc_search_last_frame.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.orc_search_for_initialization.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_int]
        L.orc_stereo_match.argtypes = [C.c_void_p, C.c_void_p, C.c_float, C.c_float] + [C.c_void_p] * 4
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class OracleExtractor:
    """CPU oracle of ORBextractor (reference ORBextractor.cpp); keeps every intermediate stage."""

    def __init__(self, nfeatures=1000, scale=1.2, nlevels=8, ini_th=20, min_th=7, tie_mode=0, legacy_blur=False):
        self.L = lib()
        self.nlevels = nlevels
        self.h = self.L.orc_create(nfeatures, scale, nlevels, ini_th, min_th, tie_mode, int(legacy_blur))

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orc_destroy(self.h)
            self.h = None

    def extract(self, img):
        img = np.ascontiguousarray(img, np.uint8)
        h, w = img.shape
        n = self.L.orc_extract(self.h, _p(img), w, h, w)
        kps = np.zeros(n, KP_DTYPE)
        desc = np.zeros((n, 32), np.uint8)
        if n:
            self.L.orc_get_result(self.h, _p(kps), _p(desc), n)
        return kps, desc

    def tables(self):
        n = self.nlevels
        sc, isc, s2, is2 = (np.zeros(n, np.float32) for _ in range(4))
        per = np.zeros(n, np.int32)
        um = np.zeros(16, np.int32)
        self.L.orc_tables(self.h, _p(sc), _p(isc), _p(s2), _p(is2), _p(per), _p(um))
        return dict(scale=sc, inv_scale=isc, sigma2=s2, inv_sigma2=is2, per_level=per, umax=um)

    def level_size(self, level):
        w, h = C.c_int(), C.c_int()
        self.L.orc_level_size(self.h, level, C.byref(w), C.byref(h))
        return w.value, h.value

    def level_image(self, level, blurred=False):
        w, h = self.level_size(level)
        out = np.zeros((h, w), np.uint8)
        self.L.orc_level_image(self.h, level, int(blurred), _p(out))
        return out

    def level_raw(self, level):
        n = self.L.orc_level_raw_count(self.h, level)
        out = np.zeros((n, 3), np.float32)
        if n:
            self.L.orc_level_raw(self.h, level, _p(out))
        return out

    def level_kps(self, level):
        n = self.L.orc_level_kp_count(self.h, level)
        out = np.zeros(n, KP_DTYPE)
        if n:
            self.L.orc_level_kps(self.h, level, _p(out))
        return out


def resize(src, dw, dh):
    src = np.ascontiguousarray(src, np.uint8)
    dst = np.zeros((dh, dw), np.uint8)
    lib().orc_resize(_p(src), src.shape[1], src.shape[0], _p(dst), dw, dh)
    return dst


def remap(src, mapx, mapy):
    """cv::remap(src, dst, mapx, mapy, INTER_LINEAR) (BORDER_CONSTANT 0) of a u8 image with float32 maps."""
    src = np.ascontiguousarray(src, np.uint8)
    mapx = np.ascontiguousarray(mapx, np.float32); mapy = np.ascontiguousarray(mapy, np.float32)
    dh, dw = mapx.shape
    out = np.zeros((dh, dw), np.uint8)
    lib().orc_remap(_p(src), src.shape[1], src.shape[0], _p(mapx), _p(mapy), _p(out), dw, dh)
    return out


def cvt_gray(src, rgb_order=True):
    """cv::cvtColor RGB/BGR/RGBA/BGRA -> GRAY of an (h, w, 3|4) u8 image."""
    src = np.ascontiguousarray(src, np.uint8)
    h, w, ch = src.shape
    out = np.zeros((h, w), np.uint8)
    lib().orc_cvt_gray(_p(src), w, h, ch, 1 if rgb_order else 0, _p(out))
    return out


def blur(src, legacy=False):
    src = np.ascontiguousarray(src, np.uint8)
    dst = np.zeros_like(src)
    lib().orc_blur(_p(src), src.shape[1], src.shape[0], _p(dst), int(legacy))
    return dst


def fast(img, th):
    img = np.ascontiguousarray(img, np.uint8)
    cap = img.size
    out = np.zeros((cap, 3), np.int32)
    n = lib().orc_fast(_p(img), img.shape[1], img.shape[0], th, _p(out), cap)
    return out[:n]


def fast_atan2(y, x):
    y = np.ascontiguousarray(y, np.float32)
    x = np.ascontiguousarray(x, np.float32)
    out = np.zeros_like(y)
    lib().orc_fast_atan2(_p(y), _p(x), _p(out), y.size)
    return out


def distance_pairs(a, b):
    a = np.ascontiguousarray(a, np.uint8)
    b = np.ascontiguousarray(b, np.uint8)
    out = np.zeros(a.shape[0], np.int32)
    lib().orc_distance_pairs(_p(a), _p(b), a.shape[0], _p(out))
    return out


def distinctive_descriptors(desc, offsets):
    """MapPoint::ComputeDistinctiveDescriptors per CSR segment -> (best_row, best_median)."""
    desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    offsets = np.ascontiguousarray(offsets, np.int64)
    n = offsets.size - 1
    row = np.zeros(n, np.int32)
    med = np.zeros(n, np.int32)
    lib().orc_distinctive(_p(desc), _p(offsets), n, _p(row), _p(med))
    return row, med


def bf_top2(db, q, nthreads=1):
    db = np.ascontiguousarray(db, np.uint8)
    q = np.ascontiguousarray(q, np.uint8)
    nq = q.shape[0]
    d1 = np.zeros(nq, np.int32)
    i1 = np.zeros(nq, np.int64)
    d2 = np.zeros(nq, np.int32)
    lib().orc_bf_top2(_p(db), db.shape[0], _p(q), nq, _p(d1), _p(i1), _p(d2), nthreads)
    return d1, i1, d2


class OracleVocabulary:
    """DBoW2 TemplatedVocabulary (tree + transform) on flattened node arrays, nodes in loadFromTextFile order (oracle)."""

    def __init__(self, k, L, parent, is_leaf, desc, weight, weighting=0, norm=1):
        self.L_ = lib()
        self.parent = np.ascontiguousarray(parent, np.int32); self.is_leaf = np.ascontiguousarray(is_leaf, np.uint8)
        self.desc = np.ascontiguousarray(desc, np.uint8); self.weight = np.ascontiguousarray(weight, np.float64)
        self.h = self.L_.orc_vocab_create(k, L, len(self.parent), _p(self.parent), _p(self.is_leaf), _p(self.desc), _p(self.weight),
                                          weighting, norm)

    def __del__(self):
        if getattr(self, "h", None):
            self.L_.orc_vocab_destroy(self.h)
            self.h = None

    def transform(self, feats, levelsup=4):
        feats = np.ascontiguousarray(feats, np.uint8).reshape(-1, 32)
        n = len(feats)
        wid = np.zeros(n, np.int32); w = np.zeros(n, np.float64); nid = np.zeros(n, np.int32)
        self.L_.orc_vocab_transform(self.h, _p(feats), n, levelsup, _p(wid), _p(w), _p(nid))
        return wid, w, nid

    def bow(self, feats, levelsup=4):
        """-> (bow_ids, bow_vals, fv_nodes, fv_offsets, fv_features)"""
        feats = np.ascontiguousarray(feats, np.uint8).reshape(-1, 32)
        n = len(feats)
        ids = np.zeros(n, np.int32); vals = np.zeros(n, np.float64); nb = np.zeros(1, np.int32)
        fn = np.zeros(n, np.int32); fo = np.zeros(n + 1, np.int32); ff = np.zeros(n, np.int32); nf = np.zeros(1, np.int32)
        self.L_.orc_vocab_bow(self.h, _p(feats), n, levelsup, _p(ids), _p(vals), _p(nb), _p(fn), _p(fo), _p(ff), _p(nf))
        return ids[:nb[0]].copy(), vals[:nb[0]].copy(), fn[:nf[0]].copy(), fo[:nf[0] + 1].copy(), ff[:fo[nf[0]]].copy()


def search_by_bow(desc_kf, angle_kf, kf_has_mp, fv_kf, desc_f, angle_f, fv_f, nnratio=0.7, check_ori=True):
    """ORBmatcher::SearchByBoW(KeyFrame, Frame): fv_* = (nodes, offsets, features) -> (matches[F.N] of KF indices, nmatches)."""
    desc_kf = np.ascontiguousarray(desc_kf, np.uint8); desc_f = np.ascontiguousarray(desc_f, np.uint8)
    angle_kf = np.ascontiguousarray(angle_kf, np.float32); angle_f = np.ascontiguousarray(angle_f, np.float32)
    kf_has_mp = np.ascontiguousarray(kf_has_mp, np.uint8)
    kn, ko, kf = [np.ascontiguousarray(a, np.int32) for a in fv_kf]
    fn, fo, ff = [np.ascontiguousarray(a, np.int32) for a in fv_f]
    out = np.zeros(len(desc_f), np.int32)
    n = lib().orc_search_by_bow(_p(desc_kf), _p(angle_kf), _p(kf_has_mp), _p(kn), _p(ko), _p(kf), len(kn), _p(desc_f), _p(angle_f), len(desc_f),
                                _p(fn), _p(fo), _p(ff), len(fn), nnratio, 1 if check_ori else 0, _p(out))
    return out, n


def search_by_bow_kf(desc1, angle1, has1, fv1, desc2, angle2, has2, fv2, nnratio=0.75, check_ori=True):
    """ORBmatcher::SearchByBoW(KeyFrame, KeyFrame) -> (matches12[n1] of KF2 indices, nmatches)."""
    desc1 = np.ascontiguousarray(desc1, np.uint8); desc2 = np.ascontiguousarray(desc2, np.uint8)
    angle1 = np.ascontiguousarray(angle1, np.float32); angle2 = np.ascontiguousarray(angle2, np.float32)
    has1 = np.ascontiguousarray(has1, np.uint8); has2 = np.ascontiguousarray(has2, np.uint8)
    a = [np.ascontiguousarray(x, np.int32) for x in fv1]; b = [np.ascontiguousarray(x, np.int32) for x in fv2]
    out = np.zeros(len(desc1), np.int32)
    n = lib().orc_search_by_bow_kf(_p(desc1), _p(angle1), _p(has1), len(desc1), _p(a[0]), _p(a[1]), _p(a[2]), len(a[0]),
                                   _p(desc2), _p(angle2), _p(has2), len(desc2), _p(b[0]), _p(b[1]), _p(b[2]), len(b[0]),
                                   nnratio, 1 if check_ori else 0, _p(out))
    return out, n


def search_for_triangulation(desc1, kps1, ur1, has1, fv1, desc2, kps2, ur2, has2, fv2, F12, ex, ey, sf2, ls2, only_stereo=False, check_ori=True):
    """ORBmatcher::SearchForTriangulation -> (vMatches12[n1], nmatches)."""
    desc1 = np.ascontiguousarray(desc1, np.uint8); desc2 = np.ascontiguousarray(desc2, np.uint8)
    kps1 = np.ascontiguousarray(kps1, KP_DTYPE); kps2 = np.ascontiguousarray(kps2, KP_DTYPE)
    ur1 = None if ur1 is None else np.ascontiguousarray(ur1, np.float32); ur2 = None if ur2 is None else np.ascontiguousarray(ur2, np.float32)
    has1 = np.ascontiguousarray(has1, np.uint8); has2 = np.ascontiguousarray(has2, np.uint8)
    a = [np.ascontiguousarray(x, np.int32) for x in fv1]; b = [np.ascontiguousarray(x, np.int32) for x in fv2]
    F12 = np.ascontiguousarray(F12, np.float32); sf2 = np.ascontiguousarray(sf2, np.float32); ls2 = np.ascontiguousarray(ls2, np.float32)
    out = np.zeros(len(desc1), np.int32)
    n = lib().orc_search_for_triangulation(_p(desc1), _p(kps1), _p(ur1), _p(has1), len(desc1), _p(a[0]), _p(a[1]), _p(a[2]), len(a[0]),
                                           _p(desc2), _p(kps2), _p(ur2), _p(has2), len(desc2), _p(b[0]), _p(b[1]), _p(b[2]), len(b[0]),
                                           _p(F12), float(ex), float(ey), _p(sf2), _p(ls2), int(only_stereo), int(check_ori), _p(out))
    return out, n


def undistort_points(xy, K4, dist):
    """cv::undistortPoints(xy, K, dist, noArray(), K) on float32 (n, 2)."""
    xy = np.ascontiguousarray(xy, np.float32); K4 = np.ascontiguousarray(K4, np.float32); dist = np.ascontiguousarray(dist, np.float32)
    out = np.zeros_like(xy)
    lib().orc_undistort_points(_p(xy), len(xy), _p(K4), _p(dist), len(dist), _p(out))
    return out


def undistort_keypoints(kps, K4, dist):
    """Frame::UndistortKeyPoints -> mvKeysUn."""
    kps = np.ascontiguousarray(kps, KP_DTYPE); K4 = np.ascontiguousarray(K4, np.float32); dist = np.ascontiguousarray(dist, np.float32)
    out = np.zeros_like(kps)
    lib().orc_undistort_keypoints(_p(kps), len(kps), _p(K4), _p(dist), len(dist), _p(out))
    return out


def image_bounds(cols, rows, K4, dist):
    """Frame::ComputeImageBounds -> (mnMinX, mnMinY, mnMaxX, mnMaxY) float32."""
    K4 = np.ascontiguousarray(K4, np.float32); dist = np.ascontiguousarray(dist, np.float32)
    b = np.zeros(4, np.float32)
    lib().orc_image_bounds(int(cols), int(rows), _p(K4), _p(dist), len(dist), _p(b))
    return b


class OracleFrame:
    """Frame grid + GetFeaturesInArea + the Search* loops on flattened arrays (oracle)."""

    def __init__(self, kps, desc, bounds):
        self.L = lib()
        self.kps = np.ascontiguousarray(kps, KP_DTYPE)
        self.desc = np.ascontiguousarray(desc, np.uint8)
        self.n = len(self.kps)
        self.h = self.L.orc_frame_create(_p(self.kps), _p(self.desc), self.n, *[float(b) for b in bounds])

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orc_frame_destroy(self.h)
            self.h = None

    def features_in_area(self, x, y, r, min_level=-1, max_level=-1):
        out = np.zeros(max(self.n, 1), np.int32)
        n = self.L.orc_features_in_area(self.h, x, y, r, min_level, max_level, _p(out), self.n)
        return out[:n].copy()

    def window_search(self, queries):
        q = np.ascontiguousarray(queries, QUERY_DTYPE)
        nq = len(q)
        offs = np.zeros(nq + 1, np.int32)
        total = self.L.orc_window_search(self.h, _p(q), nq, _p(offs), None, None, 0)
        idx = np.zeros(max(total, 1), np.int32)
        dist = np.zeros(max(total, 1), np.int32)
        self.L.orc_window_search(self.h, _p(q), nq, _p(offs), _p(idx), _p(dist), total)
        return offs, idx[:total], dist[:total]

    def search_local_points(self, queries, nnratio, frame_mp, frame_obs, uright=None):
        q = np.ascontiguousarray(queries, QUERY_DTYPE)
        mp = np.ascontiguousarray(frame_mp, np.int32).copy()
        ob = np.ascontiguousarray(frame_obs, np.int32).copy()
        ur = None if uright is None else np.ascontiguousarray(uright, np.float32)
        n = self.L.orc_search_local_points(self.h, _p(ur), _p(q), len(q), nnratio, _p(mp), _p(ob))
        return n, mp, ob

    def search_last_frame(self, queries, check_ori, frame_mp, frame_obs, uright=None):
        q = np.ascontiguousarray(queries, QUERY_DTYPE)
        mp = np.ascontiguousarray(frame_mp, np.int32).copy()
        ob = np.ascontiguousarray(frame_obs, np.int32).copy()
        ur = None if uright is None else np.ascontiguousarray(uright, np.float32)
        n = self.L.orc_search_last_frame(self.h, _p(ur), _p(q), len(q), int(check_ori), _p(mp), _p(ob))
        return n, mp, ob


    def search_reloc(self, queries, orbdist, check_ori, frame_mp):
        q = np.ascontiguousarray(queries, QUERY_DTYPE)
        mp = np.ascontiguousarray(frame_mp, np.int32).copy()
        n = self.L.orc_search_reloc(self.h, _p(q), len(q), orbdist, int(check_ori), _p(mp))
        return n, mp

    def search_loop(self, queries, matched):
        q = np.ascontiguousarray(queries, QUERY_DTYPE)
        m = np.ascontiguousarray(matched, np.int32).copy()
        n = self.L.orc_search_loop(self.h, _p(q), len(q), _p(m))
        return n, m

    def fuse_best(self, queries, uright, inv_sigma2, gate, max_dist):
        q = np.ascontiguousarray(queries, QUERY_DTYPE)
        ur = None if uright is None else np.ascontiguousarray(uright, np.float32)
        s2 = None if inv_sigma2 is None else np.ascontiguousarray(inv_sigma2, np.float32)
        bi = np.zeros(len(q), np.int32); bd = np.zeros(len(q), np.int32)
        self.L.orc_fuse_best(self.h, _p(q), len(q), _p(ur), _p(s2), int(gate), max_dist, _p(bi), _p(bd))
        return bi, bd


def search_for_initialization(f1, f2, prev_matched, window=100, nnratio=0.9, check_ori=True):
    pm = np.ascontiguousarray(prev_matched, np.float32).copy()
    m12 = np.zeros(f1.n, np.int32)
    n = lib().orc_search_for_initialization(f1.h, f2.h, _p(pm), _p(m12), window, nnratio, int(check_ori))
    return n, m12, pm


def stereo_match(exL, exR, mbf, mb=0.0):
    """exL/exR: OracleExtractor objects after extract()."""
    n = lib().orc_get_result(exL.h, None, None, 0)
    ur = np.zeros(n, np.float32)
    dp = np.zeros(n, np.float32)
    bi = np.zeros(n, np.int32)
    bd = np.zeros(n, np.int32)
    lib().orc_stereo_match(exL.h, exR.h, mbf, mb, _p(ur), _p(dp), _p(bi), _p(bd))
    return ur, dp, bi, bd


# ------------------------------------------------------------------------------------------
# oracle/_ref: the reference's own ORBextractor.cpp (unmodified) over the mini-OpenCV shim.
# ------------------------------------------------------------------------------------------
_ref = None


def ref_lib():
    global _ref
    if _ref is None:
        build()
        path = os.path.join(_HERE, "_ref", "liborb_ref.so")
        if not os.path.exists(path):
            return None
        _ref = C.CDLL(path)
        _ref.ref_create.restype = C.c_void_p
        _ref.ref_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        _ref.ref_destroy.argtypes = [C.c_void_p]
        _ref.ref_extract.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_long, C.c_void_p, C.c_void_p, C.c_int]
        _ref.ref_level_size.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        _ref.ref_level_image.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
    return _ref


class RefExtractor:
    """ORB_SLAM2::ORBextractor compiled from /root/reference (oracle/_ref)."""

    def __init__(self, nfeatures=1000, scale=1.2, nlevels=8, ini_th=20, min_th=7):
        self.L = ref_lib()
        if self.L is None:
            raise RuntimeError("oracle/_ref/liborb_ref.so not built")
        self.nlevels = nlevels
        self.h = self.L.ref_create(nfeatures, scale, nlevels, ini_th, min_th)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.ref_destroy(self.h)
            self.h = None

    def extract(self, img, cap=None):
        img = np.ascontiguousarray(img, np.uint8)
        h, w = img.shape
        cap = cap or 200000
        kps = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        n = self.L.ref_extract(self.h, _p(img), w, h, w, _p(kps), _p(desc), cap)
        return kps[:n].copy(), desc[:n].copy()

    def level_image(self, level):
        w, h = C.c_int(), C.c_int()
        self.L.ref_level_size(self.h, level, C.byref(w), C.byref(h))
        out = np.zeros((h.value, w.value), np.uint8)
        self.L.ref_level_image(self.h, level, _p(out))
        return out
