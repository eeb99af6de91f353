"""ctypes harness over liborbb200.so (the C ABI of include/orbb200.h) for tests and bench.py.

The product is the CUDA library and the C++ host mirror in host/; this module is plumbing that lets pytest and the
benchmark drive the same C entry points a C++ caller binds.  It never computes anything on the CPU and raises
if the library is missing or a call fails (no fallback).
"""
import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(_HERE, "liborbb200.so")

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])
QUERY_DTYPE = np.dtype([("u", "<f4"), ("v", "<f4"), ("r", "<f4"), ("min_level", "<i4"), ("max_level", "<i4"),
                        ("desc", "u1", (32,))])
PROJ_QUERY_DTYPE = np.dtype([("q", QUERY_DTYPE), ("u_right", "<f4"), ("max_err_right", "<f4"), ("id", "<i4"),
                             ("has_obs", "<i4"), ("angle", "<f4")])
assert KP_DTYPE.itemsize == 28 and QUERY_DTYPE.itemsize == 52 and PROJ_QUERY_DTYPE.itemsize == 72

OK, E_INVALID, E_CUDA, E_CAPACITY, E_BAD_SHAPE, E_NOMEM, E_STATE = 0, -1, -2, -3, -4, -5, -6
NOALLOC_MAX_QUERIES, NOALLOC_MAX_KEYS = 4096, 1 << 19      # ORBM_NOALLOC_MAX_QUERIES / _KEYS of include/orbb200.h
STAGE_BLUR, STAGE_FASTKPS, STAGE_LEVELKPS = 1, 2, 3

# every symbol include/orbb200.h declares (tests check the .so exports all of them)
SYMBOLS = ["orbx_create", "orbx_destroy", "orbx_get_tables", "orbx_extract", "orbx_extract_batch_device",
           "orbx_extract_batch", "orbx_graph_capture", "orbx_graph_launch", "orbx_sync", "orbx_set_profiling", "orbx_get_profile", "orbx_stream", "orbx_pyramid_level", "orbx_debug_stage",
           "orbx_extract_batch_color", "orbx_cvt_gray_device", "orbx_set_rectification", "orbx_extract_batch_rectified", "orbx_remap_device", "orbm_distance_pairs", "orbm_distinctive_descriptors", "orbm_distinctive_descriptors_device", "orbm_frame_create", "orbm_frame_destroy", "orbm_window_search",
           "orbm_search_by_projection_points", "orbm_search_by_projection_frame", "orbm_search_for_initialization", "orbm_search_by_bow", "orbm_search_by_bow_kf", "orbm_search_for_triangulation", "orbm_frame_create_device", "orbm_undistort_keypoints", "orbm_undistort_keypoints_device", "orbm_image_bounds",
           "orbm_search_by_projection_points_device", "orbm_search_by_projection_frame_device", "orbm_search_by_projection_reloc", "orbm_search_by_projection_reloc_device", "orbm_search_by_projection_loop", "orbm_fuse_best",
           "orbm_stereo_match", "orbm_stereo_frame_create", "orbm_mono_frame_create", "orbm_db_create", "orbm_db_append", "orbm_db_append_device", "orbm_db_destroy",
           "orbm_db_rows", "orbm_bf_top2", "orbm_bf_top2_device", "orbm_bf_top2_noalloc", "orbm_merge_top2_device", "orbm_ratio_test",
           "orbv_vocab_create", "orbv_vocab_destroy", "orbv_vocab_words", "orbv_transform", "orbv_transform_device",
           "orbv_compute_bow", "orbb200_strerror", "orbb200_last_cuda_error", "orbb200_device_count", "orbb200_measure_popc_peak", "orbb200_measure_imma_peak", "orbb200_measure_tcgen05_i8_peak",
           "orbb200_launch_count"]


class OrbConfig(C.Structure):
    _fields_ = [("nfeatures", C.c_int32), ("scale_factor", C.c_float), ("nlevels", C.c_int32),
                ("ini_th_fast", C.c_int32), ("min_th_fast", C.c_int32), ("max_width", C.c_int32),
                ("max_height", C.c_int32), ("max_batch", C.c_int32), ("device", C.c_int32), ("blur_legacy", C.c_int32)]


class OrbError(RuntimeError):
    def __init__(self, code, where):
        self.code = code
        L = lib()
        msg = L.orbb200_strerror(code).decode()
        if code == E_CUDA:
            msg += ": " + L.orbb200_last_cuda_error().decode()
        super().__init__(f"{where}: {msg} ({code})")


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(SO_PATH):
            raise RuntimeError(f"{SO_PATH} is missing: build it with __graft_entry__.build() (there is no CPU fallback)")
        L = C.CDLL(SO_PATH)
        vp, i32, f32, sz = C.c_void_p, C.c_int, C.c_float, C.c_size_t
        L.orbx_create.argtypes = [C.POINTER(OrbConfig), C.POINTER(vp)]
        L.orbx_destroy.argtypes = [vp]
        L.orbx_get_tables.argtypes = [vp] + [vp] * 6
        L.orbx_extract.argtypes = [vp, vp, i32, i32, sz, vp, vp, i32, C.POINTER(i32)]
        L.orbx_extract_batch_device.argtypes = [vp, vp, i32, i32, i32, sz, sz, vp, vp, vp, i32]
        L.orbx_graph_capture.argtypes = [vp, vp, i32, i32, i32, sz, sz, vp, vp, vp, i32]
        L.orbx_graph_launch.argtypes = [vp]
        L.orbx_extract_batch.argtypes = [vp, vp, i32, i32, i32, sz, sz, vp, vp, vp, i32]
        L.orbx_extract_batch_color.argtypes = [vp, vp, i32, i32, i32, sz, sz, i32, i32, vp, vp, vp, i32]
        L.orbx_cvt_gray_device.argtypes = [vp, i32, i32, i32, sz, sz, i32, i32, vp, sz, sz, vp]
        L.orbx_set_rectification.argtypes = [vp, vp, vp, i32, i32, i32, i32]
        L.orbx_extract_batch_rectified.argtypes = [vp, vp, i32, sz, sz, vp, vp, vp, i32]
        L.orbx_remap_device.argtypes = [vp, i32, i32, i32, sz, sz, vp, vp, i32, i32, vp, sz, sz, vp]
        L.orbx_sync.argtypes = [vp]
        L.orbx_set_profiling.argtypes = [vp, i32]
        L.orbx_get_profile.argtypes = [vp, vp, C.POINTER(i32)]
        L.orbx_stream.argtypes = [vp]
        L.orbx_stream.restype = vp
        L.orbx_pyramid_level.argtypes = [vp, i32, i32, vp, sz, C.POINTER(i32), C.POINTER(i32)]
        L.orbx_debug_stage.argtypes = [vp, i32, i32, i32, vp, sz, C.POINTER(sz)]
        L.orbm_distance_pairs.argtypes = [vp, vp, i32, vp, i32]
        L.orbm_distinctive_descriptors.argtypes = [vp, vp, i32, vp, vp, i32]
        L.orbm_distinctive_descriptors_device.argtypes = [vp, vp, i32, vp, vp, vp]
        L.orbm_frame_create.argtypes = [C.POINTER(vp), vp, vp, i32, f32, f32, f32, f32, i32]
        L.orbv_vocab_create.argtypes = [C.POINTER(vp), i32, i32, i32, vp, vp, vp, vp, i32, i32, i32]
        L.orbv_vocab_destroy.argtypes = [vp]
        L.orbv_vocab_words.argtypes = [vp]
        L.orbv_transform.argtypes = [vp, vp, i32, i32, vp, vp, vp]
        L.orbv_transform_device.argtypes = [vp, vp, i32, i32, vp, vp, vp, vp]
        L.orbv_compute_bow.argtypes = [vp, vp, i32, i32] + [vp] * 7
        L.orbm_frame_destroy.argtypes = [vp]
        L.orbm_window_search.argtypes = [vp, vp, i32, vp, vp, vp, i32]
        L.orbm_undistort_keypoints.argtypes = [vp, i32, vp, vp, i32, vp, i32]
        L.orbm_undistort_keypoints_device.argtypes = [vp, i32, vp, vp, i32, vp, i32, vp]
        L.orbm_image_bounds.argtypes = [i32, i32, vp, vp, i32, vp, i32]
        L.orbm_frame_create_device.argtypes = [C.POINTER(vp), vp, vp, i32, f32, f32, f32, f32, i32, vp]
        L.orbm_search_by_projection_points_device.argtypes = [vp, vp, vp, i32, f32, vp, vp, vp, vp, i32, vp]
        L.orbm_search_by_projection_frame_device.argtypes = [vp, vp, vp, i32, i32, vp, vp, vp, vp, i32, vp]
        L.orbm_search_by_projection_points.argtypes = [vp, vp, vp, i32, f32, vp, vp, C.POINTER(i32)]
        L.orbm_search_by_projection_frame.argtypes = [vp, vp, vp, i32, i32, vp, vp, C.POINTER(i32)]
        L.orbm_search_by_bow.argtypes = [vp, vp, vp, i32, vp, vp, vp, i32, vp, vp, i32, vp, vp, vp, i32, f32, i32, vp, C.POINTER(i32), i32]
        L.orbm_search_by_bow_kf.argtypes = [vp, vp, vp, i32, vp, vp, vp, i32, vp, vp, vp, i32, vp, vp, vp, i32, f32, i32, vp, C.POINTER(i32), i32]
        L.orbm_search_by_projection_reloc.argtypes = [vp, vp, i32, i32, i32, vp, C.POINTER(i32)]
        L.orbm_search_by_projection_reloc_device.argtypes = [vp, vp, i32, i32, i32, vp, vp, vp, i32, vp]
        L.orbm_search_by_projection_loop.argtypes = [vp, vp, i32, vp, C.POINTER(i32)]
        L.orbm_fuse_best.argtypes = [vp, vp, i32, vp, vp, i32, vp, vp]
        L.orbm_search_for_triangulation.argtypes = ([vp] * 4 + [i32] + [vp] * 3 + [i32]) * 2 + [vp, f32, f32, vp, vp, i32, i32, vp, C.POINTER(i32), i32]
        L.orbm_search_for_initialization.argtypes = [vp, vp, vp, vp, i32, f32, i32, C.POINTER(i32)]
        L.orbm_stereo_match.argtypes = [vp, vp, vp, vp, i32, vp, vp, i32, f32, f32, vp, vp]
        L.orbm_mono_frame_create.argtypes = [vp, vp, i32, i32, C.c_size_t, vp, vp, i32, f32, f32, f32, f32, vp, vp, vp, vp, vp, i32]
        L.orbm_stereo_frame_create.argtypes = [vp, vp, vp, i32, i32, C.c_size_t, f32, f32, f32, f32, f32, f32, vp, vp, vp, vp, vp, vp, vp, i32, vp, vp]
        L.orbm_db_create.argtypes = [C.POINTER(vp), C.c_int64, C.c_int64, i32]
        L.orbm_db_append.argtypes = [vp, vp, C.c_int64]
        L.orbm_db_append_device.argtypes = [vp, vp, C.c_int64]
        L.orbm_db_destroy.argtypes = [vp]
        L.orbm_db_rows.argtypes = [vp]
        L.orbm_db_rows.restype = C.c_int64
        L.orbm_bf_top2.argtypes = [vp, vp, i32, vp, vp, vp]
        L.orbm_bf_top2_device.argtypes = [vp, vp, i32, vp, vp, vp, vp]
        L.orbm_bf_top2_noalloc.argtypes = [vp, C.c_int64, C.c_int64, vp, i32, vp, vp, vp, i32, vp]
        L.orbm_merge_top2_device.argtypes = [vp, vp, vp, i32, i32, vp, vp, vp, vp]
        L.orbm_ratio_test.argtypes = [vp, vp, i32, i32, f32, vp]
        L.orbb200_strerror.argtypes = [i32]
        L.orbb200_strerror.restype = C.c_char_p
        L.orbb200_last_cuda_error.restype = C.c_char_p
        L.orbb200_measure_popc_peak.argtypes = [i32, C.POINTER(C.c_double)]
        L.orbb200_measure_imma_peak.argtypes = [i32, C.POINTER(C.c_double)]
        L.orbb200_measure_tcgen05_i8_peak.argtypes = [i32, C.POINTER(C.c_double)]
        L.orbb200_launch_count.restype = C.c_uint64
        _lib = L
    return _lib


def _p(a):
    if a is None:
        return None
    if isinstance(a, int):
        return C.c_void_p(a)
    if isinstance(a, C.c_void_p):
        return a
    return a.ctypes.data_as(C.c_void_p)


def _check(code, where):
    if code != OK:
        raise OrbError(code, where)


def launch_count():
    return int(lib().orbb200_launch_count())


def popc_peak(device=0):
    v = C.c_double()
    _check(lib().orbb200_measure_popc_peak(device, C.byref(v)), "orbb200_measure_popc_peak")
    return v.value


def imma_peak(device=0):
    v = C.c_double()
    _check(lib().orbb200_measure_imma_peak(device, C.byref(v)), "orbb200_measure_imma_peak")
    return v.value


def tcgen05_i8_peak(device=0):
    v = C.c_double()
    _check(lib().orbb200_measure_tcgen05_i8_peak(device, C.byref(v)), "orbb200_measure_tcgen05_i8_peak")
    return v.value


class ORBextractor:
    """Mirror of ORB_SLAM2::ORBextractor (reference include/ORBextractor.h:45-111) over the C ABI.

    ex = ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, max_size=(w, h))
    keypoints, descriptors = ex(image)        # == operator()(image, mask, keypoints, descriptors)
    """

    def __init__(self, nfeatures=1000, scaleFactor=1.2, nlevels=8, iniThFAST=20, minThFAST=7,
                 max_size=(640, 480), max_batch=1, device=0, blur_legacy=False):
        self.L = lib()
        self.cfg = OrbConfig(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, max_size[0], max_size[1],
                             max_batch, device, int(blur_legacy))
        self.h = C.c_void_p()
        _check(self.L.orbx_create(C.byref(self.cfg), C.byref(self.h)), "orbx_create")
        self.nlevels = nlevels
        self.nfeatures = nfeatures
        self.device = device
        self.default_capacity = nfeatures + 64 * nlevels + 256

    def close(self):
        if getattr(self, "h", None) and self.h.value:
            self.L.orbx_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # --- getters (ORBextractor.h:63-83) ---
    def _tables(self):
        n = self.nlevels
        sc, isc, s2, is2 = (np.zeros(n, np.float32) for _ in range(4))
        per = np.zeros(n, np.int32)
        um = np.zeros(16, np.int32)
        _check(self.L.orbx_get_tables(self.h, _p(sc), _p(isc), _p(s2), _p(is2), _p(per), _p(um)), "orbx_get_tables")
        return dict(scale=sc, inv_scale=isc, sigma2=s2, inv_sigma2=is2, per_level=per, umax=um)

    def GetLevels(self):
        return self.nlevels

    def GetScaleFactor(self):
        return float(self.cfg.scale_factor)

    def GetScaleFactors(self):
        return self._tables()["scale"]

    def GetInverseScaleFactors(self):
        return self._tables()["inv_scale"]

    def GetScaleSigmaSquares(self):
        return self._tables()["sigma2"]

    def GetInverseScaleSigmaSquares(self):
        return self._tables()["inv_sigma2"]

    # --- operator() ---
    def __call__(self, image, mask=None, capacity=None):
        image = np.asarray(image)
        if image.size == 0:
            return np.zeros(0, KP_DTYPE), np.zeros((0, 32), np.uint8)
        assert image.dtype == np.uint8 and image.ndim == 2, "image.type() == CV_8UC1 (ORBextractor.cpp:1060)"
        if image.strides[1] != 1:
            image = np.ascontiguousarray(image)
        h, w = image.shape
        cap = capacity or self.default_capacity
        kps = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        n = C.c_int()
        rc = self.L.orbx_extract(self.h, _p(image), w, h, image.strides[0], _p(kps), _p(desc), cap, C.byref(n))
        if rc == E_CAPACITY and capacity is None and n.value > cap:
            return self.__call__(image, mask, capacity=n.value + 16)
        _check(rc, "orbx_extract")
        return kps[:n.value].copy(), desc[:n.value].copy()

    def extract_batch(self, images, capacity=None):
        """images: (F, H, W) uint8 host array (pinned or pageable). Returns list of (kps, desc)."""
        images = np.ascontiguousarray(images, np.uint8)
        F, h, w = images.shape
        cap = capacity or self.default_capacity
        kps = np.zeros((F, cap), KP_DTYPE)
        desc = np.zeros((F, cap, 32), np.uint8)
        counts = np.zeros(F, np.int32)
        _check(self.L.orbx_extract_batch(self.h, _p(images), F, w, h, w, w * h, _p(kps), _p(desc), _p(counts), cap),
               "orbx_extract_batch")
        return [(kps[f, :counts[f]].copy(), desc[f, :counts[f]].copy()) for f in range(F)]

    def extract_batch_color(self, images, rgb_order=True, capacity=None):
        """images: (F, H, W, 3|4) uint8 host array; cv::cvtColor to grey on the device (Tracking.cpp:202-227), then extract."""
        images = np.ascontiguousarray(images, np.uint8)
        F, h, w, ch = images.shape
        cap = capacity or self.default_capacity
        kps = np.zeros((F, cap), KP_DTYPE)
        desc = np.zeros((F, cap, 32), np.uint8)
        counts = np.zeros(F, np.int32)
        _check(self.L.orbx_extract_batch_color(self.h, _p(images), F, w, h, w * ch, w * h * ch, ch, 1 if rgb_order else 0,
                                               _p(kps), _p(desc), _p(counts), cap), "orbx_extract_batch_color")
        return [(kps[f, :counts[f]].copy(), desc[f, :counts[f]].copy()) for f in range(F)]

    def set_rectification(self, mapx, mapy, src_size):
        """Store cv::initUndistortRectifyMap's CV_32F maps (dst_h x dst_w each) for raw frames of src_size = (w, h); None clears."""
        if mapx is None:
            _check(self.L.orbx_set_rectification(self.h, None, None, 0, 0, 0, 0), "orbx_set_rectification")
            return
        mapx = np.ascontiguousarray(mapx, np.float32); mapy = np.ascontiguousarray(mapy, np.float32)
        dh, dw = mapx.shape
        _check(self.L.orbx_set_rectification(self.h, _p(mapx), _p(mapy), dw, dh, src_size[0], src_size[1]), "orbx_set_rectification")

    def extract_batch_rectified(self, raw, capacity=None):
        """raw: (F, src_h, src_w) uint8 host frames; cv::remap INTER_LINEAR on the device (stereo_euroc.cc:174-185), then extract."""
        raw = np.ascontiguousarray(raw, np.uint8)
        F, h, w = raw.shape
        cap = capacity or self.default_capacity
        kps = np.zeros((F, cap), KP_DTYPE)
        desc = np.zeros((F, cap, 32), np.uint8)
        counts = np.zeros(F, np.int32)
        _check(self.L.orbx_extract_batch_rectified(self.h, _p(raw), F, w, w * h, _p(kps), _p(desc), _p(counts), cap), "orbx_extract_batch_rectified")
        return [(kps[f, :counts[f]].copy(), desc[f, :counts[f]].copy()) for f in range(F)]

    def extract_batch_device(self, imgs_ptr, F, w, h, stride, fstride, kps_ptr, desc_ptr, counts_ptr, capacity):
        """Raw device-pointer entry (ints): asynchronous on the handle's stream."""
        _check(self.L.orbx_extract_batch_device(self.h, _p(imgs_ptr), F, w, h, stride, fstride, _p(kps_ptr),
                                                _p(desc_ptr), _p(counts_ptr), capacity), "orbx_extract_batch_device")

    def graph_capture(self, imgs_ptr, F, w, h, stride, fstride, kps_ptr, desc_ptr, counts_ptr, capacity):
        """Capture the launch sequence of extract_batch_device on these buffers as a CUDA graph (orbx_graph_capture)."""
        _check(self.L.orbx_graph_capture(self.h, _p(imgs_ptr), F, w, h, stride, fstride, _p(kps_ptr), _p(desc_ptr), _p(counts_ptr), capacity),
               "orbx_graph_capture")

    def graph_launch(self):
        _check(self.L.orbx_graph_launch(self.h), "orbx_graph_launch")

    def sync(self):
        _check(self.L.orbx_sync(self.h), "orbx_sync")

    def stream(self):
        return int(self.L.orbx_stream(self.h) or 0)

    def set_profiling(self, on):
        _check(self.L.orbx_set_profiling(self.h, int(on)), "orbx_set_profiling")

    def get_profile(self):
        ms = np.zeros(3, np.float64)
        calls = C.c_int()
        _check(self.L.orbx_get_profile(self.h, _p(ms), C.byref(calls)), "orbx_get_profile")
        return dict(level_ms=ms[0], octree_ms=ms[1], desc_ms=ms[2], calls=calls.value)

    # --- mvImagePyramid (ORBextractor.h:85) and parity stages ---
    def pyramid_level(self, level, frame=0):
        w, h = C.c_int(), C.c_int()
        _check(self.L.orbx_pyramid_level(self.h, frame, level, None, 0, C.byref(w), C.byref(h)), "orbx_pyramid_level")
        out = np.zeros((h.value, w.value), np.uint8)
        _check(self.L.orbx_pyramid_level(self.h, frame, level, _p(out), w.value, C.byref(w), C.byref(h)), "orbx_pyramid_level")
        return out

    @property
    def mvImagePyramid(self):
        return [self.pyramid_level(l) for l in range(self.nlevels)]

    def _stage(self, level, stage, frame=0):
        need = C.c_size_t()
        _check(self.L.orbx_debug_stage(self.h, frame, level, stage, None, 0, C.byref(need)), "orbx_debug_stage")
        buf = np.zeros(max(need.value, 1), np.uint8)
        _check(self.L.orbx_debug_stage(self.h, frame, level, stage, _p(buf), need.value, C.byref(need)), "orbx_debug_stage")
        return buf[:need.value]

    def blurred_level(self, level, frame=0):
        w, h = C.c_int(), C.c_int()
        _check(self.L.orbx_pyramid_level(self.h, frame, level, None, 0, C.byref(w), C.byref(h)), "orbx_pyramid_level")
        return self._stage(level, STAGE_BLUR, frame).reshape(h.value, w.value)

    def fast_keypoints(self, level, frame=0):
        return self._stage(level, STAGE_FASTKPS, frame).view(np.int32).reshape(-1, 3)

    def level_keypoints(self, level, frame=0):
        return self._stage(level, STAGE_LEVELKPS, frame).view(KP_DTYPE)


def DescriptorDistance(a, b, device=0):
    """ORBmatcher::DescriptorDistance for n pairs (rows of a, b)."""
    a = np.ascontiguousarray(a, np.uint8).reshape(-1, 32)
    b = np.ascontiguousarray(b, np.uint8).reshape(-1, 32)
    out = np.zeros(a.shape[0], np.int32)
    _check(lib().orbm_distance_pairs(_p(a), _p(b), a.shape[0], _p(out), device), "orbm_distance_pairs")
    return out


def search_by_bow(desc_kf, angle_kf, kf_has_mp, fv_kf, desc_f, angle_f, fv_f, nnratio=0.7, check_ori=True, device=0):
    """ORBmatcher::SearchByBoW(KeyFrame*, Frame&, ...) (ORBmatcher.cpp:205-348): fv_* = (nodes, offsets, features) as
    Vocabulary.bow() returns them.  -> (matches[F.N]: KeyFrame feature index or -1, nmatches)."""
    desc_kf = np.ascontiguousarray(desc_kf, np.uint8); desc_f = np.ascontiguousarray(desc_f, np.uint8)
    angle_kf = np.ascontiguousarray(angle_kf, np.float32); angle_f = np.ascontiguousarray(angle_f, np.float32)
    kf_has_mp = np.ascontiguousarray(kf_has_mp, np.uint8)
    kn, ko, kf = [np.ascontiguousarray(a, np.int32) for a in fv_kf]
    fn, fo, ff = [np.ascontiguousarray(a, np.int32) for a in fv_f]
    out = np.zeros(len(desc_f), np.int32)
    n = C.c_int32()
    _check(lib().orbm_search_by_bow(_p(desc_kf), _p(angle_kf), _p(kf_has_mp), len(desc_kf), _p(kn), _p(ko), _p(kf), len(kn),
                                    _p(desc_f), _p(angle_f), len(desc_f), _p(fn), _p(fo), _p(ff), len(fn),
                                    nnratio, 1 if check_ori else 0, _p(out), C.byref(n), device), "orbm_search_by_bow")
    return out, n.value


def search_by_bow_kf(desc1, angle1, has1, fv1, desc2, angle2, has2, fv2, nnratio=0.75, check_ori=True, device=0):
    """ORBmatcher::SearchByBoW(KeyFrame*, KeyFrame*, ...) (ORBmatcher.cpp:604-743) -> (matches12[n1]: KF2 feature or -1, nmatches)."""
    desc1 = np.ascontiguousarray(desc1, np.uint8); desc2 = np.ascontiguousarray(desc2, np.uint8)
    angle1 = np.ascontiguousarray(angle1, np.float32); angle2 = np.ascontiguousarray(angle2, np.float32)
    has1 = np.ascontiguousarray(has1, np.uint8); has2 = np.ascontiguousarray(has2, np.uint8)
    a = [np.ascontiguousarray(x, np.int32) for x in fv1]; b = [np.ascontiguousarray(x, np.int32) for x in fv2]
    out = np.zeros(len(desc1), np.int32)
    n = C.c_int32()
    _check(lib().orbm_search_by_bow_kf(_p(desc1), _p(angle1), _p(has1), len(desc1), _p(a[0]), _p(a[1]), _p(a[2]), len(a[0]),
                                       _p(desc2), _p(angle2), _p(has2), len(desc2), _p(b[0]), _p(b[1]), _p(b[2]), len(b[0]),
                                       nnratio, 1 if check_ori else 0, _p(out), C.byref(n), device), "orbm_search_by_bow_kf")
    return out, n.value


def search_for_triangulation(desc1, kps1, ur1, has1, fv1, desc2, kps2, ur2, has2, fv2, F12, ex, ey, sf2, ls2, only_stereo=False, check_ori=True, device=0):
    """ORBmatcher::SearchForTriangulation (ORBmatcher.cpp:755-927) -> (vMatches12[n1], nmatches)."""
    desc1 = np.ascontiguousarray(desc1, np.uint8); desc2 = np.ascontiguousarray(desc2, np.uint8)
    kps1 = np.ascontiguousarray(kps1, KP_DTYPE); kps2 = np.ascontiguousarray(kps2, KP_DTYPE)
    ur1 = None if ur1 is None else np.ascontiguousarray(ur1, np.float32); ur2 = None if ur2 is None else np.ascontiguousarray(ur2, np.float32)
    has1 = np.ascontiguousarray(has1, np.uint8); has2 = np.ascontiguousarray(has2, np.uint8)
    a = [np.ascontiguousarray(x, np.int32) for x in fv1]; b = [np.ascontiguousarray(x, np.int32) for x in fv2]
    F12 = np.ascontiguousarray(F12, np.float32); sf2 = np.ascontiguousarray(sf2, np.float32); ls2 = np.ascontiguousarray(ls2, np.float32)
    out = np.zeros(len(desc1), np.int32)
    n = C.c_int32()
    _check(lib().orbm_search_for_triangulation(_p(desc1), _p(kps1), _p(ur1), _p(has1), len(desc1), _p(a[0]), _p(a[1]), _p(a[2]), len(a[0]),
                                               _p(desc2), _p(kps2), _p(ur2), _p(has2), len(desc2), _p(b[0]), _p(b[1]), _p(b[2]), len(b[0]),
                                               _p(F12), float(ex), float(ey), _p(sf2), _p(ls2), int(only_stereo), int(check_ori), _p(out),
                                               C.byref(n), device), "orbm_search_for_triangulation")
    return out, n.value


def undistort_keypoints(kps, K4, dist, device=0):
    """Frame::UndistortKeyPoints (Frame.cpp:462-496) -> mvKeysUn."""
    kps = np.ascontiguousarray(kps, KP_DTYPE); K4 = np.ascontiguousarray(K4, np.float32); dist = np.ascontiguousarray(dist, np.float32)
    out = np.zeros_like(kps)
    _check(lib().orbm_undistort_keypoints(_p(kps), len(kps), _p(K4), _p(dist), len(dist), _p(out), device), "orbm_undistort_keypoints")
    return out


def undistort_keypoints_device(kps_ptr, n, K4, dist, out_ptr, device=0, stream=0):
    K4 = np.ascontiguousarray(K4, np.float32); dist = np.ascontiguousarray(dist, np.float32)
    _check(lib().orbm_undistort_keypoints_device(C.c_void_p(kps_ptr), n, _p(K4), _p(dist), len(dist), C.c_void_p(out_ptr), device, C.c_void_p(stream)),
           "orbm_undistort_keypoints_device")


def image_bounds(cols, rows, K4, dist, device=0):
    """Frame::ComputeImageBounds (Frame.cpp:498-536) -> float32 (mnMinX, mnMinY, mnMaxX, mnMaxY)."""
    K4 = np.ascontiguousarray(K4, np.float32); dist = np.ascontiguousarray(dist, np.float32)
    b = np.zeros(4, np.float32)
    _check(lib().orbm_image_bounds(int(cols), int(rows), _p(K4), _p(dist), len(dist), _p(b), device), "orbm_image_bounds")
    return b


def distinctive_descriptors(desc, offsets, device=0):
    """MapPoint::ComputeDistinctiveDescriptors (MapPoint.cpp:302-380) for a batch of map points in CSR form:
    rows offsets[p]..offsets[p+1] of desc are point p's observed descriptors.  Returns (best_row, best_median)."""
    desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    offsets = np.ascontiguousarray(offsets, np.int64)
    n = offsets.size - 1
    row = np.zeros(max(n, 0), np.int32); med = np.zeros(max(n, 0), np.int32)
    _check(lib().orbm_distinctive_descriptors(_p(desc) if desc.size else None, _p(offsets), n, _p(row), _p(med), device),
           "orbm_distinctive_descriptors")
    return row, med


class Vocabulary:
    """Device image of a DBoW2 vocabulary tree; transform() / bow() = TemplatedVocabulary::transform (Frame::ComputeBoW)."""

    def __init__(self, k, L, parent, is_leaf, desc, weight, weighting=0, norm=1, device=0):
        self.L = lib()
        self.h = C.c_void_p()
        parent = np.ascontiguousarray(parent, np.int32); is_leaf = np.ascontiguousarray(is_leaf, np.uint8)
        desc = np.ascontiguousarray(desc, np.uint8); weight = np.ascontiguousarray(weight, np.float64)
        _check(self.L.orbv_vocab_create(C.byref(self.h), k, L, len(parent), _p(parent), _p(is_leaf), _p(desc), _p(weight),
                                        weighting, norm, device), "orbv_vocab_create")

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orbv_vocab_destroy(self.h)
            self.h = None

    @property
    def words(self):
        return self.L.orbv_vocab_words(self.h)

    def transform(self, feats, levelsup=4):
        feats = np.ascontiguousarray(feats, np.uint8).reshape(-1, 32)
        n = len(feats)
        wid = np.zeros(n, np.int32); w = np.zeros(n, np.float64); nid = np.zeros(n, np.int32)
        _check(self.L.orbv_transform(self.h, _p(feats) if n else None, n, levelsup, _p(wid), _p(w), _p(nid)), "orbv_transform")
        return wid, w, nid

    def bow(self, feats, levelsup=4):
        """-> (bow_ids, bow_vals, fv_nodes, fv_offsets, fv_features)"""
        feats = np.ascontiguousarray(feats, np.uint8).reshape(-1, 32)
        n = len(feats)
        ids = np.zeros(max(n, 1), np.int32); vals = np.zeros(max(n, 1), np.float64); nb = np.zeros(1, np.int32)
        fn = np.zeros(max(n, 1), np.int32); fo = np.zeros(n + 1, np.int32); ff = np.zeros(max(n, 1), np.int32); nf = np.zeros(1, np.int32)
        _check(self.L.orbv_compute_bow(self.h, _p(feats) if n else None, n, levelsup, _p(ids), _p(vals), _p(nb), _p(fn), _p(fo), _p(ff), _p(nf)),
               "orbv_compute_bow")
        return ids[:nb[0]].copy(), vals[:nb[0]].copy(), fn[:nf[0]].copy(), fo[:nf[0] + 1].copy(), ff[:fo[nf[0]]].copy()


class FrameMatcher:
    """Device image of a Frame's matching state (mvKeysUn, mDescriptors, 64x48 grid)."""

    def __init__(self, kps_un, desc, bounds, device=0):
        self.L = lib()
        self.kps = np.ascontiguousarray(kps_un, KP_DTYPE)
        self.desc = np.ascontiguousarray(desc, np.uint8)
        self.n = len(self.kps)
        self.h = C.c_void_p()
        _check(self.L.orbm_frame_create(C.byref(self.h), _p(self.kps), _p(self.desc), self.n,
                                        *[float(b) for b in bounds], device), "orbm_frame_create")

    @classmethod
    def from_device(cls, kps_un_ptr, desc_ptr, n, bounds, device=0, stream=0):
        """orbm_frame_create_device: kps_un_ptr / desc_ptr are device addresses (ints); nothing visits the host."""
        self = cls.__new__(cls)
        self.L = lib(); self.n = int(n); self.kps = None; self.desc = None
        self.h = C.c_void_p()
        _check(self.L.orbm_frame_create_device(C.byref(self.h), C.c_void_p(kps_un_ptr), C.c_void_p(desc_ptr), self.n,
                                               *[float(b) for b in bounds], device, C.c_void_p(stream)), "orbm_frame_create_device")
        return self

    def close(self):
        if getattr(self, "h", None) and self.h.value:
            self.L.orbm_frame_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def window_search(self, queries, cap=None):
        q = np.ascontiguousarray(queries, QUERY_DTYPE)
        nq = len(q)
        cap = cap or max(1024, 64 * nq)
        offs = np.zeros(nq + 1, np.int32)
        idx = np.zeros(cap, np.int32)
        dist = np.zeros(cap, np.int32)
        rc = self.L.orbm_window_search(self.h, _p(q), nq, _p(offs), _p(idx), _p(dist), cap)
        if rc == E_CAPACITY:
            return self.window_search(queries, cap=int(offs[nq]) + 16)
        _check(rc, "orbm_window_search")
        return offs, idx[:offs[nq]].copy(), dist[:offs[nq]].copy()


class ORBmatcher:
    """Mirror of ORB_SLAM2::ORBmatcher (reference include/ORBmatcher.h:37-141) on flattened Frame state."""
    TH_LOW, TH_HIGH, HISTO_LENGTH = 50, 100, 30

    def __init__(self, nnratio=0.6, checkOri=True):
        self.mfNNratio = float(nnratio)
        self.mbCheckOrientation = bool(checkOri)
        self.L = lib()

    @staticmethod
    def DescriptorDistance(a, b, device=0):
        return int(DescriptorDistance(a, b, device)[0])

    def SearchByProjection_points(self, frame, proj_queries, frame_mp, frame_mp_obs, mvuRight=None):
        q = np.ascontiguousarray(proj_queries, PROJ_QUERY_DTYPE)
        mp = np.ascontiguousarray(frame_mp, np.int32).copy(); ob = np.ascontiguousarray(frame_mp_obs, np.int32).copy()
        ur = mvuRight if mvuRight is None or isinstance(mvuRight, C.c_void_p) else np.ascontiguousarray(mvuRight, np.float32)
        n = C.c_int()
        _check(self.L.orbm_search_by_projection_points(frame.h, _p(ur), _p(q), len(q), self.mfNNratio, _p(mp), _p(ob), C.byref(n)),
               "orbm_search_by_projection_points")
        return n.value, mp, ob

    def SearchByProjection_frame(self, frame, proj_queries, frame_mp, frame_mp_obs, mvuRight=None):
        q = np.ascontiguousarray(proj_queries, PROJ_QUERY_DTYPE)
        mp = np.ascontiguousarray(frame_mp, np.int32).copy(); ob = np.ascontiguousarray(frame_mp_obs, np.int32).copy()
        ur = mvuRight if mvuRight is None or isinstance(mvuRight, C.c_void_p) else np.ascontiguousarray(mvuRight, np.float32)
        n = C.c_int()
        _check(self.L.orbm_search_by_projection_frame(frame.h, _p(ur), _p(q), len(q), int(self.mbCheckOrientation), _p(mp), _p(ob), C.byref(n)),
               "orbm_search_by_projection_frame")
        return n.value, mp, ob

    def SearchByProjection_device(self, mode, frame, q_ptr, nq, mp_ptr, obs_ptr, nmatches_ptr, status_ptr, uright_ptr=0, cand_capacity=0, stream=0):
        """orbm_search_by_projection_{points,frame}_device on raw device addresses (mode "points" | "frame"); asynchronous."""
        vp = C.c_void_p
        if mode == "points":
            _check(self.L.orbm_search_by_projection_points_device(frame.h, vp(uright_ptr), vp(q_ptr), nq, self.mfNNratio, vp(mp_ptr), vp(obs_ptr),
                                                                  vp(nmatches_ptr), vp(status_ptr), cand_capacity, vp(stream)),
                   "orbm_search_by_projection_points_device")
        else:
            _check(self.L.orbm_search_by_projection_frame_device(frame.h, vp(uright_ptr), vp(q_ptr), nq, int(self.mbCheckOrientation), vp(mp_ptr),
                                                                 vp(obs_ptr), vp(nmatches_ptr), vp(status_ptr), cand_capacity, vp(stream)),
                   "orbm_search_by_projection_frame_device")

    def SearchByProjection_reloc(self, frame, proj_queries, frame_mp, ORBdist):
        """SearchByProjection(Frame&, KeyFrame*, sAlreadyFound, th, ORBdist) (ORBmatcher.cpp:1661-1789)."""
        q = np.ascontiguousarray(proj_queries, PROJ_QUERY_DTYPE)
        mp = np.ascontiguousarray(frame_mp, np.int32).copy()
        n = C.c_int()
        _check(self.L.orbm_search_by_projection_reloc(frame.h, _p(q), len(q), ORBdist, int(self.mbCheckOrientation), _p(mp), C.byref(n)),
               "orbm_search_by_projection_reloc")
        return n.value, mp

    def SearchByProjection_loop(self, frame, proj_queries, matched):
        """SearchByProjection(KeyFrame*, Scw, vpPoints, vpMatched, th) (ORBmatcher.cpp:352-472)."""
        q = np.ascontiguousarray(proj_queries, PROJ_QUERY_DTYPE)
        m = np.ascontiguousarray(matched, np.int32).copy()
        n = C.c_int()
        _check(self.L.orbm_search_by_projection_loop(frame.h, _p(q), len(q), _p(m), C.byref(n)), "orbm_search_by_projection_loop")
        return n.value, m

    def FuseBest(self, frame, proj_queries, mvuRight, inv_level_sigma2, max_dist):
        """Per-point search of Fuse / Fuse(Scw) / SearchBySim3 (ORBmatcher.cpp:962-1500): (best_idx, best_dist)."""
        q = np.ascontiguousarray(proj_queries, PROJ_QUERY_DTYPE)
        ur = mvuRight if mvuRight is None or isinstance(mvuRight, C.c_void_p) else np.ascontiguousarray(mvuRight, np.float32)
        s2 = None if inv_level_sigma2 is None else np.ascontiguousarray(inv_level_sigma2, np.float32)
        bi = np.zeros(len(q), np.int32); bd = np.zeros(len(q), np.int32)
        _check(self.L.orbm_fuse_best(frame.h, _p(q), len(q), _p(ur), _p(s2), max_dist, _p(bi), _p(bd)), "orbm_fuse_best")
        return bi, bd

    def SearchForInitialization(self, F1, F2, vbPrevMatched, windowSize=10):
        pm = np.ascontiguousarray(vbPrevMatched, np.float32).copy()
        m12 = np.zeros(max(F1.n, 1), np.int32)
        n = C.c_int()
        _check(self.L.orbm_search_for_initialization(F1.h, F2.h, _p(pm), _p(m12), windowSize, self.mfNNratio,
                                                     int(self.mbCheckOrientation), C.byref(n)), "orbm_search_for_initialization")
        return n.value, m12[:F1.n], pm


def stereo_match(exL, exR, kpsL, descL, kpsR, descR, mbf, mb=0.0):
    """Frame::ComputeStereoMatches on the two extractors' device pyramids."""
    kL = np.ascontiguousarray(kpsL, KP_DTYPE); dL = np.ascontiguousarray(descL, np.uint8)
    kR = np.ascontiguousarray(kpsR, KP_DTYPE); dR = np.ascontiguousarray(descR, np.uint8)
    ur = np.zeros(len(kL), np.float32)
    dp = np.zeros(len(kL), np.float32)
    with np.errstate(divide="ignore"):
        _check(lib().orbm_stereo_match(exL.h, exR.h, _p(kL), _p(dL), len(kL), _p(kR), _p(dR), len(kR), mbf, mb, _p(ur), _p(dp)),
               "orbm_stereo_match")
    return ur, dp


URIGHT_OF_FRAME = C.c_void_p(1)      # ORBM_URIGHT_OF_FRAME


def stereo_frame(ex, left, right, mbf, bounds, mb=0.0, capacity=None):
    """Frame::Frame (stereo constructor, Frame.cpp:61-115) as one call (orbm_stereo_frame_create): both images extracted as one batch of
    `ex` (max_batch >= 2), stereo matching and the grid on the device.  Returns (FrameMatcher, kL, dL, kR, dR, mvuRight, mvDepth)."""
    left = np.ascontiguousarray(left, np.uint8); right = np.ascontiguousarray(right, np.uint8)
    assert left.shape == right.shape and left.ndim == 2
    ht, w = left.shape
    cap = int(capacity or ex.default_capacity)
    kL = np.zeros(cap, KP_DTYPE); dL = np.zeros((cap, 32), np.uint8); kR = np.zeros(cap, KP_DTYPE); dR = np.zeros((cap, 32), np.uint8)
    ur = np.zeros(cap, np.float32); dp = np.zeros(cap, np.float32)
    nL, nR = C.c_int(), C.c_int()
    F = FrameMatcher.__new__(FrameMatcher)
    F.L = lib(); F.kps = None; F.desc = None; F.h = C.c_void_p(); F.n = 0
    with np.errstate(divide="ignore"):
        _check(F.L.orbm_stereo_frame_create(ex.h, _p(left), _p(right), w, ht, w, mbf, mb, *[float(b) for b in bounds], C.byref(F.h),
                                            _p(kL), _p(dL), C.byref(nL), _p(kR), _p(dR), C.byref(nR), cap, _p(ur), _p(dp)), "orbm_stereo_frame_create")
    F.n = nL.value
    F.kps = kL[:F.n]; F.desc = dL[:F.n]
    return F, kL[:nL.value], dL[:nL.value], kR[:nR.value], dR[:nR.value], ur[:nL.value], dp[:nL.value]


def mono_frame(ex, image, bounds, K4=None, dist=None, capacity=None):
    """Frame::Frame (monocular constructor, Frame.cpp:171-224) as one call (orbm_mono_frame_create).  Returns (FrameMatcher, mvKeys, mvKeysUn, mDescriptors)."""
    image = np.ascontiguousarray(image, np.uint8)
    ht, w = image.shape
    cap = int(capacity or ex.default_capacity)
    k = np.zeros(cap, KP_DTYPE); ku = np.zeros(cap, KP_DTYPE); d = np.zeros((cap, 32), np.uint8)
    n = C.c_int()
    K = None if K4 is None else np.ascontiguousarray(K4, np.float32); D = None if dist is None else np.ascontiguousarray(dist, np.float32)
    F = FrameMatcher.__new__(FrameMatcher)
    F.L = lib(); F.kps = None; F.desc = None; F.h = C.c_void_p(); F.n = 0
    _check(F.L.orbm_mono_frame_create(ex.h, _p(image), w, ht, w, _p(K), _p(D), 0 if D is None else len(D), *[float(b) for b in bounds], C.byref(F.h),
                                      _p(k), _p(ku), _p(d), C.byref(n), cap), "orbm_mono_frame_create")
    F.n = n.value; F.kps = ku[:F.n]; F.desc = d[:F.n]
    return F, k[:n.value], ku[:n.value], d[:n.value]


class DescriptorDB:
    """One shard of a keyframe-descriptor database for brute-force top-2 (rows [row_offset, row_offset+n))."""

    def __init__(self, capacity_rows, row_offset=0, device=0):
        self.L = lib()
        self.h = C.c_void_p()
        _check(self.L.orbm_db_create(C.byref(self.h), capacity_rows, row_offset, device), "orbm_db_create")

    def close(self):
        if getattr(self, "h", None) and self.h.value:
            self.L.orbm_db_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def append(self, rows):
        rows = np.ascontiguousarray(rows, np.uint8).reshape(-1, 32)
        _check(self.L.orbm_db_append(self.h, _p(rows), rows.shape[0]), "orbm_db_append")

    def append_device(self, ptr, nrows):
        _check(self.L.orbm_db_append_device(self.h, _p(ptr), nrows), "orbm_db_append_device")

    def rows(self):
        return int(self.L.orbm_db_rows(self.h))

    def top2(self, queries):
        q = np.ascontiguousarray(queries, np.uint8).reshape(-1, 32)
        nq = q.shape[0]
        d1 = np.zeros(nq, np.int32); i1 = np.zeros(nq, np.int64); d2 = np.zeros(nq, np.int32)
        _check(self.L.orbm_bf_top2(self.h, _p(q), nq, _p(d1), _p(i1), _p(d2)), "orbm_bf_top2")
        return d1, i1, d2

    def top2_device(self, q_ptr, nq, d1_ptr, i1_ptr, d2_ptr, stream=0):
        _check(self.L.orbm_bf_top2_device(self.h, _p(q_ptr), nq, _p(d1_ptr), _p(i1_ptr), _p(d2_ptr), _p(stream) if stream else None),
               "orbm_bf_top2_device")


def bf_top2_noalloc(rows_ptr, nrows, row_offset, q_ptr, nq, d1_ptr, i1_ptr, d2_ptr, device=0, stream=0):
    """orbm_bf_top2_noalloc: device pointers in and out, scratch in the library's static device array (no allocation)."""
    _check(lib().orbm_bf_top2_noalloc(_p(rows_ptr), nrows, row_offset, _p(q_ptr), nq, _p(d1_ptr), _p(i1_ptr), _p(d2_ptr), device,
                                      _p(stream) if stream else None), "orbm_bf_top2_noalloc")


def merge_top2_device(d1_all, i1_all, d2_all, nshards, nq, d1, i1, d2, stream=0):
    _check(lib().orbm_merge_top2_device(_p(d1_all), _p(i1_all), _p(d2_all), nshards, nq, _p(d1), _p(i1), _p(d2),
                                        _p(stream) if stream else None), "orbm_merge_top2_device")


def ratio_test(d1, d2, th_low=50, ratio=0.75):
    d1 = np.ascontiguousarray(d1, np.int32); d2 = np.ascontiguousarray(d2, np.int32)
    acc = np.zeros(len(d1), np.uint8)
    _check(lib().orbm_ratio_test(_p(d1), _p(d2), len(d1), th_low, ratio, _p(acc)), "orbm_ratio_test")
    return acc.astype(bool)
