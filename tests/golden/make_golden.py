#!/usr/bin/env python3
"""Generates the golden fixtures under tests/golden/ (run in the build container, where cv2 4.13 and /root/reference exist).

The reference (SFXiang/ORB-SLAM2-Annotation) ships no golden vectors or tests for this path, so the oracle is
pinned against outputs of the real implementations, generated here and committed:

  cv2_prims.npz   OpenCV 4.13.0 itself (cv2): the third-party primitives ORBextractor.cpp calls —
                  cv::resize INTER_LINEAR (ORBextractor.cpp:1136), cv::GaussianBlur 7x7 s=2 REFLECT_101 (:1098),
                  cv::FAST 9-16 + NMS at thresholds 20 and 7 (:818-825), cv::fastAtan2 (:105), BFMatcher Hamming knn (k=2).
  ref_extract.npz the reference's own ORBextractor.cpp, compiled unmodified from /root/reference (oracle/_ref), run on
                  seeded synthetic frames: full keypoint + descriptor output, pyramid checksums.

  cv2_cvtcolor.npz  cv2.cvtColor RGB/BGR/RGBA/BGRA -> GRAY (Tracking.cpp:202-227, the SURVEY 8(f)-4 row): a small seeded image
                  per code plus the sha256 of the gray image of all 2^24 RGB colours.  (python make_golden.py --cvtcolor
                  regenerates only this file.)

  cv2_remap.npz   cv2.remap INTER_LINEAR with float maps (stereo rectification, Examples/Stereo/stereo_euroc.cc:174-185): a
                  rectification-like map (cv2.initUndistortRectifyMap of a distorted pinhole camera) and a hostile map (fractions
                  on every 1/32 step, out-of-range and negative coordinates).  (python make_golden.py --remap regenerates it.)

  cv2_undistort.npz  cv2.undistortPoints(pts, K, D, None, K) (Frame::UndistortKeyPoints / ComputeImageBounds, Frame.cpp:462-536): the
                  EuRoC, TUM1 and a strong 5/8-coefficient calibration on seeded points inside and far outside the image (the
                  icdist < 0 branch included).  (python make_golden.py --undistort regenerates it.)

Inputs are regenerated from seeds by synth.py (numpy only), so only outputs are stored.  Usage: python tests/golden/make_golden.py
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [os.path.join(ROOT, "orb-slam2-annotation_b200"), os.path.join(ROOT, "oracle")]

from synth import synth, synth_descriptor_db, synth_queries  # noqa: E402

PRIM_SHAPES = [(160, 120, 11), (97, 61, 12), (203, 77, 13)]     # (w, h, seed): small, odd widths, ragged cell grids
ATAN_N = 4096


def level_sizes(w, h, nlevels=8, sf=1.2):
    sc = np.float32(1.0)
    out = []
    for _ in range(nlevels):
        inv = np.float32(1.0) / sc
        out.append((int(np.rint(np.float32(w) * inv)), int(np.rint(np.float32(h) * inv))))
        sc = np.float32(np.float64(sc) * np.float64(np.float32(sf)))
    return out


def kp_exact_bytes(k):
    """The bit-exact part of a keypoint array (everything but the angle), field by field."""
    return b"".join(np.ascontiguousarray(k[f]).tobytes() for f in ("x", "y", "size", "response", "octave", "class_id"))


def atan_inputs():
    rng = np.random.default_rng(99)
    y = rng.integers(-200000, 200000, ATAN_N).astype(np.float32)
    x = rng.integers(-200000, 200000, ATAN_N).astype(np.float32)
    y[:8] = [0, 0, 1, -1, 5, -5, 0, 3]
    x[:8] = [0, 1, 0, 0, 5, 5, -7, -3]
    return y, x


def color_inputs():
    rng = np.random.default_rng(77)
    return rng.integers(0, 256, (29, 37, 3), dtype=np.uint8), rng.integers(0, 256, (31, 41, 4), dtype=np.uint8)


def all_colours():
    c = np.arange(1 << 24, dtype=np.uint32)
    return np.stack([(c >> 16) & 255, (c >> 8) & 255, c & 255], 1).astype(np.uint8).reshape(4096, 4096, 3)


def main_cvtcolor():
    import cv2
    c3, c4 = color_inputs()
    out = {"cv2_version": np.array(cv2.__version__),
           "rgb2gray": cv2.cvtColor(c3, cv2.COLOR_RGB2GRAY), "bgr2gray": cv2.cvtColor(c3, cv2.COLOR_BGR2GRAY),
           "rgba2gray": cv2.cvtColor(c4, cv2.COLOR_RGBA2GRAY), "bgra2gray": cv2.cvtColor(c4, cv2.COLOR_BGRA2GRAY),
           "all_rgb_sha": np.array(hashlib.sha256(cv2.cvtColor(all_colours(), cv2.COLOR_RGB2GRAY).tobytes()).hexdigest()),
           "all_bgr_sha": np.array(hashlib.sha256(cv2.cvtColor(all_colours(), cv2.COLOR_BGR2GRAY).tobytes()).hexdigest())}
    np.savez_compressed(os.path.join(HERE, "cv2_cvtcolor.npz"), **out)
    print("cv2_cvtcolor.npz", os.path.getsize(os.path.join(HERE, "cv2_cvtcolor.npz")), "bytes")


def remap_inputs():
    """(image, [(mapx, mapy), ...]) — maps are regenerated here, not stored: pure numpy / float32 formulas."""
    img = synth(188, 120, 31)
    h, w = img.shape
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    rng = np.random.default_rng(32)
    hostile_x = (xx + rng.integers(-40, 40, (h, w)).astype(np.float32) / np.float32(32) * rng.integers(0, 33, (h, w)).astype(np.float32) - np.float32(3)).astype(np.float32)
    hostile_y = (yy + rng.integers(-40, 40, (h, w)).astype(np.float32) / np.float32(32) * rng.integers(0, 33, (h, w)).astype(np.float32) + np.float32(2)).astype(np.float32)
    hostile_x[::7, ::5] = np.float32(-1.5); hostile_y[::11, ::3] = np.float32(h + 0.25); hostile_x[3::9, 1::4] = np.float32(w - 1 + 0.5)
    return img, hostile_x, hostile_y


def rectify_maps(cv2, w, h):
    K = np.array([[0.61 * w, 0, 0.49 * w], [0, 0.61 * w, 0.52 * h], [0, 0, 1]], np.float64)
    D = np.array([-0.28, 0.07, 0.0002, 0.00002, 0.0], np.float64)
    R = cv2.Rodrigues(np.array([0.01, -0.02, 0.005]))[0]
    P = np.array([[0.58 * w, 0, 0.5 * w, 0], [0, 0.58 * w, 0.5 * h, 0], [0, 0, 1, 0]], np.float64)
    return cv2.initUndistortRectifyMap(K, D, R, P[:3, :3], (w, h), cv2.CV_32F)


def main_remap():
    import cv2
    img, hx, hy = remap_inputs()
    h, w = img.shape
    rx, ry = rectify_maps(cv2, w, h)
    out = {"cv2_version": np.array(cv2.__version__), "rect_mapx": rx, "rect_mapy": ry,
           "rect_out": cv2.remap(img, rx, ry, cv2.INTER_LINEAR), "hostile_out": cv2.remap(img, hx, hy, cv2.INTER_LINEAR)}
    np.savez_compressed(os.path.join(HERE, "cv2_remap.npz"), **out)
    print("cv2_remap.npz", os.path.getsize(os.path.join(HERE, "cv2_remap.npz")), "bytes")


def undistort_cases():
    """(name, K4 = fx fy cx cy, dist, w, h): ORB-SLAM2's shipped calibrations (Examples/Monocular/EuRoC.yaml, TUM1.yaml) + hostile ones."""
    return [("euroc", [458.654, 457.296, 367.215, 248.375], [-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05], 752, 480),
            ("tum1", [517.306408, 516.469215, 318.643040, 255.313989], [0.262383, -0.953104, -0.005358, 0.002628, 1.163314], 640, 480),
            ("strong8", [300.0, 310.0, 320.0, 240.0], [-0.45, 0.31, 0.004, -0.003, -0.12, 0.02, -0.01, 0.005], 640, 480),
            ("barrel", [200.0, 200.0, 320.0, 240.0], [-0.9, 0.2, 0.0, 0.0], 640, 480)]


def undistort_points_in(w, h, seed):
    rng = np.random.default_rng(seed)
    inside = np.stack([rng.uniform(0, w, 3000), rng.uniform(0, h, 3000)], 1)
    grid = np.stack(np.meshgrid(np.arange(0, w + 1, 40.0), np.arange(0, h + 1, 40.0)), -1).reshape(-1, 2)     # integer pixels, corners included
    far = np.stack([rng.uniform(-3 * w, 4 * w, 500), rng.uniform(-3 * h, 4 * h, 500)], 1)                   # far outside: icdist < 0 happens
    return np.concatenate([inside, grid, far]).astype(np.float32)


def main_undistort():
    import cv2
    out = {"cv2_version": np.array(cv2.__version__)}
    for i, (name, K4, D, w, h) in enumerate(undistort_cases()):
        K = np.array([[K4[0], 0, K4[2]], [0, K4[1], K4[3]], [0, 0, 1]], np.float32)
        pts = undistort_points_in(w, h, 300 + i)
        out[name] = cv2.undistortPoints(pts.reshape(-1, 1, 2), K, np.array(D, np.float32), None, K).reshape(-1, 2)
    np.savez_compressed(os.path.join(HERE, "cv2_undistort.npz"), **out)
    print("cv2_undistort.npz", os.path.getsize(os.path.join(HERE, "cv2_undistort.npz")), "bytes")


def main():
    import cv2
    cv2.setNumThreads(1)
    out = {"cv2_version": np.array(cv2.__version__)}
    for (w, h, seed) in PRIM_SHAPES:
        img = synth(w, h, seed)
        tag = f"{w}x{h}_{seed}"
        cur = img
        for L, (lw, lh) in enumerate(level_sizes(w, h)):
            if L > 0:
                cur = cv2.resize(cur, (lw, lh), interpolation=cv2.INTER_LINEAR)
                out[f"pyr_{tag}_L{L}"] = cur
            out[f"blur_{tag}_L{L}"] = cv2.GaussianBlur(cur, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101)
        for th in (20, 7):
            det = cv2.FastFeatureDetector_create(th, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
            kps = det.detect(img)
            out[f"fast{th}_{tag}"] = np.array([(int(k.pt[0]), int(k.pt[1]), int(k.response)) for k in kps], np.int32).reshape(-1, 3)
    tiny = synth(13, 9, 14)
    out["blur_13x9_14"] = cv2.GaussianBlur(tiny, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101)
    out["resize_13x9_14_to_11x8"] = cv2.resize(tiny, (11, 8), interpolation=cv2.INTER_LINEAR)
    y, x = atan_inputs()
    out["atan2"] = np.array([cv2.fastAtan2(float(a), float(b)) for a, b in zip(y, x)], np.float32)
    # Hamming: cv2.BFMatcher knn (k=2) distances on a small random DB; indices only where the best is unique
    db = synth_descriptor_db(3000, seed=5)
    q = synth_queries(db, 64, seed=6)
    m = cv2.BFMatcher(cv2.NORM_HAMMING).knnMatch(q, db, k=2)
    out["bf_d1"] = np.array([p[0].distance for p in m], np.int32)
    out["bf_d2"] = np.array([p[1].distance for p in m], np.int32)
    out["bf_i1"] = np.array([p[0].trainIdx for p in m], np.int64)
    np.savez_compressed(os.path.join(HERE, "cv2_prims.npz"), **out)

    import pyoracle
    ref = {}
    for (w, h, nf, seed) in [(320, 240, 500, 21), (203, 177, 300, 22)]:
        ex = pyoracle.RefExtractor(nf)
        k, d = ex.extract(synth(w, h, seed))
        tag = f"{w}x{h}_{nf}_{seed}"
        ref[f"kps_{tag}"] = k
        ref[f"desc_{tag}"] = d
        ref[f"pyrsha_{tag}"] = np.array([hashlib.sha256(ex.level_image(L).tobytes()).hexdigest() for L in range(8)])
    for (w, h, nf, seed) in [(640, 480, 1000, 1), (1241, 376, 2000, 2), (752, 480, 1200, 3)]:   # config shapes: checksums only
        ex = pyoracle.RefExtractor(nf)
        k, d = ex.extract(synth(w, h, seed))
        tag = f"{w}x{h}_{nf}_{seed}"
        ref[f"count_{tag}"] = np.array(len(k))
        ref[f"kpsha_{tag}"] = np.array(hashlib.sha256(kp_exact_bytes(k)).hexdigest())
        ref[f"angles_{tag}"] = k["angle"].copy()
        ref[f"descsha_{tag}"] = np.array(hashlib.sha256(d.tobytes()).hexdigest())
    np.savez_compressed(os.path.join(HERE, "ref_extract.npz"), **ref)
    for f in ("cv2_prims.npz", "ref_extract.npz"):
        print(f, os.path.getsize(os.path.join(HERE, f)), "bytes")


if __name__ == "__main__":
    only = [a for a in ("--cvtcolor", "--remap", "--undistort") if a in sys.argv]
    if not only:
        main()
    if not only or "--cvtcolor" in only:
        main_cvtcolor()
    if not only or "--remap" in only:
        main_remap()
    if not only or "--undistort" in only:
        main_undistort()
