"""Generates tests/golden/ref_match.npz: inputs and outputs of the REFERENCE's own matching code (src/ORBmatcher.cpp, src/Frame.cpp,
src/KeyFrame.cpp, src/MapPoint.cpp, Thirdparty/DBoW2 — compiled unmodified into oracle/_ref/liborb_refmatch.so, see oracle/Makefile
`refmatch` and oracle/refmatch_capi.cpp) on seeded synthetic scenes.  Run here (needs /root/reference); the .npz is committed and is
what tests/test_ref_match_golden.py checks the CPU oracle (-m "not gpu") and the CUDA path (-m gpu) against.

How a case is built: Frames come from the reference's Frame constructors (its extractor, UndistortKeyPoints, ComputeStereoMatches,
AssignFeaturesToGrid) on synth images; MapPoints get world positions that project near their features; then the reference function
runs with call tracing on, and the trace (every GetFeaturesInArea / PredictScale call it made, in order) gives the exact projected
window of every point — u, v, radius, level bounds as the reference computed them.  Those are stored as the flattened
orbm_proj_query records the C ABI takes, next to the reference's outputs.  The only values recomputed here are u_right (u - bf/z)
and the pass/fail of the early-exit gates; both are cross-checked against the trace (the recomputed u, v must equal the traced ones
bit for bit, and the set of points that reached GetFeaturesInArea must equal the set that passed the recomputed gates).

    python tests/golden/make_golden_match.py
"""
import os
import sys
import tempfile
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (os.path.join(ROOT, "oracle"), os.path.join(ROOT, "orb-slam2-annotation_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)
import pyrefmatch as R                  # noqa: E402
from synth import synth, synth_stereo, synth_vocabulary, synth_vocabulary_features   # noqa: E402

f32 = np.float32
QUERY_DTYPE = np.dtype([("u", "<f4"), ("v", "<f4"), ("r", "<f4"), ("min_level", "<i4"), ("max_level", "<i4"), ("desc", "u1", (32,))])
PROJ_QUERY_DTYPE = np.dtype([("q", QUERY_DTYPE), ("u_right", "<f4"), ("max_err_right", "<f4"), ("id", "<i4"), ("has_obs", "<i4"), ("angle", "<f4")])
assert PROJ_QUERY_DTYPE.itemsize == 72

MONO_K4 = [517.306408, 516.469215, 318.643040, 255.313989]                       # Examples/Monocular/TUM1.yaml
MONO_DIST = [0.262383, -0.953104, -0.005358, 0.002628, 1.163314]
STEREO_K4 = [435.2046959714599, 435.2046959714599, 367.4517211914062, 252.2008514404297]   # Examples/Stereo/EuRoC.yaml
STEREO_BF = 47.90639384423901
OUT = {}


def put(case, **kw):
    for k, v in kw.items():
        OUT[f"{case}/{k}"] = np.asarray(v)


def rot(rx, ry, rz):
    cx, sx, cy, sy, cz, sz = np.cos(rx), np.sin(rx), np.cos(ry), np.sin(ry), np.cos(rz), np.sin(rz)
    Rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]]); Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
    return Rz @ Ry @ Rx


def pose(rx, ry, rz, t):
    T = np.eye(4)
    T[:3, :3] = rot(rx, ry, rz); T[:3, 3] = t
    return T.astype(f32)


def world_points(T, K4, uv, z):
    """World positions (float32) that the camera with pose T (Tcw) sees at pixels uv with depths z."""
    fx, fy, cx, cy = [float(v) for v in K4]
    T = T.astype(np.float64)
    Xc = np.stack([(uv[:, 0] - cx) / fx * z, (uv[:, 1] - cy) / fy * z, z], 1)
    return ((Xc - T[:3, 3]) @ T[:3, :3]).astype(f32)           # R^T (Xc - t)


def mat3(R_, x):
    """Row-major 3x3 (float32) times points (n, 3) + nothing, in the shim's order: ((r0*x + r1*y) + r2*z), every step rounded to float32."""
    R_ = R_.astype(f32); x = x.astype(f32)
    return np.stack([(R_[i, 0] * x[:, 0] + R_[i, 1] * x[:, 1]) + R_[i, 2] * x[:, 2] for i in range(3)], 1).astype(f32)


def project_frame(T, K4f, X):
    """x3Dc = Rcw*x3Dw+tcw; invzc = 1.0/z (double divide, ORBmatcher.cpp:1537, :1685); u = fx*xc*invzc+cx — float32 throughout."""
    T = T.astype(f32)
    Xc = (mat3(T[:3, :3], X) + T[:3, 3][None, :]).astype(f32)
    invz = (1.0 / Xc[:, 2].astype(np.float64)).astype(f32)
    u = (K4f[0] * Xc[:, 0] * invz + K4f[2]).astype(f32)
    v = (K4f[1] * Xc[:, 1] * invz + K4f[3]).astype(f32)
    return Xc, invz, u, v


def project_kf(Rcw, tcw, K4f, X):
    """p3Dc = Rcw*p3Dw+tcw; invz = 1/z (float divide); x = X*invz; u = fx*x+cx (ORBmatcher.cpp:398-405, :1006-1012)."""
    Xc = (mat3(Rcw, X) + tcw.astype(f32)[None, :]).astype(f32)
    with np.errstate(divide="ignore"):
        invz = (f32(1) / Xc[:, 2]).astype(f32)
    x = (Xc[:, 0] * invz).astype(f32); y = (Xc[:, 1] * invz).astype(f32)
    u = (K4f[0] * x + K4f[2]).astype(f32); v = (K4f[1] * y + K4f[3]).astype(f32)
    return Xc, invz, u, v


def trace_queries(events, mps, point_kind, gfa_obj=None):
    """(point index, u, v, r, a, b, predicted level) for every GetFeaturesInArea event, the point being the object of the last
    `point_kind` event (3 GetWorldPos / 4 isBad) before it; a / b are the level bounds of Frame::GetFeaturesInArea (-1, -1 for
    KeyFrame's); the level is the result of the last PredictScale before it (or -1)."""
    ptr2idx = {int(p): i for i, p in enumerate(mps.ptrs())}
    out = []
    cur = -1; lvl = -1
    for e in events:
        k = int(e["kind"])
        if k == point_kind:
            cur = ptr2idx.get(int(e["obj"]), -2)
        elif k == 5:
            lvl = int(e["a"])
        elif k in (0, 1) and (gfa_obj is None or int(e["obj"]) == gfa_obj):
            out.append((cur, e["x"], e["y"], e["r"], int(e["a"]), int(e["b"]), lvl))
    return out


def best_idx_from_trace(events, mps, kf_obj):
    """Fuse: per point, the index passed to KeyFrame::GetMapPoint right after its window query (= bestIdx, ORBmatcher.cpp:1094), else -1."""
    ptr2idx = {int(p): i for i, p in enumerate(mps.ptrs())}
    best = np.full(mps.n, -1, np.int32)
    cur = -1; armed = False
    for e in events:
        k = int(e["kind"])
        if k == 3:
            cur = ptr2idx.get(int(e["obj"]), -2); armed = False
        elif k == 1:
            armed = True
        elif k == 2 and armed and int(e["obj"]) == kf_obj:
            best[cur] = int(e["a"]); armed = False
    return best


def make_queries(tq, desc, u_right, max_err_right, ids, has_obs, angle):
    q = np.zeros(len(tq), PROJ_QUERY_DTYPE)
    for j, (pi, u, v, r, a, b, lvl) in enumerate(tq):
        q["q"]["u"][j] = u; q["q"]["v"][j] = v; q["q"]["r"][j] = r; q["q"]["min_level"][j] = a; q["q"]["max_level"][j] = b
        q["q"]["desc"][j] = desc[pi]
        q["u_right"][j] = u_right[pi]; q["max_err_right"][j] = r if max_err_right is None else max_err_right[pi]
        q["id"][j] = ids[pi]; q["has_obs"][j] = has_obs[pi]; q["angle"][j] = angle[pi]
    return q


def point_descriptors(rng, desc, flip_bits=12):
    """MapPoint descriptors: the feature's descriptor with a few random bit flips (a different view of the same corner)."""
    d = desc.copy()
    for i in range(len(d)):
        for b in rng.integers(0, 256, int(rng.integers(0, flip_bits + 1))):
            d[i, b >> 3] ^= np.uint8(1 << (b & 7))
    return d


# ---- scenes ------------------------------------------------------------------------------------------------------------------
def mono_scene(seed, nf=1000, shift=(3, 1)):
    """Two overlapping 640x480 views through the reference's monocular Frame constructor (TUM1 calibration with distortion)."""
    R.reset()
    big = synth(640 + 16, 480 + 16, seed)
    cur = big[8:488, 8:648].copy(); last = big[8 + shift[1]:488 + shift[1], 8 + shift[0]:648 + shift[0]].copy()
    ex = R.Extractor(nf)
    Fc = R.Frame.mono(cur, ex, None, MONO_K4, MONO_DIST)
    Fl = R.Frame.mono(last, ex, None, MONO_K4, MONO_DIST)
    return cur, last, Fc, Fl, R.Frame.statics()


def stereo_scene(seed, nf=1200, shift=(4, 0)):
    """Two stereo frames (752x480, EuRoC calibration, rectified: no distortion) through the reference's stereo constructor."""
    R.reset()
    L, Rt = synth_stereo(752 + 16, 480 + 16, seed)
    crop = lambda im, dx, dy: im[8 + dy:488 + dy, 8 + dx:760 + dx].copy()
    exL, exR = R.Extractor(nf), R.Extractor(nf)
    dist = [0, 0, 0, 0]
    Fc = R.Frame.stereo(crop(L, 0, 0), crop(Rt, 0, 0), exL, exR, None, STEREO_K4, dist, STEREO_BF)
    Fl = R.Frame.stereo(crop(L, *shift), crop(Rt, *shift), exL, exR, None, STEREO_K4, dist, STEREO_BF)
    return Fc, Fl, R.Frame.statics()


def put_frame(case, name, F, st, raw=False):
    put(case, **{f"{name}_kps_un": F.kps_un, f"{name}_desc": F.desc, f"{name}_u_right": F.u_right})
    if raw:
        put(case, **{f"{name}_kps": F.kps})
    put(case, bounds=np.array([st["min_x"], st["min_y"], st["max_x"], st["max_y"]], f32), scale=F.scale, inv_sigma2=F.inv_sigma2,
        sigma2=F.sigma2)


def setup_points(rng, Fl, Fc, st, Tc, shift, K4, n_take=None, noise=0.3, map_=None, refkf=None, nobs_p=0.85):
    """One MapPoint per Last-frame feature (a subset): world position = what the Current camera (pose Tc) sees at the feature's
    undistorted position moved by `shift` (+ noise), descriptor = the feature's with a few bit flips, normal / distance range as
    MapPoint::UpdateNormalAndDepth would leave them for an observation from the Current camera at the feature's octave."""
    n = Fl.N
    take = np.arange(n) if n_take is None else np.sort(rng.choice(n, n_take, replace=False))
    uv = np.stack([Fl.kps_un["x"][take] + shift[0], Fl.kps_un["y"][take] + shift[1]], 1).astype(np.float64)
    uv += rng.normal(0, noise, uv.shape)
    z = rng.uniform(2.0, 15.0, len(take))
    X = world_points(Tc, K4, uv, z)
    Tc64 = Tc.astype(np.float64)
    Ow = -Tc64[:3, :3].T @ Tc64[:3, 3]
    PO = X.astype(np.float64) - Ow[None, :]
    dist = np.linalg.norm(PO, axis=1)
    normal = (PO / dist[:, None]).astype(f32)
    oct_ = Fl.kps["octave"][take]
    max_d = (dist * Fl.scale[oct_]).astype(f32)
    min_d = (max_d / Fl.scale[Fl.nlevels - 1]).astype(f32)
    desc = point_descriptors(rng, Fl.desc[take])
    nobs = np.where(rng.random(len(take)) < nobs_p, rng.integers(1, 6, len(take)), 0).astype(np.int32)
    mps = R.MapPoints(map_, refkf, X, desc, normal, min_d, max_d, nobs)
    return take, X, desc, normal, min_d, max_d, nobs, mps


def case_frame_basics():
    """f1 / a15: UndistortKeyPoints, ComputeImageBounds, AssignFeaturesToGrid, GetFeaturesInArea of the reference's Frame."""
    cur, last, Fc, Fl, st = mono_scene(101)
    rng = np.random.default_rng(1)
    put_frame("frame", "f", Fc, st, raw=True)
    off, idx = Fc.grid()
    nwin = 400
    win = np.zeros((nwin, 5), f32)
    win[:, 0] = rng.uniform(-60, 700, nwin); win[:, 1] = rng.uniform(-60, 540, nwin); win[:, 2] = rng.uniform(0.5, 90, nwin)
    win[:, 3] = rng.integers(-1, 6, nwin); win[:, 4] = rng.integers(-1, 8, nwin)
    win[:8, 2] = [0, 1e4, 1, 1, 50, 50, 3, 3]; win[:8, 0] = [320, 320, st["min_x"], st["max_x"], -500, 5000, 320, 320]
    win[1, 3:5] = -1; win[6:8, 3] = [3, 5]; win[6:8, 4] = [2, -1]                      # min > max: empty; max < 0: no upper bound
    res, roff = [], [0]
    for x, y, r, a, b in win:
        v = Fc.features_in_area(float(x), float(y), float(r), int(a), int(b))
        res.append(v); roff.append(roff[-1] + len(v))
    put("frame", image_seed=101, K4=np.asarray(MONO_K4, f32), dist=np.asarray(MONO_DIST, f32), grid_off=off, grid_idx=idx,
        windows=win, win_off=np.asarray(roff, np.int32), win_idx=np.concatenate(res).astype(np.int32), statics=np.array(list(st.values()), f32))


def case_stereo():
    """a14: Frame::ComputeStereoMatches as run by the reference's stereo Frame constructor (mb == 0 at the call, Frame.cpp:79-130)."""
    R.reset()
    L, Rt = synth_stereo(752, 480, 7)
    exL, exR = R.Extractor(1200), R.Extractor(1200)
    F = R.Frame.stereo(L, Rt, exL, exR, None, STEREO_K4, [0, 0, 0, 0], STEREO_BF)
    put("stereo", seed=7, w=752, h=480, nfeatures=1200, mbf=f32(STEREO_BF), kps_left=F.kps, desc_left=F.desc, kps_right=F.kps_right,
        desc_right=F.desc_right, u_right=F.u_right, depth=F.depth)
    assert (F.u_right >= 0).sum() > 300


def case_local_points():
    """a10: SearchByProjection(Frame&, const vector<MapPoint*>&, th) with the tracking fields filled by Frame::isInFrustum."""
    cur, last, Fc, Fl, st = mono_scene(102)
    rng = np.random.default_rng(2)
    Tc = pose(0.01, -0.02, 0.005, [0.05, -0.02, 0.1])
    Fc.set_pose(Tc)
    Fl.set_pose(pose(0, 0, 0, [0, 0, 0]))                      # a KeyFrame needs a pose (KeyFrame.cpp:59)
    map_ = R.Map(); kf0 = R.KeyFrame(Fl, map_)
    take, X, desc, normal, min_d, max_d, nobs, mps = setup_points(rng, Fl, Fc, st, Tc, (3, 1), MONO_K4, map_=map_, refkf=kf0)
    # half of the points are seen obliquely (viewCos < 0.998 -> radius 4.0, ORBmatcher.cpp:157-163): tilt their mean normal
    tilt = rng.random(len(take)) < 0.5
    ax = rng.normal(size=(len(take), 3)); ax -= (ax * normal).sum(1, keepdims=True) * normal; ax /= np.linalg.norm(ax, axis=1, keepdims=True)
    ang = rng.uniform(0.1, 0.6, len(take))
    n2 = (normal * np.cos(ang)[:, None] + ax * np.sin(ang)[:, None]).astype(f32)
    normal = np.where(tilt[:, None], n2, normal).astype(f32)
    mps = R.MapPoints(map_, kf0, X, desc, normal, min_d, max_d, nobs)
    mps.set_bad(7); mps.set_bad(100)
    iv, px, py, pxr, lv, vc = mps.in_frustum(Fc, 0.5)
    # frame slots that already hold a point (with and without observations)
    pre = np.full(Fc.N, -1, np.int32)
    pre_pts = R.MapPoints(map_, kf0, np.zeros((40, 3), f32), np.zeros((40, 32), np.uint8), None, None, None,
                          np.where(np.arange(40) % 3 == 0, 0, 2).astype(np.int32))
    for th, nnratio, name in [(1.0, 0.8, "local_points_th1"), (3.0, 0.8, "local_points_th3"), (5.0, 0.6, "local_points_th5")]:
        allp = R.MapPoints(map_, kf0, np.concatenate([X, np.zeros((40, 3), f32)]), np.concatenate([desc, np.zeros((40, 32), np.uint8)]),
                           np.concatenate([normal, np.zeros((40, 3), f32)]), np.concatenate([min_d, np.zeros(40, f32)]),
                           np.concatenate([max_d, np.zeros(40, f32)]), np.concatenate([nobs, np.where(np.arange(40) % 3 == 0, 0, 2)]).astype(np.int32))
        allp.set_bad(7); allp.set_bad(100)
        allp.set_track(np.concatenate([iv, np.zeros(40, np.uint8)]), np.concatenate([px, np.zeros(40, f32)]), np.concatenate([py, np.zeros(40, f32)]),
                       np.concatenate([pxr, np.zeros(40, f32)]), np.concatenate([lv, np.zeros(40, np.int32)]), np.concatenate([vc, np.zeros(40, f32)]))
        npt = len(take)
        pre[:] = -1; pre[5::25] = npt + np.arange(len(pre[5::25])) % 40
        Fc.set_mappoints(allp, pre)
        all_nobs = allp.nobs()
        with R.Trace() as t:
            n = R.search_local_points(Fc, allp, th, nnratio)
        tq = trace_queries(t.events, allp, 4)
        assert all(0 <= p < npt for p, *_ in tq) and len(tq) == int(((iv != 0) & (np.arange(npt) != 7) & (np.arange(npt) != 100)).sum())
        alldesc = np.concatenate([desc, np.zeros((40, 32), np.uint8)])
        q = make_queries(tq, alldesc, np.concatenate([pxr, np.zeros(40, f32)]), None, np.arange(allp.n), (all_nobs > 0).astype(np.int32), np.zeros(allp.n, f32))
        out = Fc.get_mappoints(allp)
        put(name, th=f32(th), nnratio=f32(nnratio), queries=q, mp0=pre.copy(), obs0=np.where(pre >= 0, all_nobs[np.maximum(pre, 0)] > 0, 0).astype(np.int32),
            mp_out=out, obs_out=np.where(out >= 0, all_nobs[np.maximum(out, 0)] > 0, 0).astype(np.int32), nmatches=n)
        assert n > 150, n
    put_frame("local_points", "f", Fc, st)
    put("local_points", in_view=iv, proj_x=px, proj_y=py, level=lv, view_cos=vc)


def stereo_depths(rng, Fc, uv, frac=0.7):
    """Depths that agree with the stereo disparity of the Current feature nearest to each target pixel (for `frac` of the points),
    so that the u_right consistency gates (ORBmatcher.cpp:116-121, :1595-1601) pass for some candidates and cut others."""
    z = rng.uniform(2.0, 15.0, len(uv))
    xy = np.stack([Fc.kps_un["x"], Fc.kps_un["y"]], 1)
    for i in range(len(uv)):
        j = int(np.argmin(((xy - uv[i]) ** 2).sum(1)))
        if Fc.u_right[j] > 0 and rng.random() < frac:
            d = float(Fc.kps_un["x"][j] - Fc.u_right[j])
            if d > 0.5:
                z[i] = float(Fc.mbf) / d
    return z


def combined_points(map_, kf0, X, desc, normal, min_d, max_d, nobs, n_extra, extra_nobs):
    """The scene's points followed by n_extra featureless ones (what frame slots hold before the call)."""
    z3 = np.zeros((n_extra, 3), f32)
    return R.MapPoints(map_, kf0, np.concatenate([X, z3]), np.concatenate([desc, np.zeros((n_extra, 32), np.uint8)]), np.concatenate([normal, z3]),
                       np.concatenate([min_d, np.zeros(n_extra, f32)]), np.concatenate([max_d, np.zeros(n_extra, f32)]),
                       np.concatenate([nobs, extra_nobs]).astype(np.int32))


def K4_of(st):
    return np.array([st["fx"], st["fy"], st["cx"], st["cy"]], f32)


def case_last_frame():
    """a11: SearchByProjection(Frame& Current, const Frame& Last, th, bMono): monocular scene, and stereo scenes moving forward / backward."""
    for name, stereo, tz, th, check in [("last_frame_mono", False, 0.1, 15.0, True), ("last_frame_mono_th7", False, 0.1, 7.0, False),
                                        ("last_frame_fwd", True, -0.5, 7.0, True), ("last_frame_bwd", True, 0.5, 15.0, True),
                                        ("last_frame_still", True, 0.02, 15.0, True)]:
        rng = np.random.default_rng(300 + len(name) * 7 + int(th))
        if stereo:
            Fc, Fl, st = stereo_scene(104); shift = (4, 0); K4 = STEREO_K4
        else:
            _, _, Fc, Fl, st = mono_scene(103); shift = (3, 1); K4 = MONO_K4
        Tc = pose(0.004, -0.006, 0.003, [0.03, -0.01, tz]); Tl = pose(0, 0, 0, [0, 0, 0])
        Fc.set_pose(Tc); Fl.set_pose(Tl)
        map_ = R.Map(); kf0 = R.KeyFrame(Fl, map_)
        n = Fl.N
        take = np.sort(rng.choice(n, int(0.85 * n), replace=False))
        uv = np.stack([Fl.kps_un["x"][take] + shift[0], Fl.kps_un["y"][take] + shift[1]], 1).astype(np.float64) + rng.normal(0, 0.3, (len(take), 2))
        z = stereo_depths(rng, Fc, uv) if stereo else rng.uniform(2.0, 15.0, len(take))
        z[:3] = [-4.0, -1.0, 1e-3]                                  # behind the camera / absurdly close: invzc < 0 and out-of-bounds gates
        X = world_points(Tc, K4, uv, z)
        desc = point_descriptors(rng, Fl.desc[take])
        npt = len(take)
        nobs = np.where(rng.random(npt) < 0.85, rng.integers(1, 6, npt), 0).astype(np.int32)
        n_extra = 40
        extra_nobs = np.where(np.arange(n_extra) % 3 == 0, 0, 2)
        zf = np.zeros(npt, f32)
        allp = combined_points(map_, kf0, X, desc, np.zeros((npt, 3), f32), zf, zf, nobs, n_extra, extra_nobs)
        all_nobs = allp.nobs()
        idx_l = np.full(n, -1, np.int32); idx_l[take] = np.arange(npt)
        outl = np.zeros(n, np.uint8); outl[take[10::37]] = 1
        Fl.set_mappoints(allp, idx_l, outl)
        pre = np.full(Fc.N, -1, np.int32); pre[5::25] = npt + np.arange(len(pre[5::25])) % n_extra
        Fc.set_mappoints(allp, pre)
        with R.Trace() as t:
            nm = R.search_last_frame(Fc, Fl, th, not stereo, 0.9, check)
        tq = trace_queries(t.events, allp, 3)
        K4f = K4_of(st)
        Xc, invz, u, v = project_frame(Tc, K4f, X)
        ok = (outl[take] == 0) & ~(invz < 0) & ~(u < st["min_x"]) & ~(u > st["max_x"]) & ~(v < st["min_y"]) & ~(v > st["max_y"])
        assert [p for p, *_ in tq] == list(np.flatnonzero(ok)), name
        tu = np.array([e[1] for e in tq], f32); tv = np.array([e[2] for e in tq], f32)
        assert np.array_equal(tu.view(np.uint32), u[ok].view(np.uint32)) and np.array_equal(tv.view(np.uint32), v[ok].view(np.uint32)), name
        u_right = (u - Fc.mbf * invz).astype(f32)
        ang = Fl.kps_un["angle"][take]
        q = make_queries(tq, desc, u_right, None, np.arange(npt), (nobs > 0).astype(np.int32), ang)
        out = Fc.get_mappoints(allp)
        put(name, th=f32(th), check_ori=int(check), mono=int(not stereo), queries=q, mp0=pre.copy(),
            obs0=np.where(pre >= 0, all_nobs[np.maximum(pre, 0)] > 0, 0).astype(np.int32), mp_out=out,
            obs_out=np.where(out >= 0, all_nobs[np.maximum(out, 0)] > 0, 0).astype(np.int32), nmatches=nm,
            level_modes=np.unique(np.array([(e[4] >= 0) * 2 + (e[5] >= 0) for e in tq])))
        put_frame(name, "f", Fc, st)
        assert nm > 200, (name, nm)


def kf_scene(seed, stereo, rng, Tc, Tl, nobs_p=0.85, noise=0.3, frac_points=0.85):
    """Current / Last frames with poses, a KeyFrame of each, and one MapPoint per (most) Last features placed where the Current
    camera sees the feature's position + shift.  Returns everything the KeyFrame-based searches need."""
    if stereo:
        Fc, Fl, st = stereo_scene(seed); shift = (4, 0); K4 = STEREO_K4
    else:
        _, _, Fc, Fl, st = mono_scene(seed); shift = (3, 1); K4 = MONO_K4
    Fc.set_pose(Tc); Fl.set_pose(Tl)
    map_ = R.Map()
    n = Fl.N
    take = np.sort(rng.choice(n, int(frac_points * n), replace=False))
    uv = np.stack([Fl.kps_un["x"][take] + shift[0], Fl.kps_un["y"][take] + shift[1]], 1).astype(np.float64) + rng.normal(0, noise, (len(take), 2))
    z = stereo_depths(rng, Fc, uv) if stereo else rng.uniform(2.0, 15.0, len(take))
    X = world_points(Tc, K4, uv, z)
    Tc64 = Tc.astype(np.float64)
    Ow = -Tc64[:3, :3].T @ Tc64[:3, 3]
    PO = X.astype(np.float64) - Ow[None, :]
    dist = np.linalg.norm(PO, axis=1)
    normal = (PO / dist[:, None]).astype(f32)
    oct_ = Fl.kps["octave"][take]
    max_d = (dist * Fl.scale[oct_]).astype(f32)
    min_d = (max_d / Fl.scale[Fl.nlevels - 1]).astype(f32)
    # gates: some points out of their scale-invariance range, some seen from behind (normal test of :419-421, :1021-1023)
    max_d[3::41] *= f32(0.3); normal[5::43] *= f32(-1)
    desc = point_descriptors(rng, Fl.desc[take])
    nobs = np.where(rng.random(len(take)) < nobs_p, rng.integers(1, 6, len(take)), 0).astype(np.int32)
    return dict(Fc=Fc, Fl=Fl, st=st, K4=K4, map=map_, take=take, X=X, desc=desc, normal=normal, min_d=min_d, max_d=max_d, nobs=nobs, shift=shift)


def case_reloc():
    """a12: SearchByProjection(Frame& Current, KeyFrame*, const set<MapPoint*>& sAlreadyFound, th, ORBdist)."""
    for name, th, orbdist, check in [("reloc_th10", 10.0, 100, True), ("reloc_th3", 3.0, 64, True), ("reloc_nocheck", 10.0, 100, False)]:
        rng = np.random.default_rng(400 + int(th) + orbdist)
        Tc = pose(0.006, 0.004, -0.003, [0.02, 0.01, 0.08]); Tl = pose(0, 0, 0, [0, 0, 0])
        S = kf_scene(105, False, rng, Tc, Tl)
        Fc, Fl, st = S["Fc"], S["Fl"], S["st"]
        npt = len(S["take"]); n_extra = 40
        kf = R.KeyFrame(Fl, S["map"])
        allp = combined_points(S["map"], kf, S["X"], S["desc"], S["normal"], S["min_d"], S["max_d"], S["nobs"], n_extra, np.ones(n_extra))
        allp.set_bad(11); allp.set_bad(200)
        idx_l = np.full(Fl.N, -1, np.int32); idx_l[S["take"]] = np.arange(npt)
        kf.set_mappoints(allp, idx_l, add_observation=False)
        already = np.arange(4, npt, 29).astype(np.int32)
        pre = np.full(Fc.N, -1, np.int32); pre[2::13] = npt + np.arange(len(pre[2::13])) % n_extra
        Fc.set_mappoints(allp, pre)
        with R.Trace() as t:
            nm = R.search_reloc(Fc, kf, allp, already, th, orbdist, 0.9, check)
        tq = trace_queries(t.events, allp, 3)
        K4f = K4_of(st)
        Xc, invz, u, v = project_frame(Tc, K4f, S["X"])
        sel = np.array([p for p, *_ in tq])
        tu = np.array([e[1] for e in tq], f32); tv = np.array([e[2] for e in tq], f32)
        assert np.array_equal(tu.view(np.uint32), u[sel].view(np.uint32)) and np.array_equal(tv.view(np.uint32), v[sel].view(np.uint32))
        assert all(e[4] == e[6] - 1 and e[5] == e[6] + 1 for e in tq)            # GetFeaturesInArea(u, v, radius, p-1, p+1)
        assert len(set(sel) & set(already.tolist())) == 0 and 11 not in sel and len(sel) < npt - len(already) - 2
        ang = Fl.kps_un["angle"][S["take"]]
        q = make_queries(tq, S["desc"], np.full(npt, -1, f32), None, np.arange(npt), np.ones(npt, np.int32), ang)
        out = Fc.get_mappoints(allp)
        put(name, th=f32(th), orb_dist=orbdist, check_ori=int(check), queries=q, mp0=pre.copy(), mp_out=out, nmatches=nm)
        put_frame(name, "f", Fc, st)
        assert nm > 150, (name, nm)


def sim3_of(T, s):
    S = T.astype(np.float64).copy()
    S[:3, :3] *= s; S[:3, 3] *= s
    return S.astype(f32)


def case_loop():
    """a12: SearchByProjection(KeyFrame*, cv::Mat Scw, const vector<MapPoint*>&, vector<MapPoint*>& vpMatched, th)."""
    for name, th, s in [("loop_th10", 10, 1.0), ("loop_th4_scaled", 4, 1.13)]:
        rng = np.random.default_rng(500 + th)
        Tc = pose(-0.005, 0.007, 0.002, [0.01, -0.03, 0.05]); Tl = pose(0, 0, 0, [0, 0, 0])
        S = kf_scene(106, False, rng, Tc, Tl)
        Fc, st = S["Fc"], S["st"]
        npt = len(S["take"]); n_extra = 40
        kf = R.KeyFrame(Fc, S["map"])
        allp = combined_points(S["map"], kf, S["X"], S["desc"], S["normal"], S["min_d"], S["max_d"], S["nobs"], n_extra, np.ones(n_extra))
        allp.set_bad(9)
        m0 = np.full(kf.N, -1, np.int32); m0[3::11] = npt + np.arange(len(m0[3::11])) % n_extra
        m0[7::53] = np.arange(len(m0[7::53])) * 17 % npt                    # some scene points are already matched: skipped (:375-377, :385)
        Scw = sim3_of(Tc, s)
        with R.Trace() as t:
            nm, m = R.search_loop(kf, Scw, allp, m0, th)
        tq = trace_queries(t.events, allp, 3)
        assert all(p < npt for p, *_ in tq) and len(tq) < npt
        tq = [(p, u, v, r, lvl - 1, lvl, lvl) for p, u, v, r, a, b, lvl in tq]   # the [p-1, p] gate of :440-441
        q = make_queries(tq, S["desc"], np.full(npt, -1, f32), None, np.arange(npt), np.ones(npt, np.int32), np.zeros(npt, f32))
        put(name, th=th, Scw=Scw, queries=q, matched0=m0, matched_out=m, nmatches=nm)
        put_frame(name, "f", Fc, st)
        assert nm > 150, (name, nm)


def case_fuse():
    """a12b: Fuse(KeyFrame*, const vector<MapPoint*>&, th) on a stereo KeyFrame (chi-square gates with and without u_right) and
    Fuse(KeyFrame*, cv::Mat Scw, ...) on a monocular one.  Expected bestIdx per point = the index the reference passes to
    KeyFrame::GetMapPoint right after the search (ORBmatcher.cpp:1094, :1231)."""
    for name, stereo, th, s in [("fuse_stereo", True, 3.0, None), ("fuse_mono", False, 3.0, None), ("fuse_sim3", False, 4.0, 1.07)]:
        rng = np.random.default_rng(600 + len(name))
        Tc = pose(0.003, -0.004, 0.006, [-0.02, 0.02, 0.06]); Tl = pose(0, 0, 0, [0, 0, 0])
        S = kf_scene(107 if stereo else 108, stereo, rng, Tc, Tl, noise=0.8)
        Fc, st = S["Fc"], S["st"]
        npt = len(S["take"])
        kf = R.KeyFrame(Fc, S["map"])
        pts = R.MapPoints(S["map"], kf, S["X"], S["desc"], S["normal"], S["min_d"], S["max_d"], S["nobs"])
        pts.set_bad(13)
        # the KeyFrame's own points on every third feature (Replace paths), with real observations
        own_idx = np.full(kf.N, -1, np.int32); own_idx[::3] = np.arange(len(own_idx[::3]))
        nown = int((own_idx >= 0).sum())
        own = R.MapPoints(S["map"], kf, np.zeros((nown, 3), f32), Fc.desc[::3], None, None, None, None)
        kf.set_mappoints(own, own_idx, add_observation=True)
        K4f = K4_of(st)
        with R.Trace() as t:
            if s is None:
                nm = R.fuse(kf, pts, th)
            else:
                nm, rep = R.fuse_sim3(kf, sim3_of(Tc, s), pts, own, th)
        tq = trace_queries(t.events, pts, 3)
        best = best_idx_from_trace(t.events, pts, kf.h)
        assert int((best >= 0).sum()) == nm, (name, nm, int((best >= 0).sum()))
        sel = np.array([p for p, *_ in tq])
        if s is None:
            Rcw, tcw, _ = kf.pose()
            Xc, invz, u, v = project_kf(Rcw, tcw, K4f, S["X"])
            tu = np.array([e[1] for e in tq], f32); tv = np.array([e[2] for e in tq], f32)
            assert np.array_equal(tu.view(np.uint32), u[sel].view(np.uint32)) and np.array_equal(tv.view(np.uint32), v[sel].view(np.uint32)), name
            u_right = (u - Fc.mbf * invz).astype(f32)
        else:
            u_right = np.full(npt, -1, f32)
            assert np.array_equal(np.flatnonzero(rep >= 0), np.flatnonzero((best >= 0) & (own_idx[np.maximum(best, 0)] >= 0)))
            assert np.array_equal(rep[rep >= 0], own_idx[best[rep >= 0]])
        tq = [(p, u_, v_, r, lvl - 1, lvl, lvl) for p, u_, v_, r, a, b, lvl in tq]
        q = make_queries(tq, S["desc"], u_right, None, np.arange(npt), np.ones(npt, np.int32), np.zeros(npt, f32))
        put(name, th=f32(th), gate=int(s is None), max_dist=50, queries=q, best_idx=best[sel], nfused=nm)
        put_frame(name, "f", Fc, st)
        assert nm > 100, (name, nm)


def case_sim3():
    """a12b: SearchBySim3(pKF1, pKF2, vpMatches12, s12, R12, t12, th): both directions + the mutual check."""
    rng = np.random.default_rng(700)
    T1 = pose(0, 0, 0, [0, 0, 0]); T2 = pose(0.004, -0.003, 0.002, [0.03, 0.01, 0.04])
    S = kf_scene(109, False, rng, T2, T1, noise=0.3)           # "Current" = KF2, "Last" = KF1
    F2, F1, st = S["Fc"], S["Fl"], S["st"]
    kf1 = R.KeyFrame(F1, S["map"]); kf2 = R.KeyFrame(F2, S["map"])
    n1p = len(S["take"])
    # KF1's points: the scene points (seen by KF2 at +shift).  KF2's points: one per (most) KF2 features, placed where KF1 sees
    # that feature's position - shift.
    take2 = np.sort(rng.choice(F2.N, int(0.85 * F2.N), replace=False))
    uv2 = np.stack([F2.kps_un["x"][take2] - S["shift"][0], F2.kps_un["y"][take2] - S["shift"][1]], 1).astype(np.float64) + rng.normal(0, 0.3, (len(take2), 2))
    X2 = world_points(T1, S["K4"], uv2, rng.uniform(2.0, 15.0, len(take2)))
    d2 = np.linalg.norm(X2.astype(np.float64), axis=1)          # KF1 sits at the origin
    oct2 = F2.kps["octave"][take2]
    max2 = (d2 * F2.scale[oct2]).astype(f32); min2 = (max2 / F2.scale[F2.nlevels - 1]).astype(f32)
    desc2 = point_descriptors(rng, F2.desc[take2])
    n2p = len(take2)
    zf3 = np.zeros((n2p, 3), f32)
    allp = R.MapPoints(S["map"], kf1, np.concatenate([S["X"], X2]), np.concatenate([S["desc"], desc2]), np.concatenate([S["normal"], zf3]),
                       np.concatenate([S["min_d"], min2]), np.concatenate([S["max_d"], max2]), None)
    allp.set_bad(21); allp.set_bad(n1p + 17)
    i1 = np.full(kf1.N, -1, np.int32); i1[S["take"]] = np.arange(n1p)
    i2 = np.full(kf2.N, -1, np.int32); i2[take2] = n1p + np.arange(n2p)
    kf1.set_mappoints(allp, i1, True); kf2.set_mappoints(allp, i2, True)
    # a few pairs are already matched (vbAlreadyMatched1 / 2, :1280-1291)
    m0 = np.full(kf1.N, -1, np.int32)
    pre1 = S["take"][30::97]; m0[pre1] = n1p + (np.arange(len(pre1)) * 13) % n2p
    T12 = T1.astype(np.float64) @ np.linalg.inv(T2.astype(np.float64))
    s12 = 1.0; th = 7.5
    with R.Trace() as t:
        nm, m = R.search_by_sim3(kf1, kf2, allp, m0, s12, T12[:3, :3].astype(f32), T12[:3, 3].astype(f32), th)
    alldesc = np.concatenate([S["desc"], desc2])
    z = np.zeros(allp.n, f32)
    qs = {}
    for tag, obj in (("q12", kf2.h), ("q21", kf1.h)):
        tq = trace_queries(t.events, allp, 3, gfa_obj=obj)
        tq = [(p, u, v, r, lvl - 1, lvl, lvl) for p, u, v, r, a, b, lvl in tq]
        qs[tag] = make_queries(tq, alldesc, np.full(allp.n, -1, f32), None, np.arange(allp.n), np.ones(allp.n, np.int32), z)
    # point index -> feature index maps so the test can run the mutual check of :1478-1495 on flattened arrays
    feat_of_point = np.full(allp.n, -1, np.int32); feat_of_point[i1[i1 >= 0]] = np.flatnonzero(i1 >= 0); feat_of_point[i2[i2 >= 0]] = np.flatnonzero(i2 >= 0)
    put("sim3", th=f32(th), q12=qs["q12"], q21=qs["q21"], feat_of_point=feat_of_point, matches0=m0, matches_out=m, nmatches=nm, n1=kf1.N, n2=kf2.N,
        kf1_kps_un=F1.kps_un, kf1_desc=F1.desc, kf2_kps_un=F2.kps_un, kf2_desc=F2.desc,
        bounds=np.array([st["min_x"], st["min_y"], st["max_x"], st["max_y"]], f32), n1_points=n1p)
    assert nm > 100, nm


def case_initialization():
    """a13: SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize)."""
    for name, win, nnratio, check in [("init_w100", 100, 0.9, True), ("init_w30_nocheck", 30, 0.8, False)]:
        R.reset()
        big = synth(640 + 16, 480 + 16, 110)
        ex = R.Extractor(2000)
        F1 = R.Frame.mono(big[8:488, 8:648].copy(), ex, None, MONO_K4, MONO_DIST)
        F2 = R.Frame.mono(big[4:484, 14:654].copy(), ex, None, MONO_K4, MONO_DIST)
        st = R.Frame.statics()
        prev = np.stack([F1.kps_un["x"], F1.kps_un["y"]], 1).astype(f32)
        nm, m12, pm = R.search_for_initialization(F1, F2, prev, win, nnratio, check)
        put(name, window=win, nnratio=f32(nnratio), check_ori=int(check), f1_kps_un=F1.kps_un, f1_desc=F1.desc, f2_kps_un=F2.kps_un, f2_desc=F2.desc,
            bounds=np.array([st["min_x"], st["min_y"], st["max_x"], st["max_y"]], f32), prev_in=prev, matches12=m12, prev_out=pm, nmatches=nm)
        assert nm > 100, (name, nm)



# ---- vocabulary, BoW, triangulation, distinctive descriptors -----------------------------------------------------------------------
def build_vocab(pool, k, L, rng, stop_fraction=0.03):
    """A ragged vocabulary tree over real descriptors, in loadFromTextFile order: a node's children are k of the descriptors assigned
    to it (padded with perturbed copies), the pool is split by nearest child, nodes with < 2 descriptors become leaves early."""
    bits = np.unpackbits(pool, axis=1).astype(np.int16)
    parent = [0]; leaf = [0]; desc = [np.zeros(32, np.uint8)]; weight = [0.0]
    frontier = [(0, np.arange(len(pool)), 0)]
    while frontier:
        nxt = []
        for node, members, lvl in frontier:
            cent = []
            pick = rng.permutation(members)[:k] if len(members) else []
            for m in pick:
                cent.append(pool[m].copy())
            while len(cent) < k:
                d = (cent[0] if cent else desc[node]).copy()
                for b in rng.integers(0, 256, 20):
                    d[b >> 3] ^= np.uint8(1 << (b & 7))
                cent.append(d)
            cb = np.unpackbits(np.stack(cent), axis=1).astype(np.int16)
            assign = np.argmin((bits[members][:, None, :] != cb[None, :, :]).sum(2), axis=1) if len(members) else np.zeros(0, np.int64)
            for c in range(k):
                nid = len(parent)
                mem = members[assign == c]
                is_leaf = lvl + 1 == L or (lvl + 1 >= 2 and len(mem) < 2)
                parent.append(node); desc.append(cent[c]); leaf.append(1 if is_leaf else 0)
                weight.append(0.0 if not is_leaf else (0.0 if rng.random() < stop_fraction else float(rng.uniform(0.5, 9.5))))
                if not is_leaf:
                    nxt.append((nid, mem, lvl + 1))
        frontier = nxt
    return np.asarray(parent, np.int32), np.asarray(leaf, np.uint8), np.stack(desc).astype(np.uint8), np.asarray(weight, np.float64)


def write_vocab_txt(path, k, L, parent, leaf, desc, weight):
    """The text format of TemplatedVocabulary::loadFromTextFile (:1362-1448): header "k L scoring weighting", then one line per node
    "parent isLeaf d0 .. d31 weight"; no trailing newline (the loader's eof loop would add a bogus node)."""
    lines = [f"{k} {L} 0 0"]
    for i in range(1, len(parent)):
        lines.append(f"{parent[i]} {leaf[i]} " + " ".join(str(int(b)) for b in desc[i]) + f" {float(weight[i])!r}")
    with open(path, "w") as f:
        f.write("\n".join(lines))


def bow_scene(seed, stereo, k=4, L=6):
    """Frames with a vocabulary built from their own descriptors; returns the frames (BoW computed by the reference) and the tree."""
    rng = np.random.default_rng(seed)
    if stereo:
        Fc0, Fl0, _ = stereo_scene(seed)
    else:
        _, _, Fc0, Fl0, _ = mono_scene(seed)
    pool = np.concatenate([Fc0.desc, Fl0.desc])
    parent, leaf, desc, weight = build_vocab(pool, k, L, rng)
    tmp = os.path.join(tempfile.gettempdir(), f"orbb_voc_{seed}.txt")
    write_vocab_txt(tmp, k, L, parent, leaf, desc, weight)
    voc = R.Vocabulary(tmp)
    assert voc.size() == int(leaf.sum())
    return voc, (k, L, parent, leaf, desc, weight), rng


def frames_with_voc(seed, stereo, voc):
    R.reset()
    if stereo:
        L_, Rt = synth_stereo(752 + 16, 480 + 16, seed)
        crop = lambda im, dx, dy: im[8 + dy:488 + dy, 8 + dx:760 + dx].copy()
        exL, exR = R.Extractor(1200), R.Extractor(1200)
        Fc = R.Frame.stereo(crop(L_, 0, 0), crop(Rt, 0, 0), exL, exR, voc, STEREO_K4, [0, 0, 0, 0], STEREO_BF)
        Fl = R.Frame.stereo(crop(L_, 4, 0), crop(Rt, 4, 0), exL, exR, voc, STEREO_K4, [0, 0, 0, 0], STEREO_BF)
    else:
        big = synth(640 + 16, 480 + 16, seed)
        ex = R.Extractor(1000)
        Fc = R.Frame.mono(big[8:488, 8:648].copy(), ex, voc, MONO_K4, MONO_DIST)
        Fl = R.Frame.mono(big[9:489, 11:651].copy(), ex, voc, MONO_K4, MONO_DIST)
    return Fc, Fl, R.Frame.statics()


def put_fv(case, name, bow):
    ids, vals, fn, fo, ff = bow
    put(case, **{f"{name}_bow_ids": ids, f"{name}_bow_vals": vals, f"{name}_fv_nodes": fn, f"{name}_fv_off": fo, f"{name}_fv_feat": ff})


def case_vocab():
    """f2: TemplatedVocabulary::transform (single feature, several levelsup) and Frame::ComputeBoW (transform of a whole frame)."""
    voc, (k, L, parent, leaf, desc, weight), rng = bow_scene(111, False)
    Fc, Fl, st = frames_with_voc(111, False, voc)
    put("vocab", k=k, L=L, parent=parent, is_leaf=leaf, desc=desc, weight=weight, feats=Fc.desc)
    for lu in (0, 2, 4, 7):
        wid, w, nid = voc.transform(Fc.desc, lu)
        put("vocab", **{f"word_lu{lu}": wid, f"weight_lu{lu}": w, f"node_lu{lu}": nid})
    put_fv("vocab", "frame", Fc.compute_bow())
    put_fv("vocab", "direct_lu2", voc.bow(Fc.desc, 2))
    assert len(OUT["vocab/frame_fv_nodes"]) >= 4


def case_bow():
    """f2: SearchByBoW(KeyFrame*, Frame&, ...) and SearchByBoW(KeyFrame*, KeyFrame*, ...) with BoW vectors computed by the reference."""
    voc, (k, L, parent, leaf, desc, weight), rng = bow_scene(112, False)
    Fc, Fl, st = frames_with_voc(112, False, voc)
    bow_c = Fc.compute_bow(); bow_l = Fl.compute_bow()
    Fc.set_pose(pose(0, 0, 0, [0, 0, 0])); Fl.set_pose(pose(0, 0, 0, [0, 0, 0]))
    map_ = R.Map()
    kfl = R.KeyFrame(Fl, map_); kfc = R.KeyFrame(Fc, map_)
    nl, nc = Fl.N, Fc.N
    pts = R.MapPoints(map_, kfl, np.zeros((nl + nc, 3), f32), np.concatenate([Fl.desc, Fc.desc]), None, None, None, None)
    il = np.where(rng.random(nl) < 0.8, np.arange(nl), -1).astype(np.int32)
    ic = np.where(rng.random(nc) < 0.8, nl + np.arange(nc), -1).astype(np.int32)
    kfl.set_mappoints(pts, il, True); kfc.set_mappoints(pts, ic, True)
    for b in (il[il >= 0][5::40]).tolist() + (ic[ic >= 0][7::40]).tolist():
        pts.set_bad(b)
    bad = pts.is_bad()
    has_l = ((il >= 0) & (bad[np.maximum(il, 0)] == 0)).astype(np.uint8)
    has_c = ((ic >= 0) & (bad[np.maximum(ic, 0)] == 0)).astype(np.uint8)
    put("bow", k=k, L=L, parent=parent, is_leaf=leaf, desc=desc, weight=weight, kf_desc=Fl.desc, kf_angle=Fl.kps_un["angle"], kf_has_mp=has_l,
        f_desc=Fc.desc, f_angle=Fc.kps["angle"], f_angle_un=Fc.kps_un["angle"], f_has_mp=has_c)
    put_fv("bow", "kf", bow_l); put_fv("bow", "f", bow_c)
    for nnratio, check in ((0.7, True), (0.9, False)):
        n, m = R.search_by_bow_kf_frame(kfl, Fc, pts, nnratio, check)
        assert np.all((m == -1) | ((m >= 0) & (m < nl)))             # KeyFrame Fl's points are 0 .. nl-1 = its feature indices
        put("bow", **{f"kf_f_matches_{int(nnratio * 10)}_{int(check)}": m, f"kf_f_n_{int(nnratio * 10)}_{int(check)}": n})
        assert n > 100, n
        n, m = R.search_by_bow_kf_kf(kfl, kfc, pts, nnratio, check)
        m = np.where(m >= 0, m - nl, -1).astype(np.int32)           # KeyFrame Fc's point nl + j sits on its feature j
        put("bow", **{f"kf_kf_matches_{int(nnratio * 10)}_{int(check)}": m, f"kf_kf_n_{int(nnratio * 10)}_{int(check)}": n})
        assert n > 100, n


def skew(t):
    return np.array([[0, -t[2], t[1]], [t[2], 0, -t[0]], [-t[1], t[0], 0]])


def case_triangulation():
    """f2: SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo) (:755-927) — monocular KeyFrames, and stereo ones
    with and without bOnlyStereo.  F12 as LocalMapping::ComputeF12 builds it (K1^-T [t12]x R12 K2^-1)."""
    for name, stereo, only in [("triang_mono", False, False), ("triang_stereo", True, False), ("triang_only_stereo", True, True)]:
        voc, (k, L, parent, leaf, desc, weight), rng = bow_scene(113 if not stereo else 114, stereo)
        Fc, Fl, st = frames_with_voc(113 if not stereo else 114, stereo, voc)
        b1 = Fl.compute_bow(); b2 = Fc.compute_bow()
        K4 = STEREO_K4 if stereo else MONO_K4
        shift = (4, 0) if stereo else (3, 1)
        Z = 8.0
        T1 = pose(0, 0, 0, [0, 0, 0])
        # Current (KF2) sees every Last (KF1) feature at +shift: camera 2 sits at -shift*Z/f of camera 1, plus a small advance so the epipole is finite
        T2 = pose(0, 0, 0, [shift[0] * Z / K4[0], shift[1] * Z / K4[1], -0.02])
        Fl.set_pose(T1); Fc.set_pose(T2)
        map_ = R.Map()
        kf1 = R.KeyFrame(Fl, map_); kf2 = R.KeyFrame(Fc, map_)
        n1, n2 = Fl.N, Fc.N
        pts = R.MapPoints(map_, kf1, np.zeros((n1 + n2, 3), f32), np.concatenate([Fl.desc, Fc.desc]), None, None, None, None)
        i1 = np.where(rng.random(n1) < 0.3, np.arange(n1), -1).astype(np.int32)       # features that already have a MapPoint are skipped
        i2 = np.where(rng.random(n2) < 0.3, n1 + np.arange(n2), -1).astype(np.int32)
        kf1.set_mappoints(pts, i1, True); kf2.set_mappoints(pts, i2, True)
        R1, t1, O1 = kf1.pose(); R2, t2, O2 = kf2.pose()
        R12 = R1.astype(np.float64) @ R2.astype(np.float64).T
        t12 = -R12 @ t2.astype(np.float64) + t1.astype(np.float64)
        Km = np.array([[K4[0], 0, K4[2]], [0, K4[1], K4[3]], [0, 0, 1]])
        F12 = (np.linalg.inv(Km).T @ skew(t12) @ R12 @ np.linalg.inv(Km)).astype(f32)
        n, pairs = R.search_for_triangulation(kf1, kf2, F12, only, 0.6, False)
        # the epipole of KeyFrame 1 in KeyFrame 2 (:763-769), float32 in the reference's order
        C2 = (mat3(R2, O1[None, :])[0] + t2).astype(f32)
        invz = f32(1.0) / C2[2]
        ex = f32(f32(f32(st["fx"] * C2[0]) * invz) + st["cx"]); ey = f32(f32(f32(st["fy"] * C2[1]) * invz) + st["cy"])
        m12 = np.full(n1, -1, np.int32); m12[pairs[:, 0]] = pairs[:, 1]
        put(name, k=k, L=L, only_stereo=int(only), F12=F12, ex=ex, ey=ey, match12=m12, pairs=pairs, nmatches=n,
            kf1_desc=Fl.desc, kf1_kps_un=Fl.kps_un, kf1_u_right=Fl.u_right, kf1_has_mp=(i1 >= 0).astype(np.uint8),
            kf2_desc=Fc.desc, kf2_kps_un=Fc.kps_un, kf2_u_right=Fc.u_right, kf2_has_mp=(i2 >= 0).astype(np.uint8),
            scale=Fc.scale, sigma2=Fc.sigma2)
        put_fv(name, "kf1", b1); put_fv(name, "kf2", b2)
        assert n > 40, (name, n)


def case_distinctive():
    """f3: MapPoint::ComputeDistinctiveDescriptors (MapPoint.cpp:302-380) on points observed in 1 .. 9 KeyFrames; rows are listed in
    the order the reference walks mObservations (a std::map keyed by KeyFrame*, i.e. by address)."""
    R.reset()
    rng = np.random.default_rng(115)
    ex = R.Extractor(500)
    big = synth(640 + 32, 480 + 32, 115)
    map_ = R.Map()
    kfs = []
    for i in range(9):
        F = R.Frame.mono(big[8 + i:488 + i, 8 + 2 * i:648 + 2 * i].copy(), ex, None, MONO_K4, [0, 0, 0, 0])
        F.set_pose(pose(0, 0, 0, [0, 0, 0]))
        kfs.append((R.KeyFrame(F, map_), F))
    npts = 300
    pts = R.MapPoints(map_, kfs[0][0], np.zeros((npts, 3), f32), np.zeros((npts, 32), np.uint8), None, None, None, None)
    for p in range(npts):
        nk = int(rng.integers(1, 10))
        for j in rng.permutation(9)[:nk]:
            kf, F = kfs[j]
            # the feature nearest to where the point would be in that view (views are shifted crops), or a random one (a wrong association)
            idx = int(rng.integers(0, F.N)) if rng.random() < 0.15 else int(np.argmin((F.kps["x"] - (F.kps["x"][p % F.N] if j == 0 else kfs[0][1].kps["x"][p % kfs[0][1].N] - 2 * j)) ** 2 +
                                                                                          (F.kps["y"] - (kfs[0][1].kps["y"][p % kfs[0][1].N] - j)) ** 2))
            pts.add_observation(p, kf, idx)
    kf_of = {kf.h: j for j, (kf, F) in enumerate(kfs)}
    rows, off, want = [], [0], []
    for p in range(npts):
        okfs, oidx = pts.observations(p)
        for kfp, fi in zip(okfs, oidx):
            rows.append(kfs[kf_of[int(kfp)]][1].desc[fi])
        off.append(len(rows))
        want.append(pts.distinctive(p))
    rows = np.stack(rows); want = np.stack(want)
    put("distinctive", rows=rows, offsets=np.asarray(off, np.int64), best_desc=want)


CASES = [case_frame_basics, case_stereo, case_local_points, case_last_frame, case_reloc, case_loop, case_fuse, case_sim3, case_initialization, case_vocab, case_bow,
         case_triangulation, case_distinctive]


def main():
    for c in CASES:
        c()
        print(f"{c.__name__}: ok")
    path = os.path.join(HERE, "ref_match.npz")
    np.savez_compressed(path, **OUT)
    print(f"wrote {path}: {len(OUT)} arrays, {os.path.getsize(path) / 1024:.0f} KiB")


if __name__ == "__main__":
    main()
