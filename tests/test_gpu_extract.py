"""GPU parity of the extractor: every stage of ORBextractor::operator() against the CPU oracle, through the C ABI."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SHAPES = [  # (w, h, nfeatures, seed)  C1 / C2 / C3 shapes of BASELINE.json
    (640, 480, 1000, 1),
    (1241, 376, 2000, 2),
    (752, 480, 1200, 3),
]


def _sorted_rows(a):
    a = np.asarray(a).reshape(-1, 3)
    return a[np.lexsort((a[:, 2], a[:, 0], a[:, 1]))]


def run_pair(w, h, nf, seed, **kw):
    import orbb200, pyoracle
    from synth import synth
    img = synth(w, h, seed)
    orc = pyoracle.OracleExtractor(nf, **kw)
    ok, od = orc.extract(img)
    ex = orbb200.ORBextractor(nf, max_size=(w, h), **{k: v for k, v in kw.items() if k in ()})
    gk, gd = ex(img)
    return img, orc, ok, od, ex, gk, gd


@pytest.mark.parametrize("w,h,nf,seed", SHAPES)
def test_stage_parity(gpu, w, h, nf, seed):
    img, orc, ok, od, ex, gk, gd = run_pair(w, h, nf, seed)
    for L in range(8):
        assert np.array_equal(ex.pyramid_level(L), orc.level_image(L)), f"pyramid level {L}"
        assert np.array_equal(ex.blurred_level(L), orc.level_image(L, blurred=True)), f"blur level {L}"
        raw_o = orc.level_raw(L).astype(np.int64)
        raw_g = ex.fast_keypoints(L).astype(np.int64)
        assert raw_o.shape == raw_g.shape, f"FAST count level {L}: {raw_o.shape} vs {raw_g.shape}"
        assert np.array_equal(_sorted_rows(raw_o), _sorted_rows(raw_g)), f"FAST keypoint set level {L}"
        lo, lg = orc.level_kps(L), ex.level_keypoints(L)
        assert len(lo) == len(lg), f"octree count level {L}"
        for f in ("x", "y", "response", "octave", "size"):
            assert np.array_equal(lo[f], lg[f]), f"octree {f} level {L} (set and ORDER)"
        assert np.max(np.abs(lo["angle"] - lg["angle"]), initial=0) <= 1e-3, f"angle level {L}"


@pytest.mark.parametrize("w,h,nf,seed", SHAPES)
def test_final_parity(gpu, w, h, nf, seed):
    img, orc, ok, od, ex, gk, gd = run_pair(w, h, nf, seed)
    assert len(ok) == len(gk)
    for f in ("x", "y", "size", "response", "octave", "class_id"):
        assert np.array_equal(ok[f], gk[f]), f
    # north_star allows 1e-3 deg on angles and 0.1 % differing descriptor rows (float orientation); on these fixed frames the
    # kernels reproduce the oracle's float arithmetic exactly (ordered _rn operations, no FMA contraction), so assert equality
    assert np.max(np.abs(ok["angle"] - gk["angle"])) <= 1e-3
    assert np.array_equal(ok["angle"], gk["angle"]), "angles are bit-equal on the fixed parity frames"
    bad = np.any(od != gd, axis=1).sum()
    assert bad == 0, f"{bad} of {len(od)} descriptor rows differ"


def test_edge_cases(gpu):
    import orbb200, pyoracle
    from synth import synth
    ex = orbb200.ORBextractor(1000, max_size=(640, 480))
    # flat image: zero keypoints, descriptors released (ORBextractor.cpp:1076-1077)
    k, d = ex(np.full((480, 640), 77, np.uint8))
    assert len(k) == 0 and d.shape == (0, 32)
    # empty image: silent return (ORBextractor.cpp:1056-1057)
    k, d = ex(np.zeros((0, 0), np.uint8))
    assert len(k) == 0
    # non-contiguous rows (stride > width)
    big = synth(700, 480, 31)
    view = big[:, 30:670]
    k1, d1 = ex(view)
    k2, d2 = ex(np.ascontiguousarray(view))
    assert k1.tobytes() == k2.tobytes() and np.array_equal(d1, d2)
    ok, od = pyoracle.OracleExtractor(1000).extract(np.ascontiguousarray(view))
    assert k1.tobytes() == ok.tobytes() and np.array_equal(d1, od)
    # a smaller image through the same handle (geometry switch), then back
    small = synth(333, 257, 32)
    ks, ds = ex(small)
    oks, ods = pyoracle.OracleExtractor(1000).extract(small)
    assert ks.tobytes() == oks.tobytes() and np.array_equal(ds, ods)
    k3, d3 = ex(view)
    assert k3.tobytes() == k1.tobytes()
    # larger than the handle's maximum
    with pytest.raises(orbb200.OrbError) as e:
        ex(synth(800, 480, 1))
    assert e.value.code == orbb200.E_BAD_SHAPE
    # a shape whose upper levels round to nothing is rejected half-way through the geometry set-up: the handle must
    # rebuild its tables for the previous, valid shape instead of running on the rejected shape's
    with pytest.raises(orbb200.OrbError) as e:
        ex(np.full((1, 640), 9, np.uint8))
    assert e.value.code == orbb200.E_BAD_SHAPE
    k4, d4 = ex(view)
    assert k4.tobytes() == k1.tobytes() and np.array_equal(d4, d1)
    # capacity too small: count reported, error returned, first `capacity` rows valid
    import ctypes as C
    cap = 100
    kps = np.zeros(cap, orbb200.KP_DTYPE); desc = np.zeros((cap, 32), np.uint8); n = C.c_int()
    img = np.ascontiguousarray(view)
    rc = ex.L.orbx_extract(ex.h, img.ctypes.data_as(C.c_void_p), 640, 480, 640, kps.ctypes.data_as(C.c_void_p),
                           desc.ctypes.data_as(C.c_void_p), cap, C.byref(n))
    assert rc == orbb200.E_CAPACITY and n.value == len(k1) and kps.tobytes() == k1[:cap].tobytes()
    # invalid configuration
    with pytest.raises(orbb200.OrbError):
        orbb200.ORBextractor(1000, iniThFAST=5, minThFAST=7)
    # constructor tables
    t = ex._tables(); o = pyoracle.OracleExtractor(1000).tables()
    for key in t:
        assert np.array_equal(t[key], o[key]), key


def test_noise_and_extreme_textures(gpu):
    """Uniform noise (10x the usual corner density, stresses the candidate buffers) and a hard checkerboard (score ties)."""
    import orbb200, pyoracle
    rng = np.random.default_rng(0)
    ex = orbb200.ORBextractor(1500, max_size=(640, 480))
    noise = rng.integers(0, 256, (480, 640), dtype=np.uint8)
    yy, xx = np.mgrid[0:480, 0:640]
    checker = ((((xx // 9) + (yy // 7)) & 1) * 200 + 20).astype(np.uint8)
    for img in (noise, checker):
        k, d = ex(img)
        ok, od = pyoracle.OracleExtractor(1500).extract(img)
        assert len(k) == len(ok)
        for f in ("x", "y", "size", "response", "octave"):
            assert np.array_equal(k[f], ok[f]), f
        assert np.max(np.abs(k["angle"] - ok["angle"]), initial=0) <= 1e-3
        assert np.any(d != od, axis=1).sum() <= max(1, 0.001 * len(d))


def test_batch_equals_single(gpu):
    """orbx_extract_batch (one launch sequence for F frames) == F single-frame calls; device-resident API too."""
    import torch, orbb200
    from synth import synth
    F, w, h = 6, 752, 480
    imgs = np.stack([synth(w, h, 100 + i) for i in range(F)])
    ex1 = orbb200.ORBextractor(1200, max_size=(w, h))
    exb = orbb200.ORBextractor(1200, max_size=(w, h), max_batch=F)
    single = [ex1(imgs[i]) for i in range(F)]
    batch = exb.extract_batch(imgs)
    for (k1, d1), (k2, d2) in zip(single, batch):
        assert k1.tobytes() == k2.tobytes() and np.array_equal(d1, d2)
    cap = 1500
    dimg = torch.from_numpy(imgs).cuda()
    kps = torch.zeros((F, cap, 7), dtype=torch.float32, device="cuda")
    desc = torch.zeros((F, cap, 32), dtype=torch.uint8, device="cuda")
    cnt = torch.zeros(F, dtype=torch.int32, device="cuda")
    exb.extract_batch_device(dimg.data_ptr(), F, w, h, w, w * h, kps.data_ptr(), desc.data_ptr(), cnt.data_ptr(), cap)
    exb.sync()
    cnt = cnt.cpu().numpy()
    for i in range(F):
        k = kps[i, :cnt[i]].cpu().numpy().view(orbb200.KP_DTYPE).reshape(-1)
        assert k.tobytes() == single[i][0].tobytes()
        assert np.array_equal(desc[i, :cnt[i]].cpu().numpy(), single[i][1])


def test_4k_frame(gpu):
    """C4 shape: 3840x2160, 8000 keypoints, against the oracle."""
    import orbb200, pyoracle
    from synth import synth
    img = synth(3840, 2160, 5000)
    ex = orbb200.ORBextractor(8000, max_size=(3840, 2160))
    k, d = ex(img)
    ok, od = pyoracle.OracleExtractor(8000).extract(img)
    assert len(k) == len(ok) and len(k) >= 8000
    for f in ("x", "y", "size", "response", "octave", "class_id"):
        assert np.array_equal(k[f], ok[f]), f
    assert np.max(np.abs(k["angle"] - ok["angle"])) <= 1e-3
    assert np.any(d != od, axis=1).sum() <= 0.001 * len(d)


def test_legacy_blur_and_large_scale_factor(gpu):
    """blur_legacy (OpenCV <= 3.3 kernel, needs the clamp path) and scaleFactor 1.5 (resize falls back to the plain kernel)."""
    import orbb200, pyoracle
    from synth import synth
    img = synth(640, 480, 77)
    ex = orbb200.ORBextractor(1000, max_size=(640, 480), blur_legacy=True)
    k, d = ex(img)
    orc = pyoracle.OracleExtractor(1000, legacy_blur=True)
    ok, od = orc.extract(img)
    for L in range(8):
        assert np.array_equal(ex.blurred_level(L), orc.level_image(L, blurred=True)), f"legacy blur level {L}"
    assert k.tobytes() == ok.tobytes() or (np.array_equal(k["x"], ok["x"]) and np.max(np.abs(k["angle"] - ok["angle"])) <= 1e-3)
    assert np.any(d != od, axis=1).sum() <= 0.001 * len(d)
    for sf, nl in ((1.5, 5), (1.1, 8), (1.25, 6)):
        ex2 = orbb200.ORBextractor(800, scaleFactor=sf, nlevels=nl, max_size=(640, 480))
        k2, d2 = ex2(img)
        orc2 = pyoracle.OracleExtractor(800, sf, nl)
        ok2, od2 = orc2.extract(img)
        for L in range(nl):
            assert np.array_equal(ex2.pyramid_level(L), orc2.level_image(L)), f"scale {sf} pyramid level {L}"
        assert len(k2) == len(ok2)
        for f in ("x", "y", "size", "response", "octave"):
            assert np.array_equal(k2[f], ok2[f]), (sf, f)
        assert np.any(d2 != od2, axis=1).sum() <= 0.001 * len(d2)


@pytest.mark.parametrize("w,h", [(641, 481), (123, 97), (1000, 37), (36, 300), (119, 64), (121, 65), (240, 33)])
def test_odd_shapes(gpu, w, h):
    """Widths that are not multiples of 4, tiles cut at every offset, images smaller than one FAST cell row."""
    import orbb200, pyoracle
    from synth import synth
    img = synth(w, h, w * 7 + h)
    ex = orbb200.ORBextractor(500, max_size=(w, h))
    k, d = ex(img)
    orc = pyoracle.OracleExtractor(500)
    ok, od = orc.extract(img)
    for L in range(8):
        assert np.array_equal(ex.pyramid_level(L), orc.level_image(L)), f"pyramid level {L}"
        assert np.array_equal(ex.blurred_level(L), orc.level_image(L, blurred=True)), f"blur level {L}"
    assert len(k) == len(ok)
    for f in ("x", "y", "size", "response", "octave"):
        assert np.array_equal(k[f], ok[f]), f
    assert np.any(d != od, axis=1).sum() <= max(1, 0.001 * len(d))


def _colour_frames(F, w, h, ch, seed):
    """Colour frames whose channels are differently textured (so the channel order matters)."""
    from synth import synth
    out = np.zeros((F, h, w, ch), np.uint8)
    for f in range(F):
        for c in range(ch):
            out[f, :, :, c] = synth(w, h, seed + 10 * f + c)
    return out


@pytest.mark.parametrize("w,h,ch", [(640, 480, 3), (640, 480, 4), (203, 177, 3), (201, 99, 4)])
def test_colour_frames(gpu, w, h, ch):
    """cv::cvtColor (Tracking.cpp:202-227) fused in front of the extractor.  The device conversion is bit-exact
    against the oracle restatement (pinned to cv2 4.13 by tests/golden/cv2_cvtcolor.npz) and the colour entry point returns
    exactly what the grey entry point returns on the converted frames, for both channel orders."""
    import torch, orbb200, pyoracle
    F = 3
    col = _colour_frames(F, w, h, ch, 300)
    ex = orbb200.ORBextractor(600, max_size=(w, h), max_batch=F)
    for rgb in (True, False):
        grey = np.stack([pyoracle.cvt_gray(col[f], rgb_order=rgb) for f in range(F)])
        # the stand-alone device conversion
        d_src = torch.from_numpy(col).cuda()
        d_dst = torch.zeros((F, h, w), dtype=torch.uint8, device="cuda")
        rc = orbb200.lib().orbx_cvt_gray_device(d_src.data_ptr(), F, w, h, w * ch, w * h * ch, ch, 1 if rgb else 0,
                                                d_dst.data_ptr(), w, w * h, None)
        assert rc == 0
        torch.cuda.synchronize()
        assert np.array_equal(d_dst.cpu().numpy(), grey)
        # colour batch == grey batch of the converted frames
        want = ex.extract_batch(grey)
        got = ex.extract_batch_color(col, rgb_order=rgb)
        for (k1, d1), (k2, d2) in zip(want, got):
            assert len(k1) > 50 and k1.tobytes() == k2.tobytes() and np.array_equal(d1, d2)
    # the two channel orders really differ
    assert not np.array_equal(pyoracle.cvt_gray(col[0], True), pyoracle.cvt_gray(col[0], False))


def test_colour_frames_strided_rows(gpu):
    """Row-padded colour frames (stride > w*channels) take the per-frame copy path."""
    import orbb200, pyoracle
    w, h, ch, F = 320, 240, 3, 2
    col = _colour_frames(F, w, h, ch, 400)
    padded = np.zeros((F, h, w * ch + 24), np.uint8)
    padded[:, :, :w * ch] = col.reshape(F, h, w * ch)
    ex = orbb200.ORBextractor(500, max_size=(w, h), max_batch=F)
    kps = np.zeros((F, 800), orbb200.KP_DTYPE); desc = np.zeros((F, 800, 32), np.uint8); cnt = np.zeros(F, np.int32)
    rc = orbb200.lib().orbx_extract_batch_color(ex.h, padded.ctypes.data, F, w, h, padded.shape[2], padded.shape[1] * padded.shape[2], ch, 1,
                                                kps.ctypes.data, desc.ctypes.data, cnt.ctypes.data, 800)
    assert rc == 0
    want = ex.extract_batch(np.stack([pyoracle.cvt_gray(col[f], True) for f in range(F)]))
    for f in range(F):
        assert kps[f, :cnt[f]].tobytes() == want[f][0].tobytes() and np.array_equal(desc[f, :cnt[f]], want[f][1])
    # invalid channel counts are rejected
    assert orbb200.lib().orbx_extract_batch_color(ex.h, padded.ctypes.data, F, w, h, padded.shape[2], 0, 2, 1,
                                                  kps.ctypes.data, desc.ctypes.data, cnt.ctypes.data, 800) != 0


def _golden():
    import importlib.util, os
    gold = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    spec = importlib.util.spec_from_file_location("make_golden", os.path.join(gold, "make_golden.py"))
    gen = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(gen)
    return gen, np.load(os.path.join(gold, "ref_extract.npz")), np.load(os.path.join(gold, "cv2_prims.npz"))


@pytest.mark.parametrize("w,h,nf,seed", [(320, 240, 500, 21), (203, 177, 300, 22)])
def test_gpu_equals_reference_golden_full_output(gpu, w, h, nf, seed):
    """The CUDA path against the committed outputs of the reference's own ORBextractor.cpp (tests/golden/ref_extract.npz,
    compiled unmodified as oracle/_ref when the fixture was made): keypoints incl. order and angles, descriptors, pyramid."""
    import hashlib, orbb200
    from synth import synth
    gen, refg, _ = _golden()
    tag = f"{w}x{h}_{nf}_{seed}"
    ex = orbb200.ORBextractor(nf, max_size=(w, h))
    k, d = ex(synth(w, h, seed), None)
    rk, rd = refg[f"kps_{tag}"], refg[f"desc_{tag}"]
    assert len(k) == len(rk) and len(k) > 100
    assert gen.kp_exact_bytes(k) == gen.kp_exact_bytes(rk)
    assert k["angle"].tobytes() == rk["angle"].tobytes()
    assert np.array_equal(d, rd)
    for L in range(8):
        assert hashlib.sha256(ex.pyramid_level(L).tobytes()).hexdigest() == str(refg[f"pyrsha_{tag}"][L]), f"pyramid level {L}"


@pytest.mark.parametrize("w,h,nf,seed", [(640, 480, 1000, 1), (1241, 376, 2000, 2), (752, 480, 1200, 3)])
def test_gpu_equals_reference_golden_config_shapes(gpu, w, h, nf, seed):
    """C1 / C2 / C3 shapes of BASELINE.json: the CUDA output hashes to the compiled reference's committed checksums."""
    import hashlib, orbb200
    from synth import synth
    gen, refg, _ = _golden()
    tag = f"{w}x{h}_{nf}_{seed}"
    k, d = orbb200.ORBextractor(nf, max_size=(w, h))(synth(w, h, seed), None)
    assert len(k) == int(refg[f"count_{tag}"])
    assert hashlib.sha256(gen.kp_exact_bytes(k)).hexdigest() == str(refg[f"kpsha_{tag}"])
    assert k["angle"].tobytes() == refg[f"angles_{tag}"].tobytes()
    assert hashlib.sha256(d.tobytes()).hexdigest() == str(refg[f"descsha_{tag}"])


def test_gpu_pyramid_and_blur_equal_cv2_golden(gpu):
    """Pyramid levels and blurred levels of the CUDA path against cv2 4.13's own outputs (tests/golden/cv2_prims.npz)."""
    import orbb200
    from synth import synth
    gen, _, cvg = _golden()
    for (w, h, seed) in gen.PRIM_SHAPES:
        tag = f"{w}x{h}_{seed}"
        ex = orbb200.ORBextractor(200, max_size=(w, h))
        ex(synth(w, h, seed), None)
        for L in range(8):
            if L > 0:
                assert np.array_equal(ex.pyramid_level(L), cvg[f"pyr_{tag}_L{L}"]), f"resize level {L} of {tag}"
            assert np.array_equal(ex.blurred_level(L), cvg[f"blur_{tag}_L{L}"]), f"blur level {L} of {tag}"


def test_multi_chunk_pipeline(gpu):
    """orbx_extract_batch / _color with enough frames for the ramped multi-chunk, multi-stream pipeline (head 16+32, full 64s,
    tail 32+16 for 300 small frames): every frame equals its single-frame result, pinned and pageable hosts alike."""
    import torch, orbb200, pyoracle
    from synth import synth
    w, h, F, ND = 160, 128, 300, 7
    distinct = np.stack([synth(w, h, 500 + i) for i in range(ND)])
    ex1 = orbb200.ORBextractor(150, max_size=(w, h))
    single = [ex1(distinct[i], None) for i in range(ND)]
    imgs = distinct[np.arange(F) % ND]
    exb = orbb200.ORBextractor(150, max_size=(w, h), max_batch=F)
    for host in (np.ascontiguousarray(imgs), torch.from_numpy(np.ascontiguousarray(imgs)).pin_memory().numpy()):
        out = exb.extract_batch(host)
        assert len(out) == F
        for f in range(F):
            k, d = out[f]
            assert k.tobytes() == single[f % ND][0].tobytes() and np.array_equal(d, single[f % ND][1]), f"frame {f}"
    col = np.stack([imgs, imgs, imgs], axis=3)            # R = G = B = v  ->  grey == v exactly ((9798+19235+3735) v + 2^14) >> 15 = v
    assert np.array_equal(pyoracle.cvt_gray(col[0], True), imgs[0])
    out = exb.extract_batch_color(np.ascontiguousarray(col), rgb_order=False)
    for f in range(F):
        assert out[f][0].tobytes() == single[f % ND][0].tobytes() and np.array_equal(out[f][1], single[f % ND][1]), f"colour frame {f}"


def test_rectified_frames(gpu):
    """SURVEY 8(f)-4, second half: cv::remap (stereo rectification, stereo_euroc.cc:174-185) on the device.  The device remap is
    bit-equal to cv2's committed output and to the oracle restatement; the rectified entry point returns exactly what the grey
    entry point returns on the oracle-rectified frames (raw frames larger than the rectified ones, several frames)."""
    import torch, orbb200, pyoracle
    from synth import synth
    gen, _, _ = _golden()
    import os
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "cv2_remap.npz"))
    img, hx, hy = gen.remap_inputs()
    h, w = img.shape
    for mx, my, want in ((g["rect_mapx"], g["rect_mapy"], g["rect_out"]), (hx, hy, g["hostile_out"])):
        d_src = torch.from_numpy(img).cuda(); d_dst = torch.zeros((h, w), dtype=torch.uint8, device="cuda")
        rc = orbb200.lib().orbx_remap_device(d_src.data_ptr(), 1, w, h, w, w * h, mx.ctypes.data, my.ctypes.data, w, h, d_dst.data_ptr(), w, w * h, None)
        assert rc == 0
        assert np.array_equal(d_dst.cpu().numpy(), want)
    # a camera-sized case: raw 800x520 frames rectified into 752x480
    sw, sh, dw, dh, F = 800, 520, 752, 480, 3
    raw = np.stack([synth(sw, sh, 600 + i) for i in range(F)])
    yy, xx = np.mgrid[0:dh, 0:dw].astype(np.float32)
    cx, cy = np.float32(dw / 2), np.float32(dh / 2)
    r2 = ((xx - cx) ** 2 + (yy - cy) ** 2) / np.float32(dw * dw)
    mapx = (xx + 24 + (xx - cx) * (np.float32(0.08) * r2)).astype(np.float32)
    mapy = (yy + 20 + (yy - cy) * (np.float32(0.08) * r2)).astype(np.float32)
    rect = np.stack([pyoracle.remap(raw[f], mapx, mapy) for f in range(F)])
    ex = orbb200.ORBextractor(1200, max_size=(dw, dh), max_batch=F)
    want = ex.extract_batch(rect)
    ex.set_rectification(mapx, mapy, (sw, sh))
    got = ex.extract_batch_rectified(raw)
    for (k1, d1), (k2, d2) in zip(want, got):
        assert len(k1) > 500 and k1.tobytes() == k2.tobytes() and np.array_equal(d1, d2)
    assert np.array_equal(ex.pyramid_level(0, frame=1), rect[1])       # mImGray the caller keeps = the rectified frame
    ex.set_rectification(None, None, None)
    kps = np.zeros((F, 10), orbb200.KP_DTYPE); desc = np.zeros((F, 10, 32), np.uint8); cnt = np.zeros(F, np.int32)
    assert orbb200.lib().orbx_extract_batch_rectified(ex.h, raw.ctypes.data, F, sw, sw * sh, kps.ctypes.data, desc.ctypes.data, cnt.ctypes.data, 10) == orbb200.E_STATE


def test_cuda_graph_replay_equals_plain_launches(gpu):
    """orbx_graph_capture / orbx_graph_launch: the captured launch sequence gives the same keypoints and descriptors as
    orbx_extract_batch_device on the same buffers, on replays with new image content, and reports its state errors."""
    import torch, orbb200
    from synth import synth
    W, H, CAP, B = 640, 480, 1280, 2
    ex = orbb200.ORBextractor(1000, max_size=(W, H), max_batch=B)
    assert ex.L.orbx_graph_launch(ex.h) == orbb200.E_STATE                  # nothing captured yet
    imgs = torch.from_numpy(np.stack([synth(W, H, 71), synth(W, H, 72)])).cuda()
    k = torch.zeros((B, CAP, 7), device="cuda"); d = torch.zeros((B, CAP, 32), dtype=torch.uint8, device="cuda"); c = torch.zeros(B, dtype=torch.int32, device="cuda")
    ex.graph_capture(imgs.data_ptr(), B, W, H, W, W * H, k.data_ptr(), d.data_ptr(), c.data_ptr(), CAP)
    for seeds in ((71, 72), (73, 74)):
        imgs.copy_(torch.from_numpy(np.stack([synth(W, H, s) for s in seeds])).cuda())
        torch.cuda.synchronize()
        ex.extract_batch_device(imgs.data_ptr(), B, W, H, W, W * H, k.data_ptr(), d.data_ptr(), c.data_ptr(), CAP)
        ex.sync()
        rk, rd, rc = k.clone(), d.clone(), c.clone()
        k.zero_(); d.zero_(); c.zero_()
        torch.cuda.synchronize()
        ex.graph_launch(); ex.sync()
        assert torch.equal(c, rc) and int(c.min()) >= 1000
        for f in range(B):
            n = int(c[f])
            assert torch.equal(k[f, :n].view(torch.int32), rk[f, :n].view(torch.int32)) and torch.equal(d[f, :n], rd[f, :n])      # bit patterns: class_id -1 reads as NaN
    ex.set_profiling(True)
    assert ex.L.orbx_graph_capture(ex.h, orbb200._p(imgs.data_ptr()), B, W, H, W, W * H, orbb200._p(k.data_ptr()), orbb200._p(d.data_ptr()),
                                   orbb200._p(c.data_ptr()), CAP) == orbb200.E_STATE


@pytest.mark.parametrize("env", [{"ORBB_L3_OW": "32"}, {"ORBB_L3_OW": "30"}, {"ORBB_L3_OW": "32", "ORBB_L3_SPLIT": "1"}, {"ORBB_L3_OW": "30", "ORBB_L3_SPLIT": "0"},
                                 {"ORBB_L3_OW": "32", "ORBB_NO_TMA": "1"}, {"ORBB_OCT_FAN": "0", "ORBB_L3_PDL": "0"}],
                         ids=lambda e: ",".join(f"{k[5:]}={v}" for k, v in e.items()))
def test_every_level_kernel_variant(gpu, env):
    """The level kernel has a tile-width (120 / 128 pixels), a latency (SPLIT + dependent launch + per-level octree streams) and a staging
    (TMA / ordinary loads) variant; the library picks them per level and per launch.  The knobs are read once per process, so each forced
    combination runs the stage-by-stage and the final parity tests (single frames, batches, odd shapes, noise) in a process of its own."""
    import os, subprocess, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.join(root, "tests", "test_gpu_extract.py"), "-m", "gpu", "-x", "-q", "-k",
                        "stage_parity or final_parity or noise or batch_equals_single or odd_shapes or multi_chunk or graph_replay"],
                       env={**os.environ, **env}, capture_output=True, text=True, timeout=900, cwd=root)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]


def test_random_shapes_and_parameters(gpu):
    """Thirty seeded random configurations (image size 90 .. 1000 x 70 .. 700, 3 .. 10 levels, scale factor 1.1 .. 1.25, FAST thresholds,
    feature budget), each as a single frame and as a batch of three through one handle: every pyramid / blur level, every keypoint field
    and (to the angle tolerance) every descriptor against the oracle.  The sizes fall on both sides of every tile-geometry decision the
    library takes per level (tile width, latency mode, TMA / ordinary staging through odd pitches)."""
    import orbb200, pyoracle
    from synth import synth
    rng = np.random.default_rng(20261105)
    for case in range(30):
        w = int(rng.integers(90, 1001)); h = int(rng.integers(70, 701))
        nl = int(rng.integers(3, 11)); sf = float(np.float32(rng.uniform(1.1, 1.25)))
        ini = int(rng.integers(8, 40)); mn = int(rng.integers(0, ini + 1)); nf = int(rng.integers(50, 1500))
        if min(w, h) / sf ** (nl - 1) < 40:                      # the reference needs every level larger than its 2 x 19 px border
            nl = max(1, int(np.log(min(w, h) / 40.0) / np.log(sf)))
        tag = f"case {case}: {w}x{h} nlevels={nl} scale={sf:.4f} ini={ini} min={mn} nf={nf}"
        imgs = [synth(w, h, 1000 + 3 * case + i) for i in range(3)]
        orc = pyoracle.OracleExtractor(nf, sf, nl, ini, mn)
        ex = orbb200.ORBextractor(nf, sf, nl, ini, mn, max_size=(w, h), max_batch=3)
        ok, od = orc.extract(imgs[0])
        k, d = ex(imgs[0])
        for L in range(nl):
            assert np.array_equal(ex.pyramid_level(L), orc.level_image(L)), f"{tag}: pyramid level {L}"
            assert np.array_equal(ex.blurred_level(L), orc.level_image(L, blurred=True)), f"{tag}: blur level {L}"
        assert len(k) == len(ok), tag
        for f in ("x", "y", "size", "response", "octave"):
            assert np.array_equal(k[f], ok[f]), f"{tag}: {f}"
        if len(k):
            assert np.max(np.abs(k["angle"] - ok["angle"])) <= 1e-3, tag
            assert np.any(d != od, axis=1).sum() <= max(1, 0.001 * len(d)), tag
        res = ex.extract_batch(np.stack(imgs))
        for i in range(3):
            oki, odi = orc.extract(imgs[i])
            kb, db = res[i]
            assert len(kb) == len(oki), f"{tag}: frame {i}"
            for f in ("x", "y", "response", "octave"):
                assert np.array_equal(kb[f], oki[f]), f"{tag}: frame {i} {f}"
        ex.close()
