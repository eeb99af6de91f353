"""Randomized parity sweep (not part of the test suite: minutes of oracle time): extractor configurations as in
tests/test_gpu_extract.py::test_random_shapes_and_parameters with another seed, plus for every third case the fused stereo constructor
and a tracked-frame search against the oracle.  Prints one line per failing case and a summary.
usage: fuzz_sweep.py ncases seed [max_width max_height max_features]"""
import sys, os, traceback
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "orb-slam2-annotation_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
import numpy as np, orbb200, pyoracle
from synth import synth
ncases = int(sys.argv[1]) if len(sys.argv) > 1 else 60
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 7)
MAXW = int(sys.argv[3]) if len(sys.argv) > 3 else 1300; MAXH = int(sys.argv[4]) if len(sys.argv) > 4 else 800
NFMAX = int(sys.argv[5]) if len(sys.argv) > 5 else 2500
bad = 0
for case in range(ncases):
    w = int(rng.integers(90, MAXW + 1)); h = int(rng.integers(70, MAXH + 1))
    nl = int(rng.integers(1, 11)); sf = float(np.float32(rng.uniform(1.05, 1.4)))
    ini = int(rng.integers(1, 60)); mn = int(rng.integers(0, ini + 1)); nf = int(rng.integers(20, NFMAX))
    while nl > 1 and min(w, h) / sf ** (nl - 1) < 40:
        nl -= 1
    seed = int(rng.integers(0, 1 << 30))
    tag = f"{w} {h} {nl} {sf:.6f} {ini} {mn} {nf} {seed}"
    try:
        img = synth(w, h, seed)
        legacy = bool(rng.random() < 0.25)                      # the Q8 blur kernel of OpenCV <= 3.3 (sums to 257: clamped)
        orc = pyoracle.OracleExtractor(nf, sf, nl, ini, mn, 0, legacy)
        ex = orbb200.ORBextractor(nf, sf, nl, ini, mn, max_size=(w, h), max_batch=2, blur_legacy=legacy)
        tag += f" legacy={int(legacy)}"
        ok, od = orc.extract(img)
        k, d = ex(img)
        for L in range(nl):
            assert np.array_equal(ex.pyramid_level(L), orc.level_image(L)), f"pyramid level {L}"
            assert np.array_equal(ex.blurred_level(L), orc.level_image(L, blurred=True)), f"blur level {L}"
        assert len(k) == len(ok), f"count {len(k)} vs {len(ok)}"
        for f in ("x", "y", "size", "response", "octave"):
            assert np.array_equal(k[f], ok[f]), f
        if len(k):
            assert np.max(np.abs(k["angle"] - ok["angle"])) <= 1e-3, "angle"
            assert np.any(d != od, axis=1).sum() <= max(1, 0.002 * len(d)), "descriptors"
        if case % 3 == 0 and len(k) > 20:
            right = np.roll(img, -7, axis=1)
            oL, oR = pyoracle.OracleExtractor(nf, sf, nl, ini, mn, 0, legacy), pyoracle.OracleExtractor(nf, sf, nl, ini, mn, 0, legacy)
            oL.extract(img); okR, _ = oR.extract(right)
            F, kL, dL, kR, dR, ur, dp = orbb200.stereo_frame(ex, img, right, 40.0, (0, 0, w, h))
            with np.errstate(divide="ignore"):
                our, odp, _, _ = pyoracle.stereo_match(oL, oR, 40.0, 0.0)
            assert kR.tobytes() == okR.tobytes(), "right keypoints"
            exact = np.all(kL["angle"] == ok["angle"]) and np.array_equal(dL, od)
            if exact:
                assert np.array_equal(ur, our) and np.array_equal(dp, odp), "stereo"
            F.close()
        if case % 4 == 1:
            # device-resident entry point on two frames with an arbitrary base offset, row stride and frame stride (aligned: bulk tensor copies;
            # otherwise the ordinary-load staging of level 0), results in caller buffers of arbitrary capacity
            import torch
            offb = int(rng.choice([0, 16, 1, 4, 7])); stride = w + int(rng.choice([0, 0, 16, 3, 37])); fstride = stride * h + int(rng.choice([0, 16, 5]))
            buf = torch.zeros(offb + 2 * fstride + 64, dtype=torch.uint8, device="cuda")
            img2 = synth(w, h, seed + 1)
            for f_, im in enumerate((img, img2)):
                view = buf[offb + f_ * fstride: offb + f_ * fstride + stride * h].view(h, stride)
                view[:, :w] = torch.from_numpy(im).cuda()
            capd = len(ok) + int(rng.integers(0, 300))
            dk = torch.zeros((2, capd, 7), dtype=torch.float32, device="cuda"); dd = torch.zeros((2, capd, 32), dtype=torch.uint8, device="cuda"); dc = torch.zeros(2, dtype=torch.int32, device="cuda")
            ex.extract_batch_device(buf.data_ptr() + offb, 2, w, h, stride, fstride, dk.data_ptr(), dd.data_ptr(), dc.data_ptr(), capd)
            ex.sync()
            ok2, od2 = orc.extract(img2)
            cnt = dc.cpu().numpy()
            if len(ok2) <= capd:
                assert cnt[0] == len(ok) and cnt[1] == len(ok2), f"device counts {cnt} vs {len(ok)}, {len(ok2)}"
                for f_, (okf, odf) in enumerate(((ok, od), (ok2, od2))):
                    kk = dk[f_, :cnt[f_]].cpu().numpy().view(orbb200.KP_DTYPE).reshape(-1)
                    for fld in ("x", "y", "response", "octave"):
                        assert np.array_equal(kk[fld], okf[fld]), f"device frame {f_} {fld}"
                    assert np.any(dd[f_, :cnt[f_]].cpu().numpy() != odf, axis=1).sum() <= max(1, 0.002 * len(odf)), f"device frame {f_} descriptors"
        ex.close()
    except Exception as e:
        bad += 1
        print("FAIL", tag, "|", type(e).__name__, str(e)[:160], flush=True)
print("cases", ncases, "failures", bad)
