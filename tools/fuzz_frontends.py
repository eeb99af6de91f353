"""Randomized parity sweep of the two front ends in front of the extractor: cv::cvtColor (3 / 4 channels, both channel orders, odd sizes) and
cv::remap rectification (random smooth maps incl. samples outside the raw frame, raw size != rectified size) against the oracle's restatements,
and the colour / rectified entry points against the grey entry point on the oracle-converted frames.  usage: fuzz_frontends.py ncases seed"""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "orb-slam2-annotation_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
import numpy as np, torch, orbb200, pyoracle
from synth import synth
ncases = int(sys.argv[1]) if len(sys.argv) > 1 else 40
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 13)
bad = 0
for case in range(ncases):
    w = int(rng.integers(80, 900)); h = int(rng.integers(70, 600)); F = int(rng.integers(1, 4)); nf = int(rng.integers(100, 1200))
    tag = f"case {case}: {w}x{h} F={F} nf={nf}"
    try:
        ex = orbb200.ORBextractor(nf, max_size=(w, h), max_batch=F)
        if case % 2 == 0:
            ch = int(rng.choice([3, 4])); rgb = bool(rng.random() < 0.5); tag += f" colour ch={ch} rgb={rgb}"
            col = np.zeros((F, h, w, ch), np.uint8)
            for f in range(F):
                for c in range(ch): col[f, :, :, c] = synth(w, h, 7 * case + 3 * f + c)
            grey = np.stack([pyoracle.cvt_gray(col[f], rgb_order=rgb) for f in range(F)])
            want = ex.extract_batch(grey); got = ex.extract_batch_color(col, rgb_order=rgb)
            for f, ((k1, d1), (k2, d2)) in enumerate(zip(want, got)):
                assert k1.tobytes() == k2.tobytes() and np.array_equal(d1, d2), f"colour frame {f}"
                assert np.array_equal(ex.pyramid_level(0, frame=f), grey[f]), f"grey image of frame {f}"
        else:
            sw = w + int(rng.integers(-20, 80)); sh = h + int(rng.integers(-20, 80)); tag += f" rectify from {sw}x{sh}"
            raw = np.stack([synth(sw, sh, 11 * case + f) for f in range(F)])
            yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
            cx, cy = np.float32(w / 2), np.float32(h / 2)
            r2 = ((xx - cx) ** 2 + (yy - cy) ** 2) / np.float32(w * w)
            kk = np.float32(rng.uniform(-0.15, 0.15)); ox = np.float32(rng.uniform(-15, 40)); oy = np.float32(rng.uniform(-15, 40))
            mapx = (xx + ox + (xx - cx) * (kk * r2)).astype(np.float32); mapy = (yy + oy + (yy - cy) * (kk * r2)).astype(np.float32)
            rect = np.stack([pyoracle.remap(raw[f], mapx, mapy) for f in range(F)])
            want = ex.extract_batch(rect)
            ex.set_rectification(mapx, mapy, (sw, sh))
            got = ex.extract_batch_rectified(raw)
            for f, ((k1, d1), (k2, d2)) in enumerate(zip(want, got)):
                assert k1.tobytes() == k2.tobytes() and np.array_equal(d1, d2), f"rectified frame {f}"
                assert np.array_equal(ex.pyramid_level(0, frame=f), rect[f]), f"rectified image of frame {f}"
        ex.close()
    except Exception as e:
        bad += 1
        print("FAIL", tag, "|", type(e).__name__, str(e)[:200], flush=True)
print("cases", ncases, "failures", bad)
