"""How far from bit-exact are the float outputs? angle (fastAtan2) and descriptor rows vs the CPU oracle on several frames."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "orb-slam2-annotation_b200"), os.path.join(ROOT, "oracle")]
import numpy as np, orbb200, pyoracle
from synth import synth
tot = bad = 0; amax = 0.0; aneq = 0
for (w, h, nf, seed) in [(640, 480, 1000, 1), (752, 480, 1200, 3), (1241, 376, 2000, 2), (640, 480, 1000, 7), (640, 480, 1000, 8), (320, 240, 500, 21)]:
    img = synth(w, h, seed)
    ex = orbb200.ORBextractor(nf, max_size=(w, h))
    k, d = ex(img, None)
    ok, od = pyoracle.OracleExtractor(nf).extract(img)
    assert len(k) == len(ok)
    tot += len(k); bad += int(np.any(d != od, axis=1).sum()); aneq += int((k["angle"] != ok["angle"]).sum())
    amax = max(amax, float(np.max(np.abs(k["angle"] - ok["angle"]))))
print(f"keypoints {tot}: angles differing {aneq} (max |diff| {amax:g} deg), descriptor rows differing {bad}")
