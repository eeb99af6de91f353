import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "orb-slam2-annotation_b200"), os.path.join(ROOT, "oracle")]
import numpy as np, orbb200
from synth import synth
img = synth(640, 480, 1)
ex = orbb200.ORBextractor(1000, max_size=(640, 480))
gk, gd = ex(img)
out = {"kps": gk, "desc": gd}
for L in range(8):
    out[f"fast{L}"] = ex.fast_keypoints(L)
    out[f"lvl{L}"] = ex.level_keypoints(L)
np.savez(os.path.join(ROOT, "gpurun_out", "dump.npz"), **out)
print("dumped", len(gk))
