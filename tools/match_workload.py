"""Guided-matching workload at the C2 shape (KITTI 1241x376, 2000 kp): stereo matching, tracked-frame search (window search + device
resolve), grid build.  For ncu (tools/gpu.sh ncu-match) and for CUDA-event timing of the kernel families (prints one JSON object)."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "orb-slam2-annotation_b200")]
import numpy as np, torch, orbb200
from synth import synth
w, h, nf = 1241, 376, 2000
big = synth(w + 80, h + 16, 8100)
pairs = []
for (ox, oy) in ((8, 8), (11, 9)):
    pairs.append((big[oy:oy + h, ox + 40:ox + 40 + w].copy(), big[oy:oy + h, ox + 28:ox + 28 + w].copy()))
exL = orbb200.ORBextractor(nf, max_size=(w, h)); exR = orbb200.ORBextractor(nf, max_size=(w, h))
scale = exL.GetScaleFactors()
out = {}
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
for it in range(reps):
    res = []
    for (L, R) in pairs:
        kL, dL = exL(L); kR, dR = exR(R)
        ur, dp = orbb200.stereo_match(exL, exR, kL, dL, kR, dR, 386.1448, 0.0)
        res.append((kL, dL, ur))
    (k0, d0, u0), (k1, d1, u1) = res
    q = np.zeros(len(k0), orbb200.PROJ_QUERY_DTYPE)
    q["q"]["u"] = k0["x"] - 3.0; q["q"]["v"] = k0["y"] - 1.0; q["q"]["r"] = 15.0 * scale[k0["octave"]]
    q["q"]["min_level"] = k0["octave"] - 1; q["q"]["max_level"] = k0["octave"] + 1; q["q"]["desc"] = d0
    q["u_right"] = q["q"]["u"] - 12.0; q["max_err_right"] = 15.0 * scale[k0["octave"]]
    q["id"] = np.arange(len(k0)); q["has_obs"] = 1; q["angle"] = k0["angle"]
    F = orbb200.FrameMatcher(k1, d1, (0.0, 0.0, float(w), float(h)))
    m = orbb200.ORBmatcher(0.9, True)
    n, mp = m.SearchByProjection_frame(F, q, np.full(len(k1), -1, np.int32), np.zeros(len(k1), np.int32), u1)[:2]
    F.close()
out["keypoints"] = int(len(k0)); out["stereo_matches"] = int((u0 > 0).sum()); out["projection_matches"] = int(n)
print(json.dumps(out))
