"""Phase times of resolve_projection_kernel (build with -DRESOLVE_TIMING as liborbb200_rt.so): runs tools/pair_timing's search and reads the stamps."""
import ctypes as C, os, sys, shutil
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
P = os.path.join(ROOT, "orb-slam2-annotation_b200")
shutil.copy(os.path.join(P, "liborbb200.so"), "/tmp/orig.so"); shutil.copy(os.path.join(P, "liborbb200_rt.so"), os.path.join(P, "liborbb200.so"))
try:
    sys.path[:0] = [P]
    import numpy as np, orbb200
    from synth import synth
    w, h, nf = 1241, 376, 2000
    big = synth(w + 80, h + 16, 8100)
    a = big[8:8 + h, 48:48 + w].copy(); b = big[9:9 + h, 51:51 + w].copy()
    ex = orbb200.ORBextractor(nf, max_size=(w, h))
    k0, d0 = ex(a); k1, d1 = ex(b)
    scale = ex.GetScaleFactors()
    q = np.zeros(len(k0), orbb200.PROJ_QUERY_DTYPE)
    q["q"]["u"] = k0["x"] - 3.0; q["q"]["v"] = k0["y"] - 1.0; q["q"]["r"] = 15.0 * scale[k0["octave"]]
    q["q"]["min_level"] = k0["octave"] - 1; q["q"]["max_level"] = k0["octave"] + 1; q["q"]["desc"] = d0
    q["u_right"] = -1; q["max_err_right"] = 1; q["id"] = np.arange(len(k0)); q["has_obs"] = 1; q["angle"] = k0["angle"]
    F = orbb200.FrameMatcher(k1, d1, (0.0, 0.0, float(w), float(h)))
    m = orbb200.ORBmatcher(0.9, True)
    L = orbb200.lib()
    out = (C.c_ulonglong * 8)()
    for it in range(5):
        n = m.SearchByProjection_frame(F, q, np.full(len(k1), -1, np.int32), np.zeros(len(k1), np.int32), None)[0]
        L.orbb200__resolve_timing(out)
        t = list(out)
        print("matches", n, "rounds", t[5], "us: init %.1f loop %.1f finalize %.1f rot %.1f total %.1f" % (tuple((t[i + 1] - t[i]) / 1e3 for i in range(4)) + ((t[4] - t[0]) / 1e3,)))
finally:
    shutil.copy("/tmp/orig.so", os.path.join(P, "liborbb200.so"))
