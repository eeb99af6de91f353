"""One configuration of tests/test_gpu_extract.py::test_random_shapes_and_parameters, stage by stage against the oracle.
usage: fuzz_case.py w h nlevels scale ini min nf seed"""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "orb-slam2-annotation_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
import numpy as np, orbb200, pyoracle
from synth import synth
w, h, nl, sf, ini, mn, nf, seed = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), float(np.float32(float(sys.argv[4]))), int(sys.argv[5]), int(sys.argv[6]), int(sys.argv[7]), int(sys.argv[8])
img = synth(w, h, seed)
orc = pyoracle.OracleExtractor(nf, sf, nl, ini, mn); ex = orbb200.ORBextractor(nf, sf, nl, ini, mn, max_size=(w, h))
ok, od = orc.extract(img); k, d = ex(img)
print("final", len(k), len(ok))
def srt(a):
    a = np.asarray(a).reshape(-1, 3).astype(np.int64); return a[np.lexsort((a[:, 2], a[:, 0], a[:, 1]))]
for L in range(nl):
    ro, rg = srt(orc.level_raw(L)), srt(ex.fast_keypoints(L))
    same = ro.shape == rg.shape and np.array_equal(ro, rg)
    lo, lg = orc.level_kps(L), ex.level_keypoints(L)
    print("level", L, "FAST", len(ro), len(rg), "equal" if same else "DIFFER", "octree", len(lo), len(lg))
    if not same:
        so = set(map(tuple, ro.tolist())); sg = set(map(tuple, rg.tolist()))
        print("  only oracle:", sorted(so - sg)[:12]); print("  only gpu:", sorted(sg - so)[:12])
