"""Development probe run on the GPU box: prints mismatch diagnostics instead of stopping at the first assert."""
import sys, os, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "orb-slam2-annotation_b200"), os.path.join(ROOT, "oracle")]
import numpy as np
import orbb200, pyoracle
from synth import synth, synth_descriptor_db, synth_queries

def srt(a):
    a = np.asarray(a).reshape(-1, 3).astype(np.int64)
    return a[np.lexsort((a[:, 2], a[:, 0], a[:, 1]))]

for (w, h, nf, seed) in [(640, 480, 1000, 1), (1241, 376, 2000, 2), (752, 480, 1200, 3), (100, 80, 500, 4)]:
    img = synth(w, h, seed)
    orc = pyoracle.OracleExtractor(nf); ok, od = orc.extract(img)
    ex = orbb200.ORBextractor(nf, max_size=(w, h))
    t = time.time(); gk, gd = ex(img); t1 = time.time() - t
    t = time.time(); gk, gd = ex(img); t2 = time.time() - t
    print(f"== {w}x{h} nf={nf}: oracle n={len(ok)} gpu n={len(gk)}  first call {t1*1e3:.2f} ms, second {t2*1e3:.2f} ms")
    for L in range(8):
        pe = np.array_equal(ex.pyramid_level(L), orc.level_image(L))
        bg, bo = ex.blurred_level(L), orc.level_image(L, True)
        be = np.array_equal(bg, bo)
        ro, rg = srt(orc.level_raw(L)), srt(ex.fast_keypoints(L))
        fe = ro.shape == rg.shape and np.array_equal(ro, rg)
        lo, lg = orc.level_kps(L), ex.level_keypoints(L)
        oe = len(lo) == len(lg) and all(np.array_equal(lo[f], lg[f]) for f in ("x", "y", "response"))
        ae = len(lo) == len(lg) and (np.max(np.abs(lo["angle"] - lg["angle"]), initial=0))
        print(f"  L{L}: pyr {pe} blur {be} fast {fe} ({len(ro)} vs {len(rg)}) octree {oe} ({len(lo)} vs {len(lg)}) max|dangle| {ae}")
        if not be:
            d = np.argwhere(bg != bo); print("     blur diffs", len(d), d[:5].tolist())
        if not fe:
            so = set(map(tuple, ro.tolist())); sg = set(map(tuple, rg.tolist()))
            print("     only oracle", sorted(so - sg)[:8], " only gpu", sorted(sg - so)[:8])
        if not oe and len(lo) and len(lg):
            so = set(zip(lo["x"].tolist(), lo["y"].tolist())); sg = set(zip(lg["x"].tolist(), lg["y"].tolist()))
            print("     octree set diff", len(so - sg), len(sg - so), "first order mismatch at",
                  next((i for i in range(min(len(lo), len(lg))) if (lo["x"][i], lo["y"][i]) != (lg["x"][i], lg["y"][i])), None))
    if len(ok) == len(gk):
        same = all(np.array_equal(ok[f], gk[f]) for f in ("x", "y", "size", "response", "octave", "class_id"))
        print("  final kps fields equal:", same, " angle max diff", np.max(np.abs(ok["angle"] - gk["angle"]), initial=0),
              " desc rows differing", int(np.any(od != gd, axis=1).sum()), "of", len(od))

# matcher quick checks
rng = np.random.default_rng(0)
a = rng.integers(0, 256, (1000, 32), dtype=np.uint8); b = rng.integers(0, 256, (1000, 32), dtype=np.uint8)
print("distance pairs equal:", np.array_equal(orbb200.DescriptorDistance(a, b), pyoracle.distance_pairs(a, b)))
db = synth_descriptor_db(200_000); q = synth_queries(db, 777)
for variant in ("0", "1"):
    os.environ["ORBB_BF_VARIANT"] = variant
    D = orbb200.DescriptorDB(len(db)); D.append(db)
    t = time.time(); d1, i1, d2 = D.top2(q); dt = time.time() - t
    o1, oi, o2 = pyoracle.bf_top2(db, q, nthreads=8)
    print(f"bf_top2 variant {variant}: d1 {np.array_equal(d1, o1)} idx {np.array_equal(i1, oi)} d2 {np.array_equal(d2, o2)}  {dt*1e3:.1f} ms")
    D.close()
print("popc peak Gpopc/s:", orbb200.popc_peak())
print("launches:", orbb200.launch_count())
