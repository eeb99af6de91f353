"""Small workload touching every kernel with tight buffers (max_batch == frames, max_size == frame size): a plain smoke run."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "orb-slam2-annotation_b200")]
import numpy as np, orbb200
from synth import synth, synth_vocabulary, synth_vocabulary_features, synth_descriptor_db, synth_queries
for (w, h, nf, F) in [(640, 480, 1000, 2), (203, 177, 300, 1), (752, 480, 1200, 3), (97, 61, 100, 1)]:
    ex = orbb200.ORBextractor(nf, max_size=(w, h), max_batch=F)
    imgs = np.stack([synth(w, h, 40 + i) for i in range(F)])
    out = ex.extract_batch(imgs)
    print(w, h, [len(k) for k, _ in out])
    col = np.stack([imgs, imgs[:, ::-1].copy(), imgs], axis=3)
    out2 = ex.extract_batch_color(np.ascontiguousarray(col), rgb_order=True)
    print("colour", [len(k) for k, _ in out2])
k, d = out[0]
rng = np.random.default_rng(1)
sizes = rng.integers(0, 40, 200); off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
desc = rng.integers(0, 256, (int(off[-1]), 32), dtype=np.uint8)
print("distinctive", orbb200.distinctive_descriptors(desc, off)[0][:5])
parent, leaf, vd, wt = synth_vocabulary(10, 3, 9)
voc = orbb200.Vocabulary(10, 3, parent, leaf, vd, wt)
feats = synth_vocabulary_features(vd, leaf, 500, 3)
b = voc.bow(feats, 1)
print("bow", len(b[0]), len(b[2]))
m, n = orbb200.search_by_bow(feats, np.zeros(500, np.float32), np.ones(500, np.uint8), b[2:], feats[::-1].copy(), np.zeros(500, np.float32),
                             voc.bow(feats[::-1].copy(), 1)[2:], 0.9, False)
print("search_by_bow", n)
db = synth_descriptor_db(50_000); q = synth_queries(db, 100)
D = orbb200.DescriptorDB(len(db)); D.append(db)
print("top2", D.top2(q)[0][:4])
print("done")
