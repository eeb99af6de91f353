"""Randomized parity sweep of the DBoW2 transform (Frame::ComputeBoW): random k-ary trees (regular and ragged), feature counts, levelsup,
weighting / normalisation modes against the oracle.  usage: fuzz_vocab.py ncases seed"""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "orb-slam2-annotation_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
import numpy as np, orbb200, pyoracle
from synth import synth_vocabulary, synth_vocabulary_features
ncases = int(sys.argv[1]) if len(sys.argv) > 1 else 50
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 9)
bad = 0
for case in range(ncases):
    k = int(rng.integers(2, 21)); L = int(rng.integers(1, 7))
    while k ** L > 60000: L -= 1
    ragged = bool(rng.random() < 0.4); levelsup = int(rng.integers(0, L + 2)); nfeat = int(rng.choice([0, 1, 31, 500, 3000]))
    seed = int(rng.integers(0, 1 << 30)); weighting = int(rng.integers(0, 4)); norm = int(rng.integers(0, 3))
    tag = f"case {case}: k={k} L={L} ragged={ragged} levelsup={levelsup} nfeat={nfeat} seed={seed} weighting={weighting} norm={norm}"
    try:
        parent, leaf, desc, weight = synth_vocabulary(k, L, seed, ragged=ragged)
        feats = synth_vocabulary_features(desc, leaf, nfeat, seed + 1) if nfeat else np.zeros((0, 32), np.uint8)
        o = pyoracle.OracleVocabulary(k, L, parent, leaf, desc, weight, weighting, norm); g = orbb200.Vocabulary(k, L, parent, leaf, desc, weight, weighting, norm)
        if nfeat:
            for a, b in zip(o.transform(feats, levelsup), g.transform(feats, levelsup)):
                assert np.array_equal(a, b), "transform"
        for a, b in zip(o.bow(feats, levelsup), g.bow(feats, levelsup)):
            assert len(a) == len(b) and (len(a) == 0 or (a.dtype == b.dtype and np.array_equal(a, b))), "bow"
    except Exception as e:
        bad += 1
        print("FAIL", tag, "|", type(e).__name__, str(e)[:200], flush=True)
print("cases", ncases, "failures", bad)
