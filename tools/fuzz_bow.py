"""Randomized parity sweep of the BoW-guided searches on synthetic inputs (random descriptors, random FeatureVectors with shared and
unshared nodes, random MapPoint flags): SearchByBoW (KeyFrame-Frame and KeyFrame-KeyFrame), SearchForTriangulation and
ComputeDistinctiveDescriptors against the oracle.  usage: fuzz_bow.py ncases seed"""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "orb-slam2-annotation_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
import numpy as np, orbb200, pyoracle
ncases = int(sys.argv[1]) if len(sys.argv) > 1 else 50
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 5)


def feature_vector(n, node_ids, assign, keep):
    """CSR FeatureVector: features i with keep[i] grouped by node id assign[i], nodes ascending, features ascending inside a node."""
    nodes, off, feats = [], [0], []
    for nid in node_ids:
        m = np.flatnonzero((assign == nid) & keep)
        if len(m):
            nodes.append(nid); feats.extend(m.tolist()); off.append(len(feats))
    return np.asarray(nodes, np.int32), np.asarray(off, np.int32), np.asarray(feats, np.int32)


bad = 0
for case in range(ncases):
    n1 = int(rng.choice([0, 1, 5, 300, 2000])) if rng.random() < 0.25 else int(rng.integers(1, 2500))
    n2 = int(rng.choice([0, 1, 5, 300, 2000])) if rng.random() < 0.25 else int(rng.integers(1, 2500))
    nnodes = int(rng.choice([1, 3, 40, 300, 4000]))
    tag = f"case {case}: n1={n1} n2={n2} nodes={nnodes}"
    try:
        d1 = rng.integers(0, 256, (n1, 32), dtype=np.uint8)
        d2 = rng.integers(0, 256, (n2, 32), dtype=np.uint8)
        a1 = rng.integers(0, nnodes, n1); a2 = rng.integers(0, nnodes, n2)
        m = min(n1, n2)
        if m:                                                     # true correspondences: noisy copies in the same node
            pick = rng.permutation(m)[: int(m * rng.uniform(0.2, 1.0))]
            noise = rng.integers(0, 256, (len(pick), 32), dtype=np.uint8) & rng.integers(0, 256, (len(pick), 32), dtype=np.uint8) & rng.integers(0, 256, (len(pick), 32), dtype=np.uint8) & rng.integers(0, 256, (len(pick), 32), dtype=np.uint8)
            d2[pick] = d1[pick] ^ noise; a2[pick] = a1[pick]
        ang1 = rng.uniform(0, 360, n1).astype(np.float32); ang2 = rng.uniform(0, 360, n2).astype(np.float32)
        if m: ang2[pick] = ang1[pick]
        node_ids = np.sort(rng.permutation(nnodes * 3)[:nnodes]).astype(np.int32)     # sparse ascending node ids
        fv1 = feature_vector(n1, node_ids, node_ids[a1] if n1 else a1, rng.random(n1) < 0.95)
        fv2 = feature_vector(n2, node_ids, node_ids[a2] if n2 else a2, rng.random(n2) < 0.95)
        has1 = (rng.random(n1) < rng.choice([0.0, 0.5, 1.0])).astype(np.uint8); has2 = (rng.random(n2) < rng.choice([0.0, 0.5, 1.0])).astype(np.uint8)
        chk = bool(rng.random() < 0.5); ratio = float(rng.choice([0.6, 0.75, 0.9]))
        want, nw = pyoracle.search_by_bow(d1, ang1, has1, fv1, d2, ang2, fv2, ratio, chk)
        got, ng = orbb200.search_by_bow(d1, ang1, has1, fv1, d2, ang2, fv2, ratio, chk)
        assert ng == nw and np.array_equal(got, want), "SearchByBoW(KF, F)"
        want, nw = pyoracle.search_by_bow_kf(d1, ang1, has1, fv1, d2, ang2, has2, fv2, ratio, chk)
        got, ng = orbb200.search_by_bow_kf(d1, ang1, has1, fv1, d2, ang2, has2, fv2, ratio, chk)
        assert ng == nw and np.array_equal(got, want), "SearchByBoW(KF, KF)"
        if n1 and n2:
            k1 = np.zeros(n1, orbb200.KP_DTYPE); k2 = np.zeros(n2, orbb200.KP_DTYPE)
            k1["x"] = rng.uniform(20, 620, n1); k1["y"] = rng.uniform(20, 460, n1); k1["octave"] = rng.integers(0, 8, n1); k1["angle"] = ang1
            k2["x"] = rng.uniform(20, 620, n2); k2["y"] = rng.uniform(20, 460, n2); k2["octave"] = rng.integers(0, 8, n2); k2["angle"] = ang2
            if m: k2["x"][pick] = k1["x"][pick] + 3; k2["y"][pick] = k1["y"][pick] + 1
            ur1 = np.where(rng.random(n1) < 0.5, k1["x"] - 10, -1).astype(np.float32); ur2 = np.where(rng.random(n2) < 0.5, k2["x"] - 10, -1).astype(np.float32)
            F12 = np.asarray([[0, 0, 1], [0, 0, -3], [-1, 3, 0]], np.float32)
            sf = (np.float32(1.2) ** np.arange(8)).astype(np.float32); ls2 = (sf * sf).astype(np.float32)
            only = bool(rng.random() < 0.3)
            want, nw = pyoracle.search_for_triangulation(d1, k1, ur1, has1, fv1, d2, k2, ur2, has2, fv2, F12, 200.0, 150.0, sf, ls2, only, chk)
            got, ng = orbb200.search_for_triangulation(d1, k1, ur1, has1, fv1, d2, k2, ur2, has2, fv2, F12, 200.0, 150.0, sf, ls2, only, chk)
            assert ng == nw and np.array_equal(got, want), "SearchForTriangulation"
        if case % 3 == 0:
            npts = int(rng.integers(1, 400)); sizes = rng.integers(0, 40, npts); sizes[rng.integers(0, npts)] = int(rng.choice([1, 2, 150]))
            off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
            rows = rng.integers(0, 256, (int(off[-1]), 32), dtype=np.uint8)
            gi, gm = orbb200.distinctive_descriptors(rows, off); oi, om = pyoracle.distinctive_descriptors(rows, off)
            assert np.array_equal(gi, oi) and np.array_equal(gm, om), "distinctive descriptors"
    except Exception as e:
        bad += 1
        print("FAIL", tag, "|", type(e).__name__, str(e)[:200], flush=True)
print("cases", ncases, "failures", bad)
