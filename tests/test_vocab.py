"""DBoW2 vocabulary descent of Frame::ComputeBoW (TemplatedVocabulary::transform).
CPU: the oracle restatement against an independent numpy statement on seeded synthetic trees.
GPU: orbv_transform / orbv_compute_bow bit-equal (ids, doubles) to the oracle through the C ABI."""
import numpy as np
import pytest


def _numpy_transform(k, L, parent, is_leaf, desc, weight, feats, levelsup):
    """Independent statement: children lists in file order, argmin of popcount distances with first-wins ties."""
    n = len(parent)
    children = [[] for _ in range(n)]
    for i in range(1, n):
        children[parent[i]].append(i)
    word = np.full(n, -1, np.int64)
    word[np.flatnonzero(is_leaf)] = np.arange(int(is_leaf.sum()))
    bits = np.unpackbits(desc, axis=1).astype(np.int16)
    fb = np.unpackbits(feats, axis=1).astype(np.int16)
    wid = np.zeros(len(feats), np.int32); w = np.zeros(len(feats)); nid = np.zeros(len(feats), np.int32)
    for f in range(len(feats)):
        cur, lvl, node_at = 0, 0, 0
        while True:
            lvl += 1
            ch = children[cur]
            d = (bits[ch] != fb[f]).sum(1)
            cur = ch[int(np.argmin(d))]
            if lvl == L - levelsup:
                node_at = cur
            if not children[cur]:
                break
        wid[f], w[f], nid[f] = word[cur], weight[cur], node_at
    return wid, w, nid


def _numpy_bow(wid, w, nid, weighting, norm):
    bow = {}
    fv = {}
    for i in range(len(wid)):
        if not w[i] > 0:
            continue
        if wid[i] not in bow:
            bow[int(wid[i])] = float(w[i])
        elif weighting in (0, 1):
            bow[int(wid[i])] += float(w[i])
        fv.setdefault(int(nid[i]), []).append(i)
    ids = sorted(bow)
    vals = [bow[i] for i in ids]
    if weighting in (0, 1) and norm == 0 and ids:
        vals = [v / float(len(ids)) for v in vals]
    if norm == 1:
        s = 0.0
        for v in vals:
            s += abs(v)
        vals = [v / s for v in vals] if s > 0 else vals
    elif norm == 2:
        s = 0.0
        for v in vals:
            s += v * v
        s = float(np.sqrt(s))
        vals = [v / s for v in vals] if s > 0 else vals
    return ids, vals, fv


CASES = [(10, 3, 9, False, 1), (10, 4, 11, False, 2), (6, 5, 12, True, 3), (20, 2, 13, False, 0), (3, 6, 14, True, 4)]


@pytest.mark.parametrize("k,L,seed,ragged,levelsup", CASES)
def test_oracle_transform_vs_numpy(k, L, seed, ragged, levelsup):
    import pyoracle
    from synth import synth_vocabulary, synth_vocabulary_features
    parent, leaf, desc, weight = synth_vocabulary(k, L, seed, ragged=ragged)
    feats = synth_vocabulary_features(desc, leaf, 150, seed + 100)
    voc = pyoracle.OracleVocabulary(k, L, parent, leaf, desc, weight)
    wid, w, nid = voc.transform(feats, levelsup)
    nw, nwt, nn = _numpy_transform(k, L, parent, leaf, desc, weight, feats, levelsup)
    assert np.array_equal(wid, nw) and np.array_equal(w, nwt) and np.array_equal(nid, nn)
    for weighting, norm in ((0, 1), (1, 2), (2, 1), (3, 0), (0, 0)):
        v = pyoracle.OracleVocabulary(k, L, parent, leaf, desc, weight, weighting, norm)
        ids, vals, fn, fo, ff = v.bow(feats, levelsup)
        eids, evals, efv = _numpy_bow(wid, w, nid, weighting, norm)
        assert list(ids) == eids and list(vals) == evals          # doubles: same additions in the same order
        assert list(fn) == sorted(efv)
        for j, node in enumerate(fn):
            assert list(ff[fo[j]:fo[j + 1]]) == efv[int(node)]


@pytest.mark.gpu
@pytest.mark.parametrize("k,L,seed,ragged,levelsup", CASES)
def test_gpu_transform_and_bow(gpu, k, L, seed, ragged, levelsup):
    import orbb200, pyoracle
    from synth import synth_vocabulary, synth_vocabulary_features
    parent, leaf, desc, weight = synth_vocabulary(k, L, seed, ragged=ragged)
    feats = synth_vocabulary_features(desc, leaf, 2000, seed + 200)
    orc = pyoracle.OracleVocabulary(k, L, parent, leaf, desc, weight)
    voc = orbb200.Vocabulary(k, L, parent, leaf, desc, weight)
    assert voc.words == int(leaf.sum())
    want = orc.transform(feats, levelsup)
    got = voc.transform(feats, levelsup)
    for a, b in zip(want, got):
        assert np.array_equal(a, b)
    for weighting, norm in ((0, 1), (1, 2), (2, 1), (3, 0), (0, 0)):
        o = pyoracle.OracleVocabulary(k, L, parent, leaf, desc, weight, weighting, norm).bow(feats, levelsup)
        g = orbb200.Vocabulary(k, L, parent, leaf, desc, weight, weighting, norm).bow(feats, levelsup)
        for a, b in zip(o, g):
            assert a.dtype == b.dtype and np.array_equal(a, b)
    # no features: empty vectors
    e = voc.bow(np.zeros((0, 32), np.uint8), levelsup)
    assert all(len(x) == 0 for x in e[:3])


@pytest.mark.gpu
def test_gpu_vocabulary_orbslam_shape(gpu):
    """The ORB-SLAM2 vocabulary shape (k = 10, L = 6 -> 1.1 M nodes, 1 M words), levelsup = 4 as Frame::ComputeBoW calls it,
    on a real frame's descriptors: ids and node ids against the oracle, BowVector doubles bit-equal."""
    import orbb200, pyoracle
    from synth import synth, synth_vocabulary
    parent, leaf, desc, weight = synth_vocabulary(10, 6, 21, stop_fraction=0.01)
    assert len(parent) == 1_111_111
    ex = orbb200.ORBextractor(2000, max_size=(1241, 376))
    _, feats = ex(synth(1241, 376, 2), None)
    o = pyoracle.OracleVocabulary(10, 6, parent, leaf, desc, weight).bow(feats, 4)
    g = orbb200.Vocabulary(10, 6, parent, leaf, desc, weight).bow(feats, 4)
    for a, b in zip(o, g):
        assert np.array_equal(a, b)
    assert len(g[0]) > 100 and abs(float(np.sum(np.abs(g[1]))) - 1.0) < 1e-12


def test_vocab_create_rejects_inconsistent_trees():
    """Argument checks run before any device work: they hold on a CPU-only box as well."""
    import ctypes as C
    import orbb200
    L = orbb200.lib()
    h = C.c_void_p()
    parent = np.asarray([0, 0, 0, 1], np.int32); leaf = np.asarray([0, 0, 1, 1], np.uint8)
    desc = np.zeros((4, 32), np.uint8); w = np.ones(4)
    bad_parent = np.asarray([0, 0, 3, 1], np.int32)           # parent after child
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    assert L.orbv_vocab_create(C.byref(h), 2, 2, 4, p(bad_parent), p(leaf), p(desc), p(w), 0, 1, 0) == orbb200.E_INVALID
    assert L.orbv_vocab_create(C.byref(h), 2, 2, 4, p(parent), p(leaf), p(desc), p(w), 7, 1, 0) == orbb200.E_INVALID
    assert L.orbv_vocab_create(C.byref(h), 2, 2, 1, p(parent), p(leaf), p(desc), p(w), 0, 1, 0) == orbb200.E_INVALID


def _bow_pair(seed=31, shift=(3, 1), nf=1000):
    """Two overlapping views extracted with the CPU oracle, a vocabulary trained around their descriptors (k = 8, L = 4),
    and both FeatureVectors at levelsup = 2."""
    import pyoracle
    from synth import synth
    w, h = 640, 480
    big = synth(w + 16, h + 16, seed)
    cur = big[8:8 + h, 8:8 + w].copy()
    last = big[8 + shift[1]:8 + shift[1] + h, 8 + shift[0]:8 + shift[0] + w].copy()
    ex = pyoracle.OracleExtractor(nf)
    kc, dc = ex.extract(cur)
    kl, dl = ex.extract(last)
    # a vocabulary whose nodes are real descriptors: k-ary tree by recursive random seeding (deterministic)
    rng = np.random.default_rng(seed)
    pool = np.concatenate([dc, dl])
    k, L = 8, 4
    parent, leaf, desc, weight = [0], [0], [np.zeros(32, np.uint8)], [0.0]
    frontier = [(0, np.arange(len(pool)), 0)]
    while frontier:
        nxt = []
        for nid, members, lvl in frontier:
            centres = pool[rng.choice(members, k, replace=len(members) < k)]
            d = (np.unpackbits(pool[members][:, None, :] ^ centres[None, :, :], axis=2)).sum(2)
            assign = d.argmin(1)
            for c in range(k):
                cid = len(parent)
                parent.append(nid); desc.append(centres[c]); is_leaf = lvl + 1 == L
                leaf.append(1 if is_leaf else 0); weight.append(float(rng.uniform(0.5, 5.0)) if is_leaf else 0.0)
                sub = members[assign == c]
                if not is_leaf:
                    nxt.append((cid, sub if len(sub) else members[:1], lvl + 1))
        frontier = nxt
    voc = (k, L, np.asarray(parent, np.int32), np.asarray(leaf, np.uint8), np.stack(desc), np.asarray(weight))
    return (kc, dc), (kl, dl), voc


def test_oracle_search_by_bow_properties():
    """The SearchByBoW restatement: every match joins two features of the same node, passes TH_LOW, and no Frame feature is
    used twice; with all map points bad nothing matches."""
    import pyoracle
    (kc, dc), (kl, dl), (k, L, parent, leaf, desc, weight) = _bow_pair()
    voc = pyoracle.OracleVocabulary(k, L, parent, leaf, desc, weight)
    fv_kf = voc.bow(dl, 2)[2:]
    fv_f = voc.bow(dc, 2)[2:]
    has = np.ones(len(dl), np.uint8)
    m, n = pyoracle.search_by_bow(dl, kl["angle"], has, fv_kf, dc, kc["angle"], fv_f, 0.7, True)
    assert n == int((m >= 0).sum()) and n > 100
    node_of_kf = np.full(len(dl), -1); node_of_f = np.full(len(dc), -1)
    for j, node in enumerate(fv_kf[0]):
        node_of_kf[fv_kf[2][fv_kf[1][j]:fv_kf[1][j + 1]]] = node
    for j, node in enumerate(fv_f[0]):
        node_of_f[fv_f[2][fv_f[1][j]:fv_f[1][j + 1]]] = node
    idx = np.flatnonzero(m >= 0)
    assert np.array_equal(node_of_f[idx], node_of_kf[m[idx]])
    assert np.all(np.unpackbits(dc[idx] ^ dl[m[idx]], axis=1).sum(1) <= 50)
    assert len(set(m[idx].tolist())) == len(idx)                      # a KeyFrame feature matches at most one Frame feature
    m0, n0 = pyoracle.search_by_bow(dl, kl["angle"], np.zeros(len(dl), np.uint8), fv_kf, dc, kc["angle"], fv_f, 0.7, True)
    assert n0 == 0 and np.all(m0 == -1)


@pytest.mark.gpu
@pytest.mark.parametrize("check_ori,ratio", [(True, 0.7), (False, 0.9), (True, 0.6)])
def test_gpu_search_by_bow(gpu, check_ori, ratio):
    import orbb200, pyoracle
    (kc, dc), (kl, dl), (k, L, parent, leaf, desc, weight) = _bow_pair()
    voc = orbb200.Vocabulary(k, L, parent, leaf, desc, weight)
    orc = pyoracle.OracleVocabulary(k, L, parent, leaf, desc, weight)
    fv_kf = voc.bow(dl, 2)[2:]
    fv_f = voc.bow(dc, 2)[2:]
    for a, b in zip(fv_kf, orc.bow(dl, 2)[2:]):
        assert np.array_equal(a, b)
    rng = np.random.default_rng(5)
    has = (rng.random(len(dl)) < 0.8).astype(np.uint8)                # some features without a (good) MapPoint
    want, nw = pyoracle.search_by_bow(dl, kl["angle"], has, fv_kf, dc, kc["angle"], fv_f, ratio, check_ori)
    got, ng = orbb200.search_by_bow(dl, kl["angle"], has, fv_kf, dc, kc["angle"], fv_f, ratio, check_ori)
    assert ng == nw and np.array_equal(got, want) and ng > 50
    # empty feature vectors: no matches, no device work
    e = (np.zeros(0, np.int32), np.zeros(1, np.int32), np.zeros(0, np.int32))
    g0, n0 = orbb200.search_by_bow(dl, kl["angle"], has, e, dc, kc["angle"], fv_f, ratio, check_ori)
    assert n0 == 0 and np.all(g0 == -1)


@pytest.mark.gpu
@pytest.mark.parametrize("check_ori,ratio", [(True, 0.75), (False, 0.9)])
def test_gpu_search_by_bow_keyframes(gpu, check_ori, ratio):
    """ORBmatcher::SearchByBoW(KeyFrame, KeyFrame) (ORBmatcher.cpp:604-743): MapPoint flags on both sides, vbMatched2."""
    import orbb200, pyoracle
    (kc, dc), (kl, dl), (k, L, parent, leaf, desc, weight) = _bow_pair(seed=33)
    voc = orbb200.Vocabulary(k, L, parent, leaf, desc, weight)
    fv1 = voc.bow(dl, 2)[2:]
    fv2 = voc.bow(dc, 2)[2:]
    rng = np.random.default_rng(6)
    has1 = (rng.random(len(dl)) < 0.7).astype(np.uint8); has2 = (rng.random(len(dc)) < 0.7).astype(np.uint8)
    want, nw = pyoracle.search_by_bow_kf(dl, kl["angle"], has1, fv1, dc, kc["angle"], has2, fv2, ratio, check_ori)
    got, ng = orbb200.search_by_bow_kf(dl, kl["angle"], has1, fv1, dc, kc["angle"], has2, fv2, ratio, check_ori)
    assert ng == nw and np.array_equal(got, want) and ng > 30
    idx = np.flatnonzero(got >= 0)
    assert np.all(has1[idx] == 1) and np.all(has2[got[idx]] == 1) and len(set(got[idx].tolist())) == len(idx)


@pytest.mark.gpu
@pytest.mark.parametrize("only_stereo,check_ori", [(False, True), (True, True), (False, False)])
def test_gpu_search_for_triangulation(gpu, only_stereo, check_ori):
    """ORBmatcher::SearchForTriangulation (ORBmatcher.cpp:755-927).  The second view is the first translated by (3, 1) px, so the
    fundamental matrix of a pure image-plane translation F12 = [t]x, t = (3, 1, 0), puts true matches on their epipolar lines."""
    import orbb200, pyoracle
    (kc, dc), (kl, dl), (k, L, parent, leaf, desc, weight) = _bow_pair(seed=35)
    voc = orbb200.Vocabulary(k, L, parent, leaf, desc, weight)
    fv1 = voc.bow(dl, 2)[2:]; fv2 = voc.bow(dc, 2)[2:]
    rng = np.random.default_rng(7)
    has1 = (rng.random(len(dl)) < 0.3).astype(np.uint8); has2 = (rng.random(len(dc)) < 0.3).astype(np.uint8)   # already triangulated
    ur1 = np.where(rng.random(len(dl)) < 0.5, kl["x"] - 10, -1).astype(np.float32)
    ur2 = np.where(rng.random(len(dc)) < 0.5, kc["x"] - 10, -1).astype(np.float32)
    F12 = np.asarray([[0, 0, 1], [0, 0, -3], [-1, 3, 0]], np.float32)
    sf = (np.float32(1.2) ** np.arange(8)).astype(np.float32); ls2 = (sf * sf).astype(np.float32)
    ex, ey = 200.0, 150.0                                  # an epipole inside the image: nearby mono candidates are rejected
    want, nw = pyoracle.search_for_triangulation(dl, kl, ur1, has1, fv1, dc, kc, ur2, has2, fv2, F12, ex, ey, sf, ls2, only_stereo, check_ori)
    got, ng = orbb200.search_for_triangulation(dl, kl, ur1, has1, fv1, dc, kc, ur2, has2, fv2, F12, ex, ey, sf, ls2, only_stereo, check_ori)
    assert ng == nw and np.array_equal(got, want) and ng > 30
    idx = np.flatnonzero(got >= 0)
    assert np.all(has1[idx] == 0) and np.all(has2[got[idx]] == 0)
    if only_stereo:
        assert np.all(ur1[idx] >= 0) and np.all(ur2[got[idx]] >= 0)
    # most accepted pairs are the true correspondences (x2 = x1 + (3, 1))
    dx = kc["x"][got[idx]] - kl["x"][idx]; dy = kc["y"][got[idx]] - kl["y"][idx]
    assert np.mean((np.abs(dx - 3) < 1.5) & (np.abs(dy - 1) < 1.5)) > 0.5
