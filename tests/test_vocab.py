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
