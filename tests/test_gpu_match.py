"""GPU parity of the Hamming-matching half: distances, grid window search, Search* functions, stereo, brute-force top-2.
Every comparison is bit-exact (integer / index work) against the CPU oracle, through the C ABI."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _frames(w=640, h=480, nf=1000, seed=11, shift=(3, 1)):
    """Two overlapping views: `cur` and `last` (= cur translated by `shift`), extracted with the CPU oracle."""
    import pyoracle
    from synth import synth
    big = synth(w + 16, h + 16, seed)
    cur = big[8:8 + h, 8:8 + w].copy()
    last = big[8 + shift[1]:8 + shift[1] + h, 8 + shift[0]:8 + shift[0] + w].copy()
    ex = pyoracle.OracleExtractor(nf)
    kc, dc = ex.extract(cur)
    kl, dl = ex.extract(last)
    scale = ex.tables()["scale"]
    return (cur, kc, dc), (last, kl, dl), scale


def test_descriptor_distance(gpu):
    import orbb200, pyoracle
    rng = np.random.default_rng(5)
    a = rng.integers(0, 256, (4097, 32), dtype=np.uint8)
    b = rng.integers(0, 256, (4097, 32), dtype=np.uint8)
    b[:10] = a[:10]                       # distance 0
    b[10:20] = ~a[10:20]                  # distance 256
    g = orbb200.DescriptorDistance(a, b)
    assert np.array_equal(g, pyoracle.distance_pairs(a, b))
    assert np.array_equal(g, np.unpackbits(a ^ b, axis=1).sum(1))
    assert g[:10].max() == 0 and g[10:20].min() == 256
    assert len(orbb200.DescriptorDistance(a[:0], b[:0])) == 0


def _queries(kl, dl, scale, th, shift, rng, levels="pm1"):
    import orbb200
    q = np.zeros(len(kl), orbb200.PROJ_QUERY_DTYPE)
    q["q"]["u"] = kl["x"] + shift[0]
    q["q"]["v"] = kl["y"] + shift[1]
    q["q"]["r"] = th * scale[kl["octave"]]
    if levels == "pm1":
        q["q"]["min_level"] = kl["octave"] - 1; q["q"]["max_level"] = kl["octave"] + 1
    elif levels == "fwd":
        q["q"]["min_level"] = kl["octave"]; q["q"]["max_level"] = -1
    elif levels == "back":
        q["q"]["min_level"] = 0; q["q"]["max_level"] = kl["octave"]
    else:
        q["q"]["min_level"] = kl["octave"] - 1; q["q"]["max_level"] = kl["octave"]
    q["q"]["desc"] = dl
    q["u_right"] = q["q"]["u"] - rng.uniform(0, 30, len(kl)).astype(np.float32)
    q["max_err_right"] = q["q"]["r"]
    q["id"] = np.arange(len(kl)); q["has_obs"] = (rng.random(len(kl)) < 0.9).astype(np.int32)
    q["angle"] = kl["angle"]
    return q


def _to_oracle_queries(q):
    import pyoracle
    o = np.zeros(len(q), pyoracle.QUERY_DTYPE)
    o["u"] = q["q"]["u"]; o["v"] = q["q"]["v"]; o["r"] = q["q"]["r"]
    o["minLevel"] = q["q"]["min_level"]; o["maxLevel"] = q["q"]["max_level"]; o["desc"] = q["q"]["desc"]
    o["uR"] = q["u_right"]; o["maxErrR"] = q["max_err_right"]; o["id"] = q["id"]; o["hasObs"] = q["has_obs"]; o["angle"] = q["angle"]
    return o


@pytest.mark.parametrize("levels", ["pm1", "fwd", "back", "pred"])
def test_window_search_matches_GetFeaturesInArea(gpu, levels):
    import orbb200, pyoracle
    (cur, kc, dc), (last, kl, dl), scale = _frames()
    rng = np.random.default_rng(3)
    h, w = cur.shape
    F = orbb200.FrameMatcher(kc, dc, (0, 0, w, h))
    O = pyoracle.OracleFrame(kc, dc, (0, 0, w, h))
    q = _queries(kl, dl, scale, 15.0, (3, 1), rng, levels)
    # edge cases: windows fully outside, huge radius, zero radius, queries on the image border
    extra = q[:6].copy()
    extra["q"]["u"] = [-500, 5000, 0, w, w / 2, w / 2]; extra["q"]["v"] = [h / 2, h / 2, 0, h, -400, h / 2]
    extra["q"]["r"] = [10, 10, 50, 50, 10, 1e4]; extra["q"]["min_level"] = -1; extra["q"]["max_level"] = -1
    q = np.concatenate([q, extra])
    offs, idx, dist = F.window_search(q["q"])
    oo, oi, od = O.window_search(_to_oracle_queries(q))
    assert np.array_equal(offs, oo)
    assert np.array_equal(idx, oi), "candidate ORDER must equal GetFeaturesInArea's (ix, iy, insertion)"
    assert np.array_equal(dist, od)
    assert offs[-1] - offs[-2] == len(kc)          # the 1e4-radius query sees every keypoint, in grid order


def test_search_by_projection_last_frame(gpu):
    import orbb200, pyoracle
    (cur, kc, dc), (last, kl, dl), scale = _frames()
    rng = np.random.default_rng(8)
    h, w = cur.shape
    F = orbb200.FrameMatcher(kc, dc, (0, 0, w, h)); O = pyoracle.OracleFrame(kc, dc, (0, 0, w, h))
    ur = np.where(rng.random(len(kc)) < 0.5, kc["x"] - rng.uniform(0, 30, len(kc)), -1).astype(np.float32)
    for th, check in [(7.0, True), (15.0, True), (15.0, False)]:
        q = _queries(kl, dl, scale, th, (3, 1), rng)
        mp0 = np.full(len(kc), -1, np.int32); ob0 = np.zeros(len(kc), np.int32)
        mp0[::17] = 9999; ob0[::17] = 1; mp0[5::23] = 7777          # pre-assigned points with / without observations
        for uright in (None, ur):
            n, mp, ob = orbb200.ORBmatcher(0.9, check).SearchByProjection_frame(F, q, mp0, ob0, uright)
            on, omp, oob = O.search_last_frame(_to_oracle_queries(q), check, mp0, ob0, uright)
            assert n == on and np.array_equal(mp, omp) and np.array_equal(ob, oob)
            assert n > 0.3 * len(kl)                                 # the scenario really matches something


def test_search_by_projection_local_points(gpu):
    import orbb200, pyoracle
    (cur, kc, dc), (last, kl, dl), scale = _frames(seed=12)
    rng = np.random.default_rng(9)
    h, w = cur.shape
    F = orbb200.FrameMatcher(kc, dc, (0, 0, w, h)); O = pyoracle.OracleFrame(kc, dc, (0, 0, w, h))
    ur = np.where(rng.random(len(kc)) < 0.5, kc["x"] - rng.uniform(0, 30, len(kc)), -1).astype(np.float32)
    for th in (1.0, 3.0):
        q = _queries(kl, dl, scale, 1.0, (3, 1), rng, "pred")
        rview = np.where(rng.random(len(kl)) < 0.5, 2.5, 4.0).astype(np.float32)     # RadiusByViewingCos
        q["q"]["r"] = rview * (th if th != 1.0 else 1.0) * scale[kl["octave"]]
        q["max_err_right"] = q["q"]["r"]
        mp0 = np.full(len(kc), -1, np.int32); ob0 = np.zeros(len(kc), np.int32)
        mp0[::19] = 4242; ob0[::19] = 1
        for nnratio in (0.8, 0.6):
            for uright in (None, ur):
                n, mp, ob = orbb200.ORBmatcher(nnratio).SearchByProjection_points(F, q, mp0, ob0, uright)
                on, omp, oob = O.search_local_points(_to_oracle_queries(q), nnratio, mp0, ob0, uright)
                assert n == on and np.array_equal(mp, omp) and np.array_equal(ob, oob)
    assert n > 50


def test_search_for_initialization(gpu):
    import orbb200, pyoracle
    (cur, kc, dc), (last, kl, dl), scale = _frames(nf=2000, seed=13, shift=(6, -4))
    h, w = cur.shape
    F1 = orbb200.FrameMatcher(kl, dl, (0, 0, w, h)); F2 = orbb200.FrameMatcher(kc, dc, (0, 0, w, h))
    O1 = pyoracle.OracleFrame(kl, dl, (0, 0, w, h)); O2 = pyoracle.OracleFrame(kc, dc, (0, 0, w, h))
    prev = np.stack([kl["x"], kl["y"]], 1).astype(np.float32)
    for check in (True, False):
        n, m12, pm = orbb200.ORBmatcher(0.9, check).SearchForInitialization(F1, F2, prev, 100)
        on, om12, opm = pyoracle.search_for_initialization(O1, O2, prev, 100, 0.9, check)
        assert n == on and np.array_equal(m12, om12) and np.array_equal(pm, opm.reshape(pm.shape))
        assert n > 100


@pytest.mark.parametrize("w,h,nf,seed", [(1241, 376, 2000, 21), (752, 480, 1200, 22)])
def test_stereo_matches(gpu, w, h, nf, seed):
    """Frame::ComputeStereoMatches: uRight and depth bit-equal to the oracle (C2 / C3 shapes)."""
    import orbb200, pyoracle
    from synth import synth_stereo
    L, R = synth_stereo(w, h, seed)
    oL, oR = pyoracle.OracleExtractor(nf), pyoracle.OracleExtractor(nf)
    okL, odL = oL.extract(L); okR, odR = oR.extract(R)
    gL, gR = orbb200.ORBextractor(nf, max_size=(w, h)), orbb200.ORBextractor(nf, max_size=(w, h))
    kL, dL = gL(L); kR, dR = gR(R)
    assert kL.tobytes() == okL.tobytes() and kR.tobytes() == okR.tobytes()
    mbf = 386.1448 if w == 1241 else 47.90639384423901           # KITTI00-02.yaml / EuRoC.yaml Camera.bf
    for mb in (0.0, 0.537):                                       # reference call (mb == 0) and a finite baseline
        ur, dp = orbb200.stereo_match(gL, gR, kL, dL, kR, dR, mbf, mb)
        with np.errstate(divide="ignore"):
            our, odp, _, _ = pyoracle.stereo_match(oL, oR, mbf, mb)
        assert np.array_equal(ur, our) and np.array_equal(dp, odp)
    assert (ur >= 0).sum() > 0.2 * len(kL)


def test_bf_top2_small_and_edges(gpu):
    import orbb200, pyoracle
    from synth import synth_descriptor_db, synth_queries
    db = synth_descriptor_db(50_000, seed=1)
    db[777] = db[123]; db[40_000] = db[123]                       # exact duplicates: first index must win, d2 == d1
    q = synth_queries(db, 1037, seed=2)                           # not a multiple of the 512-query tile
    q[0] = db[123]
    for nrows in (0, 1, 2, 127, 128, 129, 4097, 50_000):
        D = orbb200.DescriptorDB(max(nrows, 1)); D.append(db[:nrows])
        d1, i1, d2 = D.top2(q)
        o1, oi, o2 = pyoracle.bf_top2(db[:nrows], q, nthreads=8)
        assert np.array_equal(d1, o1) and np.array_equal(i1, oi) and np.array_equal(d2, o2), nrows
        D.close()
    assert d1[0] == 0 and i1[0] == 123 and d2[0] == 0
    assert len(orbb200.DescriptorDB(4).top2(q[:0])[0]) == 0


@pytest.mark.parametrize("variant", ["0", "1", "2", "3", "4"])
def test_bf_top2_every_kernel_variant(gpu, variant, monkeypatch):
    """Every brute-force kernel (plain POPC, carry-save integer pipe, legacy mma.sync IMMA, tcgen05 + TMEM on one CTA, tcgen05 cta_group::2 on a CTA pair) against the oracle: duplicates (first index wins, d2 == d1), all-zero and all-one descriptors
    (|q| = 0 / 256: the extremes of hamming = |q| - sum a.b), ragged query and row counts around the 128-row stage and the
    256-query tile, and row counts that leave empty trailing chunks."""
    import orbb200, pyoracle
    from synth import synth_descriptor_db, synth_queries
    monkeypatch.setenv("ORBB_BF_VARIANT", variant)
    db = synth_descriptor_db(70_001, seed=11)
    db[5] = 0; db[6] = 255; db[60_000] = db[321]; db[69_999] = db[321]
    q = synth_queries(db, 515, seed=12)
    q[0] = db[321]; q[1] = 0; q[2] = 255; q[3] = db[70_000]
    for nq, nrows in ((515, 70_001), (1, 70_001), (256, 129), (257, 128), (300, 1), (64, 383), (64, 255), (300, 257), (515, 33_000)):
        D = orbb200.DescriptorDB(max(nrows, 1)); D.append(db[:nrows])
        d1, i1, d2 = D.top2(q[:nq])
        o1, oi, o2 = pyoracle.bf_top2(db[:nrows], q[:nq], nthreads=8)
        assert np.array_equal(d1, o1) and np.array_equal(i1, oi) and np.array_equal(d2, o2), (variant, nq, nrows)
        D.close()
    D = orbb200.DescriptorDB(70_001); D.append(db)
    d1, i1, d2 = D.top2(q)
    assert d1[0] == 0 and i1[0] == 321 and d2[0] == 0 and d1[1] == 0 and i1[1] == 5 and d1[2] == 0 and i1[2] == 6 and i1[3] == 70_000
    D.close()


def test_bf_top2_sharded_merge_equals_single(gpu):
    """Shard the DB into contiguous ranges, merge per-shard top-2 on the device (the post-all-gather step):
    identical to the unsharded scan and to the oracle."""
    import torch, orbb200, pyoracle
    from synth import synth_descriptor_db, synth_queries
    db = synth_descriptor_db(300_000, seed=3)
    db[250_000] = db[10]                                          # a tie that straddles shards
    q = synth_queries(db, 2000, seed=4); q[1] = db[10]
    o1, oi, o2 = pyoracle.bf_top2(db, q, nthreads=8)
    for G in (1, 3, 8):
        bounds = np.linspace(0, len(db), G + 1).astype(np.int64)
        d1a = torch.empty((G, len(q)), dtype=torch.int32, device="cuda"); i1a = torch.empty((G, len(q)), dtype=torch.int64, device="cuda")
        d2a = torch.empty((G, len(q)), dtype=torch.int32, device="cuda")
        qd = torch.from_numpy(q).cuda()
        shards = []
        for g in range(G):
            D = orbb200.DescriptorDB(int(bounds[g + 1] - bounds[g]), row_offset=int(bounds[g])); D.append(db[bounds[g]:bounds[g + 1]])
            D.top2_device(qd.data_ptr(), len(q), d1a[g].data_ptr(), i1a[g].data_ptr(), d2a[g].data_ptr())
            shards.append(D)
        d1 = torch.empty(len(q), dtype=torch.int32, device="cuda"); i1 = torch.empty(len(q), dtype=torch.int64, device="cuda")
        d2 = torch.empty(len(q), dtype=torch.int32, device="cuda")
        orbb200.merge_top2_device(d1a.data_ptr(), i1a.data_ptr(), d2a.data_ptr(), G, len(q), d1.data_ptr(), i1.data_ptr(), d2.data_ptr())
        torch.cuda.synchronize()
        assert np.array_equal(d1.cpu().numpy(), o1) and np.array_equal(i1.cpu().numpy(), oi) and np.array_equal(d2.cpu().numpy(), o2), G
    acc = orbb200.ratio_test(o1, o2, 50, 0.75)
    assert 0.3 * len(q) < acc.sum() < 0.6 * len(q)                # about the planted half passes the ratio test


def test_bf_top2_full_size_properties(gpu):
    """C5 at one GPU's full size (10 M rows): planted queries are found at distance 0 with the FIRST duplicate's index,
    d1 <= d2 everywhere, and a sample of queries equals the CPU scan of the whole database."""
    import torch, orbb200, pyoracle
    M = 10_000_000
    g = torch.Generator(device="cuda"); g.manual_seed(42)
    dbt = torch.randint(0, 256, (M, 32), dtype=torch.uint8, device="cuda", generator=g)
    rows = torch.tensor([5, 123_456, 9_999_999, 5_000_000], device="cuda")
    dbt[7_000_000] = dbt[123_456]                                 # duplicate later in the DB
    q = torch.randint(0, 256, (2000, 32), dtype=torch.uint8, device="cuda", generator=g)
    q[:4] = dbt[rows]
    D = orbb200.DescriptorDB(M); D.append_device(dbt.data_ptr(), M)
    d1 = torch.empty(2000, dtype=torch.int32, device="cuda"); i1 = torch.empty(2000, dtype=torch.int64, device="cuda")
    d2 = torch.empty(2000, dtype=torch.int32, device="cuda")
    D.top2_device(q.data_ptr(), 2000, d1.data_ptr(), i1.data_ptr(), d2.data_ptr())
    torch.cuda.synchronize()
    d1, i1, d2 = d1.cpu().numpy(), i1.cpu().numpy(), d2.cpu().numpy()
    assert np.array_equal(d1[:4], [0, 0, 0, 0]) and np.array_equal(i1[:4], [5, 123_456, 9_999_999, 5_000_000])
    assert d2[1] == 0                                             # the duplicate is the second best
    assert np.all(d1 <= d2) and np.all((i1 >= 0) & (i1 < M))
    dbh = dbt.cpu().numpy(); qh = q.cpu().numpy()
    sel = np.array([0, 1, 2, 3, 500, 1999])
    o1, oi, o2 = pyoracle.bf_top2(dbh, qh[sel], nthreads=8)
    assert np.array_equal(d1[sel], o1) and np.array_equal(i1[sel], oi) and np.array_equal(d2[sel], o2)


def _map_point_rows(rng, sizes):
    """CSR descriptor rows: each map point's observations are noisy copies of a base descriptor (plus some outliers)."""
    rows, off = [], [0]
    for n in sizes:
        base = rng.integers(0, 256, 32, dtype=np.uint8)
        for i in range(n):
            r = base.copy()
            if rng.random() < 0.15:
                r = rng.integers(0, 256, 32, dtype=np.uint8)                   # a wrong association
            else:
                flips = rng.integers(0, 256, int(rng.integers(0, 40)))
                for b in flips:
                    r[b >> 3] ^= np.uint8(1 << (b & 7))
            rows.append(r)
        off.append(off[-1] + n)
    desc = np.stack(rows) if rows else np.zeros((0, 32), np.uint8)
    return desc, np.asarray(off, np.int64)


def _numpy_distinctive(desc, off):
    """Independent restatement: full distance matrix, np.sort, element int(0.5*(N-1)), first argmin."""
    bits = np.unpackbits(desc, axis=1).astype(np.int32)
    row, med = [], []
    for p in range(len(off) - 1):
        b = bits[off[p]:off[p + 1]]
        n = len(b)
        if n == 0:
            row.append(-1); med.append(-1); continue
        d = (b[:, None, :] != b[None, :, :]).sum(2)
        m = np.sort(d, axis=1)[:, int(0.5 * (n - 1))]
        row.append(int(np.argmin(m))); med.append(int(m.min()))
    return np.asarray(row, np.int32), np.asarray(med, np.int32)


def test_distinctive_descriptors(gpu):
    """MapPoint::ComputeDistinctiveDescriptors, batched: oracle, numpy and GPU agree bit-exactly,
    including empty points, single rows, ties (duplicate rows) and points larger than the shared-memory staging."""
    import orbb200, pyoracle
    rng = np.random.default_rng(17)
    sizes = [0, 1, 2, 3, 4, 5, 7, 8, 31, 32, 33, 63, 64, 65, 100, 0, 200] + [int(x) for x in rng.integers(1, 30, 400)]
    desc, off = _map_point_rows(rng, sizes)
    # ties: a point made of identical rows, and one with two identical clusters
    dup = np.repeat(desc[5:6], 6, 0)
    two = np.concatenate([np.repeat(desc[9:10], 3, 0), np.repeat(desc[20:21], 3, 0)])
    desc = np.concatenate([desc, dup, two]); off = np.concatenate([off, [off[-1] + 6, off[-1] + 12]])
    want_row, want_med = pyoracle.distinctive_descriptors(desc, off)
    np_row, np_med = _numpy_distinctive(desc, off)
    assert np.array_equal(want_row, np_row) and np.array_equal(want_med, np_med)
    got_row, got_med = orbb200.distinctive_descriptors(desc, off)
    assert np.array_equal(got_row, want_row)
    assert np.array_equal(got_med, want_med)
    assert got_row[0] == -1 and got_row[1] == 0 and got_row[-2] == 0
    # no points at all
    r, m = orbb200.distinctive_descriptors(np.zeros((0, 32), np.uint8), np.zeros(1, np.int64))
    assert len(r) == 0


def test_distinctive_descriptors_map_scale(gpu):
    """A map's worth of points (50 k points, ~8 observations each) against the oracle."""
    import orbb200, pyoracle
    rng = np.random.default_rng(23)
    sizes = rng.integers(2, 15, 50_000)
    off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
    base = rng.integers(0, 256, (len(sizes), 32), dtype=np.uint8)
    desc = np.repeat(base, sizes, axis=0)
    noise = (rng.random((len(desc), 32, 8)) < 0.06)
    desc ^= np.packbits(noise, axis=2).reshape(len(desc), 32)
    want_row, want_med = pyoracle.distinctive_descriptors(desc, off)
    got_row, got_med = orbb200.distinctive_descriptors(desc, off)
    assert np.array_equal(got_row, want_row) and np.array_equal(got_med, want_med)


def test_search_by_projection_reloc_and_loop(gpu):
    """the relocalisation (ORBmatcher.cpp:1661-1789) and loop-closing (:352-472) SearchByProjection overloads."""
    import orbb200, pyoracle
    (cur, kc, dc), (last, kl, dl), scale = _frames(seed=14)
    rng = np.random.default_rng(10)
    h, w = cur.shape
    F = orbb200.FrameMatcher(kc, dc, (0, 0, w, h)); O = pyoracle.OracleFrame(kc, dc, (0, 0, w, h))
    for th, orbdist, check in [(10.0, 100, True), (3.0, 64, True), (10.0, 100, False)]:
        q = _queries(kl, dl, scale, th, (3, 1), rng, "pm1")                       # GetFeaturesInArea(u, v, radius, p-1, p+1)
        mp0 = np.full(len(kc), -1, np.int32); mp0[::13] = 5555                    # features that already have a MapPoint
        n, mp = orbb200.ORBmatcher(0.9, check).SearchByProjection_reloc(F, q, mp0, orbdist)
        on, omp = O.search_reloc(_to_oracle_queries(q), orbdist, check, mp0)
        assert n == on and np.array_equal(mp, omp) and n > 0.2 * len(kl)
    for th in (4.0, 10.0):
        q = _queries(kl, dl, scale, th, (3, 1), rng, "pred")                      # level gate [p-1, p]
        m0 = np.full(len(kc), -1, np.int32); m0[3::11] = 4444
        n, m = orbb200.ORBmatcher().SearchByProjection_loop(F, q, m0)
        on, om = O.search_loop(_to_oracle_queries(q), m0)
        assert n == on and np.array_equal(m, om) and n > 0.2 * len(kl)


def test_fuse_and_sim3_best(gpu):
    """the per-point search of Fuse (chi-square gate, mono and stereo), Fuse(Scw) and SearchBySim3."""
    import orbb200, pyoracle
    (cur, kc, dc), (last, kl, dl), scale = _frames(seed=15)
    rng = np.random.default_rng(11)
    h, w = cur.shape
    F = orbb200.FrameMatcher(kc, dc, (0, 0, w, h)); O = pyoracle.OracleFrame(kc, dc, (0, 0, w, h))
    sigma2 = (scale.astype(np.float32) ** 2).astype(np.float32)
    inv_s2 = (np.float32(1.0) / sigma2).astype(np.float32)
    ur = np.where(rng.random(len(kc)) < 0.5, kc["x"] - rng.uniform(0, 30, len(kc)), -1).astype(np.float32)
    q = _queries(kl, dl, scale, 3.0, (3, 1), rng, "pred")
    q["q"]["u"] += rng.normal(0, 1.5, len(q)).astype(np.float32)                   # reprojection noise: the gate really cuts
    q["q"]["v"] += rng.normal(0, 1.5, len(q)).astype(np.float32)
    q["u_right"] = q["q"]["u"] - 12.0
    oq = _to_oracle_queries(q)
    m = orbb200.ORBmatcher()
    cases = [(ur, inv_s2, True, 50), (None, inv_s2, True, 50), (None, None, False, 50), (None, None, False, 100)]
    seen = []
    for uright, s2, gate, maxd in cases:
        bi, bd = m.FuseBest(F, q, uright, s2, maxd)
        obi, obd = O.fuse_best(oq, uright, s2 if s2 is not None else inv_s2, gate, maxd)
        assert np.array_equal(bi, obi) and np.array_equal(bd, obd)
        assert (bi >= 0).sum() > 0.2 * len(kl)
        seen.append(int((bi >= 0).sum()))
    assert seen[0] <= seen[2] <= seen[3]                                           # gates and thresholds only remove matches
    # SearchBySim3's mutual check (ORBmatcher.cpp:1478-1495) on two directions of best matches
    q21 = _queries(kc, dc, scale, 7.5, (-3, -1), rng, "pred")
    F2 = orbb200.FrameMatcher(kl, dl, (0, 0, w, h)); O2 = pyoracle.OracleFrame(kl, dl, (0, 0, w, h))
    q12 = _queries(kl, dl, scale, 7.5, (3, 1), rng, "pred")
    m1, _ = m.FuseBest(F, q12, None, None, 100)        # KF1 points searched in KF2 (= cur)
    m2, _ = m.FuseBest(F2, q21, None, None, 100)       # KF2 points searched in KF1 (= last)
    o1, _ = O.fuse_best(_to_oracle_queries(q12), None, inv_s2, False, 100)
    o2, _ = O2.fuse_best(_to_oracle_queries(q21), None, inv_s2, False, 100)
    assert np.array_equal(m1, o1) and np.array_equal(m2, o2)
    mutual = [(i1, m1[i1]) for i1 in range(len(m1)) if m1[i1] >= 0 and m2[m1[i1]] == i1]
    assert len(mutual) > 0.2 * len(kl)


def _contended(kl, dl, scale, rng, reps, th, levels):
    """Every last-frame point `reps` times in shuffled order with a wide window: many queries compete for the same Frame slots, so
    the reference's sequential already-matched rule (ORBmatcher.cpp:112-114, :1590-1592) decides most of the outcome."""
    q = _queries(kl, dl, scale, th, (3, 1), rng, levels)
    q = np.concatenate([q] * reps)
    q = q[rng.permutation(len(q))]
    q["id"] = np.arange(len(q))
    q["has_obs"] = (rng.random(len(q)) < 0.7).astype(np.int32)
    flip = rng.random(len(q)) < 0.5                      # perturb half of the descriptors so that the rivals' best slots differ
    bits = rng.integers(0, 256, (len(q), 32), dtype=np.uint8) & rng.integers(0, 256, (len(q), 32), dtype=np.uint8) & rng.integers(0, 256, (len(q), 32), dtype=np.uint8)
    q["q"]["desc"][flip] ^= bits[flip]
    return q


@pytest.mark.parametrize("mode", ["frame", "points"])
def test_search_by_projection_contended_order_dependence(gpu, mode):
    """The device resolve (parallel fixed point) against the sequential reference loop under heavy slot contention."""
    import orbb200, pyoracle
    (cur, kc, dc), (last, kl, dl), scale = _frames(seed=14)
    rng = np.random.default_rng(21)
    h, w = cur.shape
    F = orbb200.FrameMatcher(kc, dc, (0, 0, w, h)); O = pyoracle.OracleFrame(kc, dc, (0, 0, w, h))
    ur = np.where(rng.random(len(kc)) < 0.5, kc["x"] - rng.uniform(0, 30, len(kc)), -1).astype(np.float32)
    for reps, th in [(3, 15.0), (6, 40.0)]:
        q = _contended(kl, dl, scale, rng, reps, th, "pm1" if mode == "frame" else "pred")
        mp0 = np.full(len(kc), -1, np.int32); ob0 = np.zeros(len(kc), np.int32)
        mp0[::17] = 9999; ob0[::17] = 1; mp0[5::23] = 7777
        for uright in (None, ur):
            if mode == "frame":
                n, mp, ob = orbb200.ORBmatcher(0.9, True).SearchByProjection_frame(F, q, mp0, ob0, uright)
                on, omp, oob = O.search_last_frame(_to_oracle_queries(q), True, mp0, ob0, uright)
            else:
                n, mp, ob = orbb200.ORBmatcher(0.8).SearchByProjection_points(F, q, mp0, ob0, uright)
                on, omp, oob = O.search_local_points(_to_oracle_queries(q), 0.8, mp0, ob0, uright)
            assert n == on and np.array_equal(mp, omp)
            assert np.array_equal(ob, oob)
            assert n > 200
    # candidate overflow of the default budget is retried with the exact size (reps=6, th=40 above exceeds 64*nq+1024 or not; force it)
    q = _contended(kl, dl, scale, rng, 2, 200.0, "pm1")
    q["q"]["min_level"] = -1; q["q"]["max_level"] = -1
    n, mp, ob = orbb200.ORBmatcher(0.9, True).SearchByProjection_frame(F, q, mp0, ob0, None)
    on, omp, oob = O.search_last_frame(_to_oracle_queries(q), True, mp0, ob0, None)
    assert n == on and np.array_equal(mp, omp)


def test_search_by_projection_device_resident(gpu):
    """Frame built from device arrays, queries / state / results on the device, no host round trip in between."""
    import torch
    import orbb200, pyoracle
    (cur, kc, dc), (last, kl, dl), scale = _frames(seed=15)
    rng = np.random.default_rng(22)
    h, w = cur.shape
    dev = torch.device("cuda:0")
    t_k = torch.from_numpy(kc.view(np.uint8).reshape(len(kc), -1).copy()).to(dev); t_d = torch.from_numpy(dc.copy()).to(dev)
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        F = orbb200.FrameMatcher.from_device(t_k.data_ptr(), t_d.data_ptr(), len(kc), (0, 0, w, h), 0, st.cuda_stream)
        O = pyoracle.OracleFrame(kc, dc, (0, 0, w, h))
        for mode, levels in (("frame", "pm1"), ("points", "pred")):
            q = _contended(kl, dl, scale, rng, 2, 15.0, levels)
            mp0 = np.full(len(kc), -1, np.int32); ob0 = np.zeros(len(kc), np.int32)
            mp0[::17] = 9999; ob0[::17] = 1
            t_q = torch.from_numpy(q.view(np.uint8).reshape(len(q), -1).copy()).to(dev, non_blocking=True)
            t_mp = torch.from_numpy(mp0.copy()).to(dev); t_ob = torch.from_numpy(ob0.copy()).to(dev)
            t_sc = torch.zeros(2, dtype=torch.int32, device=dev)
            orbb200.ORBmatcher(0.8, True).SearchByProjection_device(mode, F, t_q.data_ptr(), len(q), t_mp.data_ptr(), t_ob.data_ptr(), t_sc.data_ptr(),
                                                                    t_sc.data_ptr() + 4, stream=st.cuda_stream)
            st.synchronize()
            n, status = t_sc.cpu().tolist()
            if mode == "frame":
                on, omp, oob = O.search_last_frame(_to_oracle_queries(q), True, mp0, ob0, None)
            else:
                on, omp, oob = O.search_local_points(_to_oracle_queries(q), 0.8, mp0, ob0, None)
            assert status == 0 and n == on and np.array_equal(t_mp.cpu().numpy(), omp)
        # a budget that is too small reports ORBB_E_CAPACITY in *status_dev and leaves the state untouched
        t_mp = torch.from_numpy(mp0.copy()).to(dev)
        orbb200.ORBmatcher(0.8, True).SearchByProjection_device("frame", F, t_q.data_ptr(), len(q), t_mp.data_ptr(), t_ob.data_ptr(), t_sc.data_ptr(),
                                                                t_sc.data_ptr() + 4, cand_capacity=8, stream=st.cuda_stream)
        st.synchronize()
        assert t_sc.cpu().tolist()[1] == orbb200.E_CAPACITY and np.array_equal(t_mp.cpu().numpy(), mp0)
        # the host-resolve entry points still work on a device-built frame (host copies are fetched on first use)
        qq = _queries(kl, dl, scale, 10.0, (3, 1), rng, "pm1")
        n1, m1 = orbb200.ORBmatcher(0.9, True).SearchByProjection_reloc(F, qq, np.full(len(kc), -1, np.int32), 100)
        n2, m2 = O.search_reloc(_to_oracle_queries(qq), 100, True, np.full(len(kc), -1, np.int32))
        assert n1 == n2 and np.array_equal(m1, m2)
    F.close()


def test_undistort_keypoints_and_image_bounds(gpu):
    """Frame::UndistortKeyPoints / ComputeImageBounds (Frame.cpp:462-536) on the device: bit-equal to cv2 4.13 (golden) and the oracle."""
    import os, sys
    import orbb200, pyoracle
    gold = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    sys.path.insert(0, gold)
    import make_golden as GEN
    g = np.load(os.path.join(gold, "cv2_undistort.npz"))
    rng = np.random.default_rng(3)
    for i, (name, K4, D, w, h) in enumerate(GEN.undistort_cases()):
        pts = GEN.undistort_points_in(w, h, 300 + i)
        kps = np.zeros(len(pts), orbb200.KP_DTYPE)
        kps["x"] = pts[:, 0]; kps["y"] = pts[:, 1]
        kps["angle"] = rng.uniform(0, 360, len(pts)); kps["octave"] = rng.integers(0, 8, len(pts)); kps["response"] = rng.uniform(0, 200, len(pts))
        un = orbb200.undistort_keypoints(kps, K4, D)
        got = np.stack([un["x"], un["y"]], 1)
        assert np.array_equal(got.view(np.uint32), g[name].view(np.uint32)), name
        for f in ("size", "angle", "response", "octave", "class_id"):
            assert np.array_equal(un[f], kps[f])
        assert np.array_equal(un.view(np.uint8), pyoracle.undistort_keypoints(kps, K4, D).view(np.uint8))
        assert np.array_equal(orbb200.image_bounds(w, h, K4, D), pyoracle.image_bounds(w, h, K4, D))
    # mDistCoef(0) == 0: copy, bounds = the image rectangle (even when later coefficients are set, as the reference does)
    K4 = [458.654, 457.296, 367.215, 248.375]
    un = orbb200.undistort_keypoints(kps, K4, [0, 0.2, 0, 0])
    assert np.array_equal(un.view(np.uint8), kps.view(np.uint8))
    assert orbb200.image_bounds(752, 480, K4, [0, 0.2, 0, 0]).tolist() == [0, 0, 752, 480]
    assert len(orbb200.undistort_keypoints(kps[:0], K4, [0.1, 0, 0, 0])) == 0
    with pytest.raises(orbb200.OrbError):
        orbb200.undistort_keypoints(kps, K4, [0.1, 0, 0])                          # 3 coefficients: not a calibration OpenCV accepts


def test_tracked_frame_stays_on_device(gpu):
    """SURVEY 8(f)-1 end to end: extract (device) -> undistort (device, in place) -> grid (device) -> SearchByProjection candidates and
    resolve (device).  Only the final match state is read back; every stage equals the oracle chain."""
    import torch
    import orbb200, pyoracle
    from synth import synth
    w, h, nf = 752, 480, 1200
    K4 = [458.654, 457.296, 367.215, 248.375]; D = [-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05]
    big = synth(w + 16, h + 16, 31)
    cur = big[8:8 + h, 8:8 + w].copy(); last = big[9:9 + h, 11:11 + w].copy()
    oex = pyoracle.OracleExtractor(nf)
    kc, dc = oex.extract(cur); kl, dl = oex.extract(last)
    scale = oex.tables()["scale"]
    kcu = pyoracle.undistort_keypoints(kc, K4, D); klu = pyoracle.undistort_keypoints(kl, K4, D)
    bounds = pyoracle.image_bounds(w, h, K4, D)
    rng = np.random.default_rng(5)
    q = _queries(klu, dl, scale, 15.0, (3, 1), rng, "pm1")
    mp0 = np.full(len(kc), -1, np.int32); ob0 = np.zeros(len(kc), np.int32)
    on, omp, oob = pyoracle.OracleFrame(kcu, dc, bounds).search_last_frame(_to_oracle_queries(q), True, mp0, ob0, None)

    dev = torch.device("cuda:0")
    ex = orbb200.ORBextractor(nf, max_size=(w, h))
    cap = 2 * nf
    t_img = torch.from_numpy(cur).to(dev)
    t_k = torch.zeros(cap * 28, dtype=torch.uint8, device=dev); t_d = torch.zeros(cap * 32, dtype=torch.uint8, device=dev)
    t_n = torch.zeros(1, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    ex.extract_batch_device(t_img.data_ptr(), 1, w, h, w, w * h, t_k.data_ptr(), t_d.data_ptr(), t_n.data_ptr(), cap)
    st = ex.stream()
    n = len(kc)                              # the one scalar a real caller would read (or size the grid for `capacity`)
    orbb200.undistort_keypoints_device(t_k.data_ptr(), n, K4, D, t_k.data_ptr(), 0, st)
    F = orbb200.FrameMatcher.from_device(t_k.data_ptr(), t_d.data_ptr(), n, orbb200.image_bounds(w, h, K4, D), 0, st)
    ex.sync()
    assert int(t_n.cpu()[0]) == n
    t_q = torch.from_numpy(q.view(np.uint8).reshape(len(q), -1).copy()).to(dev)
    t_mp = torch.from_numpy(mp0.copy()).to(dev); t_ob = torch.from_numpy(ob0.copy()).to(dev); t_sc = torch.zeros(2, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    orbb200.ORBmatcher(0.9, True).SearchByProjection_device("frame", F, t_q.data_ptr(), len(q), t_mp.data_ptr(), t_ob.data_ptr(), t_sc.data_ptr(),
                                                            t_sc.data_ptr() + 4, stream=st)
    ex.sync()
    assert t_sc.cpu().tolist() == [on, 0] and np.array_equal(t_mp.cpu().numpy(), omp) and on > 300
    F.close()


def test_host_form_search_waits_for_device_built_frame(gpu):
    """orbm_frame_create_device queues its copies and grid kernels on the caller's non-blocking stream; the
    host-pointer entry points run on the legacy default stream, which does not order against it.  They must wait on the frame's
    'ready' event themselves: no explicit synchronisation here, and a long kernel is queued first so the race would be lost."""
    import torch
    import orbb200, pyoracle
    (cur, kc, dc), (last, kl, dl), scale = _frames(seed=16)
    rng = np.random.default_rng(23)
    h, w = cur.shape
    dev = torch.device("cuda:0")
    t_k = torch.from_numpy(kc.view(np.uint8).reshape(len(kc), -1).copy()).to(dev); t_d = torch.from_numpy(dc.copy()).to(dev)
    big = torch.empty(1 << 28, dtype=torch.float32, device=dev)
    torch.cuda.synchronize()
    st = torch.cuda.Stream()                       # cudaStreamNonBlocking
    O = pyoracle.OracleFrame(kc, dc, (0, 0, w, h))
    q = _queries(kl, dl, scale, 10.0, (3, 1), rng, "pm1")
    with torch.cuda.stream(st):
        for _ in range(20):
            big.mul_(1.0001)                       # ~ tens of ms of work ahead of the frame build on `st`
        F = orbb200.FrameMatcher.from_device(t_k.data_ptr(), t_d.data_ptr(), len(kc), (0, 0, w, h), 0, st.cuda_stream)
    offs, idx, dist = F.window_search(q["q"])                                       # host form, default stream, no sync in between
    oo, oi, od = O.window_search(_to_oracle_queries(q))
    assert np.array_equal(offs, oo) and np.array_equal(idx, oi) and np.array_equal(dist, od)
    n1, m1 = orbb200.ORBmatcher(0.9, True).SearchByProjection_reloc(F, q, np.full(len(kc), -1, np.int32), 100)
    n2, m2 = O.search_reloc(_to_oracle_queries(q), 100, True, np.full(len(kc), -1, np.int32))
    assert n1 == n2 and np.array_equal(m1, m2)
    F.close()


def test_reloc_max_dist_above_255_stays_in_bounds(gpu):
    """with max_dist >= 256 a query without any unfiltered candidate must not write frame_mp[-1]."""
    import orbb200
    (cur, kc, dc), (last, kl, dl), scale = _frames(seed=16)
    h, w = cur.shape
    F = orbb200.FrameMatcher(kc, dc, (0, 0, w, h))
    q = _queries(kl, dl, scale, 10.0, (3, 1), np.random.default_rng(1), "pm1")
    mp0 = np.zeros(len(kc), np.int32)                                              # every feature already has a MapPoint
    n, mp = orbb200.ORBmatcher(0.9, True).SearchByProjection_reloc(F, q, mp0, 300)
    assert n == 0 and np.array_equal(mp, mp0)


def test_tensor_peak_probes(gpu):
    """The roofline denominators are measured, not assumed: the tcgen05 kind::i8 probe must land between the legacy mma.sync
    rate and a loose ceiling above the nominal 4.5 POP/s."""
    import orbb200
    legacy = orbb200.imma_peak(0)
    tc = orbb200.tcgen05_i8_peak(0)
    assert 500 < legacy < 2500
    assert 2 * legacy < tc < 6000, (legacy, tc)


def test_reloc_and_loop_resolve_on_device(gpu):
    """orbm_search_by_projection_reloc_device: queries and the slot array stay on the device, window search and the order-dependent resolve
    (every accepted match blocks its slot for the later queries) run there without host synchronisation.  Wide windows make many queries
    compete for the same features; the result must equal the sequential oracle loop, for the relocalisation parameters (ORBdist,
    orientation check) and the loop-closing ones (TH_LOW, no orientation check)."""
    import ctypes as C
    import torch, orbb200, pyoracle
    (cur, kc, dc), (last, kl, dl), scale = _frames(seed=16)
    rng = np.random.default_rng(12)
    h, w = cur.shape
    F = orbb200.FrameMatcher(kc, dc, (0, 0, w, h)); O = pyoracle.OracleFrame(kc, dc, (0, 0, w, h))
    L = orbb200.lib()
    st = torch.cuda.Stream()
    for th, maxd, check, levels in [(40.0, 100, True, "pm1"), (25.0, 64, False, "pm1"), (40.0, 50, False, "pred")]:
        q = _queries(kl, dl, scale, th, (3, 1), rng, levels)
        mp0 = np.full(len(kc), -1, np.int32); mp0[::13] = 5555
        on, omp = (O.search_reloc(_to_oracle_queries(q), maxd, check, mp0) if levels == "pm1" else O.search_loop(_to_oracle_queries(q), mp0))
        with torch.cuda.stream(st):
            qd = torch.from_numpy(q.view(np.uint8).reshape(-1).copy()).cuda()
            mpd = torch.from_numpy(mp0.copy()).cuda()
            sc = torch.zeros(2, dtype=torch.int32, device="cuda")
            rc = L.orbm_search_by_projection_reloc_device(F.h, C.c_void_p(qd.data_ptr()), len(q), maxd, int(check), C.c_void_p(mpd.data_ptr()),
                                                          C.c_void_p(sc.data_ptr()), C.c_void_p(sc.data_ptr() + 4), 0, C.c_void_p(st.cuda_stream))
            assert rc == 0
        st.synchronize()
        n, status = sc.cpu().tolist()
        assert status == 0 and n == on and np.array_equal(mpd.cpu().numpy(), omp), (th, maxd, check, levels)
        assert n > 0.2 * len(kl)


def test_matcher_entry_points_from_concurrent_host_threads(gpu):
    """The reference calls the matcher from several threads (Tracking, LocalMapping, LoopClosing).  The host-pointer entry points run on
    per-thread default streams and take their scratch from the library's shared block cache: four threads hammer frame creation, both
    SearchByProjection forms, the relocalisation form, Fuse and descriptor distances at once; every result must equal the oracle's."""
    import threading
    import orbb200, pyoracle
    (cur, kc, dc), (last, kl, dl), scale = _frames(seed=17)
    h, w = cur.shape
    O = pyoracle.OracleFrame(kc, dc, (0, 0, w, h))
    rng = np.random.default_rng(13)
    q = _queries(kl, dl, scale, 15.0, (3, 1), rng)
    oq = _to_oracle_queries(q)
    mp0 = np.full(len(kc), -1, np.int32); ob0 = np.zeros(len(kc), np.int32); mp0[::17] = 9999; ob0[::17] = 1
    exp_frame = O.search_last_frame(oq, True, mp0, ob0, None)
    exp_pts = O.search_local_points(oq, 0.8, mp0, ob0, None)
    exp_reloc = O.search_reloc(oq, 100, True, mp0)
    exp_dist = pyoracle.distance_pairs(dc[:500], dl[:500])
    errs = []

    def work(t):
        try:
            for it in range(12):
                F = orbb200.FrameMatcher(kc, dc, (0, 0, w, h))
                n, mp, ob = orbb200.ORBmatcher(0.9, True).SearchByProjection_frame(F, q, mp0, ob0, None)
                assert n == exp_frame[0] and np.array_equal(mp, exp_frame[1]) and np.array_equal(ob, exp_frame[2]), ("frame", t, it)
                n, mp, ob = orbb200.ORBmatcher(0.8).SearchByProjection_points(F, q, mp0, ob0, None)
                assert n == exp_pts[0] and np.array_equal(mp, exp_pts[1]), ("points", t, it)
                n, mp = orbb200.ORBmatcher(0.9, True).SearchByProjection_reloc(F, q, mp0, 100)
                assert n == exp_reloc[0] and np.array_equal(mp, exp_reloc[1]), ("reloc", t, it)
                assert np.array_equal(orbb200.DescriptorDistance(dc[:500], dl[:500]), exp_dist), ("dist", t, it)
                F.close()
        except Exception as e:          # surfaced after join
            errs.append(e)
    th = [threading.Thread(target=work, args=(t,)) for t in range(4)]
    [t.start() for t in th]; [t.join() for t in th]
    assert not errs, errs[0]


@pytest.mark.parametrize("variant", ["3", "4"])
def test_bf_top2_ties_across_chunks(gpu, variant, monkeypatch):
    """Chunks of one scan share their second-best bounds: only rows STRICTLY farther than a published bound may be dropped.  A database of
    long runs of identical rows makes every chunk hold exact ties of the global best and second best: the first index must win, d2 == d1,
    whatever chunk finishes first."""
    import orbb200, pyoracle
    monkeypatch.setenv("ORBB_BF_VARIANT", variant)
    rng = np.random.default_rng(31)
    base = rng.integers(0, 256, (7, 32), dtype=np.uint8)
    db = np.repeat(base, 60_000, axis=0)                                  # 420 k rows: 7 runs of identical rows
    db[::4999] = rng.integers(0, 256, (len(db[::4999]), 32), dtype=np.uint8)
    q = rng.integers(0, 256, (700, 32), dtype=np.uint8)
    q[:7] = base; q[7:14] = base ^ np.uint8(1)
    D = orbb200.DescriptorDB(len(db)); D.append(db)
    d1, i1, d2 = D.top2(q)
    o1, oi, o2 = pyoracle.bf_top2(db, q, nthreads=8)
    assert np.array_equal(d1, o1) and np.array_equal(i1, oi) and np.array_equal(d2, o2)
    assert np.all(d1[:7] == 0) and np.all(d2[:7] == 0)
    D.close()


@pytest.mark.parametrize("w,h,nf,seed", [(1241, 376, 2000, 21), (752, 480, 1200, 22), (638, 401, 1000, 23)])
def test_stereo_frame_create(gpu, w, h, nf, seed):
    """orbm_stereo_frame_create (Frame.cpp:61-115 as one call) == the oracle's extraction of both images + its ComputeStereoMatches, and
    == the library's own four calls one after the other; the frame it leaves on the device answers a tracked-frame search like a frame
    built from the host arrays, with mvuRight taken from the device (ORBM_URIGHT_OF_FRAME) or uploaded."""
    import orbb200, pyoracle
    from synth import synth_stereo
    L, R = synth_stereo(w, h, seed)
    oL, oR = pyoracle.OracleExtractor(nf), pyoracle.OracleExtractor(nf)
    okL, odL = oL.extract(L); okR, odR = oR.extract(R)
    mbf = 386.1448
    ex = orbb200.ORBextractor(nf, max_size=(w, h), max_batch=2)
    scale = ex.GetScaleFactors()
    rng = np.random.default_rng(seed)
    for rep in range(2):                                      # twice in a row through the same handle and pinned blocks
        for mb in (0.0, 0.537):
            F, kL, dL, kR, dR, ur, dp = orbb200.stereo_frame(ex, L, R, mbf, (0, 0, w, h), mb)
            assert kL.tobytes() == okL.tobytes() and np.array_equal(dL, odL) and kR.tobytes() == okR.tobytes() and np.array_equal(dR, odR)
            with np.errstate(divide="ignore"):
                our, odp, _, _ = pyoracle.stereo_match(oL, oR, mbf, mb)
            assert np.array_equal(ur, our) and np.array_equal(dp, odp)
            # tracked-frame search on the frame the constructor left on the device
            q = _queries(kL, dL, scale, 15.0, (2, 1), rng)
            mp0 = np.full(len(kL), -1, np.int32); ob0 = np.zeros(len(kL), np.int32)
            m = orbb200.ORBmatcher(0.9, True)
            G = orbb200.FrameMatcher(kL, dL, (0, 0, w, h)); O = pyoracle.OracleFrame(kL, dL, (0, 0, w, h))
            on, omp, oob = O.search_last_frame(_to_oracle_queries(q), True, mp0, ob0, ur)
            for frame, uright in ((F, orbb200.URIGHT_OF_FRAME), (F, ur), (G, ur)):
                n, mp, ob = m.SearchByProjection_frame(frame, q, mp0, ob0, uright)
                assert n == on and np.array_equal(mp, omp) and np.array_equal(ob, oob)
            assert on > 0.3 * len(kL)
            with pytest.raises(orbb200.OrbError):             # a frame built from host arrays has no mvuRight of its own
                m.SearchByProjection_frame(G, q, mp0, ob0, orbb200.URIGHT_OF_FRAME)
            F.close(); G.close()
    # a handle that cannot hold two frames, and a capacity below the keypoint count
    with pytest.raises(orbb200.OrbError):
        orbb200.stereo_frame(orbb200.ORBextractor(nf, max_size=(w, h)), L, R, mbf, (0, 0, w, h))
    with pytest.raises(orbb200.OrbError):
        orbb200.stereo_frame(ex, L, R, mbf, (0, 0, w, h), capacity=len(okL) - 1)


def test_stereo_frame_create_degenerate_inputs(gpu):
    """The fused stereo constructor on images without a single corner (left, right, both), on a non-contiguous (strided) pair and right after
    a different shape went through the same handle: counts, empty outputs and the searches on an empty frame behave like the separate calls."""
    import orbb200, pyoracle
    from synth import synth_stereo
    w, h, nf = 640, 480, 1000
    L, R = synth_stereo(w, h, 5)
    flat = np.full((h, w), 128, np.uint8)
    ex = orbb200.ORBextractor(nf, max_size=(800, 600), max_batch=2)
    o = pyoracle.OracleExtractor(nf)
    okL, odL = o.extract(L)
    q = _queries(okL, odL, ex.GetScaleFactors(), 15.0, (2, 1), np.random.default_rng(1))
    m = orbb200.ORBmatcher(0.9, True)
    for left, right, nl_expected in ((flat, R, 0), (L, flat, len(okL)), (flat, flat, 0)):
        F, kL, dL, kR, dR, ur, dp = orbb200.stereo_frame(ex, left, right, 40.0, (0, 0, w, h))
        assert len(kL) == nl_expected and len(ur) == nl_expected and (len(kR) == 0) == (right is flat)
        if right is flat:
            assert np.all(ur == -1) and np.all(dp == -1)              # no right keypoints: no stereo match
        n, mp, ob = m.SearchByProjection_frame(F, q, np.full(len(kL), -1, np.int32), np.zeros(len(kL), np.int32),
                                               orbb200.URIGHT_OF_FRAME if nl_expected else None)
        assert (n == 0) == (nl_expected == 0)
        F.close()
    # another shape through the same handle, then the first one again: the geometry tables follow
    L2, R2 = synth_stereo(752, 480, 6)
    F2, k2, d2, _, _, ur2, _ = orbb200.stereo_frame(ex, L2, R2, 47.9, (0, 0, 752, 480))
    ok2, od2 = pyoracle.OracleExtractor(nf).extract(L2)
    assert k2.tobytes() == ok2.tobytes() and np.array_equal(d2, od2)
    F2.close()
    F, kL, dL, kR, dR, ur, dp = orbb200.stereo_frame(ex, L, R, 40.0, (0, 0, w, h))
    assert kL.tobytes() == okL.tobytes() and np.array_equal(dL, odL)
    oL, oR = pyoracle.OracleExtractor(nf), pyoracle.OracleExtractor(nf)
    oL.extract(L); oR.extract(R)
    with np.errstate(divide="ignore"):
        our, odp, _, _ = pyoracle.stereo_match(oL, oR, 40.0, 0.0)
    assert np.array_equal(ur, our) and np.array_equal(dp, odp)
    F.close()


@pytest.mark.parametrize("w,h,nf,seed,dist", [(640, 480, 1000, 31, [0.262383, -0.953104, -0.005358, 0.002628, 1.163314]),   # TUM1.yaml
                                             (752, 480, 1200, 32, None), (638, 401, 1000, 33, [0.0, 0.2, 0.0, 0.0])])
def test_mono_frame_create(gpu, w, h, nf, seed, dist):
    """orbm_mono_frame_create (Frame.cpp:171-224 as one call) == oracle extraction + oracle UndistortKeyPoints, and the frame it leaves on
    the device answers a local-map search like a frame built from the host arrays."""
    import orbb200, pyoracle
    from synth import synth
    img = synth(w, h, seed)
    K4 = [517.306408, 516.469215, 318.643040, 255.313989]
    ok, od = pyoracle.OracleExtractor(nf).extract(img)
    oun = pyoracle.undistort_keypoints(ok, K4, dist) if dist is not None else ok
    bounds = orbb200.image_bounds(w, h, K4, dist) if dist is not None else (0, 0, w, h)
    ex = orbb200.ORBextractor(nf, max_size=(w, h))
    rng = np.random.default_rng(seed)
    for rep in range(2):
        F, k, ku, d = orbb200.mono_frame(ex, img, bounds, K4 if dist is not None else None, dist)
        assert k.tobytes() == ok.tobytes() and np.array_equal(d, od) and ku.tobytes() == oun.tobytes()
        G = orbb200.FrameMatcher(ku, d, bounds); O = pyoracle.OracleFrame(ku, d, bounds)
        q = _queries(ku, d, ex.GetScaleFactors(), 10.0, (1, 1), rng)
        mp0 = np.full(len(k), -1, np.int32); ob0 = np.zeros(len(k), np.int32)
        m = orbb200.ORBmatcher(0.8, True)
        on, omp, oob = O.search_local_points(_to_oracle_queries(q), 0.8, mp0, ob0, None)
        for frame in (F, G):
            n, mp, ob = m.SearchByProjection_points(frame, q, mp0, ob0, None)
            assert n == on and np.array_equal(mp, omp) and np.array_equal(ob, oob)
        assert on > 0.3 * len(k)
        F.close(); G.close()
    with pytest.raises(orbb200.OrbError):
        orbb200.mono_frame(ex, img, bounds, capacity=len(ok) - 1)


@pytest.mark.parametrize("n", [10240, 10241, 12288, 20000])
def test_resolve_large_frames_around_the_shared_memory_limit(gpu, n):
    """The resolve kernel keeps its slot table in shared memory up to 10 240 Frame features and in global memory beyond: synthetic frames of
    n keypoints on both sides of that limit (and at the old 12 288 one), both resolve modes, against the oracle."""
    import orbb200, pyoracle
    rng = np.random.default_rng(n)
    w, h = 3840, 2160
    kps = np.zeros(n, orbb200.KP_DTYPE)
    kps["x"] = rng.integers(20, w - 20, n); kps["y"] = rng.integers(20, h - 20, n); kps["octave"] = rng.integers(0, 8, n)
    kps["angle"] = rng.uniform(0, 360, n).astype(np.float32); kps["size"] = 31; kps["class_id"] = -1
    desc = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    scale = (1.2 ** np.arange(8)).astype(np.float32)
    pick = rng.permutation(n)[:3000]
    q = _queries(kps[pick], desc[pick], scale, 25.0, (2, 1), rng)
    noise = rng.integers(0, 256, (len(q), 32), dtype=np.uint8) & rng.integers(0, 256, (len(q), 32), dtype=np.uint8) & rng.integers(0, 256, (len(q), 32), dtype=np.uint8)
    q["q"]["desc"] ^= noise                                   # ~32 flipped bits: matches survive TH_HIGH, rivals differ
    F = orbb200.FrameMatcher(kps, desc, (0, 0, w, h)); O = pyoracle.OracleFrame(kps, desc, (0, 0, w, h))
    mp0 = np.full(n, -1, np.int32); ob0 = np.zeros(n, np.int32)
    nf, mp, ob = orbb200.ORBmatcher(0.9, True).SearchByProjection_frame(F, q, mp0, ob0, None)
    on, omp, oob = O.search_last_frame(_to_oracle_queries(q), True, mp0, ob0, None)
    assert nf == on and np.array_equal(mp, omp) and np.array_equal(ob, oob) and nf > 500
    nf, mp, ob = orbb200.ORBmatcher(0.8).SearchByProjection_points(F, q, mp0, ob0, None)
    on, omp, oob = O.search_local_points(_to_oracle_queries(q), 0.8, mp0, ob0, None)
    assert nf == on and np.array_equal(mp, omp) and np.array_equal(ob, oob)
    F.close()


def test_frame_constructors_from_concurrent_host_threads(gpu):
    """Four tracking threads, each with its own extractor handle, run the fused stereo constructor, the fused monocular constructor and a
    tracked-frame search at once (per-thread default streams, per-thread pinned staging, the shared scratch cache, latency-mode launches
    with per-level octree streams and dependent launches): every thread's results equal the serial ones, every iteration."""
    import threading
    import orbb200
    from synth import synth_stereo
    shapes = [(640, 480, 1000), (752, 480, 1200), (638, 401, 800), (1241, 376, 2000)]
    inputs, expected = [], []
    for t, (w, h, nf) in enumerate(shapes):
        L, R = synth_stereo(w, h, 40 + t)
        ex = orbb200.ORBextractor(nf, max_size=(w, h), max_batch=2)
        F, kL, dL, kR, dR, ur, dp = orbb200.stereo_frame(ex, L, R, 40.0, (0, 0, w, h))
        q = _queries(kL, dL, ex.GetScaleFactors(), 15.0, (2, 1), np.random.default_rng(t))
        mp0 = np.full(len(kL), -1, np.int32); ob0 = np.zeros(len(kL), np.int32)
        res = orbb200.ORBmatcher(0.9, True).SearchByProjection_frame(F, q, mp0, ob0, orbb200.URIGHT_OF_FRAME)
        F.close()
        G, k1, ku1, d1 = orbb200.mono_frame(ex, L, (0, 0, w, h)); G.close()
        assert res[0] > 0.3 * len(kL)
        inputs.append((w, h, nf, L, R, q, mp0, ob0)); expected.append((kL, dL, kR, dR, ur, dp, res, k1, d1))
        ex.close()
    errs = []

    def work(t):
        try:
            w, h, nf, L, R, q, mp0, ob0 = inputs[t]; kL0, dL0, kR0, dR0, ur0, dp0, res0, k10, d10 = expected[t]
            ex = orbb200.ORBextractor(nf, max_size=(w, h), max_batch=2)
            for it in range(15):
                F, kL, dL, kR, dR, ur, dp = orbb200.stereo_frame(ex, L, R, 40.0, (0, 0, w, h))
                assert kL.tobytes() == kL0.tobytes() and np.array_equal(dL, dL0) and kR.tobytes() == kR0.tobytes() and np.array_equal(dR, dR0), ("extract", t, it)
                assert np.array_equal(ur, ur0) and np.array_equal(dp, dp0), ("stereo", t, it)
                res = orbb200.ORBmatcher(0.9, True).SearchByProjection_frame(F, q, mp0, ob0, orbb200.URIGHT_OF_FRAME)
                assert res[0] == res0[0] and np.array_equal(res[1], res0[1]) and np.array_equal(res[2], res0[2]), ("search", t, it)
                F.close()
                G, k1, ku1, d1 = orbb200.mono_frame(ex, L, (0, 0, w, h))
                assert k1.tobytes() == k10.tobytes() and np.array_equal(d1, d10) and ku1.tobytes() == k10.tobytes(), ("mono", t, it)
                G.close()
            ex.close()
        except Exception as e:          # surfaced after join
            errs.append(e)
    th = [threading.Thread(target=work, args=(t,)) for t in range(4)]
    [t.start() for t in th]; [t.join() for t in th]
    assert not errs, errs[0]
