"""Parity against the REFERENCE's own matching code.  tests/golden/ref_match.npz holds inputs and outputs of ORBmatcher.cpp /
Frame.cpp / KeyFrame.cpp / MapPoint.cpp / DBoW2 compiled unmodified (oracle/_ref/liborb_refmatch.so; generator:
tests/golden/make_golden_match.py).  Every case runs twice: the CPU oracle against the fixture (-m "not gpu": pins the oracle) and the
CUDA path through the C ABI against the fixture (-m gpu).  All comparisons are bit-exact."""
import os
import sys
import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden", "ref_match.npz")


@pytest.fixture(scope="module")
def G():
    return np.load(GOLD)


class OracleBackend:
    name = "oracle"

    def __init__(self):
        import pyoracle
        self.O = pyoracle

    def _q(self, q):
        o = np.zeros(len(q), self.O.QUERY_DTYPE)
        o["u"] = q["q"]["u"]; o["v"] = q["q"]["v"]; o["r"] = q["q"]["r"]
        o["minLevel"] = q["q"]["min_level"]; o["maxLevel"] = q["q"]["max_level"]; o["desc"] = q["q"]["desc"]
        o["uR"] = q["u_right"]; o["maxErrR"] = q["max_err_right"]; o["id"] = q["id"]; o["hasObs"] = q["has_obs"]; o["angle"] = q["angle"]
        return o

    def frame(self, kps_un, desc, bounds):
        return self.O.OracleFrame(kps_un, desc, bounds)

    def features_in_area(self, F, x, y, r, a, b):
        return F.features_in_area(x, y, r, a, b)

    def local_points(self, F, q, nnratio, mp0, ob0, ur):
        return F.search_local_points(self._q(q), nnratio, mp0, ob0, ur)

    def last_frame(self, F, q, check, mp0, ob0, ur):
        return F.search_last_frame(self._q(q), check, mp0, ob0, ur)

    def reloc(self, F, q, orbdist, check, mp0):
        return F.search_reloc(self._q(q), orbdist, check, mp0)

    def loop(self, F, q, m0):
        return F.search_loop(self._q(q), m0)

    def fuse_best(self, F, q, ur, inv_s2, gate, max_dist):
        return F.fuse_best(self._q(q), ur, inv_s2, gate, max_dist)[0]

    def initialization(self, F1, F2, prev, win, nnratio, check):
        return self.O.search_for_initialization(F1, F2, prev, win, nnratio, check)

    def undistort(self, kps, K4, dist):
        return self.O.undistort_keypoints(kps, K4, dist)

    def image_bounds(self, w, h, K4, dist):
        return self.O.image_bounds(w, h, K4, dist)

    def stereo(self, L, R, nf, mbf):
        oL, oR = self.O.OracleExtractor(nf), self.O.OracleExtractor(nf)
        kl, dl = oL.extract(L); kr, dr = oR.extract(R)
        with np.errstate(divide="ignore"):
            ur, dp, _, _ = self.O.stereo_match(oL, oR, mbf, 0.0)
        return kl, dl, kr, dr, ur, dp


class CudaBackend:
    name = "cuda"

    def __init__(self):
        import orbb200
        self.B = orbb200

    def frame(self, kps_un, desc, bounds):
        return self.B.FrameMatcher(kps_un, desc, bounds)

    def features_in_area(self, F, x, y, r, a, b):
        q = np.zeros(1, self.B.QUERY_DTYPE)
        q["u"] = x; q["v"] = y; q["r"] = r; q["min_level"] = a; q["max_level"] = b
        off, idx, _ = F.window_search(q)
        return idx[:off[1]]

    def local_points(self, F, q, nnratio, mp0, ob0, ur):
        return self.B.ORBmatcher(nnratio).SearchByProjection_points(F, q, mp0, ob0, ur)

    def last_frame(self, F, q, check, mp0, ob0, ur):
        return self.B.ORBmatcher(0.9, check).SearchByProjection_frame(F, q, mp0, ob0, ur)

    def reloc(self, F, q, orbdist, check, mp0):
        return self.B.ORBmatcher(0.9, check).SearchByProjection_reloc(F, q, mp0, orbdist)

    def loop(self, F, q, m0):
        return self.B.ORBmatcher().SearchByProjection_loop(F, q, m0)

    def fuse_best(self, F, q, ur, inv_s2, gate, max_dist):
        return self.B.ORBmatcher().FuseBest(F, q, ur, inv_s2 if gate else None, max_dist)[0]

    def initialization(self, F1, F2, prev, win, nnratio, check):
        return self.B.ORBmatcher(nnratio, check).SearchForInitialization(F1, F2, prev, win)

    def undistort(self, kps, K4, dist):
        return self.B.undistort_keypoints(kps, K4, dist)

    def image_bounds(self, w, h, K4, dist):
        return self.B.image_bounds(w, h, K4, dist)

    def stereo(self, L, R, nf, mbf):
        h, w = L.shape
        gL, gR = self.B.ORBextractor(nf, max_size=(w, h)), self.B.ORBextractor(nf, max_size=(w, h))
        kl, dl = gL(L); kr, dr = gR(R)
        ur, dp = self.B.stereo_match(gL, gR, kl, dl, kr, dr, mbf, 0.0)
        return kl, dl, kr, dr, ur, dp


@pytest.fixture(scope="module", params=["oracle", pytest.param("cuda", marks=pytest.mark.gpu)])
def be(request):
    if request.param == "cuda":
        import orbb200
        if orbb200.lib().orbb200_device_count() < 1:
            pytest.fail("the CUDA path needs a GPU: liborbb200 has no CPU fallback")
        return CudaBackend()
    return OracleBackend()


def _case(G, c):
    return {k.split("/", 1)[1]: G[k] for k in G.files if k.startswith(c + "/")}


def _frame(be, g, name="f"):
    return be.frame(g[name + "_kps_un"], g[name + "_desc"], g["bounds"])


def test_fixture_is_reference_generated(G):
    """The generator script and the harness that produced the fixture are committed; the fixture covers every dagger row of VERDICT r01."""
    assert os.path.exists(os.path.join(HERE, "golden", "make_golden_match.py"))
    assert os.path.exists(os.path.join(HERE, "..", "oracle", "refmatch_capi.cpp"))
    cases = {k.split("/")[0] for k in G.files}
    for need in ("frame", "stereo", "local_points_th1", "last_frame_mono", "last_frame_fwd", "last_frame_bwd", "reloc_th10", "loop_th10",
                 "fuse_stereo", "fuse_mono", "fuse_sim3", "sim3", "init_w100", "vocab", "bow", "triang_mono", "triang_only_stereo", "distinctive"):
        assert need in cases, need


def test_frame_undistort_bounds_grid_windows(be, G):
    """f1 / a15: Frame::UndistortKeyPoints, ComputeImageBounds (Frame.cpp:462-536), AssignFeaturesToGrid + GetFeaturesInArea
    (:241-257, :379-444) as run by the reference's Frame constructor."""
    g = _case(G, "frame")
    un = be.undistort(g["f_kps"], g["K4"], g["dist"])
    assert np.array_equal(un.view(np.uint8), g["f_kps_un"].view(np.uint8))
    assert np.array_equal(np.asarray(be.image_bounds(640, 480, g["K4"], g["dist"]), np.float32).view(np.uint32), g["bounds"].view(np.uint32))
    F = _frame(be, g)
    for i, (x, y, r, a, b) in enumerate(g["windows"]):
        want = g["win_idx"][g["win_off"][i]:g["win_off"][i + 1]]
        got = be.features_in_area(F, float(x), float(y), float(r), int(a), int(b))
        assert np.array_equal(got, want), (i, x, y, r, a, b)
    # the window with r = 1e4 enumerates the whole grid in (column, row, insertion) order = the reference's mGrid
    assert np.array_equal(g["win_idx"][g["win_off"][1]:g["win_off"][2]], g["grid_idx"])


def test_stereo_matches(be, G):
    """a14: Frame::ComputeStereoMatches (Frame.cpp:540-748) through the reference's stereo Frame constructor."""
    sys.path.insert(0, os.path.join(HERE, "..", "orb-slam2-annotation_b200"))
    from synth import synth_stereo
    g = _case(G, "stereo")
    L, R = synth_stereo(int(g["w"]), int(g["h"]), int(g["seed"]))
    kl, dl, kr, dr, ur, dp = be.stereo(L, R, int(g["nfeatures"]), float(g["mbf"]))
    assert kl.tobytes() == g["kps_left"].tobytes() and kr.tobytes() == g["kps_right"].tobytes()
    assert np.array_equal(dl, g["desc_left"]) and np.array_equal(dr, g["desc_right"])
    assert np.array_equal(ur.view(np.uint32), g["u_right"].view(np.uint32))
    assert np.array_equal(dp.view(np.uint32), g["depth"].view(np.uint32))


@pytest.mark.parametrize("case", ["local_points_th1", "local_points_th3", "local_points_th5"])
def test_search_by_projection_local_points(be, G, case):
    """a10: ORBmatcher::SearchByProjection(Frame&, const vector<MapPoint*>&, th) (ORBmatcher.cpp:63-155)."""
    g = _case(G, case); f = _case(G, "local_points")
    F = _frame(be, f)
    n, mp, ob = be.local_points(F, g["queries"], float(g["nnratio"]), g["mp0"], g["obs0"], f["f_u_right"])
    assert n == int(g["nmatches"]) and np.array_equal(mp, g["mp_out"])
    assert np.array_equal(ob[mp >= 0], g["obs_out"][mp >= 0])


@pytest.mark.parametrize("case", ["last_frame_mono", "last_frame_mono_th7", "last_frame_fwd", "last_frame_bwd", "last_frame_still"])
def test_search_by_projection_last_frame(be, G, case):
    """a11: ORBmatcher::SearchByProjection(Frame& Current, const Frame& Last, th, bMono) (:1506-1659): monocular, and stereo
    moving forward (levels >= last octave), backward (0 .. last octave) and standing still (octave +- 1), rotation histogram on / off."""
    g = _case(G, case)
    F = _frame(be, g)
    n, mp, ob = be.last_frame(F, g["queries"], bool(g["check_ori"]), g["mp0"], g["obs0"], g["f_u_right"])
    assert n == int(g["nmatches"]) and np.array_equal(mp, g["mp_out"])
    assert np.array_equal(ob[mp >= 0], g["obs_out"][mp >= 0])


@pytest.mark.parametrize("case", ["reloc_th10", "reloc_th3", "reloc_nocheck"])
def test_search_by_projection_reloc(be, G, case):
    """a12: SearchByProjection(Frame&, KeyFrame*, const set<MapPoint*>&, th, ORBdist) (:1661-1789)."""
    g = _case(G, case)
    n, mp = be.reloc(_frame(be, g), g["queries"], int(g["orb_dist"]), bool(g["check_ori"]), g["mp0"])
    assert n == int(g["nmatches"]) and np.array_equal(mp, g["mp_out"])


@pytest.mark.parametrize("case", ["loop_th10", "loop_th4_scaled"])
def test_search_by_projection_loop(be, G, case):
    """a12: SearchByProjection(KeyFrame*, cv::Mat Scw, const vector<MapPoint*>&, vector<MapPoint*>&, th) (:352-472)."""
    g = _case(G, case)
    n, m = be.loop(_frame(be, g), g["queries"], g["matched0"])
    assert n == int(g["nmatches"]) and np.array_equal(m, g["matched_out"])


@pytest.mark.parametrize("case", ["fuse_stereo", "fuse_mono", "fuse_sim3"])
def test_fuse_best(be, G, case):
    """a12b: the per-point search of Fuse(KeyFrame*, const vector<MapPoint*>&, th) (:962-1115, chi-square gates with / without u_right)
    and of Fuse(KeyFrame*, cv::Mat Scw, ...) (:1119-1251): bestIdx as the reference hands it to KeyFrame::GetMapPoint."""
    g = _case(G, case)
    gate = bool(g["gate"])
    ur = g["f_u_right"] if gate else None
    bi = be.fuse_best(_frame(be, g), g["queries"], ur, g["inv_sigma2"], gate, int(g["max_dist"]))
    assert np.array_equal(bi, g["best_idx"])
    assert int((bi >= 0).sum()) == int(g["nfused"])


def test_search_by_sim3(be, G):
    """a12b: SearchBySim3 (:1253-1500): each direction's search through fuse_best (TH_HIGH, no gate), then the mutual check (:1478-1495)."""
    g = _case(G, "sim3")
    F1 = be.frame(g["kf1_kps_un"], g["kf1_desc"], g["bounds"]); F2 = be.frame(g["kf2_kps_un"], g["kf2_desc"], g["bounds"])
    b12 = be.fuse_best(F2, g["q12"], None, None, False, 100)
    b21 = be.fuse_best(F1, g["q21"], None, None, False, 100)
    fop = g["feat_of_point"]
    vn1 = np.full(int(g["n1"]), -1, np.int32); vn2 = np.full(int(g["n2"]), -1, np.int32)
    vn1[fop[g["q12"]["id"]]] = b12; vn2[fop[g["q21"]["id"]]] = b21
    want = g["matches_out"]
    got = g["matches0"].copy()
    point_of_feat2 = np.full(int(g["n2"]), -1, np.int32)
    p2 = np.flatnonzero(np.arange(len(fop)) >= int(g["n1_points"]))
    point_of_feat2[fop[p2]] = p2
    found = 0
    for i1 in range(len(vn1)):
        idx2 = vn1[i1]
        if idx2 >= 0 and vn2[idx2] == i1:
            got[i1] = point_of_feat2[idx2]; found += 1
    assert found == int(g["nmatches"]) and np.array_equal(got, want)


@pytest.mark.parametrize("case", ["init_w100", "init_w30_nocheck"])
def test_search_for_initialization(be, G, case):
    """a13: SearchForInitialization (:474-590)."""
    g = _case(G, case)
    F1 = be.frame(g["f1_kps_un"], g["f1_desc"], g["bounds"]); F2 = be.frame(g["f2_kps_un"], g["f2_desc"], g["bounds"])
    n, m12, pm = be.initialization(F1, F2, g["prev_in"], int(g["window"]), float(g["nnratio"]), bool(g["check_ori"]))
    assert n == int(g["nmatches"]) and np.array_equal(m12, g["matches12"])
    assert np.array_equal(np.asarray(pm, np.float32).reshape(-1, 2).view(np.uint32), g["prev_out"].view(np.uint32))


# ---- f2 / f3: vocabulary, BoW searches, triangulation, distinctive descriptors ---------------------------------------------------
def _mod(be):
    return be.O if be.name == "oracle" else be.B


def _voc(be, g):
    m = _mod(be)
    cls = m.OracleVocabulary if be.name == "oracle" else m.Vocabulary
    return cls(int(g["k"]), int(g["L"]), g["parent"], g["is_leaf"], g["desc"], g["weight"])


def _fv(g, name):
    return (g[name + "_fv_nodes"], g[name + "_fv_off"], g[name + "_fv_feat"])


def test_vocabulary_transform_and_compute_bow(be, G):
    """f2: TemplatedVocabulary::transform (:1242-1285) per feature for levelsup 0 / 2 / 4 / 7 and Frame::ComputeBoW (Frame.cpp:452-459:
    BowVector doubles and FeatureVector as the reference's loadFromTextFile'd vocabulary produced them)."""
    g = _case(G, "vocab")
    voc = _voc(be, g)
    for lu in (0, 2, 4, 7):
        wid, w, nid = voc.transform(g["feats"], lu)
        assert np.array_equal(wid, g[f"word_lu{lu}"]) and np.array_equal(w.view(np.uint64), g[f"weight_lu{lu}"].view(np.uint64))
        assert np.array_equal(nid, g[f"node_lu{lu}"]), lu
    for name, lu in (("frame", 4), ("direct_lu2", 2)):
        ids, vals, fn, fo, ff = voc.bow(g["feats"], lu)
        assert np.array_equal(ids, g[name + "_bow_ids"]) and np.array_equal(np.asarray(vals).view(np.uint64), g[name + "_bow_vals"].view(np.uint64))
        assert np.array_equal(fn, g[name + "_fv_nodes"]) and np.array_equal(fo, g[name + "_fv_off"]) and np.array_equal(ff, g[name + "_fv_feat"])


@pytest.mark.parametrize("nnratio,check", [(0.7, True), (0.9, False)])
def test_search_by_bow(be, G, nnratio, check):
    """f2: SearchByBoW(KeyFrame*, Frame&, ...) (:205-348) and SearchByBoW(KeyFrame*, KeyFrame*, ...) (:604-743)."""
    g = _case(G, "bow")
    m = _mod(be)
    tag = f"{int(nnratio * 10)}_{int(check)}"
    out, n = m.search_by_bow(g["kf_desc"], g["kf_angle"], g["kf_has_mp"], _fv(g, "kf"), g["f_desc"], g["f_angle"], _fv(g, "f"), nnratio, check)
    assert n == int(g["kf_f_n_" + tag]) and np.array_equal(out, g["kf_f_matches_" + tag])
    out, n = m.search_by_bow_kf(g["kf_desc"], g["kf_angle"], g["kf_has_mp"], _fv(g, "kf"), g["f_desc"], g["f_angle_un"], g["f_has_mp"], _fv(g, "f"),
                                nnratio, check)
    assert n == int(g["kf_kf_n_" + tag]) and np.array_equal(out, g["kf_kf_matches_" + tag])


@pytest.mark.parametrize("case", ["triang_mono", "triang_stereo", "triang_only_stereo"])
def test_search_for_triangulation(be, G, case):
    """f2: SearchForTriangulation (:755-927) with CheckDistEpipolarLine (:166-190)."""
    g = _case(G, case)
    m = _mod(be)
    out, n = m.search_for_triangulation(g["kf1_desc"], g["kf1_kps_un"], g["kf1_u_right"], g["kf1_has_mp"], _fv(g, "kf1"),
                                        g["kf2_desc"], g["kf2_kps_un"], g["kf2_u_right"], g["kf2_has_mp"], _fv(g, "kf2"),
                                        g["F12"], float(g["ex"]), float(g["ey"]), g["scale"], g["sigma2"], bool(g["only_stereo"]), False)
    assert n == int(g["nmatches"]) and np.array_equal(out, g["match12"])


def test_distinctive_descriptors(be, G):
    """f3: MapPoint::ComputeDistinctiveDescriptors (MapPoint.cpp:302-380): the descriptor the reference stored per point."""
    g = _case(G, "distinctive")
    row, med = _mod(be).distinctive_descriptors(g["rows"], g["offsets"])
    off = g["offsets"]
    got = g["rows"][off[:-1] + row]
    assert np.all(row >= 0) and np.array_equal(got, g["best_desc"])
