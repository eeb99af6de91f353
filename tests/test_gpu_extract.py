"""GPU parity of the extractor: every stage of ORBextractor::operator() against the CPU oracle, through the C ABI."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SHAPES = [  # (w, h, nfeatures, seed)  C1 / C2 / C3 shapes of BASELINE.json
    (640, 480, 1000, 1),
    (1241, 376, 2000, 2),
    (752, 480, 1200, 3),
]


def _sorted_rows(a):
    a = np.asarray(a).reshape(-1, 3)
    return a[np.lexsort((a[:, 2], a[:, 0], a[:, 1]))]


def run_pair(w, h, nf, seed, **kw):
    import orbb200, pyoracle
    from synth import synth
    img = synth(w, h, seed)
    orc = pyoracle.OracleExtractor(nf, **kw)
    ok, od = orc.extract(img)
    ex = orbb200.ORBextractor(nf, max_size=(w, h), **{k: v for k, v in kw.items() if k in ()})
    gk, gd = ex(img)
    return img, orc, ok, od, ex, gk, gd


@pytest.mark.parametrize("w,h,nf,seed", SHAPES)
def test_stage_parity(gpu, w, h, nf, seed):
    img, orc, ok, od, ex, gk, gd = run_pair(w, h, nf, seed)
    for L in range(8):
        assert np.array_equal(ex.pyramid_level(L), orc.level_image(L)), f"pyramid level {L}"
        assert np.array_equal(ex.blurred_level(L), orc.level_image(L, blurred=True)), f"blur level {L}"
        raw_o = orc.level_raw(L).astype(np.int64)
        raw_g = ex.fast_keypoints(L).astype(np.int64)
        assert raw_o.shape == raw_g.shape, f"FAST count level {L}: {raw_o.shape} vs {raw_g.shape}"
        assert np.array_equal(_sorted_rows(raw_o), _sorted_rows(raw_g)), f"FAST keypoint set level {L}"
        lo, lg = orc.level_kps(L), ex.level_keypoints(L)
        assert len(lo) == len(lg), f"octree count level {L}"
        for f in ("x", "y", "response", "octave", "size"):
            assert np.array_equal(lo[f], lg[f]), f"octree {f} level {L} (set and ORDER)"
        assert np.max(np.abs(lo["angle"] - lg["angle"]), initial=0) <= 1e-3, f"angle level {L}"


@pytest.mark.parametrize("w,h,nf,seed", SHAPES)
def test_final_parity(gpu, w, h, nf, seed):
    img, orc, ok, od, ex, gk, gd = run_pair(w, h, nf, seed)
    assert len(ok) == len(gk)
    for f in ("x", "y", "size", "response", "octave", "class_id"):
        assert np.array_equal(ok[f], gk[f]), f
    assert np.max(np.abs(ok["angle"] - gk["angle"])) <= 1e-3      # tolerance of BASELINE.json north_star
    bad = np.any(od != gd, axis=1).sum()
    assert bad <= 0.001 * len(od), f"{bad} of {len(od)} descriptors differ (> 0.1 %)"
