"""Randomized parity sweep of the guided-matching entry points on synthetic frames (random keypoints, descriptors, query sets, windows,
contention, pre-assigned slots): SearchByProjection (tracked frame / local points / relocalisation / loop closing), the fusion search,
window search, brute-force top-2 and the distinctive-descriptor pick against the oracle.  usage: fuzz_match.py ncases seed"""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "orb-slam2-annotation_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
import numpy as np, orbb200, pyoracle
ncases = int(sys.argv[1]) if len(sys.argv) > 1 else 50
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 3)
scale = (1.2 ** np.arange(8)).astype(np.float32)


def oq(q):
    o = np.zeros(len(q), pyoracle.QUERY_DTYPE)
    o["u"] = q["q"]["u"]; o["v"] = q["q"]["v"]; o["r"] = q["q"]["r"]; o["minLevel"] = q["q"]["min_level"]; o["maxLevel"] = q["q"]["max_level"]
    o["desc"] = q["q"]["desc"]; o["uR"] = q["u_right"]; o["maxErrR"] = q["max_err_right"]; o["id"] = q["id"]; o["hasObs"] = q["has_obs"]; o["angle"] = q["angle"]
    return o


bad = 0
for case in range(ncases):
    n = int(rng.choice([0, 1, 7, 200, 1500, 4000, 11000])) if rng.random() < 0.3 else int(rng.integers(1, 5000))
    nq = int(rng.choice([0, 1, 33, 1025, 3000])) if rng.random() < 0.3 else int(rng.integers(1, 3500))
    w = int(rng.integers(200, 2000)); h = int(rng.integers(150, 1200))
    th = float(rng.choice([2.0, 7.0, 15.0, 40.0, 120.0])); reps = int(rng.integers(1, 4))
    tag = f"case {case}: n={n} nq={nq} {w}x{h} th={th} reps={reps}"
    try:
        kps = np.zeros(n, orbb200.KP_DTYPE)
        kps["x"] = rng.integers(0, w, n); kps["y"] = rng.integers(0, h, n); kps["octave"] = rng.integers(0, 8, n)
        kps["angle"] = rng.uniform(0, 360, n).astype(np.float32); kps["size"] = 31; kps["class_id"] = -1
        desc = rng.integers(0, 256, (n, 32), dtype=np.uint8)
        q = np.zeros(nq, orbb200.PROJ_QUERY_DTYPE)
        if n > 0 and nq > 0:
            src = rng.integers(0, n, nq)
            if reps > 1: src = np.repeat(src[: max(1, nq // reps)], reps)[:nq]; src = np.concatenate([src, rng.integers(0, n, nq - len(src))])
            q["q"]["u"] = kps["x"][src] + rng.uniform(-3, 3, nq); q["q"]["v"] = kps["y"][src] + rng.uniform(-3, 3, nq)
            q["q"]["desc"] = desc[src] ^ (rng.integers(0, 256, (nq, 32), dtype=np.uint8) & rng.integers(0, 256, (nq, 32), dtype=np.uint8) & rng.integers(0, 256, (nq, 32), dtype=np.uint8))
            lv = kps["octave"][src]
            q["angle"] = ((kps["angle"][src] + rng.choice([0.0, 0.0, 0.0, 90.0], nq)) % 360.0).astype(np.float32)      # keypoint angles live in [0, 360)
        else:
            q["q"]["u"] = rng.uniform(0, w, nq); q["q"]["v"] = rng.uniform(0, h, nq); lv = rng.integers(0, 8, nq)
            q["q"]["desc"] = rng.integers(0, 256, (nq, 32), dtype=np.uint8)
        q["q"]["r"] = th * scale[lv]
        mode = int(rng.integers(0, 3))
        q["q"]["min_level"] = [lv - 1, lv, np.full(nq, -1)][mode]; q["q"]["max_level"] = [lv + 1, np.full(nq, -1), lv][mode]
        q["u_right"] = q["q"]["u"] - rng.uniform(0, 30, nq); q["max_err_right"] = q["q"]["r"]
        q["id"] = np.arange(nq); q["has_obs"] = (rng.random(nq) < rng.choice([0.0, 0.5, 0.9, 1.0])).astype(np.int32)
        F = orbb200.FrameMatcher(kps, desc, (0, 0, w, h)); O = pyoracle.OracleFrame(kps, desc, (0, 0, w, h))
        mp0 = np.full(n, -1, np.int32); ob0 = np.zeros(n, np.int32)
        if n:
            pre = rng.random(n) < 0.1; mp0[pre] = 9000 + np.arange(pre.sum()); ob0[pre] = rng.random(pre.sum()) < 0.5
        ur = np.where(rng.random(n) < 0.5, kps["x"] - rng.uniform(0, 30, n), -1).astype(np.float32) if rng.random() < 0.5 else None
        O_q = oq(q)
        chk = bool(rng.random() < 0.5)
        if nq:
            r1 = orbb200.ORBmatcher(0.9, chk).SearchByProjection_frame(F, q, mp0, ob0, ur); o1 = O.search_last_frame(O_q, chk, mp0, ob0, ur)
            assert r1[0] == o1[0] and np.array_equal(r1[1], o1[1]) and np.array_equal(r1[2], o1[2]), "frame search"
            nn = float(rng.choice([0.6, 0.8, 0.95]))
            r2 = orbb200.ORBmatcher(nn).SearchByProjection_points(F, q, mp0, ob0, ur); o2 = O.search_local_points(O_q, nn, mp0, ob0, ur)
            assert r2[0] == o2[0] and np.array_equal(r2[1], o2[1]) and np.array_equal(r2[2], o2[2]), "local points search"
            od_ = int(rng.choice([50, 64, 100, 255]))      # (>= 256 makes the reference, and the oracle that restates it, write slot -1 for a query without candidates)
            r3 = orbb200.ORBmatcher(0.9, chk).SearchByProjection_reloc(F, q, mp0, od_); o3 = O.search_reloc(O_q, od_, chk, mp0)
            assert r3[0] == o3[0] and np.array_equal(r3[1], o3[1]), "reloc search"
            ql = q.copy(); ql["q"]["min_level"] = lv - 1; ql["q"]["max_level"] = lv          # the loop overload's window is [p - 1, p] (include/orbb200.h)
            r4 = orbb200.ORBmatcher().SearchByProjection_loop(F, ql, mp0); o4 = O.search_loop(oq(ql), mp0)
            assert r4[0] == o4[0] and np.array_equal(r4[1], o4[1]), "loop search"
        if nq and n:
            # fusion / Sim3 search: level window [p - 1, p] like the loop overload, chi-square gate with or without mvuRight
            sigma2 = (scale ** 2).astype(np.float32); inv_s2 = (np.float32(1.0) / sigma2).astype(np.float32)
            gate = bool(rng.random() < 0.5); maxd = int(rng.choice([50, 100]))
            urf = ur if gate else None
            bi, bd = orbb200.ORBmatcher().FuseBest(F, ql, urf, inv_s2 if gate else None, maxd)
            obi, obd = O.fuse_best(oq(ql), urf, inv_s2, gate, maxd)
            assert np.array_equal(bi, obi) and np.array_equal(bd, obd), "fuse best"
            # GetFeaturesInArea order and distances
            wq = np.zeros(min(nq, 400), orbb200.QUERY_DTYPE)
            for f_ in ("u", "v", "r", "min_level", "max_level", "desc"): wq[f_] = q["q"][f_][: len(wq)]
            offs, idx, dist = F.window_search(wq)
            ow = np.zeros(len(wq), pyoracle.QUERY_DTYPE)
            ow["u"] = wq["u"]; ow["v"] = wq["v"]; ow["r"] = wq["r"]; ow["minLevel"] = wq["min_level"]; ow["maxLevel"] = wq["max_level"]; ow["desc"] = wq["desc"]
            ooffs, oidx, odist = O.window_search(ow)
            assert np.array_equal(offs, ooffs) and np.array_equal(idx[: offs[-1]], oidx[: ooffs[-1]]) and np.array_equal(dist[: offs[-1]], odist[: ooffs[-1]]), "window search"
        if n > 1 and case % 3 == 0:
            # SearchForInitialization: the frame against a jittered copy of itself
            k2 = kps.copy(); k2["x"] = np.clip(kps["x"] + rng.integers(-4, 5, n), 0, w - 1); k2["y"] = np.clip(kps["y"] + rng.integers(-4, 5, n), 0, h - 1)
            d2 = desc ^ (rng.integers(0, 256, (n, 32), dtype=np.uint8) & rng.integers(0, 256, (n, 32), dtype=np.uint8) & rng.integers(0, 256, (n, 32), dtype=np.uint8))
            F2 = orbb200.FrameMatcher(k2, d2, (0, 0, w, h)); O2 = pyoracle.OracleFrame(k2, d2, (0, 0, w, h))
            prev = np.stack([kps["x"], kps["y"]], 1).astype(np.float32)
            win = int(rng.choice([10, 50, 100]))
            r5 = orbb200.ORBmatcher(0.9, chk).SearchForInitialization(F, F2, prev, win)
            o5 = pyoracle.search_for_initialization(O, O2, prev, win, 0.9, chk)
            assert r5[0] == o5[0] and np.array_equal(r5[1], o5[1]) and np.array_equal(r5[2], o5[2].reshape(r5[2].shape)), "initialization search"
            F2.close()
        F.close()
        if case % 4 == 0:
            nr = int(rng.integers(1, 70000)); nqq = int(rng.integers(1, 600))
            db = rng.integers(0, 256, (nr, 32), dtype=np.uint8); qq = rng.integers(0, 256, (nqq, 32), dtype=np.uint8)
            if nr > 10: qq[: nqq // 2] = db[rng.integers(0, nr, nqq // 2)]; db[rng.integers(0, nr, 5)] = db[0]
            D = orbb200.DescriptorDB(nr); D.append(db)
            d1, i1, d2 = D.top2(qq); e1, j1, e2 = pyoracle.bf_top2(db, qq)
            assert np.array_equal(d1, e1) and np.array_equal(i1, j1) and np.array_equal(d2, e2), "bf top2"
            D.close()
    except Exception as e:
        bad += 1
        print("FAIL", tag, "|", type(e).__name__, str(e)[:200], flush=True)
print("cases", ncases, "failures", bad)
