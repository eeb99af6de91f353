"""CPU (-m "not gpu"): host-side logic of the N>1 path — shard plan, frame round-robin and the all-gather layout —
run with world_size 2 over gloo.  The per-shard scans come from the oracle here (the checker); on the GPU box the same
plumbing carries the CUDA kernel's results (tests/test_gpu_match.py::test_bf_top2_sharded_merge_equals_single)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def merge_rule(d1a, i1a, d2a):
    """SURVEY A.9 restated in numpy (test-side): sequential-scan semantics over ascending shards."""
    G, nq = d1a.shape
    b1 = np.full(nq, 256, np.int64); b2 = np.full(nq, 256, np.int64); bi = np.full(nq, -1, np.int64)
    for g in range(G):
        c1, c2, ci = d1a[g].astype(np.int64), d2a[g].astype(np.int64), i1a[g]
        ok = ci >= 0
        better = ok & (c1 < b1)
        b2 = np.where(better, np.minimum(b1, c2), np.where(ok, np.minimum(b2, c1), b2))
        bi = np.where(better, ci, bi)
        b1 = np.where(better, c1, b1)
    return b1, bi, b2


def test_shard_rows_and_frames():
    import sharding
    for total in (0, 1, 7, 10_000_000, 10_000_003):
        for world in (1, 2, 3, 8):
            cover = []
            for r in range(world):
                b, e = sharding.shard_rows(total, world, r)
                assert 0 <= b <= e <= total and (e - b) in (total // world, total // world + 1)
                cover.append((b, e))
            assert cover[0][0] == 0 and cover[-1][1] == total
            assert all(cover[i][1] == cover[i + 1][0] for i in range(world - 1))      # contiguous ascending ranges
    assert sharding.shard_rows(10_000_000, 8, 3) == (3_750_000, 5_000_000)            # C5: 1.25 M rows per GPU
    assert sharding.frames_for_rank(10, 4, 1) == [1, 5, 9]
    assert sorted(sum((sharding.frames_for_rank(4096, 8, r) for r in range(8)), [])) == list(range(4096))
    with pytest.raises(ValueError):
        sharding.shard_rows(10, 2, 2)


def _worker(rank, world, port, ret):
    for p in (os.path.join(ROOT, "orb-slam2-annotation_b200"), os.path.join(ROOT, "oracle")):
        sys.path.insert(0, p)
    import pyoracle
    import sharding
    from synth import synth_descriptor_db, synth_queries
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        M, nq = 6001, 96
        db = synth_descriptor_db(M, seed=9)
        db[4000] = db[100]; db[5000] = db[100]                    # exact duplicates across the shard boundary: tie rules
        q = synth_queries(db, nq, seed=10)
        q[0] = db[100]
        b, e = sharding.shard_rows(M, world, rank)
        d1, i1, d2 = pyoracle.bf_top2(db[b:e], q)
        i1 = np.where(i1 >= 0, i1 + b, i1)
        g1, gi, g2 = sharding.gather_top2(torch.from_numpy(d1), torch.from_numpy(i1), torch.from_numpy(d2))
        assert g1.shape == (world, nq) and gi.dtype == torch.int64
        assert np.array_equal(g1[rank].numpy(), d1) and np.array_equal(gi[rank].numpy(), i1) and np.array_equal(g2[rank].numpy(), d2)
        m1, mi, m2 = merge_rule(g1.numpy(), gi.numpy(), g2.numpy())
        f1, fi, f2 = pyoracle.bf_top2(db, q)                       # the sequential scan over the whole database
        assert np.array_equal(m1, f1) and np.array_equal(mi, fi) and np.array_equal(m2, f2)
        assert fi[0] == 100 and f1[0] == 0 and f2[0] == 0         # first index wins, second may equal best
        ret[rank] = "ok"
    finally:
        dist.destroy_process_group()


def test_gather_layout_world2_gloo():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(2, port, ret), nprocs=2, join=True)
    assert dict(ret) == {0: "ok", 1: "ok"}


def test_gather_single_process_passthrough():
    import sharding
    d1 = torch.arange(5, dtype=torch.int32); i1 = torch.arange(5, dtype=torch.int64); d2 = d1 + 1
    a, b, c = sharding.gather_top2(d1, i1, d2)
    assert a.shape == (1, 5) and torch.equal(a[0], d1) and torch.equal(b[0], i1) and torch.equal(c[0], d2)
