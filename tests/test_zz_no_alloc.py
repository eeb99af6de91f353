"""GPU: orbm_bf_top2_noalloc (scratch in a file-scope static device array, no allocation) against orbm_bf_top2_device, through the
C ABI.  Bit-exact (integer distances and row indices).  Collected last on purpose."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _both(db_t, q_t, row_offset=0):
    """(d1, idx1, d2) of the existing entry point and of the no-allocation one on the same device arrays."""
    import torch, orbb200
    nrows, nq = db_t.shape[0], q_t.shape[0]
    D = orbb200.DescriptorDB(max(nrows, 1), row_offset=row_offset)
    D.append_device(db_t.data_ptr(), nrows)
    outs = []
    for _ in range(2):
        outs.append((torch.full((nq,), -7, dtype=torch.int32, device="cuda"), torch.full((nq,), -7, dtype=torch.int64, device="cuda"),
                     torch.full((nq,), -7, dtype=torch.int32, device="cuda")))
    D.top2_device(q_t.data_ptr(), nq, *(t.data_ptr() for t in outs[0]))
    orbb200.bf_top2_noalloc(db_t.data_ptr(), nrows, row_offset, q_t.data_ptr(), nq, *(t.data_ptr() for t in outs[1]))
    torch.cuda.synchronize()
    D.close()
    return [tuple(t.cpu().numpy() for t in o) for o in outs]


def test_noalloc_equals_existing_entry_point_two_calls_in_a_row(gpu):
    import torch
    from synth import synth_descriptor_db, synth_queries
    db = synth_descriptor_db(300_001, seed=21)
    db[5] = 0; db[6] = 255; db[250_000] = db[321]
    q = synth_queries(db, 2000, seed=22)
    q[0] = db[321]; q[1] = 0; q[2] = 255
    dbt = torch.from_numpy(db).cuda(); qt = torch.from_numpy(q).cuda()
    # two calls in a row share the static array: the second (fewer queries, fewer rows, another offset) must not see the first's keys
    for nq, nrows, off in ((2000, 300_001, 0), (515, 70_001, 1_000_000), (2000, 300_001, 0), (1, 129, 0), (300, 1, 7)):
        ref, got = _both(dbt[:nrows], qt[:nq], off)
        for a, b, name in zip(ref, got, ("d1", "idx1", "d2")):
            assert np.array_equal(a, b), (name, nq, nrows, off)
        assert np.all(ref[0] <= ref[2]) and np.all(got[1] >= off)
    ref, got = _both(dbt, qt)
    assert got[0][0] == 0 and got[1][0] == 321 and got[2][0] == 0 and got[1][1] == 5 and got[1][2] == 6


def test_noalloc_limits_and_arguments(gpu):
    import torch, orbb200
    L = orbb200.lib()
    dbt = torch.zeros((1024, 32), dtype=torch.uint8, device="cuda")
    qt = torch.zeros((orbb200.NOALLOC_MAX_QUERIES + 1, 32), dtype=torch.uint8, device="cuda")
    d1 = torch.empty(len(qt), dtype=torch.int32, device="cuda"); i1 = torch.empty(len(qt), dtype=torch.int64, device="cuda")
    d2 = torch.empty(len(qt), dtype=torch.int32, device="cuda")
    args = lambda nrows, nq, rows=dbt.data_ptr(): (orbb200._p(rows), nrows, 0, orbb200._p(qt.data_ptr()), nq, orbb200._p(d1.data_ptr()),
                                                   orbb200._p(i1.data_ptr()), orbb200._p(d2.data_ptr()), 0, None)
    assert L.orbm_bf_top2_noalloc(*args(1024, orbb200.NOALLOC_MAX_QUERIES + 1)) == orbb200.E_CAPACITY      # too many queries
    assert b"4096" in L.orbb200_last_cuda_error()
    # 4096 queries leave 128 chunks of <= 2^23 rows: one row more is beyond the static scratch (sizes are checked before any access)
    assert L.orbm_bf_top2_noalloc(*args(128 * (1 << 23) + 1, 4096)) == orbb200.E_CAPACITY
    assert L.orbm_bf_top2_noalloc(*args(1024, 16, dbt.data_ptr() + 4)) == orbb200.E_INVALID                 # misaligned rows
    assert L.orbm_bf_top2_noalloc(*args(-1, 16)) == orbb200.E_INVALID
    assert L.orbm_bf_top2_noalloc(*args(1024, 0)) == orbb200.OK
    # the maximum query count itself works, and an empty shard gives (256, -1, 256)
    assert L.orbm_bf_top2_noalloc(*args(1024, orbb200.NOALLOC_MAX_QUERIES)) == orbb200.OK
    torch.cuda.synchronize()
    assert int(d1[:4096].max()) == 0 and int(i1[:4096].max()) == 0 and int(d2[:4096].max()) == 0
    assert L.orbm_bf_top2_noalloc(*args(0, 16)) == orbb200.OK
    torch.cuda.synchronize()
    assert d1[:16].tolist() == [256] * 16 and i1[:16].tolist() == [-1] * 16 and d2[:16].tolist() == [256] * 16
