"""Brute-force top-2 timing, 2000 queries x 10 M rows, and a bit-equality check against the integer-pipe variant (ORBB_BF_VARIANT=1)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "orb-slam2-annotation_b200")]
import numpy as np, torch, orbb200
NQ, ROWS = int(os.environ.get("NQ", 2000)), int(os.environ.get("ROWS", 10_000_000))
g = torch.Generator(device="cuda"); g.manual_seed(42)
db = torch.randint(0, 256, (ROWS, 32), dtype=torch.uint8, device="cuda", generator=g)
q = torch.randint(0, 256, (NQ, 32), dtype=torch.uint8, device="cuda", generator=g)
q[: NQ // 2] = db[torch.randint(0, ROWS, (NQ // 2,), device="cuda", generator=g)]
q[: NQ // 2, :3] ^= 0x55
res = {}
for var in (os.environ.get("ORBB_BF_VARIANT", "3"), "1"):
    os.environ["ORBB_BF_VARIANT"] = var
    D = orbb200.DescriptorDB(ROWS); D.append_device(db.data_ptr(), ROWS)
    d1 = torch.empty(NQ, dtype=torch.int32, device="cuda"); i1 = torch.empty(NQ, dtype=torch.int64, device="cuda"); d2 = torch.empty(NQ, dtype=torch.int32, device="cuda")
    st = torch.cuda.current_stream()
    for _ in range(2): D.top2_device(q.data_ptr(), NQ, d1.data_ptr(), i1.data_ptr(), d2.data_ptr(), st.cuda_stream)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): D.top2_device(q.data_ptr(), NQ, d1.data_ptr(), i1.data_ptr(), d2.data_ptr(), st.cuda_stream)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    res[var] = (d1.cpu(), i1.cpu(), d2.cpu())
    print(f"variant {var}: {ms:.3f} ms, {NQ * ROWS / ms / 1e6:.1f} Gcmp/s")
    del D
a, b = list(res.values())
print("bit-equal to variant 1:", all(torch.equal(x, y) for x, y in zip(a, b)))
