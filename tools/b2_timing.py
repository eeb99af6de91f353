"""Diagnostic (needs a -DB2_TIMING build as liborbb200.so): clock stamps of the accumulator hand-over chain of bf_top2_tc2_kernel."""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "orb-slam2-annotation_b200")]
import numpy as np, torch, orbb200
os.environ["ORBB_BF_VARIANT"] = "4"
NQ, ROWS = 2000, 10_000_000
db = torch.randint(0, 256, (ROWS, 32), dtype=torch.uint8, device="cuda")
q = torch.randint(0, 256, (NQ, 32), dtype=torch.uint8, device="cuda")
D = orbb200.DescriptorDB(ROWS); D.append_device(db.data_ptr(), ROWS)
d1 = torch.empty(NQ, dtype=torch.int32, device="cuda"); i1 = torch.empty(NQ, dtype=torch.int64, device="cuda"); d2 = torch.empty(NQ, dtype=torch.int32, device="cuda")
for _ in range(3): D.top2_device(q.data_ptr(), NQ, d1.data_ptr(), i1.data_ptr(), d2.data_ptr())
torch.cuda.synchronize()
t = np.zeros((32, 2, 8), np.int64)
assert orbb200.lib().orbb200__b2_timing(t.ctypes.data_as(C.c_void_p)) == 0
t0 = t[4, 0, 2]
names = ["iss_wait", "iss_free", "iss_done", "con_wait", "con_full", "con_ld0", "con_ld15"]
for it in range(4, 12):
    for mt in range(2):
        print(it, mt, {n: int(t[it, mt, i] - t0) for i, n in enumerate(names)})
stage = (t[30, 0, 2] - t[4, 0, 2]) / 26
print("cycles per stage", stage)
# hand-over of set (it, mt): MMAs issued at iss_done(it) -> consumer sees full -> loads done -> issuer sees buffer free at iss_free(it+1)
full_after_issue = np.mean([t[it, mt, 4] - t[it, mt, 2] for it in range(4, 30) for mt in range(2)])
ld = np.mean([max(t[it, mt, 5], t[it, mt, 6]) - t[it, mt, 4] for it in range(4, 30) for mt in range(2)])
free_after_ld = np.mean([t[it + 1, mt, 1] - max(t[it, mt, 5], t[it, mt, 6]) for it in range(4, 30) for mt in range(2)])
wait_issuer = np.mean([t[it, mt, 1] - t[it, mt, 0] for it in range(4, 30) for mt in range(2)])
con_wait = np.mean([t[it, mt, 4] - t[it, mt, 3] for it in range(4, 30) for mt in range(2)])
print(f"issue done -> consumer sees accfull {full_after_issue:.0f}; accfull -> 64 columns loaded {ld:.0f}; loaded -> issuer sees accempty {free_after_ld:.0f}; issuer waits {wait_issuer:.0f}; consumer waits {con_wait:.0f}")
g = np.zeros((2, 32, 2, 8), np.int64)
assert orbb200.lib().orbb200__b2_gtiming(g.ctypes.data_as(C.c_void_p)) == 0
print("globaltimer (ns), relative to the leader's issue-done of the set:")
for it in range(8, 14):
    for mt in range(2):
        b = g[0, it, mt, 2]
        print(it, mt, "cta0 full/ld0/ld15", [int(g[0, it, mt, i] - b) for i in (4, 5, 6)], "cta1 full/ld0/ld15", [int(g[1, it, mt, i] - b) for i in (4, 5, 6)],
              "leader sees free", int(g[0, it + 1, mt, 1] - b))
