"""Small fixed workload for ncu: 64-frame C1 batches (--frames N for the bench's 512) through the device-resident API, plus one 2 M-row top-2."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "orb-slam2-annotation_b200")]
import numpy as np, torch, orbb200
from synth import synth
B, W, H, CAP = 64, 640, 480, 1280
if "--frames" in sys.argv: B = int(sys.argv[sys.argv.index("--frames") + 1])
frames = torch.from_numpy(np.stack([synth(W, H, 1 + i % 16) for i in range(B)])).cuda()
ex = orbb200.ORBextractor(1000, max_size=(W, H), max_batch=B)
kps = torch.zeros((B, CAP, 7), device="cuda"); desc = torch.zeros((B, CAP, 32), dtype=torch.uint8, device="cuda"); cnt = torch.zeros(B, dtype=torch.int32, device="cuda")
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
for _ in range(steps):
    ex.extract_batch_device(frames.data_ptr(), B, W, H, W, W * H, kps.data_ptr(), desc.data_ptr(), cnt.data_ptr(), CAP)
ex.sync()
if "--match" in sys.argv:
    M = 2_000_000
    db = torch.randint(0, 256, (M, 32), dtype=torch.uint8, device="cuda")
    q = torch.randint(0, 256, (2000, 32), dtype=torch.uint8, device="cuda")
    D = orbb200.DescriptorDB(M); D.append_device(db.data_ptr(), M)
    d1 = torch.empty(2000, dtype=torch.int32, device="cuda"); i1 = torch.empty(2000, dtype=torch.int64, device="cuda"); d2 = torch.empty(2000, dtype=torch.int32, device="cuda")
    for _ in range(2):
        D.top2_device(q.data_ptr(), 2000, d1.data_ptr(), i1.data_ptr(), d2.data_ptr())
    torch.cuda.synchronize()
print("ok", int(cnt.sum()))
