"""Host-side cost of enqueueing one chunk (orbx_extract_batch_device, no sync) vs its GPU time. usage: python tools/enqueue_cost.py [frames]"""
import sys, os, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "orb-slam2-annotation_b200")]
import numpy as np, torch, orbb200
from synth import synth
B = int(sys.argv[1]) if len(sys.argv) > 1 else 54
W, H, CAP = 640, 480, 1280
frames = torch.from_numpy(np.stack([synth(W, H, 1 + i % 16) for i in range(B)])).cuda()
ex = orbb200.ORBextractor(1000, max_size=(W, H), max_batch=B)
kps = torch.zeros((B, CAP, 7), device="cuda"); desc = torch.zeros((B, CAP, 32), dtype=torch.uint8, device="cuda"); cnt = torch.zeros(B, dtype=torch.int32, device="cuda")
def call(): ex.extract_batch_device(frames.data_ptr(), B, W, H, W, W * H, kps.data_ptr(), desc.data_ptr(), cnt.data_ptr(), CAP)
for _ in range(5): call()
ex.sync()
N = 20
t0 = time.perf_counter()
for _ in range(N): call()
t1 = time.perf_counter()
ex.sync()
t2 = time.perf_counter()
print(f"frames {B}: host enqueue {1e3 * (t1 - t0) / N:.3f} ms/call, total {1e3 * (t2 - t0) / N:.3f} ms/call")
