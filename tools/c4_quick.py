"""Device-resident extraction timing at a config shape: c4_quick.py W H NFEAT FRAMES CAP"""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "orb-slam2-annotation_b200")]
import numpy as np, torch, orbb200
from synth import synth
w, h, nf, F, cap = (int(x) for x in sys.argv[1:6])
nd = 4 if w > 2000 else 16
distinct = np.stack([synth(w, h, 5000 + i) for i in range(nd)])
frames = torch.from_numpy(np.ascontiguousarray(np.tile(distinct, (F // nd, 1, 1)))).cuda()
ex = orbb200.ORBextractor(nf, 1.2, 8, 20, 7, max_size=(w, h), max_batch=F)
st = torch.cuda.ExternalStream(ex.stream())
kps = torch.zeros((F, cap, 7), device="cuda"); desc = torch.zeros((F, cap, 32), dtype=torch.uint8, device="cuda"); cnt = torch.zeros(F, dtype=torch.int32, device="cuda")
step = lambda: ex.extract_batch_device(frames.data_ptr(), F, w, h, w, w * h, kps.data_ptr(), desc.data_ptr(), cnt.data_ptr(), cap)
for _ in range(3): step()
ex.sync()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(st)
for _ in range(10): step()
e1.record(st); ex.sync()
ms = e0.elapsed_time(e1) / 10
print(f"{w}x{h} nf={nf} F={F}: {ms:.3f} ms/step, {F / ms * 1e3:.0f} frames/s, kp/frame {float(cnt.float().mean()):.1f}, env OCT_STREAMS={os.environ.get('ORBB_OCT_STREAMS')} NO_OVERLAP={os.environ.get('ORBB_NO_OCT_OVERLAP')}")
