import sys, os
sys.path[:0] = ["orb-slam2-annotation_b200"]
import numpy as np, torch, orbb200
from synth import synth
w, h, cap = 640, 480, 1280
fr = torch.from_numpy(synth(w, h, 4242)[None]).cuda()
ex = orbb200.ORBextractor(1000, 1.2, 8, 20, 7, max_size=(w, h), max_batch=1)
k = torch.zeros((1, cap, 7), device="cuda"); d = torch.zeros((1, cap, 32), dtype=torch.uint8, device="cuda"); c = torch.zeros(1, dtype=torch.int32, device="cuda")
for _ in range(6):
    ex.extract_batch_device(fr.data_ptr(), 1, w, h, w, w * h, k.data_ptr(), d.data_ptr(), c.data_ptr(), cap); ex.sync()
