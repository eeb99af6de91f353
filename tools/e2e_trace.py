"""Per-chunk timeline of orbx_extract_batch (ORBB_TRACE=1) on 512 pinned C1 frames. usage: ORBB_TRACE=1 python tools/e2e_trace.py"""
import sys, os, time, ctypes as C
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "orb-slam2-annotation_b200")]
import numpy as np, torch, orbb200
from synth import synth
B, W, H, CAP = 512, 640, 480, 1280
distinct = np.stack([synth(W, H, 1 + i) for i in range(32)])
host = torch.from_numpy(np.ascontiguousarray(np.tile(distinct, (B // 32, 1, 1)))).pin_memory()
ex = orbb200.ORBextractor(1000, max_size=(W, H), max_batch=B)
hk = torch.zeros((B, CAP, 7), dtype=torch.float32).pin_memory(); hd = torch.zeros((B, CAP, 32), dtype=torch.uint8).pin_memory(); hc = np.zeros(B, np.int32)
L = orbb200.lib()
for i in range(4):
    t0 = time.perf_counter()
    rc = L.orbx_extract_batch(ex.h, C.c_void_p(host.data_ptr()), B, W, H, W, W * H, C.c_void_p(hk.data_ptr()), C.c_void_p(hd.data_ptr()), hc.ctypes.data_as(C.c_void_p), CAP)
    print(f"call {i}: rc {rc} wall {1e3 * (time.perf_counter() - t0):.3f} ms", file=sys.stderr)
