"""Pinned host<->device copy bandwidth on this box (ceiling for the e2e leg of bench.py). usage: python tools/pcie_bw.py"""
import torch, time
dev = torch.device("cuda:0")
for mb in (23, 185):
    n = mb << 20
    h = torch.empty(n, dtype=torch.uint8).pin_memory(); d = torch.empty(n, dtype=torch.uint8, device=dev)
    h2 = torch.empty(n // 6, dtype=torch.uint8).pin_memory(); d2 = torch.empty(n // 6, dtype=torch.uint8, device=dev)
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    for both in (False, True):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        for _ in range(20):
            with torch.cuda.stream(s1): d.copy_(h, non_blocking=True)
            if both:
                with torch.cuda.stream(s2): h2.copy_(d2, non_blocking=True)
        torch.cuda.synchronize(); dt = time.perf_counter() - t0
        print(f"{mb} MB H2D{' + D2H/6' if both else ''}: {20 * n / dt / 1e9:.1f} GB/s")
