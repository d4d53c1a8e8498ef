"""Pinned host -> device copy time as a function of size (what bounds bench.py's e2e number on a given box)."""
import time
import torch

for mb in (0.002, 0.25, 1.0, 3.15, 6.3, 12.6, 50.0):
    n = int(mb * 1e6 / 4)
    h = torch.randn(n).pin_memory(); d = torch.empty(n, device="cuda")
    for _ in range(5): d.copy_(h, non_blocking=True)
    torch.cuda.synchronize(); t = time.perf_counter()
    for _ in range(50): d.copy_(h, non_blocking=True)
    torch.cuda.synchronize(); dt = (time.perf_counter() - t) / 50
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(); d.copy_(h, non_blocking=True); ev1.record(); torch.cuda.synchronize()
    print(f"{n*4/1e6:8.3f} MB: back-to-back {dt*1e6:7.1f} us/copy ({n*4/dt/1e9:5.1f} GB/s), one copy between events {ev0.elapsed_time(ev1)*1e3:7.1f} us")
