"""bench.py's own e2e section, repeated, with and without the things that precede it in bench.py."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ddp_cifar10_b200 as nrd
B = 256
dev = torch.device("cuda:0")
torch.manual_seed(0)
m = nrd.NetResDeep(32, 10, dtype="bf16").to(dev)
step = nrd.TrainStep(m, B, lr=1e-2)
imgs, labels = nrd.SyntheticCifar10Loader(batch_size=B, device=dev).batch(0)
step.imgs.copy_(imgs); step.labels.copy_(labels)
h_imgs = imgs.cpu().pin_memory(); h_labels = labels.cpu().pin_memory()
class HostBatches:
    def __init__(self, n): self.n = n
    def __len__(self): return self.n
    def __iter__(self): return ((h_imgs, h_labels) for _ in range(self.n))
def epoch(n):
    tot = 0.0
    for loss in step.iterate(HostBatches(n), lookahead=1):
        tot += float(loss)
    return tot
def timed(n):
    torch.cuda.synchronize(); t0 = time.perf_counter(); epoch(n); torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e6
epoch(3)
print("e2e epochs of 200:", [round(timed(200), 1) for _ in range(5)])
if "--flush" in sys.argv:
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(200)]
    for s, e in evs:
        flush.zero_(); s.record(); step.run(); e.record()
    torch.cuda.synchronize()
    print("after a flushed timed loop:", [round(timed(200), 1) for _ in range(5)])
    from bench import ClockSampler
    sm = ClockSampler(0); sm.start(); time.sleep(0.3); print("clock sampler:", sm.stop())
    print("after the clock sampler:", [round(timed(200), 1) for _ in range(5)])
