"""In-graph timeline of the training step: microseconds per main-stream kernel inside the replayed CUDA graph (event-record
nodes, NRD_OPT_TIMELINE), median over replays, L2 flushed between replays like bench.py.
    NRD_TIMELINE=1 python tools/timeline.py [B] [C] [NB]"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["NRD_TIMELINE"] = "1"
import torch
import ddp_cifar10_b200 as nrd

B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
C = int(sys.argv[2]) if len(sys.argv) > 2 else 32
NB = int(sys.argv[3]) if len(sys.argv) > 3 else 10
dev = torch.device("cuda:0")
torch.manual_seed(0)
m = nrd.NetResDeep(C, NB, dtype="bf16").to(dev)
step = nrd.TrainStep(m, B, lr=1e-2)
imgs, labels = nrd.SyntheticCifar10Loader(batch_size=B, device=dev).batch(0)
step.imgs.copy_(imgs); step.labels.copy_(labels)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
rows = []
for i in range(30):
    if os.environ.get("NO_FLUSH") != "1":
        flush.zero_()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); step.run(); b.record()
    tl = step.timeline()
    if i >= 5:
        rows.append((a.elapsed_time(b) * 1e3, tl))
names = [n for n, _ in rows[0][1]]
med = lambda v: sorted(v)[len(v) // 2]
tot = med([r[0] for r in rows])
print(f"B={B} C={C} NB={NB}: step (with marks) {tot:.1f} us")
s = 0.0
for i, n in enumerate(names):
    u = med([r[1][i][1] for r in rows]); s += u
    print(f"  {n:24s} {u:8.2f} us")
print(f"  {'sum of intervals':24s} {s:8.2f} us")
