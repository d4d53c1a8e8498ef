"""In-kernel stamps of the small kernels (NRD_OPT_TRACE): where the head / reduce kernels spend their time inside the replayed graph."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ddp_cifar10_b200 as nrd
from ddp_cifar10_b200._lib import lib, check
B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
dev = torch.device("cuda:0")
torch.manual_seed(0)
m = nrd.NetResDeep(32, 10, dtype="bf16").to(dev)
step = nrd.TrainStep(m, B, lr=1e-2, use_graph=False)
check(lib.nrd_plan_set_option(step.plan.handle, 2, 1))
step._capture(0); step.use_graph = True
imgs, labels = nrd.SyntheticCifar10Loader(batch_size=B, device=dev).batch(0)
step.imgs.copy_(imgs); step.labels.copy_(labels)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for i in range(5):
    flush.zero_(); step.run()
torch.cuda.synchronize()
d = step.plan.tensor_view(11).view(torch.int64).view(64, 16, 2).cpu()
def show(row, n, names):
    c = d[row, :n, 0]; g = d[row, :n, 1]
    print(f"  row {row}: clock64 deltas", [int(c[i] - c[0]) for i in range(n)], "globaltimer deltas ns", [int(g[i] - g[0]) for i in range(n)], names)
print("head kernel, CTAs (0,0) (0,KS-1) (last,0) (last,KS-1): start, rg loop top, pooled, partial stored, barrier passed, logits, dh1, dp stored")
for r in range(4): show(r, 8, "")
print("body_reduce CTA 0 / last: entry, after pdl wait, loads reduced, exit")
for r in (8, 9): show(r, 4, "")
g = d[:, :, 1]
print("head start -> reduce exit (globaltimer, us):", (int(g[9, 3]) - int(g[0, 0])) / 1e3, " head span", (max(int(g[r, 7]) for r in range(4)) - min(int(g[r, 0]) for r in range(4))) / 1e3,
      "reduce span", (max(int(g[8, 3]), int(g[9, 3])) - min(int(g[8, 0]), int(g[9, 0]))) / 1e3)

print("stem fwd CTA 0 / last: entry, prologue done, image 0 landed, tiles h0, tiles h1, acc h0 ready, epi h0 done, acc h1 ready, epi h1 done, exit")
for r in (16, 17): show(r, 10, "")
print("stem bwd CTA 0 / last: entry, setup done, pdl wait done, image 0 landed, A tiles it0..3, G tiles it0..3, accumulator done, exit")
for r in (24, 25): show(r, 14, "")
