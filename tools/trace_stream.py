"""Where the streaming persistent kernels spend their time: clock64 stamps of one epilogue thread per CTA (NRD_OPT_TRACE).
    python tools/trace_stream.py [B]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ddp_cifar10_b200 as nrd
from ddp_cifar10_b200._lib import lib, check

B = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
dev = torch.device("cuda:0")
torch.manual_seed(0)
m = nrd.NetResDeep(32, 10, dtype="bf16").to(dev)
x = torch.randn(B, 3, 32, 32, device=dev); y = torch.randint(0, 10, (B,), device=dev)
step = nrd.TrainStep(m, B, lr=1e-3, use_graph=False)
pl = step.plan
for _ in range(3):
    step(x, y)
check(lib.nrd_plan_set_option(pl.handle, 2, 1))
pe = m.engine.plan(B, dev)
m.engine.forward(x)
check(lib.nrd_plan_set_option(pe.handle, 2, 1))
m.engine.forward(x)
torch.cuda.synchronize()
grid = min(B, 148)
tr = pe.tensor_view(6, 0).view(torch.int64)[: grid * 10 * 8].view(grid, 10, 8).cpu()
print(f"B={B}: streaming forward, cycles since block start [phase 1 end, atomics sent, exchange read, p1done+barrier, phase 2 end | waited for MMAs, for staging tiles]")
for cta in (0, grid - 1):
    for k in (0, 1, 5, 9):
        t = tr[cta, k]
        print(f"  CTA {cta} k={k}: " + " ".join(str(int(t[i] - t[0])) for i in range(1, 6)) + f" | {int(t[6])} {int(t[7])}" + (f"  next block +{int(tr[cta, k + 1, 0] - t[0])}" if k < 9 else ""))
step(x, y)
torch.cuda.synchronize()
trb = pl.tensor_view(2, 0).view(torch.int64)[: grid * 10 * 16].view(grid, 10, 16).cpu()
print("streaming backward, cycles since block start [atomics sent, exchange done, loop end | waited for loads, slots, MMAs; total in B, in C]")
for cta in (0, grid - 1):
    for it in (0, 1, 5, 9):
        t = trb[cta, it]
        print(f"  CTA {cta} it={it}: " + " ".join(str(int(t[i] - t[0])) for i in range(1, 4)) + " | " + " ".join(str(int(t[i])) for i in range(4, 9)) + (f"  next block +{int(trb[cta, it + 1, 0] - t[0])}" if it < 9 else ""))
