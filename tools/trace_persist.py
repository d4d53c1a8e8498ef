"""Phase timeline of the persistent forward kernel (clock64 stamps, see NRD_OPT_TRACE)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ddp_cifar10_b200 as nrd
from ddp_cifar10_b200._lib import lib, check

B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
dev = torch.device("cuda:0")
torch.manual_seed(0)
m = nrd.NetResDeep(32, 10, dtype="bf16").to(dev)
x = torch.randn(B, 3, 32, 32, device=dev)
pl = m.engine.plan(B, dev)
for _ in range(3):
    m.engine.forward(x)
check(lib.nrd_plan_set_option(pl.handle, 2, int(os.environ.get('TRACE_MODE', '1'))))
m.engine.forward(x)
torch.cuda.synchronize()
grid = min(B, 148)
tr = pl.tensor_view(6, 0).view(torch.int64)[: grid * 10 * 8].view(grid, 10, 8).cpu()
names = ["start", "tfull0", "tfull1", "pre-barrier", "post-barrier", "combined", "pass2 done"]
for cta in (0, 1, grid - 1):
    print(f"CTA {cta}: cycles since block start (k rows)")
    for k in range(10):
        t = tr[cta, k]
        print("  k=%d " % k + " ".join(f"{names[i]}={int(t[i]-t[0]):6d}" for i in range(1, 7)) + (f"  next_start=+{int(tr[cta,k+1,0]-t[0])}" if k < 9 else ""))

# ---- backward ----
y = torch.randint(0, 10, (B,), device=dev)
out = m(x); loss = nrd.CrossEntropyLoss()(out, y); loss.backward()
torch.cuda.synchronize()
trb = pl.tensor_view(2, 0).view(torch.int64)[: grid * 10 * 16].view(grid, 10, 16).cpu()
namesb = ["start", "phaseA", "pre-barrier", "post-barrier", "combined", "phaseB(dy ready)", "phaseC(dx)"]
for cta in (0, grid - 1):
    print(f"BWD CTA {cta}:")
    for it in range(10):
        t = trb[cta, it]
        print("  it=%d " % it + f"sync={int(t[7]-t[0]):6d} " + " ".join(f"{namesb[i]}={int(t[i]-t[0]):6d}" for i in range(1, 7)) + (f"  next=+{int(trb[cta,it+1,0]-t[0])}" if it < 9 else "") + "  A:" + ",".join(str(int(t[i]-t[7])) for i in range(8, 14)))
