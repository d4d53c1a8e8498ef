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
    print(f"  kernel entry -> first block start {int(tr[cta,0,0]-tr[cta,0,7])}, last pass 2 -> kernel exit {int(tr[cta,9,7]-tr[cta,9,6])}, whole kernel {int(tr[cta,9,7]-tr[cta,0,7])} cycles")

st = pl.tensor_view(6, 0).view(torch.int64)[grid * 10 * 8: grid * 10 * 8 + grid * 16].view(grid, 16).cpu()
if int(st[0, 0]) != 0:
    e = int(tr[0, 0, 7])
    print("fused stem, CTA 0 (cycles since kernel entry): images in smem", int(st[0, 0]) - e,
          " per half [built, accumulators ready, epilogue done]:", [[int(st[0, 1 + i * 3 + j]) - e for j in range(3)] for i in range(4)])

# ---- backward ----
y = torch.randint(0, 10, (B,), device=dev)
step = nrd.TrainStep(m, B, lr=1e-3, use_graph=False)          # the graph path's backward (fused stem backward when on)
check(lib.nrd_plan_set_option(step.plan.handle, 2, 1))
pl = step.plan
step(x, y)
torch.cuda.synchronize()
trb = pl.tensor_view(2, 0).view(torch.int64)[: grid * 10 * 16].view(grid, 10, 16).cpu()
namesb = ["start", "phaseA", "pre-barrier", "post-barrier", "combined", "phaseB(dy ready)", "phaseC(dx)"]
for cta in (0, grid - 1):
    print(f"BWD CTA {cta}:")
    for it in range(10):
        t = trb[cta, it]
        print("  it=%d " % it + f"sync={int(t[7]-t[0]):6d} " + " ".join(f"{namesb[i]}={int(t[i]-t[0]):6d}" for i in range(1, 7)) + (f"  next=+{int(trb[cta,it+1,0]-t[0])}" if it < 9 else "") + "  A:" + ",".join(str(int(t[i]-t[7])) for i in range(8, 14)))
    print(f"  kernel entry -> first block start {int(trb[cta,0,0]-trb[cta,0,14])}, last phase C -> kernel exit {int(trb[cta,9,15]-trb[cta,9,6])}, whole kernel {int(trb[cta,9,15]-trb[cta,0,14])} cycles")

# ---- who is late?  wait = post-barrier - pre-barrier per CTA (the last arriver waits least) ----
if os.environ.get("TRACE_WAITS"):
    import numpy as np
    for name, t, i0, i1, seg in (("fwd", tr, 3, 4, [(0, 2, "start->pass1"), (2, 3, "pass1->atomics"), (4, 5, "combine"), (5, 6, "pass2")]),
                                 ("bwd", trb, 2, 3, [(0, 2, "start->atomics"), (3, 4, "combine"), (4, 5, "phaseB"), (5, 6, "phaseC+A")])):
        w = (t[:, 2:9, i1] - t[:, 2:9, i0]).float().mean(1).numpy()
        order = np.argsort(w)
        wb = (t[:108, :, i1] - t[:108, :, i0]).numpy()
        print("   per block, 2-image CTAs: min wait", [int(x) for x in wb.min(0)], "argmin", [int(x) for x in wb.argmin(0)], "median", [int(x) for x in np.median(wb, 0)])
        print(f"{name}: mean wait per CTA over blocks 2..8: min {w.min():.0f} (CTA {order[0]}), median {np.median(w):.0f}, max {w.max():.0f} (CTA {order[-1]})")
        print("   least-waiting CTAs:", [(int(c), int(w[c])) for c in order[:12]])
        two = w[: 2 * 148 - 256 if False else 108]
        print(f"   2-image CTAs (0..107): mean wait {w[:108].mean():.0f}; 1-image CTAs: mean wait {w[108:].mean():.0f}")
        for a_, b_, nm in seg:
            d_ = (t[:, 2:9, b_] - t[:, 2:9, a_]).float().mean(1).numpy()
            print(f"   {nm}: 2-image CTAs mean {d_[:108].mean():.0f} min {d_[:108].min():.0f} max {d_[:108].max():.0f} (CTA {int(np.argmax(d_[:108]))})")
