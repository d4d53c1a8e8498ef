"""Does the number of CTAs of the persistent kernels change the distance to the rounding-matched oracle?  (ReLU-mask flips make
the body gradients' error a chaotic function of every fp32 summation order: this prints the spread.)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from oracle import netresdeep_oracle as O
from util import *
B, C, NB, lr = 256, 32, 10, 1e-2
torch.set_num_threads(16)
for seed in (4242 + B, 7):
    gen = torch.Generator().manual_seed(seed)
    x = torch.randn(B, 3, 32, 32, generator=gen); y = torch.randint(0, 10, (B,), generator=gen)
    p, b_rm = O.init_params(C, NB, seed=0)
    p_rm, l_rm, loss_rm, g_rm, masks_rm = O.train_step_rm(p, b_rm, x, y, NB, lr, keep_masks=True)
    for R in (0, 4, 8, 12, 16):
        os.environ["NRD_RESERVE_SMS"] = str(R)
        pm, bm = O.init_params(C, NB, seed=0)
        m = model_from_oracle(pm, bm, NB, "bf16")
        step = nrd.TrainStep(m, B, lr=lr)
        loss = float(step(x.to(DEV), y.to(DEV))); torch.cuda.synchronize()
        eng = m.engine
        g = {k: eng.flat_grads[eng.offsets[i]:eng.offsets[i + 1]].detach().float().cpu().view(-1) for i, k in enumerate(O.PARAM_ORDER)}
        print(f"seed {seed} reserve {R:2d}: logits {rell2(step.logits.cpu(), l_rm):.2e} " + " ".join(f"{k}={rell2(g[k], g_rm[k].flatten()):.2e}" for k in ("conv1.weight", "rb.conv.weight", "rb.bn.weight", "rb.bn.bias", "fc1.weight")), flush=True)
        step.close()
