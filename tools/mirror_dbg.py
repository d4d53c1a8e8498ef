import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ddp_cifar10_b200 as nrd
dev = torch.device("cuda:0")
torch.manual_seed(0)
m = nrd.NetResDeep(32, 10, dtype="bf16").to(dev)
step = nrd.TrainStep(m, 64, lr=1e-3)
print("after ctor: steps_run", step.steps_run, "seq", int(step._mirror_seq_np[0]), "ring", step._mirror_np[:8])
imgs, labels = nrd.SyntheticCifar10Loader(batch_size=64, device=dev).batch(0)
for i in range(4):
    l = step(imgs, labels); torch.cuda.synchronize()
    print("step", i, "device loss", float(l), "steps_run", step.steps_run, "seq", int(step._mirror_seq_np[0]), "ring", step._mirror_np[:8])
loader = nrd.SyntheticCifar10Loader(batch_size=64, steps_per_epoch=3, device=dev)
print([float(x) for x in step.iterate(loader, lookahead=1)], "steps_run", step.steps_run, "seq", int(step._mirror_seq_np[0]), "ring", step._mirror_np[:8])
