"""Per-rank timeline of the step's tail under DDP (peer-memory exchange): in-kernel globaltimer stamps of k_body_reduce
(entry, predecessor done, partials reduced, pushed + published, world's flags arrived, exit) and of the head kernel's start,
so that the exposed part of the exchange splits into wire time (push) and waiting for the slowest rank (skew).
    python -m torch.distributed.run --nproc-per-node N tools/stamps_ddp.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
import ddp_cifar10_b200 as nrd
from ddp_cifar10_b200._lib import lib, check
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
B = 256
torch.manual_seed(0)
m = nrd.NetResDeep(32, 10, dtype="bf16").to(dev)
ddp = nrd.DistributedDataParallel(m, device_ids=[local], output_device=local)
step = nrd.TrainStep(ddp, B, lr=1e-2, use_graph=False)
check(lib.nrd_plan_set_option(step.plan.handle, 2, 1))
step._capture(0); step.use_graph = True
imgs, labels = nrd.SyntheticCifar10Loader(batch_size=B, device=dev, rank=rank).batch(0)
step.imgs.copy_(imgs); step.labels.copy_(labels)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
rows = []
for i in range(40):
    flush.zero_(); dist.barrier(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); step.run(); b.record(); torch.cuda.synchronize()
    d = step.plan.tensor_view(11).view(torch.int64).view(64, 16, 2).cpu()
    g = d[:, :, 1]
    if i >= 10:
        rd = [max(int(g[8, j]), int(g[9, j])) for j in range(6)]
        rows.append([a.elapsed_time(b) * 1e3, (rd[1] - int(g[0, 0])) / 1e3, (rd[2] - rd[1]) / 1e3, (rd[4] - rd[2]) / 1e3, (rd[5] - rd[4]) / 1e3, (rd[3] - rd[5]) / 1e3])
t = torch.tensor(rows, dtype=torch.float64, device=dev).median(0).values
out = [torch.empty_like(t) for _ in range(world)]
dist.all_gather(out, t)
if rank == 0:
    print(f"N={world} B={B}: per rank, median over 30 steps (us): step | head start -> reduce's predecessor done | reduce partials | push+publish | wait for the world's flags | sum + update")
    for r, o in enumerate(out):
        print(f"  rank {r}: " + " | ".join(f"{float(v):7.2f}" for v in o))
step.close()
dist.barrier(); os._exit(0)
