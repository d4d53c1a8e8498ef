"""Generate tests/golden/*.npz from the UNMODIFIED reference (run in the build container only).

    python oracle/make_golden.py            # needs /root/reference (read-only)

Imports /root/reference/model/resnet.py as-is, restates the 6-line step of
main_no_ddp.py:45-50 around it (main_no_ddp.py itself trains at import time and needs the
CIFAR-10 files, so it cannot be imported) and records inputs + outputs.  The W=2 fixture wraps
the same reference model in the real torch DistributedDataParallel over gloo/CPU, exactly as
main.py:63 does over NCCL.  Nothing under tests/ or bench.py reads /root/reference at run time;
they read these fixtures.
"""
import os
import sys

import numpy as np
import torch

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")
NAMES = {"conv1.weight": "conv1.weight", "conv1.bias": "conv1.bias", "rb.conv.weight": "resblocks.0.conv.weight",
         "rb.bn.weight": "resblocks.0.batch_norm.weight", "rb.bn.bias": "resblocks.0.batch_norm.bias",
         "fc1.weight": "fc1.weight", "fc1.bias": "fc1.bias", "fc2.weight": "fc2.weight", "fc2.bias": "fc2.bias"}


def _ref_model(C, NB, seed):
    sys.path.insert(0, REF)
    from model.resnet import NetResDeep  # the reference, unmodified
    torch.manual_seed(seed)
    return NetResDeep(n_chans1=C, n_blocks=NB)


def _data(B, seed=1234):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(B, 3, 32, 32, generator=g), torch.randint(0, 10, (B,), generator=g)


def single(C, NB, B, steps, tag, lr=1e-2, light=False):
    torch.set_num_threads(1)  # SURVEY 8(c) tier 5: fixed thread count => bitwise reproducible
    m = _ref_model(C, NB, 0)
    named = dict(m.named_parameters())
    rec = {"C": C, "NB": NB, "B": B, "lr": lr, "steps": steps}
    for k, rk in NAMES.items():
        w = named[rk].detach().numpy()
        rec[f"w0_sum/{k}"] = np.float64(w.astype(np.float64).sum())
        rec[f"w0_abs/{k}"] = np.float64(np.abs(w.astype(np.float64)).sum())
    opt = torch.optim.SGD(m.parameters(), lr=lr)                 # main_no_ddp.py:37
    loss_fn = torch.nn.CrossEntropyLoss()                        # main_no_ddp.py:38
    x, y = _data(B)
    rec["x_sum"] = np.float64(x.double().sum().item())
    if not light:            # light fixtures regenerate x from the seed and check x_sum
        rec["x"] = x.numpy()
    rec["labels"] = y.numpy()
    losses = []
    for s in range(steps):
        out = m(x)                                               # main_no_ddp.py:45
        loss = loss_fn(out, y)                                   # :46
        opt.zero_grad()                                          # :48
        loss.backward()                                          # :49
        if s == 0 and not light:
            rec["logits"] = out.detach().numpy().copy()
            for k, rk in NAMES.items():
                rec[f"grad/{k}"] = named[rk].grad.numpy().copy()
            bn = m.resblocks[0].batch_norm
            rec["running_mean"], rec["running_var"] = bn.running_mean.numpy().copy(), bn.running_var.numpy().copy()
            rec["num_batches_tracked"] = np.int64(bn.num_batches_tracked.item())
        opt.step()                                               # :50
        losses.append(loss.item())
    rec["losses"] = np.array(losses, np.float64)
    for k, rk in NAMES.items():
        w = named[rk].detach().numpy()
        rec[f"wN_sum/{k}"] = np.float64(w.astype(np.float64).sum())
    rec["state_dict_keys"] = np.array(list(m.state_dict().keys()))
    np.savez_compressed(os.path.join(OUT, f"{tag}.npz"), **rec)
    print(tag, "loss0", losses[0], "lossN", losses[-1])


def _ddp_worker(rank, W, C, NB, B, port, q):
    import torch.distributed as dist
    from torch.nn.parallel import DistributedDataParallel as DDP
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=W)
    torch.set_num_threads(1)
    m = _ref_model(C, NB, rank)          # ranks init DIFFERENTLY (main.py:62 is unseeded); DDP broadcasts rank 0
    ddp = DDP(m)                         # main.py:63 (device_ids omitted: CPU)
    named = dict(m.named_parameters())
    opt = torch.optim.SGD(ddp.parameters(), lr=1e-2)
    loss_fn = torch.nn.CrossEntropyLoss()
    x, y = _data(B, seed=1234 + rank)
    rec = {}
    for s in range(2):
        out = ddp(x); loss = loss_fn(out, y)
        opt.zero_grad(); loss.backward()
        bn = m.resblocks[0].batch_norm
        rec[f"s{s}/logits"] = out.detach().numpy().copy()
        rec[f"s{s}/loss"] = np.float64(loss.item())
        rec[f"s{s}/running_mean"] = bn.running_mean.numpy().copy()
        rec[f"s{s}/running_var"] = bn.running_var.numpy().copy()
        for k, rk in NAMES.items():
            rec[f"s{s}/grad/{k}"] = named[rk].grad.numpy().copy()
        opt.step()
    rec["x"], rec["labels"] = x.numpy(), y.numpy()
    q.put((rank, rec))
    dist.barrier()
    dist.destroy_process_group()


def ddp(C, NB, B, W, tag):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    procs = [ctx.Process(target=_ddp_worker, args=(r, W, C, NB, B, 29533, q)) for r in range(W)]
    [p.start() for p in procs]
    got = dict(q.get() for _ in range(W))
    [p.join() for p in procs]
    rec = {"C": C, "NB": NB, "B": B, "W": W}
    for r in range(W):
        for k, v in got[r].items():
            rec[f"r{r}/{k}"] = v
    np.savez_compressed(os.path.join(OUT, f"{tag}.npz"), **rec)
    print(tag, "losses", [got[r]["s0/loss"] for r in range(W)])


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    single(32, 10, 8, 3, "step_c32_nb10_b8")            # BASELINE config 1 shape, small batch
    single(64, 3, 4, 2, "step_c64_nb3_b4")              # wide-channel variant (config 4 family)
    single(32, 10, 64, 12, "curve_c32_nb10_b64_lr1e-2", light=True)  # degenerate lr=1e-2 curve, config 1
    ddp(32, 10, 8, 2, "ddp_w2_c32_nb10_b8")
    ddp(32, 10, 4, 4, "ddp_w4_c32_nb10_b4")             # four ranks: the mean over more than two contributions
