"""Data-parallel semantics on real GPUs (needs >= 2; skipped otherwise): our DistributedDataParallel wrapper + bucketed NCCL
all-reduce against the oracle's sequential simulation of torch DDP (which tests/test_oracle.py pins to real torch DDP over
gloo): rank-0 parameter broadcast at wrap time, rank-0 BN-buffer broadcast before every forward, mean of per-rank grads,
identical weights on every rank after the step; and the whole-step CUDA graph (all-reduce captured inside) == eager."""
import os

import pytest
import torch

pytestmark = pytest.mark.gpu


def _worker(rank, W, port, q):
    import torch.distributed as dist
    import ddp_cifar10_b200 as nrd
    from oracle import netresdeep_oracle as O
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=W, device_id=torch.device("cuda", rank))
    dev = f"cuda:{rank}"
    B, NB = 8, 10
    gen = torch.Generator().manual_seed(1234 + rank)
    x = torch.randn(B, 3, 32, 32, generator=gen); y = torch.randint(0, 10, (B,), generator=gen)
    out = {}
    for mode in ("eager", "graph"):
        torch.manual_seed(rank)                                  # ranks start DIFFERENT, as main.py:62 does
        m = nrd.NetResDeep(32, NB, dtype="fp32").to(dev)
        ddp = nrd.DistributedDataParallel(m, device_ids=[rank], output_device=rank)
        losses = []
        if mode == "eager":
            opt = nrd.SGD(ddp.parameters(), lr=1e-2); loss_fn = nrd.CrossEntropyLoss()
            for s in range(2):
                o = ddp(x.to(dev)); loss = loss_fn(o, y.to(dev)); opt.zero_grad(); loss.backward()
                if s == 0:
                    out["grads"] = m.engine.flat_grads.detach().cpu().numpy().copy()
                    out["logits"] = o.detach().cpu().numpy().copy()
                opt.step(); losses.append(float(loss))
        else:
            step = nrd.TrainStep(ddp, B, lr=1e-2)
            for s in range(2):
                losses.append(float(step(x.to(dev), y.to(dev))))
            step.graph = None; del step          # the graph holds captured NCCL kernels: drop it before the communicator
        out[mode] = (losses, m.engine.flat_params.detach().cpu().numpy().copy(), m.engine.bn_running.detach().cpu().numpy().copy())
    torch.cuda.synchronize()
    q.put((rank, out, x.numpy(), y.numpy()))     # numpy pickles inline: nothing refers back to this process after the put
    dist.barrier()
    os._exit(0)                                  # skip communicator teardown (can block behind captured collectives)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_ddp_two_gpus_matches_oracle_and_graph_equals_eager():
    import torch.multiprocessing as mp
    from oracle import netresdeep_oracle as O
    from util import rell2
    ctx = mp.get_context("spawn"); q = ctx.SimpleQueue(); W = 2
    ps = [ctx.Process(target=_worker, args=(r, W, 29561, q)) for r in range(W)]
    [p.start() for p in ps]
    got = {}
    for _ in range(W):
        r, out, x, y = q.get(); got[r] = (out, x, y)
    [p.join(60) for p in ps]
    [p.kill() for p in ps if p.is_alive()]
    # oracle: rank 0's init wins; sequential DDP simulation
    p, b0 = O.init_params(32, 10, seed=0)
    bufs = [dict((k, v.clone()) for k, v in b0.items()) for _ in range(W)]
    xs = [torch.from_numpy(got[r][1]) for r in range(W)]; ys = [torch.from_numpy(got[r][2]) for r in range(W)]
    p1, logits, losses, gavg = O.ddp_train_step(p, bufs, xs, ys, 10, 1e-2)
    import ddp_cifar10_b200 as nrd
    off = nrd._lib.param_offsets(32)
    order = ["conv1.weight", "conv1.bias", "rb.conv.weight", "rb.bn.weight", "rb.bn.bias", "fc1.weight", "fc1.bias", "fc2.weight", "fc2.bias"]
    for r in range(W):
        out = got[r][0]
        out = {k: (torch.from_numpy(v) if not isinstance(v, tuple) else (v[0], torch.from_numpy(v[1]), torch.from_numpy(v[2]))) for k, v in out.items()}
        got[r] = (out, got[r][1], got[r][2])
        assert rell2(out["logits"], logits[r]) < 1e-5
        assert abs(out["eager"][0][0] - float(losses[r])) < 1e-5 * float(losses[r])
        for i, k in enumerate(order):
            tol = 2e-5 if k.startswith("fc") else 5e-3
            assert rell2(out["grads"][off[i]:off[i + 1]], gavg[k].flatten()) < tol, (r, k)
        # graph path == eager path, bit for bit; every rank ends with the same weights
        assert out["graph"][0] == out["eager"][0]
        assert torch.equal(out["graph"][1], out["eager"][1]) and torch.equal(out["graph"][2], out["eager"][2])
        assert torch.equal(out["eager"][1], torch.as_tensor(got[0][0]["eager"][1]))
