"""Data-parallel semantics on real GPUs, W = 2, 4, 8 ranks (each case is skipped when the box has fewer GPUs): our
DistributedDataParallel wrapper and the whole-step graph path against the oracle's sequential simulation of torch DDP (which
tests/test_oracle.py pins to REAL torch DDP over gloo at W = 2 and 4):

  * rank-0 parameter broadcast at wrap time (the ranks start from different seeds, as main.py:62 does),
  * rank 0's BN buffers in front of every forward (step 2 starts from them on every rank),
  * mean of the per-rank gradients, identical weights on every rank after every step,
  * fp32: eager wrapper (bucketed NCCL all-reduce) vs the fp32 oracle to 1e-5; graph path == eager path (bit for bit at
    W = 2, where a two-term sum has one order; to 1e-6 beyond), for BOTH exchanges of the graph path: our kernels over NVLink
    peer memory ("p2p") and NCCL all-reduces captured in the graph ("nccl");
  * bf16 (the path bench.py --gpus N times): graph path, both exchanges, against the rounding-matched oracle inside the
    single-GPU envelopes of tests/test_gpu_parity_bench.py; p2p == nccl bit for bit at W = 2; the same with the streaming
    persistent kernels forced (NRD_STREAM=1: what per-rank batches beyond two images per CTA run);
  * the literal reference loop (stock torch.optim.SGD + nn.CrossEntropyLoss, main.py:27-41) around our wrapper.

One process group per W runs every combination (the exchange is chosen per TrainStep by NRD_P2P), so a case costs one spawn.
"""
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

B, NB, LR = 8, 10, 1e-2
B_BF16 = 40            # bf16: enough pixels per channel for the A.7 envelopes to apply (B*256 = 10k per statistic)
LR_BF16 = 1e-3         # bf16 runs two steps: at 1e-2 the second one sits in the loss spike of SURVEY fact 9 (noise amplifier)


def _worker(rank, W, port, q):
    import torch.distributed as dist
    import ddp_cifar10_b200 as nrd
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=W, device_id=torch.device("cuda", rank))
    dev = f"cuda:{rank}"
    out = {}

    def data(b):
        gen = torch.Generator().manual_seed(1234 + rank)
        return torch.randn(b, 3, 32, 32, generator=gen), torch.randint(0, 10, (b,), generator=gen)

    def fresh(dtype):
        torch.manual_seed(rank)                                  # ranks start DIFFERENT, as main.py:62 does
        m = nrd.NetResDeep(32, NB, dtype=dtype).to(dev)
        return m, nrd.DistributedDataParallel(m, device_ids=[rank], output_device=rank)

    def snap(m):
        return (m.engine.flat_params.detach().cpu().numpy().copy(), m.engine.bn_running.detach().cpu().numpy().copy())

    # ---- fp32 eager: our SGD/CE, then the literal reference loop with stock torch SGD/CE ----
    x, y = data(B)
    for mode in ("eager", "stock"):
        m, ddp = fresh("fp32")
        if mode == "eager":
            opt = nrd.SGD(ddp.parameters(), lr=LR); loss_fn = nrd.CrossEntropyLoss()
        else:
            opt = torch.optim.SGD(ddp.parameters(), lr=LR); loss_fn = torch.nn.CrossEntropyLoss()       # main.py:27-28
        losses = []
        for s in range(2):
            o = ddp(x.to(dev)); loss = loss_fn(o, y.to(dev)); opt.zero_grad(); loss.backward()
            if s == 0 and mode == "eager":
                out["grads"] = m.engine.flat_grads.detach().cpu().numpy().copy()
                out["logits"] = o.detach().cpu().numpy().copy()
            opt.step(); losses.append(float(loss))
        out[mode] = (losses,) + snap(m)
    # ---- graph path, both exchanges, fp32 and bf16 ----
    for dtype, b in (("fp32", B), ("bf16", B_BF16)):
        xb, yb = data(b)
        for xchg in ("p2p", "nccl", "p2p-unfused") + (("p2p-stream",) if dtype == "bf16" else ()):
            os.environ["NRD_P2P"] = "0" if xchg == "nccl" else "1"
            os.environ["NRD_FUSED_UPDATE"] = "0" if xchg == "p2p-unfused" else "1"       # unfused: flags per bucket + a separate optimizer kernel
            os.environ["NRD_STREAM"] = "1" if xchg == "p2p-stream" else "0"              # the streaming persistent kernels (large per-rank batches)
            m, ddp = fresh(dtype)
            step = nrd.TrainStep(ddp, b, lr=LR if dtype == "fp32" else LR_BF16)
            losses, first = [], None
            for s in range(2):
                losses.append(float(step(xb.to(dev), yb.to(dev))))
                if s == 0:
                    first = (step.logits.detach().cpu().numpy().copy(), m.engine.flat_grads.detach().cpu().numpy().copy() / W)   # .grad = world SUM
            step.check_status()
            out[f"graph/{dtype}/{xchg}"] = (losses,) + snap(m) + first + (bool(step.p2p),)
            step.close(); del step          # graphs hold captured NCCL kernels / peer mappings: drop them before the communicator
    os.environ.pop("NRD_P2P", None); os.environ.pop("NRD_FUSED_UPDATE", None); os.environ.pop("NRD_STREAM", None)
    torch.cuda.synchronize()
    xb, yb = data(B_BF16)
    q.put((rank, out, x.numpy(), y.numpy(), xb.numpy(), yb.numpy()))     # numpy pickles inline: nothing refers back to this process
    dist.barrier()
    os._exit(0)                                  # skip communicator teardown (can block behind captured collectives)


@pytest.mark.parametrize("W", [2, 4, 8])
def test_ddp_matches_oracle_and_graph_equals_eager(W):
    if torch.cuda.device_count() < W:
        pytest.skip(f"needs {W} GPUs")
    import numpy as np
    import torch.multiprocessing as mp
    from oracle import netresdeep_oracle as O
    from util import rell2
    import ddp_cifar10_b200 as nrd
    ctx = mp.get_context("spawn"); q = ctx.SimpleQueue()
    ps = [ctx.Process(target=_worker, args=(r, W, 29561 + W, q)) for r in range(W)]
    [p.start() for p in ps]
    got = {}
    for _ in range(W):
        r, out, x, y, xb, yb = q.get(); got[r] = (out, x, y, xb, yb)
    [p.join(120) for p in ps]
    [p.kill() for p in ps if p.is_alive()]
    off = nrd._lib.param_offsets(32)
    T = torch.from_numpy

    def simulate(fn, xs, ys, steps=2, lr=LR):
        p, b0 = O.init_params(32, NB, seed=0)       # rank 0's init wins (wrapper ctor broadcast); torch.manual_seed(0) == rank 0's seed
        bufs = [dict((k, v.clone()) for k, v in b0.items()) for _ in range(W)]
        res = []
        for _ in range(steps):
            p, logits, losses, gavg = fn(p, bufs, xs, ys, NB, lr)
            res.append((p, logits, losses, gavg, [torch.cat([b["running_mean"], b["running_var"]]) for b in bufs]))
        return res

    def flat(p):
        return torch.cat([p[k].flatten() for k in O.PARAM_ORDER])

    # ---------------- fp32 ----------------
    sim = simulate(O.ddp_train_step, [T(got[r][1]) for r in range(W)], [T(got[r][2]) for r in range(W)])
    for r in range(W):
        out = got[r][0]
        assert rell2(T(out["logits"]), sim[0][1][r]) < 1e-5
        for mode in ("eager", "stock"):
            losses, w, bn = out[mode]
            for s in range(2):
                assert abs(losses[s] - float(sim[s][2][r])) < 1e-5 * float(sim[s][2][r]) * (1 if s == 0 else 20), (mode, r, s)
            assert rell2(T(w), flat(sim[1][0])) < 1e-5, (mode, r)
            assert rell2(T(bn), sim[1][4][r]) < 1e-5, (mode, r)            # each rank's own statistics on top of rank 0's buffers
            assert np.array_equal(w, got[0][0][mode][1]), (mode, r)        # every rank ends with the same weights
        for i, k in enumerate(O.PARAM_ORDER):
            tol = 2e-5 if k.startswith("fc") else 5e-3
            assert rell2(T(out["grads"][off[i]:off[i + 1]]), sim[0][3][k].flatten()) < tol, (r, k)
        # torch's foreach SGD and NCCL-mean + our fused update are the same arithmetic up to one rounding of lr * g
        assert np.allclose(out["stock"][0], out["eager"][0], rtol=1e-6) and rell2(T(out["stock"][1]), T(out["eager"][1])) < 1e-6
        for xchg in ("p2p", "nccl"):
            losses, w, bn, logits, gavg, used_p2p = out[f"graph/fp32/{xchg}"]
            assert used_p2p == (xchg == "p2p"), "peer access between the GPUs of one box is expected"
            if W == 2:                                                     # one summation order: bit for bit
                assert losses == out["eager"][0] and np.array_equal(w, out["eager"][1]) and np.array_equal(bn, out["eager"][2]), xchg
            else:
                assert np.allclose(losses, out["eager"][0], rtol=1e-6) and rell2(T(w), T(out["eager"][1])) < 1e-6 and rell2(T(bn), T(out["eager"][2])) < 1e-6
            assert np.array_equal(w, got[0][0][f"graph/fp32/{xchg}"][1]), (xchg, r)
    # ---------------- bf16: the benchmarked path ----------------
    simb = simulate(O.ddp_train_step_rm, [T(got[r][3]) for r in range(W)], [T(got[r][4]) for r in range(W)], lr=LR_BF16)
    HEAD, BNK = ("fc1.weight", "fc1.bias", "fc2.weight", "fc2.bias"), ("rb.bn.weight", "rb.bn.bias")
    for r in range(W):
        out = got[r][0]
        for xchg in ("p2p", "nccl", "p2p-stream"):
            losses, w, bn, logits, gavg, used_p2p = out[f"graph/bf16/{xchg}"]
            assert rell2(T(logits), simb[0][1][r]) < 2e-3, (xchg, r, rell2(T(logits), simb[0][1][r]))
            assert abs(losses[0] - float(simb[0][2][r])) < 2e-3 * float(simb[0][2][r])
            for i, k in enumerate(O.PARAM_ORDER):
                # small per-rank batch (40): a flipped fc1 ReLU moves the fc1 gradients by ~1/sqrt(32*40) = 3e-2, body gradients
                # carry more flip noise than at B = 256 (tests/test_gpu_parity_bench.py has the tight bars)
                tol = (4e-3 if k.startswith("fc2") else 6e-2) if k in HEAD else (2.5e-2 if k in BNK else 6e-2)
                e = rell2(T(gavg[off[i]:off[i + 1]]), simb[0][3][k].flatten())
                assert e < tol, (xchg, r, k, e)
            assert rell2(T(w), flat(simb[1][0])) < 2e-4, (xchg, r)
            assert rell2(T(bn), simb[1][4][r]) < 2e-3, (xchg, r)
            assert np.array_equal(w, got[0][0][f"graph/bf16/{xchg}"][1]), (xchg, r)
        if W == 2:
            a, b_ = out["graph/bf16/p2p"], out["graph/bf16/nccl"]
            assert a[0] == b_[0] and np.array_equal(a[1], b_[1]) and np.array_equal(a[2], b_[2])
        # the two peer-memory protocols (flag-in-data slots with the optimizer step fused into the reduce kernel; flags per bucket
        # and a separate optimizer kernel) sum the world's slots in the same rank order: bit for bit, any W, both dtypes
        for dtype in ("fp32", "bf16"):
            a, b_ = out[f"graph/{dtype}/p2p"], out[f"graph/{dtype}/p2p-unfused"]
            assert a[0] == b_[0] and np.array_equal(a[1], b_[1]) and np.array_equal(a[2], b_[2]) and np.array_equal(a[4], b_[4]), dtype
