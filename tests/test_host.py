"""CPU-only tests: the C-ABI library loads and exports what include/nrd.h declares, the drop-in
surface matches the reference's, and the multi-rank host logic works over gloo (world_size 2)."""
import os
import re

import numpy as np
import pytest
import torch

from conftest import GOLDEN, ROOT
import ddp_cifar10_b200 as nrd
from ddp_cifar10_b200 import _lib
from ddp_cifar10_b200._lib import lib


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "nrd.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(nrd_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 25
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/nrd.h but not exported by libnrd.so"
    assert declared == set(_lib.EXPORTS), declared ^ set(_lib.EXPORTS)
    assert lib.nrd_version() >= 100


def test_option_numbers_match_the_header():
    """`_lib.OPT_*` (what engine.py / graph.py pass to nrd_plan_set_option) against the `#define NRD_OPT_*` values of include/nrd.h."""
    hdr = open(os.path.join(ROOT, "include", "nrd.h")).read()
    defs = {k: int(v) for k, v in re.findall(r"#define\s+NRD_OPT_([A-Z_0-9]+)\s+(\d+)", hdr)}
    assert len(defs) >= 14 and sorted(defs.values()) == list(range(len(defs))), defs
    for name, val in defs.items():
        assert getattr(_lib, "OPT_" + name) == val, name


def test_param_layout_and_workspace_sizes():
    off = _lib.param_offsets(32)
    assert off[-1] == 76074 and off[5] == 10176            # SURVEY fact 2; body bucket = 40,704 B
    assert _lib.param_offsets(64)[-1] == 170218
    a = lib.nrd_workspace_bytes(64, 32, 10, _lib.NRD_F32); b = lib.nrd_workspace_bytes(64, 32, 10, _lib.NRD_BF16)
    assert 0 < b < a and lib.nrd_workspace_bytes(256, 32, 10, _lib.NRD_BF16) > b
    assert lib.nrd_workspace_bytes(64, 48, 10, 0) == 0 and "32 or 64" in _lib.last_error()
    assert lib.nrd_workspace_bytes(0, 32, 10, 0) == 0 and lib.nrd_workspace_bytes(8, 32, 40, 0) == 0
    assert lib.nrd_workspace_bytes(8, 32, 10, 7) == 0


def test_dropin_module_surface_matches_reference():
    g = np.load(os.path.join(GOLDEN, "step_c32_nb10_b8.npz"))
    torch.manual_seed(0)
    m = nrd.NetResDeep()
    assert list(m.state_dict().keys()) == list(g["state_dict_keys"])          # 66 keys, reference order
    assert [n for n, _ in m.named_parameters()] == list(nrd.engine.PARAM_NAMES)
    assert all(b is m.resblocks[0] for b in m.resblocks) and len(m.resblocks) == 10
    assert m.n_chans1 == 32 and m.fc1.in_features == 2048
    # same seed -> bitwise the reference's initial weights (checksums recorded from the reference)
    named = dict(m.named_parameters())
    from util import ORACLE_TO_MODULE
    for k, mk in ORACLE_TO_MODULE.items():
        assert float(named[mk].double().sum()) == float(g[f"w0_sum/{k}"]), k
    wide = nrd.NetResDeep(n_chans1=64, n_blocks=20)
    assert sum(p.numel() for p in wide.parameters()) == 170218


def test_cpu_has_no_fallback_and_arguments_are_validated():
    m = nrd.NetResDeep()
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        m(torch.zeros(2, 3, 32, 32))
    with pytest.raises(NotImplementedError):
        nrd.SGD(m.parameters(), lr=0.1, momentum=0.9, nesterov=True)
    with pytest.raises(NotImplementedError):
        nrd.SGD(m.parameters(), lr=0.1, weight_decay=1e-4)
    with pytest.raises(NotImplementedError):
        nrd.CrossEntropyLoss(label_smoothing=0.1)
    with pytest.raises(ValueError):
        nrd.NetResDeep(dtype="fp8").engine
    opt = nrd.SGD(m.parameters(), lr=1e-2)
    assert opt.step() is None                     # no grads yet: a no-op like torch's SGD


def test_host_generator_restatement():
    from ddp_cifar10_b200.data import host_batch
    a, la = host_batch(8, 1234, 0, 0); b, lb = host_batch(8, 1234, 0, 0); c, _ = host_batch(8, 1234, 1, 0)
    assert np.array_equal(a, b) and np.array_equal(la, lb) and not np.array_equal(a, c)
    big, lab = host_batch(64, 1234, 0, 3)
    assert abs(big.mean()) < 0.01 and abs(big.std() - 1) < 0.01 and set(np.unique(lab)) <= set(range(10))


def _gloo_worker(rank, W, port, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=W)
    from ddp_cifar10_b200.parallel import flat_allreduce_mean_cpu
    n_body = 10176
    flat = torch.arange(76074, dtype=torch.float32) * (rank + 1)
    flat_allreduce_mean_cpu(flat, n_body)
    bufs = torch.full((64,), float(rank))
    dist.broadcast(bufs, src=0)                    # C2: rank 0's BN buffers win
    # the graph path's folded buffer sync: [2C slot | body bucket] as one SUM all-reduce, rank 0 contributes W * stats
    torch.manual_seed(100 + rank)
    running = torch.rand(64) * 3 + 1e-3
    xchg = torch.zeros(64 + n_body)
    xchg[:64] = running * float(W) if rank == 0 else 0.0
    xchg[64:] = rank + 1.0
    dist.all_reduce(xchg, op=dist.ReduceOp.SUM)
    folded = xchg[:64] * (1.0 / W)
    ref = running.clone(); dist.broadcast(ref, src=0)
    assert torch.equal(folded, ref) and float(xchg[64]) == 3.0      # bit-exact, body bucket untouched by the slot
    q.put((rank, flat[[0, 10175, 10176, 76073]].tolist(), bufs.sum().item()))
    dist.barrier(); dist.destroy_process_group()


def test_bucketed_mean_allreduce_world_size_2_gloo():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn"); q = ctx.SimpleQueue()
    ps = [ctx.Process(target=_gloo_worker, args=(r, 2, 29541, q)) for r in range(2)]
    [p.start() for p in ps]
    got = dict((r, (v, s)) for r, v, s in (q.get() for _ in range(2)))
    [p.join() for p in ps]
    for r in range(2):
        assert got[r][0] == [0.0, 10175 * 1.5, 10176 * 1.5, 76073 * 1.5] and got[r][1] == 0.0


def test_sampler_order_equals_torch_distributed_sampler():
    """The device loader's sample order is torch's DistributedSampler order (main.py:60), the same in every epoch (the reference
    never calls set_epoch), with the reference's batch boundaries: 196 batches of 32 at 8 ranks, the last one of 10."""
    from torch.utils.data import DistributedSampler
    from ddp_cifar10_b200.data import distributed_sampler_indices

    class D:
        def __init__(self, n): self.n = n
        def __len__(self): return self.n

    for n in (50000, 1003, 7):
        for W in (1, 2, 8):
            for r in range(W):
                smp = DistributedSampler(D(n), num_replicas=W, rank=r)
                e0 = list(iter(smp)); e1 = list(iter(smp))                     # second epoch without set_epoch
                ours = distributed_sampler_indices(n, W, r)
                assert ours == e0 == e1, (n, W, r)
    per_rank = len(distributed_sampler_indices(50000, 8, 3))
    assert per_rank == 6250 and (per_rank + 31) // 32 == 196 and per_rank - 195 * 32 == 10
    assert (len(distributed_sampler_indices(50000, 1, 0)) + 31) // 32 == 1563
