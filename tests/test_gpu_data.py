"""The reference's real input path on the device (main.py:53-61): uint8 dataset resident in HBM, ToTensor +
Normalize fused into the batch gather, DistributedSampler order, batch_size 32 with the ragged last batch of drop_last=False.
Oracle: torchvision's own transforms + torch's DataLoader / DistributedSampler on the same synthetic uint8 data (no CIFAR files
exist on the box), and the CPU oracle's training step on those batches."""
import numpy as np
import pytest
import torch

from oracle import netresdeep_oracle as O
from util import *  # noqa: F401,F403
from ddp_cifar10_b200.data import Cifar10DeviceLoader, synthetic_cifar10_uint8, CIFAR10_MEAN, CIFAR10_STD

pytestmark = pytest.mark.gpu


class _RefDataset(torch.utils.data.Dataset):
    """what torchvision's CIFAR10 hands the DataLoader: transform(HWC uint8 image), label (main.py:54-58)"""

    def __init__(self, data, labels):
        import torchvision.transforms as T
        self.data, self.labels = data.numpy(), labels
        self.tf = T.Compose([T.ToTensor(), T.Normalize(CIFAR10_MEAN, CIFAR10_STD)])

    def __len__(self):
        return len(self.labels)

    def __getitem__(self, i):
        return self.tf(self.data[i]), int(self.labels[i])


@pytest.mark.parametrize("W,rank", [(1, 0), (2, 1), (8, 5)])
def test_device_loader_equals_torch_dataloader_bitwise(W, rank):
    from torch.utils.data import DataLoader, DistributedSampler
    data, labels = synthetic_cifar10_uint8(2 * 330 + 7, seed=3)
    ref = _RefDataset(data, labels)
    ref_loader = DataLoader(ref, batch_size=32, sampler=DistributedSampler(ref, num_replicas=W, rank=rank))     # main.py:60-61
    ours = Cifar10DeviceLoader(data, labels, batch_size=32, num_replicas=W, rank=rank, device=DEV)
    assert len(ours) == len(ref_loader)
    sizes = []
    for (xi, yi), (xr, yr) in zip(ours, ref_loader):
        assert torch.equal(xi.cpu(), xr) and torch.equal(yi.cpu(), yr)          # same samples, same order, same bits
        sizes.append(xi.shape[0])
    assert len(sizes) == len(ref_loader) and sizes[-1] == (len(ours.index_list) - 32 * (len(ours) - 1))
    # a second epoch is the same epoch (no set_epoch in the reference)
    first = next(iter(ours)); again = next(iter(ours))
    assert torch.equal(first[0], again[0])


def test_ragged_last_batch_of_ten_step_vs_oracle():
    """BN over a batch of 10 (the reference's last batch at 8 ranks), fp32, against the fp64 oracle."""
    B, NB = 10, 10
    g = torch.Generator().manual_seed(12)
    x = torch.randn(B, 3, 32, 32, generator=g); y = torch.randint(0, 10, (B,), generator=g)
    p64, b64 = O.init_params(32, NB, seed=2, dtype=torch.float64)
    _, lref, lossref, gref = O.train_step(p64, b64, x.double(), y, NB)
    pm, bm = O.init_params(32, NB, seed=2)
    m = model_from_oracle(pm, bm, NB, "fp32")
    opt = nrd.SGD(m.parameters(), lr=1e-2); loss_fn = nrd.CrossEntropyLoss()
    out = m(x.to(DEV)); loss = loss_fn(out, y.to(DEV)); opt.zero_grad(); loss.backward()
    assert rell2(out.detach(), lref) < 1e-5 and abs(float(loss) - float(lossref)) < 1e-5 * float(lossref)
    gr = module_grads(m)
    for k in ("fc1.weight", "fc1.bias", "fc2.weight", "fc2.bias"):
        assert rell2(gr[k], gref[k]) < 2e-5, k
    bn = m.resblocks[0].batch_norm
    assert rell2(bn.running_var, b64["running_var"]) < 1e-5
    # and in bf16 through the graph path (a TrainStep of size 10)
    pm, bm = O.init_params(32, NB, seed=2)
    pr, br = O.init_params(32, NB, seed=2)
    _, l_rm, loss_rm, _ = O.train_step_rm(pr, br, x, y, NB)
    mb = model_from_oracle(pm, bm, NB, "bf16")
    step = nrd.TrainStep(mb, B, lr=1e-2)
    lb = float(step(x.to(DEV), y.to(DEV)))
    assert rell2(step.logits, l_rm) < 5e-3 and abs(lb - float(loss_rm)) < 5e-3 * float(loss_rm)


@pytest.mark.parametrize("graph", [False, True])
def test_epoch_over_the_device_loader_tracks_the_cpu_reference(graph):
    """main.train_loop over the device loader (10 full batches + a ragged one of 10), fp32, two epochs: the per-epoch mean
    losses against the CPU oracle stepping through torch's own DataLoader order on the same data."""
    from torch.utils.data import DataLoader, DistributedSampler
    data, labels = synthetic_cifar10_uint8(330, seed=5)
    ref = _RefDataset(data, labels)
    ref_loader = DataLoader(ref, batch_size=32, sampler=DistributedSampler(ref, num_replicas=1, rank=0))
    p, bufs = O.init_params(32, 10, seed=0, dtype=torch.float64)
    want = []
    for _ in range(2):
        tot = 0.0
        for xr, yr in ref_loader:
            p, _, l, _ = O.train_step(p, bufs, xr.double(), yr, 10, 1e-3)
            tot += float(l)
        want.append(tot / len(ref_loader))
    pm, bm = O.init_params(32, 10, seed=0)
    m = model_from_oracle(pm, bm, 10, "fp32")
    loader = Cifar10DeviceLoader(data, labels, batch_size=32, device=DEV)
    assert len(loader) == 11
    hist = nrd.main.train_loop(m, loader, 0, epochs=2, lr=1e-3, save=False, log=lambda *_: None, graph=graph)
    assert np.allclose(hist, want, rtol=2e-4), (hist, want)
    assert int(m.resblocks[0].batch_norm.num_batches_tracked) == 2 * 11 * 10
