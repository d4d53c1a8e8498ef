"""On-device synthetic CIFAR-10-shaped loader (replaces torchvision CIFAR10 + DataLoader of
main.py:53-61 for benchmarking: there is no dataset on the box).  Iterating yields
`(imgs fp32 [B,3,32,32], labels int64 [B])` already on the device, so `.to(rank)` is free."""
import numpy as np
import torch

from ._lib import lib, check
from .engine import _stream_ptr


class SyntheticCifar10Loader:
    def __init__(self, batch_size=32, steps_per_epoch=196, device="cuda", seed=1234, rank=0):
        self.batch_size, self.steps, self.seed, self.rank = batch_size, steps_per_epoch, seed, rank
        self.device = torch.device(device)
        self._step = 0

    def __len__(self):
        return self.steps

    def batch(self, step: int):
        imgs = torch.empty(self.batch_size, 3, 32, 32, dtype=torch.float32, device=self.device)
        labels = torch.empty(self.batch_size, dtype=torch.int64, device=self.device)
        with torch.cuda.device(self.device):
            check(lib.nrd_gen_batch(imgs.data_ptr(), labels.data_ptr(), self.batch_size, self.seed, self.rank, step, _stream_ptr()),
                  "nrd_gen_batch")
        return imgs, labels

    def __iter__(self):
        for _ in range(self.steps):
            yield self.batch(self._step)
            self._step += 1


# ---- the reference's real input path (main.py:53-61) on the device ---------------------------------------------------------
CIFAR10_MEAN = (0.4915, 0.4823, 0.4468)        # main.py:56
CIFAR10_STD = (0.2470, 0.2435, 0.2616)         # main.py:57


def distributed_sampler_indices(n: int, num_replicas: int, rank: int, seed: int = 0, epoch: int = 0, shuffle: bool = True,
                                drop_last: bool = False):
    """The index list `DistributedSampler(dataset_of_n, num_replicas, rank)` yields (main.py:60): `randperm(n)` from a CPU
    generator seeded with seed + epoch, wrap-padded to a multiple of num_replicas (or truncated with drop_last), strided
    [rank::num_replicas].  The reference never calls `set_epoch`, so it is the epoch-0 list in every epoch.  Same generator,
    same calls as torch's sampler: identical by construction (tests/test_host.py compares the two)."""
    import math
    if shuffle:
        g = torch.Generator()
        g.manual_seed(seed + epoch)
        indices = torch.randperm(n, generator=g).tolist()
    else:
        indices = list(range(n))
    if drop_last and n % num_replicas != 0:
        num_samples = math.ceil((n - num_replicas) / num_replicas)
    else:
        num_samples = math.ceil(n / num_replicas)
    total = num_samples * num_replicas
    if not drop_last:
        pad = total - len(indices)
        if pad <= len(indices):
            indices += indices[:pad]
        else:
            indices += (indices * math.ceil(pad / len(indices)))[:pad]
    else:
        indices = indices[:total]
    return indices[rank:total:num_replicas]


def load_cifar10_python_batches(root: str):
    """(uint8 [50000,32,32,3], int64 [50000]) from torchvision's on-disk layout `root/cifar-10-batches-py/data_batch_1..5`
    (what `datasets.CIFAR10(data_path, train=True)` reads at main.py:54), without torchvision."""
    import os
    import pickle
    d = os.path.join(root, "cifar-10-batches-py")
    xs, ys = [], []
    for i in range(1, 6):
        with open(os.path.join(d, f"data_batch_{i}"), "rb") as f:
            e = pickle.load(f, encoding="latin1")
        xs.append(np.asarray(e["data"], dtype=np.uint8).reshape(-1, 3, 32, 32).transpose(0, 2, 3, 1))
        ys.append(np.asarray(e.get("labels", e.get("fine_labels")), dtype=np.int64))
    return torch.from_numpy(np.ascontiguousarray(np.concatenate(xs))), torch.from_numpy(np.concatenate(ys))


def synthetic_cifar10_uint8(n: int = 50000, seed: int = 1234):
    """A CIFAR-10-shaped uint8 dataset for boxes without the files: class templates + noise, so that training on it learns."""
    g = torch.Generator().manual_seed(seed)
    labels = torch.randint(0, 10, (n,), generator=g)
    templates = torch.rand(10, 32, 32, 3, generator=g)
    imgs = (templates[labels] * 0.6 + torch.rand(n, 32, 32, 3, generator=g) * 0.4) * 255.0
    return imgs.round().clamp_(0, 255).to(torch.uint8), labels


class Cifar10DeviceLoader:
    """`DataLoader(cifar10, batch_size=32, sampler=DistributedSampler(cifar10, num_replicas, rank))` of main.py:60-61 with the
    dataset resident in HBM: uint8 [N,32,32,3] + labels; every batch is ONE kernel (gather by the sampler's indices + ToTensor +
    Normalize, nrd_gather_normalize) writing the fp32 NCHW batch the model takes.  Same sample order as the reference in every
    epoch (it never calls set_epoch), `drop_last=False` as there: the last batch is ragged (10 samples at 8 ranks).

    Iterating yields `(imgs fp32 [b,3,32,32], labels int64 [b])` on the device, so `.to(rank)` (main.py:33) is free."""

    def __init__(self, data_u8: torch.Tensor, labels: torch.Tensor, batch_size: int = 32, num_replicas: int = 1, rank: int = 0,
                 device="cuda", shuffle: bool = True, seed: int = 0, drop_last: bool = False,
                 mean=CIFAR10_MEAN, std=CIFAR10_STD):
        if data_u8.dtype != torch.uint8 or data_u8.dim() != 4 or tuple(data_u8.shape[1:]) != (32, 32, 3):
            raise ValueError(f"expected a uint8 [N,32,32,3] dataset (torchvision's CIFAR10.data), got {data_u8.dtype} {tuple(data_u8.shape)}")
        if labels.shape[0] != data_u8.shape[0]:
            raise ValueError("one label per image")
        self.device = torch.device(device)
        self.data = data_u8.contiguous().to(self.device)
        self.labels = labels.to(torch.int64).contiguous().to(self.device)
        self.batch_size, self.drop_last = batch_size, drop_last
        self.n = int(data_u8.shape[0])
        self.index_list = distributed_sampler_indices(self.n, num_replicas, rank, seed=seed, epoch=0, shuffle=shuffle)
        self.indices = torch.tensor(self.index_list, dtype=torch.int64, device=self.device)
        import ctypes as C
        self._mean = (C.c_float * 3)(*mean); self._std = (C.c_float * 3)(*std)

    def __len__(self):
        m = len(self.index_list)
        return m // self.batch_size if self.drop_last else (m + self.batch_size - 1) // self.batch_size

    def batch(self, i: int):
        lo = i * self.batch_size
        hi = min(lo + self.batch_size, len(self.index_list))
        b = hi - lo
        imgs = torch.empty(b, 3, 32, 32, dtype=torch.float32, device=self.device)
        labels = torch.empty(b, dtype=torch.int64, device=self.device)
        with torch.cuda.device(self.device):
            check(lib.nrd_gather_normalize(self.data.data_ptr(), self.labels.data_ptr(), self.n, self.indices.data_ptr() + 8 * lo, b,
                                           self._mean, self._std, imgs.data_ptr(), labels.data_ptr(), _stream_ptr()), "nrd_gather_normalize")
        return imgs, labels

    def __iter__(self):
        for i in range(len(self)):
            yield self.batch(i)


class DevicePrefetcher:
    """Wraps a loader that yields HOST batches `(imgs, labels)` (main.py:61's DataLoader; pinned memory makes the copies
    asynchronous) and yields them on the device, with batch t+1's host-to-device copy issued on a copy stream BEFORE batch
    t is handed out -- so it runs beside step t's kernels.  The training loop keeps the reference's shape (main.py:31-33:
    `for imgs, labels in loader: imgs, labels = imgs.to(rank), labels.to(rank)`; `.to` of a resident tensor is free).

    Two device slots are reused in turn: a yielded batch stays valid until the NEXT-BUT-ONE `next()`; work that reads it
    must have been enqueued on the current stream by then (a training loop does exactly that)."""

    def __init__(self, loader, device, slots: int = 2):
        self.loader, self.device, self.n = loader, torch.device(device), max(2, slots)
        self.copy_stream = torch.cuda.Stream(device=self.device)
        self.bufs = [None] * self.n
        self.ready = [torch.cuda.Event() for _ in range(self.n)]

    def __len__(self):
        return len(self.loader)

    def _issue(self, slot, batch):
        imgs, labels = batch
        if self.bufs[slot] is None or self.bufs[slot][0].shape != imgs.shape:
            self.bufs[slot] = (torch.empty(imgs.shape, dtype=imgs.dtype, device=self.device),
                               torch.empty(labels.shape, dtype=labels.dtype, device=self.device))
        # the slot's previous contents were handed out two batches ago: everything enqueued on the consumer's stream so
        # far (its readers included) goes first
        self.copy_stream.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(self.copy_stream):
            self.bufs[slot][0].copy_(imgs, non_blocking=True)
            self.bufs[slot][1].copy_(labels, non_blocking=True)
            self.ready[slot].record(self.copy_stream)

    def __iter__(self):
        it = iter(self.loader)
        try:
            nxt = next(it)
        except StopIteration:
            return
        slot = 0
        self._issue(slot, nxt)
        while True:
            cur = slot
            try:
                nxt = next(it)
                slot = (slot + 1) % self.n
                self._issue(slot, nxt)
            except StopIteration:
                nxt = None
            torch.cuda.current_stream(self.device).wait_event(self.ready[cur])
            yield self.bufs[cur]
            if nxt is None:
                return


# ---- host restatement of the generator (tests compare the device batch with it) -------------
def _philox4x32_10(c0, c1, c2, c3, k0, k1):
    M0, M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
    c0, c1, c2, c3 = (np.asarray(v, dtype=np.uint64) for v in (c0, c1, c2, c3))
    k0, k1 = np.uint64(k0), np.uint64(k1)
    mask = np.uint64(0xFFFFFFFF)
    for _ in range(10):
        p0, p1 = M0 * c0, M1 * c2
        hi0, lo0, hi1, lo1 = p0 >> np.uint64(32), p0 & mask, p1 >> np.uint64(32), p1 & mask
        c0, c1, c2, c3 = (hi1 ^ c1 ^ k0) & mask, lo1, (hi0 ^ c3 ^ k1) & mask, lo0
        k0, k1 = (k0 + np.uint64(0x9E3779B9)) & mask, (k1 + np.uint64(0xBB67AE85)) & mask
    return c0, c1, c2, c3


def host_batch(batch_size, seed=1234, rank=0, step=0):
    """numpy restatement of k_gen_batch: same counters, same Box-Muller."""
    n4 = batch_size * 768
    k0, k1 = seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF
    i = np.arange(n4, dtype=np.uint64)
    z = np.zeros_like(i)
    r = _philox4x32_10(i, z + np.uint64(rank), z + np.uint64(step & 0xFFFFFFFF), z + np.uint64(step >> 32), k0, k1)
    u = [(x >> np.uint64(8)).astype(np.float32) for x in r]
    u0, u1 = (u[0] + np.float32(1)) * np.float32(1 / 16777216), u[1] * np.float32(1 / 16777216)
    u2, u3 = (u[2] + np.float32(1)) * np.float32(1 / 16777216), u[3] * np.float32(1 / 16777216)
    ra, rb = np.sqrt(np.float32(-2) * np.log(u0)), np.sqrt(np.float32(-2) * np.log(u2))
    a0, a1 = np.float32(2 * np.pi) * u1, np.float32(2 * np.pi) * u3
    img = np.stack([ra * np.cos(a0), ra * np.sin(a0), rb * np.cos(a1), rb * np.sin(a1)], axis=1).astype(np.float32)
    b = np.arange(batch_size, dtype=np.uint64)
    zb = np.zeros_like(b)
    rl = _philox4x32_10(b, zb + np.uint64(rank | 0x80000000), zb + np.uint64(step & 0xFFFFFFFF), zb + np.uint64(step >> 32), k0, k1)
    return img.reshape(batch_size, 3, 32, 32), (rl[0] % np.uint64(10)).astype(np.int64)
