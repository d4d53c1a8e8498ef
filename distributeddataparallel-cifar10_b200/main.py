"""`setup` / `train_loop` / `main(rank, world_size)` with the reference's signatures (main.py:21-65,
80-84), on the B200 engine.  Differences, all forced by the box: the loader is the on-device
synthetic CIFAR-shaped generator (no dataset, no network), the rendezvous address is 127.0.0.1 and
the port comes from MASTER_PORT when set, epochs/steps are arguments with the reference's values
as defaults, and the checkpoint is written by rank 0 only (the reference lets every rank race on
one file, main.py:45)."""
import os
import time

import torch
import torch.distributed as dist
import torch.multiprocessing as mp
from torch.distributed import destroy_process_group

from . import nn as nrd_nn
from . import optim
from .data import SyntheticCifar10Loader
from .model.resnet import NetResDeep
from .parallel import DistributedDataParallel as DDP

data_path = 'data/CIFAR-10/'


def setup(rank, world_size):
    os.environ.setdefault('MASTER_ADDR', '127.0.0.1')      # reference: 'localhost' (main.py:22)
    os.environ.setdefault('MASTER_PORT', '12355')          # main.py:23
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world_size)   # main.py:24


def train_loop(model, train_loader, rank, epochs=99, lr=1e-2, save=True, log=print, graph=False):
    """`graph=True`: the same loop as ONE captured CUDA graph per step (graph.TrainStep.iterate: next batch's host-to-device
    copy beside the running step); needs a fixed batch size (drop_last)."""
    optimizer = optim.SGD(model.parameters(), lr=lr)       # main.py:27
    loss_fn = nrd_nn.CrossEntropyLoss()                    # main.py:28
    start_time = time.time()
    history = []
    step = None
    for epoch in range(1, epochs + 1):                     # main.py:30
        loss_train = torch.zeros((), device=f"cuda:{rank}")  # device-side accumulation: one sync per epoch, not per step
        if graph:
            if step is None:
                from .graph import TrainStep
                step = TrainStep(model, train_loader.batch_size, lr=lr)
            host_sum = 0.0
            for loss in step.iterate(train_loader, lookahead=1):   # main.py:31-41 (no early exit here: look-ahead is free)
                host_sum += float(loss)                    # main.py:41: the loss is already on the host
            loss_train += host_sum
        for imgs, labels in (() if graph else train_loader):
            imgs, labels = imgs.to(rank), labels.to(rank)  # main.py:33
            outputs = model(imgs)                          # main.py:34
            loss = loss_fn(outputs, labels)                # main.py:35
            optimizer.zero_grad()                          # main.py:37
            loss.backward()                                # main.py:38
            optimizer.step()                               # main.py:39
            loss_train += loss.detach()                    # main.py:41 (.item() deferred to the epoch end)
        mean_loss = loss_train.item() / len(train_loader)
        history.append(mean_loss)
        # the one host sync of the epoch has just happened: also read the kernels' sticky status (clamped / non-finite BN sums)
        getattr(model, "module", model).engine.check_status()
        if step is not None:
            step.check_status()
        if epoch == 1 or epoch % 10 == 0:                  # main.py:43
            log('Epoch {}, Training loss {}'.format(epoch, mean_loss))
            if save and rank == 0:
                os.makedirs(data_path, exist_ok=True)
                torch.save(model.module.state_dict(), data_path + 'birds_vs_airplanes.pt')   # main.py:45
    total_time = time.time() - start_time
    log(f"training time: {total_time:.3f} seconds")
    return history


def make_loader(rank, world_size, batch_size=32, steps_per_epoch=196, dataset=None):
    """main.py:53-61.  dataset=None: the on-device synthetic generator (no files on the box).  "cifar10": the reference's own
    input path -- torchvision's files under data_path, resident in HBM as uint8, DistributedSampler order, ToTensor+Normalize
    fused into the batch gather (raises like the reference's download=False when the files are missing).  "synthetic-uint8[:N]":
    the same path over a generated uint8 dataset of N images.  Or a (uint8 [N,32,32,3], labels [N]) pair."""
    from .data import Cifar10DeviceLoader, load_cifar10_python_batches, synthetic_cifar10_uint8
    if dataset is None:
        return SyntheticCifar10Loader(batch_size=batch_size, steps_per_epoch=steps_per_epoch, device=f"cuda:{rank}", rank=rank)
    if isinstance(dataset, str):
        if dataset == "cifar10":
            data, labels = load_cifar10_python_batches(data_path)                  # main.py:54 (download=False)
        elif dataset.startswith("synthetic-uint8"):
            n = int(dataset.split(":")[1]) if ":" in dataset else 50000
            data, labels = synthetic_cifar10_uint8(n)
        else:
            raise ValueError(f"unknown dataset {dataset!r}")
    else:
        data, labels = dataset
    return Cifar10DeviceLoader(data, labels, batch_size=batch_size, num_replicas=world_size, rank=rank, device=f"cuda:{rank}")   # main.py:60-61


def main(rank, world_size, epochs=99, batch_size=32, steps_per_epoch=196, dtype=None, save=True, dataset=None, graph=False):
    setup(rank, world_size)
    train_loader = make_loader(rank, world_size, batch_size, steps_per_epoch, dataset)   # main.py:53-61
    model = NetResDeep(dtype=dtype).to(rank)               # main.py:62
    model = DDP(model, device_ids=[rank], output_device=rank)   # main.py:63
    train_loop(model, train_loader, rank, epochs=epochs, save=save, graph=graph)
    destroy_process_group()                                # main.py:65


if __name__ == '__main__':
    if torch.cuda.device_count() > 1:
        print("Using", torch.cuda.device_count(), "GPUs")
    world_size = torch.cuda.device_count()
    mp.spawn(main, args=(world_size,), nprocs=world_size)  # main.py:84
