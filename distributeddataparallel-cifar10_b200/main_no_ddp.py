"""Single-process loop with the reference's shape (main_no_ddp.py:36-59) on the B200 engine.
Unlike the reference file nothing runs at import time."""
import time

import torch

from . import nn as nrd_nn
from . import optim
from .data import SyntheticCifar10Loader
from .model.resnet import NetResDeep


def prepare(batch_size=32, steps_per_epoch=782, device="cuda"):
    # the reference hard-codes batch 64 here whatever the argument says (main_no_ddp.py:31)
    return SyntheticCifar10Loader(batch_size=64, steps_per_epoch=steps_per_epoch, device=device)


def training_loop(model, train_loader, epochs=99, lr=1e-2, log=print):
    device = next(model.parameters()).device
    optimizer = optim.SGD(model.parameters(), lr=lr)       # main_no_ddp.py:37
    loss_fn = nrd_nn.CrossEntropyLoss()                    # main_no_ddp.py:38
    start_time = time.time()
    history = []
    for epoch in range(1, epochs + 1):
        loss_train = torch.zeros((), device=device)
        for imgs, labels in train_loader:
            imgs = imgs.to(device=device); labels = labels.to(device=device)
            outputs = model(imgs)                          # :45
            loss = loss_fn(outputs, labels)                # :46
            optimizer.zero_grad()                          # :48
            loss.backward()                                # :49
            optimizer.step()                               # :50
            loss_train += loss.detach()
        history.append(loss_train.item() / len(train_loader))
        model.engine.check_status()
        if epoch == 1 or epoch % 10 == 0:
            log('Epoch {}, Training loss {}'.format(epoch, history[-1]))
    log(f"training time: {time.time() - start_time:.3f} seconds")
    return history


if __name__ == '__main__':
    model = NetResDeep().to("cuda")
    training_loop(model, prepare())
