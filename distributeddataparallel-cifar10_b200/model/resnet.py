"""Drop-in `NetResDeep` / `ResBlock` (reference: model/resnet.py:5-37) running on libnrd.so.

Same constructor, attributes, parameter names/shapes/order, initialisation and `state_dict()`
(66 keys, the `n_blocks` entries of `resblocks` alias ONE block -- model/resnet.py:10-11) as the
reference; `forward` launches the hand-written sm_100a kernels instead of ATen ops.
"""
import torch
import torch.nn as nn

from ..engine import Engine
from ..functional import netresdeep_forward, resblock_forward


class NetResDeep(nn.Module):
    def __init__(self, n_chans1=32, n_blocks=10, dtype=None):
        """`dtype` (extension): "bf16" (default; tcgen05 tensor cores, fp32 accumulate) or "fp32"
        (validation mode, CUDA cores).  Env NRD_DTYPE overrides the default."""
        super().__init__()
        self.n_chans1 = n_chans1
        self.conv1 = nn.Conv2d(3, n_chans1, kernel_size=3, padding=1)           # resnet.py:9
        self.resblocks = nn.Sequential(*(n_blocks * [ResBlock(n_chans=n_chans1)]))  # resnet.py:10-11 (shared block)
        self.fc1 = nn.Linear(8 * 8 * n_chans1, 32)                              # resnet.py:12
        self.fc2 = nn.Linear(32, 10)                                            # resnet.py:13
        object.__setattr__(self, "_engine", Engine(self, dtype))                # not a submodule / not in state_dict
        for p in self.parameters():
            p._nrd_engine = self._engine                                        # lets optim.SGD find the flat vectors

    @property
    def engine(self) -> Engine:
        return self._engine

    def forward(self, x):
        """model/resnet.py:15-22.  x: fp32 NCHW [B,3,32,32] on the model's CUDA device -> logits [B,10]."""
        return netresdeep_forward(self, x)


class ResBlock(nn.Module):
    def __init__(self, n_chans):
        super().__init__()
        self.conv = nn.Conv2d(n_chans, n_chans, kernel_size=3, padding=1, bias=False)   # resnet.py:27
        self.batch_norm = nn.BatchNorm2d(num_features=n_chans)                          # resnet.py:28
        torch.nn.init.kaiming_normal_(self.conv.weight, nonlinearity='relu')            # resnet.py:29
        torch.nn.init.constant_(self.batch_norm.weight, 0.5)                            # resnet.py:30
        torch.nn.init.zeros_(self.batch_norm.bias)                                      # resnet.py:31

    def forward(self, x):
        """model/resnet.py:33-37 on its own: relu(bn(conv(x))) + x, x: fp32 NCHW [B,C,16,16]."""
        return resblock_forward(self, x)
