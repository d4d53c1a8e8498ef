"""Helpers shared by the GPU parity tests."""
import ctypes as C

import torch

import ddp_cifar10_b200 as nrd
from ddp_cifar10_b200 import _lib
from ddp_cifar10_b200._lib import lib, check

DEV = "cuda:0"
TDT = {"fp32": torch.float32, "bf16": torch.bfloat16}
CODE = {"fp32": _lib.NRD_F32, "bf16": _lib.NRD_BF16}


def st():
    return torch.cuda.current_stream().cuda_stream


def nhwc(x, dt):
    """NCHW fp32 (cpu or cuda) -> NHWC `dt` on the device."""
    return x.to(DEV).permute(0, 2, 3, 1).contiguous().to(TDT[dt])


def nchw(x):
    """NHWC device tensor -> NCHW fp32 on the CPU."""
    return x.float().permute(0, 3, 1, 2).contiguous().cpu()


def relmax(a, b):
    a = a.double().cpu(); b = b.double().cpu()
    return float((a - b).abs().max() / (b.abs().max() + 1e-30))


def rell2(a, b):
    a = a.double().cpu(); b = b.double().cpu()
    return float((a - b).norm() / (b.norm() + 1e-30))


def cosine(a, b):
    a = a.double().cpu().flatten(); b = b.double().cpu().flatten()
    return float((a @ b) / (a.norm() * b.norm() + 1e-30))


def dev_f32(t):
    return t.to(DEV, torch.float32).contiguous()


ORACLE_TO_MODULE = {"conv1.weight": "conv1.weight", "conv1.bias": "conv1.bias", "rb.conv.weight": "resblocks.0.conv.weight",
                    "rb.bn.weight": "resblocks.0.batch_norm.weight", "rb.bn.bias": "resblocks.0.batch_norm.bias",
                    "fc1.weight": "fc1.weight", "fc1.bias": "fc1.bias", "fc2.weight": "fc2.weight", "fc2.bias": "fc2.bias"}


def model_from_oracle(p, bufs, n_blocks, dtype):
    """Our NetResDeep on the device carrying the oracle's parameters and BN buffers."""
    Cc = p["conv1.weight"].shape[0]
    m = nrd.NetResDeep(n_chans1=Cc, n_blocks=n_blocks, dtype=dtype)
    sd = {}
    for k, mk in ORACLE_TO_MODULE.items():
        sd[mk] = p[k].float()
    bn = "resblocks.0.batch_norm."
    sd[bn + "running_mean"] = bufs["running_mean"].float(); sd[bn + "running_var"] = bufs["running_var"].float()
    sd[bn + "num_batches_tracked"] = bufs["num_batches_tracked"]
    full = m.state_dict()
    for k in list(full):
        base = k
        for i in range(1, n_blocks):
            base = base.replace(f"resblocks.{i}.", "resblocks.0.")
        full[k] = sd[base]
    m.load_state_dict(full)
    return m.to(DEV)


def module_grads(m):
    named = dict(m.named_parameters())
    return {k: named[mk].grad.detach().float().cpu() for k, mk in ORACLE_TO_MODULE.items()}
