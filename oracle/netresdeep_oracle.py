"""CPU oracle for the NetResDeep data-parallel training step.  TEST INFRASTRUCTURE ONLY.

This file restates, on the CPU, the algorithm of the reference hot path so the CUDA
path can be checked against it.  Only `tests/`, `__graft_entry__.smoke()` and the
`cpu_baseline` / `--impl reference` legs of `bench.py` may import it; nothing in the
product package does (the product fails loudly when its CUDA library is missing).

Where the arithmetic lives: the reference is 4 Python files that dispatch every
operation to PyTorch (un-pinned third-party dependency; this image: torch 2.11.0+cu128,
oneDNN 3.10.2).  The oracle therefore restates the *model and step* -- explicit forward
AND explicit closed-form backward, no autograd -- with torch CPU tensor ops in a chosen
dtype (fp32 or fp64), each function citing the reference line it follows.

Pinning: the reference has no tests and no golden vectors.  The
oracle is pinned instead against outputs of the reference itself:
`oracle/make_golden.py` imports the unmodified `/root/reference/model/resnet.py`,
runs the step of `main_no_ddp.py:45-50` under torch autograd and writes
`tests/golden/*.npz`; `tests/test_oracle.py` checks this file against those fixtures
(and, when `/root/reference` is present, against the live reference).

Layout conventions here are the reference's: activations NCHW, conv weights OIHW,
fc1 columns in (c, h, w) order.
"""
from __future__ import annotations

import math
from typing import Dict, List, Tuple

import torch
import torch.nn.functional as F

BN_EPS = 1e-5        # nn.BatchNorm2d default, model/resnet.py:28
BN_MOMENTUM = 0.1    # nn.BatchNorm2d default, model/resnet.py:28

PARAM_ORDER = ("conv1.weight", "conv1.bias", "rb.conv.weight", "rb.bn.weight", "rb.bn.bias",
               "fc1.weight", "fc1.bias", "fc2.weight", "fc2.bias")


# --------------------------------------------------------------------------- init
def init_params(n_chans1: int = 32, n_blocks: int = 10, seed: int | None = 0,
                dtype=torch.float32) -> Tuple[Dict[str, torch.Tensor], Dict[str, torch.Tensor]]:
    """Parameters and BN buffers initialised as `NetResDeep.__init__` does.

    Draw order follows module construction order (model/resnet.py:9-13, 27-31):
    conv1 (Conv2d default: kaiming_uniform(a=sqrt 5) weight, U(+-1/sqrt(fan_in)) bias),
    the ONE shared ResBlock (conv default init, then overwritten by kaiming_normal_(relu),
    gamma=0.5, beta=0), fc1, fc2 (Linear default init).  With the same torch seed this
    reproduces the reference's weights bit for bit (checked in tests/test_oracle.py).
    """
    if seed is not None:
        torch.manual_seed(seed)
    C = n_chans1

    def default_wb(shape_w, fan_in, bias=True):
        w = torch.empty(shape_w)
        torch.nn.init.kaiming_uniform_(w, a=math.sqrt(5))
        b = None
        if bias:
            bound = 1.0 / math.sqrt(fan_in)
            b = torch.empty(shape_w[0])
            torch.nn.init.uniform_(b, -bound, bound)
        return w, b

    p: Dict[str, torch.Tensor] = {}
    p["conv1.weight"], p["conv1.bias"] = default_wb((C, 3, 3, 3), 27)          # resnet.py:9
    w, _ = default_wb((C, C, 3, 3), 9 * C, bias=False)                          # resnet.py:27
    torch.nn.init.kaiming_normal_(w, nonlinearity="relu")                       # resnet.py:29
    p["rb.conv.weight"] = w
    p["rb.bn.weight"] = torch.full((C,), 0.5)                                   # resnet.py:30
    p["rb.bn.bias"] = torch.zeros(C)                                            # resnet.py:31
    p["fc1.weight"], p["fc1.bias"] = default_wb((32, 64 * C), 64 * C)           # resnet.py:12
    p["fc2.weight"], p["fc2.bias"] = default_wb((10, 32), 32)                   # resnet.py:13
    p = {k: v.to(dtype) for k, v in p.items()}
    bufs = {"running_mean": torch.zeros(C, dtype=dtype), "running_var": torch.ones(C, dtype=dtype),
            "num_batches_tracked": torch.zeros((), dtype=torch.int64)}
    return p, bufs


# --------------------------------------------------------------------------- op pieces
def maxpool2_fwd(x: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """2x2/2 max pool with the index (0..3, row-major in the window) of the FIRST max,
    which is where `F.max_pool2d` (model/resnet.py:16,18) routes the gradient."""
    B, C, H, W = x.shape
    win = x.reshape(B, C, H // 2, 2, W // 2, 2).permute(0, 1, 2, 4, 3, 5).reshape(B, C, H // 2, W // 2, 4)
    val, idx = win.max(dim=-1)
    # torch.max returns *an* index of the max; force first-max semantics explicitly
    first = (win == val.unsqueeze(-1)).to(torch.int8).argmax(dim=-1)
    return val, first.to(torch.uint8)


def maxpool2_bwd(g: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
    B, C, H2, W2 = g.shape
    out = torch.zeros(B, C, H2, W2, 4, dtype=g.dtype)
    out.scatter_(-1, idx.long().unsqueeze(-1), g.unsqueeze(-1))
    return out.reshape(B, C, H2, W2, 2, 2).permute(0, 1, 2, 4, 3, 5).reshape(B, C, H2 * 2, W2 * 2)


def bn_train_fwd(y, gamma, beta, bufs):
    """Training-mode BatchNorm2d (model/resnet.py:35): biased variance normalises, eps
    inside the sqrt; running_var takes the UNBIASED variance; nbt += 1."""
    n = y.shape[0] * y.shape[2] * y.shape[3]
    mean = y.mean(dim=(0, 2, 3))
    var = ((y - mean[None, :, None, None]) ** 2).mean(dim=(0, 2, 3))
    invstd = 1.0 / torch.sqrt(var + BN_EPS)
    xhat = (y - mean[None, :, None, None]) * invstd[None, :, None, None]
    z = xhat * gamma[None, :, None, None] + beta[None, :, None, None]
    bufs["running_mean"] = (1 - BN_MOMENTUM) * bufs["running_mean"] + BN_MOMENTUM * mean
    bufs["running_var"] = (1 - BN_MOMENTUM) * bufs["running_var"] + BN_MOMENTUM * var * (n / (n - 1))
    bufs["num_batches_tracked"] = bufs["num_batches_tracked"] + 1
    return z, xhat, mean, invstd


def bn_train_bwd(dz, xhat, gamma, invstd):
    """Closed form of native_batch_norm_backward."""
    n = dz.shape[0] * dz.shape[2] * dz.shape[3]
    dbeta = dz.sum(dim=(0, 2, 3))
    dgamma = (dz * xhat).sum(dim=(0, 2, 3))
    dy = (gamma * invstd)[None, :, None, None] * (
        dz - dbeta[None, :, None, None] / n - xhat * dgamma[None, :, None, None] / n)
    return dy, dgamma, dbeta


# --------------------------------------------------------------------------- forward
def forward(p, bufs, x, n_blocks: int = 10, train: bool = True):
    """`NetResDeep.forward` (model/resnet.py:15-22) with every intermediate kept."""
    C = p["conv1.weight"].shape[0]
    c = {"x": x}
    a = F.conv2d(x, p["conv1.weight"], p["conv1.bias"], padding=1)               # resnet.py:16
    r = torch.relu(a)
    x0, idx0 = maxpool2_fwd(r)
    c["idx0"] = idx0
    xs, ys, xhats, means, invstds = [x0], [], [], [], []
    out = x0
    for _ in range(n_blocks):                                                    # resnet.py:17 (same block NB times)
        y = F.conv2d(out, p["rb.conv.weight"], None, padding=1)                  # resnet.py:34
        if train:
            z, xhat, mean, invstd = bn_train_fwd(y, p["rb.bn.weight"], p["rb.bn.bias"], bufs)  # resnet.py:35
        else:
            mean, invstd = bufs["running_mean"], 1.0 / torch.sqrt(bufs["running_var"] + BN_EPS)
            xhat = (y - mean[None, :, None, None]) * invstd[None, :, None, None]
            z = xhat * p["rb.bn.weight"][None, :, None, None] + p["rb.bn.bias"][None, :, None, None]
        out = torch.relu(z) + out                                                # resnet.py:36-37
        ys.append(y); xhats.append(xhat); means.append(mean); invstds.append(invstd); xs.append(out)
    pooled, idx2 = maxpool2_fwd(out)                                             # resnet.py:18
    flat = pooled.reshape(-1, 8 * 8 * C)                                         # resnet.py:19 (c,h,w order)
    h1pre = flat @ p["fc1.weight"].t() + p["fc1.bias"]                           # resnet.py:20
    h1 = torch.relu(h1pre)
    logits = h1 @ p["fc2.weight"].t() + p["fc2.bias"]                            # resnet.py:21
    c.update(xs=xs, ys=ys, xhats=xhats, means=means, invstds=invstds, idx2=idx2, flat=flat,
             h1=h1, logits=logits, n_blocks=n_blocks)
    return logits, c


def cross_entropy(logits, labels):
    """`nn.CrossEntropyLoss()` (main.py:28,35): mean NLL of log-softmax; grad (softmax-onehot)/B."""
    B = logits.shape[0]
    lse = torch.logsumexp(logits, dim=1)
    loss = (lse - logits.gather(1, labels.long()[:, None])[:, 0]).mean()
    dlogits = torch.softmax(logits, dim=1)
    dlogits[torch.arange(B), labels.long()] -= 1.0
    return loss, dlogits / B


# --------------------------------------------------------------------------- backward
def backward(p, c, dlogits, keep_intermediates: bool = False):
    """`loss.backward()` (main.py:38) written out.  Shared-weight grads are summed over the
    NB applications of the one ResBlock."""
    C = p["conv1.weight"].shape[0]
    g: Dict[str, torch.Tensor] = {}
    inter: Dict[str, list] = {"dx": [], "dy": []}
    g["fc2.weight"] = dlogits.t() @ c["h1"]
    g["fc2.bias"] = dlogits.sum(0)
    dh1 = (dlogits @ p["fc2.weight"]) * (c["h1"] > 0).to(dlogits.dtype)          # relu'(0) = 0
    g["fc1.weight"] = dh1.t() @ c["flat"]
    g["fc1.bias"] = dh1.sum(0)
    dflat = dh1 @ p["fc1.weight"]
    dout = maxpool2_bwd(dflat.reshape(-1, C, 8, 8), c["idx2"])
    dW = torch.zeros_like(p["rb.conv.weight"]); dgam = torch.zeros_like(p["rb.bn.weight"])
    dbet = torch.zeros_like(p["rb.bn.bias"])
    if keep_intermediates:
        inter["dx"].append(dout)
    for k in reversed(range(c["n_blocks"])):
        xhat, invstd, xk = c["xhats"][k], c["invstds"][k], c["xs"][k]
        z = xhat * p["rb.bn.weight"][None, :, None, None] + p["rb.bn.bias"][None, :, None, None]
        dz = dout * (z > 0).to(dout.dtype)
        dy, dga, dbe = bn_train_bwd(dz, xhat, p["rb.bn.weight"], invstd)
        dgam += dga; dbet += dbe
        dW += torch.nn.grad.conv2d_weight(xk, p["rb.conv.weight"].shape, dy, padding=1)
        dout = torch.nn.grad.conv2d_input(xk.shape, p["rb.conv.weight"], dy, padding=1) + dout
        if keep_intermediates:
            inter["dy"].append(dy); inter["dx"].append(dout)
    g["rb.conv.weight"], g["rb.bn.weight"], g["rb.bn.bias"] = dW, dgam, dbet
    dr = maxpool2_bwd(dout, c["idx0"])
    x0_full = maxpool2_bwd(c["xs"][0], c["idx0"])      # value at argmax position, 0 elsewhere
    da = dr * (x0_full > 0).to(dr.dtype)
    g["conv1.weight"] = torch.nn.grad.conv2d_weight(c["x"], p["conv1.weight"].shape, da, padding=1)
    g["conv1.bias"] = da.sum(dim=(0, 2, 3))
    if keep_intermediates:
        return g, inter
    return g


def sgd_step(p, g, lr: float = 1e-2):
    """`optim.SGD(lr=1e-2).step()` (main.py:27,39): p <- p - lr*g, no momentum / weight decay."""
    return {k: p[k] - lr * g[k] for k in p}


def train_step(p, bufs, x, labels, n_blocks: int = 10, lr: float = 1e-2):
    """The 6-line step of main_no_ddp.py:45-50 (= main.py:34-39)."""
    logits, c = forward(p, bufs, x, n_blocks)
    loss, dlogits = cross_entropy(logits, labels)
    g = backward(p, c, dlogits)
    return sgd_step(p, g, lr), logits, loss, g


def ddp_train_step(p, bufs_per_rank: List[dict], xs, labels, n_blocks: int = 10, lr: float = 1e-2):
    """Sequential simulation of W DDP ranks (main.py:63 + torch DDP defaults):
    rank-0 BN buffers overwrite every rank's before forward; each rank uses its own batch
    statistics; gradients are averaged; every rank applies the same SGD update."""
    W = len(xs)
    grads, losses, logits_all = [], [], []
    src = {k: v.clone() for k, v in bufs_per_rank[0].items()}       # rank 0's buffers BEFORE this forward
    for r in range(W):
        bufs_per_rank[r].update({k: v.clone() for k, v in src.items()})
        logits, c = forward(p, bufs_per_rank[r], xs[r], n_blocks)
        loss, dlogits = cross_entropy(logits, labels[r])
        grads.append(backward(p, c, dlogits)); losses.append(loss); logits_all.append(logits)
    gavg = {k: sum(g[k] for g in grads) / W for k in p}
    return sgd_step(p, gavg, lr), logits_all, losses, gavg


# --------------------------------------------------------------------------- rounding-matched (bf16) oracle
# SURVEY 8(c) tier 3(ii): the same explicit forward/backward with a quantiser `q` applied at exactly the
# points where the bf16 CUDA path rounds: conv operands (images, conv1 weight, block
# conv weight), every stored activation (x_k, y_k), dL/d(pooled), dy_k, dL/dx_k; BN statistics, gamma/beta,
# the head (fc1/fc2 weights, h1, logits, loss), weight-gradient accumulators and master weights stay in the
# arithmetic dtype.  With `q = identity` it IS `forward`/`backward` above (tests/test_oracle.py pins that),
# which in turn is pinned to the reference's own outputs.
def q_bf16(t: torch.Tensor) -> torch.Tensor:
    """round-to-nearest-even to bf16, back in the arithmetic dtype"""
    return t.to(torch.bfloat16).to(t.dtype)


def q_none(t: torch.Tensor) -> torch.Tensor:
    return t


def forward_rm(p, bufs, x, n_blocks: int = 10, train: bool = True, q=q_bf16):
    """`NetResDeep.forward` (model/resnet.py:15-22) with bf16 rounding where the kernels round.

    Differences to `forward` that are NOT rounding and do not change values: the stem pools the
    pre-bias convolution and adds bias + ReLU afterwards (max commutes with the monotone bias+ReLU;
    the arg-max can differ only where the output is 0 and the gradient is masked anyway); BN is
    applied as y*scale + shift with scale = gamma*invstd, shift = beta - mean*scale."""
    C = p["conv1.weight"].shape[0]
    c = {"x": x, "xq": q(x)}
    a = F.conv2d(c["xq"], q(p["conv1.weight"]), None, padding=1)                 # resnet.py:16, operands rounded
    amax, idx0 = maxpool2_fwd(a)
    x0 = q(torch.relu(amax + p["conv1.bias"][None, :, None, None]))
    c["idx0"] = idx0
    wq = q(p["rb.conv.weight"])
    gamma, beta = p["rb.bn.weight"], p["rb.bn.bias"]
    xs, ybs, scs, shs, means, invstds = [x0], [], [], [], [], []
    out = x0
    for _ in range(n_blocks):                                                    # resnet.py:17
        y = F.conv2d(out, wq, None, padding=1)                                   # resnet.py:34 (fp32 accumulate)
        if train:
            n = y.shape[0] * y.shape[2] * y.shape[3]
            mean = y.mean(dim=(0, 2, 3))                                         # statistics of the UNROUNDED accumulators
            var = (y * y).mean(dim=(0, 2, 3)) - mean * mean
            var = torch.clamp(var, min=0.0)
            invstd = 1.0 / torch.sqrt(var + BN_EPS)
            bufs["running_mean"] = (1 - BN_MOMENTUM) * bufs["running_mean"] + BN_MOMENTUM * mean
            bufs["running_var"] = (1 - BN_MOMENTUM) * bufs["running_var"] + BN_MOMENTUM * var * (n / (n - 1))
            bufs["num_batches_tracked"] = bufs["num_batches_tracked"] + 1
        else:
            mean, invstd = bufs["running_mean"], 1.0 / torch.sqrt(bufs["running_var"] + BN_EPS)
        sc = gamma * invstd
        sh = beta - mean * sc
        yb = q(y)                                                                # y_k as stored; BN sees the stored value
        z = yb * sc[None, :, None, None] + sh[None, :, None, None]               # resnet.py:35
        out = q(torch.relu(z) + out)                                             # resnet.py:36-37
        ybs.append(yb); scs.append(sc); shs.append(sh); means.append(mean); invstds.append(invstd); xs.append(out)
    pooled, idx2 = maxpool2_fwd(out)                                             # resnet.py:18 (exact on bf16 values)
    flat = pooled.reshape(-1, 8 * 8 * C)                                         # resnet.py:19
    h1 = torch.relu(flat @ p["fc1.weight"].t() + p["fc1.bias"])                  # resnet.py:20 (fp32 weights)
    logits = h1 @ p["fc2.weight"].t() + p["fc2.bias"]                            # resnet.py:21
    c.update(xs=xs, ybs=ybs, scs=scs, shs=shs, means=means, invstds=invstds, idx2=idx2, flat=flat, h1=h1,
             logits=logits, n_blocks=n_blocks)
    return logits, c


def backward_rm(p, c, dlogits, q=q_bf16, keep_masks: bool = False):
    """`loss.backward()` (main.py:38) with bf16 rounding where the kernels round: d(pooled), dy_k and
    dL/dx_k are stored/kept as bf16; the BN-backward sums, dgamma/dbeta and all weight gradients
    accumulate in the arithmetic dtype."""
    C = p["conv1.weight"].shape[0]
    g: Dict[str, torch.Tensor] = {}
    g["fc2.weight"] = dlogits.t() @ c["h1"]
    g["fc2.bias"] = dlogits.sum(0)
    dh1 = (dlogits @ p["fc2.weight"]) * (c["h1"] > 0).to(dlogits.dtype)
    g["fc1.weight"] = dh1.t() @ c["flat"]
    g["fc1.bias"] = dh1.sum(0)
    dflat = q(dh1 @ p["fc1.weight"])                                             # dL/d(pooled), stored bf16
    dout = maxpool2_bwd(dflat.reshape(-1, C, 8, 8), c["idx2"])
    wq = q(p["rb.conv.weight"])
    gamma = p["rb.bn.weight"]
    dW = torch.zeros_like(p["rb.conv.weight"]); dgam = torch.zeros_like(gamma); dbet = torch.zeros_like(gamma)
    masks = []
    for k in reversed(range(c["n_blocks"])):
        yb, sc, sh, mean, invstd, xk = c["ybs"][k], c["scs"][k], c["shs"][k], c["means"][k], c["invstds"][k], c["xs"][k]
        n = yb.shape[0] * yb.shape[2] * yb.shape[3]
        mask = (yb * sc[None, :, None, None] + sh[None, :, None, None]) > 0
        if keep_masks:
            masks.append(mask)
        dz = dout * mask.to(dout.dtype)
        db = dz.sum(dim=(0, 2, 3))
        dgy = (dz * yb).sum(dim=(0, 2, 3))
        dg = invstd * (dgy - mean * db)                                          # sum dz*xhat
        dgam += dg; dbet += db
        a_ = gamma * invstd
        A2 = -a_ * (dg / n) * invstd
        A3 = a_ * (dg / n) * invstd * mean - a_ * (db / n)
        dy = q(dz * a_[None, :, None, None] + yb * A2[None, :, None, None] + A3[None, :, None, None])
        dW += torch.nn.grad.conv2d_weight(xk, p["rb.conv.weight"].shape, dy, padding=1)
        dout = q(torch.nn.grad.conv2d_input(xk.shape, wq, dy, padding=1) + dout)
    g["rb.conv.weight"], g["rb.bn.weight"], g["rb.bn.bias"] = dW, dgam, dbet
    dr = maxpool2_bwd(dout, c["idx0"])
    x0_full = maxpool2_bwd(c["xs"][0], c["idx0"])
    da = dr * (x0_full > 0).to(dr.dtype)
    g["conv1.weight"] = torch.nn.grad.conv2d_weight(c["xq"], p["conv1.weight"].shape, da, padding=1)
    g["conv1.bias"] = da.sum(dim=(0, 2, 3))
    if keep_masks:
        return g, masks[::-1]
    return g


def train_step_rm(p, bufs, x, labels, n_blocks: int = 10, lr: float = 1e-2, q=q_bf16, keep_masks: bool = False):
    """main.py:34-39 with the bf16 path's rounding points; master weights stay in the arithmetic dtype."""
    logits, c = forward_rm(p, bufs, x, n_blocks, q=q)
    loss, dlogits = cross_entropy(logits, labels)
    out = backward_rm(p, c, dlogits, q=q, keep_masks=keep_masks)
    g, masks = out if keep_masks else (out, None)
    res = (sgd_step(p, g, lr), logits, loss, g)
    # keep_masks: the ReLU masks of the NB blocks (list) and of fc1's output, for flip-aware comparisons
    return res + ({"blocks": masks, "h1": c["h1"] > 0},) if keep_masks else res


def relu_masks(p, c):
    """per-block ReLU masks of a plain `forward` cache (for flip counts against `backward_rm(keep_masks=True)`)"""
    return [(xh * p["rb.bn.weight"][None, :, None, None] + p["rb.bn.bias"][None, :, None, None]) > 0 for xh in c["xhats"]]


def ddp_train_step_rm(p, bufs_per_rank: List[dict], xs, labels, n_blocks: int = 10, lr: float = 1e-2, q=q_bf16):
    """`ddp_train_step` with the bf16 path's rounding points on every rank."""
    W = len(xs)
    grads, losses, logits_all = [], [], []
    src = {k: v.clone() for k, v in bufs_per_rank[0].items()}
    for r in range(W):
        bufs_per_rank[r].update({k: v.clone() for k, v in src.items()})
        logits, c = forward_rm(p, bufs_per_rank[r], xs[r], n_blocks, q=q)
        loss, dlogits = cross_entropy(logits, labels[r])
        grads.append(backward_rm(p, c, dlogits, q=q)); losses.append(loss); logits_all.append(logits)
    gavg = {k: sum(g[k] for g in grads) / W for k in p}
    return sgd_step(p, gavg, lr), logits_all, losses, gavg


# --------------------------------------------------------------------------- timed CPU baseline
class TorchModulePort(torch.nn.Module):
    """nn.Module port of NetResDeep used ONLY as the timed CPU baseline (`cpu_baseline.kind =
    "port"`): it dispatches the same ATen ops under autograd as the reference model does
    (model/resnet.py:15-22, 33-37), so its CPU step time is the reference's CPU step time."""

    def __init__(self, p, bufs, n_blocks=10):
        super().__init__()
        self.n_blocks = n_blocks
        self.params = torch.nn.ParameterDict({k.replace(".", "_"): torch.nn.Parameter(v.clone()) for k, v in p.items()})
        self.register_buffer("rm", bufs["running_mean"].clone())
        self.register_buffer("rv", bufs["running_var"].clone())

    def forward(self, x):
        q = self.params
        out = F.max_pool2d(torch.relu(F.conv2d(x, q["conv1_weight"], q["conv1_bias"], padding=1)), 2)
        for _ in range(self.n_blocks):
            y = F.conv2d(out, q["rb_conv_weight"], None, padding=1)
            y = F.batch_norm(y, self.rm, self.rv, q["rb_bn_weight"], q["rb_bn_bias"], True, BN_MOMENTUM, BN_EPS)
            out = torch.relu(y) + out
        out = F.max_pool2d(out, 2).view(x.shape[0], -1)
        out = torch.relu(F.linear(out, q["fc1_weight"], q["fc1_bias"]))
        return F.linear(out, q["fc2_weight"], q["fc2_bias"])


def time_cpu_step(batch: int = 64, n_chans1: int = 32, n_blocks: int = 10, steps: int = 20, warmup: int = 3,
                  threads: int | None = None, seed: int = 0):
    """Time the restated main_no_ddp.py:45-50 step on CPU; returns (best_s, median_s, threads)."""
    import time
    if threads:
        torch.set_num_threads(threads)
    p, bufs = init_params(n_chans1, n_blocks, seed)
    m = TorchModulePort(p, bufs, n_blocks)
    opt = torch.optim.SGD(m.parameters(), lr=1e-2)
    loss_fn = torch.nn.CrossEntropyLoss()
    gen = torch.Generator().manual_seed(1234)
    x = torch.randn(batch, 3, 32, 32, generator=gen)
    y = torch.randint(0, 10, (batch,), generator=gen)
    ts = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        out = m(x); loss = loss_fn(out, y)
        opt.zero_grad(); loss.backward(); opt.step()
        _ = loss.item()
        if i >= warmup:
            ts.append(time.perf_counter() - t0)
    ts.sort()
    return ts[0], ts[len(ts) // 2], torch.get_num_threads()
