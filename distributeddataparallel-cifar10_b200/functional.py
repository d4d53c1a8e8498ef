"""Custom autograd Functions: the seam between torch autograd and libnrd.so.

`netresdeep_forward` is ONE Function for the whole model (per-op Python overhead would exceed the
~70 us step budget); `resblock_forward` and `cross_entropy` are the single-op seams.
"""
import ctypes as C

import torch

from . import _lib
from ._lib import lib, check
from .engine import _stream_ptr, cross_entropy_fwd_bwd


class _NetResDeepFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, module, *params):
        eng = module.engine
        logits = eng.forward(x, training=module.training)
        ctx.eng, ctx.plan, ctx.module = eng, eng._last_plan, module
        return logits

    @staticmethod
    def backward(ctx, dlogits):
        eng, pl = ctx.eng, ctx.plan
        params = eng.params()
        views = eng.grad_views()
        # accumulate semantics if the caller did not zero_grad (the reference always does, main.py:37)
        keep = [None if p.grad is None else p.grad.detach().clone() for p in params]
        hooks = getattr(ctx.module, "_nrd_bucket_hooks", None)      # DDP wrapper: (head, body) bucket all-reduces
        eng.backward(pl, dlogits, *(hooks or (None, None)))
        for p, v, k in zip(params, views, keep):
            if k is not None:
                v.add_(k)
            p.grad = v
        return (None, None) + (None,) * len(params)


def netresdeep_forward(module, x):
    if not x.is_cuda:
        raise RuntimeError("NetResDeep (B200) needs a CUDA input: there is no CPU fallback")
    if not module.training or not torch.is_grad_enabled():
        return module.engine.forward(x, training=module.training)
    return _NetResDeepFn.apply(x, module, *module.engine.params())


class _CrossEntropyFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, labels):
        loss, dlogits = cross_entropy_fwd_bwd(logits, labels)
        ctx.save_for_backward(dlogits)
        return loss

    @staticmethod
    def backward(ctx, gout):
        (dlogits,) = ctx.saved_tensors
        return dlogits * gout, None


def cross_entropy(logits, labels):
    """mean softmax cross-entropy, `nn.CrossEntropyLoss()(logits, labels)` (main.py:28,35)."""
    return _CrossEntropyFn.apply(logits, labels)


# ---- single ResBlock (per-op entries) ------------------------------------------------------
def _code(dtype):
    return _lib.NRD_F32 if dtype == torch.float32 else _lib.NRD_BF16


class _ResBlockFn(torch.autograd.Function):
    """relu(bn(conv(x))) + x via nrd_conv3x3_fwd / nrd_bn_finalize / nrd_bn_relu_add and their
    backward entries; NHWC inside, NCHW fp32 at the module boundary."""

    @staticmethod
    def forward(ctx, x, w, gamma, beta, bn, act_dtype):
        B, Cc, H, W = x.shape
        if (H, W) != (16, 16) or Cc not in (32, 64):
            raise ValueError("ResBlock (B200) supports [B,32|64,16,16] inputs (the resolution NetResDeep runs it at)")
        dev, st, code = x.device, _stream_ptr(), _code(act_dtype)
        xh = x.permute(0, 2, 3, 1).contiguous().to(act_dtype)
        wf = torch.empty(9 * Cc * Cc, dtype=torch.float32, device=dev); wd = torch.empty_like(wf)
        check(lib.nrd_pack_conv_weight(w.contiguous().data_ptr(), Cc, wf.data_ptr(), wd.data_ptr(), st))
        y = torch.empty_like(xh)
        partial = torch.empty(B * 2 * 2 * Cc, dtype=torch.float32, device=dev)
        nblk = C.c_int(0)
        check(lib.nrd_conv3x3_fwd(xh.data_ptr(), wf.data_ptr(), y.data_ptr(), B, Cc, code, partial.data_ptr(), C.byref(nblk), None, st))
        saved = torch.empty(2 * Cc, dtype=torch.float32, device=dev); ss = torch.empty_like(saved)
        run = torch.cat([bn.running_mean, bn.running_var]).float().contiguous() if bn.training else None
        check(lib.nrd_bn_finalize(partial.data_ptr(), nblk.value, B * 256, Cc, gamma.contiguous().data_ptr(), beta.contiguous().data_ptr(),
                                  saved.data_ptr(), ss.data_ptr(), run.data_ptr() if run is not None else None,
                                  bn.num_batches_tracked.data_ptr() if bn.training else None, st))
        if run is not None:
            bn.running_mean.copy_(run[:Cc]); bn.running_var.copy_(run[Cc:])
        out = torch.empty_like(xh)
        check(lib.nrd_bn_relu_add(y.data_ptr(), xh.data_ptr(), ss.data_ptr(), out.data_ptr(), B, Cc, code, st))
        ctx.save_for_backward(xh, y, saved, ss, wd, gamma)
        ctx.code = code
        return out.float().permute(0, 3, 1, 2).contiguous()

    @staticmethod
    def backward(ctx, gout):
        xh, y, saved, ss, wd, gamma = ctx.saved_tensors
        B, _, _, Cc = xh.shape
        dev, st, code = xh.device, _stream_ptr(), ctx.code
        g = gout.permute(0, 2, 3, 1).contiguous().to(xh.dtype)
        dy = torch.empty_like(xh)
        dgb = torch.empty(2 * Cc, dtype=torch.float32, device=dev)
        scratch = torch.empty(lib.nrd_bn_bwd_scratch_bytes(B, Cc) // 4, dtype=torch.float32, device=dev)
        check(lib.nrd_bn_relu_bwd(g.data_ptr(), y.data_ptr(), ss.data_ptr(), saved.data_ptr(), gamma.contiguous().data_ptr(), dy.data_ptr(),
                                  dgb.data_ptr(), 0, B, Cc, code, scratch.data_ptr(), st))
        dw = torch.empty(Cc, Cc, 3, 3, dtype=torch.float32, device=dev)
        part = torch.empty(lib.nrd_wgrad_partial_bytes(B, Cc) // 4, dtype=torch.float32, device=dev)
        check(lib.nrd_conv3x3_wgrad(xh.data_ptr(), dy.data_ptr(), dw.data_ptr(), B, Cc, code, part.data_ptr(), st))
        dx = torch.empty_like(xh)
        check(lib.nrd_conv3x3_fwd(dy.data_ptr(), wd.data_ptr(), dx.data_ptr(), B, Cc, code, None, None, g.data_ptr(), st))
        return dx.float().permute(0, 3, 1, 2).contiguous(), dw, dgb[:Cc].clone(), dgb[Cc:].clone(), None, None


def resblock_forward(block, x, act_dtype=torch.float32):
    if not x.is_cuda:
        raise RuntimeError("ResBlock (B200) needs a CUDA input: there is no CPU fallback")
    if not block.training:
        raise NotImplementedError("standalone ResBlock eval mode: use NetResDeep.eval() (whole-model entry)")
    bn = block.batch_norm
    return _ResBlockFn.apply(x, block.conv.weight, bn.weight, bn.bias, bn, act_dtype)
