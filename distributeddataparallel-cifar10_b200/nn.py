"""`CrossEntropyLoss` with the reference's call signature (main.py:28,35), fused fwd+grad kernel."""
import torch

from .functional import cross_entropy


class CrossEntropyLoss(torch.nn.Module):
    """Drop-in for `nn.CrossEntropyLoss()` as used by the reference: mean reduction, class-index
    targets, no weights / label smoothing (anything else raises instead of silently differing)."""

    def __init__(self, weight=None, size_average=None, ignore_index=-100, reduce=None, reduction="mean", label_smoothing=0.0):
        super().__init__()
        if weight is not None or reduction != "mean" or label_smoothing != 0.0 or size_average is not None or reduce is not None:
            raise NotImplementedError("only nn.CrossEntropyLoss() defaults (mean reduction) are on the reference's path")

    def forward(self, input, target):
        if input.dim() != 2 or input.shape[1] != 10:
            raise ValueError(f"expected logits [B,10], got {tuple(input.shape)}")
        return cross_entropy(input, target)


class _BCEWithLogitsFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, targets):
        from ._lib import lib, check
        from .engine import _stream_ptr
        x = logits.contiguous().float(); t = targets.contiguous().float()
        if not x.is_cuda or x.shape != t.shape:
            raise ValueError("binary_cross_entropy_with_logits (B200): CUDA tensors of equal shape")
        loss = torch.empty((), dtype=torch.float32, device=x.device); dx = torch.empty_like(x)
        check(lib.nrd_bce_with_logits(x.data_ptr(), t.data_ptr(), x.numel(), 1.0, loss.data_ptr(), dx.data_ptr(), _stream_ptr()), "nrd_bce_with_logits")
        ctx.save_for_backward(dx)
        return loss

    @staticmethod
    def backward(ctx, gout):
        (dx,) = ctx.saved_tensors
        return dx * gout, None


def binary_cross_entropy_with_logits(input, target):
    """`F.binary_cross_entropy_with_logits(input, target)` (mean reduction) as the reference's second script calls it
    (ppe_main_ddp.py:147), fused forward + gradient kernel."""
    return _BCEWithLogitsFn.apply(input, target)


class BCEWithLogitsLoss(torch.nn.Module):
    def forward(self, input, target):
        return binary_cross_entropy_with_logits(input, target)
