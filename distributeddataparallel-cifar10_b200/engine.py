"""Host-side engine: owns the flat parameter/gradient vectors, per-batch-size plans and workspaces
and drives libnrd.so on torch's current CUDA stream.  torch is plumbing here (device memory,
streams); every kernel that runs is ours.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, List, Optional

import torch

from . import _lib
from ._lib import lib, check

PARAM_NAMES = ("conv1.weight", "conv1.bias", "resblocks.0.conv.weight", "resblocks.0.batch_norm.weight",
               "resblocks.0.batch_norm.bias", "fc1.weight", "fc1.bias", "fc2.weight", "fc2.bias")
N_BODY = 5   # first 5 tensors = "body" bucket, last 4 = "head" bucket

_DTYPES = {"fp32": _lib.NRD_F32, "float32": _lib.NRD_F32, "bf16": _lib.NRD_BF16, "bfloat16": _lib.NRD_BF16}


def default_dtype() -> str:
    return os.environ.get("NRD_DTYPE", "bf16")


def _stream_ptr() -> int:
    return torch.cuda.current_stream().cuda_stream


class _Plan:
    def __init__(self, eng: "Engine", B: int):
        self.B = B
        nbytes = lib.nrd_workspace_bytes(B, eng.C, eng.NB, eng.dtype_code)
        if nbytes == 0:
            raise _lib.NrdError("nrd_workspace_bytes: " + _lib.last_error())
        self.workspace = torch.empty(nbytes + 256, dtype=torch.uint8, device=eng.device)
        base = (self.workspace.data_ptr() + 255) & ~255
        h = C.c_void_p()
        check(lib.nrd_plan_create(C.byref(h), eng.device.index, B, eng.C, eng.NB, eng.dtype_code, base, nbytes,
                                  eng.flat_params.data_ptr(), eng.flat_grads.data_ptr(), eng.bn_running.data_ptr(),
                                  eng.nbt.data_ptr()), "nrd_plan_create")
        self.handle = h
        for env, opt in (("NRD_HEAD_TAIL", _lib.OPT_HEAD_TAIL), ("NRD_FUSE_POOL", _lib.OPT_FUSE_POOL), ("NRD_FUSE_STEM", _lib.OPT_FUSE_STEM), ("NRD_ONE_COPY", _lib.OPT_ONE_COPY),
                         ("NRD_FUSE_STEM_BWD", _lib.OPT_FUSE_STEM_BWD), ("NRD_STREAM", _lib.OPT_STREAM)):   # A/B switches
            if env in os.environ:
                check(lib.nrd_plan_set_option(h, opt, int(os.environ[env])), "nrd_plan_set_option")
        if "NRD_PDL" in os.environ:                        # A/B switch for measurements (bit mask, see nrd.h)
            check(lib.nrd_plan_set_option(h, _lib.OPT_PDL, int(os.environ["NRD_PDL"])), "nrd_plan_set_option")
        self.dlogits = torch.empty(B, 10, dtype=torch.float32, device=eng.device)
        self.loss_rows = torch.empty(B, dtype=torch.float32, device=eng.device)
        self.packed_version = -1

    def tensor(self, what: int, k: int = 0):
        p, n = C.c_void_p(), C.c_size_t()
        check(lib.nrd_plan_tensor(self.handle, what, k, C.byref(p), C.byref(n)), "nrd_plan_tensor")
        return p.value, n.value

    def tensor_view(self, what: int, k: int = 0) -> torch.Tensor:
        """uint8 torch view of a saved tensor inside the workspace (tests / debugging)."""
        ptr, n = self.tensor(what, k)
        off = ptr - self.workspace.data_ptr()
        return self.workspace[off:off + n]

    def __del__(self):
        try:
            lib.nrd_plan_destroy(self.handle)
        except Exception:
            pass


class Engine:
    """Binds one NetResDeep module (its 9 unique parameters + 3 BN buffers) to libnrd plans."""

    def __init__(self, module, dtype: Optional[str] = None):
        self.module = module
        self.C = module.n_chans1
        self.NB = len(module.resblocks)
        self.dtype_name = dtype or default_dtype()
        if self.dtype_name not in _DTYPES:
            raise ValueError(f"dtype must be one of {sorted(_DTYPES)}, got {self.dtype_name!r}")
        self.dtype_code = _DTYPES[self.dtype_name]
        self.offsets = _lib.param_offsets(self.C)
        self.device: Optional[torch.device] = None
        self.plans: Dict[int, _Plan] = {}
        self.flat_params = self.flat_grads = self.bn_running = self.nbt = None
        self.conv_impl: Optional[int] = None
        self._last_plan: Optional[_Plan] = None
        self.weights_epoch = 0       # bumped by our own fused update (torch's _version cannot see it)

    # -- parameter adoption ----------------------------------------------------------------
    def params(self) -> List[torch.nn.Parameter]:
        named = dict(self.module.named_parameters())
        return [named[n] for n in PARAM_NAMES]

    def _adopted(self) -> bool:
        if self.flat_params is None:
            return False
        base = self.flat_params.data_ptr()
        bn = self.module.resblocks[0].batch_norm
        return (all(p.data_ptr() == base + 4 * o for p, o in zip(self.params(), self.offsets))
                and bn.running_mean.data_ptr() == self.bn_running.data_ptr()
                and bn.num_batches_tracked.data_ptr() == self.nbt.data_ptr())

    def adopt(self, device: torch.device):
        """Re-home the module's parameters and BN buffers as views of flat vectors (same Parameter
        objects, same values), so buckets/updates need no gather.  Redone after `.to(device)`."""
        ps = self.params()
        if device.type == "cuda" and device.index is None:
            device = torch.device("cuda", torch.cuda.current_device())
        if any(p.device != device for p in ps):
            raise RuntimeError(f"NetResDeep parameters live on {ps[0].device} but the input is on {device}; call model.to(device) first")
        if device.type != "cuda":
            raise RuntimeError("NetResDeep (B200) runs on CUDA devices only: there is no CPU fallback")
        n = self.offsets[9]
        flat = torch.empty(n, dtype=torch.float32, device=device)
        # gradient exchange buffer: [2C BN-sync slot | body bucket | head bucket]; the slot lets rank 0's running stats
        # ride the body bucket's all-reduce (graph.TrainStep) instead of a broadcast in front of every forward
        self.xchg = torch.zeros(2 * self.C + n, dtype=torch.float32, device=device)
        grads = self.xchg[2 * self.C:]
        for p, o in zip(ps, self.offsets):
            view = flat[o:o + p.numel()].view(p.shape)
            view.copy_(p.data.float())
            p.data = view
        bn = self.module.resblocks[0].batch_norm
        run = torch.empty(2 * self.C, dtype=torch.float32, device=device)
        run[:self.C].copy_(bn.running_mean); run[self.C:].copy_(bn.running_var)
        bn.running_mean.data = run[:self.C]; bn.running_var.data = run[self.C:]
        self.flat_params, self.flat_grads, self.bn_running = flat, grads, run
        self.nbt = bn.num_batches_tracked
        if self.nbt.device != device or self.nbt.dtype != torch.int64:
            raise RuntimeError("num_batches_tracked must be an int64 scalar on the model's device")
        self.device = device
        self.plans.clear()
        self._last_plan = None

    def grad_views(self) -> List[torch.Tensor]:
        return [self.flat_grads[o:o + p.numel()].view(p.shape) for p, o in zip(self.params(), self.offsets)]

    def body_bucket_with_bn_slot(self) -> torch.Tensor:
        """BN-sync slot + body bucket as ONE contiguous all-reduce operand."""
        return self.xchg[:2 * self.C + self.offsets[N_BODY]]

    def bucket_slices(self):
        """(body, head) element ranges of the flat vectors."""
        return (0, self.offsets[N_BODY]), (self.offsets[N_BODY], self.offsets[9])

    # -- plans ------------------------------------------------------------------------------
    def plan(self, B: int, device: torch.device) -> _Plan:
        if self.device != device or not self._adopted():
            self.adopt(device)
        pl = self.plans.get(B)
        if pl is None:
            pl = _Plan(self, B)
            if self.conv_impl is not None:
                check(lib.nrd_plan_set_conv_impl(pl.handle, self.conv_impl), "nrd_plan_set_conv_impl")
            self.plans[B] = pl
        return pl

    def check_status(self, clear: bool = True):
        """Raise if a kernel of any cached plan clamped or saw non-finite BatchNorm sums since the last check
        (nrd_plan_status).  Synchronises the current stream: call it where the loop reads the loss on the host."""
        for pl in self.plans.values():
            flags = C.c_uint32(0)
            rc = lib.nrd_plan_status(pl.handle, C.byref(flags), 1 if clear else 0, _stream_ptr())
            if rc != 0:
                raise _lib.NrdError(f"training step status {flags.value:#x} (batch {pl.B}): {_lib.last_error()}")

    def set_conv_impl(self, impl: int):
        self.conv_impl = impl
        for pl in self.plans.values():
            check(lib.nrd_plan_set_conv_impl(pl.handle, impl), "nrd_plan_set_conv_impl")

    def _weights_version(self) -> int:
        return sum(p._version for p in self.params()) + (self.weights_epoch << 20)

    def ensure_packed(self, pl: _Plan):
        v = self._weights_version()
        if pl.packed_version != v:
            check(lib.nrd_pack_weights(pl.handle, _stream_ptr()), "nrd_pack_weights")
            pl.packed_version = v

    # -- the step pieces --------------------------------------------------------------------
    def forward(self, x: torch.Tensor, training: bool = True) -> torch.Tensor:
        if x.dim() != 4 or tuple(x.shape[1:]) != (3, 32, 32):
            raise ValueError(f"expected input of shape [B,3,32,32], got {tuple(x.shape)}")
        if x.dtype != torch.float32:
            raise TypeError(f"expected float32 input, got {x.dtype}")
        x = x.contiguous()
        pl = self.plan(x.shape[0], x.device)
        self.ensure_packed(pl)
        logits = torch.empty(x.shape[0], 10, dtype=torch.float32, device=x.device)
        check(lib.nrd_forward(pl.handle, x.data_ptr(), logits.data_ptr(), 1 if training else 0, _stream_ptr()), "nrd_forward")
        pl.x_ref = x            # stem backward reads the images again
        self._last_plan = pl
        return logits

    def backward(self, pl: _Plan, dlogits: torch.Tensor, after_head=None, after_body=None):
        """loss.backward() below the logits.  The head weight gradients (fc1/fc2: 86.6 % of the gradient bytes) run on a
        side stream beside the body backward; `after_head` / `after_body` are called with that half's stream current right
        after its kernels are enqueued (the DDP wrapper launches each bucket's all-reduce there).  Both halves are joined
        on the caller's stream before returning."""
        dlogits = dlogits.contiguous().float()
        main = torch.cuda.current_stream()
        side = self._side_stream()
        fork = torch.cuda.Event(); join = torch.cuda.Event()
        fork.record(main)
        side.wait_event(fork)
        with torch.cuda.stream(side):
            check(lib.nrd_backward_head(pl.handle, dlogits.data_ptr(), side.cuda_stream), "nrd_backward_head")
            if after_head is not None:
                after_head()
            join.record(side)
        check(lib.nrd_backward_body(pl.handle, dlogits.data_ptr(), main.cuda_stream), "nrd_backward_body")
        if after_body is not None:
            after_body()
        main.wait_event(join)
        if not torch.cuda.is_current_stream_capturing():
            dlogits.record_stream(side)

    def _side_stream(self):
        if getattr(self, "_side", None) is None or self._side.device != self.device:
            self._side = torch.cuda.Stream(device=self.device)
        return self._side

    def sgd_update(self, lr: float, inv_world: float = 1.0, lo: int = 0, hi: Optional[int] = None):
        if lo == 0 and hi is None and self._last_plan is not None:
            # whole vector: update and weight re-pack in one launch
            check(lib.nrd_sgd_update_pack(self._last_plan.handle, lr, inv_world, _stream_ptr()), "nrd_sgd_update_pack")
            self.weights_epoch += 1
            self._last_plan.packed_version = self._weights_version()
            return
        hi = self.offsets[9] if hi is None else hi
        check(lib.nrd_sgd_update(self.flat_params.data_ptr() + 4 * lo, self.flat_grads.data_ptr() + 4 * lo, hi - lo,
                                 lr, inv_world, _stream_ptr()), "nrd_sgd_update")
        self.weights_epoch += 1


def cross_entropy_fwd_bwd(logits: torch.Tensor, labels: torch.Tensor, grad_scale: float = 1.0):
    """(loss scalar tensor, dlogits) from one fused launch pair (main.py:28,35)."""
    B = logits.shape[0]
    logits = logits.contiguous()
    labels = labels.contiguous()
    if labels.dtype != torch.int64:
        raise TypeError("labels must be int64 (as the reference DataLoader yields)")
    loss = torch.empty((), dtype=torch.float32, device=logits.device)
    dlogits = torch.empty_like(logits)
    rows = torch.empty(B, dtype=torch.float32, device=logits.device)
    check(lib.nrd_cross_entropy(logits.data_ptr(), labels.data_ptr(), B, grad_scale, loss.data_ptr(), dlogits.data_ptr(),
                                rows.data_ptr(), _stream_ptr()), "nrd_cross_entropy")
    return loss, dlogits
