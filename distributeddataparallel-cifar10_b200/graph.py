"""Whole-step CUDA graph: forward, loss, backward, gradient exchange, fused mean+SGD and weight
re-pack replayed as ONE graph launch (the ~70 us roofline step cannot afford ~60 separate
launches from Python).  This is main.py:34-39 for one rank."""
import os
from typing import Optional

import torch
import torch.distributed as dist

from ._lib import lib, check, OPT_RESERVE_SMS
from .engine import _stream_ptr


class TrainStep:
    """step = TrainStep(model_or_ddp, batch_size, lr); loss = step(imgs, labels)

    `imgs`/`labels` may be device tensors (copied device-to-device into the graph's static inputs)
    or pinned host tensors (async H2D on the same stream).  Returns the device loss tensor of the
    step (read it with .item() only when you need the number, main.py:41)."""

    def __init__(self, model, batch_size: int, lr: float = 1e-2, use_graph: bool = True):
        from .parallel import DistributedDataParallel
        self.ddp = model if isinstance(model, DistributedDataParallel) else None
        self.module = model.module if self.ddp is not None else model
        self.eng = self.module.engine
        self.B, self.lr = batch_size, lr
        dev = next(self.module.parameters()).device
        if dev.type != "cuda":
            raise RuntimeError("TrainStep needs the model on a CUDA device")
        self.device = dev
        self.world = self.ddp.world_size if self.ddp is not None else 1
        self.imgs = torch.zeros(batch_size, 3, 32, 32, dtype=torch.float32, device=dev)
        self.labels = torch.zeros(batch_size, dtype=torch.int64, device=dev)
        self.logits = torch.empty(batch_size, 10, dtype=torch.float32, device=dev)
        self.loss = torch.zeros((), dtype=torch.float32, device=dev)
        self.plan = self.eng.plan(batch_size, dev)
        self.eng.ensure_packed(self.plan)
        # DDP on the graph path: (1) rank 0's BN running stats ride the body bucket's all-reduce of the PREVIOUS step
        # instead of a broadcast in front of every forward (same values: the buffers do not change in between; exact
        # for power-of-two world sizes, otherwise the broadcast stays); (2) the persistent kernels leave a few SMs
        # free so the head bucket's all-reduce really runs beside the body backward.
        self.fold_bn_sync = (self.world > 1 and self.ddp.broadcast_buffers and (self.world & (self.world - 1)) == 0
                             and os.environ.get("NRD_FOLD_BN_SYNC", "1") != "0")
        if self.world > 1:
            rank = dist.get_rank(self.ddp.process_group)
            if self.fold_bn_sync:
                check(lib.nrd_plan_set_bn_sync(self.plan.handle, self.eng.xchg.data_ptr(), float(self.world) if rank == 0 else 0.0,
                                               1.0 / self.world), "nrd_plan_set_bn_sync")
                self.resync_buffers()
            check(lib.nrd_plan_set_option(self.plan.handle, OPT_RESERVE_SMS, int(os.environ.get("NRD_RESERVE_SMS", "8"))),
                  "nrd_plan_set_option")
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self.kernel_launches = None
        if use_graph:
            self._capture()

    def resync_buffers(self):
        """Rank 0's BN running stats to every rank, and into the slot the next forward starts from.  Called at construction;
        call it again after changing the buffers outside the step (load_state_dict)."""
        if self.world > 1 and self.ddp.broadcast_buffers:
            self.ddp._sync_buffers()
            if self.fold_bn_sync:
                torch.mul(self.eng.bn_running, float(self.world), out=self.eng.xchg[:2 * self.eng.C])

    # the step body, launched on the current stream
    def _body(self):
        eng, pl, st = self.eng, self.plan, _stream_ptr()
        if self.world > 1 and not self.fold_bn_sync:
            self.ddp._sync_buffers()                                   # C2 (folded into the body bucket otherwise)
        check(lib.nrd_forward(pl.handle, self.imgs.data_ptr(), self.logits.data_ptr(), 1, st), "nrd_forward")
        check(lib.nrd_cross_entropy(self.logits.data_ptr(), self.labels.data_ptr(), self.B, 1.0, self.loss.data_ptr(),
                                    pl.dlogits.data_ptr(), pl.loss_rows.data_ptr(), st), "nrd_cross_entropy")
        # backward: head weight grads (+ their bucket's all-reduce) on a side stream beside the body backward (C3)
        sum_head = (lambda: self.ddp._allreduce_bucket(1, average=False)) if self.world > 1 else None
        sum_body = None
        if self.world > 1:
            sum_body = ((lambda: dist.all_reduce(eng.body_bucket_with_bn_slot(), op=dist.ReduceOp.SUM, group=self.ddp.process_group))
                        if self.fold_bn_sync else (lambda: self.ddp._allreduce_bucket(0, average=False)))
        eng.backward(pl, pl.dlogits, sum_head, sum_body)
        # fused mean + SGD (K16), then re-derive the packed weight copies
        check(lib.nrd_sgd_update_pack(pl.handle, self.lr, 1.0 / self.world, st), "nrd_sgd_update_pack")

    def _capture(self):
        s = torch.cuda.Stream(device=self.device)
        s.wait_stream(torch.cuda.current_stream(self.device))
        # snapshot state: warm-up steps on a side stream must not change the model
        keep = (self.eng.flat_params.clone(), self.eng.bn_running.clone(), self.eng.nbt.clone())
        with torch.cuda.stream(s):
            for _ in range(2):
                self._body()
        torch.cuda.current_stream(self.device).wait_stream(s)
        torch.cuda.synchronize(self.device)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph, stream=s):
            self._body()
        self.eng.flat_params.copy_(keep[0]); self.eng.bn_running.copy_(keep[1]); self.eng.nbt.copy_(keep[2])
        if self.fold_bn_sync:
            torch.mul(keep[1], float(self.world), out=self.eng.xchg[:2 * self.eng.C])
        check(lib.nrd_pack_weights(self.plan.handle, _stream_ptr()), "nrd_pack_weights")
        torch.cuda.synchronize(self.device)

    def __call__(self, imgs: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        self.imgs.copy_(imgs, non_blocking=True)
        self.labels.copy_(labels, non_blocking=True)
        return self.run()

    def run(self) -> torch.Tensor:
        """Replay on the static inputs already in `self.imgs` / `self.labels`."""
        if self.graph is not None:
            self.graph.replay()
        else:
            self._body()
        self.eng.weights_epoch += 1
        self.plan.packed_version = self.eng._weights_version()
        return self.loss
