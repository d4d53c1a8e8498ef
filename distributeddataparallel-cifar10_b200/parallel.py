"""`DistributedDataParallel` with the reference's call shape `DDP(model, device_ids=[rank],
output_device=rank)` (main.py:63) and torch DDP's semantics for this model:

  C1  constructor: rank 0's parameters and BN buffers are broadcast to every rank;
  C2  every forward: rank 0's running_mean/var (+ num_batches_tracked) overwrite the other ranks';
  C3  backward: gradients are averaged over ranks -- here as TWO buckets of the flat gradient
      vector (head = fc1/fc2, 86.6% of the bytes; body = conv/BN) on a side stream.

`torch.distributed` (NCCL over NVLink/NVSwitch on the box, gloo in CPU tests of the host logic)
is the transport; the averaging is fused into the optimizer kernel on the graph path
(`graph.TrainStep`) and done by ncclAvg / a divide on this eager path so `.grad` holds the mean
exactly as the reference's does.
"""
import torch
import torch.distributed as dist


class DistributedDataParallel(torch.nn.Module):
    def __init__(self, module, device_ids=None, output_device=None, process_group=None, broadcast_buffers=True):
        super().__init__()
        if not dist.is_initialized():
            raise RuntimeError("init_process_group() first (main.py:24)")
        self.module = module
        self.process_group = process_group
        self.world_size = dist.get_world_size(process_group)
        self.broadcast_buffers = broadcast_buffers
        eng = module.engine
        dev = next(module.parameters()).device
        if dev.type == "cuda":
            eng.adopt(dev)
            self._sync_module_states()
        object.__setattr__(module, "_nrd_bucket_hooks", (lambda: self._allreduce_bucket(1), lambda: self._allreduce_bucket(0)))

    # C1 ------------------------------------------------------------------------------------
    def _sync_module_states(self):
        eng = self.module.engine
        dist.broadcast(eng.flat_params, src=0, group=self.process_group)
        dist.broadcast(eng.bn_running, src=0, group=self.process_group)
        dist.broadcast(eng.nbt, src=0, group=self.process_group)
        eng.weights_epoch += 1

    # C2 ------------------------------------------------------------------------------------
    def _sync_buffers(self):
        eng = self.module.engine
        if self.world_size > 1 and self.broadcast_buffers and eng.bn_running is not None:
            dist.broadcast(eng.bn_running, src=0, group=self.process_group)
            # num_batches_tracked advances by n_blocks on every rank each forward: identical by construction

    def forward(self, *inputs, **kwargs):
        if self.module.training and torch.is_grad_enabled():
            self._sync_buffers()
        return self.module(*inputs, **kwargs)

    # C3 ------------------------------------------------------------------------------------
    def _allreduce_bucket(self, which: int, average: bool = True):
        """Mean all-reduce of one bucket of the flat gradient vector (0 = body, 1 = head) on the CURRENT stream: the engine
        calls this right after the kernels that produce the bucket, on the stream they run on, so the head bucket's
        exchange overlaps the whole body backward (the reference's single bucket fires only after the last gradient)."""
        if self.world_size == 1:
            return
        eng = self.module.engine
        lo, hi = eng.bucket_slices()[which]
        bucket = eng.flat_grads[lo:hi]
        if not average:
            dist.all_reduce(bucket, op=dist.ReduceOp.SUM, group=self.process_group)
        elif dist.get_backend(self.process_group) == "nccl":
            dist.all_reduce(bucket, op=dist.ReduceOp.AVG, group=self.process_group)
        else:
            dist.all_reduce(bucket, op=dist.ReduceOp.SUM, group=self.process_group)
            bucket.div_(self.world_size)


def flat_allreduce_mean_cpu(flat: torch.Tensor, n_body: int, group=None):
    """The bucketing/averaging host logic on a plain tensor (exercised by the gloo world_size-2 test
    on CPU): head bucket first, then body, SUM then divide."""
    W = dist.get_world_size(group)
    for lo, hi in ((n_body, flat.numel()), (0, n_body)):
        b = flat[lo:hi]
        dist.all_reduce(b, op=dist.ReduceOp.SUM, group=group)
        b.div_(W)
    return flat
