"""Whole-step CUDA graph: forward, loss, backward, gradient exchange, fused mean+SGD and weight
re-pack replayed as ONE graph launch (the ~70 us roofline step cannot afford ~60 separate
launches from Python).  This is main.py:34-39 for one rank."""
import os
from typing import Optional

import torch
import torch.distributed as dist

from . import _lib
from ._lib import lib, check, OPT_RESERVE_SMS
from .engine import _stream_ptr, _Plan


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
        # two input slots: `iterate()` fills one over PCIe while the step reads the other; slot 0 is the default
        self.inputs = [(torch.zeros(batch_size, 3, 32, 32, dtype=torch.float32, device=dev),
                        torch.zeros(batch_size, dtype=torch.int64, device=dev)) for _ in range(2)]
        self.imgs, self.labels = self.inputs[0]
        self.logits = torch.empty(batch_size, 10, dtype=torch.float32, device=dev)
        self.loss = torch.zeros((), dtype=torch.float32, device=dev)
        # a plan of its OWN (not the engine's per-batch-size cache): the exchange wiring, the BN-sync slot and the SM
        # reservation set below must not leak into eager steps on the same model and batch size
        if self.eng.device != dev or not self.eng._adopted():
            self.eng.adopt(dev)
        self.plan = _Plan(self.eng, batch_size)
        if self.eng.conv_impl is not None:
            check(lib.nrd_plan_set_conv_impl(self.plan.handle, self.eng.conv_impl), "nrd_plan_set_conv_impl")
        self.eng.ensure_packed(self.plan)
        self._exchange = None
        self._tails = {}                        # TrainSteps for batches of other sizes (the ragged tail of drop_last=False)
        # DDP on the graph path:
        # (1) gradients are exchanged over NVLink peer memory by our own kernels (nrd_plan_set_exchange: the producing
        #     kernels store into every rank's buffer, the optimizer kernel sums the world's slots in rank order; the head
        #     bucket leaves on the side stream under the body backward).  Without peer access (or beyond 8 ranks) the two
        #     buckets are NCCL all-reduces, the head one on the side stream;
        # (2) rank 0's BN running stats ride that exchange of the PREVIOUS step instead of a broadcast in front of every
        #     forward (same values: the buffers do not change in between);
        # (3) the persistent kernels leave a few SMs free so the head bucket's all-reduce really runs beside them.
        self.p2p = False
        self.fold_bn_sync = False
        self.slot_scale = 1.0
        if self.world > 1:
            rank = dist.get_rank(self.ddp.process_group)
            if os.environ.get("NRD_P2P", "1") != "0" and self.world <= 8:
                self.p2p = self._setup_exchange(rank)
            pow2 = (self.world & (self.world - 1)) == 0
            self.fold_bn_sync = (self.ddp.broadcast_buffers and (self.p2p or pow2) and os.environ.get("NRD_FOLD_BN_SYNC", "1") != "0")
            if self.fold_bn_sync:
                # all-reduce: rank 0 contributes world * stats (exact for powers of two); peer exchange: the stats themselves
                self.slot_scale = 1.0 if self.p2p else float(self.world)
                check(lib.nrd_plan_set_bn_sync(self.plan.handle, self.eng.xchg.data_ptr(), self.slot_scale if rank == 0 else 0.0,
                                               1.0 / self.slot_scale), "nrd_plan_set_bn_sync")
                self.resync_buffers()
            # the head bucket's kernel (NCCL all-reduce or our push kernel) has to fit beside the persistent kernels
            reserve = self._default_reserve(batch_size) if self.p2p else 8
            check(lib.nrd_plan_set_option(self.plan.handle, OPT_RESERVE_SMS, int(os.environ.get("NRD_RESERVE_SMS", str(max(reserve, 1))))),
                  "nrd_plan_set_option")
        else:
            # single rank: the side stream (head weight gradients, head optimizer step) and the programmatically launched successors
            # of the persistent kernels need SMs to run UNDER them -- with all 148 SMs holding a persistent CTA they queue behind
            # the backward stack.  Measured at B = 256: 195.0 us with 0 reserved SMs, 188.5 with 2, 184.3-184.7 with 4-16.
            # Only where the persistent kernels run (bf16, C = 32): the per-block kernels of the other configurations are
            # throughput-bound and leave the side stream its turn between launches (C = 64, B = 512: 2313 us with 0, 2446 with 8).
            persistent = self.eng.C == 32 and self.eng.dtype_code == _lib.NRD_BF16
            reserve = self._default_reserve(batch_size) if persistent else 0
            check(lib.nrd_plan_set_option(self.plan.handle, OPT_RESERVE_SMS, int(os.environ.get("NRD_RESERVE_SMS", str(reserve)))),
                  "nrd_plan_set_option")
        # optimizer.step() rides the backward (nrd_plan_set_fused_update): the head bucket is updated on the side stream under
        # the body backward, the body bucket by the kernel that finishes its gradients -- no optimizer launch at the end
        self.fused_update = os.environ.get("NRD_FUSED_UPDATE", "1") != "0" and os.environ.get("NRD_DIAG_NO_EXCHANGE") != "1" and (self.world == 1 or self.p2p)
        if self.fused_update:
            check(lib.nrd_plan_set_fused_update(self.plan.handle, self.lr, 1.0 / self.world), "nrd_plan_set_fused_update")
        # main.py:41 `loss.item()`: the head kernel also writes every step's loss into a small ring in pinned host memory and then
        # the step count (nrd_plan_set_loss_mirror), so reading it needs no copy node and no stream synchronisation
        self._mirror = torch.zeros(_lib.LOSS_RING + 2, dtype=torch.float32).pin_memory()      # ring | seq (as int32 view)
        self._mirror_seq = self._mirror[_lib.LOSS_RING:].view(torch.int32)
        self._mirror_np = self._mirror.numpy(); self._mirror_seq_np = self._mirror_seq.numpy()
        self.steps_run = 0
        if os.environ.get("NRD_LOSS_MIRROR", "1") != "0":
            check(lib.nrd_plan_set_loss_mirror(self.plan.handle, self._mirror.data_ptr(), self._mirror_seq.data_ptr()), "nrd_plan_set_loss_mirror")
            self.loss_mirror = True
        else:
            self.loss_mirror = False
        if os.environ.get("NRD_TIMELINE") == "1":   # measurement aid: event marks behind every main-stream kernel (see timeline())
            check(lib.nrd_plan_set_option(self.plan.handle, _lib.OPT_TIMELINE, 1), "nrd_plan_set_option")
        self.graphs = [None, None]              # one captured step per input slot (slot 1 on first use)
        self.use_graph = use_graph
        self.kernel_launches = None
        if use_graph:
            self._capture(0)

    def _default_reserve(self, batch_size: int) -> int:
        """SMs the persistent kernels leave to the side stream: 8, or fewer when 8 would put one more image on some CTA (the
        kernels' time goes with the LARGEST number of images per CTA).  Measured on one B200 (us/step with 0 / 1 / 8 reserved SMs):
        B = 256: 195.0 / 190.0 / 184.3; 512: 362.9 / 355.1 / 349.2; 768: 502.6 / 488.3 / 474.1 (same images per CTA either way);
        B = 1024: 571.8 / 566.9 / 589.2 (7 images per CTA on 147-148 CTAs, 8 on 140)."""
        sms = torch.cuda.get_device_properties(self.eng.device).multi_processor_count
        per_cta = -(-batch_size // sms)                       # images on the busiest CTA with every SM in use
        return max(0, min(8, sms - -(-batch_size // per_cta)))

    def _setup_exchange(self, rank: int) -> bool:
        """Exchange buffers for the body bucket: allocate ours, swap CUDA IPC handles over the process group, map the peers'.
        Returns False (on every rank alike) when any rank cannot, e.g. no peer access between the devices."""
        import ctypes as C
        W, dev = self.world, self.device
        nbytes = lib.nrd_exchange_bytes(self.eng.C, W)
        ptr, handle = C.c_void_p(), (C.c_ubyte * 64)()
        ok = nbytes > 0 and lib.nrd_exchange_alloc(dev.index, nbytes, C.byref(ptr), handle) == 0
        handles = [None] * W
        dist.all_gather_object(handles, bytes(handle) if ok else None, group=self.ddp.process_group)
        ptrs = (C.c_void_p * W)()
        opened = []
        if all(h is not None for h in handles):
            for r in range(W):
                if r == rank:
                    ptrs[r] = ptr.value
                    continue
                q = C.c_void_p()
                hb = (C.c_ubyte * 64).from_buffer_copy(handles[r])
                if lib.nrd_exchange_open(dev.index, hb, C.byref(q)) != 0:
                    ok = False
                    break
                ptrs[r] = q.value
                opened.append(q.value)
        else:
            ok = False
        flag = torch.tensor([1 if ok else 0], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=self.ddp.process_group)
        if int(flag) == 0:
            for q in opened:
                lib.nrd_exchange_release(q, 1)
            if ptr.value:
                lib.nrd_exchange_release(ptr, 0)
            return False
        check(lib.nrd_plan_set_exchange(self.plan.handle, rank, W, ptrs), "nrd_plan_set_exchange")
        self._exchange = (ptr, opened)          # kept mapped for the life of the process (captured graphs point into them)
        return True

    def timeline(self):
        """[(kernel name, microseconds)] of the main stream in the last step, when the step was built with NRD_TIMELINE=1
        (nrd_plan_timeline: event-record nodes inside the captured graph)."""
        import ctypes as C
        torch.cuda.synchronize(self.device)
        us = (C.c_float * 40)(); names = (C.c_char_p * 40)()
        n = lib.nrd_plan_timeline(self.plan.handle, 40, us, names)
        if n < 0:
            raise _lib.NrdError("nrd_plan_timeline: " + _lib.last_error())
        return [(names[i].decode(), float(us[i])) for i in range(n)]

    def check_status(self, clear: bool = True):
        """Raise if a kernel of this step clamped or saw non-finite BatchNorm sums since the last check (nrd_plan_status:
        the persistent kernels' fixed-point statistics have a finite range).  Synchronises the current stream; call it
        where the loop reads the loss on the host anyway (main.py:41,44)."""
        import ctypes as C
        flags = C.c_uint32(0)
        rc = lib.nrd_plan_status(self.plan.handle, C.byref(flags), 1 if clear else 0, _stream_ptr())
        if rc != 0:
            raise _lib.NrdError(f"training step status {flags.value:#x}: {_lib.last_error()}")

    def close(self):
        """Drop the captured graphs and release the peer-memory exchange buffers (ours and the mappings of the peers').
        Every rank must call it at the same point (a peer may still be storing into our buffer until its last step is done)."""
        self.graphs = [None, None]
        for t in getattr(self, "_tails", {}).values():
            t.close()
        ex, self._exchange = getattr(self, "_exchange", None), None
        if ex is not None:
            torch.cuda.synchronize(self.device)
            lib.nrd_plan_set_exchange(self.plan.handle, 0, 0, None)
            own, opened = ex
            for q in opened:
                lib.nrd_exchange_release(q, 1)
            lib.nrd_exchange_release(own, 0)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def resync_buffers(self):
        """Rank 0's BN running stats to every rank, and into the slot the next forward starts from.  Called at construction;
        call it again after changing the buffers outside the step (load_state_dict)."""
        if self.world > 1 and self.ddp.broadcast_buffers:
            self.ddp._sync_buffers()
            if self.fold_bn_sync:
                torch.mul(self.eng.bn_running, self.slot_scale, out=self.eng.xchg[:2 * self.eng.C])

    # the step body, launched on the current stream
    @property
    def graph(self):
        return self.graphs[0]

    @graph.setter
    def graph(self, g):                     # `step.graph = None` drops every captured step (before communicator teardown)
        self.graphs = [g, None]

    def _body(self, slot: int = 0):
        eng, pl, st = self.eng, self.plan, _stream_ptr()
        imgs, labels = self.inputs[slot]
        if self.world > 1 and not self.fold_bn_sync:
            self.ddp._sync_buffers()                                   # C2 (folded into the body bucket otherwise)
        check(lib.nrd_forward_loss(pl.handle, imgs.data_ptr(), labels.data_ptr(), 1.0, self.logits.data_ptr(), self.loss.data_ptr(),
                                   pl.dlogits.data_ptr(), pl.loss_rows.data_ptr(), st), "nrd_forward_loss")
        # backward: head weight grads (+ their bucket's all-reduce) on a side stream beside the body backward (C3)
        # peer-memory exchange: both buckets travel inside nrd_backward_head / nrd_backward_body, no collective here
        sum_head = (lambda: self.ddp._allreduce_bucket(1, average=False)) if (self.world > 1 and not self.p2p) else None
        sum_body = None
        if self.world > 1 and not self.p2p:
            sum_body = ((lambda: dist.all_reduce(eng.body_bucket_with_bn_slot(), op=dist.ReduceOp.SUM, group=self.ddp.process_group))
                        if self.fold_bn_sync else (lambda: self.ddp._allreduce_bucket(0, average=False)))
        if os.environ.get("NRD_DIAG_NO_EXCHANGE") == "1":          # timing diagnosis only: every rank trains alone
            sum_head = sum_body = None
        eng.backward(pl, pl.dlogits, sum_head, sum_body)
        # fused mean + SGD (K16) and the re-derived packed weight copies: inside the backward's own kernels, or one launch here
        if not self.fused_update:
            check(lib.nrd_sgd_update_pack(pl.handle, self.lr, 1.0 / self.world, st), "nrd_sgd_update_pack")

    def _capture(self, slot: int = 0):
        s = torch.cuda.Stream(device=self.device)
        s.wait_stream(torch.cuda.current_stream(self.device))
        # snapshot state: warm-up steps on a side stream must not change the model
        keep = (self.eng.flat_params.clone(), self.eng.bn_running.clone(), self.eng.nbt.clone())
        with torch.cuda.stream(s):
            for _ in range(2):
                self._body(slot)
        torch.cuda.current_stream(self.device).wait_stream(s)
        torch.cuda.synchronize(self.device)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            self._body(slot)
        self.graphs[slot] = g
        self.eng.flat_params.copy_(keep[0]); self.eng.bn_running.copy_(keep[1]); self.eng.nbt.copy_(keep[2])
        if self.fold_bn_sync:
            torch.mul(keep[1], self.slot_scale, out=self.eng.xchg[:2 * self.eng.C])
        check(lib.nrd_pack_weights(self.plan.handle, _stream_ptr()), "nrd_pack_weights")
        torch.cuda.synchronize(self.device)
        if self.loss_mirror:
            self.steps_run = int(self._mirror_seq_np[0])      # the warm-up steps above were counted by the device as well

    def __call__(self, imgs: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        self.imgs.copy_(imgs, non_blocking=True)
        self.labels.copy_(labels, non_blocking=True)
        return self.run()

    def run(self, slot: int = 0) -> torch.Tensor:
        """Replay on the static inputs already in `self.inputs[slot]` (slot 0 = `self.imgs` / `self.labels`)."""
        if self.use_graph and self.graphs[slot] is None and self.graphs[0] is not None:
            self._capture(slot)
        # load_state_dict / in-place edits since the last step: re-derive the packed weight copies first (host-side version
        # compare; a no-op in the steady state because the step's own update re-packs them)
        self.eng.ensure_packed(self.plan)
        if self.graphs[slot] is not None:
            self.graphs[slot].replay()
        else:
            self._body(slot)
        self.eng.weights_epoch += 1
        self.plan.packed_version = self.eng._weights_version()
        self.steps_run += 1
        return self.loss

    def host_loss(self, n: int, spin_s: float = 30.0) -> float:
        """The loss of the n-th step run since this TrainStep was (re)captured (1-based), read from the pinned-host ring the head
        kernel writes: no copy, no stream synchronisation; spins (briefly) only if that step's head kernel has not run yet.
        Valid while fewer than LOSS_RING further steps have been launched."""
        import time
        seq = self._mirror_seq_np
        if int(seq[0]) < n:
            t0 = time.perf_counter()
            while int(seq[0]) < n:
                if time.perf_counter() - t0 > spin_s:
                    torch.cuda.synchronize(self.device)
                    if int(seq[0]) < n:
                        raise RuntimeError(f"loss of step {n} never reached the host mirror (count {int(seq[0])})")
        return float(self._mirror_np[(n - 1) % _lib.LOSS_RING])

    def iterate(self, loader, lookahead: int = 0):
        """One pass over `loader` (HOST batches `(imgs, labels)`; pinned memory makes the copies asynchronous), one step per
        batch: the loop of main.py:31-41.  Yields the step's loss as a host value (the reference's per-step `loss.item()`,
        main.py:41): a Python float read from the pinned-host ring the head kernel writes (no copy node, no stream
        synchronisation), or, with NRD_LOSS_MIRROR=0, a pinned host scalar tensor filled by a D2H copy.  Order per step: launch the step
        on the slot that is ready, THEN fetch the next host batch and start its copy into the other slot on a copy stream
        -- loader work and PCIe traffic run beside the step's kernels instead of in front of them.

        `lookahead=1`: step t's loss is handed out after step t+1 has been LAUNCHED (same values, same order, one step
        later), so the GPU never waits for the host between steps; the price is that breaking out of the loop at step t
        leaves step t+1 applied as well.

        A batch of another size -- the ragged last batch of the reference's `drop_last=False` loader (main.py:61: 10 samples at
        8 ranks) -- runs through a second TrainStep of that size on the same model (built on first use, kept), after the
        steps in flight have been handed out; every rank must see it at the same point of the epoch, which
        DistributedSampler's equal-length index lists guarantee."""
        dev = self.device
        main = torch.cuda.current_stream(dev)
        if self.use_graph and self.graphs[1] is None:
            self._capture(1)
        if getattr(self, "_copy_stream", None) is None:
            self._copy_stream = torch.cuda.Stream(device=dev)
            self._h2d_chunks = int(os.environ.get("NRD_H2D_CHUNKS", "1"))
            self._ready = [torch.cuda.Event(), torch.cuda.Event()]
            self._consumed = [torch.cuda.Event(), torch.cuda.Event()]
            self._h_loss = [torch.zeros((), dtype=torch.float32).pin_memory() for _ in range(2)]
            self._loss_done = [torch.cuda.Event(), torch.cuda.Event()]
            for e in self._consumed:
                e.record(main)
        copy = self._copy_stream

        def fits(batch):
            imgs = batch[0]
            if imgs.dim() != 4 or tuple(imgs.shape[1:]) != (3, 32, 32) or imgs.shape[0] < 1:
                raise ValueError(f"expected batches of shape [b,3,32,32], got {tuple(imgs.shape)}")
            return imgs.shape[0] == self.B

        def run_other_size(batch):
            """a batch of another size (the ragged tail): its own TrainStep on the same model; returns the loss on the host"""
            imgs, labels = batch
            tail = self._tails.get(imgs.shape[0])
            if tail is None:
                tail = self._tails[imgs.shape[0]] = TrainStep(self.ddp if self.ddp is not None else self.module, imgs.shape[0], lr=self.lr,
                                                              use_graph=self.use_graph)
            loss = tail(imgs.to(dev, non_blocking=True), labels.to(dev, non_blocking=True))
            return tail.host_loss(tail.steps_run) if tail.loss_mirror else float(loss)

        def issue(slot, batch):
            imgs, labels = batch
            if imgs.is_cuda or labels.is_cuda:
                # device-resident batch (e.g. SyntheticCifar10Loader): its producer runs on the current stream, so the copy
                # goes there too (nothing to overlap: there is no PCIe transfer)
                self.inputs[slot][0].copy_(imgs, non_blocking=True)
                self.inputs[slot][1].copy_(labels, non_blocking=True)
                self._ready[slot].record(main)
                return
            copy.wait_event(self._consumed[slot])        # the step that last read this slot
            with torch.cuda.stream(copy):
                nch = self._h2d_chunks
                if nch > 1 and imgs.shape[0] % nch == 0:
                    # several ~1 MB copies: measured faster than one 3 MB copy on a busy host (tools/h2d_rate.py)
                    step = imgs.shape[0] // nch
                    for c in range(nch):
                        self.inputs[slot][0][c * step:(c + 1) * step].copy_(imgs[c * step:(c + 1) * step], non_blocking=True)
                else:
                    self.inputs[slot][0].copy_(imgs, non_blocking=True)
                self.inputs[slot][1].copy_(labels, non_blocking=True)
                self._ready[slot].record(copy)

        it = iter(loader)
        nxt = next(it, None)
        if nxt is None:
            return
        slot = 0
        k, pending = 0, None
        while nxt is not None and not fits(nxt):            # (a loader that starts with an odd-sized batch)
            yield run_other_size(nxt)
            nxt = next(it, None)
        if nxt is None:
            return
        issue(slot, nxt)
        while nxt is not None:
            main.wait_event(self._ready[slot])
            self.run(slot)
            self._consumed[slot].record(main)
            if self.loss_mirror:
                tok = self.steps_run                                  # this step's number in the device's count
                fetch = self.host_loss
            else:
                self._h_loss[k].copy_(self.loss, non_blocking=True)
                self._loss_done[k].record(main)
                tok = k

                def fetch(j):
                    self._loss_done[j].synchronize()
                    return self._h_loss[j]
            nxt = next(it, None)
            odd = []
            while nxt is not None and not fits(nxt):        # ragged batch(es): after the steps in flight
                odd.append(nxt)
                nxt = next(it, None)
            if nxt is not None and not odd:
                issue(slot ^ 1, nxt)
            if odd:
                if pending is not None:
                    yield fetch(pending)
                    pending = None
                yield fetch(tok)
                for ob in odd:
                    yield run_other_size(ob)
                if nxt is not None:
                    issue(slot ^ 1, nxt)
            elif lookahead and nxt is not None:
                if pending is not None:
                    yield fetch(pending)
                pending = tok
            else:
                if pending is not None:
                    yield fetch(pending)
                    pending = None
                yield fetch(tok)
            slot ^= 1
            k ^= 1
