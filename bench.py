#!/usr/bin/env python
"""bench.py -- NetResDeep data-parallel training step on B200: train images/sec.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
  (N > 1: launched by `python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...`)

A "step" is main.py:34-39 for one rank on one synthetic CIFAR-shaped batch (forward, loss, backward,
gradient exchange, SGD update), replayed as one CUDA graph.  Workload = BASELINE.json configs[1]
(NetResDeep C=32 / 10 blocks, per-rank batch 256, bf16 storage + fp32 accumulate), weak scaling.

  value     images/s, whole job, batch resident in HBM, device-timed (CUDA events per step, summed;
            max over ranks); L2 is flushed between steps (256 MiB write, untimed).
  e2e       same metric through the public API from PINNED HOST buffers: per step an H2D copy of the
            batch (TrainStep.iterate: issued one step ahead on a copy stream), the step, and a D2H read
            of the loss with a host sync (wall clock around K steps, max over ranks).
  roofline  dominant kernel class of the step, timed alone with CUDA events on the launching stream
            (nrd_plan_run_op), algorithmic HBM bytes / duration against MEASURED_PEAKS.json.
  cpu_baseline  the reference CPU step (main_no_ddp.py:45-50, B=64, fp32) timed on this box's host cores (rank 0,
            N=1 only): the UNMODIFIED reference NetResDeep from baseline/_ref when that git-ignored copy travelled with
            the snapshot (kind "reference"), else the oracle's nn.Module port of it (kind "port").
  parity    N > 1 only, outside every timed region: two steps of the SAME graph path that is timed (peer-memory or NCCL
            exchange, bf16) on per-rank batches, gathered on rank 0 and compared with the oracle's DDP simulation
            (rounding-matched), one fp32 step against the fp32 oracle, and after the timed loop a bitwise comparison of
            every rank's parameters.
  gpu_eager_reference  secondary, outside every timed region: the reference model + stock torch.optim.SGD /
            nn.CrossEntropyLoss under stock torch eager on the same GPU (real torch DDP when N > 1), same batch shape.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FALLBACK_HBM_GBS = 6650.0       # B200_PROFILING.md fallback when MEASURED_PEAKS.json is absent


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), float(d.get("bf16_tflops_sustained", 1402.0)), float(d.get("bf16_tflops", 1590.0)), "measured"
    except Exception:
        return FALLBACK_HBM_GBS, 1400.0, 1590.0, "fallback"


def source_sha16():
    """sha256 over the kernel sources: ties profiles/ncu_traffic.json (written by tools/ncu_traffic.py from an ncu
    --set full capture) to the build it was captured from."""
    import hashlib
    d = os.path.join(ROOT, "distributeddataparallel-cifar10_b200", "csrc")
    h = hashlib.sha256()
    for f in sorted(os.listdir(d)):
        if f.endswith((".cu", ".cuh")):
            h.update(f.encode()); h.update(open(os.path.join(d, f), "rb").read())
    return h.hexdigest()[:16]


def ncu_traffic(kernel_key, B, C, NB, dtype):
    """DRAM bytes per launch of the dominant kernel from the tracked ncu capture -- only when that capture was taken from
    THIS source and THIS configuration; otherwise None (never a stale constant)."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            d = json.load(f)
        if d.get("source_sha16") != source_sha16() or d.get("config") != {"B": B, "C": C, "NB": NB, "dtype": dtype}:
            return None
        k = d["kernels"].get(kernel_key)
        return int(k["dram_read_bytes"] + k["dram_write_bytes"]) if k else None
    except Exception:
        return None


def reference_netresdeep():
    """The unmodified reference NetResDeep class from the git-ignored baseline/_ref copy, or None."""
    path = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.exists(os.path.join(path, "model", "resnet.py")):
        return None
    sys.path.insert(0, path)
    try:
        from model.resnet import NetResDeep
        return NetResDeep
    except Exception:
        return None
    finally:
        sys.path.remove(path)


def bytes_per_image(C, NB, s):
    """BASELINE.md section 4 / SURVEY 8(d): compulsory layer-boundary HBM bytes per image."""
    T = 256 * C
    return (2 * 3072 * s + (T * s + T) + NB * 4 * T * s + (T * s + 64 * C * s + 64 * C) + 3 * 64 * C * s
            + (64 * C * s + 64 * C + T * s) + NB * 6 * T * s + (T * s + T + 3072 * s))


def flops_per_image(C, NB):
    return 2 * 55296 * C + 3 * NB * 4608 * C * C + 3 * 4096 * C + 3 * 640


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "25",
                                          "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = sorted(float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit())
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 9 for n, v in zip(names, r[5:9]) if v.lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


def cpu_reference(batch, C, NB, steps, warmup):
    """The reference CPU step (main_no_ddp.py:45-50 restated around the model; that file trains at import time and needs
    the CIFAR files, so it cannot be imported) on all host cores."""
    from oracle import netresdeep_oracle as O
    import torch
    cores = len(os.sched_getaffinity(0))
    torch.set_num_threads(cores)
    Ref = reference_netresdeep()
    if Ref is None:
        best, med, thr = O.time_cpu_step(batch=batch, n_chans1=C, n_blocks=NB, steps=steps, warmup=warmup, threads=cores)
        return {"best_s": best, "median_s": med, "cores": thr, "batch": batch, "kind": "port"}
    torch.manual_seed(0)
    m = Ref(n_chans1=C, n_blocks=NB)                                   # model/resnet.py, unmodified
    opt = torch.optim.SGD(m.parameters(), lr=1e-2)                     # main_no_ddp.py:37
    loss_fn = torch.nn.CrossEntropyLoss()                              # main_no_ddp.py:38
    gen = torch.Generator().manual_seed(1234)
    x = torch.randn(batch, 3, 32, 32, generator=gen); y = torch.randint(0, 10, (batch,), generator=gen)
    ts = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        out = m(x); loss = loss_fn(out, y)                             # :45-46
        opt.zero_grad(); loss.backward(); opt.step()                   # :48-50
        _ = loss.item()                                                # :52
        if i >= warmup:
            ts.append(time.perf_counter() - t0)
    ts.sort()
    return {"best_s": ts[0], "median_s": ts[len(ts) // 2], "cores": torch.get_num_threads(), "batch": batch, "kind": "reference"}


def gpu_eager_reference(B, C, NB, dev, world, steps=20, warmup=5):
    """Secondary number, never inside a timed region of ours: the reference model and its literal loop
    (main.py:27-41) under stock torch eager on this GPU -- cuDNN/ATen kernels, torch DDP over NCCL when world > 1."""
    import torch
    import torch.distributed as dist
    Ref = reference_netresdeep()
    kind = "reference"
    if Ref is None:
        from oracle import netresdeep_oracle as O
        p, bufs = O.init_params(C, NB, seed=0)
        m = O.TorchModulePort(p, bufs, NB).to(dev); kind = "port"
    else:
        torch.manual_seed(0)
        m = Ref(n_chans1=C, n_blocks=NB).to(dev)
    if world > 1:
        m = torch.nn.parallel.DistributedDataParallel(m, device_ids=[dev.index], output_device=dev.index)     # main.py:63
    opt = torch.optim.SGD(m.parameters(), lr=1e-2); loss_fn = torch.nn.CrossEntropyLoss()
    gen = torch.Generator().manual_seed(1234)
    x = torch.randn(B, 3, 32, 32, generator=gen).to(dev); y = torch.randint(0, 10, (B,), generator=gen).to(dev)
    evs = []
    for i in range(warmup + steps):
        a, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        out = m(x); loss = loss_fn(out, y); opt.zero_grad(); loss.backward(); opt.step()
        b_.record()
        if i >= warmup:
            evs.append((a, b_))
    torch.cuda.synchronize(dev)
    ms = sorted(a.elapsed_time(b_) for a, b_ in evs)
    t = torch.tensor([ms[len(ms) // 2]], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    del m, opt
    return {"ms_per_step": float(t), "images_per_sec": world * B / (float(t) / 1e3), "kind": kind, "dtype": "fp32",
            "cudnn_allow_tf32": bool(torch.backends.cudnn.allow_tf32), "matmul_precision": torch.get_float32_matmul_precision(),
            "steps": steps, "timing": "CUDA events per step, median, max over ranks; no L2 flush; loss not read back",
            "ddp": "torch.nn.parallel.DistributedDataParallel" if world > 1 else None}


def ddp_parity(nrd, dist, dev, rank, world, B, C, NB, exchange_env):
    """N > 1: the timed graph path against the oracle's DDP simulation, outside every timed region.  bf16: two steps at the
    bench's own per-rank batch (rounding-matched oracle; step 2 also proves rank 0's BN buffers reach every rank); fp32: one
    step at a small batch against the plain fp32 oracle.  Everything is gathered on rank 0, which runs the CPU oracle.
    lr = 1e-3 here: at the reference's 1e-2 the second step of this net is inside the loss spike of SURVEY fact 9, where
    the fp32 oracle's own run-to-run noise already reaches 1e-3 on the body gradients; the exchange, the mean and the update
    under test do not depend on the value of lr."""
    PLR = 1e-3
    import torch
    res = {}
    for dtype, Bp, nsteps in (("bf16", B, 2), ("fp32", min(B, 32), 1)):
        torch.manual_seed(100 + rank)                        # ranks start DIFFERENT (main.py:62); the wrapper broadcasts rank 0's
        m = nrd.NetResDeep(C, NB, dtype=dtype).to(dev)
        ddp = nrd.DistributedDataParallel(m, device_ids=[dev.index], output_device=dev.index)
        step = nrd.TrainStep(ddp, Bp, lr=PLR)
        eng = m.engine
        w0 = eng.flat_params.detach().clone()
        loader = nrd.SyntheticCifar10Loader(batch_size=Bp, device=dev, rank=rank, seed=4321)
        rec = []
        for s in range(nsteps):
            imgs, labels = loader.batch(s)
            loss = step(imgs, labels)
            torch.cuda.synchronize(dev)
            rec.append(dict(imgs=imgs.clone(), labels=labels.clone(), logits=step.logits.clone(), loss=loss.clone().reshape(1),
                            gsum=eng.flat_grads.detach().clone(), bn=eng.bn_running.detach().clone(), w=eng.flat_params.detach().clone()))
        def gather(t):
            out = [torch.empty_like(t) for _ in range(world)]
            dist.all_gather(out, t.contiguous())
            return [o.cpu() for o in out]
        got = [{k: gather(v) for k, v in r.items()} for r in rec]
        w0s = gather(w0)
        mode = "p2p" if step.p2p else "nccl"
        step.close(); del step
        if rank != 0:
            continue
        from oracle import netresdeep_oracle as O
        torch.set_num_threads(len(os.sched_getaffinity(0)))
        off = eng.offsets
        shapes = {"conv1.weight": (C, 3, 3, 3), "conv1.bias": (C,), "rb.conv.weight": (C, C, 3, 3), "rb.bn.weight": (C,), "rb.bn.bias": (C,),
                  "fc1.weight": (32, 64 * C), "fc1.bias": (32,), "fc2.weight": (10, 32), "fc2.bias": (10,)}
        p = {k: w0s[0][off[i]:off[i + 1]].view(shapes[k]).clone() for i, k in enumerate(O.PARAM_ORDER)}
        b0 = {"running_mean": torch.zeros(C), "running_var": torch.ones(C), "num_batches_tracked": torch.zeros((), dtype=torch.int64)}
        bufs = [dict((k, v.clone()) for k, v in b0.items()) for _ in range(world)]
        rel = lambda a, b_: float((a.double() - b_.double()).norm() / (b_.double().norm() + 1e-30))
        worst = {"logits": 0.0, "loss": 0.0, "head_grads": 0.0, "bn_grads": 0.0, "conv_grads": 0.0, "bn_buffers": 0.0, "update": 0.0}
        flat = lambda d: torch.cat([d[k].flatten() for k in O.PARAM_ORDER])
        init_bcast = all(torch.equal(w0s[0], w) for w in w0s)
        sim = O.ddp_train_step_rm if dtype == "bf16" else O.ddp_train_step
        per_step = []
        for s in range(nsteps):
            g = got[s]
            before = dict(worst)
            for k in worst:
                worst[k] = 0.0
            p_new, logits, losses, gavg = sim(p, bufs, g["imgs"], g["labels"], NB, PLR)
            for r in range(world):
                worst["logits"] = max(worst["logits"], rel(g["logits"][r], logits[r]))
                worst["loss"] = max(worst["loss"], abs(float(g["loss"][r]) - float(losses[r])) / float(losses[r]))
                bn_ref = torch.cat([bufs[r]["running_mean"], bufs[r]["running_var"]])
                worst["bn_buffers"] = max(worst["bn_buffers"], rel(g["bn"][r], bn_ref))
                for i, k in enumerate(O.PARAM_ORDER):
                    e = rel(g["gsum"][r][off[i]:off[i + 1]] / world, gavg[k].flatten())       # .grad holds the world's SUM on the graph path
                    key = "head_grads" if k.startswith("fc") else ("bn_grads" if ".bn." in k else "conv_grads")
                    worst[key] = max(worst[key], e)
                # the update this rank applied to its whole parameter vector (w_new - w_old = -lr * mean gradient) vs the oracle's
                w_prev = w0s[r] if s == 0 else got[s - 1]["w"][r]
                worst["update"] = max(worst["update"], rel(g["w"][r] - w_prev, flat(p_new) - flat(p)))
            per_step.append({k: float(f"{v:.3g}") for k, v in worst.items()})
            for k in worst:
                worst[k] = max(worst[k], before[k])
            p = p_new
        # bf16, step 1: the single-GPU envelopes of tests/test_gpu_parity_bench.py (profiles/r02_parity.md; conv1 included in
        # "conv"; head: room for two flipped fc1 ReLUs at ~6e-3 each).  Step 2 starts from weights that differ by lr * (gradient
        # noise), and re-rounding them to bf16 amplifies that: the rounding-matched oracle run in fp32 vs fp64 between identical
        # rounding points differs at its OWN second step by logits 4.7e-3, head 2e-2, BN 3e-2, conv 6-9e-2 (profiles/r02_parity.md),
        # so step 2 is held to 4x that and mainly proves the buffer broadcast and that the update reached the kernels' weights.
        lim = ({"logits": 2e-3, "loss": 2e-3, "head_grads": 1.5e-2, "bn_grads": 1e-2, "conv_grads": 4e-2, "bn_buffers": 1e-3, "update": 1.5e-2} if dtype == "bf16"
               else {"logits": 1e-5, "loss": 1e-5, "head_grads": 2e-5, "bn_grads": 5e-3, "conv_grads": 5e-3, "bn_buffers": 1e-5, "update": 1e-3})
        lim2 = {"logits": 2e-2, "loss": 5e-3, "head_grads": 8e-2, "bn_grads": 1.2e-1, "conv_grads": 3e-1, "bn_buffers": 1e-3, "update": 1e-1}
        ok = init_bcast and all(per_step[0][k] <= lim[k] for k in lim) and all(ps[k] <= lim2[k] for ps in per_step[1:] for k in lim2)
        res[dtype] = {"ok": bool(ok), "exchange": mode, "batch": Bp, "steps": nsteps, "limits_later_steps": lim2 if nsteps > 1 else None,
                      "init_broadcast": init_bcast, "max_rel_err": {k: float(f"{v:.3g}") for k, v in worst.items()}, "per_step": per_step, "limits": lim,
                      "oracle": "ddp_train_step_rm (bf16 rounding-matched)" if dtype == "bf16" else "ddp_train_step (fp32)"}
    return res if rank == 0 else None



def run_reference(args):
    """--impl reference: the reference's CPU path (oracle port of main_no_ddp.py:45-50 around the
    restated model; the reference is pure Python/torch and cannot be compiled or shipped) on the
    box's host cores, same metric/config.  Rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    r = cpu_reference(args.batch, args.chans, args.blocks, args.steps, args.warmup)
    val = args.batch / r["median_s"]
    sample = f"{args.steps} steps of B={args.batch} fp32 on CPU (median step {r['median_s'] * 1e3:.1f} ms, best {r['best_s'] * 1e3:.1f} ms)"
    line = {"impl": "reference", "metric": "train_images_per_sec", "value": val, "unit": "images/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["median_s"] * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"NetResDeep C={args.chans} NB={args.blocks} train step, batch {args.batch}, CPU", "device": "cpu"},
            "cpu_baseline": {"value": val, "unit": "images/s", "cores": r["cores"], "kind": r["kind"], "sample": sample},
            "e2e": {"value": val, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=256, help="per-rank batch (BASELINE configs[1])")
    ap.add_argument("--chans", type=int, default=32)
    ap.add_argument("--blocks", type=int, default=10)
    ap.add_argument("--dtype", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the N>1 DDP-vs-oracle check (A/B timing runs)")
    ap.add_argument("--no-eager-ref", action="store_true", help="skip the stock-torch-eager secondary number")
    args = ap.parse_args()
    if args.impl == "reference":
        if args.steps > 30:
            args.steps = 30                      # bounded sample: keeps the CPU arm within minutes
        return run_reference(args)

    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.gpus != world:
        if world == 1 and args.gpus > 1:
            sys.exit(f"--gpus {args.gpus} needs torchrun: python -m torch.distributed.run --nproc-per-node {args.gpus} bench.py --gpus {args.gpus}")
    if args.warmup < 3:
        args.warmup = 3
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        r = cpu_reference(64, args.chans, args.blocks, steps=30, warmup=5)           # BASELINE configs[0]
        cpu = {"value": 64 / r["median_s"], "unit": "images/s", "cores": r["cores"], "kind": r["kind"],
               "sample": f"30 steps of configs[0] (B=64, fp32, main_no_ddp.py step) after 5 warm-up; median {r['median_s'] * 1e3:.1f} ms, best {r['best_s'] * 1e3:.1f} ms/step"}

    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    import ddp_cifar10_b200 as nrd
    from ddp_cifar10_b200._lib import lib, check
    B, C, NB = args.batch, args.chans, args.blocks
    parity = None
    if world > 1 and not args.no_parity:
        parity = ddp_parity(nrd, dist, dev, rank, world, B, C, NB, os.environ.get("NRD_P2P", "1"))
        dist.barrier()
    torch.manual_seed(0)
    model = nrd.NetResDeep(C, NB, dtype=args.dtype).to(dev)
    wrapped = nrd.DistributedDataParallel(model, device_ids=[local], output_device=local) if world > 1 else model
    n0 = lib.nrd_launch_count()
    step = nrd.TrainStep(wrapped, B, lr=1e-2, use_graph=not args.no_graph)
    launches_per_step = (lib.nrd_launch_count() - n0 - 2) // 3 if not args.no_graph else None   # 2 warm-up bodies + 1 captured, +2 weight packs
    loader = nrd.SyntheticCifar10Loader(batch_size=B, device=dev, rank=rank)
    imgs, labels = loader.batch(0)
    step.imgs.copy_(imgs); step.labels.copy_(labels)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    for _ in range(args.warmup):
        step.run()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    if args.no_graph:
        n0 = lib.nrd_launch_count()
    barrier()
    for s, e in evs:
        flush.zero_()                      # L2 flush, outside the timed events
        s.record(); step.run(); e.record()
    barrier()
    if args.no_graph:
        launches_per_step = (lib.nrd_launch_count() - n0) // args.steps
    clocks = sampler.stop() if rank == 0 else None
    ms = sum(s.elapsed_time(e) for s, e in evs)
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t)
    value = world * B * args.steps / (ms_total / 1e3)
    loss_last = float(step.loss)
    if world > 1:
        # the one invariant the exchange must keep: after any number of steps every rank holds bit-identical parameters
        mine = model.engine.flat_params.detach().clone()
        allw = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(allw, mine)
        ranks_identical = all(torch.equal(allw[0], w) for w in allw)
        if parity is not None:
            parity["ranks_identical"] = bool(ranks_identical)
            parity["ddp_vs_oracle"] = bool(all(v["ok"] for k, v in parity.items() if isinstance(v, dict)))
        elif rank == 0:
            parity = {"ranks_identical": bool(ranks_identical), "ddp_vs_oracle": None}

    # ---- end to end from pinned host memory -------------------------------------------------
    h_imgs = imgs.cpu().pin_memory(); h_labels = labels.cpu().pin_memory()
    h_loss = torch.zeros((), dtype=torch.float32).pin_memory()
    class HostBatches:                                      # stands in for main.py:61's DataLoader (pinned host batches)
        def __init__(self, n): self.n = n
        def __len__(self): return self.n
        def __iter__(self): return ((h_imgs, h_labels) for _ in range(self.n))

    def epoch(n):
        # main.py:31-41 as TrainStep.iterate: launch step t, then start batch t+1's H2D copy beside it, then read the loss of
        # step t-1 (look-ahead 1: every step's loss still reaches the host, in order, one launch later)
        tot = 0.0
        for loss in step.iterate(HostBatches(n), lookahead=1):
            tot += float(loss)                                  # host value of the step's loss (main.py:41)
        return tot

    # Warm-up of the END-TO-END path, untimed: the first ~1000-1500 steps after a pause without host<->device traffic run 30-50 %
    # slower than the steady state (measured: 294, 275, 260, 251, 243 us/step over five 200-step epochs, then 189.5 for good --
    # the PCIe link and the host cores leave their power-saving states only under sustained traffic; profiles/r02_e2e.md).  The
    # decay is slow (consecutive 50-step blocks differ by < 3 %), so the warm-up is a fixed 2500 steps (~0.5 s), the same on every
    # rank; the first and last 250-step block are reported so that the line shows what the warm-up did.
    e2e_warm, warm_blocks = 0, []
    while e2e_warm < 2500:
        barrier(); tb = time.perf_counter(); epoch(250); barrier()
        warm_blocks.append((time.perf_counter() - tb) / 250 * 1e6); e2e_warm += 250
    barrier()
    t0 = time.perf_counter()
    epoch(args.steps)                                           # every step's H2D copy (the first one too) is inside
    barrier()
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e = {"value": world * B * args.steps / float(t), "unit": "images/s",
           "h2d_bytes_per_step": h_imgs.numel() * 4 + h_labels.numel() * 8, "d2h_bytes_per_step": 8 if getattr(step, "loss_mirror", False) else 4,
           "loss_readback": ("head kernel writes loss + step count into pinned host memory; the host reads every step's loss from there"
                             if getattr(step, "loss_mirror", False) else "D2H copy + event synchronise per step"), "loss_lookahead": 1,
           "ms_per_step": float(t) / args.steps * 1e3, "warmup_steps": e2e_warm,
           "warmup_us_per_step_first_last_block": [round(warm_blocks[0], 1), round(warm_blocks[-1], 1)]}

    # ---- per-kernel-class timing + roofline (rank 0) ------------------------------------------
    hbm, tflops, tflops_burst, peak_kind = peaks()
    roof, breakdown = None, None
    if rank == 0:
        s_act = 2 if args.dtype == "bf16" else 4
        T = B * 256 * C * s_act
        persist = lib.nrd_plan_run_op(step.plan.handle, 8, 0, torch.cuda.current_stream().cuda_stream) == 0
        torch.cuda.synchronize(dev)
        # algorithmic HBM bytes per launch = SURVEY 8(d) per-image figures x the images (and blocks) one launch covers
        grid = min(B, 148)
        head_bytes = T + T // 4 + T // (4 * s_act) + T // 4 + 64 * C * 32 * 4       # x_NB in; pooled, arg-max, dL/d(pooled) out; fc1 weight
        reduce_bytes = grid * (9 * C * C + 28 * C) * 4 + (9 * C * C + 28 * C) * 4   # per-CTA partials in, gradients out
        if persist:
            # every kernel on the step's critical path, in launch order (the head weight-gradient kernels run beside the backward)
            ops = {"stem_fwd": (5, 1, B * 3072 * 4 + T + T // s_act), "fwd_blocks(persistent)": (8, 1, NB * 4 * T), "head": (10, 1, head_bytes),
                   "bwd_blocks(persistent)": (9, 1, NB * 6 * T), "stem_bwd": (6, 1, B * 3072 * 4 + 2 * T + T // s_act),
                   "body_reduce(+sgd)": (11, 1, reduce_bytes)}
            if os.environ.get("NRD_FUSE_STEM_BWD", "0") == "1":
                ops.pop("stem_bwd")                       # conv1's gradient is computed in the persistent backward's tail: no such launch
            if world == 1 and not getattr(step, "fused_update", False):
                ops["sgd_pack"] = (12, 1, 3 * 4 * (64 * C * 32 + 9 * C * C + 27 * C))
        else:
            ops = {"conv_fwd": (0, NB, 2 * T), "bn_apply": (1, NB, 3 * T), "bn_bwd": (2, NB, 5 * T), "wgrad": (3, NB, 2 * T),
                   "dgrad": (4, NB, 3 * T), "stem_fwd": (5, 1, B * 3072 * 4 + T + T // s_act), "stem_bwd": (6, 1, B * 3072 * 4 + 2 * T + T // s_act)}
        stream = torch.cuda.current_stream().cuda_stream
        breakdown = {}
        for name, (op, count, nbytes) in ops.items():
            for _ in range(3):
                check(lib.nrd_plan_run_op(step.plan.handle, op, NB // 2, stream), name)
            tt = []
            for _ in range(10):
                flush.zero_()
                a, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(); check(lib.nrd_plan_run_op(step.plan.handle, op, NB // 2, stream), name); b_.record()
                torch.cuda.synchronize(dev)
                tt.append(a.elapsed_time(b_) * 1e3)
            us = sum(tt) / len(tt)
            breakdown[name] = {"us": round(us, 2), "per_step": count, "algo_bytes": nbytes, "gbs": round(nbytes / us / 1e3, 1)}
        tot = sum(v["us"] * v["per_step"] for v in breakdown.values())
        for v in breakdown.values():
            v["share"] = round(v["us"] * v["per_step"] / tot, 3)
        dom = max(breakdown, key=lambda k: breakdown[k]["us"] * breakdown[k]["per_step"])
        d = breakdown[dom]
        # DRAM bytes per launch: from the tracked ncu --set full capture of THIS source and configuration
        # (profiles/ncu_traffic.json, written by tools/ncu_traffic.py), else null
        traffic = ncu_traffic(dom, B, C, NB, args.dtype)
        # FLOPs of the dominant kernel: forward = one conv per block, backward = dgrad + wgrad per block (2 * 4608 * C^2 each)
        kflops = {"fwd_blocks(persistent)": NB * 4608 * C * C * B, "bwd_blocks(persistent)": 2 * NB * 4608 * C * C * B}.get(dom)
        roof = {"bound": "hbm", "kernel": dom, "achieved": d["gbs"], "peak": hbm, "unit": "GB/s", "frac": round(d["gbs"] / hbm, 4),
                "traffic": traffic, "peak_source": peak_kind, "launch_us": d["us"], "algo_bytes_per_launch": d["algo_bytes"],
                # physical fractions: the fused kernel moves LESS than the algorithmic bytes (activations stay on chip between
                # blocks), so `frac` says how far the kernel is from the layer-by-layer HBM bound, these say what it really uses
                "dram_frac": round(traffic / (d["us"] * 1e-6) / (hbm * 1e9), 4) if traffic else None,
                "tensor_frac": round(kflops / (d["us"] * 1e-6) / (tflops_burst * 1e12), 4) if kflops else None,
                "tensor_peak_tflops": tflops_burst,
                "impl": "tcgen05" if lib.nrd_plan_conv_impl(step.plan.handle) == 1 else "cuda_core"}
        t_hbm_us = B * bytes_per_image(C, NB, s_act) / (hbm * 1e9) * 1e6
        t_tc_us = B * flops_per_image(C, NB) / (tflops * 1e12) * 1e6
        t_roof_us = max(t_hbm_us, t_tc_us)
        t_step_us = ms_total / args.steps * 1e3
        step_roof = {"t_roof_us": round(t_roof_us, 1), "t_step_us": round(t_step_us, 1),
                     "frac": round(t_roof_us / t_step_us, 4), "bytes_per_image": bytes_per_image(C, NB, s_act),
                     "flops_per_image": flops_per_image(C, NB),
                     # a cross-layer-fused step can beat the layer-boundary HBM bound; the bound it cannot beat:
                     "compute_bound_us": round(t_tc_us, 1), "compute_bound_frac": round(t_tc_us / t_step_us, 4),
                     "tensor_peak_tflops_sustained": tflops}
    eager_ref = None
    if not args.no_eager_ref:
        try:
            eager_ref = gpu_eager_reference(B, C, NB, dev, world)
        except Exception as e:                                   # secondary number: never let it take the line down
            eager_ref = {"error": f"{type(e).__name__}: {e}"[:200]}
    if rank == 0:
        line = {"metric": "train_images_per_sec", "value": value, "unit": "images/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": args.dtype, "data": "synthetic",
                "config": {"workload": f"NetResDeep C={C} NB={NB} DDP train step (BASELINE configs[1]), per-rank batch {B}",
                           "global_batch": B * world, "parallelism": f"dp{world}", "l2": "flushed between steps (256 MiB write, untimed)",
                           "graph": not args.no_graph, "lr": 1e-2,
                           "body_exchange": ("none (NRD_DIAG_NO_EXCHANGE: ranks train alone, timing diagnosis only)" if (world > 1 and os.environ.get("NRD_DIAG_NO_EXCHANGE") == "1")
                                             else "p2p" if getattr(step, "p2p", False) else ("nccl" if world > 1 else None)),
                           "value_inputs": "one synthetic batch (nrd_gen_batch), resident in HBM, replayed every step: the generator is outside the timed step",
                           "source_sha16": source_sha16()},
                "e2e": e2e, "gpu_launches": int(launches_per_step * args.steps) if launches_per_step else None,
                "launches_per_step": launches_per_step, "clocks": clocks, "roofline": roof, "step_roofline": step_roof,
                "breakdown": breakdown, "loss_last": loss_last}
        if cpu is not None:
            line["cpu_baseline"] = cpu
        if parity is not None:
            line["parity"] = parity
        if eager_ref is not None:
            line["gpu_eager_reference"] = eager_ref
        print(json.dumps(line), flush=True)
    if world > 1:
        # release the captured graph (it holds NCCL kernels) before the communicator goes away; then leave without the
        # communicator teardown, which can block behind captured collectives
        step.graph = None
        del step
        torch.cuda.synchronize(dev)
        dist.barrier()
        sys.stdout.flush()
        os._exit(0)


if __name__ == "__main__":
    main()
