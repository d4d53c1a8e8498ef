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
  cpu_baseline  the oracle's port of the reference CPU step (main_no_ddp.py:45-50, B=64, fp32) timed
            on this box's host cores (rank 0, N=1 only).
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
        return float(d["hbm_gbs"]), float(d.get("bf16_tflops_sustained", 1402.0)), "measured"
    except Exception:
        return FALLBACK_HBM_GBS, 1400.0, "fallback"


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
    from oracle import netresdeep_oracle as O
    import torch
    cores = len(os.sched_getaffinity(0))
    torch.set_num_threads(cores)
    best, med, thr = O.time_cpu_step(batch=batch, n_chans1=C, n_blocks=NB, steps=steps, warmup=warmup, threads=cores)
    return {"best_s": best, "median_s": med, "cores": thr, "batch": batch}


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
            "cpu_baseline": {"value": val, "unit": "images/s", "cores": r["cores"], "kind": "port", "sample": sample},
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
        cpu = {"value": 64 / r["median_s"], "unit": "images/s", "cores": r["cores"], "kind": "port",
               "sample": f"30 steps of configs[0] (B=64, fp32, main_no_ddp.py step) after 5 warm-up; median {r['median_s'] * 1e3:.1f} ms, best {r['best_s'] * 1e3:.1f} ms/step"}

    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    import ddp_cifar10_b200 as nrd
    from ddp_cifar10_b200._lib import lib, check
    B, C, NB = args.batch, args.chans, args.blocks
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

    epoch(3)
    barrier()
    t0 = time.perf_counter()
    epoch(args.steps)                                           # every step's H2D copy (the first one too) is inside
    barrier()
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e = {"value": world * B * args.steps / float(t), "unit": "images/s",
           "h2d_bytes_per_step": h_imgs.numel() * 4 + h_labels.numel() * 8, "d2h_bytes_per_step": 4, "loss_lookahead": 1,
           "ms_per_step": float(t) / args.steps * 1e3}

    # ---- per-kernel-class timing + roofline (rank 0) ------------------------------------------
    hbm, tflops, peak_kind = peaks()
    roof, breakdown = None, None
    if rank == 0:
        s_act = 2 if args.dtype == "bf16" else 4
        T = B * 256 * C * s_act
        persist = lib.nrd_plan_run_op(step.plan.handle, 8, 0, torch.cuda.current_stream().cuda_stream) == 0
        torch.cuda.synchronize(dev)
        # algorithmic HBM bytes per launch = SURVEY 8(d) per-image figures x the images (and blocks) one launch covers
        if persist:
            ops = {"stem_fwd": (5, 1, B * 3072 * 4 + T + T // s_act), "fwd_blocks(persistent)": (8, 1, NB * 4 * T),
                   "bwd_blocks(persistent)": (9, 1, NB * 6 * T), "stem_bwd": (6, 1, B * 3072 * 4 + 2 * T + T // s_act)}
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
        # DRAM bytes per launch from the committed ncu --set full capture (profiles/r01_step_v10_final.md), same config only
        ncu_traffic = {"bwd_blocks(persistent)": 85645056 + 3553280, "fwd_blocks(persistent)": 4365056 + 32791552}
        traffic = ncu_traffic.get(dom) if (B, C, NB, args.dtype) == (256, 32, 10, "bf16") else None
        roof = {"bound": "hbm", "kernel": dom, "achieved": d["gbs"], "peak": hbm, "unit": "GB/s", "frac": round(d["gbs"] / hbm, 4),
                "traffic": traffic, "peak_source": peak_kind, "launch_us": d["us"], "algo_bytes_per_launch": d["algo_bytes"],
                "impl": "tcgen05" if lib.nrd_plan_conv_impl(step.plan.handle) == 1 else "cuda_core"}
        t_roof_us = B * max(flops_per_image(C, NB) / (tflops * 1e12), bytes_per_image(C, NB, s_act) / (hbm * 1e9)) * 1e6
        step_roof = {"t_roof_us": round(t_roof_us, 1), "t_step_us": round(ms_total / args.steps * 1e3, 1),
                     "frac": round(t_roof_us / (ms_total / args.steps * 1e3), 4), "bytes_per_image": bytes_per_image(C, NB, s_act),
                     "flops_per_image": flops_per_image(C, NB)}
    if rank == 0:
        line = {"metric": "train_images_per_sec", "value": value, "unit": "images/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": args.dtype, "data": "synthetic",
                "config": {"workload": f"NetResDeep C={C} NB={NB} DDP train step (BASELINE configs[1]), per-rank batch {B}",
                           "global_batch": B * world, "parallelism": f"dp{world}", "l2": "flushed between steps (256 MiB write, untimed)",
                           "graph": not args.no_graph, "body_exchange": ("p2p" if getattr(step, "p2p", False) else ("nccl" if world > 1 else None)), "lr": 1e-2},
                "e2e": e2e, "gpu_launches": int(launches_per_step * args.steps) if launches_per_step else None,
                "launches_per_step": launches_per_step, "clocks": clocks, "roofline": roof, "step_roofline": step_roof,
                "breakdown": breakdown, "loss_last": loss_last}
        if cpu is not None:
            line["cpu_baseline"] = cpu
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
