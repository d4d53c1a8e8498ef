"""Where the end-to-end step loses time against the device-timed step: variants of the host loop, wall-clock per step."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ddp_cifar10_b200 as nrd
B, N = 256, 300
dev = torch.device("cuda:0")
torch.manual_seed(0)
m = nrd.NetResDeep(32, 10, dtype="bf16").to(dev)
step = nrd.TrainStep(m, B, lr=1e-2)
imgs, labels = nrd.SyntheticCifar10Loader(batch_size=B, device=dev).batch(0)
h_imgs = imgs.cpu().pin_memory(); h_labels = labels.cpu().pin_memory()
step.imgs.copy_(imgs); step.labels.copy_(labels)
def wall(fn, n=N):
    fn(20); torch.cuda.synchronize(); t0 = time.perf_counter(); fn(n); torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e6
def replay_only(n):
    for _ in range(n): step.run()
def replay_graph_only(n):
    g = step.graphs[0]
    for _ in range(n): g.replay()
class HB:
    def __init__(s, n): s.n = n
    def __len__(s): return s.n
    def __iter__(s): return ((h_imgs, h_labels) for _ in range(s.n))
def it0(n):
    for l in step.iterate(HB(n), lookahead=0): float(l)
def it1(n):
    for l in step.iterate(HB(n), lookahead=1): float(l)
print("graph.replay() only          %.1f us/step" % wall(replay_graph_only))
print("step.run() only              %.1f us/step" % wall(replay_only))
print("iterate lookahead=0          %.1f us/step" % wall(it0))
print("iterate lookahead=1          %.1f us/step" % wall(it1))
# hand-rolled: copies on copy stream + replay, loss read every step with lookahead 1, minimal python
copy = torch.cuda.Stream(); main = torch.cuda.current_stream()
step._capture(1) if step.graphs[1] is None else None
ready = [torch.cuda.Event(), torch.cuda.Event()]; cons = [torch.cuda.Event(), torch.cuda.Event()]
hl = [torch.zeros((), dtype=torch.float32).pin_memory() for _ in range(2)]; ld = [torch.cuda.Event(), torch.cuda.Event()]
def hand(n, h2d=True, d2h=True):
    for e in cons: e.record(main)
    s = 0
    if h2d:
        with torch.cuda.stream(copy):
            step.inputs[0][0].copy_(h_imgs, non_blocking=True); step.inputs[0][1].copy_(h_labels, non_blocking=True); ready[0].record(copy)
    tot = 0.0
    for k in range(n):
        if h2d: main.wait_event(ready[s])
        step.graphs[s].replay()
        cons[s].record(main)
        if d2h:
            hl[k & 1].copy_(step.loss, non_blocking=True); ld[k & 1].record(main)
        if h2d and k + 1 < n:
            copy.wait_event(cons[s ^ 1])
            with torch.cuda.stream(copy):
                step.inputs[s ^ 1][0].copy_(h_imgs, non_blocking=True); step.inputs[s ^ 1][1].copy_(h_labels, non_blocking=True); ready[s ^ 1].record(copy)
        if d2h and k > 0:
            ld[(k - 1) & 1].synchronize(); tot += float(hl[(k - 1) & 1])
        s ^= 1
    return tot
print("hand loop h2d+d2h            %.1f us/step" % wall(lambda n: hand(n, True, True)))
print("hand loop h2d only           %.1f us/step" % wall(lambda n: hand(n, True, False)))
print("hand loop d2h only           %.1f us/step" % wall(lambda n: hand(n, False, True)))
print("hand loop neither            %.1f us/step" % wall(lambda n: hand(n, False, False)))
def it1m(n):
    tot = 0.0
    for l in step.iterate(HB(n), lookahead=1): tot += float(l)
    return tot
print("iterate lookahead=1 (mirror=%s) %.1f us/step" % (step.loss_mirror, wall(it1m)))
# values: mirror == device loss every step
ls = [float(l) for l in step.iterate(HB(5), lookahead=1)]
print("losses via iterate:", ls, " device loss now:", float(step.loss))
