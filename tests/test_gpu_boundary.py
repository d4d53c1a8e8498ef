"""The drop-in boundary exercised the way the reference itself uses it:

  * the literal loop of /root/reference/main.py:27-41 -- STOCK `torch.optim.SGD(model.parameters(), lr=1e-2)` and STOCK
    `torch.nn.CrossEntropyLoss()` around our NetResDeep -- against the fp32 oracle over 3 steps;
  * the raw C ABI as INTEGRATION.md section 3 shows it: `nrd_train_step` and `nrd_forward` + `nrd_cross_entropy` +
    `nrd_backward` through ctypes on caller-owned buffers, against the oracle;
  * `ResBlock.forward` on its own (model/resnet.py:33-37) against the oracle's block, forward and backward;
  * `mp.spawn(main, args=(world_size,), nprocs=world_size)` (main.py:80-84) for one short epoch with the checkpoint
    branch of main.py:43-45 taken, and the written `birds_vs_airplanes.pt` loaded into the reference's own NetResDeep
    (the unmodified file under baseline/_ref when it travelled with the snapshot; keys and shapes from the golden
    fixture otherwise) and back into ours;
  * eval-mode forward in bf16 and an accuracy loop over the BN buffers the training forward maintains (8(f3));
  * the sticky saturation status (clamped BatchNorm sums must be loud).
"""
import ctypes as C
import os
import sys

import numpy as np
import pytest
import torch

from conftest import GOLDEN, ROOT
from oracle import netresdeep_oracle as O
from util import *  # noqa: F401,F403
from ddp_cifar10_b200 import _lib
from ddp_cifar10_b200._lib import lib, check

pytestmark = pytest.mark.gpu


def _data(B, seed):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(B, 3, 32, 32, generator=g), torch.randint(0, 10, (B,), generator=g)


def test_literal_reference_loop_with_stock_torch_sgd_and_cross_entropy():
    """main.py:27-41 verbatim around the drop-in model: torch's optimizer updates the parameters in place (they are views
    of the engine's flat vector; the engine re-derives its packed copies when it sees their version change)."""
    import torch.nn as nn
    import torch.optim as optim
    B, NB = 32, 10
    x, y = _data(B, 21)
    p, bufs = O.init_params(32, NB, seed=7, dtype=torch.float64)
    pm, bm = O.init_params(32, NB, seed=7)
    model = model_from_oracle(pm, bm, NB, "fp32")
    optimizer = optim.SGD(model.parameters(), lr=1e-2)        # main.py:27
    loss_fn = nn.CrossEntropyLoss()                           # main.py:28
    for s in range(3):
        imgs, labels = x.to(0), y.to(0)                       # main.py:33
        outputs = model(imgs)                                 # main.py:34
        loss = loss_fn(outputs, labels)                       # main.py:35
        optimizer.zero_grad()                                 # main.py:37
        loss.backward()                                       # main.py:38
        optimizer.step()                                      # main.py:39
        p, lref, lossref, gref = O.train_step(p, bufs, x.double(), y, NB, 1e-2)
        assert rell2(outputs.detach(), lref) < 1e-5, (s, rell2(outputs.detach(), lref))
        assert abs(loss.item() - float(lossref)) < 1e-5 * float(lossref), s          # main.py:41
        named = dict(model.named_parameters())
        for k, mk in ORACLE_TO_MODULE.items():
            if k.startswith("fc"):
                assert rell2(named[mk].grad, gref[k]) < 2e-5, (s, k)
            # post-step weights; body tensors carry ReLU-flip noise of the gradient times lr (the fp32 oracle itself is
            # 5e-5 from the fp64 one on conv1.weight after the third, spiking step of this seed)
            assert rell2(named[mk].detach(), p[k]) < (1e-5 if k.startswith("fc") else 3e-4), (s, k)
    assert int(model.resblocks[0].batch_norm.num_batches_tracked) == 3 * NB


def _raw_plan(B, Cc, NB, dt, p, bufs):
    """caller-owned flat vectors + workspace, plan handle."""
    off = (C.c_int64 * 10)(); check(lib.nrd_param_offsets(Cc, off))
    params = torch.empty(off[9], device=DEV); grads = torch.zeros(off[9], device=DEV)
    for i, k in enumerate(O.PARAM_ORDER):
        params[off[i]:off[i + 1]] = p[k].float().flatten().to(DEV)
    bn = torch.cat([bufs["running_mean"].float(), bufs["running_var"].float()]).to(DEV)
    nbt = torch.zeros((), dtype=torch.int64, device=DEV)
    nbytes = lib.nrd_workspace_bytes(B, Cc, NB, dt)
    ws = torch.empty(nbytes + 256, dtype=torch.uint8, device=DEV); base = (ws.data_ptr() + 255) & ~255
    plan = C.c_void_p()
    check(lib.nrd_plan_create(C.byref(plan), 0, B, Cc, NB, dt, base, nbytes, params.data_ptr(), grads.data_ptr(), bn.data_ptr(), nbt.data_ptr()))
    check(lib.nrd_pack_weights(plan, st()))
    return plan, off, params, grads, bn, nbt, ws


@pytest.mark.parametrize("dtype", ["fp32", "bf16"])
def test_raw_c_abi_train_step_and_backward(dtype):
    B, Cc, NB, lr = 32, 32, 10, 1e-2
    x, y = _data(B, 33)
    p, bufs = O.init_params(Cc, NB, seed=8)
    ref = O.train_step if dtype == "fp32" else O.train_step_rm
    pr, br = O.init_params(Cc, NB, seed=8)
    p_new, lref, lossref, gref = ref(pr, br, x, y, NB, lr)
    tol_l, tol_h, tol_b = (1e-5, 2e-5, 5e-3) if dtype == "fp32" else (2e-3, 1e-2, 4e-2)     # bf16 at B=32: 8x fewer samples per statistic than the parity test
    imgs, labels = x.to(DEV), y.to(DEV)
    # (1) nrd_train_step: main.py:34-39 in one call
    plan, off, params, grads, bn, nbt, ws = _raw_plan(B, Cc, NB, CODE[dtype], p, bufs)
    logits = torch.empty(B, 10, device=DEV); loss = torch.zeros((), device=DEV)
    check(lib.nrd_train_step(plan, imgs.data_ptr(), labels.data_ptr(), C.c_float(lr), logits.data_ptr(), loss.data_ptr(), st()), "nrd_train_step")
    torch.cuda.synchronize()
    assert rell2(logits, lref) < tol_l and abs(float(loss) - float(lossref)) < 10 * tol_l * float(lossref)
    for i, k in enumerate(O.PARAM_ORDER):
        tol = tol_h if k.startswith("fc") else tol_b
        assert rell2(grads[off[i]:off[i + 1]], gref[k].flatten()) < tol, (k, rell2(grads[off[i]:off[i + 1]], gref[k].flatten()))
        # post-step weights: the update each tensor took (w_new - w_old = -lr * g) at its gradient's tolerance
        w_old = p[k].float().flatten()
        assert rell2(params[off[i]:off[i + 1]].cpu() - w_old, p_new[k].flatten() - w_old) < tol + 1e-4, k
    assert int(nbt) == NB and rell2(bn[:Cc], br["running_mean"]) < 1e-4
    flags = C.c_uint32(7)
    check(lib.nrd_plan_status(plan, C.byref(flags), 1, st()), "nrd_plan_status")
    assert flags.value == 0
    # (2) the same step from its pieces: nrd_forward, nrd_cross_entropy, nrd_backward, nrd_sgd_update_pack
    plan2, off, params2, grads2, bn2, nbt2, ws2 = _raw_plan(B, Cc, NB, CODE[dtype], p, bufs)
    logits2 = torch.empty(B, 10, device=DEV); loss2 = torch.zeros((), device=DEV)
    dlogits = torch.empty(B, 10, device=DEV); rows = torch.empty(B, device=DEV)
    check(lib.nrd_forward(plan2, imgs.data_ptr(), logits2.data_ptr(), 1, st()), "nrd_forward")
    check(lib.nrd_cross_entropy(logits2.data_ptr(), labels.data_ptr(), B, C.c_float(1.0), loss2.data_ptr(), dlogits.data_ptr(), rows.data_ptr(), st()), "nrd_cross_entropy")
    check(lib.nrd_backward(plan2, dlogits.data_ptr(), st()), "nrd_backward")
    check(lib.nrd_sgd_update_pack(plan2, C.c_float(lr), C.c_float(1.0), st()), "nrd_sgd_update_pack")
    torch.cuda.synchronize()
    # same kernels in the same order on the same inputs: bit-identical to the one-call entry
    assert torch.equal(logits2, logits) and torch.equal(loss2, loss) and torch.equal(grads2, grads) and torch.equal(params2, params)
    assert torch.equal(bn2, bn) and int(nbt2) == NB
    check(lib.nrd_plan_destroy(plan)); check(lib.nrd_plan_destroy(plan2))


@pytest.mark.parametrize("dtype,Cc", [("fp32", 32), ("bf16", 32), ("fp32", 64)])
def test_standalone_resblock_forward_backward(dtype, Cc):
    """`ResBlock.forward` alone (model/resnet.py:33-37): relu(bn(conv(x))) + x, training-mode BN with its buffer updates."""
    B = 6
    g = torch.Generator().manual_seed(3)
    x = torch.randn(B, Cc, 16, 16, generator=g); gout = torch.randn(B, Cc, 16, 16, generator=g)
    torch.manual_seed(11)
    blk = nrd.ResBlock(Cc).to(DEV)
    w, gamma, beta = (t.detach().cpu().double() for t in (blk.conv.weight, blk.batch_norm.weight, blk.batch_norm.bias))
    # reference: the oracle's block arithmetic with the quantiser at the kernels' rounding points (identity in fp32): the
    # per-op path rounds the block input and the stored y / dy / output to bf16 and keeps the conv weight in fp32
    quant = (lambda t: t) if dtype == "fp32" else O.q_bf16
    xd = quant(x.double())
    y = torch.nn.functional.conv2d(xd, w, None, padding=1)
    n = B * 256
    mean = y.mean(dim=(0, 2, 3)); var = (y * y).mean(dim=(0, 2, 3)) - mean * mean
    invstd = 1.0 / torch.sqrt(var + O.BN_EPS)
    sc = gamma * invstd; sh = beta - mean * sc
    yb = quant(y)
    z = yb * sc[None, :, None, None] + sh[None, :, None, None]
    ref = quant(torch.relu(z) + xd)
    rm_ref = 0.1 * mean; rv_ref = 0.9 + 0.1 * var * (n / (n - 1))
    from ddp_cifar10_b200.functional import resblock_forward
    xin = x.to(DEV).requires_grad_(True)
    out = resblock_forward(blk, xin, TDT[dtype])
    tol = 1e-5 if dtype == "fp32" else 4e-3                  # bf16: one rounding step of the output is 2^-9
    assert rell2(out.detach(), ref) < tol, rell2(out.detach(), ref)
    assert rell2(blk.batch_norm.running_mean, rm_ref) < 1e-5 and rell2(blk.batch_norm.running_var, rv_ref) < 1e-5
    assert int(blk.batch_norm.num_batches_tracked) == 1
    out.backward(gout.to(DEV))
    go = quant(gout.double())
    dz = go * (z > 0).double()
    db = dz.sum(dim=(0, 2, 3)); dgy = (dz * yb).sum(dim=(0, 2, 3))
    dg = invstd * (dgy - mean * db)
    a_ = gamma * invstd
    dy = quant(dz * a_[None, :, None, None] + yb * (-a_ * (dg / n) * invstd)[None, :, None, None]
               + (a_ * (dg / n) * invstd * mean - a_ * (db / n))[None, :, None, None])
    dW = torch.nn.grad.conv2d_weight(xd, w.shape, dy, padding=1)
    dx = torch.nn.grad.conv2d_input(xd.shape, w, dy, padding=1) + go
    gt = 1e-4 if dtype == "fp32" else 1e-2                   # bf16: a handful of masks / dy roundings at the edge of a rounding step
    assert rell2(blk.conv.weight.grad, dW) < gt, rell2(blk.conv.weight.grad, dW)
    assert rell2(blk.batch_norm.weight.grad, dg) < gt, rell2(blk.batch_norm.weight.grad, dg)
    assert rell2(blk.batch_norm.bias.grad, db) < gt, rell2(blk.batch_norm.bias.grad, db)
    assert rell2(xin.grad, dx) < gt, rell2(xin.grad, dx)


def _ref_netresdeep():
    """The UNMODIFIED reference NetResDeep when its file travelled with the snapshot (baseline/_ref, git-ignored), else None."""
    path = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.exists(os.path.join(path, "model", "resnet.py")):
        return None
    sys.path.insert(0, path)
    try:
        for k in [k for k in sys.modules if k == "model" or k.startswith("model.")]:
            del sys.modules[k]
        from model.resnet import NetResDeep
        return NetResDeep
    finally:
        sys.path.remove(path)


def test_mp_spawn_main_writes_checkpoint_the_reference_model_loads(tmp_path):
    """main.py:80-84 + the save branch main.py:43-45, one epoch of 3 steps; world_size = the GPUs present (at most 2)."""
    import torch.multiprocessing as mp
    world_size = min(torch.cuda.device_count(), 2)
    cwd = os.getcwd()
    os.chdir(tmp_path)
    os.environ["MASTER_PORT"] = "29583"
    try:
        mp.spawn(nrd.main.main, args=(world_size, 1, 32, 3, "fp32", True), nprocs=world_size)       # main.py:84
    finally:
        os.chdir(cwd); os.environ.pop("MASTER_PORT", None)
    ckpt = tmp_path / "data" / "CIFAR-10" / "birds_vs_airplanes.pt"                                  # main.py:45
    assert ckpt.exists()
    sd = torch.load(ckpt, map_location="cpu")
    g = np.load(os.path.join(GOLDEN, "step_c32_nb10_b8.npz"))
    assert list(sd.keys()) == list(g["state_dict_keys"]) and len(sd) == 66
    assert sd["resblocks.9.conv.weight"].shape == (32, 32, 3, 3) and sd["fc1.weight"].shape == (32, 2048)
    assert int(sd["resblocks.0.batch_norm.num_batches_tracked"]) == 30                             # 3 steps x 10 applications
    assert torch.equal(sd["resblocks.0.conv.weight"], sd["resblocks.9.conv.weight"])
    x, _ = _data(8, 2)
    ours = nrd.NetResDeep(dtype="fp32"); ours.load_state_dict(sd); ours = ours.to(DEV).eval()
    with torch.no_grad():
        out = ours(x.to(DEV)).cpu()
    Ref = _ref_netresdeep()
    if Ref is not None:                       # the reference's own class accepts the file, strictly, and computes the same function
        ref = Ref(); ref.load_state_dict(sd, strict=True); ref.eval()
        with torch.no_grad():
            assert rell2(out, ref(x)) < 1e-5
        # and the reverse: a checkpoint written by the reference model loads into ours
        torch.manual_seed(5); r2 = Ref()
        torch.save(r2.state_dict(), tmp_path / "ref.pt")
        o2 = nrd.NetResDeep(dtype="fp32"); o2.load_state_dict(torch.load(tmp_path / "ref.pt")); o2 = o2.to(DEV).eval(); r2.eval()
        with torch.no_grad():
            assert rell2(o2(x.to(DEV)).cpu(), r2(x)) < 1e-5
    else:                                     # no reference file on this box: the oracle's forward on the loaded tensors
        p = {k: sd[mk].double() for k, mk in ORACLE_TO_MODULE.items()}
        b = {"running_mean": sd["resblocks.0.batch_norm.running_mean"].double(), "running_var": sd["resblocks.0.batch_norm.running_var"].double()}
        lref, _ = O.forward(p, b, x.double(), 10, train=False)
        assert rell2(out, lref) < 1e-5


def test_eval_mode_bf16_and_accuracy_loop():
    """Eval-mode forward (BN with the running statistics the training forward maintained 10x per step) in bf16 against the
    rounding-matched oracle, and an accuracy loop: class-template data, train at lr=1e-3, then accuracy in eval mode."""
    gen = torch.Generator().manual_seed(1234)
    T = torch.randn(10, 3, 32, 32, generator=gen); yy = torch.randint(0, 10, (64,), generator=gen)
    xx = T[yy] + torch.randn(64, 3, 32, 32, generator=gen)
    p, bufs = O.init_params(32, 10, seed=0)
    m = model_from_oracle(p, bufs, 10, "bf16")
    step = nrd.TrainStep(m, 64, lr=1e-3)
    for _ in range(120):
        step(xx.to(DEV), yy.to(DEV))
    step.check_status()
    m.eval()
    sd = {k: v.detach().cpu() for k, v in m.state_dict().items()}
    po = {k: sd[mk].float() for k, mk in ORACLE_TO_MODULE.items()}
    bo = {"running_mean": sd["resblocks.0.batch_norm.running_mean"].float(), "running_var": sd["resblocks.0.batch_norm.running_var"].float()}
    with torch.no_grad():
        train_acc = float((m(xx.to(DEV)).argmax(1).cpu() == yy).float().mean())
    assert train_acc >= 0.95, train_acc                      # eval-mode BN (running statistics) reproduces what training fitted
    correct = total = 0
    for s in range(4):                                       # fresh noise around the same templates
        g2 = torch.Generator().manual_seed(99 + s)
        yb = torch.randint(0, 10, (64,), generator=g2); xb = T[yb] + torch.randn(64, 3, 32, 32, generator=g2)
        with torch.no_grad():
            out = m(xb.to(DEV)).cpu()
        lref, _ = O.forward_rm(po, dict(bo), xb, 10, train=False)
        assert rell2(out, lref) < 2e-2, rell2(out, lref)
        agree = (out.argmax(1) == lref.argmax(1)).float().mean()
        assert agree >= 0.95
        correct += int((out.argmax(1) == yb).sum()); total += 64
    assert int(m.resblocks[0].batch_norm.num_batches_tracked) == 1200      # eval forwards leave the buffers alone
    assert correct / total > 0.4, correct / total            # 64 training samples: far above the 0.1 of chance, not a benchmark


def test_saturated_bn_sums_are_loud():
    """Activations x 1e4: the per-CTA sums of y^2 leave the fixed-point range of the persistent forward; the step still
    runs (clamped) but the status check raises -- and a normal step afterwards is clean again."""
    torch.manual_seed(0)
    m = nrd.NetResDeep(32, 10, dtype="bf16").to(DEV)
    x, y = _data(64, 5)
    step = nrd.TrainStep(m, 64, lr=1e-3)
    step(x.to(DEV), y.to(DEV)); step.check_status()
    step((x * 1e4).to(DEV), y.to(DEV))
    with pytest.raises(_lib.NrdError, match="fixed-point range|NaN"):
        step.check_status()
    step.check_status()                                      # the failed check cleared the word
    # eager path: the engine's cached plans carry the same word (fresh model: the step above poisoned that one's weights,
    # exactly as Inf gradients would in the reference)
    torch.manual_seed(0)
    m2 = nrd.NetResDeep(32, 10, dtype="bf16").to(DEV)
    out = m2(x.to(DEV)); nrd.CrossEntropyLoss()(out, y.to(DEV)).backward(); m2.engine.check_status()
    out = m2((x * 1e4).to(DEV)); nrd.CrossEntropyLoss()(out, y.to(DEV)).backward()
    with pytest.raises(_lib.NrdError):
        m2.engine.check_status()


def test_eager_step_after_graph_steps_on_the_same_model():
    """TrainStep owns a private plan: eager steps on the same model and batch size afterwards behave like eager steps."""
    B = 16
    x, y = _data(B, 8)
    finals = []
    for mode in ("eager", "mixed"):
        p, bufs = O.init_params(32, 10, seed=5)
        m = model_from_oracle(p, bufs, 10, "bf16")
        opt = nrd.SGD(m.parameters(), lr=1e-3); loss_fn = nrd.CrossEntropyLoss()
        losses = []
        def eager():
            loss = loss_fn(m(x.to(DEV)), y.to(DEV)); opt.zero_grad(); loss.backward(); opt.step(); losses.append(float(loss))
        if mode == "eager":
            for _ in range(4):
                eager()
        else:
            step = nrd.TrainStep(m, B, lr=1e-3)
            losses.append(float(step(x.to(DEV), y.to(DEV))))
            eager()
            losses.append(float(step(x.to(DEV), y.to(DEV))))
            eager()
            assert step.plan is not m.engine.plans[B]
        finals.append((losses, m.engine.flat_params.clone(), m.engine.bn_running.clone(), int(m.engine.nbt)))
    assert finals[0][0] == finals[1][0]
    assert torch.equal(finals[0][1], finals[1][1]) and torch.equal(finals[0][2], finals[1][2]) and finals[0][3] == finals[1][3] == 40


def test_sgd_momentum_and_bce_with_logits_variants():
    """the optimizer / loss of the reference's second script (ppe_main_ddp.py:133,147) around the same model:
    SGD(lr=0.001, momentum=0.9) against torch.optim.SGD on the same gradients over 3 steps, binary_cross_entropy_with_logits
    (value and gradient) against torch's."""
    import torch.nn.functional as F
    x, y = _data(16, 41)
    pm, bm = O.init_params(32, 10, seed=9)
    m = model_from_oracle(pm, bm, 10, "fp32")
    opt = nrd.SGD(m.parameters(), lr=1e-3, momentum=0.9)
    m.engine.adopt(next(m.parameters()).device)             # (normally done by the first forward)
    ref_p = m.engine.flat_params.detach().clone().requires_grad_(True)
    ref_opt = torch.optim.SGD([ref_p], lr=1e-3, momentum=0.9)
    for s in range(3):
        out = m(x.to(DEV))
        onehot = F.one_hot(y, 10).float().to(DEV)
        loss = nrd.binary_cross_entropy_with_logits(out, onehot)
        want = F.binary_cross_entropy_with_logits(out.detach(), onehot)
        assert abs(float(loss) - float(want)) < 1e-6 * abs(float(want))
        o2 = out.detach().clone().requires_grad_(True)
        F.binary_cross_entropy_with_logits(o2, onehot).backward()
        opt.zero_grad(); loss.backward()
        ref_p.grad = m.engine.flat_grads.detach().clone()
        opt.step(); ref_opt.step()
        assert rell2(m.engine.flat_params, ref_p.detach()) < 1e-7, s
    # dlogits of the fused kernel vs autograd through torch's own loss
    lg = torch.randn(64, 10, device=DEV, requires_grad=True); tg = torch.rand(64, 10, device=DEV)
    nrd.binary_cross_entropy_with_logits(lg, tg).backward()
    l2 = lg.detach().clone().requires_grad_(True); F.binary_cross_entropy_with_logits(l2, tg).backward()
    assert rell2(lg.grad, l2.grad) < 1e-6
