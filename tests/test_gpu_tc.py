"""tcgen05 tensor-core convolutions (fwd / dgrad / wgrad) against (a) our CUDA-core kernels on the
same bf16 operands and (b) torch fp32 conv on the same bf16-rounded operands.  Tolerance: fp32 side
outputs (BN partial sums, weight-grad partials) <= 1e-3 relative (north_star; measured ~1e-6: only the
accumulation order differs); bf16 outputs within one bf16 rounding step (2^-8) of the fp32 result."""
import pytest
import torch
import torch.nn.functional as F

from util import *  # noqa: F401,F403
from ddp_cifar10_b200 import _lib

pytestmark = pytest.mark.gpu
OP_CONV_FWD, OP_WGRAD, OP_DGRAD = 0, 3, 4


def _put(pl, what, k, t):
    pl.tensor_view(what, k).view(torch.bfloat16).copy_(t.to(DEV, torch.bfloat16).flatten())


@pytest.mark.parametrize("B,Cc", [(3, 32), (80, 32), (256, 32), (2, 64), (75, 64)])
def test_tc_conv_fwd_dgrad_wgrad(B, Cc):
    g = torch.Generator().manual_seed(B + Cc)
    m = nrd.NetResDeep(Cc, 2, dtype="bf16").to(DEV)
    eng = m.engine
    pl = eng.plan(B, torch.device(DEV))
    eng.ensure_packed(pl)
    x = torch.randn(B, 16, 16, Cc, generator=g).bfloat16()
    dy = torch.randn(B, 16, 16, Cc, generator=g).bfloat16()
    res = torch.randn(B, 16, 16, Cc, generator=g).bfloat16()
    _put(pl, 0, 0, x); _put(pl, 2, 0, dy); _put(pl, 5, 0, res)
    w = m.resblocks[0].conv.weight.detach().bfloat16().float().cpu()      # operands as the tensor cores see them
    got = {}
    for name, impl in (("cc", _lib.CONV_CUDA_CORE), ("tc", _lib.CONV_TCGEN05)):
        eng.set_conv_impl(impl)
        assert lib.nrd_plan_conv_impl(pl.handle) == impl
        for what in (1, 6, 7, 8):
            pl.tensor_view(what, 0).zero_()
        for op in (OP_CONV_FWD, OP_DGRAD, OP_WGRAD):
            check(lib.nrd_plan_run_op(pl.handle, op, 0, st()), f"run_op {op}")
        torch.cuda.synchronize()
        got[name] = dict(
            y=pl.tensor_view(1, 0).view(torch.bfloat16).float().view(B, 16, 16, Cc).cpu(),
            dx=pl.tensor_view(6, 0).view(torch.bfloat16).float().view(B, 16, 16, Cc).cpu(),
            dw=pl.tensor_view(7, 0).view(torch.float32).view(-1, 9, Cc, Cc).double().sum(0).cpu(),        # unused CTA slots were zeroed
            stats=pl.tensor_view(8, 0).view(torch.float32).view(-1, 2, Cc).double().sum(0).cpu())
    xn, dyn, rn = (t.float().permute(0, 3, 1, 2) for t in (x, dy, res))
    if name:   # the CUDA-core path uses fp32 weights, the tensor cores bf16-rounded ones: compare each with its own oracle
        ref_y = F.conv2d(xn.double(), w.double(), padding=1)
        ref_dx = torch.nn.grad.conv2d_input(xn.shape, w.double(), dyn.double(), padding=1) + rn.double()
        ref_dw = torch.nn.grad.conv2d_weight(xn.double(), w.shape, dyn.double(), padding=1)     # [co][ci][ky][kx]
    tc = got["tc"]
    assert relmax(tc["y"].permute(0, 3, 1, 2), ref_y) < 2 ** -8 * 1.01
    assert relmax(tc["dx"].permute(0, 3, 1, 2), ref_dx) < 2 ** -8 * 1.01
    assert relmax(tc["stats"][0], ref_y.sum(dim=(0, 2, 3))) < 1e-3 and relmax(tc["stats"][1], (ref_y ** 2).sum(dim=(0, 2, 3))) < 1e-4
    # partial layout [tap][co][ci] -> OIHW
    dw = tc["dw"].view(3, 3, Cc, Cc).permute(2, 3, 0, 1)
    assert relmax(dw, ref_dw) < 1e-4, relmax(dw, ref_dw)
    # and against our own CUDA-core kernels (fp32 weights there: bf16 weight rounding is the only extra difference)
    cc = got["cc"]
    assert rell2(tc["y"], cc["y"]) < 5e-3 and rell2(tc["dx"], cc["dx"]) < 5e-3
    assert relmax(tc["dw"], cc["dw"]) < 1e-4


@pytest.mark.parametrize("B,Cc,NB", [(5, 32, 3), (256, 32, 2), (300, 32, 2), (3, 64, 2), (150, 64, 2)])
def test_fused_block_forward_bitwise_equals_unfused(B, Cc, NB):
    """conv + BN statistics + grid barrier + BN/ReLU/residual in ONE cooperative kernel (accumulators parked in TMEM)
    must give bit-identical activations, saved statistics, running statistics and gradients to the same tensor-core
    convolution followed by the separate finalize / apply kernels."""
    g = torch.Generator().manual_seed(B)
    x = torch.randn(B, 3, 32, 32, generator=g).to(DEV); y = torch.randint(0, 10, (B,), generator=g).to(DEV)
    res = []
    for impl in (_lib.CONV_TCGEN05_UNFUSED, _lib.CONV_TCGEN05):
        torch.manual_seed(0)
        m = nrd.NetResDeep(Cc, NB, dtype="bf16").to(DEV)
        m.engine.set_conv_impl(impl)
        pl = m.engine.plan(B, torch.device(DEV))
        check(lib.nrd_plan_set_option(pl.handle, 1, 0))      # per-block kernels (the persistent stacks are tested below)
        check(lib.nrd_plan_set_option(pl.handle, 3, 0))
        out = m(x); loss = nrd.CrossEntropyLoss()(out, y); loss.backward()
        bn = m.resblocks[0].batch_norm
        res.append((out.detach().clone(), m.engine.flat_grads.clone(), bn.running_mean.clone(), bn.running_var.clone(),
                    int(bn.num_batches_tracked), pl.tensor_view(0, NB).clone(), pl.tensor_view(1, NB - 1).clone(),
                    pl.tensor_view(4, NB - 1).clone()))
    a, b = res
    assert a[4] == b[4] == NB
    for i in (0, 1, 2, 3, 5, 6, 7):
        assert torch.equal(a[i], b[i]), i


@pytest.mark.parametrize("dtype,B,Cc", [("bf16", 7, 32), ("bf16", 256, 32), ("fp32", 40, 32), ("bf16", 20, 64), ("bf16", 600, 64), ("fp32", 500, 32)])
def test_fused_bn_backward_matches_three_kernel_path(dtype, B, Cc):
    """One cooperative BN+ReLU-backward kernel (sums -> grid barrier -> dy from shared memory) against the reduce /
    finalize / apply kernels: same math, different partial-sum grouping => fp32 round-off only.  B = 600 at C = 64 and
    B = 500 in fp32 are 5 / 4 images per CTA of which 3 fit in shared memory: the rest is read again (L2) in phase 2."""
    g = torch.Generator().manual_seed(B)
    x = torch.randn(B, 3, 32, 32, generator=g).to(DEV); y = torch.randint(0, 10, (B,), generator=g).to(DEV)
    res = []
    for fuse in (0, 1):
        torch.manual_seed(0)
        m = nrd.NetResDeep(Cc, 2, dtype=dtype).to(DEV)
        pl = m.engine.plan(B, torch.device(DEV))
        check(lib.nrd_plan_set_option(pl.handle, 0, fuse))
        out = m(x); loss = nrd.CrossEntropyLoss()(out, y); loss.backward()
        res.append(m.engine.flat_grads.clone())
    assert rell2(res[1], res[0]) < (1e-5 if dtype == "fp32" else 2e-3), rell2(res[1], res[0])


@pytest.mark.parametrize("B", [1, 100, 148, 149, 256, 296])
def test_persistent_forward_matches_per_block_kernels(B):
    """All residual blocks in ONE persistent cooperative kernel (activations stay in shared memory / TMEM between blocks)
    against one fused kernel per block: identical MMAs; only the grouping of the fp32 BN partial sums differs, so the
    saved statistics agree to fp32 round-off and activations to a bf16 rounding step here and there."""
    g = torch.Generator().manual_seed(B)
    x = torch.randn(B, 3, 32, 32, generator=g).to(DEV); y = torch.randint(0, 10, (B,), generator=g).to(DEV)
    res = []
    for persist in (0, 1):
        torch.manual_seed(0)
        m = nrd.NetResDeep(32, 10, dtype="bf16").to(DEV)
        pl = m.engine.plan(B, torch.device(DEV))
        check(lib.nrd_plan_set_option(pl.handle, 1, persist))
        out = m(x); loss = nrd.CrossEntropyLoss()(out, y); loss.backward()
        bn = m.resblocks[0].batch_norm
        res.append(dict(out=out.detach().clone(), grads=m.engine.flat_grads.clone(), rm=bn.running_mean.clone(), rv=bn.running_var.clone(),
                        nbt=int(bn.num_batches_tracked), x10=pl.tensor_view(0, 10).view(torch.bfloat16).float().clone(),
                        y9=pl.tensor_view(1, 9).view(torch.bfloat16).float().clone(), st9=pl.tensor_view(4, 9).view(torch.float32).clone()))
    a, b = res
    assert a["nbt"] == b["nbt"] == 10
    assert relmax(b["rm"], a["rm"]) < 1e-5 and relmax(b["rv"], a["rv"]) < 1e-5 and relmax(b["st9"], a["st9"]) < 1e-4
    assert rell2(b["y9"], a["y9"]) < 2e-3 and rell2(b["x10"], a["x10"]) < 2e-3 and rell2(b["out"], a["out"]) < 2e-3
    assert rell2(b["grads"], a["grads"]) < 4e-2      # chaotic amplification of bf16-step differences through 10 blocks


@pytest.mark.parametrize("B", [1, 100, 149, 256, 296])
def test_persistent_backward_matches_per_block_kernels(B):
    """All residual blocks' backward in ONE persistent cooperative kernel (dL/dx in registers, dy in shared memory, the
    weight-gradient accumulators in TMEM across blocks) against the per-block kernels (fused BN backward + wgrad + dgrad):
    same arithmetic per element; only the grouping of fp32 partial sums differs."""
    g = torch.Generator().manual_seed(B)
    x = torch.randn(B, 3, 32, 32, generator=g).to(DEV); y = torch.randint(0, 10, (B,), generator=g).to(DEV)
    res = []
    for persist in (0, 1):
        torch.manual_seed(0)
        m = nrd.NetResDeep(32, 10, dtype="bf16").to(DEV)
        pl = m.engine.plan(B, torch.device(DEV))
        check(lib.nrd_plan_set_option(pl.handle, 3, persist))
        out = m(x); loss = nrd.CrossEntropyLoss()(out, y); loss.backward()
        res.append((out.detach().clone(), module_grads(m)))
    assert torch.equal(res[0][0], res[1][0])
    for k in res[0][1]:
        a, b = res[0][1][k], res[1][1][k]
        assert rell2(b, a) < 1e-2 and cosine(b, a) > 0.9999, (k, rell2(b, a), cosine(b, a))


@pytest.mark.parametrize("B", [1, 37, 256, 300])
def test_tensor_core_stem_matches_torch_and_cuda_core_stem(B):
    """Stem forward / weight gradient as tcgen05 implicit GEMMs (im2col built in shared memory, pooling across the four
    accumulators of a TMEM lane) against torch on the same bf16-rounded image and weights (<= one bf16 rounding step) and
    against our CUDA-core stem (fp32 image/weights there: bf16 operand rounding is the only difference)."""
    g = torch.Generator().manual_seed(B)
    x = torch.randn(B, 3, 32, 32, generator=g).to(DEV); y = torch.randint(0, 10, (B,), generator=g).to(DEV)
    res = []
    for tc in (0, 1):
        torch.manual_seed(0)
        m = nrd.NetResDeep(32, 2, dtype="bf16").to(DEV)
        pl = m.engine.plan(B, torch.device(DEV))
        check(lib.nrd_plan_set_option(pl.handle, 4, tc))
        out = m(x); loss = nrd.CrossEntropyLoss()(out, y); loss.backward()
        res.append(dict(x0=pl.tensor_view(0, 0).view(torch.bfloat16).float().view(B, 16, 16, 32).cpu().clone(), g=module_grads(m),
                        w=m.conv1.weight.detach().cpu().clone(), b=m.conv1.bias.detach().cpu().clone()))
    cc, tc = res
    ref = F.max_pool2d(torch.relu(F.conv2d(x.cpu().bfloat16().double(), tc["w"].bfloat16().double(), tc["b"].double(), padding=1)), 2)
    assert relmax(tc["x0"].permute(0, 3, 1, 2), ref) < 2 ** -8 * 1.05
    assert rell2(tc["x0"], cc["x0"]) < 1e-2
    # weight-gradient kernel alone, both implementations on the SAME x0 / argmax / incoming gradient (left by the last run)
    part = {}
    for name, opt in (("cc", 0), ("tc", 1)):
        check(lib.nrd_plan_set_option(pl.handle, 4, opt))
        pl.tensor_view(9, 0).zero_()
        check(lib.nrd_plan_run_op(pl.handle, 6, 0, st()), "stem bwd")
        torch.cuda.synchronize()
        part[name] = pl.tensor_view(9, 0).view(torch.float32).view(-1, 28, 32).double().sum(0).cpu()
    assert rell2(part["tc"], part["cc"]) < 1e-2 and cosine(part["tc"], part["cc"]) > 0.9999, rell2(part["tc"], part["cc"])


def test_pool_fused_into_persistent_forward_is_bitwise_the_head_kernels_pool():
    """NRD_OPT_FUSE_POOL: the last block of the persistent forward max-pools its own output (warp shuffles, first maximum
    wins) and fc1 reads the pooled tensor: pooled values, arg-max bytes, logits and gradients equal the default path's."""
    B = 200
    g = torch.Generator().manual_seed(5)
    x = torch.randn(B, 3, 32, 32, generator=g).to(DEV); y = torch.randint(0, 10, (B,), generator=g).to(DEV)
    res = []
    for fuse in (0, 1):
        torch.manual_seed(0)
        m = nrd.NetResDeep(32, 10, dtype="bf16").to(DEV)
        pl = m.engine.plan(B, torch.device(DEV))
        check(lib.nrd_plan_set_option(pl.handle, 8, fuse))
        out = m(x); loss = nrd.CrossEntropyLoss()(out, y); loss.backward()
        torch.cuda.synchronize()
        res.append((out.detach().clone(), pl.tensor_view(3, 0).clone(), m.engine.flat_grads.clone()))
    assert torch.equal(res[0][0], res[1][0]) and torch.equal(res[0][1], res[1][1]) and torch.equal(res[0][2], res[1][2])


@pytest.mark.parametrize("B", [1, 100, 256, 296])
def test_stem_fused_into_persistent_forward_is_bitwise_the_stem_kernel(B):
    """NRD_OPT_FUSE_STEM: x_0 computed inside the persistent forward (same im2col tiles, same MMAs, same max / first-arg-max
    rule as k_stem_tc_fwd): x_0, logits, BN buffers and every gradient (the stem's goes through the arg-max) are equal."""
    g = torch.Generator().manual_seed(B)
    x = torch.randn(B, 3, 32, 32, generator=g).to(DEV); y = torch.randint(0, 10, (B,), generator=g).to(DEV)
    res = []
    for fuse in (0, 1):
        torch.manual_seed(0)
        m = nrd.NetResDeep(32, 10, dtype="bf16").to(DEV)
        pl = m.engine.plan(B, torch.device(DEV))
        check(lib.nrd_plan_set_option(pl.handle, 9, fuse))
        out = m(x); loss = nrd.CrossEntropyLoss()(out, y); loss.backward()
        torch.cuda.synchronize()
        bn = m.resblocks[0].batch_norm
        res.append((out.detach().clone(), pl.tensor_view(0, 0).clone(), m.engine.flat_grads.clone(), bn.running_mean.clone(), bn.running_var.clone()))
    for a_, b_ in zip(res[0], res[1]):
        assert torch.equal(a_, b_)


@pytest.mark.parametrize("B", [3, 256])
def test_one_copy_operand_layouts_match_three_copy_layouts(B):
    """NRD_OPT_ONE_COPY: one haloed operand copy (pixel-granular descriptor starts, 16 x 8 pixel tiles) in the persistent forward
    (bit 0) and backward (bit 1) against the three kx-shifted copies: the same products per output element; only the grouping
    of the fp32 BN partial sums differs (different pixels per warp)."""
    g = torch.Generator().manual_seed(40 + B)
    x = torch.randn(B, 3, 32, 32, generator=g).to(DEV); y = torch.randint(0, 10, (B,), generator=g).to(DEV)
    res = []
    for mode in (0, 1, 3):
        torch.manual_seed(0)
        m = nrd.NetResDeep(32, 10, dtype="bf16").to(DEV)
        pl = m.engine.plan(B, torch.device(DEV))
        check(lib.nrd_plan_set_option(pl.handle, 10, mode))
        out = m(x); loss = nrd.CrossEntropyLoss()(out, y); loss.backward()
        torch.cuda.synchronize()
        res.append((out.detach().clone(), module_grads(m)))
    for other in res[1:]:
        assert rell2(other[0], res[0][0]) < 2e-3
        for k in res[0][1]:
            assert rell2(other[1][k], res[0][1][k]) < 4e-2 and cosine(other[1][k], res[0][1][k]) > 0.999, k
    # forward layout fixed: the backward layouts see identical inputs and must agree closely
    for k in res[1][1]:
        assert rell2(res[2][1][k], res[1][1][k]) < 1e-2 and cosine(res[2][1][k], res[1][1][k]) > 0.9999, k


@pytest.mark.parametrize("B", [256, 149, 40])
def test_fused_stem_backward_equals_stem_kernel_bitwise(B, monkeypatch):
    """NRD_OPT_FUSE_STEM_BWD (an option kept with its numbers: measured no faster): conv1's weight / bias gradient computed in the persistent backward's tail (from dL/dx_0 in registers
    and x_0 in shared memory) against the stand-alone stem backward kernel: same tiles, same MMA order, same CTA -> image map,
    hence bit-identical partials and gradients (and an identical step)."""
    import ddp_cifar10_b200 as nrd
    g = torch.Generator().manual_seed(17)
    x = torch.randn(B, 3, 32, 32, generator=g).to(DEV); y = torch.randint(0, 10, (B,), generator=g).to(DEV)
    res = []
    monkeypatch.setenv("NRD_RESERVE_SMS", "0")       # all 148 SMs: the persistent backward then groups the images per CTA exactly as the
    #                                                  stem kernel does (bit-identical partial sums); TrainStep's default leaves 8 SMs free
    for fused in ("0", "1"):
        monkeypatch.setenv("NRD_FUSE_STEM_BWD", fused)
        torch.manual_seed(4)
        m = nrd.NetResDeep(32, 10, dtype="bf16").to(DEV)
        step = nrd.TrainStep(m, B, lr=1e-2)
        losses = [float(step(x, y)) for _ in range(2)]
        res.append((losses, m.engine.flat_grads.clone(), m.engine.flat_params.clone()))
    off = m.engine.offsets
    assert float(res[0][1][off[0]:off[2]].abs().sum()) > 0
    assert res[0][0] == res[1][0] and torch.equal(res[0][1], res[1][1]) and torch.equal(res[0][2], res[1][2])


@pytest.mark.parametrize("B", [1, 2, 149, 256, 296])
def test_streaming_kernels_are_bitwise_the_resident_persistent_kernels(B):
    """NRD_OPT_STREAM = 1: the streaming persistent forward (every image through two operand slots per block, y_k / x_k read back
    from L2 for the second pass) against the resident one (<= 2 images per CTA kept on chip): same MMAs, same image-to-CTA
    assignment, same order of the BN partial sums, same rounding points -- bit for bit."""
    g = torch.Generator().manual_seed(B)
    x = torch.randn(B, 3, 32, 32, generator=g).to(DEV); y = torch.randint(0, 10, (B,), generator=g).to(DEV)
    res = []
    for stream in (0, 1):
        torch.manual_seed(0)
        m = nrd.NetResDeep(32, 10, dtype="bf16").to(DEV)
        pl = m.engine.plan(B, torch.device(DEV))
        check(lib.nrd_plan_set_option(pl.handle, 13, stream))
        check(lib.nrd_plan_set_option(pl.handle, 10, 3))     # resident kernels in the streaming kernels' operand layout (one haloed copy:
        #                                                      the same pixels per warp, so the same grouping of the fp32 partial sums)
        out = m(x); loss = nrd.CrossEntropyLoss()(out, y); loss.backward()
        bn = m.resblocks[0].batch_norm
        res.append(dict(out=out.detach().clone(), rm=bn.running_mean.clone(), rv=bn.running_var.clone(), nbt=int(bn.num_batches_tracked),
                        xs=[pl.tensor_view(0, k).clone() for k in range(11)], ys=[pl.tensor_view(1, k).clone() for k in range(10)],
                        st=[pl.tensor_view(4, k).clone() for k in range(10)], grads=m.engine.flat_grads.clone()))
    a, b = res
    assert a["nbt"] == b["nbt"] == 10
    for k in range(10):
        assert torch.equal(a["ys"][k], b["ys"][k]), k
        assert torch.equal(a["st"][k], b["st"][k]), k
        assert torch.equal(a["xs"][k + 1], b["xs"][k + 1]), k
    assert torch.equal(a["out"], b["out"]) and torch.equal(a["rm"], b["rm"]) and torch.equal(a["rv"], b["rv"])
    # ... and the streaming backward (dL/dx in one global buffer updated in place, y_k through a shared-memory ring) against the
    # resident one (dL/dx in registers): same per-element arithmetic, same order of every partial sum
    assert torch.equal(a["grads"], b["grads"]), (B, (a["grads"] - b["grads"]).abs().max().item(), rell2(b["grads"], a["grads"]))


@pytest.mark.parametrize("B", [297, 512, 1000])
def test_streaming_kernels_beyond_two_images_per_cta_match_per_block_kernels(B):
    """Default dispatch beyond 2 images per CTA = the streaming kernels; against the per-block kernels (NRD_OPT_STREAM = 2)."""
    g = torch.Generator().manual_seed(B)
    x = torch.randn(B, 3, 32, 32, generator=g).to(DEV); y = torch.randint(0, 10, (B,), generator=g).to(DEV)
    res = []
    for stream in (2, 0):
        torch.manual_seed(0)
        m = nrd.NetResDeep(32, 10, dtype="bf16").to(DEV)
        pl = m.engine.plan(B, torch.device(DEV))
        check(lib.nrd_plan_set_option(pl.handle, 13, stream))
        out = m(x); loss = nrd.CrossEntropyLoss()(out, y); loss.backward()
        bn = m.resblocks[0].batch_norm
        res.append(dict(out=out.detach().clone(), grads=m.engine.flat_grads.clone(), rm=bn.running_mean.clone(), rv=bn.running_var.clone(),
                        nbt=int(bn.num_batches_tracked), x10=pl.tensor_view(0, 10).view(torch.bfloat16).float().clone(),
                        y9=pl.tensor_view(1, 9).view(torch.bfloat16).float().clone(), st9=pl.tensor_view(4, 9).view(torch.float32).clone()))
    a, b = res
    assert a["nbt"] == b["nbt"] == 10
    assert relmax(b["rm"], a["rm"]) < 1e-5 and relmax(b["rv"], a["rv"]) < 1e-5 and relmax(b["st9"], a["st9"]) < 1e-4
    assert rell2(b["y9"], a["y9"]) < 2e-3 and rell2(b["x10"], a["x10"]) < 2e-3 and rell2(b["out"], a["out"]) < 2e-3
    assert rell2(b["grads"], a["grads"]) < 4e-2
