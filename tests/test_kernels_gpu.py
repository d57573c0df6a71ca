"""Numerics of every sm_100a kernel against a plain PyTorch fp32 reference of the same op
(SURVEY §4 kernel tier).  Needs a B200: ``pytest -m gpu``."""
import copy
import math

import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

cuda = torch.device("cuda")


@pytest.fixture(scope="module", autouse=True)
def _lib():
    from pytorch_distributedtraining_b200.ops import require_library
    require_library()  # fail loudly: a GPU box without the kernel library is a broken build


def _params(shapes, seed=0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    out = []
    for s in shapes:
        t = torch.randn(s, generator=g).to(cuda)
        if len(s) == 4:  # conv weights live in channels-last, like the models on the GPU path
            t = t.contiguous(memory_format=torch.channels_last)
        out.append(t.requires_grad_(True))
    return out


SHAPES = [(64, 3, 7, 7), (1000,), (513,), (256, 64, 1, 1), (7,), (2048, 512), (3, 5)]


@pytest.mark.parametrize("wd,maximize", [(1e-4, False), (0.0, False), (1e-2, True)])
def test_fused_adamw_matches_torch(wd, maximize):
    from pytorch_distributedtraining_b200.ops.optim import FusedAdamW
    ps, ref = _params(SHAPES), _params(SHAPES)
    kw = dict(lr=1e-3, betas=(0.9, 0.99), eps=1e-8, weight_decay=wd, maximize=maximize)
    a, b = FusedAdamW(ps, **kw), torch.optim.AdamW(ref, **kw)
    for it in range(5):
        for p, r in zip(ps, ref):
            g = torch.randn_like(p) * (it + 1)  # randn_like preserves the channels-last layout
            p.grad, r.grad = g.clone(), g.clone()
        a.step(); b.step()
    for p, r in zip(ps, ref):
        torch.testing.assert_close(p, r, rtol=2e-5, atol=2e-6)
    sa, sb = a.state_dict(), b.state_dict()
    assert sa["state"].keys() == sb["state"].keys()
    for k in sb["state"]:
        torch.testing.assert_close(sa["state"][k]["exp_avg_sq"], sb["state"][k]["exp_avg_sq"], rtol=1e-5, atol=1e-8)
        assert float(sa["state"][k]["step"]) == float(sb["state"][k]["step"]) == 5.0
    # state dict round trip into the stock optimizer and back
    c = torch.optim.AdamW(_params(SHAPES), **kw); c.load_state_dict(copy.deepcopy(sa))
    d = FusedAdamW(_params(SHAPES), **kw); d.load_state_dict(copy.deepcopy(sb))  # load_state_dict aliases tensors
    for p, r in zip(d.param_groups[0]["params"], ref):
        p.data.copy_(r.data)
    for p, r in zip(d.param_groups[0]["params"], ref):
        g = torch.randn_like(p); p.grad, r.grad = g.clone(), g.clone()
    d.step(); b.step()
    for p, r in zip(d.param_groups[0]["params"], ref):
        torch.testing.assert_close(p, r, rtol=2e-5, atol=2e-6)


@pytest.mark.parametrize("momentum,nesterov,wd", [(0.0, False, 0.0), (0.9, False, 1e-4), (0.9, True, 1e-4)])
def test_fused_sgd_matches_torch(momentum, nesterov, wd):
    from pytorch_distributedtraining_b200.ops.optim import FusedSGD
    ps, ref = _params(SHAPES), _params(SHAPES)
    kw = dict(lr=0.1, momentum=momentum, nesterov=nesterov, weight_decay=wd)
    a, b = FusedSGD(ps, **kw), torch.optim.SGD(ref, **kw)
    for it in range(4):
        for p, r in zip(ps, ref):
            g = torch.randn_like(p); p.grad, r.grad = g.clone(), g.clone()
        a.step(); b.step()
    for p, r in zip(ps, ref):
        torch.testing.assert_close(p, r, rtol=1e-5, atol=1e-6)


def test_fused_unscale_clip_and_device_skip():
    from pytorch_distributedtraining_b200.ops.optim import FusedAdamW
    from pytorch_distributedtraining_b200.parallel import GradScaler
    ps, ref = _params(SHAPES), _params(SHAPES)
    kw = dict(lr=1e-3, betas=(0.9, 0.99), eps=1e-8, weight_decay=1e-4)
    a, b = FusedAdamW(ps, **kw), torch.optim.AdamW(ref, **kw)
    scaler = GradScaler(init_scale=1024.0, growth_interval=3)
    scaler._lazy_init(cuda)
    for it in range(4):
        for p, r in zip(ps, ref):
            g = torch.randn_like(p)
            p.grad, r.grad = g * scaler.get_scale(), g.clone()
        total = scaler.unscale_and_clip_fused(a, 0.5)
        ref_total = torch.nn.utils.clip_grad_norm_(ref, 0.5)
        torch.testing.assert_close(total, ref_total, rtol=1e-4, atol=1e-5)
        scaler.step(a); scaler.update(); b.step()
    for p, r in zip(ps, ref):
        torch.testing.assert_close(p, r, rtol=3e-5, atol=3e-6)
    assert scaler.get_scale() == 2048.0
    # an inf anywhere: the step is skipped on the device, step count does not advance, scale backs off
    before = [p.detach().clone() for p in ps]
    for p in ps:
        p.grad = torch.randn_like(p)
    ps[3].grad.view(-1)[5] = float("inf")
    scaler.unscale_and_clip_fused(a, None)
    scaler.step(a); scaler.update()
    for p, q in zip(ps, before):
        assert torch.equal(p, q)
    assert float(a.state[ps[0]]["step"]) == 4.0 and scaler.get_scale() == 1024.0
    # plain unscale_ kernel
    gs = [torch.randn(1001, device=cuda), torch.randn(64, 64, device=cuda)]
    want = [g / 8 for g in gs]
    from pytorch_distributedtraining_b200.ops import fused
    flag = torch.zeros((), device=cuda)
    fused.unscale_check_(gs, torch.tensor(0.125, device=cuda), flag)
    for g, w in zip(gs, want):
        torch.testing.assert_close(g, w)
    assert flag.item() == 0.0
    gs[1][3, 3] = float("nan")
    fused.unscale_check_(gs, torch.tensor(1.0, device=cuda), flag)
    assert flag.item() == 1.0


@pytest.mark.parametrize("n", [1, 1000, 16384, 1 << 20, (1 << 22) + 13])
def test_sumsq_matches_torch(n):
    from pytorch_distributedtraining_b200.ops import fused
    x = torch.randn(n, device=cuda)
    for _ in range(2):  # second call checks the completion counter was left at zero
        got = fused.sumsq_f32(x)
        torch.testing.assert_close(got, (x.double() ** 2).sum().float(), rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("dtype", [torch.float32, torch.float16, torch.bfloat16])
@pytest.mark.parametrize("rows,classes,smooth", [(256, 1000, 0.0), (7, 10, 0.1), (32, 5000, 0.0)])
def test_cross_entropy_matches_torch(dtype, rows, classes, smooth):
    from pytorch_distributedtraining_b200.ops.loss import cross_entropy
    x = (torch.randn(rows, classes, device=cuda) * 3).to(dtype).requires_grad_(True)
    y = torch.randint(0, classes, (rows,), device=cuda)
    xr = x.detach().float().requires_grad_(True)
    loss = cross_entropy(x, y, smooth)
    ref = F.cross_entropy(xr, y, label_smoothing=smooth)
    torch.testing.assert_close(loss, ref, rtol=1e-4, atol=1e-5)
    (loss * 3.0).backward(); (ref * 3.0).backward()
    tol = 1e-6 if dtype == torch.float32 else 2e-3
    torch.testing.assert_close(x.grad.float(), xr.grad, rtol=1e-2 if dtype != torch.float32 else 1e-4, atol=tol)


@pytest.mark.parametrize("dtype", [torch.float32, torch.float16])
def test_mse_matches_torch(dtype):
    from pytorch_distributedtraining_b200.ops.loss import mse_loss
    a = torch.randn(4, 3, 33, 17, device=cuda).to(dtype).requires_grad_(True)
    b = torch.randn(4, 3, 33, 17, device=cuda).to(dtype)
    ar = a.detach().float().requires_grad_(True)
    l, r = mse_loss(a, b), F.mse_loss(ar, b.float())
    torch.testing.assert_close(l, r, rtol=1e-4, atol=1e-6)
    l.backward(); r.backward()
    torch.testing.assert_close(a.grad.float(), ar.grad, rtol=1e-2, atol=1e-5)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("shape,res,act", [((8, 64, 14, 14), False, "relu"), ((4, 256, 7, 7), True, "relu"),
                                            ((3, 2048, 5, 5), True, "relu"), ((16, 24, 9, 9), False, None),
                                            ((2, 512, 28, 28), True, "relu")])
def test_batchnorm_act_matches_torch(dtype, shape, res, act):
    from pytorch_distributedtraining_b200.ops.layers import BatchNormAct2d
    torch.manual_seed(1)
    C = shape[1]
    bn = BatchNormAct2d(C, act=act).to(cuda)
    with torch.no_grad():
        bn.weight.uniform_(0.5, 1.5); bn.bias.uniform_(-0.5, 0.5)
    ref = copy.deepcopy(bn).float()
    x = torch.randn(shape, device=cuda).to(dtype).contiguous(memory_format=torch.channels_last).requires_grad_(True)
    r = torch.randn(shape, device=cuda).to(dtype).contiguous(memory_format=torch.channels_last).requires_grad_(True) if res else None
    xr = x.detach().float().requires_grad_(True)
    rr = r.detach().float().requires_grad_(True) if res else None
    y = bn(x, r)
    yr = F.batch_norm(xr, ref.running_mean, ref.running_var, ref.weight, ref.bias, True, 0.1, ref.eps)
    if res:
        yr = yr + rr
    if act == "relu":
        yr = F.relu(yr)
    tol = dict(rtol=2e-2, atol=2e-2) if dtype == torch.bfloat16 else dict(rtol=5e-3, atol=5e-3)
    torch.testing.assert_close(y.float(), yr, **tol)
    torch.testing.assert_close(bn.running_mean, ref.running_mean, rtol=1e-3, atol=1e-4)
    torch.testing.assert_close(bn.running_var, ref.running_var, rtol=1e-3, atol=1e-4)
    g = torch.randn(shape, device=cuda).to(dtype).contiguous(memory_format=torch.channels_last)
    # compare backward through the same mask: use the reference's own forward for the reference
    y.backward(g); yr.backward(g.float())
    gtol = dict(rtol=5e-2, atol=5e-2) if dtype == torch.bfloat16 else dict(rtol=1e-2, atol=1e-2)
    # elements whose pre-activation is within rounding of zero may flip the ReLU mask between
    # the 16-bit kernel and the fp32 reference; allow a tiny fraction of mismatches
    def close_frac(a, b, rtol, atol):
        return ((a - b).abs() <= atol + rtol * b.abs()).float().mean().item()
    assert close_frac(x.grad.float(), xr.grad, **gtol) > 0.995
    if res:
        assert close_frac(r.grad.float(), rr.grad, **gtol) > 0.995
    n = shape[0] * shape[2] * shape[3]
    torch.testing.assert_close(bn.weight.grad, ref.weight.grad, rtol=3e-2, atol=3e-2 * math.sqrt(n))
    torch.testing.assert_close(bn.bias.grad, ref.bias.grad, rtol=3e-2, atol=3e-2 * math.sqrt(n))
    # eval mode
    bn.eval()
    ye = bn(x.detach(), r.detach() if res else None)
    yer = F.batch_norm(xr.detach(), ref.running_mean, ref.running_var, ref.weight, ref.bias, False, 0.0, ref.eps)
    if res:
        yer = yer + rr.detach()
    if act == "relu":
        yer = F.relu(yer)
    torch.testing.assert_close(ye.float(), yer, **tol)


@pytest.mark.parametrize("dtype", [torch.float32, torch.float16, torch.bfloat16])
@pytest.mark.parametrize("rows,D", [(197 * 4, 768), (33, 64), (5, 2048), (1000, 384), (4099, 60), (77, 180), (9, 4)])
def test_layernorm_matches_torch(dtype, rows, D):
    from pytorch_distributedtraining_b200.ops.layers import LayerNorm
    torch.manual_seed(2)
    ln = LayerNorm(D, eps=1e-6).to(cuda)
    with torch.no_grad():
        ln.weight.uniform_(0.5, 1.5); ln.bias.uniform_(-0.5, 0.5)
    ref = copy.deepcopy(ln)
    x = torch.randn(rows, D, device=cuda).to(dtype).requires_grad_(True)
    xr = x.detach().float().requires_grad_(True)
    y = ln(x)
    yr = F.layer_norm(xr, (D,), ref.weight, ref.bias, 1e-6)
    tol = {torch.float32: 1e-5, torch.float16: 5e-3, torch.bfloat16: 3e-2}[dtype]
    torch.testing.assert_close(y.float(), yr, rtol=tol, atol=tol)
    g = torch.randn(rows, D, device=cuda).to(dtype)
    y.backward(g); yr.backward(g.float())
    torch.testing.assert_close(x.grad.float(), xr.grad, rtol=tol * 4, atol=tol * 4)
    torch.testing.assert_close(ln.weight.grad, ref.weight.grad, rtol=tol * 10, atol=tol * 4 * math.sqrt(rows))
    torch.testing.assert_close(ln.bias.grad, ref.bias.grad, rtol=tol * 10, atol=tol * 4 * math.sqrt(rows))


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("rows,D,P", [(4099, 60, 64), (77, 180, 192), (9, 4, 8), (300, 60, 60)])
def test_layernorm_on_zero_padded_rows(dtype, rows, D, P):
    """LayerNorm over the first D elements of rows that are P wide (``LayerNorm.phys``): statistics over D only, padding
    of the output and of the input gradient comes out as zeros."""
    from pytorch_distributedtraining_b200.ops.layers import LayerNorm
    torch.manual_seed(3)
    ln = LayerNorm(D, eps=1e-5).to(cuda)
    ln.phys = P
    with torch.no_grad():
        ln.weight.uniform_(0.5, 1.5); ln.bias.uniform_(-0.5, 0.5)
    ref = copy.deepcopy(ln)
    x = torch.randn(rows, P, device=cuda).to(dtype)
    x[:, D:] = 0
    x.requires_grad_(True)
    xr = x.detach().float()[:, :D].clone().requires_grad_(True)
    y = ln(x)
    yr = F.layer_norm(xr, (D,), ref.weight, ref.bias, 1e-5)
    tol = {torch.float32: 1e-5, torch.bfloat16: 3e-2}[dtype]
    torch.testing.assert_close(y.float()[:, :D], yr, rtol=tol, atol=tol)
    assert (y[:, D:] == 0).all()
    g = torch.randn(rows, P, device=cuda).to(dtype)
    y.backward(g); yr.backward(g.float()[:, :D])
    torch.testing.assert_close(x.grad.float()[:, :D], xr.grad, rtol=tol * 4, atol=tol * 4)
    assert (x.grad[:, D:] == 0).all()
    torch.testing.assert_close(ln.weight.grad, ref.weight.grad, rtol=tol * 10, atol=tol * 4 * math.sqrt(rows))
    torch.testing.assert_close(ln.bias.grad, ref.bias.grad, rtol=tol * 10, atol=tol * 4 * math.sqrt(rows))


def test_resnet18_fused_path_matches_stock_ops():
    """Whole-model check: the fused BN/ReLU/CE path must be as close to an fp32 reference as
    the same model on stock torch fp16 ops is (both carry fp16 rounding; neither is exact)."""
    from pytorch_distributedtraining_b200.models import resnet18
    from pytorch_distributedtraining_b200.ops import dispatch
    from pytorch_distributedtraining_b200.ops.loss import cross_entropy
    torch.manual_seed(0)
    m = resnet18(num_classes=10, small_input=True).to(cuda).to(memory_format=torch.channels_last)
    stock, exact = copy.deepcopy(m), copy.deepcopy(m)
    x = torch.randn(16, 3, 32, 32, device=cuda).contiguous(memory_format=torch.channels_last)
    y = torch.randint(0, 10, (16,), device=cuda)
    with torch.autocast("cuda", dtype=torch.float16):
        l1 = cross_entropy(m(x), y)
    l1.backward()
    feats = ("bn", "ce", "gemm", "conv", "ln")
    for f in feats:
        dispatch.disable(f)
    try:
        with torch.autocast("cuda", dtype=torch.float16):
            l2 = F.cross_entropy(stock(x).float(), y)
        l2.backward()
        l3 = F.cross_entropy(exact(x), y)   # fp32 everywhere
        l3.backward()
    finally:
        for f in feats:
            dispatch.enable(f)

    def rel(model):
        num = sum(((a.grad - b.grad) ** 2).sum() for a, b in zip(model.parameters(), exact.parameters()))
        den = sum((b.grad ** 2).sum() for b in exact.parameters())
        return (num / den).sqrt().item()

    e_fused, e_stock = rel(m), rel(stock)
    print(f"grad rel err vs fp32: fused {e_fused:.4f}, stock fp16 {e_stock:.4f}")
    torch.testing.assert_close(l1, l3, rtol=2e-2, atol=2e-2)
    assert e_fused < max(1.5 * e_stock, 2e-2), (e_fused, e_stock)


def test_bn_statistics_arena_and_step_counter_over_several_steps():
    """The conv -> BatchNorm statistics vectors are slices of one block zeroed once per optimizer step, and
    ``num_batches_tracked`` is incremented inside the apply kernel: several steps (with and without the parameter-epoch
    signal between them, and past the end of a block) must keep matching fp32 BatchNorm statistics."""
    from pytorch_distributedtraining_b200.models import resnet18
    from pytorch_distributedtraining_b200.ops import conv as _conv, dispatch, fused
    torch.manual_seed(0)
    m = resnet18(num_classes=10, small_input=True).to(cuda).to(memory_format=torch.channels_last)
    exact = copy.deepcopy(m)
    arena = _conv._STATS_ARENAS.get(torch.device("cuda", torch.cuda.current_device()))
    for step in range(5):
        x = torch.randn(16, 3, 32, 32, device=cuda).contiguous(memory_format=torch.channels_last)
        with torch.autocast("cuda", dtype=torch.float16):
            m(x).float().square().mean().backward()
        for f in ("bn", "ce", "gemm", "conv", "ln"):
            dispatch.disable(f)
        try:
            exact(x).square().mean().backward()
        finally:
            for f in ("bn", "ce", "gemm", "conv", "ln"):
                dispatch.enable(f)
        if step in (0, 2):
            fused.bump_param_epoch()                     # "an optimizer step happened": the block is re-zeroed and reused
        if step == 3:
            arena = _conv._STATS_ARENAS[x.device]
            arena.cursor = arena.buf.numel() - 64        # the next request does not fit: a fresh zeroed block
    bns = [(a, b) for a, b in zip(m.modules(), exact.modules()) if isinstance(a, torch.nn.BatchNorm2d)]
    assert len(bns) >= 10
    for a, b in bns:
        assert int(a.num_batches_tracked) == 5 == int(b.num_batches_tracked)
        torch.testing.assert_close(a.running_mean, b.running_mean, rtol=2e-2, atol=2e-2)
        torch.testing.assert_close(a.running_var, b.running_var, rtol=3e-2, atol=3e-2)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("shape,k,st,pd", [((4, 64, 112, 112), 3, 2, 1), ((2, 16, 9, 11), 3, 2, 1), ((3, 8, 8, 8), 2, 2, 0),
                                            ((2, 24, 13, 7), 3, 1, 1)])
def test_maxpool_matches_torch_including_ties(dtype, shape, k, st, pd):
    from pytorch_distributedtraining_b200.ops.layers import MaxPool2d
    torch.manual_seed(3)
    # post-ReLU style input: many exact ties (zeros), where the argmax convention matters
    x = torch.relu(torch.randn(shape, device=cuda)).to(dtype).contiguous(memory_format=torch.channels_last).requires_grad_(True)
    xr = x.detach().clone().requires_grad_(True)
    pool = MaxPool2d(k, st, pd)
    y = pool(x)
    yr = F.max_pool2d(xr, k, st, pd)
    assert torch.equal(y, yr)
    g = torch.randn_like(yr)
    y.backward(g); yr.backward(g)
    torch.testing.assert_close(x.grad.float(), xr.grad.float(), rtol=1e-2, atol=1e-2)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("B,S,H,d", [(3, 197, 12, 64), (2, 50, 6, 32), (1, 7, 3, 8)])
def test_attention_layout_kernels_and_autograd(dtype, B, S, H, d):
    """split_qkv / merge_heads: forward values and the packed gradient equal the plain reshape /
    permute formulation; the streaming kernels are exercised directly as well."""
    import torch.nn.functional as F
    from pytorch_distributedtraining_b200.ops import attention as A
    dev = torch.device("cuda")
    g = torch.Generator().manual_seed(B * S)
    heads = [torch.randn(B, H, S, d, generator=g).to(dtype).to(dev) for _ in range(3)]
    packed = A._heads_to_tokens(heads, B, H, S, d)
    want = torch.stack([h.permute(0, 2, 1, 3) for h in heads], dim=2)
    assert torch.equal(packed, want)
    back = A._tokens_to_heads(packed, B, H, S, d, 3)
    for a, b in zip(back, heads):
        assert torch.equal(a, b)
    x = (torch.randn(B, S, 3 * H * d, generator=g) * 0.5).to(dtype).to(dev).requires_grad_(True)
    q, k, v = A.split_qkv(x, H)
    y = A.merge_heads(F.scaled_dot_product_attention(q, k, v))
    y.float().square().sum().backward()
    g1, x.grad = x.grad.clone(), None
    qkv = x.view(B, S, 3, H, d).permute(2, 0, 3, 1, 4)
    y2 = F.scaled_dot_product_attention(qkv[0], qkv[1], qkv[2]).transpose(1, 2).reshape(B, S, H * d)
    y2.float().square().sum().backward()
    torch.testing.assert_close(y, y2, rtol=1e-2, atol=1e-2)
    torch.testing.assert_close(g1, x.grad, rtol=2e-2, atol=2e-2)


@pytest.mark.parametrize("embed_dim", [96, 60])
def test_swinir_on_kernel_path_matches_stock_ops(embed_dim):
    """SwinIR on the kernel path (tcgen05 linears / convolutions, fused LayerNorm, tcgen05 window attention) vs the same
    weights on stock torch ops.  embed_dim 96: widths inside the kernels' envelope; embed_dim 60: the reference's own
    SwinIR-S (Stoke-DDP.py:206-208) — head dimension 10, every width padded to the next multiple of 8 / 64."""
    from pytorch_distributedtraining_b200.models.swinir import SwinIR
    from pytorch_distributedtraining_b200.ops import _lib, dispatch
    torch.manual_seed(2)
    net = SwinIR(upscale=2, in_chans=3, img_size=64, window_size=8, img_range=1., depths=[2, 2], embed_dim=embed_dim,
                 num_heads=[6, 6], mlp_ratio=2, upsampler="pixelshuffledirect", resi_connection="1conv").cuda()
    x = torch.rand(4, 3, 64, 64, device="cuda")
    outs = []
    for stock in (False, True):
        net.zero_grad(set_to_none=True)
        feats = ("gemm", "ln", "attn_layout", "mlp", "conv", "attn", "pad")
        launches0 = _lib.LAUNCHES
        if stock:
            for f in feats:
                dispatch.disable(f)
        try:
            with torch.autocast("cuda", dtype=torch.bfloat16):
                y = net(x)
            y.float().square().mean().backward()
        finally:
            for f in feats:
                dispatch.enable(f)
        g = torch.cat([p.grad.flatten().float() for p in net.parameters() if p.grad is not None])
        outs.append((y.float().detach(), g))
        if not stock:
            # 4 blocks x (qkv, proj, fc1, fc2 forward + 2 backward GEMMs each, attention fwd + bwd) + convolutions
            assert _lib.LAUNCHES - launches0 >= 4 * (4 * 3 + 2), _lib.LAUNCHES - launches0
    assert torch.isfinite(outs[0][0]).all() and torch.isfinite(outs[0][1]).all()
    torch.testing.assert_close(outs[0][0], outs[1][0], rtol=5e-2, atol=5e-2)
    rel = (outs[0][1] - outs[1][1]).norm() / outs[1][1].norm()
    assert rel < 5e-2, rel
