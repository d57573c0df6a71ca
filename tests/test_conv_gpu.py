"""Implicit-GEMM convolution (tcgen05 + TMA) vs torch's convolution in fp32 on the same 16-bit data."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
cuda = torch.device("cuda")

CASES = [
    # N, H, W, Cin, Cout, R, stride, pad
    (2, 56, 56, 64, 64, 3, 1, 1),
    (3, 28, 28, 128, 128, 3, 1, 1),
    (4, 14, 14, 256, 256, 3, 1, 1),
    (5, 7, 7, 512, 512, 3, 1, 1),
    (2, 56, 56, 128, 128, 3, 2, 1),
    (2, 56, 56, 256, 512, 1, 2, 0),
    (1, 17, 23, 64, 72, 3, 1, 1),
    (2, 16, 16, 64, 40, 5, 1, 2),
    (1, 130, 130, 64, 64, 3, 1, 1),
    (2, 9, 9, 192, 264, 3, 2, 1),
    (32, 7, 7, 512, 512, 3, 1, 1),
    (2, 20, 20, 192, 136, 3, 1, 1),
]


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("N,H,W,Cin,Cout,R,stride,pad", CASES)
def test_conv_forward_matches_reference(dtype, N, H, W, Cin, Cout, R, stride, pad):
    from pytorch_distributedtraining_b200.ops.conv import conv2d_raw
    g = torch.Generator().manual_seed(N * 1000 + H)
    x = (torch.randn(N, Cin, H, W, generator=g) * 0.5).to(dtype).to(cuda).contiguous(memory_format=torch.channels_last)
    w = (torch.randn(Cout, Cin, R, R, generator=g) * 0.1).to(dtype).to(cuda).contiguous(memory_format=torch.channels_last)
    b = torch.randn(Cout, generator=g).to(cuda)
    want = F.conv2d(x.float(), w.float(), b, stride=stride, padding=pad)
    got = conv2d_raw(0, x, w, b, N, H, W, Cin, Cout, R, R, stride, pad)
    torch.cuda.synchronize()
    assert got.shape == want.shape and got.is_contiguous(memory_format=torch.channels_last)
    tol = 3e-2 if dtype == torch.bfloat16 else 5e-3
    torch.testing.assert_close(got.float(), want, rtol=tol, atol=tol * (Cin * R * R) ** 0.5 * 0.05)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("N,H,W,Cin,Cout,R,stride,pad", [c for c in CASES if c[6] == 1 and c[4] % 64 == 0])
def test_conv_dgrad_matches_reference(dtype, N, H, W, Cin, Cout, R, stride, pad):
    from pytorch_distributedtraining_b200.ops.conv import conv2d_raw
    g = torch.Generator().manual_seed(N * 77 + H)
    P = (H + 2 * pad - R) + 1
    Q = (W + 2 * pad - R) + 1
    dy = (torch.randn(N, Cout, P, Q, generator=g) * 0.5).to(dtype).to(cuda).contiguous(memory_format=torch.channels_last)
    w = (torch.randn(Cout, Cin, R, R, generator=g) * 0.1).to(dtype).to(cuda).contiguous(memory_format=torch.channels_last)
    want = torch.nn.grad.conv2d_input((N, Cin, H, W), w.float(), dy.float(), stride=1, padding=pad)
    got = conv2d_raw(1, dy, w, None, N, H, W, Cin, Cout, R, R, 1, pad)
    torch.cuda.synchronize()
    tol = 3e-2 if dtype == torch.bfloat16 else 5e-3
    torch.testing.assert_close(got.float(), want, rtol=tol, atol=tol * (Cout * R * R) ** 0.5 * 0.05)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("N,H,W,Cin,Cout,R,stride,pad", CASES)
def test_conv_wgrad_matches_reference(dtype, N, H, W, Cin, Cout, R, stride, pad):
    from pytorch_distributedtraining_b200.ops.conv import conv2d_wgrad_raw
    g = torch.Generator().manual_seed(N * 31 + H)
    P = (H + 2 * pad - R) // stride + 1
    Q = (W + 2 * pad - R) // stride + 1
    x = (torch.randn(N, Cin, H, W, generator=g) * 0.5).to(dtype).to(cuda).contiguous(memory_format=torch.channels_last)
    dy = (torch.randn(N, Cout, P, Q, generator=g) * 0.5).to(dtype).to(cuda).contiguous(memory_format=torch.channels_last)
    want = torch.nn.grad.conv2d_weight(x.float(), (Cout, Cin, R, R), dy.float(), stride=stride, padding=pad)
    got = conv2d_wgrad_raw(x, dy, N, H, W, Cin, Cout, R, R, stride, pad)
    torch.cuda.synchronize()
    assert got.shape == want.shape and got.dtype == torch.float32
    rel = ((got - want).norm() / want.norm()).item()
    assert rel < 2e-3, rel
    torch.testing.assert_close(got, want, rtol=1e-2, atol=1e-2 * (N * P * Q) ** 0.5 * 0.25)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_conv_module_autograd(dtype):
    """Conv2d modules through autocast: forward/backward vs the same modules on the library path."""
    import copy
    from pytorch_distributedtraining_b200.ops import dispatch
    from pytorch_distributedtraining_b200.ops.layers import Conv2d
    torch.manual_seed(0)
    net = torch.nn.Sequential(Conv2d(64, 128, 3, 1, 1, bias=False), torch.nn.ReLU(), Conv2d(128, 128, 3, 2, 1, bias=True),
                              torch.nn.ReLU(), Conv2d(128, 64, 1, 1, 0, bias=False)).to(cuda).to(memory_format=torch.channels_last)
    ref = copy.deepcopy(net)
    x = torch.randn(4, 64, 20, 20, device=cuda).contiguous(memory_format=torch.channels_last).requires_grad_(True)
    xr = x.detach().clone().requires_grad_(True)
    with torch.autocast("cuda", dtype=dtype):
        y = net(x)
    dispatch.disable("conv")
    try:
        with torch.autocast("cuda", dtype=dtype):
            yr = ref(xr)
    finally:
        dispatch.enable("conv")
    tol = 4e-2 if dtype == torch.bfloat16 else 8e-3
    torch.testing.assert_close(y.float(), yr.float(), rtol=tol, atol=tol)
    gr = torch.randn_like(yr)
    y.backward(gr.to(y.dtype)); yr.backward(gr)
    torch.testing.assert_close(x.grad, xr.grad, rtol=tol, atol=tol * 4)
    for a, b in zip(net.parameters(), ref.parameters()):
        # weight gradients sum thousands of 16-bit products on both sides: compare in relative L2
        rel = ((a.grad - b.grad).norm() / b.grad.norm()).item()
        assert rel < tol, rel


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("N,H,W,Cin,Cout,R,stride,pad", CASES)
def test_conv_epilogue_emits_batchnorm_statistics(dtype, N, H, W, Cin, Cout, R, stride, pad):
    """The forward kernel's optional epilogue output: per-channel sum and sum of squares of exactly the
    values it stored (ragged tiles, strided and padded shapes included)."""
    from pytorch_distributedtraining_b200.ops.conv import conv2d_raw
    g = torch.Generator().manual_seed(N * 13 + H)
    x = (torch.randn(N, Cin, H, W, generator=g) * 0.5).to(dtype).to(cuda).contiguous(memory_format=torch.channels_last)
    w = (torch.randn(Cout, Cin, R, R, generator=g) * 0.1).to(dtype).to(cuda).contiguous(memory_format=torch.channels_last)
    stats = torch.zeros(2 * Cout + 1, device=cuda)
    got = conv2d_raw(0, x, w, None, N, H, W, Cin, Cout, R, R, stride, pad, colstats=stats)
    plain = conv2d_raw(0, x, w, None, N, H, W, Cin, Cout, R, R, stride, pad)
    torch.cuda.synchronize()
    assert torch.equal(got, plain), "the statistics epilogue must not change the stored output"
    y = got.double()
    torch.testing.assert_close(stats[:Cout].double(), y.sum(dim=(0, 2, 3)), rtol=2e-4, atol=2e-3)
    torch.testing.assert_close(stats[Cout:2 * Cout].double(), (y * y).sum(dim=(0, 2, 3)), rtol=2e-4, atol=2e-3)
    assert float(stats[2 * Cout]) == 0.0


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("M,N,K", [(4096, 256, 64), (3000, 72, 128), (25088, 512, 128), (130, 1024, 256)])
def test_gemm_epilogue_emits_column_statistics(dtype, M, N, K):
    from pytorch_distributedtraining_b200.ops.gemm import gemm
    g = torch.Generator().manual_seed(M + N)
    a = (torch.randn(M, K, generator=g) * 0.5).to(dtype).to(cuda)
    b = (torch.randn(N, K, generator=g) * 0.2).to(dtype).to(cuda)
    stats = torch.zeros(2 * N + 1, device=cuda)
    got = gemm(a, b, M, N, K, colstats=stats)
    plain = gemm(a, b, M, N, K)
    torch.cuda.synchronize()
    assert torch.equal(got, plain)
    y = got.double()
    torch.testing.assert_close(stats[:N].double(), y.sum(0), rtol=2e-4, atol=2e-3)
    torch.testing.assert_close(stats[N:2 * N].double(), (y * y).sum(0), rtol=2e-4, atol=2e-3)


def test_conv_bn_pair_matches_unfused_statistics():
    """A bottleneck block trained with the conv -> BatchNorm statistics hand-over gives the same outputs,
    running statistics and gradients as with the BatchNorm's own reduction pass."""
    from pytorch_distributedtraining_b200.models.resnet import Bottleneck
    from pytorch_distributedtraining_b200.ops import dispatch
    from pytorch_distributedtraining_b200.ops.layers import BatchNormAct2d, Conv2d, pair_conv_bn
    import torch.nn as nn
    torch.manual_seed(3)
    down = nn.Sequential(Conv2d(64, 256, 1, stride=2, bias=False), BatchNormAct2d(256, act=None))
    blk = Bottleneck(64, 64, stride=2, downsample=down).to(cuda).to(memory_format=torch.channels_last)
    assert pair_conv_bn(blk) == 4
    x = torch.randn(8, 64, 30, 30, device=cuda).contiguous(memory_format=torch.channels_last)
    res = {}
    state = {k: v.clone() for k, v in blk.state_dict().items()}
    for mode in ("fused", "plain"):
        blk.load_state_dict(state)
        blk.zero_grad(set_to_none=True)
        if mode == "plain":
            dispatch.disable("convstats")
        try:
            with torch.autocast("cuda", dtype=torch.float16):
                y = blk(x)
            (y.float() ** 2).mean().backward()
        finally:
            dispatch.enable("convstats")
        res[mode] = (y.detach().float(), blk.bn2.running_var.clone(), blk.conv1.weight.grad.clone(),
                     blk.bn3.weight.grad.clone())
    # the two paths sum the statistics in a different order; a 1e-7 difference there can flip the fp16
    # rounding of an activation two layers later, so compare at a few fp16 ulps
    for a, b in zip(res["fused"], res["plain"]):
        torch.testing.assert_close(a, b, rtol=1e-2, atol=1e-2)


def test_residual_gradient_accumulates_in_the_dgrad_epilogue():
    """Bottleneck blocks route the bypass branch through conv1 (``forward_carry``): conv1's input gradient
    is reduce-added onto the bypass gradient by the GEMM epilogue.  Same gradients as the plain autograd
    formulation where the engine adds the two branches itself."""
    from pytorch_distributedtraining_b200.models.resnet import Bottleneck
    from pytorch_distributedtraining_b200.ops.layers import BatchNormAct2d, Conv2d
    import torch.nn as nn
    torch.manual_seed(5)
    blocks = [Bottleneck(256, 64).to(cuda).to(memory_format=torch.channels_last),
              Bottleneck(64, 64, stride=1, downsample=nn.Sequential(Conv2d(64, 256, 1, bias=False), BatchNormAct2d(256, act=None))
                         ).to(cuda).to(memory_format=torch.channels_last)]
    for blk in blocks:
        cin = blk.conv1.in_channels
        x0 = torch.randn(32, cin, 28, 28, device=cuda).half().contiguous(memory_format=torch.channels_last)
        res = []
        for plain in (False, True):
            x = x0.clone().requires_grad_(True)
            blk.zero_grad(set_to_none=True)
            if plain:
                blk.conv1.forward_carry = lambda t, c=blk.conv1: (c.forward(t), t)      # engine adds the branches
            with torch.autocast("cuda", dtype=torch.float16):
                y = blk(x * 1.0)
            (y.float() ** 2).mean().backward()
            if plain:
                del blk.conv1.forward_carry
            res.append((x.grad.float().clone(), blk.conv1.weight.grad.clone()))
        scale = res[1][0].abs().max()
        torch.testing.assert_close(res[0][0], res[1][0], rtol=2e-2, atol=float(scale) * 4e-3)
        torch.testing.assert_close(res[0][1], res[1][1], rtol=2e-2, atol=float(res[1][1].abs().max()) * 1e-2)


PAIR_CASES = [
    # batches large enough for the cta_group::2 kernel (>= 74 pairs of 128-pixel tiles)
    (64, 28, 28, 128, 128, 3, 1, 1),
    (128, 14, 14, 256, 256, 3, 1, 1),
    (130, 14, 14, 256, 512, 1, 2, 0),
    (67, 27, 29, 128, 192, 3, 1, 1),          # ragged pixel tiles, odd tile count, ragged channel tile
    (128, 28, 28, 256, 256, 3, 2, 1),
]


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("N,H,W,Cin,Cout,R,stride,pad", PAIR_CASES)
def test_conv_cta_pair_kernel(dtype, N, H, W, Cin, Cout, R, stride, pad):
    """Forward (+ statistics epilogue) and stride-1 input gradient on the CTA-pair implicit-GEMM kernel."""
    from pytorch_distributedtraining_b200.ops.conv import conv2d_raw
    g = torch.Generator().manual_seed(N + H)
    x = (torch.randn(N, Cin, H, W, generator=g) * 0.5).to(dtype).to(cuda).contiguous(memory_format=torch.channels_last)
    w = (torch.randn(Cout, Cin, R, R, generator=g) * 0.1).to(dtype).to(cuda).contiguous(memory_format=torch.channels_last)
    want = F.conv2d(x.float(), w.float(), None, stride=stride, padding=pad)
    stats = torch.zeros(2 * Cout + 1, device=cuda)
    got = conv2d_raw(0, x, w, None, N, H, W, Cin, Cout, R, R, stride, pad, colstats=stats)
    plain = conv2d_raw(0, x, w, None, N, H, W, Cin, Cout, R, R, stride, pad)
    torch.cuda.synchronize()
    tol = 3e-2 if dtype == torch.bfloat16 else 5e-3
    torch.testing.assert_close(got.float(), want, rtol=tol, atol=tol * (Cin * R * R) ** 0.5 * 0.05)
    assert torch.equal(got, plain)
    y = got.double()
    torch.testing.assert_close(stats[:Cout].double(), y.sum(dim=(0, 2, 3)), rtol=2e-4, atol=2e-2)
    torch.testing.assert_close(stats[Cout:2 * Cout].double(), (y * y).sum(dim=(0, 2, 3)), rtol=2e-4, atol=2e-2)
    if stride == 1 and Cout % 64 == 0:
        dy = (torch.randn(want.shape, generator=g) * 0.5).to(dtype).to(cuda).contiguous(memory_format=torch.channels_last)
        want_dx = torch.nn.grad.conv2d_input((N, Cin, H, W), w.float(), dy.float(), stride=1, padding=pad)
        got_dx = conv2d_raw(1, dy, w, None, N, H, W, Cin, Cout, R, R, 1, pad)
        torch.cuda.synchronize()
        torch.testing.assert_close(got_dx.float(), want_dx, rtol=tol, atol=tol * (Cout * R * R) ** 0.5 * 0.05)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("N,H,W,Cin,Cout,R,pad", [
    (2, 56, 56, 128, 128, 3, 1),
    (2, 56, 56, 256, 512, 1, 0),        # downsample 1x1: three of the four parity classes are empty (zero-filled)
    (3, 15, 17, 64, 128, 3, 1),         # odd sizes: the classes have different extents
    (2, 28, 28, 256, 256, 3, 1),
    (1, 16, 16, 64, 64, 5, 2),
])
def test_conv_strided_dgrad_by_parity_classes(dtype, N, H, W, Cin, Cout, R, pad):
    from pytorch_distributedtraining_b200.ops.conv import conv2d_raw
    g = torch.Generator().manual_seed(N * 5 + H)
    P, Q = (H + 2 * pad - R) // 2 + 1, (W + 2 * pad - R) // 2 + 1
    dy = (torch.randn(N, Cout, P, Q, generator=g) * 0.5).to(dtype).to(cuda).contiguous(memory_format=torch.channels_last)
    w = (torch.randn(Cout, Cin, R, R, generator=g) * 0.1).to(dtype).to(cuda).contiguous(memory_format=torch.channels_last)
    want = torch.nn.grad.conv2d_input((N, Cin, H, W), w.float(), dy.float(), stride=2, padding=pad)
    got = conv2d_raw(1, dy, w, None, N, H, W, Cin, Cout, R, R, 2, pad)
    torch.cuda.synchronize()
    tol = 3e-2 if dtype == torch.bfloat16 else 5e-3
    torch.testing.assert_close(got.float(), want, rtol=tol, atol=tol * (Cout * R * R) ** 0.5 * 0.05)


FIRST_LAYERS = [
    # N, H, W, Cin, Cout, R, stride, pad, bias
    (4, 224, 224, 3, 64, 7, 2, 3, False),      # ResNet stem
    (3, 32, 32, 3, 64, 3, 1, 1, False),        # CIFAR stem / VGG conv1_1 of the perceptual loss
    (2, 40, 36, 3, 64, 5, 1, 2, True),         # the SR net's first layer (Fairscale-DDP.py:74)
    (2, 64, 64, 3, 60, 3, 1, 1, True),         # SwinIR conv_first (Stoke-DDP.py:206): Cout padded to 64
    (2, 19, 23, 4, 24, 3, 2, 1, True),         # four channels, ragged sizes, stride 2
    (1, 16, 16, 1, 8, 3, 1, 0, False),         # one channel, no padding
]


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("N,H,W,Cin,Cout,R,stride,pad,bias", FIRST_LAYERS)
def test_first_layer_convolution_matches_reference(dtype, N, H, W, Cin, Cout, R, stride, pad, bias):
    """tcgen05 first-layer kernels (im2col tile built in shared memory): forward, weight gradient, bias gradient and —
    through the padded-Cin route — the input gradient, against fp32 torch on the same rounded data."""
    from pytorch_distributedtraining_b200.ops import _lib
    from pytorch_distributedtraining_b200.ops.layers import Conv2d
    torch.manual_seed(N * 100 + H)
    conv = Conv2d(Cin, Cout, R, stride, pad, bias=bias).to(cuda)
    x = torch.randn(N, Cin, H, W, device=cuda).to(dtype).float().requires_grad_(True)       # values exactly representable
    before = _lib.LAUNCHES
    with torch.autocast("cuda", dtype=dtype):
        y = conv(x)
    assert _lib.LAUNCHES > before, "the first layer did not run on the kernel library"
    dy = torch.randn_like(y.float())
    y.float().backward(dy)
    gw, gb, gx = conv.weight.grad.clone(), (conv.bias.grad.clone() if bias else None), x.grad.clone()
    conv.zero_grad(); x.grad = None
    w32 = conv.weight.detach().to(dtype).float().requires_grad_(True)
    b32 = conv.bias.detach().clone().requires_grad_(True) if bias else None
    ref = F.conv2d(x, w32, b32, stride=stride, padding=pad)
    ref.backward(dy.to(dtype).float())
    tol = 3e-2 if dtype == torch.bfloat16 else 5e-3
    k = Cin * R * R
    torch.testing.assert_close(y.float(), ref, rtol=tol, atol=tol * k ** 0.5 * 0.1)
    scale_w = max(1.0, float(w32.grad.abs().max()))
    torch.testing.assert_close(gw.float(), w32.grad, rtol=tol, atol=tol * scale_w)
    if bias:
        torch.testing.assert_close(gb.float(), b32.grad, rtol=tol, atol=tol * max(1.0, float(b32.grad.abs().max())))
    torch.testing.assert_close(gx, x.grad, rtol=tol, atol=tol * max(1.0, float(x.grad.abs().max())))


@pytest.mark.parametrize("cin,cout,k,stride", [(64, 32, 3, 1), (32, 12, 3, 1), (64, 40, 3, 2), (128, 72, 3, 1)])
def test_convolutions_outside_the_64_channel_grid_stay_on_the_kernel_library(cin, cout, k, stride):
    """Layers of the SR net (Fairscale-DDP.py:74: 64 -> 32 -> 12 channels) and other widths that miss the kernels'
    64-channel granularity: forward, weight gradient and INPUT gradient all run on the tcgen05 kernels (channels
    zero-padded where needed) and match fp32 torch; the profiler sees no cuDNN kernel."""
    from torch.profiler import ProfilerActivity, profile
    from pytorch_distributedtraining_b200.ops.layers import Conv2d
    torch.manual_seed(cin + cout)
    conv = Conv2d(cin, cout, k, stride, k // 2).to(cuda).to(memory_format=torch.channels_last)
    x = torch.randn(4, cin, 24, 24, device=cuda).half().float().contiguous(memory_format=torch.channels_last).requires_grad_(True)

    def run():
        with torch.autocast("cuda", dtype=torch.float16):
            y = conv(x)
        y.float().square().mean().backward()
        return y

    run(); conv.zero_grad(); x.grad = None
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        y = run()
        torch.cuda.synchronize()
    names = [e.key for e in prof.key_averages()]
    assert not [n for n in names if "cudnn" in n.lower() or "cutlass" in n.lower()], names
    gw, gx = conv.weight.grad.clone(), x.grad.clone()
    conv.zero_grad(); x.grad = None
    w32 = conv.weight.detach().half().float().requires_grad_(True)
    ref = F.conv2d(x, w32, conv.bias.detach(), stride=stride, padding=k // 2)
    ref.square().mean().backward()
    torch.testing.assert_close(y.float(), ref, rtol=5e-3, atol=5e-3 * (cin * k * k) ** 0.5 * 0.1)
    torch.testing.assert_close(gw.float(), w32.grad, rtol=2e-2, atol=2e-2 * max(1e-3, float(w32.grad.abs().max())))
    torch.testing.assert_close(gx, x.grad, rtol=2e-2, atol=2e-2 * max(1e-4, float(x.grad.abs().max())))
