"""tcgen05 attention (csrc/attn_sm100.cu) against an fp32 PyTorch reference of the same op:
softmax(q k^T * scale + bias + mask) v, forward and backward, on the shapes the models use."""
import pytest
import torch

pytestmark = pytest.mark.gpu
cuda = torch.device("cuda") if torch.cuda.is_available() else None


def _reference(qkv, heads, scale, bias, mask):
    """fp32 math on the same (16-bit rounded) inputs."""
    B, S, D3 = qkv.shape
    d = D3 // (3 * heads)
    q, k, v = qkv.float().view(B, S, 3, heads, d).permute(2, 0, 3, 1, 4).unbind(0)      # [B, H, S, d]
    s = (q @ k.transpose(-1, -2)) * scale
    if bias is not None:
        s = s + bias.float().unsqueeze(0)
    if mask is not None:
        nW = mask.shape[0]
        s = s + mask.float().view(1, nW, 1, S, S).expand(B // nW, -1, heads, -1, -1).reshape(B, heads, S, S)
    p = torch.softmax(s, dim=-1)
    return (p @ v).transpose(1, 2).reshape(B, S, heads * d)


CASES = [
    # B, H, S, d, bias, nW (0: no mask), dtype
    (4, 12, 197, 64, False, 0, torch.bfloat16),      # ViT-B/16: TMA path, two query tiles, 208 staged keys
    (3, 2, 129, 64, False, 0, torch.float16),        # just over one tile
    (2, 3, 256, 64, False, 0, torch.bfloat16),       # the envelope's edge
    (16, 6, 64, 10, True, 4, torch.bfloat16),        # SwinIR-S windows: head dim 10, bias + shift mask, stacked pairs
    (7, 6, 64, 30, True, 0, torch.float16),          # SwinIR-M head dim, odd window count (half-empty last tile)
    (8, 4, 49, 32, True, 4, torch.bfloat16),         # 7x7 windows
    (5, 2, 17, 32, False, 0, torch.bfloat16),        # vit_tiny
    (3, 2, 100, 32, False, 0, torch.bfloat16),       # single tile, thread-staged operands
]


@pytest.mark.parametrize("B,H,S,d,with_bias,nW,dtype", CASES)
def test_attention_matches_fp32_reference(B, H, S, d, with_bias, nW, dtype):
    from pytorch_distributedtraining_b200.ops.attention import attention, attention_supported
    g = torch.Generator(device="cpu").manual_seed(B * 1000 + S + d)
    qkv = torch.randn(B, S, 3 * H * d, generator=g).to(cuda).to(dtype).requires_grad_(True)
    bias = (torch.randn(H, S, S, generator=g) * 0.5).to(cuda).requires_grad_(True) if with_bias else None
    mask = None
    if nW:
        mask = torch.where(torch.rand(nW, S, S, generator=g) < 0.2, -100.0, 0.0).to(cuda)
        mask[:, torch.arange(S), torch.arange(S)] = 0.0            # never a fully masked row
    scale = d ** -0.5
    assert attention_supported(qkv, H, bias)
    out = attention(qkv, H, scale, bias, mask)
    dout = torch.randn(B, S, H * d, generator=g).to(cuda).to(dtype)
    out.backward(dout)
    dq, db = qkv.grad.clone(), (bias.grad.clone() if with_bias else None)

    q32 = qkv.detach().float().requires_grad_(True)
    b32 = bias.detach().float().clone().requires_grad_(True) if with_bias else None
    ref = _reference(q32, H, scale, b32, mask)
    ref.backward(dout.float())
    tol = 2e-2 if dtype == torch.bfloat16 else 4e-3
    torch.testing.assert_close(out.float(), ref, rtol=tol, atol=tol)
    gtol = 4e-2 if dtype == torch.bfloat16 else 8e-3
    torch.testing.assert_close(dq.float(), q32.grad, rtol=gtol, atol=gtol * max(1.0, float(q32.grad.abs().max())) * 0.25)
    if with_bias:
        # the bias gradient sums over every window: compare relative to its scale
        err = (db.float() - b32.grad).abs().max().item()
        assert err <= gtol * max(1.0, b32.grad.abs().max().item()), err


WIN_CASES = [
    # images, Hi, Wi, ws, shift, heads, d, qkv pitch pad, out pitch pad, dtype
    (3, 16, 24, 8, 0, 6, 10, 0, 0, torch.bfloat16),        # SwinIR-S block, plain windows
    (3, 16, 24, 8, 4, 6, 10, 4, 4, torch.bfloat16),        # shifted windows + mask, rows zero-padded 180 -> 184 / 60 -> 64
    (2, 64, 64, 8, 4, 6, 10, 4, 4, torch.bfloat16),        # the bench geometry
    (1, 14, 21, 7, 3, 4, 32, 0, 0, torch.float16),         # 7 x 7 windows, odd window count
    (5, 8, 8, 8, 0, 2, 16, 8, 0, torch.float16),           # one window per image
]


@pytest.mark.parametrize("B,Hi,Wi,ws,shift,H,d,qpad,opad,dtype", WIN_CASES)
def test_window_attention_matches_roll_partition_reference(B, Hi, Wi, ws, shift, H, d, qpad, opad, dtype):
    """Window attention addressed on the image grid (no roll / partition copies) against the fp32 chain
    roll -> window_partition -> softmax(q k^T * scale + bias + mask) v -> window_reverse -> roll."""
    from pytorch_distributedtraining_b200.ops.attention import window_attention, window_attention_supported
    g = torch.Generator(device="cpu").manual_seed(Hi * 100 + Wi + d)
    C, S = H * d, ws * ws
    P, OP = 3 * C + qpad, C + opad
    qkv = torch.randn(B, Hi, Wi, P, generator=g).to(cuda).to(dtype).requires_grad_(True)
    bias = (torch.randn(H, S, S, generator=g) * 0.5).to(cuda).requires_grad_(True)
    nW = (Hi // ws) * (Wi // ws)
    mask = None
    if shift:
        mask = torch.where(torch.rand(nW, S, S, generator=g) < 0.2, -100.0, 0.0).to(cuda)
        mask[:, torch.arange(S), torch.arange(S)] = 0.0
    scale = d ** -0.5
    assert window_attention_supported(qkv, H, d, ws)
    out = window_attention(qkv, H, d, ws, shift, scale, bias, mask, OP)
    assert out.shape == (B, Hi, Wi, OP)
    if opad:
        assert (out[..., C:] == 0).all()
    dout = torch.randn(B, Hi, Wi, OP, generator=g).to(cuda).to(dtype)
    out.backward(dout)
    dq, db = qkv.grad.clone(), bias.grad.clone()
    if qpad:
        assert (dq[..., 3 * C:] == 0).all()

    q32 = qkv.detach().float().requires_grad_(True)
    b32 = bias.detach().float().clone().requires_grad_(True)
    x = q32[..., :3 * C]
    if shift:
        x = torch.roll(x, shifts=(-shift, -shift), dims=(1, 2))
    xw = x.reshape(B, Hi // ws, ws, Wi // ws, ws, 3 * C).permute(0, 1, 3, 2, 4, 5).reshape(-1, S, 3 * C)
    o = _reference(xw, H, scale, b32, mask)
    o = o.reshape(B, Hi // ws, Wi // ws, ws, ws, C).permute(0, 1, 3, 2, 4, 5).reshape(B, Hi, Wi, C)
    if shift:
        o = torch.roll(o, shifts=(shift, shift), dims=(1, 2))
    o.backward(dout.float()[..., :C])
    tol = 2e-2 if dtype == torch.bfloat16 else 4e-3
    torch.testing.assert_close(out.float()[..., :C], o, rtol=tol, atol=tol)
    gtol = 4e-2 if dtype == torch.bfloat16 else 8e-3
    torch.testing.assert_close(dq.float(), q32.grad, rtol=gtol, atol=gtol * max(1.0, float(q32.grad.abs().max())) * 0.25)
    err = (db.float() - b32.grad).abs().max().item()
    assert err <= gtol * max(1.0, b32.grad.abs().max().item()), err


def test_attention_twice_is_bit_identical_and_graph_capturable():
    from pytorch_distributedtraining_b200.ops.attention import attention
    qkv = torch.randn(8, 197, 3 * 12 * 64, device=cuda).bfloat16()
    a = attention(qkv, 12)
    b = attention(qkv, 12)
    assert torch.equal(a, b)
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        attention(qkv, 12)
    torch.cuda.current_stream().wait_stream(s)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        c = attention(qkv, 12)
    graph.replay()
    torch.cuda.synchronize()
    assert torch.equal(a, c)


def test_models_use_the_attention_kernel():
    from pytorch_distributedtraining_b200.models import vit_tiny
    from pytorch_distributedtraining_b200.ops import _lib
    m = vit_tiny().to(cuda)
    x = torch.randn(4, 3, 32, 32, device=cuda)
    before = _lib.LAUNCHES
    with torch.autocast("cuda", dtype=torch.bfloat16):
        y = m(x)
    y.float().sum().backward()
    torch.cuda.synchronize()
    assert _lib.LAUNCHES > before
    import pytorch_distributedtraining_b200.ops.attention as A
    calls = {"n": 0}
    orig = A._AttentionFn.apply

    def spy(*a, **k):
        calls["n"] += 1
        return orig(*a, **k)

    A._AttentionFn.apply = spy
    try:
        import pytorch_distributedtraining_b200.models.vit as V
        V.attention = lambda *a, **k: (calls.__setitem__("n", calls["n"] + 1), A.attention(*a, **k))[1]
        with torch.autocast("cuda", dtype=torch.bfloat16):
            m(x)
    finally:
        A._AttentionFn.apply = orig
        V.attention = A.attention
    assert calls["n"] >= 2          # both encoder blocks went through the kernel
