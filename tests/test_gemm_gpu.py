"""tcgen05 GEMM vs a plain PyTorch fp32 matmul of the same 16-bit operands."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
cuda = torch.device("cuda")


@pytest.fixture(scope="module", autouse=True)
def _lib():
    from pytorch_distributedtraining_b200.ops import require_library
    require_library()


def _mk(rows, cols, dtype, seed):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(rows, cols, generator=g) * 0.5).to(dtype).to(cuda)


SHAPES = [(128, 128, 64), (256, 256, 256), (512, 384, 192), (200, 136, 72), (1000, 1000, 520), (8, 8, 8),
          (4096, 768, 3072), (197 * 8, 2304, 768), (130, 264, 4104)]


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("M,N,K", SHAPES)
@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, True), (True, False)])
def test_gemm_matches_fp32_reference(dtype, M, N, K, a_mn, b_mn):
    from pytorch_distributedtraining_b200.ops.gemm import gemm
    a = _mk(K, M, dtype, 1) if a_mn else _mk(M, K, dtype, 1)
    b = _mk(K, N, dtype, 2) if b_mn else _mk(N, K, dtype, 2)
    if (a.shape[1] % 8) or (b.shape[1] % 8):
        pytest.skip("row pitch must be a multiple of 8 elements")
    A = a.float().t() if a_mn else a.float()
    B = b.float().t() if b_mn else b.float()
    want = A @ B.t()
    got = gemm(a, b, M, N, K, a_mn=a_mn, b_mn=b_mn)
    torch.cuda.synchronize()
    tol = 2e-2 if dtype == torch.bfloat16 else 4e-3
    torch.testing.assert_close(got.float(), want, rtol=tol, atol=tol * max(1.0, K ** 0.5 * 0.25))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_gemm_bias_fp32_out_and_accumulate(dtype):
    from pytorch_distributedtraining_b200.ops.gemm import gemm
    M, N, K = 300, 264, 136
    a, b = _mk(M, K, dtype, 3), _mk(N, K, dtype, 4)
    bias = torch.randn(N, device=cuda)
    want = a.float() @ b.float().t() + bias
    got = gemm(a, b, M, N, K, bias=bias, out_dtype=torch.float32)
    torch.testing.assert_close(got, want, rtol=1e-3, atol=1e-3)
    out = torch.ones(M, N, device=cuda)
    gemm(a, b, M, N, K, out=out, accumulate=True)
    torch.testing.assert_close(out, 1.0 + a.float() @ b.float().t(), rtol=1e-3, atol=1e-3)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("shape,out_f", [((4, 197, 768), 3072), ((64, 512), 1000), ((3, 5, 7, 64), 24)])
def test_linear_layer_forward_backward(dtype, shape, out_f):
    from pytorch_distributedtraining_b200.ops.layers import Linear
    torch.manual_seed(0)
    lin = Linear(shape[-1], out_f).to(cuda)
    x = torch.randn(*shape, device=cuda, requires_grad=True)
    xr = x.detach().clone().requires_grad_(True)
    with torch.autocast("cuda", dtype=dtype):
        y = lin(x)
    assert y.dtype == dtype
    # reference: the same 16-bit rounding of inputs, fp32 math
    w16, x16 = lin.weight.detach().to(dtype).float(), xr.to(dtype).float()
    yr = F.linear(x16, w16, lin.bias.detach())
    tol = 3e-2 if dtype == torch.bfloat16 else 5e-3
    torch.testing.assert_close(y.float(), yr, rtol=tol, atol=tol)
    g = torch.randn_like(yr)
    y.backward(g.to(dtype))
    g16 = g.to(dtype).float()
    dx = g16 @ w16
    dw = g16.reshape(-1, out_f).t() @ x16.reshape(-1, shape[-1])
    db = g16.reshape(-1, out_f).sum(0)
    rows = x16.reshape(-1, shape[-1]).shape[0]
    torch.testing.assert_close(x.grad, dx.view_as(x), rtol=tol, atol=tol * out_f ** 0.5)
    torch.testing.assert_close(lin.weight.grad, dw, rtol=tol, atol=tol * rows ** 0.5)
    torch.testing.assert_close(lin.bias.grad, db, rtol=tol, atol=tol * rows ** 0.5)
    assert lin.weight.grad.dtype == torch.float32


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("M,N,K", [(64, 256, 100352), (256, 64, 50176), (768, 768, 25216), (2048, 512, 12544), (8, 16, 4096)])
def test_gemm_split_k_wgrad_shapes(dtype, M, N, K):
    """Few output tiles + long reduction: the kernel splits K and combines with TMA reduce-add."""
    from pytorch_distributedtraining_b200.ops.gemm import gemm
    a, b = _mk(K, M, dtype, 7), _mk(K, N, dtype, 8)          # both MN-major, like dW = dY^T X
    want = a.float().t() @ b.float()
    got = gemm(a, b, M, N, K, a_mn=True, b_mn=True, out_dtype=torch.float32)
    torch.testing.assert_close(got, want, rtol=2e-3, atol=2e-3 * K ** 0.5 * 0.25)
    out = torch.full((M, N), 2.0, device=cuda)
    gemm(a, b, M, N, K, a_mn=True, b_mn=True, out=out, accumulate=True)
    torch.testing.assert_close(out, want + 2.0, rtol=2e-3, atol=2e-3 * K ** 0.5 * 0.25)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("M,N,K,world", [(4096, 3072, 768, 8), (1000, 768, 3072, 4), (300, 512, 264, 2), (25216, 2304, 768, 8)])
def test_gemm_with_fused_gather_of_row_sharded_weights(dtype, M, N, K, world):
    """W4 on one GPU: the shards are separate local buffers; the kernel must assemble W tile by tile
    (flag per tile) while it computes, starting with the tiles "this rank" owns."""
    from pytorch_distributedtraining_b200.ops.gemm import gemm_gather
    x, w = _mk(M, K, dtype, 11), _mk(N, K, dtype, 12)
    bias = torch.randn(N, device=cuda)
    rpr = (N + world - 1) // world
    shards = [w[r * rpr:(r + 1) * rpr].clone() for r in range(world)]
    shards = [s if s.shape[0] == rpr else torch.cat([s, torch.zeros(rpr - s.shape[0], K, dtype=dtype, device=cuda)]) for s in shards]
    want = x.float() @ w.float().t() + bias
    for rank in (0, world - 1):
        w_full = torch.full((N, K), float("nan"), dtype=dtype, device=cuda)
        got = gemm_gather(x, shards, w_full, rank, world, rpr, bias=bias)
        torch.cuda.synchronize()
        assert torch.equal(w_full, w), "the gathered copy of W must be exact"
        tol = 2e-2 if dtype == torch.bfloat16 else 4e-3
        torch.testing.assert_close(got.float(), want, rtol=tol, atol=tol * max(1.0, K ** 0.5 * 0.25))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("M,N,K,bias", [(8192, 4096, 512, False), (25216, 3072, 768, True), (4500, 4360, 136, True),
                                        (19456, 256, 64, False)])
def test_gemm_cta_pair_kernel(dtype, M, N, K, bias):
    """Shapes large enough for the cta_group::2 kernel (256 x 256 tiles per CTA pair): ragged M / N / K
    tails, with and without the bias epilogue."""
    from pytorch_distributedtraining_b200.ops.gemm import gemm
    a, b = _mk(M, K, dtype, 5), _mk(N, K, dtype, 6)
    bv = torch.randn(N, device=cuda) if bias else None
    want = a.float() @ b.float().t()
    if bias:
        want = want + bv
    got = gemm(a, b, M, N, K, bias=bv)
    torch.cuda.synchronize()
    tol = 2e-2 if dtype == torch.bfloat16 else 4e-3
    torch.testing.assert_close(got.float(), want, rtol=tol, atol=tol * max(1.0, K ** 0.5 * 0.25))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("M,N,K,a_mn,b_mn,f32", [
    (25216, 768, 3072, False, True, False),      # dgrad of fc1 (weights read MN-major)
    (8192, 2304, 768, False, True, False),
    (3072, 768, 25216, True, True, True),        # wgrad of fc1: 36 tiles, split along K, fp32 reduce-add
    (768, 3072, 12000, True, True, True),
    (2304, 768, 4104, True, True, True),
    (4096, 4096, 520, True, False, False),
])
def test_gemm_cta_pair_kernel_operand_layouts(dtype, M, N, K, a_mn, b_mn, f32):
    from pytorch_distributedtraining_b200.ops.gemm import gemm
    a = _mk(K, M, dtype, 7) if a_mn else _mk(M, K, dtype, 7)
    b = _mk(K, N, dtype, 8) if b_mn else _mk(N, K, dtype, 8)
    A = a.float().t() if a_mn else a.float()
    B = b.float().t() if b_mn else b.float()
    want = A @ B.t()
    got = gemm(a, b, M, N, K, a_mn=a_mn, b_mn=b_mn, out_dtype=torch.float32 if f32 else None)
    torch.cuda.synchronize()
    tol = (2e-3 if f32 else (2e-2 if dtype == torch.bfloat16 else 4e-3))
    torch.testing.assert_close(got.float(), want, rtol=tol, atol=tol * max(1.0, K ** 0.5 * 0.25))
    if f32:
        # accumulate into an existing gradient
        got2 = gemm(a, b, M, N, K, a_mn=a_mn, b_mn=b_mn, out=got.clone(), out_dtype=torch.float32, accumulate=True)
        torch.cuda.synchronize()
        torch.testing.assert_close(got2, 2 * want, rtol=2 * tol, atol=2 * tol * max(1.0, K ** 0.5 * 0.25))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_gemm_fused_epilogues(dtype):
    """GELU-dual, dGELU and residual-add epilogues of the CTA-pair kernel vs the unfused fp32 formulation."""
    from pytorch_distributedtraining_b200.ops.gemm import fused_supported, gemm_fused
    M, N, K = 25216, 3072, 768
    assert fused_supported(M, N) and not fused_supported(128, 128)
    a, b = _mk(M, K, dtype, 21), (_mk(N, K, dtype, 22) * 0.2)
    bias = torch.randn(N, device=cuda) * 0.1
    z_ref = a.float() @ b.float().t() + bias
    z, g = gemm_fused(a, b, M, N, K, 1, bias=bias)
    tol = 2e-2 if dtype == torch.bfloat16 else 4e-3
    torch.testing.assert_close(z.float(), z_ref, rtol=tol, atol=tol * 4)
    torch.testing.assert_close(g.float(), F.gelu(z_ref), rtol=tol, atol=tol * 4)
    res = _mk(M, N, dtype, 23)
    y = gemm_fused(a, b, M, N, K, 3, bias=bias, aux=res)
    torch.testing.assert_close(y.float(), z_ref + res.float(), rtol=tol, atol=tol * 4)
    # dgrad form: dz[M, K'] = (dy[M, N'] W[N', K']) * gelu'(zsaved[M, K']) with W read MN-major
    dy, w = _mk(M, 768, dtype, 24), (_mk(768, N, dtype, 25) * 0.2)
    zs = _mk(M, N, dtype, 26)
    zf = zs.float().requires_grad_(True)
    F.gelu(zf).backward(dy.float() @ w.float())
    dz = gemm_fused(dy, w, M, N, 768, 2, aux=zs)
    torch.testing.assert_close(dz.float(), zf.grad, rtol=tol, atol=tol * 4)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_fused_transformer_mlp_and_residual_stream(dtype):
    """One ViT encoder block with the fused MLP / residual epilogues / LayerNorm carry vs the same block in
    fp32 torch ops: output and every gradient."""
    from pytorch_distributedtraining_b200.models.vit import EncoderBlock
    torch.manual_seed(11)
    blk = EncoderBlock(768, 12, 3072).to(cuda)
    x0 = torch.randn(128, 197, 768, device=cuda) * 0.5
    x = x0.clone().requires_grad_(True)
    with torch.autocast("cuda", dtype=dtype):
        y = blk(x)
    gy = torch.randn_like(y) * 0.1
    y.backward(gy)
    got = {n: p.grad.float().clone() for n, p in blk.named_parameters()}
    gx = x.grad.float().clone()

    def ref_block(t):
        h = F.layer_norm(t, (768,), blk.ln_1.weight, blk.ln_1.bias, 1e-6)
        B, S, D = h.shape
        qkv = F.linear(h, blk.self_attention.in_proj.weight, blk.self_attention.in_proj.bias).view(B, S, 3, 12, 64).permute(2, 0, 3, 1, 4)
        o = F.scaled_dot_product_attention(qkv[0], qkv[1], qkv[2]).transpose(1, 2).reshape(B, S, D)
        t = t + F.linear(o, blk.self_attention.out_proj.weight, blk.self_attention.out_proj.bias)
        h = F.layer_norm(t, (768,), blk.ln_2.weight, blk.ln_2.bias, 1e-6)
        return t + F.linear(F.gelu(F.linear(h, blk.mlp.fc1.weight, blk.mlp.fc1.bias)), blk.mlp.fc2.weight, blk.mlp.fc2.bias)

    blk.zero_grad(set_to_none=True)
    xr = x0.clone().requires_grad_(True)
    yr = ref_block(xr)
    yr.backward(gy.float())
    tol = 4e-2 if dtype == torch.bfloat16 else 1e-2
    torch.testing.assert_close(y.float(), yr, rtol=tol, atol=tol)
    torch.testing.assert_close(gx, xr.grad, rtol=tol, atol=tol * float(xr.grad.abs().max()))
    for n, p in blk.named_parameters():
        ref = p.grad.float()
        torch.testing.assert_close(got[n], ref, rtol=tol, atol=tol * float(ref.abs().max()) + 1e-4), n
