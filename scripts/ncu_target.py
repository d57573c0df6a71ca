"""Small single-GPU workload that launches each hot kernel a few times on a realistic shape; used plain
(must exit 0) and then under ``ncu --set full`` (B200_PROFILING.md).  Not a benchmark."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
import torch.nn.functional as F
from pytorch_distributedtraining_b200.ops.conv import conv2d_raw, conv2d_wgrad_raw
from pytorch_distributedtraining_b200.ops.gemm import gemm, gemm_fused
from pytorch_distributedtraining_b200.ops import attention as A
from pytorch_distributedtraining_b200.ops.layers import BatchNormAct2d, LayerNorm
from pytorch_distributedtraining_b200.ops.loss import cross_entropy
from pytorch_distributedtraining_b200.ops.optim import FusedAdamW

dev = torch.device("cuda")
torch.manual_seed(0)
for it in range(2):
    # BatchNorm + residual + ReLU, ResNet-50 layer1 output shape (batch 256)
    bn = BatchNormAct2d(256).to(dev)
    x = torch.randn(256, 256, 56, 56, device=dev).half().contiguous(memory_format=torch.channels_last).requires_grad_(True)
    r = torch.randn_like(x).requires_grad_(True)
    y = bn(x, r)
    y.backward(torch.randn_like(y))
    del x, r, y
    # GEMMs: square, ViT fc1 forward / dgrad / wgrad
    a = torch.randn(4096, 4096, device=dev).bfloat16(); b = torch.randn(4096, 4096, device=dev).bfloat16()
    gemm(a, b, 4096, 4096, 4096)
    xa = torch.randn(25216, 768, device=dev).bfloat16(); w = torch.randn(3072, 768, device=dev).bfloat16()
    dy = torch.randn(25216, 3072, device=dev).bfloat16()
    gemm(xa, w, 25216, 3072, 768)
    gemm(dy, w, 25216, 768, 3072, a_mn=False, b_mn=True)
    gemm(dy, xa, 3072, 768, 25216, a_mn=True, b_mn=True, out_dtype=torch.float32)
    del a, b, xa, w, dy
    # implicit-GEMM convolution, ResNet-50 layer2 3x3
    cx = torch.randn(256, 128, 28, 28, device=dev).half().contiguous(memory_format=torch.channels_last)
    cw = (torch.randn(128, 128, 3, 3, device=dev) * 0.1).half().contiguous(memory_format=torch.channels_last)
    cy = conv2d_raw(0, cx, cw, None, 256, 28, 28, 128, 128, 3, 3, 1, 1)
    conv2d_raw(1, cy, cw, None, 256, 28, 28, 128, 128, 3, 3, 1, 1)
    conv2d_wgrad_raw(cx, cy, 256, 28, 28, 128, 128, 3, 3, 1, 1)
    del cx, cw, cy
    # shallow 3x3 layer with the BatchNorm-statistics epilogue (row-grouped kernel), deep layer on the CTA-pair kernel
    sx = torch.randn(256, 64, 56, 56, device=dev).half().contiguous(memory_format=torch.channels_last)
    sw = (torch.randn(64, 64, 3, 3, device=dev) * 0.1).half().contiguous(memory_format=torch.channels_last)
    conv2d_raw(0, sx, sw, None, 256, 56, 56, 64, 64, 3, 3, 1, 1, colstats=torch.zeros(129, device=dev))
    dx_ = torch.randn(256, 256, 14, 14, device=dev).half().contiguous(memory_format=torch.channels_last)
    dw_ = (torch.randn(256, 256, 3, 3, device=dev) * 0.1).half().contiguous(memory_format=torch.channels_last)
    conv2d_raw(0, dx_, dw_, None, 256, 14, 14, 256, 256, 3, 3, 1, 1)
    conv2d_wgrad_raw(sx, sx, 256, 56, 56, 64, 64, 3, 3, 1, 1)
    del sx, sw, dx_, dw_
    # 1x1 convolution as a GEMM with the statistics epilogue; fused MLP epilogues; attention layout pack
    ga = torch.randn(256 * 56 * 56, 64, device=dev).half(); gb = (torch.randn(256, 64, device=dev) * 0.1).half()
    gemm(ga, gb, ga.shape[0], 256, 64, colstats=torch.zeros(513, device=dev))
    del ga, gb
    fa = torch.randn(25216, 768, device=dev).bfloat16(); fb = (torch.randn(3072, 768, device=dev) * 0.1).bfloat16()
    z, _ = gemm_fused(fa, fb, 25216, 3072, 768, 1, bias=torch.zeros(3072, device=dev))
    gemm_fused(fa, (torch.randn(768, 3072, device=dev) * 0.1).bfloat16(), 25216, 3072, 768, 2, aux=z)
    heads = [torch.randn(128, 12, 197, 64, device=dev).bfloat16() for _ in range(3)]
    A._heads_to_tokens(heads, 128, 12, 197, 64)
    del fa, fb, z, heads
    # LayerNorm (ViT), cross-entropy, fused AdamW with fused norm / clip
    ln = LayerNorm(768).to(dev)
    lx = torch.randn(25216, 768, device=dev).bfloat16().requires_grad_(True)
    ln(lx).backward(torch.randn(25216, 768, device=dev).bfloat16())
    logits = torch.randn(256, 1000, device=dev).half().requires_grad_(True)
    cross_entropy(logits, torch.randint(0, 1000, (256,), device=dev)).backward()
    ps = [torch.randn(25_000_000, device=dev).requires_grad_(True)]
    opt = FusedAdamW(ps, lr=1e-3, betas=(0.9, 0.99), weight_decay=1e-4)
    ps[0].grad = torch.randn_like(ps[0])
    opt.clip_grad_norm(1.0)
    opt.step()
    del ps, opt, lx, logits
torch.cuda.synchronize()
print("ok")
