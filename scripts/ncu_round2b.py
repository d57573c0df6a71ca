"""ncu launch target, second round-2 batch: window attention addressed on the image grid (SwinIR-S bench geometry:
18 images of 64 x 64 tokens, 8 x 8 windows shifted by 4, 6 heads of 10, rows padded 180 -> 184 / 60 -> 64), the
short-row LayerNorm kernels on the same token grid, and the bias-gradient column sums."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from pytorch_distributedtraining_b200.ops.attention import window_attention
from pytorch_distributedtraining_b200.ops.layers import LayerNorm
from pytorch_distributedtraining_b200.ops.norm import column_sums

dev = torch.device("cuda")
B, Hi, Wi, ws, shift, H, d = 18, 64, 64, 8, 4, 6, 10
S, C = ws * ws, H * d
qkv = torch.randn(B, Hi, Wi, 184, device=dev).bfloat16().requires_grad_(True)
bias = (torch.randn(H, S, S, device=dev) * 0.5).requires_grad_(True)
mask = torch.where(torch.rand(64, S, S, device=dev) < 0.2, -100.0, 0.0)
dout = torch.randn(B, Hi, Wi, 64, device=dev).bfloat16()
ln = LayerNorm(C).to(dev)
ln.phys = 64
x = torch.randn(B * Hi * Wi, 64, device=dev).bfloat16()
x[:, C:] = 0
x.requires_grad_(True)
g = torch.randn_like(x)
for _ in range(3):
    qkv.grad = None
    window_attention(qkv, H, d, ws, shift, d ** -0.5, bias, mask, 64).backward(dout)
    x.grad = None
    ln(x).backward(g)
    column_sums(g)
torch.cuda.synchronize()
print("ok")
