"""ncu launch target for the round-2 kernels: attention (ViT-B/16 and SwinIR-S shapes, forward + backward), first-layer
convolution (ResNet stem, forward + weight gradient), BatchNorm apply / backward-reduce on the 56x56x256 tensor."""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from pytorch_distributedtraining_b200.ops.attention import attention
from pytorch_distributedtraining_b200.ops.conv import conv2d_smallc_raw, conv2d_smallc_wgrad_raw
from pytorch_distributedtraining_b200.ops.layers import BatchNormAct2d
dev = torch.device("cuda")
for (B, H, S, d, wb) in ((128, 12, 197, 64, False), (1152, 6, 64, 10, True)):
    q = torch.randn(B, S, 3 * H * d, device=dev).bfloat16().requires_grad_(True)
    bias = (torch.randn(H, S, S, device=dev) * 0.5).requires_grad_(True) if wb else None
    dout = torch.randn(B, S, H * d, device=dev).bfloat16()
    for _ in range(2):
        q.grad = None
        attention(q, H, d ** -0.5, bias, None).backward(dout)
x = torch.randn(256, 3, 224, 224, device=dev).half().contiguous(memory_format=torch.channels_last)
w = (torch.randn(64, 3, 7, 7, device=dev) * 0.1).half().contiguous(memory_format=torch.channels_last)
dy = torch.randn(256, 64, 112, 112, device=dev).half().contiguous(memory_format=torch.channels_last)
for _ in range(2):
    conv2d_smallc_raw(x, w, None, 256, 224, 224, 3, 64, 7, 7, 2, 3)
    conv2d_smallc_wgrad_raw(x, dy, 256, 224, 224, 3, 64, 7, 7, 2, 3)
bn = BatchNormAct2d(256).to(dev)
t = torch.randn(256, 256, 56, 56, device=dev).half().contiguous(memory_format=torch.channels_last).requires_grad_(True)
r = torch.randn_like(t)
for _ in range(2):
    t.grad = None
    bn(t, residual=r).backward(r)
torch.cuda.synchronize()
print("ok")
