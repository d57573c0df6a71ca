"""ncu launch target: the attention kernels on the ViT-B/16 and SwinIR-S shapes (forward + backward, 3 rounds)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from pytorch_distributedtraining_b200.ops.attention import attention
dev = torch.device("cuda")
for (B, H, S, d, wb) in ((128, 12, 197, 64, False), (1152, 6, 64, 10, True)):
    q = torch.randn(B, S, 3 * H * d, device=dev).bfloat16().requires_grad_(True)
    bias = (torch.randn(H, S, S, device=dev) * 0.5).requires_grad_(True) if wb else None
    dout = torch.randn(B, S, H * d, device=dev).bfloat16()
    for _ in range(3):
        q.grad = None
        attention(q, H, d ** -0.5, bias, None).backward(dout)
torch.cuda.synchronize()
print("ok")
