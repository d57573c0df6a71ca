"""ncu target for the CTA-pair GEMM and its fused epilogues on the ViT fc1 / fc2 shapes (run plain first)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from pytorch_distributedtraining_b200.ops.gemm import gemm, gemm_fused

dev = torch.device("cuda")
torch.manual_seed(0)
M, N, K = 25216, 3072, 768
a = torch.randn(M, K, device=dev).bfloat16(); b = (torch.randn(N, K, device=dev) * 0.1).bfloat16()
bias = torch.randn(N, device=dev)
dy = torch.randn(M, K, device=dev).bfloat16(); w2 = (torch.randn(K, N, device=dev) * 0.1).bfloat16()
z = torch.randn(M, N, device=dev).bfloat16()
for _ in range(2):
    gemm(a, b, M, N, K, bias=bias)                      # plain forward
    gemm_fused(a, b, M, N, K, 1, bias=bias)             # GELU dual
    gemm_fused(dy, w2, M, N, K, 2, aux=z)               # dGELU
    gemm_fused(a, b, M, N, K, 3, bias=bias, aux=z)      # residual add
    gemm(z, a, N, K, M, a_mn=True, b_mn=True, out_dtype=torch.float32)   # wgrad, split-K
torch.cuda.synchronize()
print("ok")
