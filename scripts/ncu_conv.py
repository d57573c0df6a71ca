"""ncu target: 3x3 stride-1 convolutions of ResNet-50 layer1 / layer2 at batch 256 (run plain first)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from pytorch_distributedtraining_b200.ops.conv import conv2d_raw

dev = torch.device("cuda")
torch.manual_seed(0)
for (H, C) in ((56, 64), (28, 128)):
    x = torch.randn(256, C, H, H, device=dev).half().contiguous(memory_format=torch.channels_last)
    w = (torch.randn(C, C, 3, 3, device=dev) * 0.1).half().contiguous(memory_format=torch.channels_last)
    for _ in range(2):
        y = conv2d_raw(0, x, w, None, 256, H, H, C, C, 3, 3, 1, 1)
        conv2d_raw(1, y, w, None, 256, H, H, C, C, 3, 3, 1, 1)
torch.cuda.synchronize()
print("ok")
