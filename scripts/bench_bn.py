"""Per-shape timing of the fused BatchNorm kernels over the ResNet-50 layer shapes (batch 256, fp16 NHWC):
achieved bytes/s vs the measured copy bandwidth.  CUDA events, rotating buffers larger than L2."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from pytorch_distributedtraining_b200.ops.layers import BatchNormAct2d

dev = torch.device("cuda")
shapes = [(64, 112, False), (64, 56, False), (256, 56, True), (128, 28, False), (512, 28, True), (256, 14, False),
          (1024, 14, True), (512, 7, False), (2048, 7, True)]
N = 256
peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
for C, H, res in shapes:
    nset = max(2, int(400e6 // (N * C * H * H * 2)) + 1)
    bn = BatchNormAct2d(C).to(dev)
    xs = [torch.randn(N, C, H, H, device=dev).half().contiguous(memory_format=torch.channels_last).requires_grad_(True) for _ in range(nset)]
    rs = [torch.randn(N, C, H, H, device=dev).half().contiguous(memory_format=torch.channels_last).requires_grad_(True) for _ in range(nset)] if res else None
    g = torch.randn(N, C, H, H, device=dev).half().contiguous(memory_format=torch.channels_last)
    def fwd(i): return bn(xs[i % nset], rs[i % nset] if res else None)
    for i in range(3): fwd(i).backward(g)
    torch.cuda.synchronize()
    e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    iters = 10
    tf = tb = 0.0
    for i in range(iters):
        e[0].record(); y = fwd(i); e[1].record(); y.backward(g); e[2].record()
        e[2].synchronize()
        tf += e[0].elapsed_time(e[1]); tb += e[1].elapsed_time(e[2])
    nbytes = N * C * H * H * 2
    fwd_bytes = nbytes * (3 + (1 if res else 0)) + nbytes / 16
    bwd_bytes = nbytes * (2 + 2 + 1 + (1 if res else 0)) + 2 * nbytes / 16
    print(f"C={C:5d} H={H:4d} res={int(res)} tensor={nbytes/1e6:7.1f}MB fwd {tf/iters*1e3:7.1f}us {fwd_bytes/(tf/iters)/1e6:6.0f}GB/s ({fwd_bytes/(tf/iters)/1e6/peak:.2f}) "
          f"bwd {tb/iters*1e3:7.1f}us {bwd_bytes/(tb/iters)/1e6:6.0f}GB/s ({bwd_bytes/(tb/iters)/1e6/peak:.2f})", flush=True)
