"""NVLink peer bandwidth between two GPUs of the box (roofline denominator for the fused collectives): device-to-device
copy (cudaMemcpyPeer under tensor.copy_) and this library's all-gather kernel, 256 MB, CUDA-event timed on the source."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
n = 256 << 20
a = torch.empty(n, dtype=torch.uint8, device="cuda:0")
b = torch.empty(n, dtype=torch.uint8, device="cuda:1")
out = {}
for name, (src, dst) in {"copy_0_to_1": (a, b), "copy_1_to_0": (b, a)}.items():
    dev = src.device
    with torch.cuda.device(dev):
        for _ in range(3): dst.copy_(src, non_blocking=True)
        torch.cuda.synchronize(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10): dst.copy_(src, non_blocking=True)
        e1.record(); e1.synchronize()
        out[name + "_GBs"] = round(n * 10 / (e0.elapsed_time(e1) / 1e3) / 1e9, 1)
print(json.dumps(out))
