"""Convolution throughput on ResNet-50 shapes (batch 256, fp16, NHWC): implicit-GEMM tcgen05 kernel vs cuDNN."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
import torch.nn.functional as F
from pytorch_distributedtraining_b200.ops.conv import conv2d_raw, conv2d_wgrad_raw

dev = torch.device("cuda")
torch.backends.cudnn.benchmark = True
N = 256
shapes = [(56, 64, 64, 3, 1, 1), (28, 128, 128, 3, 1, 1), (14, 256, 256, 3, 1, 1), (7, 512, 512, 3, 1, 1),
          (56, 128, 128, 3, 2, 1), (28, 256, 256, 3, 2, 1), (14, 512, 512, 3, 2, 1),
          (56, 256, 512, 1, 2, 0), (28, 512, 1024, 1, 2, 0), (14, 1024, 2048, 1, 2, 0)]


def timeit(fn, iters=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); e1.synchronize()
    return e0.elapsed_time(e1) / iters


rows = []
for (H, Cin, Cout, R, st, pad) in shapes:
    xs = [torch.randn(N, Cin, H, H, device=dev).half().contiguous(memory_format=torch.channels_last) for _ in range(2)]
    w = (torch.randn(Cout, Cin, R, R, device=dev) * 0.1).half().contiguous(memory_format=torch.channels_last)
    P = (H + 2 * pad - R) // st + 1
    fl = 2.0 * N * P * P * Cout * Cin * R * R
    i = [0]
    def ours():
        i[0] += 1
        return conv2d_raw(0, xs[i[0] % 2], w, None, N, H, H, Cin, Cout, R, R, st, pad)
    def cudnn():
        i[0] += 1
        return F.conv2d(xs[i[0] % 2], w, None, stride=st, padding=pad)
    r = {"H": H, "Cin": Cin, "Cout": Cout, "R": R, "stride": st, "fwd_ours_ms": timeit(ours), "fwd_cudnn_ms": timeit(cudnn)}
    if st == 1:
        dys = [torch.randn(N, Cout, P, P, device=dev).half().contiguous(memory_format=torch.channels_last) for _ in range(2)]
        def ours_d():
            i[0] += 1
            return conv2d_raw(1, dys[i[0] % 2], w, None, N, H, H, Cin, Cout, R, R, 1, pad)
        def cudnn_d():
            i[0] += 1
            return torch.ops.aten.convolution_backward(dys[i[0] % 2], xs[0], w, None, (1, 1), (pad, pad), (1, 1), False, (0, 0), 1,
                                                       (True, False, False))[0]
        r["dgrad_ours_ms"], r["dgrad_cudnn_ms"] = timeit(ours_d), timeit(cudnn_d)
    dyw = torch.randn(N, Cout, P, P, device=dev).half().contiguous(memory_format=torch.channels_last)
    def ours_w():
        i[0] += 1
        return conv2d_wgrad_raw(xs[i[0] % 2], dyw, N, H, H, Cin, Cout, R, R, st, pad)
    def cudnn_w():
        i[0] += 1
        return torch.ops.aten.convolution_backward(dyw, xs[i[0] % 2], w, None, (st, st), (pad, pad), (1, 1), False, (0, 0), 1,
                                                   (False, True, False))[1]
    r["wgrad_ours_ms"], r["wgrad_cudnn_ms"] = timeit(ours_w), timeit(cudnn_w)
    r["fwd_ours_tflops"] = fl / r["fwd_ours_ms"] / 1e9
    r["fwd_cudnn_tflops"] = fl / r["fwd_cudnn_ms"] / 1e9
    rows.append(r)
    print(json.dumps({k: (round(v, 4) if isinstance(v, float) else v) for k, v in r.items()}), flush=True)
os.makedirs("gpurun_out", exist_ok=True)
with open("gpurun_out/conv_bench.jsonl", "w") as fh:
    for r in rows: fh.write(json.dumps(r) + "\n")
