"""First-layer convolution (ResNet stem, batch 256): tcgen05 im2col-in-smem kernels vs cuDNN, forward and weight gradient."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
import torch.nn.functional as F
from pytorch_distributedtraining_b200.ops.conv import conv2d_smallc_raw, conv2d_smallc_wgrad_raw
dev = torch.device("cuda")
torch.backends.cudnn.benchmark = True


def timeit(fn, iters=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); e1.synchronize()
    return e0.elapsed_time(e1) / iters


for (N, H, Cin, Cout, R, st, pad) in ((256, 224, 3, 64, 7, 2, 3), (18, 128, 3, 64, 3, 1, 1), (40, 256, 3, 64, 5, 1, 2)):
    xs = [torch.randn(N, Cin, H, H, device=dev).half().contiguous(memory_format=torch.channels_last) for _ in range(3)]
    w = (torch.randn(Cout, Cin, R, R, device=dev) * 0.1).half().contiguous(memory_format=torch.channels_last)
    P = (H + 2 * pad - R) // st + 1
    dy = torch.randn(N, Cout, P, P, device=dev).half().contiguous(memory_format=torch.channels_last)
    i = [0]
    def nx():
        i[0] += 1
        return xs[i[0] % 3]
    r = {"N": N, "H": H, "R": R, "stride": st,
         "fwd_ours_ms": timeit(lambda: conv2d_smallc_raw(nx(), w, None, N, H, H, Cin, Cout, R, R, st, pad)),
         "fwd_cudnn_ms": timeit(lambda: F.conv2d(nx(), w, None, stride=st, padding=pad)),
         "wgrad_ours_ms": timeit(lambda: conv2d_smallc_wgrad_raw(nx(), dy, N, H, H, Cin, Cout, R, R, st, pad)),
         "wgrad_cudnn_ms": timeit(lambda: torch.ops.aten.convolution_backward(dy, nx(), w, None, (st, st), (pad, pad), (1, 1), False,
                                                                              (0, 0), 1, (False, True, False))[1])}
    print(json.dumps({k: (round(v, 4) if isinstance(v, float) else v) for k, v in r.items()}), flush=True)
