"""Every distinct convolution of ResNet-50 (batch 256, fp16, NHWC) x {fprop, dgrad, wgrad}: this library's kernels vs
cuDNN (torch, cudnn.benchmark on), CUDA-event timed, inputs rotated through buffers larger than L2.
Writes gpurun_out/conv_vs_cudnn.md (copy to profiles/)."""
import collections, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
import torch.nn as nn
import torch.nn.functional as F
from pytorch_distributedtraining_b200.models import resnet50
from pytorch_distributedtraining_b200.ops.layers import Conv2d

dev = torch.device("cuda")
torch.backends.cudnn.benchmark = True
N = 256

# ---- enumerate the layer shapes by running the model once on the meta-free CPU path with hooks
shapes = collections.OrderedDict()
m = resnet50()
_orig_forward = Conv2d.forward
def _recording_forward(self, x, carry=False):          # forward_carry() bypasses module hooks: record in the method itself
    key = (x.shape[2], self.in_channels, self.out_channels, self.kernel_size[0], self.stride[0], self.padding[0])
    shapes[key] = shapes.get(key, 0) + 1
    return _orig_forward(self, x, carry)
Conv2d.forward = _recording_forward
with torch.no_grad():
    m.eval()(torch.zeros(1, 3, 224, 224))
Conv2d.forward = _orig_forward
del m
assert sum(shapes.values()) == 53, shapes


def timeit(fn, iters=12):
    """GPU time per call with the host out of the picture: `iters` calls captured in one CUDA graph, replayed."""
    for _ in range(3): fn()
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        fn()
    torch.cuda.current_stream().wait_stream(side)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(iters): fn()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    g.replay(); g.replay()
    e1.record(); e1.synchronize()
    return e0.elapsed_time(e1) / (2 * iters)


rows = []
for (H, Cin, Cout, R, st, pad), count in shapes.items():
    P = (H + 2 * pad - R) // st + 1
    nb = 3
    xs = [torch.randn(N, Cin, H, H, device=dev).half().contiguous(memory_format=torch.channels_last) for _ in range(nb)]
    dys = [torch.randn(N, Cout, P, P, device=dev).half().contiguous(memory_format=torch.channels_last) for _ in range(nb)]
    conv = Conv2d(Cin, Cout, R, st, pad, bias=False).to(dev).to(memory_format=torch.channels_last)
    w16 = conv.weight.detach().half().contiguous(memory_format=torch.channels_last)
    i = [0]
    def nxt(lst):
        i[0] += 1
        return lst[i[0] % nb]
    need_dx = Cin > 4

    def ours(mode):
        x = nxt(xs).detach().requires_grad_(need_dx and mode == "dgrad")
        conv.weight.requires_grad_(mode == "wgrad")
        with torch.autocast("cuda", dtype=torch.float16):
            y = conv(x)
        if mode == "fprop":
            return
        y.backward(nxt(dys))
        conv.weight.grad = None
        x.grad = None

    # forward-only time, then (forward + one gradient) minus forward, for our kernels (they run behind autograd)
    def t_ours():
        with torch.no_grad():
            f = timeit(lambda: ours("fprop"))
        d = timeit(lambda: ours("dgrad")) - f if need_dx else float("nan")
        wg = timeit(lambda: ours("wgrad")) - f
        return f, d, wg

    def cudnn_f(): F.conv2d(nxt(xs), w16, None, stride=st, padding=pad)
    def cudnn_d(): torch.ops.aten.convolution_backward(nxt(dys), xs[0], w16, None, (st, st), (pad, pad), (1, 1), False, (0, 0), 1, (True, False, False))
    def cudnn_w(): torch.ops.aten.convolution_backward(nxt(dys), nxt(xs), w16, None, (st, st), (pad, pad), (1, 1), False, (0, 0), 1, (False, True, False))
    f, d, wg = t_ours()
    cf, cw = timeit(cudnn_f), timeit(cudnn_w)
    cd = timeit(cudnn_d) if need_dx else float("nan")
    fl = 2.0 * N * P * P * Cout * Cin * R * R / 1e9
    r = dict(H=H, Cin=Cin, Cout=Cout, R=R, stride=st, layers=count, gflop=round(fl, 1), fprop_ours_ms=f, fprop_cudnn_ms=cf,
             dgrad_ours_ms=d, dgrad_cudnn_ms=cd, wgrad_ours_ms=wg, wgrad_cudnn_ms=cw)
    rows.append(r)
    print(json.dumps({k: (round(v, 4) if isinstance(v, float) else v) for k, v in r.items()}), flush=True)
    del xs, dys

os.makedirs("gpurun_out", exist_ok=True)
with open("gpurun_out/conv_vs_cudnn.md", "w") as fh:
    fh.write("# ResNet-50 convolutions (batch 256, fp16 NHWC): this library vs cuDNN, ms per call\n\n"
             "`scripts/bench_conv_table.py`; 12 calls captured in a CUDA graph (no host time), replayed twice between CUDA events; three\n"
             "rotating input sets (each larger than L2 for the big layers).  Ours is timed THROUGH the module + autograd (gradient times are forward+backward minus\n"
             "forward); cuDNN through `F.conv2d` / `aten.convolution_backward` with `cudnn.benchmark = True`.  ratio = cuDNN / ours\n"
             "(> 1: ours faster).\n\n| HxW | Cin | Cout | k | s | layers | fprop ours | cuDNN | ratio | dgrad ours | cuDNN | ratio | wgrad ours | cuDNN | ratio |\n"
             "|---|---|---|---|---|---|---|---|---|---|---|---|---|---|---|\n")
    tot = collections.Counter()
    for r in rows:
        def rt(a, b):
            return f"{b / a:.2f}" if a == a and a > 0 else "-"
        fh.write(f"| {r['H']} | {r['Cin']} | {r['Cout']} | {r['R']} | {r['stride']} | {r['layers']} | {r['fprop_ours_ms']:.3f} | {r['fprop_cudnn_ms']:.3f} | "
                 f"{rt(r['fprop_ours_ms'], r['fprop_cudnn_ms'])} | {r['dgrad_ours_ms']:.3f} | {r['dgrad_cudnn_ms']:.3f} | {rt(r['dgrad_ours_ms'], r['dgrad_cudnn_ms'])} | "
                 f"{r['wgrad_ours_ms']:.3f} | {r['wgrad_cudnn_ms']:.3f} | {rt(r['wgrad_ours_ms'], r['wgrad_cudnn_ms'])} |\n")
        for k in ("fprop", "dgrad", "wgrad"):
            a, b = r[k + "_ours_ms"], r[k + "_cudnn_ms"]
            if a == a:
                tot[k + "_ours"] += a * r["layers"]; tot[k + "_cudnn"] += b * r["layers"]
    fh.write("\nWhole network (sum over all 53 layers, ms): " + ", ".join(
        f"{k} ours {tot[k + '_ours']:.2f} / cuDNN {tot[k + '_cudnn']:.2f} ({tot[k + '_cudnn'] / tot[k + '_ours']:.2f}x)" for k in ("fprop", "dgrad", "wgrad")) + "\n")
print(open("gpurun_out/conv_vs_cudnn.md").read()[-400:])
