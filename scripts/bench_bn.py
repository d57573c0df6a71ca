"""BatchNorm kernels on the ResNet-50 shapes (batch 256, fp16, NHWC): achieved bytes/s against a plain copy."""
import ctypes, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from pytorch_distributedtraining_b200.ops._lib import require_library
from pytorch_distributedtraining_b200.ops.fused import DT, Workspace, ptr, stream_ptr
lib = require_library()
dev = torch.device("cuda")


def timeit(fn, iters=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); e1.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3


a = torch.empty(1 << 29, dtype=torch.float16, device=dev); b = torch.empty_like(a)
t = timeit(lambda: b.copy_(a))
print(json.dumps({"copy_GBs": round(2 * a.numel() * 2 / t / 1e3, 1)}))
del a, b
N = 256
for (H, C, res) in ((56, 256, True), (56, 64, False), (28, 512, True), (28, 128, False), (14, 1024, True), (112, 64, False), (7, 2048, True)):
    rows = N * H * H
    nb = 3
    xs = [torch.randn(rows, C, device=dev).half() for _ in range(nb)]
    rs = [torch.randn(rows, C, device=dev).half() for _ in range(nb)]
    ys = [torch.empty(rows, C, device=dev).half() for _ in range(nb)]
    ds = [torch.empty(rows, C, device=dev).half() for _ in range(nb)]
    mask = torch.empty(rows * (C // 8), dtype=torch.uint8, device=dev)
    w, bb = torch.ones(C, device=dev), torch.zeros(C, device=dev)
    rm, rv = torch.zeros(C, device=dev), torch.ones(C, device=dev)
    mean, invstd = torch.empty(C, device=dev), torch.empty(C, device=dev)
    sums = torch.zeros(2 * C + 1, device=dev); dsums = torch.zeros(2 * C + 1, device=dev)
    ws = Workspace.get(dev)
    scratch = ws.scratch(int(lib.b200dp_bn_scratch_floats(ctypes.c_int64(rows), ctypes.c_int(C))))
    dt = ctypes.c_int(DT[torch.float16]); st = stream_ptr(dev)
    i = [0]

    def red():
        i[0] += 1
        lib.b200dp_bn_reduce(ptr(xs[i[0] % nb]), ctypes.c_int64(rows), C, dt, ptr(scratch), ptr(ws.counters), ptr(sums), st)

    def app():
        i[0] += 1; k = i[0] % nb
        lib.b200dp_bn_apply(ptr(xs[k]), ptr(rs[k]) if res else None, ptr(ys[k]), ctypes.c_int64(rows), C, dt, ptr(sums), ctypes.c_float(rows),
                            ptr(w), ptr(bb), ctypes.c_float(1e-5), ctypes.c_float(0.1), ptr(rm), ptr(rv), ptr(mean), ptr(invstd), 1, ptr(mask), st)

    def bred():
        i[0] += 1; k = i[0] % nb
        lib.b200dp_bn_bwd_reduce(ptr(ys[k]), None, ptr(xs[k]), ctypes.c_int64(rows), C, dt, ptr(mean), ptr(invstd), 1, ptr(scratch),
                                 ptr(ws.counters), ptr(dsums), ptr(mask), st)

    def bapp():
        i[0] += 1; k = i[0] % nb
        lib.b200dp_bn_bwd_apply(ptr(ys[k]), None, ptr(xs[k]), ptr(ds[k]), ptr(rs[k]) if res else None, ctypes.c_int64(rows), C, dt, ptr(mean),
                                ptr(invstd), ptr(w), ptr(dsums), ctypes.c_float(rows), 1, ptr(mask), st)

    red(); torch.cuda.synchronize()
    T = rows * C * 2 / 1e3   # KB -> us*GB/s factor: bytes/1e3/us = GB/s
    m = rows * C / 8 / 1e3
    out = {"H": H, "C": C, "residual": res, "MB": round(rows * C * 2 / 1e6)}
    for name, fn, nbytes in (("reduce", red, T), ("apply", app, (3 if res else 2) * T + m), ("bwd_reduce", bred, 2 * T + m),
                             ("bwd_apply", bapp, (4 if res else 3) * T + m)):
        us = timeit(fn)
        out[name + "_us"] = round(us, 1); out[name + "_GBs"] = round(nbytes / us, 0)
    print(json.dumps(out), flush=True)
