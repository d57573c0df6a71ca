"""Attention kernel vs the library kernel (F.scaled_dot_product_attention) on the two model shapes, forward and
forward+backward, CUDA-event timed; inputs rotate through buffers larger than L2."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
import torch.nn.functional as F
from pytorch_distributedtraining_b200.ops.attention import attention, attention_reference

dev = torch.device("cuda")


def timeit(fn, iters=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); e1.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3


rows = []
for name, (B, H, S, d, with_bias, nW) in {"vit_b16 B=128": (128, 12, 197, 64, False, 0),
                                          "swinir_s windows B=18": (18 * 64, 6, 64, 10, True, 64),
                                          "swin d=32 windows": (18 * 64, 6, 64, 32, True, 64)}.items():
    nbuf = 6
    qs = [torch.randn(B, S, 3 * H * d, device=dev).bfloat16().requires_grad_(True) for _ in range(nbuf)]
    dout = torch.randn(B, S, H * d, device=dev).bfloat16()
    bias = (torch.randn(H, S, S, device=dev) * 0.5).requires_grad_(True) if with_bias else None
    mask = torch.zeros(nW, S, S, device=dev) if nW else None
    scale = d ** -0.5
    i = [0]

    def nxt():
        i[0] += 1
        return qs[i[0] % nbuf]

    r = {"shape": name, "B": B, "H": H, "S": S, "d": d}
    with torch.no_grad():
        r["fwd_ours_us"] = timeit(lambda: attention(nxt(), H, scale, bias, mask))
        r["fwd_lib_us"] = timeit(lambda: attention_reference(nxt(), H, scale, bias, mask))

    def fb(fn):
        q = nxt()
        q.grad = None
        fn(q, H, scale, bias, mask).backward(dout)

    r["fwdbwd_ours_us"] = timeit(lambda: fb(attention))
    r["fwdbwd_lib_us"] = timeit(lambda: fb(attention_reference))
    r["fwd_speedup"] = r["fwd_lib_us"] / r["fwd_ours_us"]
    r["fwdbwd_speedup"] = r["fwdbwd_lib_us"] / r["fwdbwd_ours_us"]
    rows.append(r)
    print(json.dumps({k: (round(v, 2) if isinstance(v, float) else v) for k, v in r.items()}), flush=True)
os.makedirs("gpurun_out", exist_ok=True)
with open("gpurun_out/attention_bench.jsonl", "w") as fh:
    for r in rows: fh.write(json.dumps(r) + "\n")
