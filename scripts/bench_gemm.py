"""GEMM throughput: tcgen05 kernel vs cuBLAS (torch.matmul), CUDA-event timed, L2 flushed between
iterations by rotating through operand sets larger than L2."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import argparse
import torch
ap = argparse.ArgumentParser(); ap.add_argument("--fwd-only", action="store_true"); ap.add_argument("--out", default="gpurun_out/gemm_bench.jsonl")
args = ap.parse_args()
from pytorch_distributedtraining_b200.ops.gemm import gemm

dev = torch.device("cuda")
shapes = [(8192, 8192, 8192), (4096, 4096, 4096), (25216, 3072, 768), (25216, 768, 3072), (25216, 2304, 768),
          (802816, 64, 256), (802816, 256, 64), (200704, 512, 128), (50176, 1024, 256), (12544, 2048, 512)]
rows = []
for (M, N, K) in shapes:
    for a_mn, b_mn, name in ((False, False, "fwd TN"), (False, True, "dgrad"), (True, True, "wgrad")):
        if args.fwd_only and (a_mn or b_mn):
            continue
        nset = max(2, int(300e6 // ((M * K + N * K + M * N) * 2)) + 1)
        nset = min(nset, 8)
        As = [torch.randn((K, M) if a_mn else (M, K), device=dev).bfloat16() for _ in range(nset)]
        Bs = [torch.randn((K, N) if b_mn else (N, K), device=dev).bfloat16() for _ in range(nset)]
        out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
        def ours(i): gemm(As[i % nset], Bs[i % nset], M, N, K, a_mn=a_mn, b_mn=b_mn, out=out)
        def ref(i):
            A = As[i % nset].t() if a_mn else As[i % nset]
            B = Bs[i % nset] if b_mn else Bs[i % nset].t()
            torch.matmul(A, B, out=out)
        res = {}
        for nm, fn in (("ours", ours), ("cublas", ref)):
            for i in range(3): fn(i)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            iters = 20
            e0.record()
            for i in range(iters): fn(i)
            e1.record(); e1.synchronize()
            ms = e0.elapsed_time(e1) / iters
            res[nm] = ms
        fl = 2.0 * M * N * K
        r = {"M": M, "N": N, "K": K, "kind": name, "ours_ms": res["ours"], "cublas_ms": res["cublas"],
             "ours_tflops": fl / res["ours"] / 1e9, "cublas_tflops": fl / res["cublas"] / 1e9}
        rows.append(r)
        print(json.dumps(r), flush=True)
os.makedirs("gpurun_out", exist_ok=True)
with open(args.out, "w") as fh:
    for r in rows: fh.write(json.dumps(r) + "\n")
