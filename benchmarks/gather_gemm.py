#!/usr/bin/env python
"""W4: all-gather of row-sharded weights fused into the consuming GEMM vs the two-call baseline
(NCCL all_gather_into_tensor + cuBLAS matmul), plus NCCL + this repo's plain tcgen05 GEMM so the gain from the
fusion itself is visible next to the gap between the two GEMM kernels.  ViT-B/16 linear shapes, bf16,
device-timed over 30 back-to-back calls (operands stay L2-resident for every arm), max over ranks.

    torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 benchmarks/gather_gemm.py
"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", rank)))
torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
from pytorch_distributedtraining_b200.ops.gemm import gemm, gemm_gather
from pytorch_distributedtraining_b200.parallel.symm import PeerArena

arena = PeerArena(dist.group.WORLD, dev, 256 << 20, timeout_s=30.0)
ctx = arena.context("main")
M = 128 * 197
rows = []
for (N, K) in ((3072, 768), (768, 3072), (2304, 768), (768, 768)):
    if N % (128 * world):
        continue
    rpr = N // world
    x = torch.randn(M, K, device=dev).bfloat16()
    shard = arena.alloc(rpr * K, torch.bfloat16)
    shard.tensor.normal_()
    w_full = torch.empty(N, K, dtype=torch.bfloat16, device=dev)
    w_nccl = torch.empty(N, K, dtype=torch.bfloat16, device=dev)
    local = shard.tensor.view(rpr, K)

    def fused():
        return gemm_gather(x, shard.ptrs, w_full, rank, world, rpr, ctx=ctx)

    def baseline():
        dist.all_gather_into_tensor(w_nccl, local)
        return x @ w_nccl.t()

    def two_call_own():
        dist.all_gather_into_tensor(w_nccl, local)
        return gemm(x, w_nccl, M, N, K)

    res = {}
    for name, fn in (("fused", fused), ("nccl_cublas", baseline), ("nccl_own", two_call_own)):
        for _ in range(5):
            fn()
        torch.cuda.synchronize(); dist.barrier(device_ids=[dev.index])
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(30):
            fn()
        e1.record(); e1.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / 30], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        res[name] = float(t)
    ya, yb = fused(), baseline()
    err = float((ya.float() - yb.float()).abs().max())
    r = {"M": M, "N": N, "K": K, "world": world, "fused_us": res["fused"] * 1e3, "nccl_cublas_us": res["nccl_cublas"] * 1e3,
         "nccl_own_gemm_us": res["nccl_own"] * 1e3, "speedup_vs_own_two_call": res["nccl_own"] / res["fused"],
         "speedup": res["nccl_cublas"] / res["fused"], "tflops_fused": 2.0 * M * N * K / res["fused"] / 1e9, "max_abs_diff": err}
    rows.append(r)
    if rank == 0:
        print(json.dumps(r), flush=True)
if rank == 0:
    os.makedirs("gpurun_out", exist_ok=True)
    with open(f"gpurun_out/gather_gemm_{world}gpu.jsonl", "w") as fh:
        for r in rows:
            fh.write(json.dumps(r) + "\n")
assert not arena.timed_out()
os._exit(0)
