#!/usr/bin/env python
"""BASELINE.json config 5: all-reduce / reduce-scatter bandwidth sweep, 1 KB - 1 GB, peer-memory
kernels vs NCCL (`dist.all_reduce`, `dist.reduce_scatter_tensor`) on the same box.

    torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 benchmarks/collectives_sweep.py

Device-timed with CUDA events, max over ranks; prints one JSON line per (op, size) on rank 0 and a
markdown table at the end.  busbw follows the nccl-tests convention
(all-reduce: 2(W-1)/W * bytes / t; reduce-scatter: (W-1)/W * bytes / t).
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist


def timed(fn, iters, warmup, dev):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize(dev)
    dist.barrier(device_ids=[dev.index])
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    e1.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())  # ms


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--max-bytes", type=int, default=1 << 30)
    ap.add_argument("--dtype", default="float32", choices=["float32", "bfloat16", "float16"])
    ap.add_argument("--out", default="gpurun_out/collectives_sweep.jsonl")
    a = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", rank)))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    from pytorch_distributedtraining_b200.ops import collectives as C
    from pytorch_distributedtraining_b200.parallel.symm import PeerArena
    dtype = getattr(torch, a.dtype)
    esz = torch.empty((), dtype=dtype).element_size()
    arena = PeerArena(dist.group.WORLD, dev, a.max_bytes + (64 << 20), max_blocks=128, timeout_s=30.0)
    ctx = arena.context("main")
    big = arena.alloc(a.max_bytes // esz, dtype)
    big.tensor.normal_()
    plain = torch.randn(a.max_bytes // esz, device=dev).to(dtype)
    rs_out = torch.empty(a.max_bytes // esz // world + 8, device=dev, dtype=torch.float32)
    rs_out_nccl = torch.empty(a.max_bytes // esz // world + 8, device=dev, dtype=dtype)
    rows = []
    size = 1024
    while size <= a.max_bytes:
        n = size // esz
        n = max(n // (8 * world) * (8 * world), 8 * world)
        nbytes = n * esz
        iters = 200 if nbytes <= (1 << 20) else (50 if nbytes <= (64 << 20) else 10)
        blocks = int(min(int(os.environ.get("SWEEP_BLOCKS", "96")), max(1, nbytes // 16384)))
        res = {"bytes": nbytes, "world": world, "dtype": a.dtype}
        res["nccl_allreduce_ms"] = timed(lambda: dist.all_reduce(plain[:n]), iters, 5, dev)
        variants = [("p2p_oneshot", 0, 0)] if nbytes <= (1 << 20) else []
        variants.append(("p2p_twoshot", 1, 0))
        if arena.mc_base:
            variants.append(("nvls", 1, big.mc))
        for name, algo, mc in variants:
            res[f"peer_{name}_allreduce_ms"] = timed(
                lambda: ctx.allreduce(big.table, mc, n, dtype, pre_mult=1.0 / world, algo=algo, blocks=blocks), iters, 5, dev)
        n_per = n // world
        res["nccl_reduce_scatter_ms"] = timed(lambda: dist.reduce_scatter_tensor(rs_out_nccl[:n_per], plain[:n]), iters, 5, dev)
        res["peer_p2p_reduce_scatter_ms"] = timed(
            lambda: ctx.reduce_scatter(big.table, 0, n_per, dtype, C.local_table(rank, rs_out), pre_mult=1.0 / world, blocks=blocks), iters, 5, dev)
        if arena.mc_base:
            res["peer_nvls_reduce_scatter_ms"] = timed(
                lambda: ctx.reduce_scatter(big.table, big.mc, n_per, dtype, C.local_table(rank, rs_out), pre_mult=1.0 / world, blocks=blocks), iters, 5, dev)
        rows.append(res)
        if rank == 0:
            print(json.dumps(res), flush=True)
        size *= 4
    if rank == 0:
        os.makedirs(os.path.dirname(a.out), exist_ok=True)
        with open(a.out, "w") as fh:
            for r in rows:
                fh.write(json.dumps(r) + "\n")
        f = 2 * (world - 1) / world
        print("\n| bytes | NCCL AR us | peer best AR us | peer/NCCL speedup | NCCL busbw GB/s | peer busbw GB/s | NCCL RS us | peer best RS us |")
        print("|---|---|---|---|---|---|---|---|")
        for r in rows:
            peer = min(v for k, v in r.items() if k.startswith("peer_") and k.endswith("allreduce_ms"))
            prs = min(v for k, v in r.items() if k.startswith("peer_") and k.endswith("reduce_scatter_ms"))
            print(f"| {r['bytes']} | {r['nccl_allreduce_ms']*1e3:.1f} | {peer*1e3:.1f} | {r['nccl_allreduce_ms']/peer:.2f}x | "
                  f"{f*r['bytes']/r['nccl_allreduce_ms']/1e6:.1f} | {f*r['bytes']/peer/1e6:.1f} | "
                  f"{r['nccl_reduce_scatter_ms']*1e3:.1f} | {prs*1e3:.1f} |")
    assert not arena.timed_out()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
