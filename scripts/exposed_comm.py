#!/usr/bin/env python
"""Exposed-communication ms per forward+backward for a bench configuration (eager mode):
    torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 scripts/exposed_comm.py --config resnet50_ddp_amp
Prints one JSON line from rank 0: step time with and without gradient synchronisation and their difference."""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
import bench
from pytorch_distributedtraining_b200.utils.timing import exposed_comm_ms

ap = argparse.ArgumentParser()
ap.add_argument("--config", default="resnet50_ddp_amp")
ap.add_argument("--iters", type=int, default=10)
a = ap.parse_args()
world = int(os.environ.get("WORLD_SIZE", "1"))
args = argparse.Namespace(cpu=not torch.cuda.is_available(), batch_size=0, impl="ours", no_graph=True, sync_bn=False)
cfg = dict(bench.CONFIGS[a.config])
stoke, step, device = bench.build_ours(cfg, args, world)
B = cfg["batch"]
x = torch.randn(B, *cfg["shape"], device=device)
if cfg["model"].startswith("resnet") and device.type == "cuda":
    x = x.contiguous(memory_format=torch.channels_last)
y = torch.randint(0, cfg["classes"], (B,), device=device)


def fwd_bwd():
    stoke.wrapped_model.zero_grad(set_to_none=False)
    loss = stoke.loss(stoke.model(x), y)
    stoke.backward(loss)


r = exposed_comm_ms(stoke.wrapped_model, fwd_bwd, iters=a.iters)
if int(os.environ.get("RANK", "0")) == 0:
    print(json.dumps({"config": a.config, "world": world, **r}), flush=True)
stoke.reset_graphs()
torch.cuda.synchronize() if torch.cuda.is_available() else None
if world > 1:
    import torch.distributed as dist
    dist.barrier()
    dist.destroy_process_group()
